"""Prints the metrics we read from an `ncu --page raw --csv` export, one block per profiled launch."""
import csv
import sys

WANT = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_warps', 'launch__waves_per_multiprocessor',
        'sm__inst_executed.sum', 'smsp__inst_executed.avg.per_cycle_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__cycles_active.avg',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum', 'lts__t_bytes.sum']
STALL = 'smsp__average_warps_issue_stalled_'


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("====")
        for w in WANT:
            if w in idx:
                print("  %-72s %s %s" % (w, r[idx[w]], units[idx[w]]))
        st = [(float(r[i].replace(',', '') or 0), h) for h, i in idx.items()
              if (h.startswith('smsp__average_warp_latency_issue_stalled') or h.startswith('smsp__average_warps_issue_stalled')) and h.endswith('per_issue_active.ratio') and r[i]]
        for v, h in sorted(st, reverse=True)[:7]:
            print("  stall %-66s %.2f" % (h.replace('smsp__average_warp_latency_issue_stalled_', '').replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v))


if __name__ == "__main__":
    main(sys.argv[1])

#!/usr/bin/env python
"""Per-source-line hot spots of one kernel from an `ncu --set full --import-source on` report:
   python tools/ncu_hotspots.py gpurun_out/x.ncu-rep [top N] > profiles/<round>_<kernel>_hotspots.txt
Reads `ncu -i rep --page source --print-source cuda,sass --csv` (the kernels are compiled with -lineinfo) and prints, per
CUDA source line, its share of the warp-stall samples and of the executed warp instructions, plus the main launch metrics."""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    lines, fname, hdr = [], "", None
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif len(r) > 8 and r[0] == "Line No":
            hdr = r
        elif hdr and len(r) > 8 and r[0].isdigit():
            s = hdr.index("Warp Stall Sampling (All Samples)"); e = hdr.index("Instructions Executed")
            x = hdr.index("L1 Wavefronts Shared Excessive") if "L1 Wavefronts Shared Excessive" in hdr else None
            try:
                lines.append((int(r[s] or 0), int(r[e] or 0), int(r[x] or 0) if x is not None else 0, fname, int(r[0]), r[1].strip()))
            except ValueError:
                pass
    ts = sum(l[0] for l in lines) or 1
    ti = sum(l[1] for l in lines) or 1
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    if len(rr) >= 3:
        h, v = rr[0], rr[-1]
        want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
                "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct", "sm__inst_executed.sum",
                "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
                "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
                "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
                "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
                "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]
        for k in want:
            if k in h:
                print("%-86s %s" % (k, v[h.index(k)]))
    print("\nstall samples %d, executed warp instructions %d; top %d source lines by stall samples" % (ts, ti, top))
    print("%7s %7s %9s  %s" % ("stall%", "instr%", "smem_exc", "file:line  source"))
    for l in sorted(lines, reverse=True)[:top]:
        print("%6.1f%% %6.1f%% %9d  %s:%d  %s" % (100.0 * l[0] / ts, 100.0 * l[1] / ti, l[2], l[3], l[4], l[5][:120]))


if __name__ == "__main__":
    main()

"""Joins an `ncu --page source --csv` SASS export with nvdisasm line info: per source line instruction and
stall-sample totals.  usage: sass_lines.py <src.csv> <lib.so> <kernel substring>"""
import csv
import re
import subprocess
import sys
import collections


def main(src_csv, so, kern):
    import os
    so = os.path.abspath(so)
    cub = "/tmp/_sl.cubin"
    names = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout.split()
    elf = [n for n in names if n.endswith(".cubin") and n.startswith("orb_kernels.sm")][0]
    subprocess.run(["cuobjdump", "-xelf", elf, so], cwd="/tmp", check=True, capture_output=True)
    subprocess.run(["mv", "/tmp/" + elf, cub], check=True)
    dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
    lines, cur, infn = [], 0, False
    for ln in dis:
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infn = kern in m.group(1)
            continue
        if not infn:
            continue
        m = re.search(r'//## File ".*?", line (\d+)', ln)
        if m:
            if "inlined at" not in ln or True:
                cur = int(m.group(1))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
            lines.append(cur)
    rows = list(csv.reader(open(src_csv)))
    h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    hdr = rows[h]
    ie, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
    end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] == "Kernel Name"), len(rows))
    body = [r for r in rows[h + 1:end] if len(r) > ie and r[ie].isdigit()]
    if len(body) != len(lines):
        print("warning: %d csv instructions vs %d disassembled" % (len(body), len(lines)))
    agg = collections.defaultdict(lambda: [0, 0])
    for r, l in zip(body, lines):
        agg[l][0] += int(r[ie]); agg[l][1] += int(r[sm] or 0)
    tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
    srcfile = open("/root/repo/my-slam_b200/csrc/orb_kernels.cu").read().splitlines()
    print("total warp instructions %d, samples %d" % (tot, ts))
    for l in sorted(agg):
        n, s = agg[l]
        if n > tot * 0.005 or s > ts * 0.005:
            print("%5d %5.1f%% inst %5.1f%% samp | %s" % (l, 100 * n / tot, 100 * s / max(ts, 1), srcfile[l - 1][:100] if 0 < l <= len(srcfile) else ""))


if __name__ == "__main__":
    main(*sys.argv[1:4])

#!/usr/bin/env python
"""Per-kernel SASS instruction mix of liborbgpu.so (cuobjdump -sass), so that the instruction-level claims of DESIGN.md
(IDP.2A / IDP.4A / PRMT in resize and blur, VIMNMX3 / VABSDIFF4 / LDGSTS in FAST, POPC / LOP3 / VIMNMX.U16x2 in the matcher,
no tensor-core or TMA mnemonics anywhere) can be checked without a GPU.  usage: python tools/sass_mix.py > profiles/<tag>_sass_mix.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "my-slam_b200", "liborbgpu.so")
KERNELS = ["k_resize_level", "k_fast_cells", "k_octree", "k_blur", "k_describe", "k_match_count_pairs", "k_match_bruteforce",
           "k_stereo_match", "k_pipe_peak"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    cur, mix = None, collections.defaultdict(collections.Counter)
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur:
            mix[cur][m.group(1).rstrip(";")] += 1
    print("SASS instruction mix per kernel (static counts), liborbgpu.so built with -gencode arch=compute_100a,code=sm_100a\n")
    special = ("HMMA", "UTC", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HGMMA", "QGMMA", "IGMMA")
    for k in KERNELS:
        for fn, c in mix.items():
            if k + "I" in fn or re.search(r"\d+" + k + r"[A-Z0-9]", fn) or (k in fn and fn.startswith("_Z") and fn[2:].lstrip("0123456789").startswith(k)):
                tot = sum(c.values())
                print("%s  (%s)  %d instructions" % (k, fn, tot))
                print("   " + "  ".join("%s %d" % kv for kv in c.most_common(18)))
                tc = {m: n for m, n in c.items() if m.startswith(special)}
                print("   tensor-core / TMA mnemonics: %s\n" % (tc if tc else "none"))
    allm = collections.Counter()
    for c in mix.values():
        allm.update(c)
    print("whole library: tensor-core / TMA mnemonics:", {m: n for m, n in allm.items() if m.startswith(special)} or "none",
          "; LDGSTS (cp.async):", allm.get("LDGSTS.E.BYPASS.LTC128B", 0) + sum(n for m, n in allm.items() if m.startswith("LDGSTS")))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""What "bit-exact against the reference" means at src/ORBextractor.cc:686 (TEST INFRASTRUCTURE; writes tie_sensitivity.json).

DistributeOctTree sorts pair<int, ExtractorNode*>: among nodes of equal size the order -- hence which node is split first
when the feature budget runs out -- follows the nodes' HEAP ADDRESSES.  The oracle pins that to "most recently created
first" by serving operator new from an ascending bump arena (oracle/ref_wrap.cc, arena mode 1); the port and the CUDA
path implement that rule.  This script runs the unmodified reference extractor three ways on the same frames:
  ascending arena (the pin) / descending arena (the opposite tie-break) / the system allocator (plain glibc malloc)
and counts (a) tie-sensitive frames (ascending != descending) and (b) frames where the pin equals the system-allocator run."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g   # noqa: E402
import oracle_lib             # noqa: E402


def main():
    pkg = g.load_package()
    out = {"what": __doc__.split("\n")[0], "generator": "tests/golden/tie_sensitivity.py", "workloads": {}}
    for name, count in (("C1_tum_640x480_nf1000", 256), ("C2_kitti_1241x376_nf2000", 48), ("C3_euroc_752x480_nf1000", 64), ("C4_hd_1920x1080_nf4000", 12)):
        w, h, nf, base, _ = pkg.synth.CONFIGS[name]
        ref = oracle_lib.OracleRef(nf, 1.2, 8, 20, 7)
        tie = same_malloc = 0
        diff_kp = []
        for i in range(count):
            img = pkg.synth.synth(w, h, base + i)
            ka, da = ref.extract(img, arena_mode=1)
            kd, dd = ref.extract(img, arena_mode=2)
            km, dm = ref.extract(img, arena_mode=0)
            if ka.tobytes() != kd.tobytes() or not np.array_equal(da, dd):
                tie += 1
                diff_kp.append(int(len(set(map(bytes, ka.view(np.uint8).reshape(len(ka), -1))) ^
                                       set(map(bytes, kd.view(np.uint8).reshape(len(kd), -1)))) // 2))
            if ka.tobytes() == km.tobytes() and np.array_equal(da, dm):
                same_malloc += 1
        out["workloads"][name] = {"frames": count, "tie_sensitive_frames": tie, "pin_equals_system_allocator_frames": same_malloc,
                                  "keypoints_differing_between_the_two_tie_breaks_median": float(np.median(diff_kp)) if diff_kp else 0.0,
                                  "keypoints_per_frame": nf}
        print(name, out["workloads"][name], flush=True)
    json.dump(out, open(os.path.join(HERE, "tie_sensitivity.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

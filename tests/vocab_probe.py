#!/usr/bin/env python
"""Driver for profiling the bag-of-words rows under ncu (uses tests/test_vocab.py's scene builders, hence under tests/):
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/x.csv python tests/vocab_probe.py
A vocabulary of ORBvoc.txt's shape (k = 10, L = 6), 1000 and 4000 descriptors through transform, 2000 resident keyframes queried."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g          # noqa: E402
import test_vocab as tv               # noqa: E402

pkg = g.load_package()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
rng = np.random.default_rng(77)
nn = (10 ** 7 - 1) // 9
parent = ((np.arange(nn, dtype=np.int64) - 1) // 10).astype(np.int32)
parent[0] = 0
leaf = (np.arange(nn) >= (10 ** 6 - 1) // 9).astype(np.uint8)
voc = pkg.ORBVocabulary(10, 6, parent, leaf, rng.integers(0, 256, (nn, 32), dtype=np.uint8), np.where(leaf > 0, rng.uniform(0.1, 9.0, nn), 0.0))
for n in (1000, 4000):
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    for _ in range(reps):
        voc.transform(d, 4)
kfs, erased, q, connected, neighbours, prev, min_score = tv.db_scene(5, 10 ** 6, n_kf=2000, reloc=True)
db = pkg.KeyFrameDatabase()
for kf in kfs:
    db.add(*kf)
for _ in range(reps):
    db.query(*q)

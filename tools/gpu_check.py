"""Developer check on a GPU box: stage-by-stage comparison of the CUDA path with the oracle port."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
import oracle_lib

pkg = g.load_package()
synth = pkg.synth


def check(img, nf, label):
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7)
    po = oracle_lib.OraclePort(nf, 1.2, 8, 20, 7)
    t = time.time(); kps, desc = ex(img); dt = time.time() - t
    okps, odesc = po.extract(img)
    ok = len(kps) == len(okps) and kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc)
    print("%s: gpu %d kps, oracle %d kps, match=%s (%.1f ms first call)" % (label, len(kps), len(okps), ok, dt * 1e3))
    if ok:
        return True
    for l in range(8):
        pg = ex.pyramid_level(l); pp = po.level_pixels(l)
        bg = ex.debug_blurred(l); bp = po.level_pixels(l, blurred=True)
        cg = ex.debug_candidates(l); cp = po.level_candidates(l)
        sg = ex.debug_selected(l); sp = po.level_selected(l)
        print("  L%d pyr diff %d | blur diff %s | cand %d vs %d equal=%s | sel %d vs %d equal=%s" % (
            l, int((pg != pp).sum()), "n/a" if bp is None else int((bg != bp).sum()),
            len(cg), len(cp), cg.shape == cp.shape and bool((cg == cp).all()),
            len(sg), len(sp), sg.shape == sp.shape and bool((sg == sp).all())))
    n = min(len(kps), len(okps))
    if n:
        for f in kps.dtype.names:
            d = np.nonzero(kps[f][:n] != okps[f][:n])[0]
            if len(d):
                print("  field %s differs at %d kps, first %d: %r vs %r" % (f, len(d), d[0], kps[f][d[0]], okps[f][d[0]]))
        dd = np.nonzero((desc[:n] != odesc[:n]).any(1))[0]
        print("  descriptors differ at %d kps" % len(dd))
    return False


if __name__ == "__main__":
    allok = True
    for name, (w, h, nf, base, _) in synth.CONFIGS.items():
        allok &= check(synth.synth(w, h, base), nf, name)
    allok &= check(synth.flat(640, 480), 1000, "flat")
    allok &= check(synth.lowtex(640, 480, 7), 1000, "lowtex")
    allok &= check(synth.noise(640, 480, 9), 1000, "noise")
    allok &= check(synth.synth(100, 100, 3), 1000, "tiny100")
    # batch timing
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    fr = np.stack([synth.synth(640, 480, 0x5EED0000 + i) for i in range(64)])
    ex.extract_batch(fr)
    ex.set_profiling(True)
    t = time.time(); k, d, n = ex.extract_batch(fr); dt = time.time() - t
    print("batch 64 pageable: %.2f ms -> %.0f frames/s; stage ms %s" % (dt * 1e3, 64 / dt, ex.stage_ms()))
    print("ALL OK" if allok else "MISMATCHES")

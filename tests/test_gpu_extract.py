"""GPU parity tests proper: the CUDA path, called through the C ABI (ctypes), against the oracle and
the committed golden results of the unmodified reference extractor.  Bit-exact bar: keypoints,
octaves, responses, angles (bitwise, which is stricter than the north-star 1e-4 deg) and descriptors."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(__file__)
GOLD = json.load(open(os.path.join(HERE, "golden", "extract_golden.json")))
PRIMS = np.load(os.path.join(HERE, "golden", "prims_golden.npz"))


def frame_for(synth, case):
    fam, w, h, seed = case["family"], case["w"], case["h"], case["seed"]
    if fam == "synth": return synth.synth(w, h, seed)
    if fam == "flat": return synth.flat(w, h)
    if fam == "lowtex": return synth.lowtex(w, h, seed)
    return synth.noise(w, h, seed)


@pytest.fixture(scope="module")
def ex1000(pkg):
    return pkg.ORBextractor(1000, 1.2, 8, 20, 7)


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: "%s-%x" % (c["name"], c["seed"]))
def test_matches_reference_golden(pkg, case):
    """Every committed golden case (all four BASELINE shapes + edge families), full image sizes."""
    img = frame_for(pkg.synth, case)
    ex = pkg.ORBextractor(case["nfeatures"], 1.2, 8, 20, 7)
    kps, desc = ex(img)
    assert len(kps) == case["n"]
    assert hashlib.sha256(kps.tobytes()).hexdigest() == case["kps_sha256"]
    assert hashlib.sha256(desc.tobytes()).hexdigest() == case["desc_sha256"]
    if case["n"]:
        assert case["n"] <= ex.max_keypoints(case["w"], case["h"])
        assert (kps["class_id"] == -1).all() and (np.diff(kps["octave"]) >= 0).all()     # level-ascending concatenation


@pytest.mark.parametrize("shape", [(640, 480, 1000), (1241, 376, 2000), (333, 217, 700)])
def test_every_stage_matches_oracle(pkg, oracle, shape):
    """Pyramid bytes, blurred bytes, FAST candidate list (order included) and octree survivors per level."""
    w, h, nf = shape
    img = pkg.synth.synth(w, h, 12345)
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7); po = oracle.OraclePort(nf, 1.2, 8, 20, 7)
    kps, desc = ex(img); okps, odesc = po.extract(img)
    for l in range(8):
        assert np.array_equal(ex.pyramid_level(l), po.level_pixels(l)), "pyramid level %d" % l
        ob = po.level_pixels(l, blurred=True)
        if ob is not None:
            assert np.array_equal(ex.debug_blurred(l), ob), "blurred level %d" % l
        assert np.array_equal(ex.debug_candidates(l), po.level_candidates(l)), "FAST candidates level %d" % l
        assert np.array_equal(ex.debug_selected(l), po.level_selected(l)), "octree survivors level %d" % l
    assert kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc)


@pytest.mark.parametrize("shape", [(641, 480), (642, 361), (643, 200), (752, 480), (1241, 376), (401, 333), (130, 131), (203, 149)])
def test_blur_row_ends_and_pyramid_for_every_width_residue(pkg, oracle, shape):
    """The row-end columns of the blur (word + PRMT path with REFLECT_101) and the resize tails: level widths of every residue
    modulo 4 and 16 (the eight levels of each shape give 64 different widths), noise images so that every byte matters."""
    w, h = shape
    img = pkg.synth.noise(w, h, 77 + w)
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7); po = oracle.OraclePort(500, 1.2, 8, 20, 7)
    ex(img); po.extract(img)
    for l in range(8):
        assert np.array_equal(ex.pyramid_level(l), po.level_pixels(l)), "pyramid level %d" % l
        ob = po.level_pixels(l, blurred=True)
        if ob is not None:
            gb = ex.debug_blurred(l)
            assert np.array_equal(gb, ob), ("blurred level %d" % l, np.argwhere(gb != ob)[:4])


def test_angles_within_north_star_tolerance_and_bitwise(pkg, oracle):
    img = pkg.synth.synth(640, 480, 999)
    kps, _ = pkg.ORBextractor(1000, 1.2, 8, 20, 7)(img)
    okps, _ = oracle.OraclePort(1000, 1.2, 8, 20, 7).extract(img)
    assert np.abs(kps["angle"].astype(np.float64) - okps["angle"].astype(np.float64)).max() <= 1e-4   # north_star tolerance
    assert kps["angle"].tobytes() == okps["angle"].tobytes()                                        # and in fact bit-exact
    assert ((kps["angle"] >= 0) & (kps["angle"] < 360)).all()


def test_other_parameter_sets_match_oracle(pkg, oracle):
    img = pkg.synth.synth(480, 360, 11)
    for (nf, sc, nl, ini, mn) in [(500, 1.2, 4, 20, 7), (800, 1.5, 5, 30, 10), (300, 1.1, 8, 12, 5), (50, 1.2, 8, 20, 7),
                                  (3000, 1.2, 8, 20, 7), (1000, 1.2, 8, 7, 20), (1000, 1.2, 1, 20, 7)]:
        kps, desc = pkg.ORBextractor(nf, sc, nl, ini, mn)(img)
        okps, odesc = oracle.OraclePort(nf, sc, nl, ini, mn).extract(img)
        assert kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc), (nf, sc, nl, ini, mn)


def test_constructor_tables(pkg, ex1000):
    t = GOLD["tables_nf1000"]
    assert [float(x).hex() for x in ex1000.GetScaleFactors()] == t["scale_hex"]
    assert [float(x).hex() for x in ex1000.GetInverseScaleFactors()] == t["inv_scale_hex"]
    assert [float(x).hex() for x in ex1000.GetScaleSigmaSquares()] == t["sigma2_hex"]
    assert [float(x).hex() for x in ex1000.GetInverseScaleSigmaSquares()] == t["inv_sigma2_hex"]
    assert ex1000.features_per_level().tolist() == t["features_per_level"]
    assert ex1000.umax().tolist() == t["umax"]
    assert ex1000.GetLevels() == 8 and abs(ex1000.GetScaleFactor() - 1.2) < 1e-6


def test_batch_equals_single_and_is_deterministic(pkg, ex1000):
    synth = pkg.synth
    frames = np.stack([synth.synth(640, 480, 0x5EED0000 + i) for i in range(7)] + [synth.flat(640, 480)] +
                      [synth.noise(640, 480, 5), synth.lowtex(640, 480, 6)])
    kps, desc, n = ex1000.extract_batch(frames)
    kps2, desc2, n2 = ex1000.extract_batch(frames)
    assert np.array_equal(n, n2)
    for i in range(len(frames)):
        k1, d1 = ex1000(frames[i])
        assert n[i] == len(k1)
        assert kps[i, :n[i]].tobytes() == k1.tobytes() and np.array_equal(desc[i, :n[i]], d1)
        assert kps2[i, :n[i]].tobytes() == k1.tobytes() and np.array_equal(desc2[i, :n[i]], d1)
    assert n[7] == 0


def test_strided_rows_and_frames(pkg, ex1000):
    """cv::Mat may be non-contiguous (Mat.step): row stride and frame stride larger than the image."""
    synth = pkg.synth
    big = np.zeros((3, 500, 700), np.uint8)
    for i in range(3):
        big[i, 10:490, 30:670] = synth.synth(640, 480, 70 + i)
    view = big[:, 10:490, 30:670]
    cap = ex1000.max_keypoints(640, 480)
    kps = np.zeros((3, cap), pkg.KP_DTYPE); desc = np.zeros((3, cap, 32), np.uint8); n = np.zeros(3, np.int32)
    ex1000.extract_batch_ptr(view.ctypes.data, 3, 640, 480, view.strides[1], view.strides[0], kps.ctypes.data, desc.ctypes.data, cap, n.ctypes.data)
    for i in range(3):
        k1, d1 = ex1000(np.ascontiguousarray(view[i]))
        assert kps[i, :n[i]].tobytes() == k1.tobytes() and np.array_equal(desc[i, :n[i]], d1)
        k2, d2 = ex1000(view[i])                       # single-frame call with a row stride
        assert k2.tobytes() == k1.tobytes() and np.array_equal(d2, d1)


def test_device_resident_and_pinned_paths(pkg, ex1000):
    import torch
    synth = pkg.synth
    frames = np.stack([synth.synth(640, 480, 300 + i) for i in range(5)])
    ref_k, ref_d, ref_n = ex1000.extract_batch(frames)
    cap = ex1000.max_keypoints(640, 480)
    # device-resident
    d_fr = torch.from_numpy(frames).cuda()
    d_k = torch.zeros((5, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((5, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(5, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ex1000.extract_batch_ptr(d_fr.data_ptr(), 5, 640, 480, 640, 640 * 480, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr(), device=True)
    n = d_n.cpu().numpy(); k = d_k.cpu().numpy(); d = d_d.cpu().numpy()
    assert np.array_equal(n, ref_n)
    for i in range(5):
        assert k[i, :n[i]].tobytes() == ref_k[i, :n[i]].tobytes() and np.array_equal(d[i, :n[i]], ref_d[i, :n[i]])
    # pinned host buffers, used in place
    p_fr = torch.from_numpy(frames).pin_memory()
    p_k = torch.zeros((5, cap, 28), dtype=torch.uint8).pin_memory(); p_d = torch.zeros((5, cap, 32), dtype=torch.uint8).pin_memory()
    p_n = torch.zeros(5, dtype=torch.int32).pin_memory()
    ex1000.extract_batch_ptr(p_fr.data_ptr(), 5, 640, 480, 640, 640 * 480, p_k.data_ptr(), p_d.data_ptr(), cap, p_n.data_ptr())
    n = p_n.numpy(); k = p_k.numpy(); d = p_d.numpy()
    assert np.array_equal(n, ref_n)
    for i in range(5):
        assert k[i, :n[i]].tobytes() == ref_k[i, :n[i]].tobytes() and np.array_equal(d[i, :n[i]], ref_d[i, :n[i]])


def test_frame_queue_submit_collect(pkg):
    synth = pkg.synth
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    depth = ex.queue_depth()
    assert depth >= 2
    frames = [synth.synth(640, 480, 500 + i) for i in range(depth + 4)]
    expect = [ex(f) for f in frames]
    tickets = [ex.submit(frames[i]) for i in range(depth)]
    with pytest.raises(pkg.OrbGpuError):
        ex.submit(frames[0])                                     # ring full: error code, no blocking, no crash
    for i in range(depth, len(frames)):                         # steady state: collect oldest, submit next
        k, d = ex.collect(tickets[i - depth], 640, 480)
        assert k.tobytes() == expect[i - depth][0].tobytes() and np.array_equal(d, expect[i - depth][1])
        tickets.append(ex.submit(frames[i]))
    for i in range(len(frames) - depth, len(frames))[::-1]:     # drain out of order
        k, d = ex.collect(tickets[i], 640, 480)
        assert k.tobytes() == expect[i][0].tobytes() and np.array_equal(d, expect[i][1])
    with pytest.raises(pkg.OrbGpuError):
        ex.collect(tickets[0], 640, 480)                        # already collected
    # mvImagePyramid of the last collected frame
    assert np.array_equal(ex.pyramid_level(0), frames[len(frames) - depth])


def test_pyramid_levels_and_border(pkg, oracle, ex1000):
    """mvImagePyramid contract (stereo matcher reads it, src/Frame.cc:563-580): level pixels + 19-px REFLECT_101 frame."""
    img = pkg.synth.synth(640, 480, 31337)
    ex1000(img)
    assert np.array_equal(ex1000.pyramid_level(0), img)
    if oracle.have_ref():
        ref = oracle.OracleRef(1000, 1.2, 8, 20, 7)
        levels = ref.pyramid(img)
        for l in range(8):
            got = ex1000.pyramid_level(l)
            assert np.array_equal(got, levels[l])
            hh, ww = levels[l].shape
            assert np.array_equal(ex1000.pyramid_level(l, bordered=True), ref.pyramid_level_bordered(l, ww, hh))


def test_edge_cases(pkg, ex1000):
    L = pkg.lib()
    # empty image: outputs untouched, n_out = -1 (the reference returns silently, src/ORBextractor.cc:1048-1049)
    n = C.c_int(7)
    kps = np.full(4, 3.0, np.float32)
    assert L.orbgpu_extract(ex1000._h, None, 0, 0, 0, kps.ctypes.data, kps.ctypes.data, 0, C.byref(n)) == 0
    assert n.value == -1 and (kps == 3.0).all()
    assert ex1000(np.zeros((0, 0), np.uint8)) == (None, None)
    # flat image: zero keypoints (descriptors.release() path, :1066-1067)
    k, d = ex1000(pkg.synth.flat(640, 480))
    assert len(k) == 0 and d.shape == (0, 32)
    # capacity too small -> error code, not a crash
    img = pkg.synth.synth(640, 480, 1)
    small_k = np.zeros(10, pkg.KP_DTYPE); small_d = np.zeros((10, 32), np.uint8)
    assert L.orbgpu_extract(ex1000._h, img.ctypes.data, 640, 480, 640, small_k.ctypes.data, small_d.ctypes.data, 10, C.byref(n)) == -2
    # images too small for one 30-px cell on some / all levels
    for (w, h) in [(64, 64), (100, 100), (61, 200), (200, 61)]:
        k, d = pkg.ORBextractor(300, 1.2, 8, 20, 7)(pkg.synth.noise(w, h, w * h))
        assert len(k) == len(d)
    # portrait aspect ratio below 0.5 gives zero octree roots: undefined behaviour in the reference, error code here
    with pytest.raises(pkg.OrbGpuError):
        pkg.ORBextractor(300, 1.2, 8, 20, 7)(pkg.synth.noise(100, 400, 1))
    # wrong dtype is rejected by the binding (the reference asserts CV_8UC1)
    with pytest.raises(ValueError):
        ex1000(np.zeros((480, 640), np.float32))


def test_shape_change_on_one_handle(pkg, oracle):
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7); po = oracle.OraclePort(1000, 1.2, 8, 20, 7)
    for (w, h) in [(640, 480), (752, 480), (320, 240), (640, 480)]:
        img = pkg.synth.synth(w, h, w + h)
        k, d = ex(img); ok, od = po.extract(img)
        assert k.tobytes() == ok.tobytes() and np.array_equal(d, od)


def test_two_handles_concurrently(pkg):
    """Distinct instances run concurrently from two threads, as the stereo Frame ctor does (src/Frame.cc:78-81)."""
    import threading
    synth = pkg.synth
    imgs = [synth.synth(640, 480, 900 + i) for i in range(6)]
    ex_ref = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    expect = [ex_ref(i) for i in imgs]
    results = {}

    def worker(tag, order):
        ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
        results[tag] = [(i, ex(imgs[i])) for i in order for _ in range(3)]

    t1 = threading.Thread(target=worker, args=("L", [0, 1, 2, 3, 4, 5])); t2 = threading.Thread(target=worker, args=("R", [5, 4, 3, 2, 1, 0]))
    t1.start(); t2.start(); t1.join(); t2.join()
    for tag in ("L", "R"):
        for i, (k, d) in results[tag]:
            assert k.tobytes() == expect[i][0].tobytes() and np.array_equal(d, expect[i][1])


def test_float_steps_match_host(pkg, oracle, ex1000):
    """cv::fastAtan2 and sincosf as the kernels evaluate them vs cv2 / glibc golden values and vs the
    host libm on a dense sweep of [0, 2*pi] (every 7th float: 155 M angles)."""
    got = ex1000.debug_atan2(PRIMS["atan2_y"], PRIMS["atan2_x"])
    assert got.tobytes() == PRIMS["atan2_deg"].tobytes()
    s, c = ex1000.debug_sincosf(PRIMS["sincos_x"])
    assert s.tobytes() == PRIMS["sincos_sin"].tobytes() and c.tobytes() == PRIMS["sincos_cos"].tobytes()
    L = oracle.OraclePort().L
    L.orbo_sincos_checksum.restype = C.c_uint64
    L.orbo_sincos_checksum.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
    two_pi_bits = int(np.float32(6.2831855).view(np.uint32))
    count = two_pi_bits // 7
    assert ex1000.debug_sincos_checksum(0, count, 7) == L.orbo_sincos_checksum(0, count, 7, 1)


def test_sincos_exhaustive_over_the_angle_domain(pkg, oracle, ex1000):
    """EVERY float in [0, 2*pi] (step 1: 1 086 918 619 bit patterns -- keypoint angles are degrees in [0, 360) times pi/180):
    the device's sincosf follower vs the host's glibc sincosf, as an order-independent 64-bit checksum of (x, sin x, cos x).
    The host side is split over the cores (ctypes releases the GIL); chunk sums add modulo 2^64."""
    from concurrent.futures import ThreadPoolExecutor
    L = oracle.OraclePort().L
    L.orbo_sincos_checksum.restype = C.c_uint64
    L.orbo_sincos_checksum.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
    total = int(np.float32(6.2831855).view(np.uint32)) + 1
    nchunk = 64
    edges = [total * i // nchunk for i in range(nchunk + 1)]
    with ThreadPoolExecutor(max(1, len(os.sched_getaffinity(0)))) as pool:
        parts = list(pool.map(lambda i: L.orbo_sincos_checksum(edges[i], edges[i + 1] - edges[i], 1, 1), range(nchunk)))
    host = sum(parts) & 0xFFFFFFFFFFFFFFFF
    dev = 0
    for i in range(0, nchunk, 8):                              # eight launches: each stays well below the watchdog-free seconds range
        dev = (dev + ex1000.debug_sincos_checksum(edges[i], edges[i + 8] - edges[i], 1)) & 0xFFFFFFFFFFFFFFFF
    assert dev == host


def test_launches_are_counted(pkg):
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    before = ex.launch_count()
    ex(pkg.synth.synth(640, 480, 2))
    assert ex.launch_count() - before == 7 + 4       # 7 resize levels + FAST + octree + blur + describe


def test_async_batches_equal_synchronous_results(pkg):
    """orbgpu_extract_batch_async + orbgpu_wait: several batches in flight, pinned and pageable buffers."""
    import ctypes as C
    synth = pkg.synth
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    w, h = 640, 480
    cap = ex.max_keypoints(w, h)
    batches = [np.stack([synth.synth(w, h, 9000 + 10 * b + i) for i in range(5 + b)]) for b in range(3)]
    want = [ex.extract_batch(f) for f in batches]
    outs = []
    for f in batches:
        n = len(f)
        outs.append((np.zeros((n, cap), pkg.KP_DTYPE), np.zeros((n, cap, 32), np.uint8), np.full(n, -7, np.int32)))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for f, (k, d, c) in zip(batches, outs):
        ex.extract_batch_async_ptr(p(f), len(f), w, h, w, w * h, p(k), p(d), cap, p(c))
    with pytest.raises(pkg.OrbGpuError):
        ex.extract_batch(batches[0])                            # synchronous calls are refused while a window is open
    ex.wait()
    ex.wait()                                                   # idempotent
    for (k, d, c), (wk, wd, wc) in zip(outs, want):
        assert np.array_equal(c, wc)
        for i in range(len(c)):
            assert k[i, :c[i]].tobytes() == wk[i, :wc[i]].tobytes() and np.array_equal(d[i, :c[i]], wd[i, :wc[i]])
    assert ex.last_batch_ms() > 0
    kps, desc = ex(batches[0][0])                               # the handle is usable synchronously again
    assert len(kps) == want[0][2][0]


def test_graph_replay_equals_direct_launches(pkg):
    """Host-buffer chunks replay a captured CUDA graph of their kernels (orbgpu_set_graphs, default on): same bytes as
    direct launches, for batch sizes that change the per-context frame count (graph cache), shapes that change the plan
    (graphs invalidated) and the single-frame call; the launch counter counts the kernels inside the graph."""
    synth = pkg.synth
    exg = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    exd = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    exd.set_graphs(False)
    for (w, h), sizes in (((640, 480), (19, 3, 19, 64, 1, 40)), ((333, 251), (9, 2, 9)), ((640, 480), (5,))):
        frames = np.stack([synth.synth(w, h, 0xBEEF00 + i) for i in range(max(sizes))])
        for n in sizes:
            lg, ld = exg.launch_count(), exd.launch_count()
            kg, dg, ng = exg.extract_batch(frames[:n])
            kd, dd, nd = exd.extract_batch(frames[:n])
            assert exg.launch_count() - lg == exd.launch_count() - ld > 0
            assert np.array_equal(ng, nd) and ng.sum() > 0
            for i in range(n):
                assert kg[i, :ng[i]].tobytes() == kd[i, :nd[i]].tobytes() and np.array_equal(dg[i, :ng[i]], dd[i, :nd[i]])
        k1, d1 = exg(frames[0]); k2, d2 = exd(frames[0])
        assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    # profiling mode takes the direct path and still agrees
    exg.set_profiling(True)
    k3, d3 = exg(frames[0])
    exg.set_profiling(False)
    assert k3.tobytes() == k2.tobytes() and np.array_equal(d3, d2)

"""GPU: the sharded entry points of the C ABI.  Single-GPU forms run everywhere; the one-process-several-GPUs forms
(two extractor handles / two matchers on devices 0 and 1, NCCL group gather) need a box with at least 2 GPUs."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _descs(pkg, K, n):
    return pkg.synth.keyframe_descriptors(K, n, seed=0x5EED4000 + K)


def test_all_pairs_sharded_single_gpu_equals_count_pairs(pkg, oracle):
    import torch
    K, n = 23, 300
    desc = _descs(pkg, K, n)
    m = pkg.ORBmatcher(0.6, True)
    m.set_keyframes(desc)
    pairs = pkg.shard.all_pairs(K)
    want = m.count_pairs(pairs[:, 0], pairs[:, 1])
    out = torch.full((len(pairs),), -1, dtype=torch.int32, device="cuda")
    k_ms, g_ms, tot = m.all_pairs_sharded(None, out.data_ptr())
    assert np.array_equal(out.cpu().numpy(), want) and k_ms > 0 and tot >= k_ms
    po = oracle.OraclePort()
    for p in (0, 7, len(pairs) - 1):
        assert want[p] == po.match_count(desc[pairs[p, 0]], desc[pairs[p, 1]], 50, 0.6)
    # second call reuses the cached pair list; a different keyframe set rebuilds it
    k2 = m.all_pairs_sharded(None, out.data_ptr())
    assert np.array_equal(out.cpu().numpy(), want)
    desc2 = _descs(pkg, 9, 120)
    m.set_keyframes(desc2)
    p2 = pkg.shard.all_pairs(9)
    out2 = torch.full((len(p2),), -1, dtype=torch.int32, device="cuda")
    m.all_pairs_sharded(None, out2.data_ptr())
    assert np.array_equal(out2.cpu().numpy(), m.count_pairs(p2[:, 0], p2[:, 1]))


def test_nccl_unique_id_and_one_rank_comm(pkg):
    import torch
    ident = pkg.Comm.unique_id()
    assert len(ident) == 128 and any(ident)
    comm = pkg.Comm(0, 0, 1, ident)                      # a world of one: the collective degenerates, the call path is the same
    K, n = 12, 200
    desc = _descs(pkg, K, n)
    m = pkg.ORBmatcher(0.6, True)
    m.set_keyframes(desc)
    pairs = pkg.shard.all_pairs(K)
    out = torch.full((len(pairs),), -1, dtype=torch.int32, device="cuda")
    m.all_pairs_sharded(comm, out.data_ptr())
    assert np.array_equal(out.cpu().numpy(), m.count_pairs(pairs[:, 0], pairs[:, 1]))
    comm.close()


def test_two_gpus_in_one_process(pkg):
    """Two extractor handles and two matchers on devices 0 and 1 of ONE process (the per-device shared-memory opt-in:
    nfeatures 2000 needs more than 48 KB in k_octree), results equal device 0 alone."""
    if pkg.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch
    synth = pkg.synth
    frames = np.stack([synth.synth(1241, 376, 0x5EED1000 + i) for i in range(10)])
    e0 = pkg.ORBextractor(2000, 1.2, 8, 20, 7, device=0)
    e1 = pkg.ORBextractor(2000, 1.2, 8, 20, 7, device=1)
    k0, d0, n0 = e0.extract_batch(frames)
    k1, d1, n1 = e1.extract_batch(frames)                 # device 1 after device 0 has already raised its own limits
    assert np.array_equal(n0, n1) and n0.sum() > 10000
    for i in range(len(frames)):
        assert k0[i, :n0[i]].tobytes() == k1[i, :n1[i]].tobytes() and np.array_equal(d0[i, :n0[i]], d1[i, :n1[i]])
    km, dm, nm = pkg.extract_batch_multi([e0, e1], frames)
    assert np.array_equal(nm, n0)
    for i in range(len(frames)):
        assert km[i, :nm[i]].tobytes() == k0[i, :n0[i]].tobytes() and np.array_equal(dm[i, :nm[i]], d0[i, :n0[i]])
    # all-pairs matching: two matchers, one NCCL group gather, every GPU ends with the full vector
    K, n = 40, 500
    desc = _descs(pkg, K, n)
    ms = [pkg.ORBmatcher(0.6, True, device=d) for d in (0, 1)]
    for m in ms:
        m.set_keyframes(desc)
    pairs = pkg.shard.all_pairs(K)
    want = ms[0].count_pairs(pairs[:, 0], pairs[:, 1])
    outs = [torch.full((len(pairs),), -1, dtype=torch.int32, device="cuda:%d" % d) for d in (0, 1)]
    t = pkg.all_pairs_multi(ms, [0, 1], [o.data_ptr() for o in outs])
    assert t[2] > 0
    for o in outs:
        assert np.array_equal(o.cpu().numpy(), want)
    # and the distinctive-descriptor kernel (65.8 KB of dynamic shared memory) on the second device
    g = np.array([0, 50, 120, 300], np.int32)
    b0 = pkg.distinctive_descriptors(desc.reshape(-1, 32)[:300], g, device=0)
    b1 = pkg.distinctive_descriptors(desc.reshape(-1, 32)[:300], g, device=1)
    assert np.array_equal(b0, b1)


def test_host_alloc_pinned_and_write_combined_inputs(pkg):
    """orbgpu_host_alloc: frames in page-locked (and write-combined) memory of the library are used in place by the batch
    entry points and give the same results as pageable numpy memory."""
    synth = pkg.synth
    frames = np.stack([synth.synth(640, 480, 0xA110C + i) for i in range(6)])
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    want = ex.extract_batch(frames)
    for wc in (False, True):
        hb = pkg.HostBuffer(frames.shape, np.uint8, write_combined=wc)
        hb.array[:] = frames
        got = ex.extract_batch(hb.array)
        assert np.array_equal(got[2], want[2])
        for i in range(len(frames)):
            n = want[2][i]
            assert got[0][i, :n].tobytes() == want[0][i, :n].tobytes() and np.array_equal(got[1][i, :n], want[1][i, :n])
        hb.close()
    L = pkg.lib()
    p = C.c_void_p()
    assert L.orbgpu_host_alloc(C.byref(p), 0, 0) == -1 and L.orbgpu_host_alloc(None, 16, 0) == -1

"""Host-side multi-GPU logic on CPU: world_size-2 gloo runs of the frame sharding and the pair-count gather."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _descriptors(K, n, seed):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    out = np.empty((K, n, 32), np.uint8)
    for k in range(K):
        noise = np.ones((n, 32), np.uint8) * 255
        for _ in range(4):
            noise &= rng.integers(0, 256, (n, 32), dtype=np.uint8)
        out[k] = base[(7 * k + 3 * np.arange(n)) % 64] ^ noise
    return out


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as g
    import oracle_lib
    shard = g.load_package().shard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # ---- pair blocks + all_gather of counts (the oracle's CPU matcher stands in for the kernel)
        K, n, block = 12, 40, 8
        desc = _descriptors(K, n, 11)
        pairs = shard.all_pairs(K)
        port_lib = oracle_lib.OraclePort()
        local = []
        for s, e in shard.pair_blocks_of_rank(len(pairs), rank, world, block):
            local += [port_lib.match_count(desc[a], desc[b], 50, 0.6) for a, b in pairs[s:e]]
        mine = shard.pairs_of_rank(pairs, rank, world, block)
        assert len(mine) == len(local)
        assert np.array_equal(mine, np.concatenate([pairs[s:e] for s, e in shard.pair_blocks_of_rank(len(pairs), rank, world, block)]))
        plan = shard.gather_plan(len(pairs), world, block)
        full = shard.gather_pair_counts(np.array(local, np.int32), plan, dist=dist).numpy()
        want = np.array([port_lib.match_count(desc[a], desc[b], 50, 0.6) for a, b in pairs], np.int32)
        assert np.array_equal(full, want), (rank, full, want)
        assert want.max() > 0
        # ---- frame sharding: per-rank results merge back into frame order
        n_frames = 7
        mine = [("frame", int(i)) for i in shard.frames_of_rank(n_frames, rank, world)]
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        merged = shard.merge_frame_results(gathered, n_frames, world)
        assert merged == [("frame", i) for i in range(n_frames)]
        # ---- max over ranks of a per-rank time, as bench.py does
        t = torch.tensor([1.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert float(t) == float(world)
        open(os.path.join(tmp, "ok%d" % rank), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world2_gloo_pair_gather_and_frame_sharding(tmp_path):
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_pair_blocks_cover_every_pair_once():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    shard = g.load_package().shard
    for K, world, block in [(5, 2, 3), (64, 8, 100), (9, 4, 1), (3, 8, 4096)]:
        P = K * (K - 1) // 2
        seen = np.zeros(P, np.int32)
        for r in range(world):
            for s, e in shard.pair_blocks_of_rank(P, r, world, block):
                seen[s:e] += 1
        assert (seen == 1).all()
        pairs = shard.all_pairs(K)
        assert len(pairs) == P and (pairs[:, 0] < pairs[:, 1]).all()
    for n, world in [(10, 3), (1024, 8), (2, 4)]:
        allf = np.sort(np.concatenate([shard.frames_of_rank(n, r, world) for r in range(world)]))
        assert np.array_equal(allf, np.arange(n))


def test_adaptive_block_and_gather_plan():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    shard = g.load_package().shard
    # config-5 sizes: every rank gets at least 16 blocks, ranks differ by at most one block
    for K, world in [(384, 8), (1024, 1), (1024, 4), (4096, 8), (6, 8)]:
        P = K * (K - 1) // 2
        block = shard.adaptive_pair_block(P, world)
        assert 1 <= block <= 4096
        sizes = [len(shard.pair_index_of_rank(P, r, world, block)) for r in range(world)]
        assert sum(sizes) == P
        if P >= 16 * world:
            assert max(sizes) - min(sizes) <= block and max(sizes) <= 1.07 * (P / world) + 1
        # gather with world 1 semantics on the CPU: concatenating every rank's padded buffer and scattering = identity
        plan = shard.gather_plan(P, world, block)
        want = (np.arange(P, dtype=np.int32) * 7 + 3) % 1000
        allc = np.full(world * plan["per_rank"], -1, np.int32)
        for r in range(world):
            idx = shard.pair_index_of_rank(P, r, world, block)
            allc[r * plan["per_rank"]: r * plan["per_rank"] + len(idx)] = want[idx]
        full = np.empty(P, np.int32)
        full[plan["dst"].numpy()] = allc[plan["src"].numpy()]
        assert np.array_equal(full, want)

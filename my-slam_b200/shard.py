"""Multi-GPU sharding of the two hot paths.  One process per GPU.

Extraction: frames are independent, so frame i goes to rank i mod G and nothing is exchanged on the
data path; results return over each GPU's own PCIe link and are re-ordered on the host by frame index.
All-pairs matching: every rank holds all descriptors; the upper-triangular keyframe pair list is cut
into contiguous blocks dealt round-robin to ranks; the only exchange is ONE all_gather of the per-pair
accepted-match counts (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import numpy as np


def frames_of_rank(n_frames, rank, world):
    """Indices of the frames rank `rank` extracts (round-robin)."""
    return np.arange(rank, n_frames, world, dtype=np.int64)


def merge_frame_results(per_rank, n_frames, world):
    """per_rank[r] = list of results for frames_of_rank(n_frames, r, world), in that order."""
    out = [None] * n_frames
    for r in range(world):
        idx = frames_of_rank(n_frames, r, world)
        assert len(per_rank[r]) == len(idx)
        for i, res in zip(idx, per_rank[r]):
            out[int(i)] = res
    return out


def all_pairs(num_keyframes):
    """Upper-triangular keyframe pairs (a < b), row-major: int32 [P, 2], P = K(K-1)/2."""
    a, b = np.triu_indices(num_keyframes, 1)
    return np.stack([a, b], 1).astype(np.int32)


def adaptive_pair_block(num_pairs, world, max_block=4096, blocks_per_rank=16):
    """Pair-block size for `world` ranks: at most max_block, and small enough that every rank owns at least
    blocks_per_rank blocks -- with round-robin dealing the ranks then differ by at most one block in blocks_per_rank,
    i.e. by about 6 % at worst instead of a whole 4096-pair block."""
    return int(max(1, min(max_block, num_pairs // (blocks_per_rank * max(world, 1)))))


def pair_blocks_of_rank(num_pairs, rank, world, block=4096):
    """[start, stop) ranges of the pair list owned by `rank`: blocks of `block` pairs, round-robin."""
    nblocks = (num_pairs + block - 1) // block
    return [(k * block, min((k + 1) * block, num_pairs)) for k in range(rank, nblocks, world)]


def pair_index_of_rank(num_pairs, rank, world, block=4096):
    """Indices into the pair list owned by `rank`, in the order pair_blocks_of_rank lists them (int64)."""
    nblocks = (num_pairs + block - 1) // block
    ks = np.arange(rank, nblocks, world, dtype=np.int64)
    idx = (ks[:, None] * block + np.arange(block, dtype=np.int64)[None, :]).reshape(-1)
    return idx[idx < num_pairs]


def pairs_of_rank(pairs, rank, world, block=4096):
    return np.ascontiguousarray(pairs[pair_index_of_rank(len(pairs), rank, world, block)])


def gather_plan(num_pairs, world, block=4096, device="cpu"):
    """What gather_pair_counts needs, built once per (num_pairs, world, block): the padded per-rank length and, for every
    slot of the concatenated all_gather output, the pair it belongs to (the scatter index)."""
    import torch
    owned = [pair_index_of_rank(num_pairs, r, world, block) for r in range(world)]
    per_rank = max(len(o) for o in owned) if owned else 0
    src = np.concatenate([r * per_rank + np.arange(len(o), dtype=np.int64) for r, o in enumerate(owned)])
    dst = np.concatenate(owned)
    return {"num_pairs": num_pairs, "world": world, "per_rank": per_rank,
            "src": torch.from_numpy(src).to(device), "dst": torch.from_numpy(dst).to(device), "device": device}


def gather_pair_counts(local_counts, plan, dist=None):
    """local_counts: int32 counts of this rank's pairs in pair_index_of_rank order.  Returns the full int32 [num_pairs]
    vector on every rank: ONE all_gather of equal-size padded buffers, then ONE index scatter into pair order."""
    import torch
    device, world, per_rank = plan["device"], plan["world"], plan["per_rank"]
    lc = local_counts if isinstance(local_counts, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(local_counts, np.int32))
    pad = torch.full((per_rank,), -1, dtype=torch.int32, device=device)
    pad[:lc.numel()] = lc.to(device)
    if world > 1:
        allc = torch.empty((world * per_rank,), dtype=torch.int32, device=device)
        dist.all_gather_into_tensor(allc, pad)
    else:
        allc = pad
    full = torch.empty((plan["num_pairs"],), dtype=torch.int32, device=device)
    full[plan["dst"]] = allc[plan["src"]]
    return full

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


def pair_blocks_of_rank(num_pairs, rank, world, block=4096):
    """[start, stop) ranges of the pair list owned by `rank`: blocks of `block` pairs, round-robin."""
    nblocks = (num_pairs + block - 1) // block
    return [(k * block, min((k + 1) * block, num_pairs)) for k in range(rank, nblocks, world)]


def gather_pair_counts(local_counts, num_pairs, rank, world, block=4096, dist=None, device="cpu"):
    """local_counts: int32 counts of this rank's pairs, concatenated in pair_blocks_of_rank order.
    Returns the full int32 [num_pairs] vector on every rank after one all_gather (equal-size padded)."""
    import torch
    nblocks = (num_pairs + block - 1) // block
    per_rank_blocks = (nblocks + world - 1) // world
    pad = torch.full((per_rank_blocks * block,), -1, dtype=torch.int32, device=device)
    lc = local_counts if isinstance(local_counts, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(local_counts, np.int32))
    pad[:lc.numel()] = lc.to(device)
    if world > 1:
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad)
    else:
        bufs = [pad]
    full = torch.empty((num_pairs,), dtype=torch.int32, device=device)
    for r in range(world):
        off = 0
        for s, e in pair_blocks_of_rank(num_pairs, r, world, block):
            full[s:e] = bufs[r][off:off + (e - s)]
            off += e - s
    return full

"""Synthetic grayscale frames for parity tests and benchmarks.

synth(w, h, seed): splitmix64 stream seeded with `seed`; mid-gray canvas; R = round(400*w*h/307200)
axis-aligned rectangles of random gray; per-pixel noise in [-8, 8]; 3x3 box blur (REFLECT_101,
round-half-up).  Identical bytes feed the CUDA path, the oracle and the CPU baseline.
The other families exercise edge cases of the reference extractor (src/ORBextractor.cc):
flat -> no keypoints (descriptors.release() path, :1066), lowtex -> levels under budget (octree
ends by size==prevSize, :671), noise -> maximum corner density.
"""
import numpy as np

_GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed: int, start: int, count: int) -> np.ndarray:
    """Outputs number start .. start+count-1 (0-based) of the splitmix64 stream seeded with seed."""
    with np.errstate(over="ignore"):
        k = np.arange(start + 1, start + count + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + k * _GAMMA
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _box3(img: np.ndarray) -> np.ndarray:
    p = np.pad(img.astype(np.int32), 1, mode="reflect")      # numpy 'reflect' == REFLECT_101
    s = np.zeros(img.shape, np.int32)
    for dy in range(3):
        for dx in range(3):
            s += p[dy:dy + img.shape[0], dx:dx + img.shape[1]]
    return ((s * 2 + 9) // 18).astype(np.uint8)


def synth(w: int, h: int, seed: int, rect_div: int = 1, noise_amp: int = 8) -> np.ndarray:
    nrect = int(round(400.0 * w * h / 307200.0)) // rect_div
    r = splitmix64(seed, 0, 5 * nrect + w * h)
    img = np.full((h, w), 128, np.int32)
    for i in range(nrect):
        x0 = int(r[5 * i] % np.uint64(w)); y0 = int(r[5 * i + 1] % np.uint64(h))
        rw = 4 + int(r[5 * i + 2] % np.uint64(76)); rh = 4 + int(r[5 * i + 3] % np.uint64(76))
        g = int(r[5 * i + 4] % np.uint64(256))
        img[y0:min(y0 + rh, h), x0:min(x0 + rw, w)] = g
    span = np.uint64(2 * noise_amp + 1)
    noise = (r[5 * nrect:] % span).astype(np.int32).reshape(h, w) - noise_amp
    return _box3(np.clip(img + noise, 0, 255).astype(np.uint8))


def stereo_pair(w: int, h: int, seed: int, disparities=(6, 14, 27), noise_amp: int = 3):
    """A rectified left/right pair: horizontal bands of one scene, band k seen with disparity disparities[k]
    (uR = uL - d), plus independent per-pixel noise in the right image."""
    pad = max(disparities) + 1
    scene = synth(w + pad, h, seed)
    left = scene[:, :w].copy()
    right = np.empty_like(left)
    nb = len(disparities)
    for k, d in enumerate(disparities):
        y0, y1 = h * k // nb, h * (k + 1) // nb
        right[y0:y1] = scene[y0:y1, d:d + w]
    r = splitmix64(seed ^ 0x57E2E0, 0, w * h)
    nz = (r % np.uint64(2 * noise_amp + 1)).astype(np.int32).reshape(h, w) - noise_amp
    right = np.clip(right.astype(np.int32) + nz, 0, 255).astype(np.uint8)
    return left, right


def flat(w: int, h: int, value: int = 128) -> np.ndarray:
    return np.full((h, w), value, np.uint8)


def lowtex(w: int, h: int, seed: int) -> np.ndarray:
    return synth(w, h, seed, rect_div=8, noise_amp=2)


def noise(w: int, h: int, seed: int) -> np.ndarray:
    return (splitmix64(seed, 0, w * h) % np.uint64(256)).astype(np.uint8).reshape(h, w)


# BASELINE.json configs -> (width, height, nfeatures, seed base, frames)
CONFIGS = {
    "C1_tum_640x480_nf1000": (640, 480, 1000, 0x5EED0000, 256),
    "C2_kitti_1241x376_nf2000": (1241, 376, 2000, 0x5EED1000, 256),
    "C3_euroc_752x480_nf1000": (752, 480, 1000, 0x5EED2000, 1024),
    "C4_hd_1920x1080_nf4000": (1920, 1080, 4000, 0x5EED3000, 64),
}


def config_frames(name: str, count: int, start: int = 0) -> np.ndarray:
    w, h, _, base, _ = CONFIGS[name]
    return np.stack([synth(w, h, base + start + i) for i in range(count)])


def keyframe_descriptors(num_keyframes: int, n_per_keyframe: int, seed: int = 0x5EED4000) -> np.ndarray:
    """BASELINE config 5 input: [K, n, 32] uint8.  A pool of 65 536 base 256-bit words;
    keyframe k, descriptor j = pool[(977k + 31j) mod 65536] XOR noise, each noise bit the AND of 4 PRNG bits
    (p = 1/16), so two copies of a base word differ by ~30 bits (< TH_LOW) and unrelated words by ~128."""
    pool = splitmix64(seed, 0, 65536 * 4).reshape(65536, 4)
    k = np.arange(num_keyframes, dtype=np.int64)[:, None]
    j = np.arange(n_per_keyframe, dtype=np.int64)[None, :]
    base = pool[(977 * k + 31 * j) % 65536]                                   # [K, n, 4] uint64
    cnt = num_keyframes * n_per_keyframe * 4
    noise = splitmix64(seed ^ 0xA5A5, 0, cnt)
    for r in range(1, 4):
        noise &= splitmix64(seed ^ 0xA5A5, r * cnt, cnt)
    out = base ^ noise.reshape(num_keyframes, n_per_keyframe, 4)
    return np.ascontiguousarray(out).view(np.uint8).reshape(num_keyframes, n_per_keyframe, 32)

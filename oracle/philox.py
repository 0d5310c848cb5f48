"""CPU restatement of the in-kernel sampler noise of the native path (csrc/small_ops.cuh unit_normal): test
infrastructure only -- imported by tests/, never by the product path.

z[step][utterance][sample] ~ N(0, 1) is a pure function of (seed, step, utterance, sample): Philox4x32-10
(Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; the generator behind cuRAND / torch.randn on
CUDA, whose draws the reference consumes at open_universe/networks/universe/universe.py:39-41) keyed by the 64-bit
seed on the counter (sample // 4, utterance, step, 'UNIV'), two Box-Muller pairs per block:
sample & 2 selects the pair, sample & 1 cosine / sine.  Uniforms are 24-bit: ((r >> 8) + 0.5) 2^-24."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
TAG = 0x554E4956
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint64 arrays holding 32-bit values; returns the four output words."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def unit_normal(seed, step, utt, n):
    """float32 [n]: the draws of samples 0..n-1 of utterance `utt` at sampler step `step`."""
    l = np.arange(n, dtype=np.uint64)
    r = philox4x32_10(l >> np.uint64(2), np.full(n, utt, np.uint64), np.full(n, step, np.uint64), np.full(n, TAG, np.uint64),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    r = np.stack(r)                                            # [4][n]
    pair = ((l >> np.uint64(1)) & np.uint64(1)).astype(np.int64)
    idx = np.arange(n)
    ra, rb = r[2 * pair, idx], r[2 * pair + 1, idx]
    scale = np.float32(2.0 ** -24)
    u1 = ((ra >> np.uint64(8)).astype(np.float32) + np.float32(0.5)) * scale
    u2 = ((rb >> np.uint64(8)).astype(np.float32) + np.float32(0.5)) * scale
    rad = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
    ang = (np.float32(2.0) * u2).astype(np.float64) * np.pi
    trig = np.where((l & np.uint64(1)) == 1, np.sin(ang), np.cos(ang)).astype(np.float32)
    return (rad * trig).astype(np.float32)


def noise(seed, n_steps, B, Lp, b_off=0):
    """[n_steps][B][Lp] float32: what the kernels draw for a call of B utterances starting at utterance b_off."""
    return np.stack([np.stack([unit_normal(seed, s, b_off + b, Lp) for b in range(B)]) for s in range(n_steps)])

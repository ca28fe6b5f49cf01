"""Host-side mirror of the landmark sharding rule of libvio360_cuda.so (csrc/ba.cu:
compute_shard): each rank owns a contiguous landmark range, ranges balanced by
sum(d^2 + 4 d) over the landmarks' degrees d.  Used by the CPU-only
multi-process tests and by bench.py to report the shard sizes; the product path computes the
same cuts in C++."""
import numpy as np


def landmark_shards(point_idx, M, world):
    """returns [(lm_begin, lm_end)] * world"""
    deg = np.bincount(np.asarray(point_idx), minlength=M).astype(np.float64)
    w = np.concatenate([[0.0], np.cumsum(deg * deg + 4.0 * deg)])
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(w, w[M] * r / world, side="left")))
    cuts.append(M)
    cuts = [min(c, M) for c in cuts]
    return [(cuts[r], max(cuts[r], cuts[r + 1])) for r in range(world)]


def sequence_shards(n_sequences, world):
    """front-end: sequence s -> rank s mod world (independent units, no collective)"""
    return [list(range(r, n_sequences, world)) for r in range(world)]


BAND_BLOCKS = 24   # csrc/ba_schur.cuh kSWin: a chunk window, hence a block row of S, spans at most 24 poses


def band_pack(S, rhs, cost, band_w=None):
    """Layout of vio360_ba_reduced_band_dev: [block row][block column - row < w][6 x 6] | rhs | cost.
    S: dense (6 n_free)^2.  Blocks beyond the band must be zero (asserted)."""
    n_free = S.shape[0] // 6
    w = min(n_free, BAND_BLOCKS) if band_w is None else band_w
    B = S.reshape(n_free, 6, n_free, 6).transpose(0, 2, 1, 3)
    out = np.zeros((n_free, w, 6, 6))
    for d in range(w):
        idx = np.arange(n_free - d)
        out[idx, d] = B[idx, idx + d]
    for d in range(w, n_free):
        idx = np.arange(n_free - d)
        assert not B[idx, idx + d].any(), "block outside the exchange band"
    return np.concatenate([out.ravel(), np.asarray(rhs, np.float64).ravel(), [float(cost)]])


def band_expand(buf, n_free, band_w=None):
    """Inverse of band_pack (what k_expand_band does on the device): dense symmetric S, rhs, cost."""
    w = min(n_free, BAND_BLOCKS) if band_w is None else band_w
    nb = n_free * w * 36
    blocks = np.asarray(buf[:nb]).reshape(n_free, w, 6, 6)
    B = np.zeros((n_free, n_free, 6, 6))
    for d in range(w):
        idx = np.arange(n_free - d)
        B[idx, idx + d] = blocks[idx, d]
        if d:
            B[idx + d, idx] = blocks[idx, d].transpose(0, 2, 1)
    S = B.transpose(0, 2, 1, 3).reshape(6 * n_free, 6 * n_free)
    return S, np.array(buf[nb:nb + 6 * n_free]), float(buf[nb + 6 * n_free])

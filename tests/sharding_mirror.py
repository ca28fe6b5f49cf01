"""Host-side mirror of the landmark sharding rule of libvio360_cuda.so (csrc/ba.cu:
compute_shard): each rank owns a contiguous landmark range, ranges balanced by
sum(d^2 + 4 d) over the landmarks' degrees d.  Used by the CPU-only
multi-process tests (it lives under tests/, not in the product package); the product path computes the
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


def peer_rows(pose_idx, point_idx, free_map, lm_begin, lm_end):
    """Block rows (free-pose index range [lo, hi)) that the landmarks [lm_begin, lm_end) of a rank touch: what
    vio360_ba_peer_export publishes so that the other ranks pull a block row only from the ranks covering it
    (csrc/ba.cu: the hull of the shard's chunk windows, counted in free poses)."""
    pose_idx = np.asarray(pose_idx); point_idx = np.asarray(point_idx); free_map = np.asarray(free_map)
    sel = (point_idx >= lm_begin) & (point_idx < lm_end)
    if not sel.any():
        return 0, 0
    p0, p1 = int(pose_idx[sel].min()), int(pose_idx[sel].max())
    rows = free_map[p0:p1 + 1]
    rows = rows[rows >= 0]
    return (int(rows.min()), int(rows.max()) + 1) if rows.size else (0, 0)


def peer_exchange_sum(bands, rows, n_free, band_w=None, pose_diag=None):
    """numpy statement of k_band_pull (csrc/ba_peer.cuh): every element of the band is the sum, in rank order,
    of the partial bands of the ranks whose row range covers its block row (rhs entries: the pose's row; the cost:
    every rank); the pose part of the LM diagonal (squared, pose_diag: 6 n_free) is added on the diagonal; the
    result is expanded to the dense symmetric system.  bands: list of band_pack() buffers, rows: list of (lo, hi)."""
    w = min(n_free, BAND_BLOCKS) if band_w is None else band_w
    nS, n = n_free * w * 36, 6 * n_free
    row_of = np.concatenate([np.repeat(np.arange(n_free), w * 36), np.repeat(np.arange(n_free), 6), [-1]])
    out = np.zeros(nS + n + 1)
    for band, (lo, hi) in zip(bands, rows):           # rank order
        cover = (row_of < 0) | ((row_of >= lo) & (row_of < hi))
        out = out + np.where(cover, np.asarray(band, np.float64), 0.0)
    S, rhs, cost = band_expand(out, n_free, w)
    if pose_diag is not None:
        S[np.arange(n), np.arange(n)] += np.asarray(pose_diag, np.float64) ** 2
    return S, rhs, cost

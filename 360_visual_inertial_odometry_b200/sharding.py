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

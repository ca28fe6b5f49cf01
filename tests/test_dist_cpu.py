"""world_size-2 (gloo, CPU) test of the multi-GPU decomposition of path B:
every rank eliminates the landmarks of its shard, one all_reduce(sum) of the packed
[S | rhs | cost] buffer gives the full reduced camera system.  The per-rank arithmetic is the
oracle's here (no GPU in this suite); tests/test_gpu_ba.py checks the same identity with the
CUDA kernels on one device."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "360_visual_inertial_odometry_b200"


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    import importlib
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ba_oracle
    syn = importlib.import_module(PKG + ".synthetic")
    import sharding_mirror as sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    win = syn.make_ba_window(K=14, M=240, degree=6, seed=123, mid_iteration=True)
    full = ba_oracle.Window(win)
    r, Jp, Jl, cost = full.evaluate(win["pose_delta"], win["points"])
    D = full.lm_diagonal(Jp, Jl, 1e4)
    lo, hi = sharding.landmark_shards(win["point_idx"], win["M"], world)[rank]
    sel = (win["point_idx"] >= lo) & (win["point_idx"] < hi)
    sub = dict(win, O=int(sel.sum()))
    for k in ("pose_idx", "point_idx", "obs_u", "obs_v", "outlier"):
        sub[k] = np.ascontiguousarray(win[k][sel])
    W = ba_oracle.Window(sub)
    rs, Jps, Jls, cs = W.evaluate(win["pose_delta"], win["points"])
    Dr = D.copy()
    if rank != 0:
        Dr[:6 * win["K"]] = 0.0          # the pose part of the LM diagonal is added once (rank 0)
    S, rhs, _ = W.schur(rs, Jps, Jls, Dr)
    packed = torch.from_numpy(np.concatenate([S.ravel(), rhs, [cs]]))
    dist.all_reduce(packed)              # the one collective of the path
    # the same exchange with the block-band payload (vio360_ba_reduced_band_dev / vio360_ba_expand_band)
    band_part = sharding.band_pack(S, rhs, cs, band_w=8)
    band = torch.from_numpy(band_part.copy())
    dist.all_reduce(band)
    # the peer-memory exchange (vio360_ba_peer_exchange) stated in numpy: every rank publishes the block rows its
    # landmarks touch, pulls a row only from the ranks covering it and sums in rank order; no rank adds the pose
    # diagonal to its partial, the exchange adds it once
    fm = W.schur(rs, Jps, Jls, Dr)[2]                  # pose -> reduced block index, -1 for constant poses
    rows_mine = sharding.peer_rows(win["pose_idx"], win["point_idx"], fm, lo, hi)
    Sn, rhsn, _ = W.schur(rs, Jps, Jls, np.concatenate([np.zeros(6 * win["K"]), D[6 * win["K"]:]]))
    mine = sharding.band_pack(Sn, rhsn, cs, band_w=8)
    n_free = Sn.shape[0] // 6
    outside = np.ones(n_free, bool); outside[rows_mine[0]:rows_mine[1]] = False
    assert not mine[:n_free * 8 * 36].reshape(n_free, -1)[outside].any()      # nothing outside the published rows
    gathered = [None] * world
    dist.all_gather_object(gathered, (mine, rows_mine))
    pose_diag = D[:6 * win["K"]].reshape(-1, 6)[fm >= 0].ravel()
    Sp, rhsp, cp = sharding.peer_exchange_sum([g[0] for g in gathered], [g[1] for g in gathered], n_free, 8, pose_diag)
    digest = torch.tensor([float(Sp.sum()), float(rhsp.sum()), cp], dtype=torch.float64)
    digests = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(digests, digest)
    assert all(torch.equal(d_, digests[0]) for d_ in digests)                 # bit-identical on every rank
    if rank == 0:
        Sf, rhsf, _ = full.schur(r, Jp, Jl, D)
        n = Sf.shape[0]
        got = packed.numpy()
        rel = lambda a, b: float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
        Sb, rhsb, cb = sharding.band_expand(band.numpy(), n // 6, band_w=8)
        np.save(out, np.array([rel(got[:n * n], Sf.ravel()), rel(got[n * n:n * n + n], rhsf), abs(got[-1] - cost) / cost,
                               rel(Sb, Sf), rel(rhsb, rhsf), abs(cb - cost) / cost, band.numel() / packed.numel(),
                               rel(Sp, Sf), rel(rhsp, rhsf), abs(cp - cost) / cost,
                               float(sum(g[1][1] - g[1][0] for g in gathered)) / (world * n_free)]))
    dist.barrier()
    dist.destroy_process_group()


def test_landmark_shards_cover_and_balance():
    sys.path.insert(0, ROOT)
    import importlib
    import sharding_mirror as sharding
    rng = np.random.default_rng(0)
    point_idx = np.repeat(np.arange(1000), rng.integers(0, 25, 1000))
    for world in (1, 2, 4, 8):
        sh = sharding.landmark_shards(point_idx, 1000, world)
        assert sh[0][0] == 0 and sh[-1][1] == 1000 and all(sh[i][1] == sh[i + 1][0] for i in range(world - 1))
        deg = np.bincount(point_idx, minlength=1000).astype(float)
        cost = np.array([np.sum(deg[a:b] ** 2 + 4 * deg[a:b]) for a, b in sh])
        assert cost.max() <= cost.sum() / world + (24 ** 2 + 96) + 1e-9
    assert sharding.sequence_shards(10, 4) == [[0, 4, 8], [1, 5, 9], [2, 6], [3, 7]]


def test_two_rank_schur_allreduce(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "res.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    res = np.load(out)
    assert res[0] < 1e-9 and res[1] < 1e-9 and res[2] < 1e-12
    assert res[3] < 1e-9 and res[4] < 1e-9 and res[5] < 1e-12 and res[6] < 0.75      # band payload: same sums, fewer bytes
    assert res[7] < 1e-9 and res[8] < 1e-9 and res[9] < 1e-12                        # peer exchange: same system
    assert res[10] <= 1.0                                                            # rows pulled per rank / all rows


def test_band_pack_expand_round_trip():
    sys.path.insert(0, ROOT)
    import importlib
    import sharding_mirror as sharding
    rng = np.random.default_rng(3)
    for n_free in (1, 5, 24, 40):
        w = min(n_free, sharding.BAND_BLOCKS)
        A = rng.standard_normal((6 * n_free, 6 * n_free)); S = A + A.T
        bi = np.arange(6 * n_free) // 6
        S[np.abs(bi[:, None] - bi[None, :]) >= w] = 0.0
        rhs = rng.standard_normal(6 * n_free)
        buf = sharding.band_pack(S, rhs, 2.5)
        assert buf.size == n_free * w * 36 + 6 * n_free + 1
        S2, rhs2, c2 = sharding.band_expand(buf, n_free)
        assert np.array_equal(S2, S) and np.array_equal(rhs2, rhs) and c2 == 2.5

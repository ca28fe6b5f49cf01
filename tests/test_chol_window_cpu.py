"""numpy statement of the data flow of k_chol_band_reg (csrc/ba_solve.cuh): the 24-block window of the band Cholesky held
as one 6x6 block per owner, owners = unordered pairs of residues mod 24, a finished block (i, k) replaced by (k + 24, i)
(a diagonal owner moves from (j, j) to (j + 24, j + 24)), panel rows handed over by residue, the right-hand side riding
along.  The kernel's index arithmetic is restated here one to one and checked against LAPACK, so that a change of the
ownership rules can be tried on the CPU first (the first device version broke exactly on the diagonal owners' successor
block for systems of more than 24 block columns)."""
import numpy as np
import pytest

WB = 24


def _band_spd(nb, rng):
    n = 6 * nb
    A = np.zeros((n, n))
    for i in range(nb):
        for k in range(max(0, i - (WB - 1)), i + 1):
            A[6 * i:6 * i + 6, 6 * k:6 * k + 6] = rng.standard_normal((6, 6)) * 0.1
    S = np.tril(A)
    return S + S.T + np.eye(n) * 20.0


def _window_cholesky(S, b0):
    n = S.shape[0]
    nb = n // 6

    def blk_of(i, k):
        return S[6 * i:6 * i + 6, 6 * k:6 * k + 6].copy() if i < nb and k < nb else np.zeros((6, 6))

    def factor(B):
        L = np.linalg.cholesky(np.tril(B) + np.tril(B, -1).T)
        return L, 1.0 / np.diag(L)

    pairs = [(a, b) for a in range(WB) for b in range(a + 1)]
    blk, pref = {}, {}
    panelT = np.zeros((6, WB, 6)); Acol = np.zeros((WB, 6, 6)); sb = b0.copy()
    L_out = np.zeros((n, n))
    sL = sInv = None
    for t, (a, b) in enumerate(pairs):          # column 0: block (a, b); the block after it
        blk[t] = blk_of(a, b)
        pref[t] = blk_of(b + WB, a if a != b else a + WB)
        if a == 0:
            sL, sInv = factor(blk[t]); L_out[0:6, 0:6] = sL
        elif b == 0:
            Acol[a] = blk[t]
    for j in range(nb):
        jm = j % WB
        # panel: one row of the column's blocks per worker, X L_jj' = A by substitution; the last worker finishes rhs block j
        y = None
        for row in range((WB - 1) * 6 + 1):
            rhs = row == (WB - 1) * 6
            bi, ar = divmod(row, 6)
            res = (jm + 1 + bi) % WB
            src = sb[6 * j:6 * j + 6] if rhs else Acol[res][ar]
            o = np.zeros(6)
            for c in range(6):
                o[c] = (src[c] - sum(o[k] * sL[c, k] for k in range(c))) * sInv[c]
            if rhs:
                y = o
            else:
                panelT[:, res, ar] = o
                if j + 1 + bi < nb:
                    L_out[6 * (j + 1 + bi) + ar, 6 * j:6 * j + 6] = o
        sb[6 * j:6 * j + 6] = y
        # update: right-hand side rows below block j, then every owner
        for row in range((WB - 1) * 6):
            g = 6 * (j + 1) + row
            if g < n:
                res = (jm + 1 + row // 6) % WB
                sb[g] -= panelT[:, res, row % 6] @ y
        next_factor, hand_over = None, {}
        for t, (a, b) in enumerate(pairs):
            da, db = (a - jm) % WB, (b - jm) % WB
            di, dk = max(da, db), min(da, db)
            ri, rk = (a, b) if da >= db else (b, a)         # residues of the block's row / column index
            if dk == 0:                                      # (j + di, j) is finished: take up the prefetched successor
                blk[t] = pref[t]
                pref[t] = blk_of(j + di + WB, j + WB) if a != b else blk_of(j + 2 * WB, j + 2 * WB)
                if a != b and di == 1:
                    hand_over[jm] = blk[t].copy()            # (j + 24, j + 1) joins the next panel
            else:
                blk[t] = blk[t] - panelT[:, ri, :].T @ panelT[:, rk, :]
                if dk == 1:
                    if di == 1:
                        if j + 1 < nb:
                            next_factor = factor(blk[t])
                            L_out[6 * (j + 1):6 * (j + 1) + 6, 6 * (j + 1):6 * (j + 1) + 6] = next_factor[0]
                    else:
                        hand_over[ri] = blk[t].copy()
        for res, B in hand_over.items():
            Acol[res] = B
        if next_factor is not None:
            sL, sInv = next_factor
    return L_out, sb


@pytest.mark.parametrize("nb", [1, 5, 24, 25, 49, 60])
def test_residue_pair_window_matches_lapack(nb):
    rng = np.random.default_rng(nb)
    S = _band_spd(nb, rng)
    b0 = rng.standard_normal(6 * nb)
    L, y = _window_cholesky(S, b0)
    L_ref = np.linalg.cholesky(S)
    assert np.abs(L - L_ref).max() < 1e-12
    assert np.abs(y - np.linalg.solve(L_ref, b0)).max() < 1e-12


def test_column_oriented_back_substitution():
    """k_chol_band_back_col: pending sums per block, all updated when an x_j becomes known; x_{j-1} from the inverse of
    the diagonal factor."""
    nb = 40
    rng = np.random.default_rng(3)
    S = _band_spd(nb, rng)
    L = np.linalg.cholesky(S)
    y = rng.standard_normal(6 * nb)
    x = np.zeros(6 * nb); acc = np.zeros(6 * nb)
    Linv = [np.linalg.inv(L[6 * j:6 * j + 6, 6 * j:6 * j + 6]) for j in range(nb)]
    x[6 * (nb - 1):] = Linv[nb - 1].T @ y[6 * (nb - 1):]
    for j in range(nb - 1, 0, -1):
        for q in range(WB - 1):                  # worker (q, .): block k = j - 1 - q takes x_j in
            k = j - 1 - q
            if k >= 0:
                acc[6 * k:6 * k + 6] += L[6 * j:6 * j + 6, 6 * k:6 * k + 6].T @ x[6 * j:6 * j + 6]
        x[6 * (j - 1):6 * j] = Linv[j - 1].T @ (y[6 * (j - 1):6 * j] - acc[6 * (j - 1):6 * j])
    assert np.abs(x - np.linalg.solve(L.T, y)).max() < 1e-12

"""Host re-implementation of the device's procedural large-map generator (ekf_fill_procedural:
csrc/ekf_util.cuh fill_procedural_sigma_kernel / fill_procedural_state_kernel).

Sigma_0 = U U^T (rank 8, U hashed from (seed, row)) + a block-diagonal SPD term on the pose and the `count`
mapped landmarks, the reference prior (1e7) on the diagonal of the unmapped ones.  A pure function of
(seed, i, j): a host oracle never has to HOLD the matrix (57.6 GB at N = 60,000) - `sigma0_rows` returns any
block of rows on demand.  External (reference) indexing throughout: row 3 + 2k / 4 + 2k = landmark k.
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
RANK = 8


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return x ^ (x >> np.uint64(31))


def hash_sym(seed, a, b):
    """uniform in [-1, 1), pure function of (seed, a, b) - ekf_util.cuh hash_sym"""
    a = np.asarray(a, dtype=np.uint64)
    with np.errstate(over="ignore"):
        inner = _splitmix64((a * np.uint64(0x100000001B3) + np.uint64(b)) & _M64)
        h = _splitmix64(np.uint64(seed) ^ inner)
    return (h >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0


def _int_index(ext):
    ext = np.asarray(ext)
    return np.where(ext < 3, ext, ext + 1)


class ProceduralMap:
    def __init__(self, seed: int, n_landmarks: int, count: int, grid_w: int, prior: float = 10e6):
        self.seed, self.N, self.count, self.grid_w, self.prior = seed, n_landmarks, count, grid_w, prior
        self.n = 3 + 2 * n_landmarks
        self.n_active = 3 + 2 * count           # external indices below this are pose / mapped landmarks
        ext = np.arange(self.n)
        ii = _int_index(ext)
        self._int = ii
        U = np.stack([0.01 * hash_sym(seed, ii, 100 + r) for r in range(RANK)], axis=1)
        U[self.n_active:] = 0.0
        self.U = U                               # n x 8
        self.diag_extra = np.where(ext < self.n_active, 1.0e-3 * (1.5 + 0.5 * hash_sym(seed, ii, 7)), prior)
        # in-landmark correlation: entry (b, a) of landmark k (internal odd row i = b, column i - 1)
        self.off_extra = np.where((ii >= 4) & (ii % 2 == 1) & (ext < self.n_active), 3.0e-4 * hash_sym(seed, ii, 9), 0.0)

    def state(self):
        """the device's initial state estimate (prefer reading it back from the filter: the device forms it
        with fused multiply-adds, this numpy version rounds each product)"""
        k = np.arange(self.N)
        x = (k % self.grid_w) + 0.2 * hash_sym(self.seed, k, 1) + 0.01 * hash_sym(self.seed, k, 3)
        y = (k // self.grid_w) + 0.2 * hash_sym(self.seed, k, 2) + 0.01 * hash_sym(self.seed, k, 4)
        s = np.zeros(self.n)
        s[3:3 + 2 * self.count:2] = x[:self.count]
        s[4:4 + 2 * self.count:2] = y[:self.count]
        return s

    def sigma0_rows(self, i0: int, i1: int, j0: int = 0, j1: int | None = None):
        """Sigma_0[i0:i1, j0:j1] (symmetric, external indices)"""
        j1 = self.n if j1 is None else j1
        blk = self.U[i0:i1] @ self.U[j0:j1].T
        rows = np.arange(i0, i1)[:, None]
        cols = np.arange(j0, j1)[None, :]
        same = rows == cols
        blk = blk + np.where(same, self.diag_extra[i0:i1][:, None], 0.0)
        # unmapped landmarks: the diagonal is EXACTLY the prior (U rows are zero there already)
        blk = np.where(same & (rows >= self.n_active), self.prior, blk)
        lower = (rows == cols + 1)               # (b, a): row b carries the term
        upper = (cols == rows + 1)
        blk = blk + np.where(lower, self.off_extra[i0:i1][:, None], 0.0)
        blk = blk + np.where(upper, self.off_extra[j0:j1][None, :], 0.0)
        return blk

    def sigma0_cols(self, c0: int, c1: int):
        """Sigma_0[:, c0:c1] (n x (c1 - c0))"""
        return self.sigma0_rows(c0, c1, 0, self.n).T.copy()

"""numpy model of the row-sharded protocol (csrc/ekf_shard.cuh + the sharded mode of csrc/ekf_seq.cuh),
one instance per rank, the exchange done by a caller-supplied all-reduce (gloo in the CPU tests).

Host-logic check without a GPU; NOT part of the product and NOT an oracle.  Rows this rank does not
own are NaN in its copy of Sigma, so any read the real kernels must not make poisons the result.
"""
from __future__ import annotations

import math

import numpy as np

from _deferred_model import DeferredModel, normalize_angle


def equal_area_bounds(n, count):
    b = [int(round(n * math.sqrt(g / count))) for g in range(count + 1)]
    b[0], b[-1] = 0, n
    return b


class ShardedModel(DeferredModel):
    def __init__(self, rank, world, allreduce, n_landmarks=50, max_round=24, drop_guess=-1, **kw):
        super().__init__(n_landmarks=n_landmarks, **kw)
        self.rank, self.world, self.allreduce = rank, world, allreduce
        b = equal_area_bounds(self.n, world)
        self.rb, self.re = b[rank], b[rank + 1]
        self.max_round, self.drop_guess = max_round, drop_guess
        self.side_valid = False
        self.rounds = 0

    # ---- fixtures -----------------------------------------------------------------------------------
    def set_sigma(self, full):
        self.L = np.full((self.n, self.n), np.nan)
        self.L[self.rb:self.re] = np.tril(full)[self.rb:self.re]
        self.side_valid = False

    def own(self, i):
        return self.rb <= i < self.re

    def sync(self):
        """side extract + exchange: Pp (n x 3) and D (N x 3), own rows only, zeros elsewhere"""
        n, N = self.n, self.N
        buf = np.zeros(3 * n + 3 * N)
        Pp, D = buf[:3 * n].reshape(n, 3), buf[3 * n:].reshape(N, 3)
        for i in range(self.rb, self.re):
            Pp[i, :min(i + 1, 3)] = self.L[i, :min(i + 1, 3)]
        for k in range(N):
            a = 3 + 2 * k
            if self.own(a):
                D[k, 0] = self.L[a, a]
            if self.own(a + 1):
                D[k, 1], D[k, 2] = self.L[a + 1, a], self.L[a + 1, a + 1]
        self.allreduce(buf)
        Pp[0, 1], Pp[0, 2], Pp[1, 2] = Pp[1, 0], Pp[2, 0], Pp[2, 1]
        self.Pp, self.D = Pp.copy(), D.copy()
        self.side_valid = True

    # ---- one scan = rounds of (begin, exchange, finish) -------------------------------------------
    def scan(self, pose, xy, ids=None, action=None):
        assert self.side_valid and ids is None
        m = len(xy)
        assoc, t0, first = [], 0, True
        while t0 < m:
            mc = min(m - t0, self.max_round)
            cols, guess = self._begin(pose if first else None, xy[t0:t0 + mc])
            self.allreduce(cols.reshape(-1))
            done, a = self._finish(xy[t0:t0 + mc], cols, guess, last=(t0 + mc == m))
            assoc += a
            t0 += done
            first = False
            self.rounds += 1
        return assoc

    def _dist(self, th, px, py, z, k, pb):
        a = 3 + 2 * k
        h0, h1, zr_hat, zb_hat = self._lin(th, px, py, self.mu[a], self.mu[a + 1])
        B = self._block5(pb, self.Pp[a], self.Pp[a + 1], self.D[k])
        H = np.stack([h0, h1])
        S = H @ (B @ H.T) + np.eye(2) * self.r
        zd = np.array([z[0] - zr_hat, normalize_angle(z[1] - zb_hat)])
        return float(zd @ np.linalg.solve(S, zd))

    def _begin(self, pose, xy):
        n, Pp = self.n, self.Pp
        if pose is not None:                                      # shard_predict_kernel
            th = normalize_angle(pose[0]); px, py = pose[1], pose[2]
            dx, dy = px - self.prev[0], py - self.prev[1]
            A = np.eye(3); A[1, 0] = -dy; A[2, 0] = dx
            Pp[3:, 1] = Pp[3:, 0] * (-dy) + Pp[3:, 1]
            Pp[3:, 2] = Pp[3:, 0] * dx + Pp[3:, 2]
            Pp[0:3] = A @ Pp[0:3] @ A.T + np.eye(3) * self.q
            self.mu[0:3] = (th, px, py)
        th, px, py = self.mu[0:3]
        pb = Pp[0:3].copy()
        guess = []
        for (mx, my) in xy:                                       # shard_guess_kernel
            z = (math.sqrt(mx * mx + my * my), math.atan2(my, mx))
            best, kb = math.inf, -1
            for k in range(self.cnt):
                d = self._dist(th, px, py, z, k, pb)
                if d < best:
                    best, kb = d, k
            guess.append(kb if best <= 4.0 * self.thr else -1)
        if 0 <= self.drop_guess < len(guess):
            guess[self.drop_guess] = -1
        cols = np.zeros((len(xy), n, 2))                           # shard_extract_cols_kernel
        for s, j in enumerate(guess):
            if j < 0:
                continue
            a = 3 + 2 * j
            lo = max(self.rb, a + 2)
            if lo < self.re:
                cols[s, lo:self.re] = self.L[lo:self.re, a:a + 2]
            if self.own(a):
                cols[s, :a, 0] = self.L[a, :a]
            if self.own(a + 1):
                cols[s, :a, 1] = self.L[a + 1, :a]
        return cols, guess

    def _finish(self, xy, cols, guess, last):
        n, N, mu, Pp, D = self.n, self.N, self.mu, self.Pp, self.D
        m = len(xy)
        cnt_start = self.cnt
        W = np.zeros((n, 2 * m)); V = np.zeros((n, 2 * m))
        assoc, done = [], m
        for t in range(m):
            mx, my = xy[t]
            z = (math.sqrt(mx * mx + my * my), math.atan2(my, mx))
            th, px, py = mu[0:3]
            pb = Pp[0:3].copy()
            best, kb = math.inf, -1
            for k in range(self.cnt):
                d = self._dist(th, px, py, z, k, pb)
                if d < best:
                    best, kb = d, k
            if best <= self.thr:
                j, is_new = kb, False
            elif self.cnt < N:
                j, is_new = self.cnt, True
            else:
                assoc.append(-1)
                continue
            slot = -1
            if not is_new and j < cnt_start:
                slot = guess.index(j) if j in guess else -2
            if slot == -2:
                done = t
                break
            assoc.append(j)
            a = 3 + 2 * j
            if is_new:
                self.cnt += 1
                mu[a] = px + z[0] * math.cos(z[1] + th)
                mu[a + 1] = py + z[0] * math.sin(z[1] + th)
            h0, h1, zr_hat, zb_hat = self._lin(th, px, py, mu[a], mu[a + 1])
            H = np.stack([h0, h1])
            B = self._block5(pb, Pp[a], Pp[a + 1], D[j])
            M = np.linalg.inv(H @ (B @ H.T) + np.eye(2) * self.r)
            zd = np.array([z[0] - zr_hat, normalize_angle(z[1] - zb_hat)])
            c = cols[slot].copy() if slot >= 0 else np.zeros((n, 2))
            c -= V[:, :2 * t] @ W[[a, a + 1], :2 * t].T
            c[a] = (D[j][0], D[j][1]); c[a + 1] = (D[j][1], D[j][2])
            c[0:3, 0] = Pp[a, 0:3]; c[0:3, 1] = Pp[a + 1, 0:3]   # pose rows come from the current pose panel
            rows5 = np.concatenate([Pp, c], axis=1)
            G = rows5 @ H.T
            K = G @ M
            mu += K @ zd
            mu[0] = normalize_angle(mu[0])
            Pp -= K @ G[0:3].T
            pbn = Pp[0:3].copy(); Pp[0:3] = np.tril(pbn) + np.tril(pbn, -1).T
            for k in range(N):
                aa = 3 + 2 * k
                D[k, 0] -= K[aa] @ G[aa]
                D[k, 1] -= K[aa + 1] @ G[aa]
                D[k, 2] -= K[aa + 1] @ G[aa + 1]
            W[:, 2 * t:2 * t + 2] = G
            V[:, 2 * t:2 * t + 2] = K
        # covariance pass over the owned rows + write-back of the side state
        rows = slice(self.rb, self.re)
        upd = self.L[rows] - V[rows] @ W.T
        self.L[rows] = np.where(np.tri(n, dtype=bool)[rows], upd, np.nan)
        for i in range(self.rb, self.re):
            self.L[i, :min(i + 1, 3)] = Pp[i, :min(i + 1, 3)]
        for k in range(N):
            a = 3 + 2 * k
            if self.own(a):
                self.L[a, a] = D[k, 0]
            if self.own(a + 1):
                self.L[a + 1, a], self.L[a + 1, a + 1] = D[k, 1], D[k, 2]
        if last and done == m:
            self.prev = np.array([mu[1], mu[2]])
        return done, assoc

    def owned_rows(self):
        """(row_begin, rows): the lower-triangle rows this rank stores"""
        return self.rb, np.nan_to_num(np.tril(self.L)[self.rb:self.re], nan=0.0)

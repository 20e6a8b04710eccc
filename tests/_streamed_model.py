"""Streamed block-wise oracle for maps whose covariance no host can hold (BASELINE config 5: N = 60,000,
n = 120,003, Sigma = 115 GB full).

The same scan semantics as oracle A / B (slam.cpp:321-508: pose injection, prediction, per measurement the
Mahalanobis sweep, index_min, threshold rule, gain, state update, `sigma = (I - K H) sigma`), evaluated with
O(n m) memory: Sigma_0 is never stored, an accessor returns any block of it on demand (oracle/procedural_map.py),
and the scan's m sequential rank-2 downdates are carried as the panels (G_t, K_t), so that

    Sigma_cur[:, c] = Sigma_0p[:, c] - sum_{s<t} K_s G_s[c, :]^T          (any column, O(n m))
    Sigma_new[r, :] = Sigma_0p[r, :] - sum_t   K_t[r, :] G_t^T            (any block of rows, afterwards)

with Sigma_0p = A Sigma_0 A^T + Qbar (the prediction touches the pose rows / columns only).  This is the SAME
algebra as `(I - K H) Sigma` applied m times (Sigma_cur H^T = G, (I - K H) Sigma = Sigma - K G^T for symmetric
Sigma); tests/test_streamed_model.py checks it against oracle B, entry for entry, at sizes oracle B can hold.
TEST INFRASTRUCTURE ONLY - a checker, vectorised numpy, no product code involved.
"""
from __future__ import annotations

import math

import numpy as np

PI = 3.14159265358979323846


def normalize_angle(rad):
    d = math.fmod(rad + PI, 2 * PI)
    return d + PI if d <= 0.0 else d - PI


def _normalize_vec(rad):
    d = np.fmod(rad + PI, 2 * PI)
    return np.where(d <= 0.0, d + PI, d - PI)


class StreamedModel:
    def __init__(self, n_landmarks, sigma0_rows, state, count, prev_xy, q=1e-2, r=1e-2, thr=2.0):
        """sigma0_rows(i0, i1, j0, j1) -> Sigma_0[i0:i1, j0:j1] of the covariance the scan starts from"""
        self.N, self.n = n_landmarks, 3 + 2 * n_landmarks
        self.q, self.r, self.thr = q, r, thr
        self.acc = sigma0_rows
        self.mu = np.array(state, dtype=np.float64)
        self.cnt = int(count)
        self.prev = np.array(prev_xy, dtype=np.float64)
        self.G = self.K = None
        self.ndy = self.pdx = 0.0

    # Sigma_0p blocks -------------------------------------------------------------------------------
    def _pred_rows(self, i0, i1, j0=0, j1=None):
        """rows [i0, i1) of Sigma_0p = A Sigma_0 A^T + Qbar, columns [j0, j1)"""
        n = self.n
        j1 = n if j1 is None else j1
        B = self.acc(i0, i1, j0, j1).copy()
        ndy, pdx = self.ndy, self.pdx
        # column part (right multiplication by A^T): columns 1, 2 += (-dy, dx) * column 0
        need0 = self.acc(i0, i1, 0, 1)[:, 0] if j0 > 0 else B[:, 0].copy()
        if j0 <= 1 < j1:
            B[:, 1 - j0] = need0 * ndy + B[:, 1 - j0]
        if j0 <= 2 < j1:
            B[:, 2 - j0] = need0 * pdx + B[:, 2 - j0]
        # row part (left multiplication by A): rows 1, 2 += (-dy, dx) * row 0 (of the column-updated matrix)
        if i0 <= 2 and i1 > 1:
            r0 = self.acc(0, 1, j0, j1)[0].copy()
            c00 = self.acc(0, 1, 0, 1)[0, 0]
            if j0 <= 1 < j1:
                r0[1 - j0] = c00 * ndy + r0[1 - j0]
            if j0 <= 2 < j1:
                r0[2 - j0] = c00 * pdx + r0[2 - j0]
            if i0 <= 1 < i1:
                B[1 - i0] = ndy * r0 + B[1 - i0]
            if i0 <= 2 < i1:
                B[2 - i0] = pdx * r0 + B[2 - i0]
        for d in range(3):
            if i0 <= d < i1 and j0 <= d < j1:
                B[d - i0, d - j0] += self.q
        return B

    # one whole scan --------------------------------------------------------------------------------
    def scan(self, pose, xy):
        n = self.n
        mu = self.mu
        th = normalize_angle(pose[0]); px, py = float(pose[1]), float(pose[2])      # slam.cpp:323-325
        self.ndy, self.pdx = -(py - self.prev[1]), (px - self.prev[0])             # slam.cpp:327-333
        mu[0:3] = (th, px, py)
        # the n x 3 pose panel and the 2x2 diagonal blocks of the current covariance (all the sweep needs)
        Pp = np.concatenate([self._pred_rows(i0, min(n, i0 + 32768), 0, 3) for i0 in range(0, n, 32768)], axis=0)
        a_idx = 3 + 2 * np.arange(self.N)
        Daa = np.empty(self.N); Dba = np.empty(self.N); Dbb = np.empty(self.N)
        for k0 in range(0, self.N, 8192):
            k1 = min(self.N, k0 + 8192)
            blk = self.acc(3 + 2 * k0, 3 + 2 * k1, 3 + 2 * k0, 3 + 2 * k1)
            i = 2 * np.arange(k1 - k0)
            Daa[k0:k1] = blk[i, i]; Dba[k0:k1] = blk[i + 1, i]; Dbb[k0:k1] = blk[i + 1, i + 1]
        m = len(xy)
        self.G = np.zeros((n, 2 * m)); self.K = np.zeros((n, 2 * m))
        G, K = self.G, self.K
        assoc, best_d, second_d = [], [], []
        for t in range(m):
            mx, my = float(xy[t][0]), float(xy[t][1])
            zr, zphi = math.sqrt(mx * mx + my * my), math.atan2(my, mx)               # :345-347
            th, px, py = mu[0], mu[1], mu[2]
            c = self.cnt
            # ---- the sweep over the c mapped landmarks (:361-404), vectorised ----
            j, is_new = -1, False
            b1 = b2 = math.inf
            if c > 0:
                lx, ly = mu[3:3 + 2 * c:2], mu[4:4 + 2 * c:2]
                dx, dy = lx - px, ly - py
                d2 = dx * dx + dy * dy
                sd = np.sqrt(d2)
                with np.errstate(all="ignore"):
                    h0 = np.stack([np.zeros(c), -dx / sd, -dy / sd, dx / sd, dy / sd], axis=1)
                    h1 = np.stack([-np.ones(c), dy / d2, -dx / d2, -dy / d2, dx / d2], axis=1)
                    zb = _normalize_vec(np.arctan2(dy, dx) - th)
                    B = np.zeros((c, 5, 5))
                    B[:, 0:3, 0:3] = Pp[0:3, 0:3]
                    pa, pb = Pp[3:3 + 2 * c:2], Pp[4:4 + 2 * c:2]
                    B[:, 3, 0:3] = pa; B[:, 0:3, 3] = pa
                    B[:, 4, 0:3] = pb; B[:, 0:3, 4] = pb
                    B[:, 3, 3] = Daa[:c]; B[:, 4, 3] = Dba[:c]; B[:, 3, 4] = Dba[:c]; B[:, 4, 4] = Dbb[:c]
                    T0 = np.einsum("kij,kj->ki", B, h0); T1 = np.einsum("kij,kj->ki", B, h1)
                    s00 = np.einsum("ki,ki->k", h0, T0) + self.r
                    s01 = np.einsum("ki,ki->k", h0, T1)
                    s11 = np.einsum("ki,ki->k", h1, T1) + self.r
                    det = s00 * s11 - s01 * s01
                    i00, i01, i11 = s11 / det, -s01 / det, s00 / det
                    z0 = zr - sd
                    z1 = _normalize_vec(zphi - zb)
                    t0 = z0 * i00 + z1 * i01; t1 = z0 * i01 + z1 * i11
                    dist = t0 * z0 + t1 * z1
                finite = np.where(np.isnan(dist), np.inf, dist)
                kb = int(np.argmin(finite))               # first minimum, NaN never wins
                order = np.partition(finite, min(1, c - 1))
                b1 = float(order[0]); b2 = float(order[1]) if c > 1 else math.inf
                if b1 <= self.thr:                        # the tentative entry sits last with d == threshold
                    j = kb
            if j < 0:
                if c >= self.N:
                    raise IndexError("landmark capacity exceeded")
                j, is_new = c, True
                self.cnt = c + 1
            assoc.append(j); best_d.append(b1); second_d.append(b2)
            a = 3 + 2 * j
            if is_new:
                mu[a] = px + zr * math.cos(zphi + th)
                mu[a + 1] = py + zr * math.sin(zphi + th)
            # ---- linearisation at the chosen landmark, S, gain (:443-476) ----
            dx, dy = mu[a] - px, mu[a + 1] - py
            d2 = dx * dx + dy * dy; sd = math.sqrt(d2)
            H5 = np.array([[0.0, -dx / sd, -dy / sd, dx / sd, dy / sd],
                           [-1.0, dy / d2, -dx / d2, -dy / d2, dx / d2]])
            zb = normalize_angle(math.atan2(dy, dx) - th)
            zd = np.array([zr - sd, normalize_angle(zphi - zb)])
            # landmark columns of the CURRENT covariance: Sigma_0p[:, {a, b}] - earlier downdates of this scan
            cols = np.concatenate([self._pred_rows(i0, min(n, i0 + 32768), a, a + 2) for i0 in range(0, n, 32768)], 0)
            cols -= K[:, :2 * t] @ G[[a, a + 1], :2 * t].T
            rows5 = np.concatenate([Pp, cols], axis=1)    # Sigma_cur[:, {0, 1, 2, a, b}]
            Gt = rows5 @ H5.T                             # Sigma_cur H^T (H has these five non-zero columns)
            S = H5 @ np.concatenate([Gt[0:3], Gt[a:a + 2]], axis=0) + np.eye(2) * self.r
            det = S[0, 0] * S[1, 1] - S[0, 1] * S[1, 0]
            Sinv = np.array([[S[1, 1] / det, -S[0, 1] / det], [-S[1, 0] / det, S[0, 0] / det]])
            Kt = Gt @ Sinv
            mu += Kt @ zd                                 # :482
            mu[0] = normalize_angle(mu[0])                # :488
            # keep the pose panel and the diagonal blocks current
            Pp -= Kt @ Gt[0:3].T
            Ka, Kb, Ga, Gb = Kt[3::2], Kt[4::2], Gt[3::2], Gt[4::2]
            Daa -= np.einsum("ki,ki->k", Ka, Ga)
            Dba -= np.einsum("ki,ki->k", Kb, Ga)
            Dbb -= np.einsum("ki,ki->k", Kb, Gb)
            G[:, 2 * t:2 * t + 2] = Gt
            K[:, 2 * t:2 * t + 2] = Kt
        self.prev = np.array([mu[1], mu[2]])              # :508
        self.best_d, self.second_d = np.array(best_d), np.array(second_d)
        return assoc

    def sigma_new_rows(self, i0, i1):
        """rows [i0, i1) of the covariance after the scan (full width)"""
        return self._pred_rows(i0, i1) - self.K[i0:i1] @ self.G.T

"""numpy model of the GPU algorithm (csrc/ekf_seq.cuh + csrc/ekf_downdate.cuh).

Host-logic check that runs without a GPU: the same deferred-update algebra the kernels use
(pose panel Pp, diagonal blocks D, landmark columns gathered from Sigma_0 and corrected with
the W/V panels, one rank-2m downdate of the lower triangle per scan) written with numpy, so
tests can confirm the algebra against the oracle before any GPU time is spent.  It is NOT
part of the product and NOT an oracle.
"""
from __future__ import annotations

import math

import numpy as np

PI = 3.14159265358979323846


def normalize_angle(rad):
    d = math.fmod(rad + PI, 2 * PI)
    return d + PI if d <= 0.0 else d - PI


class DeferredModel:
    def __init__(self, n_landmarks=50, q=1e-2, r=1e-2, thr=2.0, prior=10e6):
        self.N = n_landmarks
        self.n = 3 + 2 * n_landmarks
        self.q, self.r, self.thr = q, r, thr
        self.mu = np.zeros(self.n)
        self.L = np.zeros((self.n, self.n))          # lower triangle of Sigma (incl. diagonal)
        self.L[np.arange(3, self.n), np.arange(3, self.n)] = prior
        self.cnt = 0
        self.prev = np.zeros(2)

    def sym(self, i, j):
        return self.L[i, j] if i >= j else self.L[j, i]

    def sigma(self):
        return np.tril(self.L) + np.tril(self.L, -1).T

    def _lin(self, th, px, py, lx, ly):
        dx, dy = lx - px, ly - py
        d = dx * dx + dy * dy
        sd = math.sqrt(d)
        h0 = np.array([0.0, -dx / sd, -dy / sd, dx / sd, dy / sd])
        h1 = np.array([-1.0, dy / d, -dx / d, -dy / d, dx / d])
        return h0, h1, sd, normalize_angle(math.atan2(dy, dx) - th)

    def scan(self, pose, xy, ids=None, action=None):
        n, N = self.n, self.N
        L = self.L
        # ---- phase 0: predict on the pose panel -------------------------------------------
        th = normalize_angle(pose[0]); px, py = pose[1], pose[2]
        dx, dy = px - self.prev[0], py - self.prev[1]
        P = np.array([[self.sym(i, j) for j in range(3)] for i in range(3)])
        A = np.eye(3); A[1, 0] = -dy; A[2, 0] = dx
        P = A @ P @ A.T + np.eye(3) * self.q
        for i in range(3):
            for j in range(i + 1):
                L[i, j] = P[i, j]
        pb = np.tril(P) + np.tril(P, -1).T
        L[3:, 1] = L[3:, 0] * (-dy) + L[3:, 1]
        L[3:, 2] = L[3:, 0] * dx + L[3:, 2]
        Pp = L[:, 0:3].copy()
        Pp[0:3] = pb
        D = np.zeros((N, 3))
        for k in range(N):
            a = 3 + 2 * k
            D[k] = (L[a, a], L[a + 1, a], L[a + 1, a + 1])
        ps = np.array([th, px, py])
        mu = self.mu
        mu[0:3] = ps
        m = len(xy)
        W = np.zeros((n, 2 * m)); V = np.zeros((n, 2 * m))
        assoc = []
        for t in range(m):
            mx, my = xy[t]
            zr, zphi = math.sqrt(mx * mx + my * my), math.atan2(my, mx)
            th, px, py = ps
            j, is_new = -1, False
            if ids is None:
                best, kb = math.inf, -1
                for k in range(self.cnt):
                    a = 3 + 2 * k
                    h0, h1, zr_hat, zb_hat = self._lin(th, px, py, mu[a], mu[a + 1])
                    B = self._block5(pb, Pp[a], Pp[a + 1], D[k])
                    T = B @ np.stack([h0, h1]).T
                    S = np.stack([h0, h1]) @ T + np.eye(2) * self.r
                    zd = np.array([zr - zr_hat, normalize_angle(zphi - zb_hat)])
                    d = zd @ np.linalg.solve(S, zd)
                    if d < best:
                        best, kb = d, k
                if best <= self.thr:
                    j = kb
                elif self.cnt < N:
                    j, is_new = self.cnt, True
                    self.cnt += 1
            else:
                if action is None or action[t] != 2:
                    j = int(ids[t])
                    is_new = mu[3 + 2 * j] == 0.0 and mu[4 + 2 * j] == 0.0
            assoc.append(j)
            if j < 0:
                continue
            a = 3 + 2 * j
            if is_new:
                mu[a] = px + zr * math.cos(zphi + th)
                mu[a + 1] = py + zr * math.sin(zphi + th)
            h0, h1, zr_hat, zb_hat = self._lin(th, px, py, mu[a], mu[a + 1])
            H = np.stack([h0, h1])
            B = self._block5(pb, Pp[a], Pp[a + 1], D[j])
            T = B @ H.T
            S = H @ T + np.eye(2) * self.r
            M = np.linalg.inv(S)
            zd = np.array([zr - zr_hat, normalize_angle(zphi - zb_hat)])
            # landmark columns of the CURRENT covariance: gather from Sigma_0 + correction
            cols = np.zeros((n, 2))
            for i in range(n):
                cols[i, 0] = self.sym(i, a)
                cols[i, 1] = self.sym(i, a + 1)
            cols -= V[:, :2 * t] @ W[[a, a + 1], :2 * t].T
            cols[a] = (D[j][0], D[j][1]); cols[a + 1] = (D[j][1], D[j][2])
            rows5 = np.concatenate([Pp, cols], axis=1)       # n x 5: Sigma_cur[:, {0,1,2,a,b}]
            rows5[0:3, 0:3] = pb
            G = rows5 @ H.T
            K = G @ M
            mu += K @ zd
            ps = mu[0:3].copy()
            ps[0] = normalize_angle(ps[0]); mu[0] = ps[0]
            Pp -= K @ G[0:3].T
            pb = Pp[0:3].copy(); pb = np.tril(pb) + np.tril(pb, -1).T; Pp[0:3] = pb
            for k in range(N):
                aa = 3 + 2 * k
                D[k, 0] -= K[aa] @ G[aa]
                D[k, 1] -= K[aa + 1] @ G[aa]
                D[k, 2] -= K[aa + 1] @ G[aa + 1]
            W[:, 2 * t:2 * t + 2] = G
            V[:, 2 * t:2 * t + 2] = K
        # ---- the covariance pass: one rank-2m downdate of the lower triangle -------------------
        self.L = np.tril(L - V @ W.T)
        self.prev = np.array([ps[1], ps[2]])
        return assoc

    @staticmethod
    def _block5(pb, pa, pbrow, d):
        B = np.zeros((5, 5))
        B[0:3, 0:3] = pb
        B[3, 0:3] = B[0:3, 3] = pa
        B[4, 0:3] = B[0:3, 4] = pbrow
        B[3, 3] = d[0]; B[4, 3] = B[3, 4] = d[1]; B[4, 4] = d[2]
        return B

"""Host-side drivers of the row-sharded large map (include/ekf_slam_b200.h, ekf_shard_*).

The C ABI leaves the one exchange per scan round to the caller.  Two callers live here:

  * `ShardedEkf`   one rank of a torch.distributed job (one process per GPU): the exchange is
                   `dist.all_reduce(SUM)` over NCCL on a tensor that aliases the handle's buffer;
  * `LocalShards`  R handles on ONE GPU (how the multi-rank path is exercised on a single-GPU test
                   box): the exchange is a plain sum of the R buffers.

torch is used for device buffers, streams and the collective only.
"""
from __future__ import annotations

import numpy as np
import torch

from . import EkfSlam

MAX_ROUND = 24   # kMaxChunkM: sightings per round


class _DevBuf:
    """zero-copy view of a raw device buffer of doubles for torch.as_tensor"""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 2}


def alias(ptr: int, count: int, device) -> torch.Tensor:
    return torch.as_tensor(_DevBuf(ptr, count), device=device)


def shard_bounds(n_landmarks: int, count: int):
    """row boundaries of the equal-area split (mirrors ekf_shard_bound in csrc/ekf_layout.h)"""
    n_pad = (4 + 2 * n_landmarks + 127) // 128 * 128
    out = []
    for g in range(count + 1):
        if g == 0:
            out.append(0)
        elif g == count:
            out.append(n_pad)
        else:
            x = n_pad * n_pad * g // count
            r = int(np.floor(np.sqrt(float(x))))
            while r * r > x:
                r -= 1
            while (r + 1) * (r + 1) <= x:
                r += 1
            out.append(min((r + 64) // 128 * 128, n_pad))
    return out


class _Driver:
    """round loop shared by both callers; subclasses provide handles() and exchange(list of (ptr, count))"""

    def _setup(self, device, stream=None):
        """every handle runs on ONE torch stream (self.stream), and so does every torch op of the driver"""
        self.device = device
        self.stream = stream if stream is not None else torch.cuda.Stream(device)
        assert self.stream.cuda_stream != 0   # ekf_set_stream reads 0 as "the handle's own stream"
        for h in self.handles():
            h.set_stream(self.stream.cuda_stream)
        with torch.cuda.stream(self.stream):
            self._done = torch.zeros(1, dtype=torch.int32, device=device)
            self._cnt = torch.zeros(1, dtype=torch.int32, device=device)
            self._pose_out = torch.zeros(3, dtype=torch.float64, device=device)

    def sync(self):
        with torch.cuda.stream(self.stream):
            bufs = [h.shard_sync_begin() for h in self.handles()]
            self.exchange(bufs)
            for h in self.handles():
                h.shard_sync_finish()

    def scan_dev(self, pose_t, xy_t, assoc_t, ids_t=None, check: bool = True, done_t=None):
        """One scan from device tensors: pose_t [3], xy_t [m, 2], assoc_t int32 [m] (written).
        With check=False a single round is enqueued without reading `done` back (fully asynchronous;
        valid when m <= 24 and the speculation holds - inspect self._done afterwards)."""
        with torch.cuda.stream(self.stream):
            return self._scan_dev(pose_t, xy_t, assoc_t, ids_t, check, done_t)

    def _scan_dev(self, pose_t, xy_t, assoc_t, ids_t, check, done_t=None):
        done_t = self._done if done_t is None else done_t
        m = xy_t.shape[0]
        t0, first, rounds = 0, True, 0
        while t0 < m:
            rem = m - t0
            last = rem <= MAX_ROUND
            bufs = [h.shard_scan_begin(pose_t.data_ptr() if first else 0, rem, xy_t[t0:].data_ptr(),
                                       ids_t[t0:].data_ptr() if ids_t is not None else 0, last)
                    for h in self.handles()]
            self.exchange(bufs)
            for h in self.handles():
                h.shard_scan_finish(assoc_t[t0:].data_ptr(), done_t.data_ptr(), self._pose_out.data_ptr(),
                                    self._cnt.data_ptr())
            rounds += 1
            first = False
            if not check:
                return rounds
            done = int(done_t.item())
            if done < 0:
                raise RuntimeError("sharded scan: the sequential kernel's grid barrier timed out (EKF_ERR_CUDA)")
            if done == 0 and rounds > 4 * m + 4:
                raise RuntimeError("sharded scan makes no progress")
            t0 += done
        return rounds

    def _host_io(self, m):
        """pinned staging for scan(): one H2D of (pose | xy), one D2H of (pose_after | count | done | assoc)"""
        if getattr(self, "_io_m", -1) >= m:
            return
        self._io_m = m
        self._h_in = torch.empty(3 + 2 * m, dtype=torch.float64, pin_memory=True)
        self._d_in = torch.empty(3 + 2 * m, dtype=torch.float64, device=self.device)
        self._h_ids = torch.empty(m, dtype=torch.int32, pin_memory=True)
        self._d_ids = torch.empty(m, dtype=torch.int32, device=self.device)
        self._d_res = torch.zeros(32 + 4 * m, dtype=torch.uint8, device=self.device)
        self._h_res = torch.zeros(32 + 4 * m, dtype=torch.uint8, pin_memory=True)
        self._r_pose = self._d_res[0:24].view(torch.float64)
        self._r_cnt = self._d_res[24:28].view(torch.int32)
        self._r_done = self._d_res[28:32].view(torch.int32)
        self._r_assoc = self._d_res[32:32 + 4 * m].view(torch.int32)
        self._res_ev = torch.cuda.Event()
        self._res_ev.record(self.stream)          # creates the underlying cudaEvent_t

    def scan(self, pose, xy, ids=None):
        """host arrays in, (assoc, count, pose_after, rounds) out.  Per round: one pinned H2D copy, the
        kernels + the exchange, one pinned D2H copy, one synchronise."""
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1, 2)
        m = len(xy)
        with torch.cuda.stream(self.stream):
            self._host_io(m)
            hin = self._h_in.numpy()
            hin[:3] = np.asarray(pose, dtype=np.float64)
            hin[3:3 + 2 * m] = xy.ravel()
            self._d_in[:3 + 2 * m].copy_(self._h_in[:3 + 2 * m], non_blocking=True)
            pose_t, xy_t = self._d_in[:3], self._d_in[3:3 + 2 * m].view(m, 2)
            ids_t = None
            if ids is not None:
                self._h_ids.numpy()[:m] = np.asarray(ids, dtype=np.int32)
                self._d_ids[:m].copy_(self._h_ids[:m], non_blocking=True)
                ids_t = self._d_ids[:m]
            hres = self._h_res.numpy()
            hbase = self._h_res.data_ptr()     # pinned + mapped: the last handle's kernel stores its results here
            t0, first, rounds = 0, True, 0
            while t0 < m:
                rem = m - t0
                bufs = [h.shard_scan_begin(pose_t.data_ptr() if first else 0, rem, xy_t[t0:].data_ptr(),
                                           ids_t[t0:].data_ptr() if ids_t is not None else 0, rem <= MAX_ROUND)
                        for h in self.handles()]
                self.exchange(bufs)
                hs = self.handles()
                for q, h in enumerate(hs):
                    tail = q == len(hs) - 1       # the last handle's finish hands the results to the host
                    if tail:
                        # association / done / pose / count go straight into the mapped pinned buffer (a few posted
                        # PCIe stores from the sequential kernel): no copy operation between it and the covariance pass
                        h.shard_scan_finish(hbase + 32 + 4 * t0, hbase + 28, hbase, hbase + 24, 0, 0, 0,
                                            self._res_ev.cuda_event)
                    else:
                        h.shard_scan_finish(self._r_assoc[t0:].data_ptr(), self._r_done.data_ptr(),
                                            self._r_pose.data_ptr(), self._r_cnt.data_ptr(), 0, 0, 0, 0)
                # results are final after the sequential phase; the covariance pass keeps running while the
                # host reads them and prepares the next round / scan
                self._res_ev.synchronize()
                rounds += 1
                first = False
                done = int(hres[28:32].view(np.int32)[0])
                if done < 0:
                    raise RuntimeError("sharded scan: the sequential kernel's grid barrier timed out (EKF_ERR_CUDA)")
                if done == 0 and rounds > 4 * m + 4:
                    raise RuntimeError("sharded scan makes no progress")
                t0 += done
            assoc = hres[32:32 + 4 * m].view(np.int32).copy()
            return assoc, int(hres[24:28].view(np.int32)[0]), hres[0:24].view(np.float64).copy(), rounds


class LocalShards(_Driver):
    """R row shards of one map as R handles on one GPU (test / single-GPU emulation of R ranks)."""

    def __init__(self, n_landmarks: int, shards: int, device_index: int = 0, peer: bool = False, **kw):
        self.hs = [EkfSlam(n_landmarks=n_landmarks, device=device_index, shard_rank=r, shard_count=shards, **kw)
                   for r in range(shards)]
        self._setup(torch.device("cuda", device_index))
        self.n = self.hs[0].n
        if peer and shards > 1:
            # the peer-memory exchange with plain device pointers: every handle stores into every handle's region
            bases = [h.shard_peer_export()[1] for h in self.hs]
            for r, h in enumerate(self.hs):
                for q, b in enumerate(bases):
                    if q != r:
                        h.shard_peer_attach(q, base_ptr=b)

    def handles(self):
        return self.hs

    def exchange(self, bufs):
        if all(c == 0 for (_, c) in bufs):
            return                                   # delivered through peer memory by the begin kernels
        ts = [alias(p, c, self.device) for (p, c) in bufs]
        total = ts[0].clone()
        for t in ts[1:]:
            total += t
        for t in ts:
            t.copy_(total)

    # fixtures applied to every shard
    def fill_procedural(self, seed, count, grid_w):
        for h in self.hs:
            h.fill_procedural(seed, count, grid_w)

    def set_previous_xy(self, x, y):
        for h in self.hs:
            h.set_previous_xy(x, y)

    def state(self):
        return self.hs[0].state()

    def sigma_block(self, i0, i1, j0, j1):
        return sum(h.sigma_block(i0, i1, j0, j1) for h in self.hs)

    def sigma(self):
        return self.sigma_block(0, self.n, 0, self.n)

    def close(self):
        for h in self.hs:
            h.close()


class ShardedEkf(_Driver):
    """This process' row shard of one map in a torch.distributed job.  The per-round exchange goes through
    peer memory (CUDA IPC over NVLink: the begin kernel stores this rank's columns into every rank's buffer)
    when peer=True, else through an NCCL all-reduce; the rare side-state sync always uses the all-reduce."""

    def __init__(self, n_landmarks: int, rank: int, world: int, device_index: int, stream=None, group=None,
                 peer: bool = None, **kw):
        import os
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.h = EkfSlam(n_landmarks=n_landmarks, device=device_index, shard_rank=rank, shard_count=world, **kw)
        self._setup(torch.device("cuda", device_index), stream)
        self.n = self.h.n
        if peer is None:
            peer = os.environ.get("EKF_SHARD_PEER", "1") != "0"
        self.peer = bool(peer) and world > 1
        if self.peer:
            handle, _, _ = self.h.shard_peer_export()
            everyone = [None] * world
            dist.all_gather_object(everyone, handle, group=group)
            for q, hq in enumerate(everyone):
                if q != rank:
                    self.h.shard_peer_attach(q, ipc_handle=hq)
            dist.barrier(group=group)          # every rank has every region mapped before the first push

    def handles(self):
        return [self.h]

    def exchange(self, bufs):
        (p, c), = bufs
        if c == 0:
            return                                   # delivered through peer memory by the begin kernel
        self.dist.all_reduce(alias(p, c, self.device), op=self.dist.ReduceOp.SUM, group=self.group)

    def close(self):
        self.h.close()


def filter_range(n_filters: int, rank: int, world: int):
    """[lo, hi) of the independent filters rank `rank` steps in batched mode (no communication)."""
    return n_filters * rank // world, n_filters * (rank + 1) // world

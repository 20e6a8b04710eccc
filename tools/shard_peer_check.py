#!/usr/bin/env python
"""Under torchrun (one rank per GPU): the row-sharded scan with the peer-memory exchange against the same
scan with the NCCL all-reduce exchange - associations, state and this rank's rows of Sigma must be
bit-identical (both deliver the same doubles; only the transport differs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/shard_peer_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ekf_slam_b200 import shard  # noqa: E402
import bench  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N, m, K = 3000, 20, int(os.environ.get("SHARD_CHECK_SCANS", "6"))    # SHARD_CHECK_SCANS=3000: soak of the exchange
side = int(np.ceil(np.sqrt(N)))
out = []
for peer in (True, False):
    drv = shard.ShardedEkf(N, rank, world, local, peer=peer, max_scan=m)
    drv.h.fill_procedural(7, N - 8, side)
    drv.sync()
    poses, xy, ids = bench.make_workload(drv.h.state(), N - 8, m, K, 107)
    xy[-1, :3] += 0.45                                   # the last scan also creates landmarks
    if K > 8:
        xy[K // 2, :2] -= 0.45                           # ... and so does one in the middle of a soak
    drv.h.set_previous_xy(poses[0][1], poses[0][2])
    if K <= 8:
        assoc = [drv.scan(poses[s], xy[s])[0] for s in range(K)]
    else:
        # soak: device-resident, back to back, no host round trip between scans (what bench.py times) - the exchange's
        # buffer recycling (completion flags, credits, sentinel) is exercised at full rate
        d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
        d_assoc = torch.full((K, m), -9, dtype=torch.int32, device="cuda")
        d_done = torch.zeros(K, dtype=torch.int32, device="cuda")
        for s in range(K):
            drv.scan_dev(d_pose[s], d_xy[s], d_assoc[s], check=False, done_t=d_done[s:s + 1])
        torch.cuda.synchronize()
        assert bool((d_done == m).all()), "a round stopped early"
        assoc = d_assoc.cpu().numpy()
    rb, re, _ = drv.h.shard_info() if hasattr(drv.h, "shard_info") else (0, 0, 0)
    rows = drv.h.sigma_block(0, drv.h.n, 0, drv.h.n)
    out.append((np.array(assoc), drv.h.state().copy(), rows.copy(), drv.h.last_status()))
    drv.close()
(a1, s1, r1, st1), (a0, s0, r0, st0) = out
ok = np.array_equal(a1, a0) and np.array_equal(s1, s0) and np.array_equal(r1, r0) and st1 == 0 and st0 == 0
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("peer-memory exchange == all-reduce exchange on every rank:", bool(flag.item()),
          "| scans:", K, "| new landmarks in the last scan:", int((a1[-1] >= N - 8).sum()), "| non-zero Sigma entries held by rank 0:",
          int(np.count_nonzero(r1)))
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)

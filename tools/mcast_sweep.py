"""Fused-broadcast GEMM step vs route and announcement group, strong-scaled (torchrun, N ranks).
    torchrun --nproc-per-node N tools/mcast_sweep.py [hgemm|linear_s8|sgemm]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench
from hpc_b200 import capi, multi
local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
torch.cuda.set_device(local)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = capi.Context(local)
peers = multi.PeerExchange(dist, torch.device("cuda", local))
for name in (sys.argv[1:] or ["hgemm", "linear_s8"]):
    W = bench.make_workload(name, torch, ctx, world, rank, dist, peers, "mcast", share=world)
    def t(fn, steps=20):
        ms, _, _ = bench.time_steps(torch, fn, steps, 3, dist)
        return ms / steps
    res = {"resident": t(W.alts["b_resident"][0])}
    for k in W.alts:
        if k.startswith("bcast_"):
            res[k] = t(W.alts[k][0])
    for g in (0, 1, 2, 4, 8, 16, 64):
        ctx.debug_set("mc_group", g)
        res["mcast_group_%d" % g] = t(W.step)
    ctx.debug_set("mc_group", 0)
    v = W.verify()
    if rank == 0:
        print(name, "world", world, {k: round(v_, 4) for k, v_ in res.items()}, "verified", v.get("ok"), flush=True)
    del W
    torch.cuda.empty_cache()
dist.barrier()
dist.destroy_process_group()

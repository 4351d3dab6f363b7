"""How fast does B arrive when there is (almost) nothing to multiply?  Row-sharded GEMMs with 256 rows per rank, so the
step time is the fused broadcast itself.   torchrun --nproc-per-node N tools/bcast_only.py"""
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
st = lambda: torch.cuda.current_stream().cuda_stream
for name, n, m in (("hgemm", 16384, 16384 // world), ("linear_s8", 8192, 8192 // world)):
    if name == "hgemm":
        A = torch.randint(-10, 10, (m, n), device="cuda").half(); B = torch.randint(-10, 10, (n, n), device="cuda").half()
        C = torch.zeros((m, n), dtype=torch.float16, device="cuda")
    else:
        A = torch.randint(0, 17, (m, n), device="cuda", dtype=torch.int32).to(torch.int8)
        B = torch.randint(-8, 9, (n, n), device="cuda", dtype=torch.int32).to(torch.int8)
        C = torch.zeros((m, n), dtype=torch.int32, device="cuda")
    res = {}
    for mode in (["chain", "star"] if world > 2 else ["star"]):
        R = peers.route(ctx, B, mode)
        def step():
            R.activate()
            if name == "hgemm":
                ctx.mg_hgemm_f16(m, n, n, 1.0, A.data_ptr(), n, R.b_up, R.b_local, n, 0.0, C.data_ptr(), n, rank == 0, capi.FLOAT16, st())
            else:
                ctx.mg_linear_s8(m, n, n, A.data_ptr(), R.b_up, R.b_local, C.data_ptr(), rank == 0, st())
        for bufs in (4, 3, 2):
            ctx.debug_set("copy_bufs", bufs)
            ms, _, _ = bench.time_steps(torch, step, 20, 3, dist)
            ok = True if rank == 0 else R.local_equals(B)
            res["%s_bufs%d" % (mode, bufs)] = (round(ms / 20, 4), "%.0f GB/s" % (B.numel() * B.element_size() / (ms / 20 * 1e-3) / 1e9), ok)
        ctx.debug_set("copy_bufs", 0)
    if rank == 0:
        print(name, "rows per rank", m, "B MiB", (B.numel() * B.element_size()) >> 20, res, flush=True)
    del A, B, C
    torch.cuda.empty_cache()
dist.barrier(); dist.destroy_process_group()

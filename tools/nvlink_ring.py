"""NVLink pull bandwidth when every GPU pulls at the same time (torchrun, N ranks): ring (rank r from r-1: every
GPU receives AND serves), disjoint pairs (odd ranks from even ranks), star (everybody from GPU 0).  One row per
pattern, one column per rank, GB/s measured by cuxb_probe_nvlink on that rank."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from hpc_b200 import capi
local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
torch.cuda.set_device(local)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = capi.Context(local)
def run(name, peer, active, mode="bulk", depth=4):
    v = torch.zeros(1, device="cuda")
    dist.barrier(); torch.cuda.synchronize()
    if active:
        v[0] = ctx.probe_nvlink(peer, mode, depth, 1 << 30, 4)
    g = [torch.zeros(1, device="cuda") for _ in range(world)]
    dist.all_gather(g, v)
    if rank == 0:
        print("%-34s" % name, " ".join("%6.0f" % float(x[0]) for x in g), flush=True)
run("single pair 1<-0 (others idle)", 0, rank == 1)
run("ring r <- r-1, bulk depth 4", (rank - 1) % world, True)
run("ring r <- r-1, bulk depth 2", (rank - 1) % world, True, depth=2)
run("ring r <- r-1, bulk depth 8", (rank - 1) % world, True, depth=8)
run("ring r <- r-1, kernel pull", (rank - 1) % world, True, mode="pull")
run("ring r <- r-1, copy engine", (rank - 1) % world, True, mode="memcpy")
run("pairs odd <- even", rank - 1, rank % 2 == 1)
run("star r <- 0", 0, rank != 0)
run("chain r <- r-1 (rank 0 idle)", rank - 1, rank != 0)
dist.barrier()
dist.destroy_process_group()

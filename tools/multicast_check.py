"""Does this box give us an NVSwitch multicast mapping?  torchrun --nproc-per-node 2 tools/multicast_check.py"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = symm.empty(1 << 20, dtype=torch.float32, device="cuda")
h = symm.rendezvous(t, dist.group.WORLD)
mc = getattr(h, "multicast_ptr", None)
if dist.get_rank() == 0:
    print("symmetric memory ok; multicast_ptr =", hex(mc) if mc else mc, "| has_multicast_support:",
          getattr(symm, "has_multicast_support", lambda *a: "n/a")("cuda", local) if hasattr(symm, "has_multicast_support") else "n/a")
    import ctypes
    cu = ctypes.CDLL("libcuda.so.1")
    v = ctypes.c_int()
    # CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 132
    cu.cuInit(0)
    cu.cuDeviceGetAttribute(ctypes.byref(v), 132, local)
    print("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED =", v.value)
    # CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED = 128
    cu.cuDeviceGetAttribute(ctypes.byref(v), 128, local)
    print("CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED =", v.value)
dist.barrier()
dist.destroy_process_group()

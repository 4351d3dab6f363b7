"""Host-side logic of the multi-GPU paths (one process per GPU, torch.distributed).

Only what the hot path needs (north_star):
  * reductions: a per-GPU partial, then ONE all-reduce of the scalar            -> allreduce_sum / nrm2_combine
  * int32 scan: per-GPU totals exchanged, every rank seeds its local scan with
    the sum of the lower ranks' totals (what MPI_Exscan does in the reference's
    mpi/base_reduce_alltoall_scan.cpp:101)                                       -> scan_seed
  * GEMM / linear: rows of A and C sharded (the decomposition of
    mpi/alg_matrix_multiply.cpp:52-67); B lives on rank 0 and is pulled by the
    GEMM kernels themselves, rank to rank                                        -> shard_rows, chain_plan
The functions take tensors on whatever device the process group's backend serves (cuda for nccl,
cpu for gloo), so the same code is exercised by the world_size-2 gloo tests on CPU.
"""
import torch


def shard_range(n, world, rank):
    """Contiguous slice [lo, hi) of n items owned by `rank`; the first n % world ranks get one more."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_rows(M, world, rank):
    return shard_range(M, world, rank)


def chain_plan(world, rank, mode="chain"):
    """Who pulls B from whom in the fused broadcast.  Returns (source, flags_up, flags_down) as ranks
    (None = not applicable): `source` is the rank whose copy of B this rank's kernel reads, `flags_up` the
    rank whose readiness flags it polls (None when the source is the root, whose B is complete),
    `flags_down` the rank that pulls from this one and must finish a call before the next overwrites B.
    mode "star": everybody pulls from rank 0 and nobody needs flags.  The root never pulls."""
    if rank == 0 or world == 1:
        return None, None, None
    if mode == "star" or world == 2:
        return 0, None, None
    return rank - 1, (rank - 1 if rank >= 2 else None), (rank + 1 if rank + 1 < world else None)


def allreduce_sum(dist, partial):
    """One all-reduce of the per-GPU partial (in place).  Sum order across ranks is NCCL's; with the
    reference's integer-valued fills it is exact in any order."""
    if dist is not None and dist.get_world_size() > 1:
        dist.all_reduce(partial)
    return partial


def nrm2_combine(dist, local_norm, scratch=None):
    """Combine per-GPU Euclidean norms: sqrt(sum_r norm_r^2)."""
    if dist is None or dist.get_world_size() == 1:
        return local_norm
    sq = scratch if scratch is not None else torch.empty_like(local_norm)
    torch.mul(local_norm, local_norm, out=sq)
    dist.all_reduce(sq)
    torch.sqrt(sq, out=local_norm)
    return local_norm


def wrap_i32(v):
    """Python int -> the int32 it wraps to (two's complement)."""
    return (int(v) + 2 ** 31) % 2 ** 32 - 2 ** 31


def scan_seed(dist, local_total, gathered=None):
    """Exclusive prefix over ranks of the per-GPU int32 totals, with int32 wrap-around.
    local_total: 1-element int32 tensor.  Returns a Python int to pass as `seed`."""
    if dist is None or dist.get_world_size() == 1:
        return 0
    world, rank = dist.get_world_size(), dist.get_rank()
    if gathered is None:
        gathered = torch.empty(world, dtype=torch.int32, device=local_total.device)
    dist.all_gather_into_tensor(gathered, local_total)
    if rank == 0:
        return 0
    return wrap_i32(int(gathered[:rank].to(torch.int64).sum().item()))


_TYPESTR = {torch.float32: "<f4", torch.float16: "<f2", torch.bfloat16: None, torch.int8: "|i1", torch.int16: "<i2",
            torch.int32: "<i4", torch.int64: "<i8"}


class _RawCuda:
    """A raw device pointer dressed up for torch.as_tensor (zero copy, __cuda_array_interface__)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerExchange:
    """Per-rank exchange buffers for the fused multi-GPU kernels (cuxb_mg_*), one process per GPU.

    Plumbing only: every rank allocates its buffer with cuxb_malloc, exports it as a 64-byte handle
    (cuxb_ipc_export), the handles travel through torch.distributed's all-gather, and every rank imports the
    others' (cuxb_ipc_import) -- after which ptrs[r] is rank r's buffer as seen from this process.  The
    exchange itself is done by the kernels with plain loads and stores on these pointers.  Where the
    container does not allow CUDA IPC between the ranks, torch's symmetric memory provides the same mapping
    (self.plumbing says which one is in use); the NVSwitch multicast route always needs it, because a
    multicast address cannot be made from an IPC handle.
    """

    def __init__(self, dist, device):
        self.dist, self.device = dist, device
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.plumbing = "cuxb_ipc_export / cuxb_ipc_import (CUDA IPC handles all-gathered through torch.distributed)"
        self._keep = []
        try:
            self.buf = self._shared(torch.empty(64, dtype=torch.int64, device=device), use_ipc=True)
        except Exception as ex:
            self.plumbing = "torch symmetric memory (CUDA IPC unavailable here: %s)" % str(ex)[:120]
            self.buf = self._shared(torch.empty(64, dtype=torch.int64, device=device), use_ipc=False)
        self.ptrs = self.buf.ptrs
        torch.cuda.synchronize()
        dist.barrier()  # every buffer is zero before anybody's first exchange

    def _shared(self, like, use_ipc):
        """A zero-filled buffer of like's shape / dtype on every rank, mapped into every process."""
        from . import capi

        class _Sym:
            pass
        out = _Sym()
        dist = self.dist
        if use_ipc:
            nbytes = max(256, like.numel() * like.element_size())
            p = capi.dmalloc(nbytes)
            capi.check(capi.lib().cuxb_memset(p, 0, nbytes, None), "cuxb_memset")
            ok = torch.ones(1, device=self.device)
            ptrs, err = [0] * self.world, None
            try:
                handles = [None] * self.world
                dist.all_gather_object(handles, capi.ipc_export(p))
                for r in range(self.world):
                    ptrs[r] = p if r == self.rank else capi.ipc_import(handles[r])
            except Exception as ex:
                err = ex
                ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)    # all ranks take the same path
            if float(ok[0]) == 0.0:
                raise RuntimeError(str(err) if err else "a peer could not import the handle")
            ts = _TYPESTR.get(like.dtype)
            if ts is None:   # bf16 has no array-interface type string: view the bytes as int16
                out.tensor = torch.as_tensor(_RawCuda(p, like.shape, "<i2"), device=self.device).view(like.dtype)
            else:
                out.tensor = torch.as_tensor(_RawCuda(p, like.shape, ts), device=self.device)
            out.handle, out.ptrs = None, ptrs
        else:
            import torch.distributed._symmetric_memory as symm
            out.tensor = symm.empty(tuple(like.shape), dtype=like.dtype, device=like.device)
            out.tensor.zero_()
            out.handle = symm.rendezvous(out.tensor, dist.group.WORLD)
            out.ptrs = [int(x) for x in out.handle.buffer_ptrs]
        self._keep.append(out)
        return out

    def symmetric_like(self, t, need_multicast=False):
        """A peer-mapped tensor with t's shape/dtype on every rank: .tensor is this rank's instance, .ptrs[r]
        is rank r's instance as mapped into this process (.handle.multicast_ptr with need_multicast)."""
        return self._shared(t, use_ipc=(not need_multicast) and self.plumbing.startswith("cuxb_ipc"))

    def route(self, ctx, B, mode):
        """How B reaches this rank in the fused-broadcast GEMMs.  Allocates a symmetric copy of B (the root
        fills its instance from B) and, for "chain", the symmetric flag buffers.  Returns an object with
          .b_up / .b_local   the pointers to pass as b_root / b_local (bt_root / bt_local)
          .activate()        select this route on ctx (host-side state; call before each step)
          .clear_local()     zero this rank's copy (tests: prove that a call really delivered B)
          .local_equals(B)   this rank's copy == B
          .describe          one line for the bench output"""
        import torch
        import torch.distributed as dist
        from . import capi
        r, w = self.rank, self.world
        if mode == "mcast" and w < 3:
            raise RuntimeError("multicast all-gather needs 3+ ranks (with two, the one receiver pulls from the root)")
        R = self.symmetric_like(B, need_multicast=(mode == "mcast"))
        if r == 0:
            R.tensor.copy_(B)
        R.b_local = R.ptrs[r]
        src, up, down = chain_plan(w, r, mode)
        R.b_up = R.ptrs[src if src is not None else 0]
        if mode == "mcast":
            # scatter from the root + NVSwitch multicast all-gather between the receivers (3+ ranks)
            if w < 3:
                raise RuntimeError("multicast all-gather needs 3+ ranks (with two, the one receiver pulls from the root)")
            mc = int(getattr(R.handle, "multicast_ptr", 0) or 0)
            F = self.symmetric_like(torch.empty(capi.MG_MCAST_FLAGS_BYTES // 4, dtype=torch.int32, device=B.device),
                                    need_multicast=True)
            fmc = int(getattr(F.handle, "multicast_ptr", 0) or 0)
            if not mc or not fmc:
                raise RuntimeError("no NVSwitch multicast mapping available for symmetric memory on this box")
            F.tensor.zero_()
            R.flags = F
            R.b_up = R.ptrs[0]          # every receiver pulls its share straight from the root's B
            if r == 0:
                R.activate = lambda: (ctx.mg_set_chain(0, 0, 0), ctx.mg_set_multicast(0, 0, 0, 0, 0))
            else:
                R.activate = lambda: (ctx.mg_set_chain(0, 0, 0), ctx.mg_set_multicast(w, r, F.ptrs[r], fmc, mc))
            R.describe = ("every receiver pulls 1/%d of B from the root and stores it to an NVSwitch multicast address "
                          "(scatter + multicast all-gather: the root sends B once, no hops)" % (w - 1))
            R.clear_local = (lambda: R.tensor.zero_()) if r > 0 else (lambda: None)
            R.local_equals = lambda Bt: bool(torch.equal(R.tensor, Bt))
            torch.cuda.synchronize()
            dist.barrier()
            return R
        if mode == "chain" and w > 2:
            F = self.symmetric_like(torch.empty(capi.MG_CHAIN_FLAGS_BYTES // 4, dtype=torch.int32, device=B.device))
            F.tensor.zero_()
            R.flags = F
            chain = (F.ptrs[r], F.ptrs[up] if up is not None else 0, F.ptrs[down] if down is not None else 0)
            R.activate = (lambda: (ctx.mg_set_multicast(0, 0, 0, 0, 0), ctx.mg_set_chain(*chain))) if r > 0 else \
                (lambda: ctx.mg_set_multicast(0, 0, 0, 0, 0))
            R.describe = "from the rank before it (chain: each GPU sends B once)"
        else:
            R.activate = lambda: (ctx.mg_set_multicast(0, 0, 0, 0, 0), ctx.mg_set_chain(0, 0, 0))
            R.describe = "straight from rank 0 (star)"
        R.clear_local = (lambda: R.tensor.zero_()) if r > 0 else (lambda: None)
        R.local_equals = lambda Bt: bool(torch.equal(R.tensor, Bt))
        torch.cuda.synchronize()
        dist.barrier()
        return R

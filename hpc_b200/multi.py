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


class PeerExchange:
    """Per-rank exchange buffers for the fused multi-GPU kernels (cuxb_mg_*), one process per GPU.

    Uses torch's symmetric memory (CUDA IPC / fabric handles under the hood) purely as the plumbing
    that maps every rank's CUXB_MG_EXCHANGE_BYTES buffer into every process; the exchange itself is
    done by the reduction kernel's finishing CTA with plain stores and loads on these pointers.
    """

    def __init__(self, dist, device):
        import torch.distributed._symmetric_memory as symm
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.buf = symm.empty(64, dtype=torch.int64, device=device)  # 512 bytes >= CUXB_MG_EXCHANGE_BYTES
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, dist.group.WORLD)
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        torch.cuda.synchronize()
        dist.barrier()  # every buffer is zero before anybody's first exchange

    def symmetric_like(self, t):
        """A symmetric (peer-mapped) tensor with t's shape/dtype on every rank: .tensor is this rank's
        instance, .ptrs[r] is rank r's instance as mapped into this process."""
        import torch.distributed._symmetric_memory as symm
        import torch.distributed as dist

        class _Sym:
            pass
        out = _Sym()
        out.tensor = symm.empty(tuple(t.shape), dtype=t.dtype, device=t.device)
        out.handle = symm.rendezvous(out.tensor, dist.group.WORLD)
        out.ptrs = [int(p) for p in out.handle.buffer_ptrs]
        return out

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
        R = self.symmetric_like(B)
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
            F = self.symmetric_like(torch.empty(capi.MG_MCAST_FLAGS_BYTES // 4, dtype=torch.int32, device=B.device))
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

    def chain(self, ctx, b_like):
        """Set up the chained broadcast (cuxb_mg_set_chain) for ctx: a symmetric copy of B and symmetric
        flags; rank r pulls from rank r - 1.  Returns (bsym, upstream_ptr): bsym.tensor is this rank's B
        (the root fills it), upstream_ptr the B to pass as b_root / bt_root."""
        import torch
        import torch.distributed as dist
        from . import capi
        bsym = self.symmetric_like(b_like)
        fsym = self.symmetric_like(torch.empty(capi.MG_CHAIN_FLAGS_BYTES // 4, dtype=torch.int32, device=b_like.device))
        fsym.tensor.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        r, w = self.rank, self.world
        src, up, down = chain_plan(w, r)
        if r > 0 and w > 2:   # with two ranks the chain is the star: no flags needed
            ctx.mg_set_chain(fsym.ptrs[r], fsym.ptrs[up] if up is not None else 0,
                             fsym.ptrs[down] if down is not None else 0)
        bsym.flags = fsym
        return bsym, bsym.ptrs[src if src is not None else 0]

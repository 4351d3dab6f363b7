"""world_size-2 (and 3) CPU tests of the multi-GPU host logic over gloo: sharding, the single
all-reduce of reduction partials, and the exclusive offsets of the sharded int32 scan.  The per-rank
"kernel" is the oracle here (there is no GPU); what is tested is that shard + local result + collective
reproduces the single-device answer."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from hpc_b200 import datagen as dg
from hpc_b200 import multi


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        a, b = dg.int_fill(n, -2, 2, tag=1), dg.int_fill(n, -2, 2, tag=2)
        lo, hi = multi.shard_range(n, world, rank)
        # dot: per-rank partial + one all-reduce
        part = torch.tensor([oracle.dot_f32(a[lo:hi], b[lo:hi])], dtype=torch.float32)
        multi.allreduce_sum(dist, part)
        # nrm2: combine per-rank norms
        nrm = torch.tensor([oracle.nrm2_f32(a[lo:hi])], dtype=torch.float32)
        multi.nrm2_combine(dist, nrm)
        # scan: totals exchanged, local scan seeded with the exclusive offset (wrap-around data)
        x = np.full(n, 3000, np.int32)
        x[::7] = -5
        tot = torch.tensor([oracle.sum_i32(x[lo:hi])], dtype=torch.int32)
        seed = multi.scan_seed(dist, tot)
        local = oracle.scan_i32(x[lo:hi], seed=seed)
        # gemm: rows of C sharded, B replicated
        M, N, K = 37, 24, 16
        A, B = dg.int_fill((M, K), -2, 5, tag=3), dg.int_fill((K, N), -2, 5, tag=4)
        r0, r1 = multi.shard_rows(M, world, rank)
        Cl = oracle.gemm_f32(A[r0:r1], B, np.zeros((r1 - r0, N), np.float32), 1.0, 0.0)
        q.put((rank, float(part[0]), float(nrm[0]), lo, hi, local, r0, r1, Cl))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_paths_reproduce_single_device_results(world):
    n = 1_000_003
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=180) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    a, b = dg.int_fill(n, -2, 2, tag=1), dg.int_fill(n, -2, 2, tag=2)
    x = np.full(n, 3000, np.int32)
    x[::7] = -5
    want_scan = oracle.scan_i32(x)
    assert want_scan.min() < 0  # the data really wraps
    M, N, K = 37, 24, 16
    A, B = dg.int_fill((M, K), -2, 5, tag=3), dg.int_fill((K, N), -2, 5, tag=4)
    want_C = oracle.gemm_f32(A, B, np.zeros((M, N), np.float32), 1.0, 0.0)
    covered = 0
    for rank, dot, nrm, lo, hi, local, r0, r1, Cl in res:
        assert dot == float(oracle.dot_f32(a, b))            # integer fill: exact in any order
        assert abs(nrm - float(oracle.nrm2_f32(a))) <= 1e-6 * nrm
        assert np.array_equal(local, want_scan[lo:hi])       # bit-exact, wrap-around included
        assert np.array_equal(Cl, want_C[r0:r1])
        covered += hi - lo
    assert covered == n


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 1 << 20, (1 << 20) + 5):
        for world in (1, 2, 3, 8):
            spans = [multi.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert multi.wrap_i32(2 ** 31) == -2 ** 31 and multi.wrap_i32(-2 ** 31 - 1) == 2 ** 31 - 1


def test_chain_plan_is_a_single_path_from_the_root():
    """Fused broadcast routing (hpc_b200.multi.chain_plan): every non-root rank has exactly one source, the
    sources form one chain rooted at rank 0, and up/down flag partners are each other's mirror image."""
    for world in range(1, 9):
        plans = [multi.chain_plan(world, r) for r in range(world)]
        assert plans[0] == (None, None, None)
        for r in range(1, world):
            src, up, down = plans[r]
            assert src is not None and 0 <= src < r
            hops, at = 0, r
            while at != 0:                      # following the sources reaches the root
                at = plans[at][0]
                hops += 1
            assert hops == (r if world > 2 else 1)
            if up is not None:
                assert plans[up][2] == r        # my upstream's downstream is me
            if down is not None:
                assert plans[down][1] == r and plans[down][0] == r
            assert (up is None) == (src == 0)   # the root's B needs no flags
        star = [multi.chain_plan(world, r, "star") for r in range(world)]
        assert all(p == (0, None, None) for p in star[1:])

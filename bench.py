#!/usr/bin/env python3
"""bench.py -- the measurement contract for the cux operator hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]
    torchrun ... bench.py --gpus N ...           (one rank per GPU; NCCL for rendezvous / barriers)

Prints ONE JSON line.  The primary workload is BASELINE.json configs[1]: the cux `dot` operator, fp32,
N = 2^28 elements (2 x 1 GiB, far larger than the 126 MB L2).  A step is one pass of the hot path.

  N = 1   value = bytes read / device time, inputs resident in HBM.
  N > 1   STRONG scaling by default (the named size is the whole job: every GPU takes 1/N of the
          elements, or of the rows of A and C, and the exchange -- the scalar, the scan offset, the
          broadcast of B -- is inside the step).  The weak-scaled figure (the named size per GPU) is
          reported beside it under "weak".  --scaling weak makes that one the primary line instead.
  e2e     the same metric through the host-buffer C-ABI call (cuxb_*_host): pinned host buffers are
          copied host->device every step and the result is read back.
  also    every other BASELINE config (sum, nrm2, scan 2^30, SGEMM 8192^3, fp16 GEMM 16384^3, int8
          linear 8192^3) measured the same way at EVERY N: >= 20 steps, its own roofline object, clock
          window and -- at N = 1 -- e2e, a second (non-integer) data set, burst and sustained rates,
          the CPU baseline and the library (cuBLAS through torch) context; plus configs[0], the cux
          gemm 1024^3 on the reference's own host path.
  verified  after each timed loop, OUTSIDE the timed region, every rank checks what the timed path left
          in memory: scalars against the oracle over the concatenated data, the scan by first element +
          all differences + an oracle prefix, GEMMs by sampled rows bit-exact against the oracle, the
          delivered copy of B, and fused == resident over the whole of C.  Any false makes the exit
          code non-zero.  (The oracle is the CHECKER here, never the thing timed.)

--impl reference times the reference's own CPU implementation of the same workload (oracle/_ref when
built, else the oracle port) on the host cores: K individually timed passes after W warm-up passes.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_DOT = 1 << 28
N_SCAN = 1 << 30
SGEMM_N = 8192
HGEMM_N = 16384
LIN8_N = 8192
FP32_FFMA_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4: 148 SMs x 128 lanes x 2 x clocks.max.sm
NVLINK_GBS_PER_DIRECTION = 900.0
ALL = ["dot", "sum", "nrm2", "scan", "sgemm", "hgemm", "linear_s8"]
SEED = 20240607


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tensor_tflops": p["bf16_tflops"],
                "tensor_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tensor_tflops": 1590.0, "tensor_tflops_sustained": 1400.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed regions (recipe in B200_PROFILING.md).  One
    sampler runs for the whole process; every workload reads the samples of its own time window."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def window(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)   # let the sample that covers t1 arrive
        sm, mx, reasons, power = [], [], set(), []
        for ts, line in list(self.lines):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                inside = t0 - 0.05 <= ts <= t1 + 0.15
                if inside:
                    sm.append(float(f[1]))
                    power.append(float(f[3]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            if inside:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()


# --------------------------------------------------------------------------- descriptions (both arms)
def describe(name, world=1, scaling="strong"):
    """metric / unit / dtype / config of a workload: shared by the GPU arm and the reference arm."""
    per = " per GPU" if (world > 1 and scaling == "weak") else (" in total, 1/%d of it per GPU" % world if world > 1 else "")
    d = {   # the metric string is the same at every N (the driver lines the per-N values up by it)
        "dot": ("reduce HBM GB/s (fp32 dot-product, N=2^28)", "GB/s", "f32",
                "cux dot fp32 N=2^28%s (configs[1]), integer-valued fill [-2,2] as cux's demo", True),
        "sum": ("reduce HBM GB/s (fp32 sum, N=2^28)", "GB/s", "f32", "fp32 sum-reduce N=2^28%s (configs[1])", True),
        "nrm2": ("reduce HBM GB/s (fp32 nrm2, N=2^28)", "GB/s", "f32", "cux nrm2 fp32 N=2^28%s", True),
        "scan": ("scan HBM GB/s (int32 inclusive scan, N=2^30)", "GB/s", "i32", "int32 prefix scan N=2^30%s (configs[4])", True),
        "sgemm": ("GEMM TFLOP/s (fp32 SGEMM 8192^3, CUDA cores)", "TFLOP/s", "f32",
                  "fp32 SGEMM 8192^3%s (configs[2]), integer-valued fill [-2,5] as cux's demo", False),
        "hgemm": ("GEMM TFLOP/s (fp16 GEMM 16384^3, fp32 accumulate, tcgen05)", "TFLOP/s", "f16",
                  "fp16 GEMM 16384^3%s, fp32 accumulate, fp16 C (configs[3]), integer-valued fill [-10,9] as gemm_fp16_wmma.cu:46", False),
        "linear_s8": ("GEMM TOP/s (int8 linear 8192^3)", "TOP/s", "s8",
                      "int8 linear 8192^3%s, int32 accumulate (configs[4]), A in [0,16], Bt in [-8,8]", False),
    }[name]
    hbm = d[4]
    cfg = {"workload": d[3] % per,
           "l2": "inputs larger than L2 (no flush needed)" if hbm else "GEMM operands exceed L2; no flush",
           "parallelism": "1 GPU" if world == 1 else
           ("%d GPUs, one rank per GPU, contiguous slice each, exchange inside the step" % world if hbm else
            "%d GPUs, one rank per GPU, rows of A and C sharded, B broadcast from rank 0 inside the step" % world)}
    return {"metric": d[0], "unit": d[1], "dtype": d[2], "config": cfg, "hbm": hbm}


# --------------------------------------------------------------------------- workloads (GPU)
class Workload:
    pass


def _gen_int(torch, shape, lo, hi, dtype, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return torch.randint(lo, hi + 1, shape, device="cuda", generator=g, dtype=torch.int32).to(dtype)


def _gen_float(torch, shape, kind, dtype, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    if kind == "normal":
        return torch.randn(shape, device="cuda", generator=g, dtype=torch.float32).to(dtype)
    return (torch.rand(shape, device="cuda", generator=g, dtype=torch.float32) * 2 - 1).to(dtype)


def make_workload(name, torch, ctx, world, rank, dist, peers=None, bcast="chain", share=1, data="int"):
    """Device-resident inputs for this rank's part of `name` plus the callables that run it.
    share: how many ranks the named size is divided over (world for strong scaling, 1 for weak).
    data: "int" = the reference's integer-valued fills; "float" = the second data set of SURVEY 8(d)."""
    from hpc_b200 import capi, multi
    W = Workload()
    W.torch, W.ctx, W.world, W.rank, W.dist, W.name, W.data, W.share = torch, ctx, world, rank, dist, name, data, share
    W.collective, W.alts, W.keep = "none (1 GPU)", {}, []
    stream = lambda: torch.cuda.current_stream().cuda_stream
    seed = SEED + 17 * rank + (1000 if data == "float" else 0)
    bseed = SEED + 99 + (1000 if data == "float" else 0)
    if name in ("dot", "sum", "nrm2"):
        n = N_DOT // share
        if data == "int":
            a = _gen_int(torch, (n,), -2, 2, torch.float32, seed)        # cux fill, demo.cpp:30
            b = _gen_int(torch, (n,), -2, 2, torch.float32, seed + 1) if name == "dot" else None
        else:
            a = _gen_float(torch, (n,), "uniform", torch.float32, seed)
            b = _gen_float(torch, (n,), "uniform", torch.float32, seed + 1) if name == "dot" else None
        res = torch.zeros(1, device="cuda")
        fn = {"dot": lambda: ctx.dot_f32(n, a.data_ptr(), b.data_ptr(), res.data_ptr(), stream()),
              "sum": lambda: ctx.sum_f32(n, a.data_ptr(), res.data_ptr(), stream()),
              "nrm2": lambda: ctx.nrm2_f32(n, a.data_ptr(), res.data_ptr(), stream())}[name]
        sq = torch.zeros(1, device="cuda")

        def step_nccl():
            fn()                                   # per-GPU partial (one kernel launch)
            if world > 1:                          # then ONE NCCL all-reduce of the scalar
                if name == "nrm2":
                    multi.nrm2_combine(dist, res, sq)
                else:
                    multi.allreduce_sum(dist, res)
        W.step = step_nccl
        if world > 1:
            W.collective = "nccl all-reduce of the scalar after the kernel"
            if peers is not None:
                W.alts["nccl"] = (step_nccl, "same kernel followed by one NCCL all-reduce of the scalar")
                W.step = {"dot": lambda: ctx.mg_dot_f32(n, a.data_ptr(), b.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream()),
                          "sum": lambda: ctx.mg_sum_f32(n, a.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream()),
                          "nrm2": lambda: ctx.mg_nrm2_f32(n, a.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream())}[name]
                W.collective = "fused: partials exchanged over NVLink peer memory inside the reduction kernel (no NCCL launch)"
        W.kernel = fn
        W.units = (8 if name == "dot" else 4) * n
        W.unit, W.bound = "GB/s", "hbm"
        W.launches_per_step = 1
        W.result = res

        def verify():
            """The scalar every rank holds after W.step vs the oracle over the CONCATENATED data."""
            import oracle
            W.step()
            torch.cuda.synchronize()
            got = float(res[0])
            ha = a.cpu().numpy()
            hb = b.cpu().numpy() if b is not None else None
            if data == "int":
                # integer-valued data: every partial sum is exact below 2^24, so the reference's serial
                # result over the whole vector is the sum of its results over the slices, in any order
                if name == "dot":
                    part = float(oracle.dot_f32(ha, hb))
                elif name == "sum":
                    part = float(oracle.sum_f32(ha))
                else:
                    part = oracle.dot_f64(ha, ha)[0]     # Nrm2HostV0 itself saturates at 2^26: fp64 truth
                allp = _gather_f64(torch, dist, part, world)
                want = float(np.sqrt(sum(allp))) if name == "nrm2" else float(np.float32(sum(allp)))
                ok = abs(got - want) <= (1e-6 * want if name == "nrm2" else 0.0)
                how = "== oracle (bit-exact, integer fill)" if name != "nrm2" else "vs fp64 truth, rel 1e-6 (Nrm2HostV0 saturates at 2^26)"
                return {"ok": bool(ok), "got": got, "want": want, "how": how}
            truth, mag = oracle.dot_f64(ha, hb if name == "dot" else (ha if name == "nrm2" else None))
            allt, allm = _gather_f64(torch, dist, truth, world), _gather_f64(torch, dist, mag, world)
            want = float(np.sqrt(sum(allt))) if name == "nrm2" else float(sum(allt))
            tol = 4 * 2.0 ** -24 * sum(allm) if name != "nrm2" else 1e-6 * want
            return {"ok": bool(abs(got - want) <= tol), "got": got, "want": want,
                    "how": "|gpu - fp64 truth| <= 4 eps32 sum|terms|"}
        W.verify = verify

        def make_e2e():
            pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)
            ha, hb = pin(a), (pin(b) if b is not None else None)
            hres = torch.zeros(1, dtype=torch.float32, pin_memory=True)
            call = {"dot": lambda: ctx.dot_f32_host(n, ha.data_ptr(), hb.data_ptr(), hres.data_ptr()),
                    "sum": lambda: ctx.sum_f32_host(n, ha.data_ptr(), hres.data_ptr()),
                    "nrm2": lambda: ctx.nrm2_f32_host(n, ha.data_ptr(), hres.data_ptr())}[name]
            return {"call": call, "h2d": W.units, "d2h": 4, "api": "cuxb_%s_f32_host" % name, "result": hres,
                    "combine": None if name == "nrm2" else "sum", "keep": (ha, hb)}
        W.make_e2e = make_e2e
    elif name == "scan":
        n = N_SCAN // share
        x = _gen_int(torch, (n,), -2, 2, torch.int32, seed)
        out = torch.empty_like(x)
        tot = torch.zeros(1, dtype=torch.int32, device="cuda")
        fn = lambda: ctx.scan_i32(n, x.data_ptr(), out.data_ptr(), 0, stream())
        W.kernel = W.step = fn
        W.launches_per_step = 1
        if world > 1:
            gathered = torch.zeros(world, dtype=torch.int32, device="cuda")

            def step_nccl():
                # phase 1: local total; exchange; phase 2: scan seeded with the exclusive offset
                ctx.sum_i32(n, x.data_ptr(), tot.data_ptr(), stream())
                seed_val = multi.scan_seed(dist, tot, gathered)
                ctx.scan_i32(n, x.data_ptr(), out.data_ptr(), seed_val, stream())
            W.step = step_nccl
            W.launches_per_step = 2
            W.collective = "nccl all-gather of per-GPU totals + host-computed seed"
            if peers is not None:
                W.alts["nccl"] = (step_nccl, "sum kernel, NCCL all-gather of the totals, host-computed seed, scan kernel")
                W.step = lambda: ctx.mg_scan_i32(n, x.data_ptr(), out.data_ptr(), world, rank, peers.ptrs, stream())
                W.collective = ("fused: totals exchanged over NVLink peer memory inside the reduction kernel; seed read "
                                "from device memory (no host sync, no NCCL launch); 12 B/element moved for 8 counted")
        W.units = 8 * n
        W.unit, W.bound = "GB/s", "hbm"
        W.result = out[-1:]

        def verify():
            """first element + every difference (together a complete proof) + an oracle prefix + the tail."""
            import oracle
            out.zero_()
            W.step()
            torch.cuda.synchronize()
            local_total = int(x.sum(dtype=torch.int64))
            totals = _gather_f64(torch, dist, float(local_total), world)
            offset = multi.wrap_i32(int(sum(totals[:rank]))) if world > 1 else 0
            ok_first = int(out[0]) == multi.wrap_i32(offset + int(x[0]))
            ok_diff = bool(torch.equal(out[1:] - out[:-1], x[1:]))
            m = min(n, 1 << 22)
            ok_prefix = bool(np.array_equal(out[:m].cpu().numpy(), oracle.scan_i32(x[:m].cpu().numpy(), seed=offset)))
            ok_tail = int(out[-1]) == multi.wrap_i32(offset + local_total)
            return {"ok": ok_first and ok_diff and ok_prefix and ok_tail, "first": ok_first, "differences": ok_diff,
                    "oracle_prefix_2p22": ok_prefix, "tail": ok_tail, "how": "bit-exact"}
        W.verify = verify

        def make_e2e():
            hx = torch.empty(n, dtype=torch.int32, pin_memory=True).copy_(x)
            hout = torch.empty(n, dtype=torch.int32, pin_memory=True)
            return {"call": lambda: ctx.scan_i32_host(n, hx.data_ptr(), hout.data_ptr(), 0), "h2d": 4 * n, "d2h": 4 * n,
                    "api": "cuxb_scan_i32_host", "result": hout[-1:], "combine": None, "keep": (hx, hout)}
        W.make_e2e = make_e2e if world == 1 else None   # sharded e2e would need the seed exchange on the host
    elif name in ("sgemm", "hgemm", "linear_s8"):
        beta = 0.0
        if name == "sgemm":
            n = SGEMM_N
            m = n // share
            if data == "int":
                A = _gen_int(torch, (m, n), -2, 5, torch.float32, seed)
                B = _gen_int(torch, (n, n), -2, 5, torch.float32, bseed)   # B identical on every rank
                C0 = None
            else:       # uniform(-1,1), beta = 3 (cfg3 secondary)
                A = _gen_float(torch, (m, n), "uniform", torch.float32, seed)
                B = _gen_float(torch, (n, n), "uniform", torch.float32, bseed)
                C0 = _gen_float(torch, (m, n), "uniform", torch.float32, seed + 5)
                beta = 3.0
            C = torch.zeros((m, n), device="cuda")
            plain = lambda Bp, Cp: ctx.sgemm_f32(m, n, n, 1.0, A.data_ptr(), n, Bp, n, beta, Cp, n, stream())
        elif name == "hgemm":
            n = HGEMM_N
            m = n // share
            if data == "int":
                A = _gen_int(torch, (m, n), -10, 9, torch.float16, seed)
                B = _gen_int(torch, (n, n), -10, 9, torch.float16, bseed)
            else:
                A = _gen_float(torch, (m, n), "normal", torch.float16, seed)
                B = _gen_float(torch, (n, n), "normal", torch.float16, bseed)
            C0 = None
            C = torch.zeros((m, n), dtype=torch.float16, device="cuda")
            plain = lambda Bp, Cp: ctx.hgemm_f16(m, n, n, 1.0, A.data_ptr(), n, Bp, n, 0.0, Cp, n, capi.FLOAT16, stream())
        else:
            n = LIN8_N
            m = n // share
            A = _gen_int(torch, (m, n), 0 if data == "int" else -8, 16 if data == "int" else 8, torch.int8, seed)
            B = _gen_int(torch, (n, n), -8, 8, torch.int8, bseed)
            C0 = None
            C = torch.zeros((m, n), dtype=torch.int32, device="cuda")
            plain = lambda Bp, Cp: ctx.linear_s8(m, n, n, A.data_ptr(), Bp, Cp, stream())
        W.m, W.n, W.A, W.B, W.C, W.C0, W.beta = m, n, A, B, C, C0, beta

        def kernel():
            if C0 is not None:
                C.copy_(C0)
            plain(B.data_ptr(), C.data_ptr())
        if C0 is not None:
            # beta != 0: C must start from C0 every step; the restore is timed separately and subtracted
            W.restore = lambda: C.copy_(C0)
        W.kernel = W.step = kernel
        fused_C = C
        if world > 1 and peers is not None:
            # B lives on rank 0 only; every other rank's GEMM kernel gets it over NVLink while multiplying
            # the chosen route first, then the two pull routes for comparison (the multicast route is measured
            # only when asked for: --bcast mcast)
            routes, order = {}, [bcast] + [x for x in ("chain", "star") if x != bcast]
            if world == 2:
                order = [x for x in order if x == "star"] or ["star"]   # with one receiver every route is: pull from the root
            for mode in order:
                try:
                    routes[mode] = peers.route(ctx, B, mode)
                except Exception as ex:                          # e.g. no multicast on this box: reported below
                    routes[mode] = ex
            W.bcast_unavailable = {k: str(v)[:160] for k, v in routes.items() if isinstance(v, Exception)}
            bcast = next(m for m in order if not isinstance(routes[m], Exception))   # the first route that exists
            W.bcast = bcast
            torch.cuda.synchronize()
            dist.barrier()
            is_root = rank == 0

            def make_step(mode):
                R = routes[mode]

                def step():
                    R.activate()
                    if name == "hgemm":
                        ctx.mg_hgemm_f16(m, n, n, 1.0, A.data_ptr(), n, R.b_up, R.b_local, n, 0.0, C.data_ptr(), n,
                                         is_root, capi.FLOAT16, stream())
                    elif name == "sgemm":
                        ctx.mg_sgemm_f32(m, n, n, 1.0, A.data_ptr(), n, R.b_up, R.b_local, n, 0.0, C.data_ptr(), n,
                                         is_root, stream())
                    else:
                        ctx.mg_linear_s8(m, n, n, A.data_ptr(), R.b_up, R.b_local, C.data_ptr(), is_root, stream())
                return step
            W.step = make_step(bcast)
            W.alts["b_resident"] = (kernel, "B already resident on every GPU: no broadcast inside the step")
            for mode, R in routes.items():
                if mode != bcast and not isinstance(R, Exception):
                    W.alts["bcast_" + mode] = (make_step(mode), R.describe)
            W.routes = routes
            W.collective = ("fused: B (%d MiB) lives on rank 0; every other rank's GEMM kernel receives it over NVLink "
                            "peer memory %s, overlapped with the mainloop (no NCCL launch); rows of C sharded"
                            % ((B.numel() * B.element_size()) >> 20, routes[bcast].describe))
            W.keep.append(routes)
            W.bcast_bytes = B.numel() * B.element_size()
            if W.bcast_unavailable:
                W.collective += "; unavailable here: %s" % W.bcast_unavailable

        def verify():
            """Sampled rows bit-exact (integer fill) or within the stated tolerance (float fill) vs the oracle;
            at N > 1 also: the delivered copy of B == B, and fused C == B-resident C over the whole matrix."""
            import oracle
            rep = {"how": "sampled rows vs oracle: bit-exact" if data == "int" or name == "linear_s8" else "sampled rows vs oracle within tolerance"}
            if world > 1 and peers is not None and rank != 0:
                W.routes[W.bcast].clear_local()    # the check must see THIS call's delivery, not an earlier one
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            if C0 is not None:
                C.copy_(C0)
            else:
                C.zero_()
            W.step()
            torch.cuda.synchronize()
            rows = sorted(set([0, 1, m // 2, m - 1]))
            hB = B.cpu().numpy()
            if name == "sgemm":
                hA = A[rows].cpu().numpy()
                hC0 = C0[rows].cpu().numpy() if C0 is not None else np.zeros((len(rows), n), np.float32)
                want = oracle.gemm_f32(hA, hB, hC0, 1.0, beta, order="ikj", threads=8)
                got = C[rows].cpu().numpy()
                if data == "int":
                    rep["rows"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
                else:
                    tol = 2 * n * 2.0 ** -24 + 4 * 2.0 ** -24 * abs(beta)
                    rep["rows"] = bool(np.max(np.abs(got - want)) <= tol)
                    rep["max_abs_err"], rep["tolerance"] = float(np.max(np.abs(got - want))), tol
                if data == "int":   # checksum of checksums over ALL rows: sum_j C[i,j] == A[i,:] . rowsum(B), exact in fp64
                    rep["checksum_all_rows"] = bool(torch.equal(C.double().sum(dim=1), A.double() @ B.double().sum(dim=1)))
            elif name == "hgemm":
                hA = A[rows].view(torch.int16).cpu().numpy().view(np.uint16)
                want = oracle.hgemm_f16(hA, hB.view(np.uint16), threads=8)
                got = C[rows].view(torch.int16).cpu().numpy().view(np.uint16)
                if data == "int":
                    rep["rows"] = bool(np.array_equal(got, want))
                else:
                    # one fp16 ulp of the output + the fp32 accumulation-order noise of a K-long sum of O(1) terms
                    # (8 eps32 K: near-zero outputs carry that noise in full, far above their own ulp)
                    g, w = oracle.f16_to_f32(got), oracle.f16_to_f32(want)
                    bad = np.abs(g - w) > 2.0 ** -10 * np.abs(w) + 8 * 2.0 ** -24 * n
                    rep["rows"] = not bool(bad.any())
                    rep["mismatches"], rep["elements"] = int(bad.sum()), int(bad.size)
                    rep["max_abs_err"], rep["tolerance"] = float(np.max(np.abs(g - w))), "2^-10 |w| + 8 eps32 K"
            else:
                hA = A[rows].cpu().numpy()
                want = oracle.linear_s8(hA, hB, threads=8)
                rep["rows"] = bool(np.array_equal(C[rows].cpu().numpy(), want))
                # sum_j C[i,j] == A[i,:] . colsum(Bt), exact in fp64 at these magnitudes
                rep["checksum_all_rows"] = bool(torch.equal(C.to(torch.int64).sum(dim=1),
                                                            (A.double() @ B.double().sum(dim=0)).to(torch.int64)))
            rep["rows_checked"] = rows
            if world > 1 and peers is not None:
                rep["b_delivered"] = True if rank == 0 else bool(W.routes[W.bcast].local_equals(B))
                C2 = torch.empty_like(C)
                plain(B.data_ptr(), C2.data_ptr())
                torch.cuda.synchronize()
                rep["fused_equals_resident_all_of_C"] = bool(torch.equal(C, C2))
                del C2
            rep["ok"] = all(v for k, v in rep.items() if isinstance(v, bool))
            return rep
        W.verify = verify

        def make_e2e():
            pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)
            hA, hB = pin(A), pin(B)
            hC = torch.empty(C.shape, dtype=C.dtype, pin_memory=True)
            if C0 is not None:
                hC.copy_(C0)
            if name == "sgemm":
                call = lambda: ctx.sgemm_f32_host(m, n, n, 1.0, hA.data_ptr(), n, hB.data_ptr(), n, 0.0, hC.data_ptr(), n)
            elif name == "hgemm":
                call = lambda: ctx.hgemm_f16_host(m, n, n, 1.0, hA.data_ptr(), n, hB.data_ptr(), n, 0.0, hC.data_ptr(), n,
                                                  capi.FLOAT16)
            else:
                call = lambda: ctx.linear_s8_host(m, n, n, hA.data_ptr(), hB.data_ptr(), hC.data_ptr())
            api = {"sgemm": "cuxb_sgemm_f32_host", "hgemm": "cuxb_hgemm_f16_f32acc_host", "linear_s8": "cuxb_linear_s8_s32_host"}[name]
            return {"call": call, "h2d": (A.numel() + B.numel()) * A.element_size(), "d2h": C.numel() * C.element_size(),
                    "api": api, "result": hC[-1, -1:], "combine": None, "keep": (hA, hB, hC)}
        W.make_e2e = make_e2e if world == 1 else None
        W.units = 2 * m * n * n
        W.unit = "TOP/s" if name == "linear_s8" else "TFLOP/s"
        W.bound = "tensor" if name != "sgemm" else "ffma"
        W.result = C[-1, -1:]
        W.launches_per_step = 1
    else:
        raise SystemExit("unknown workload " + name)
    return W


def _gather_f64(torch, dist, value, world):
    """[value of rank 0, ..., value of rank world-1] on every rank (NCCL all-gather, outside any timed region)."""
    if dist is None or world == 1:
        return [float(value)]
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    g = torch.empty(world, dtype=torch.float64, device="cuda")
    dist.all_gather_into_tensor(g, t)
    return [float(v) for v in g.cpu()]


def time_steps(torch, fn, steps, warmup, dist=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        torch.cuda.synchronize()
        ms = float(t[0])
    return ms, t0, t1


def rate(units, ms, unit):
    per_s = units / (ms * 1e-3)
    return per_s / 1e9 if unit == "GB/s" else per_s / 1e12


def peak_for(name, bound, peaks, probes):
    """(peak, unit, source) for the kernel timed ALONE (burst figure where there are two)."""
    if bound == "hbm":
        return peaks["hbm_gbs"], "GB/s", "MEASURED_PEAKS.json hbm_gbs (%s)" % peaks["source"]
    if name == "sgemm":
        return FP32_FFMA_PEAK_TFLOPS, "TFLOP/s", "FP32 FFMA 148 SM x 128 x 2 x 1.965 GHz (nominal; MEASURED_PEAKS.json has no FFMA figure)"
    if name == "linear_s8":
        if probes.get("tensor_i8"):
            return probes["tensor_i8"]["burst"], "TOP/s", ("cuxb_probe_tensor(int8), measured in this process: tcgen05.mma kind::i8 on "
                                                           "resident operands, burst (MEASURED_PEAKS.json has no int8 figure)")
        return 2 * peaks["tensor_tflops"], "TOP/s", "2 x MEASURED_PEAKS.json bf16_tflops (assumption: int8 = 2x bf16; %s)" % peaks["source"]
    return peaks["tensor_tflops"], "TFLOP/s", "MEASURED_PEAKS.json bf16_tflops burst (%s)" % peaks["source"]


def traffic_for(name):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json), flagged
    stale when the kernel's source file has changed since the capture."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tpath):
        return None, "no capture"
    t = json.load(open(tpath))
    src = {"dot": "reduce.cu", "sum": "reduce.cu", "nrm2": "reduce.cu", "scan": "scan.cu", "sgemm": "sgemm.cu",
           "hgemm": "umma_gemm.cu", "linear_s8": "umma_gemm.cu"}[name]
    sha = hashlib.sha1(open(os.path.join(ROOT, "hpc_b200", "csrc", src), "rb").read()).hexdigest()[:12]
    rec = t.get("source_sha1", {}).get(src)
    note = "profiles/traffic.json (%s)" % t.get("capture", "ncu --set full")
    if rec != sha:
        note += "; STALE: %s has changed since that capture (was %s, is %s)" % (src, rec, sha)
    return t.get(name), note


# --------------------------------------------------------------------------- one workload, fully measured
def measure(name, env, steps, primary=False):
    """Time, verify and describe one workload at this world size.  Returns a dict; when `primary` the caller
    lifts its keys into the contract's top-level line."""
    torch, ctx, dist, args = env["torch"], env["ctx"], env["dist"], env["args"]
    world, rank, peers, peaks, probes, sampler = env["world"], env["rank"], env["peers"], env["peaks"], env["probes"], env["sampler"]
    from hpc_b200 import capi
    share = world if args.scaling == "strong" else 1
    W = make_workload(name, torch, ctx, world, rank, dist, peers, args.bcast, share=share)
    out = {"workload": describe(name, world, args.scaling)["config"]["workload"], "scaling": args.scaling if world > 1 else "n/a (1 GPU)"}
    warmup = max(args.warmup, 3)
    restore_ms = 0.0
    l0 = capi.launch_count()
    ms_total, t0, t1 = time_steps(torch, W.step, steps, warmup, dist)
    launches = (capi.launch_count() - l0) - warmup * W.launches_per_step
    ms_step = ms_total / steps
    out.update({"value": rate(W.units * world, ms_step, W.unit), "unit": W.unit, "ms_per_step": ms_step, "steps": steps,
                "warmup": warmup, "gpu_launches": launches, "collective": W.collective})
    if world > 1 and name in ("sgemm", "hgemm", "linear_s8") and hasattr(W, "bcast_bytes"):
        out["nvlink"] = {"bytes_into_each_non_root_gpu_per_step": W.bcast_bytes,
                         "achieved_gbs_per_gpu_if_spread_over_the_step": W.bcast_bytes / (ms_step * 1e-3) / 1e9,
                         "ceiling_gbs_per_direction": NVLINK_GBS_PER_DIRECTION}
    # ---- variants of the same step (N > 1): NCCL after the kernel / B resident / other broadcast routes
    t_last = t1
    if W.alts:
        out["variants"] = {}
        for key, (fn, what) in W.alts.items():
            try:
                ms_v, _, t_last = time_steps(torch, fn, steps, 3, dist)
                out["variants"][key] = {"ms_per_step": ms_v / steps, "value": rate(W.units * world, ms_v / steps, W.unit), "what": what}
            except Exception as ex:
                out["variants"][key] = {"error": str(ex)[:200]}
        if "b_resident" in out["variants"] and "ms_per_step" in out["variants"]["b_resident"]:
            r = out["variants"]["b_resident"]["ms_per_step"]
            out["nvlink"]["broadcast_cost_ms"] = ms_step - r
            out["nvlink"]["achieved_gbs_per_gpu_over_the_exposed_time"] = (
                W.bcast_bytes / (max(ms_step - r, 1e-6) * 1e-3) / 1e9 if ms_step > r else None)
    # ---- the dominant kernel alone on the launching stream (roofline.achieved), its single-launch spread
    ms_k, _, t_last = time_steps(torch, W.kernel, steps, 3, None)
    ms_kernel = ms_k / steps
    n_single = max(5, min(steps, 50))
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_single)]
    for a_, b_ in evs:
        a_.record()
        W.kernel()
        b_.record()
    torch.cuda.synchronize()
    singles = sorted(a_.elapsed_time(b_) for a_, b_ in evs)
    if hasattr(W, "restore"):   # beta != 0 runs copy C0 back first: time the copy alone and take it out
        ms_r, _, _ = time_steps(torch, W.restore, steps, 2, None)
        restore_ms = ms_r / steps
        ms_kernel -= restore_ms
        singles = [s - restore_ms for s in singles]
    # ---- sustained: the kernel back to back for >= 2 s (GEMMs: the board's power limit sets the clock)
    sustained = None
    if W.bound != "hbm" and world == 1 and not args.quick:
        reps = max(steps, int(2000.0 / max(ms_kernel, 1e-3)) + 1)
        ms_s, ts0, t_last = time_steps(torch, W.kernel, reps, 0, None)
        sustained = {"ms_per_launch": ms_s / reps - restore_ms, "launches": reps, "seconds": ms_s * 1e-3,
                     "clocks": sampler.window(ts0, t_last) if rank == 0 else None}
        sustained["achieved"] = rate(W.units, sustained["ms_per_launch"], W.unit)
    elif rank == 0:
        # nvidia-smi cannot sample faster than ~10 Hz: keep the GPU under the same load a little longer
        while time.time() - t0 < 0.7:
            _, _, t_last = time_steps(torch, W.kernel, max(steps, 50), 0, None)
    if rank == 0:
        out["clocks"] = sampler.window(t0, t_last)
        out["clocks"]["window"] = "this workload's timed steps and the kernel-only loops that follow them"
    achieved = rate(W.units, ms_kernel, W.unit)
    peak, punit, pnote = peak_for(name, W.bound, peaks, probes)
    roofline = {"bound": "hbm" if W.bound == "hbm" else "tensor" if W.bound == "tensor" else "ffma",
                "achieved": achieved, "peak": peak, "unit": punit, "frac": achieved / peak,
                "kernel_ms": ms_kernel, "kernel_ms_median": singles[len(singles) // 2],
                "kernel_ms_min": singles[0], "kernel_ms_max": singles[-1], "single_launches_timed": n_single,
                "burst": {"achieved": rate(W.units, singles[0], W.unit), "what": "fastest single launch"},
                "peak_source": pnote, "algorithmic_units_per_launch": W.units}
    roofline["traffic"], roofline["traffic_source"] = traffic_for(name)
    if world > 1:
        roofline["note"] = "per-GPU kernel on this rank's 1/%d share" % share if share > 1 else "per-GPU kernel"
    if sustained is not None:
        speak = {"hgemm": peaks["tensor_tflops_sustained"], "sgemm": FP32_FFMA_PEAK_TFLOPS}.get(name)
        if name == "linear_s8":
            speak = probes["tensor_i8"]["sustained"] if probes.get("tensor_i8") else 2 * peaks["tensor_tflops_sustained"]
        sustained["peak"], sustained["frac"] = speak, sustained["achieved"] / speak
        roofline["sustained"] = sustained
    mh = {"dot": "hbm_read", "sum": "hbm_read", "nrm2": "hbm_read", "scan": "hbm_copy", "sgemm": "fp32",
          "hgemm": "tensor_f16", "linear_s8": "tensor_i8"}[name]
    if probes.get(mh):
        roofline["measured_here"] = dict(probes[mh])
        roofline["measured_here"]["frac"] = achieved / probes[mh]["burst"]
    out["roofline"] = roofline
    # ---- verification of what the timed path produces (outside the timed region; every rank)
    try:
        v = W.verify()
    except Exception as ex:
        v = {"ok": False, "error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}
    oks = _gather_f64(torch, dist, 1.0 if v.get("ok") else 0.0, world)
    v["all_ranks"] = [bool(x) for x in oks]
    out["verified"] = all(v["all_ranks"])
    out["verification"] = v
    out["device_result"] = float(W.result[0])
    # ---- end to end through the host-buffer C ABI (pinned host memory, H2D + kernel + D2H per step)
    if getattr(W, "make_e2e", None) is not None and (primary or world == 1) and not args.quick:
        out["e2e"] = run_e2e(W, env, steps)
    # ---- second data set (N = 1): non-integer values, the power-hungry case for the GEMMs
    if world == 1 and not args.quick and name != "scan":
        del W
        torch.cuda.empty_cache()
        out["second_data_set"] = second_data_set(name, env, steps)
        W = None
    # ---- weak-scaled sibling of the primary line
    if primary and world > 1 and args.scaling == "strong":
        del W
        torch.cuda.empty_cache()
        Ww = make_workload(name, torch, ctx, world, rank, dist, peers, args.bcast, share=1)
        ms_w, _, _ = time_steps(torch, Ww.step, steps, 3, dist)
        out["weak"] = {"value": rate(Ww.units * world, ms_w / steps, Ww.unit), "unit": Ww.unit, "ms_per_step": ms_w / steps,
                       "what": "the named size PER GPU (per-GPU work fixed as N grows), same step"}
        W = Ww
    del W
    torch.cuda.empty_cache()
    return out


def run_e2e(W, env, steps):
    torch, dist, world = env["torch"], env["dist"], env["world"]
    E = W.make_e2e()
    part = torch.zeros(1, device="cuda")

    def e2e_step():
        E["call"]()                                  # H2D of the inputs, kernel, D2H of the result: blocking
        if world > 1 and E["combine"] == "sum":      # the scalar of every rank, combined like the device path
            part.copy_(E["result"], non_blocking=True)
            dist.all_reduce(part)
            E["result"].copy_(part)
    e2e_steps = max(3, min(steps, 10 if W.bound == "hbm" else 5))
    for _ in range(2):
        e2e_step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t_a = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t_a], device="cuda")
    if dist is not None:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_ms = float(dt[0]) * 1e3 / e2e_steps
    return {"value": rate(W.units * world, e2e_ms, W.unit), "unit": W.unit, "h2d_bytes_per_step": int(E["h2d"]),
            "d2h_bytes_per_step": int(E["d2h"]), "ms_per_step": e2e_ms, "steps": e2e_steps,
            "api": "%s (pinned host buffers; host wall clock around the blocking call%s)" % (
                E["api"], "; + one all-reduce of the scalar" if world > 1 and E["combine"] else ""),
            "result": float(E["result"][0])}


def second_data_set(name, env, steps):
    """SURVEY 8(d)'s second data set next to each integer fill: dot/sum/nrm2 and SGEMM uniform(-1,1) (SGEMM with
    beta = 3), fp16 GEMM N(0,1), int8 linear signed A.  Same sizes, same kernels; checked within the stated
    tolerance (int8: bit-exact).  GEMMs: burst and sustained."""
    torch, ctx, peaks, probes, sampler = env["torch"], env["ctx"], env["peaks"], env["probes"], env["sampler"]
    what = {"dot": "uniform(-1,1)", "sum": "uniform(-1,1)", "nrm2": "uniform(-1,1)", "sgemm": "A, B, C uniform(-1,1), beta = 3",
            "hgemm": "A, B ~ N(0,1)", "linear_s8": "A signed in [-8,8] (LinearNormalInt8 only; dpbusd needs A >= 0)"}[name]
    try:
        W = make_workload(name, torch, ctx, 1, 0, None, None, share=1, data="float")
        ms, t0, t1 = time_steps(torch, W.kernel, steps, 3, None)
        ms /= steps
        restore_ms = 0.0
        if hasattr(W, "restore"):
            ms_r, _, _ = time_steps(torch, W.restore, steps, 2, None)
            restore_ms = ms_r / steps
            ms -= restore_ms
        rep = {"data": what, "value": rate(W.units, ms, W.unit), "unit": W.unit, "ms": ms}
        peak, _, pnote = peak_for(name, W.bound, peaks, probes)
        rep["frac"], rep["peak"], rep["peak_source"] = rep["value"] / peak, peak, pnote
        if W.bound != "hbm":
            reps = max(steps, int(2000.0 / max(ms, 1e-3)) + 1)
            ms_s, ts0, ts1 = time_steps(torch, W.kernel, reps, 0, None)
            sus = rate(W.units, ms_s / reps - restore_ms, W.unit)
            speak = {"hgemm": peaks["tensor_tflops_sustained"], "sgemm": FP32_FFMA_PEAK_TFLOPS}.get(name)
            if name == "linear_s8":
                speak = probes["tensor_i8"]["sustained"] if probes.get("tensor_i8") else 2 * peaks["tensor_tflops_sustained"]
            rep["sustained"] = {"achieved": sus, "peak": speak, "frac": sus / speak, "launches": reps, "seconds": ms_s * 1e-3,
                                "clocks": sampler.window(ts0, ts1)}
        else:
            rep["clocks"] = sampler.window(t0, t1)
        v = W.verify()
        rep["verified"], rep["verification"] = bool(v.get("ok")), v
        return rep
    except Exception as ex:
        return {"data": what, "error": "%s: %s" % (type(ex).__name__, str(ex)[:300]), "verified": False}


def library_context(name, env):
    """What the reference's GPU path calls for the same shape -- cuBLAS through torch (cublasSgemm /
    cublasSgemmEx, 0-frameworks/cux/src/operator/gemm.cu:237-238,266-269) -- timed once, for context only."""
    torch = env["torch"]
    try:
        if name == "sgemm":
            n = SGEMM_N
            old = torch.backends.cuda.matmul.allow_tf32
            torch.backends.cuda.matmul.allow_tf32 = False
            a, b = torch.rand(n, n, device="cuda"), torch.rand(n, n, device="cuda")
            f, what = (lambda: torch.matmul(a, b)), "torch.matmul fp32, TF32 off (cuBLAS SGEMM)"
        elif name == "hgemm":
            n = HGEMM_N
            a, b = torch.randn(n, n, device="cuda", dtype=torch.float16), torch.randn(n, n, device="cuda", dtype=torch.float16)
            f, what = (lambda: torch.matmul(a, b)), "torch.matmul fp16 (cuBLAS), N(0,1) data"
        elif name == "linear_s8":
            n = LIN8_N
            a = torch.randint(-8, 9, (n, n), device="cuda", dtype=torch.int8)
            b = torch.randint(-8, 9, (n, n), device="cuda", dtype=torch.int8).t()
            f, what = (lambda: torch._int_mm(a, b)), "torch._int_mm int8 (cuBLASLt)"
        else:
            return None
        ms, _, _ = time_steps(torch, f, 5, 2, None)
        if name == "sgemm":
            torch.backends.cuda.matmul.allow_tf32 = old
        return {"what": what, "ms": ms / 5, "value": 2.0 * n ** 3 / (ms / 5 * 1e-3) / 1e12,
                "unit": "TOP/s" if name == "linear_s8" else "TFLOP/s", "note": "library context, not part of this repo's path"}
    except Exception as ex:
        return {"error": str(ex)[:200]}


def run_probes(env):
    """The ceilings measured in THIS process by the library's own probes (N = 1; a few seconds)."""
    ctx, P = env["ctx"], {}
    from hpc_b200 import capi
    try:
        P["hbm_read"] = {"burst": ctx.probe_hbm("read", 2 << 30, 5), "unit": "GB/s", "what": "cuxb_probe_hbm read stream over 2 GiB, best of 5"}
        P["hbm_copy"] = {"burst": ctx.probe_hbm("copy", 2 << 30, 5), "unit": "GB/s", "what": "cuxb_probe_hbm copy stream over 2 GiB, best of 5"}
        P["fp32"] = {"burst": ctx.probe_fp32(True, 5), "unit": "TFLOP/s", "what": "cuxb_probe_fp32: packed FFMA2 chains, best of 5"}
        for key, kind in (("tensor_f16", capi.FLOAT16), ("tensor_i8", capi.INT8)):
            b, m = ctx.probe_tensor(kind, 1, 400)     # 400 launches of ~4-5 ms: ~2 s back to back
            bz, _ = ctx.probe_tensor(kind, 0, 5)
            P[key] = {"burst": b, "sustained": m, "burst_zero_operands": bz, "unit": "TFLOP/s" if key == "tensor_f16" else "TOP/s",
                      "what": "cuxb_probe_tensor: tcgen05.mma cta_group::2 256x256 on shared-memory-resident pseudo-random operands; "
                              "burst = best single launch, sustained = 400 launches back to back"}
    except Exception as ex:
        P["error"] = str(ex)[:300]
    return P


def config0_leg(env):
    """BASELINE configs[0]: cux gemm fp32 M=N=K=1024 on the reference's own host path -- GemmHostV4
    (gemm.cpp:161-218, AVX + OpenMP num_threads(6)) and MatrixMulMP (openmp/alg_matrix_multiply.cpp:33-51, all
    cores) timed, verified against the naive triple loop GemmHostV0 (gemm.cpp:13-34) -- next to the GPU kernel on
    the same inputs, verified against the same rows."""
    import oracle
    torch, ctx = env["torch"], env["ctx"]
    n = 1024
    R, F = oracle.ref(), oracle.ref_fast()
    rng = np.random.default_rng(SEED)
    A = rng.integers(-2, 6, size=(n, n)).astype(np.float32)
    B = rng.integers(-2, 6, size=(n, n)).astype(np.float32)
    C0 = rng.integers(-2, 6, size=(n, n)).astype(np.float32)
    rep = {"workload": "cux gemm_fp32 M=N=K=1024, alpha=1 beta=3, integer fill [-2,5] (configs[0]; demo.cpp:77,97)", "flop": 2.0 * n ** 3}
    rows = 64
    naive = C0[:rows].copy()
    t0 = time.perf_counter()
    if R is not None:
        R.ref_cux_gemm(0, rows, n, n, 1.0, A[:rows].copy(), n, B, n, 3.0, naive, n)
        who = "GemmHostV0 from oracle/_ref (the reference's own code)"
    else:
        naive = oracle.gemm_f32(A[:rows], B, C0[:rows], 1.0, 3.0, order="v0")
        who = "GemmHostV0 restatement (oracle port)"
    rep["naive"] = {"what": who + ", first %d rows (rows are independent)" % rows,
                    "gflops": 2.0 * rows * n * n / (time.perf_counter() - t0) / 1e9, "cores": 1}
    cpu = {}
    if R is not None:
        for key, fn, cores in (("GemmHostV4", lambda L, C: L.ref_cux_gemm(4, n, n, n, 1.0, A, n, B, n, 3.0, C, n), min(6, os.cpu_count() or 1)),
                               ("MatrixMulMP", lambda L, C: L.ref_omp_gemm(1, n, n, n, 1.0, A, n, B, n, C, n), R.ref_num_threads())):
            Cp = C0.copy()
            fn(R, Cp)          # parity build: the result that is verified
            if key == "MatrixMulMP":   # no beta in that sample (memset C): compare with the beta = 0 naive loop
                nb = np.zeros((rows, n), np.float32)
                R.ref_cux_gemm(0, rows, n, n, 1.0, A[:rows].copy(), n, B, n, 0.0, nb, n)
                ok = bool(np.array_equal(Cp[:rows], nb))
            else:
                ok = bool(np.array_equal(Cp[:rows], naive))
            ts = []
            for _ in range(7):   # -O3 build: the one that is timed
                Ct = C0.copy()
                t0 = time.perf_counter()
                fn(F, Ct)
                ts.append(time.perf_counter() - t0)
            cpu[key] = {"ms": float(np.median(ts)) * 1e3, "gflops": 2.0 * n ** 3 / float(np.median(ts)) / 1e9, "cores": cores,
                        "verified_vs_GemmHostV0": ok, "passes": len(ts)}
    else:
        Cp = oracle.gemm_f32(A, B, C0, 1.0, 3.0, order="ikj", threads=os.cpu_count() or 1)
        cpu["GemmHostV1_port_openmp"] = {"verified_vs_GemmHostV0": bool(np.array_equal(Cp[:rows], naive)), "cores": os.cpu_count()}
    rep["cpu"] = cpu
    dA, dB, dC0 = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda(), torch.from_numpy(C0).cuda()
    dC = dC0.clone()
    st = torch.cuda.current_stream().cuda_stream
    ctx.sgemm_f32(n, n, n, 1.0, dA.data_ptr(), n, dB.data_ptr(), n, 3.0, dC.data_ptr(), n, st)
    torch.cuda.synchronize()
    gok = bool(np.array_equal(dC[:rows].cpu().numpy(), naive))
    f = lambda: ctx.sgemm_f32(n, n, n, 1.0, dA.data_ptr(), n, dB.data_ptr(), n, 0.0, dC.data_ptr(), n, st)
    ms, _, _ = time_steps(torch, f, 50, 5, None)
    rep["gpu"] = {"ms": ms / 50, "tflops": 2.0 * n ** 3 / (ms / 50 * 1e-3) / 1e12, "verified_vs_GemmHostV0": gok,
                  "what": "cuxb_sgemm_f32 on the same inputs (timed with beta = 0 so that repeated launches do not grow C)"}
    rep["verified"] = gok and all(v.get("verified_vs_GemmHostV0", False) for v in cpu.values())
    rep["parallelism"] = "host path: replicas only (no GPU roofline)"
    return rep


# --------------------------------------------------------------------------- CPU reference arm
def cpu_reference_case(name):
    """(f, units, unit, cores, kind, what, extra) -- one pass of the reference's own CPU implementation of
    `name` on a BOUNDED sample of the workload.  Timed from the -O3 build of the reference's sources
    (oracle/_ref/libcuxref_fast.so) when present, else the oracle port."""
    import oracle
    R, L = oracle.ref_fast(), oracle.lib()
    kind = "reference" if R is not None else "port"
    rng = np.random.default_rng(SEED)
    ncores = os.cpu_count() or 1
    if name in ("dot", "sum", "nrm2"):
        n = N_DOT
        a = rng.integers(-2, 3, size=n).astype(np.float32)
        out = np.zeros(1, np.float32)
        extra = None
        if name == "dot":
            b = rng.integers(-2, 3, size=n).astype(np.float32)
            f = (lambda: R.ref_cux_dot_v0(n, a, b, out)) if R else (lambda: L.orc_dot_f32(n, a, b, out))
            what, units = "DotHostV0 (dot_product.cpp:7-13)", 8 * n
            if R is not None and hasattr(R, "ref_std_dot_mt"):
                extra = ("all_cores", lambda: R.ref_std_dot_mt(n, a, b, ncores, out), ncores,
                         "DotProductMTv1 (std/alg_vector_dot_product.cpp:19-56): the repo's own all-cores dot product -- not a cux op")
        elif name == "sum":
            f = (lambda: R.ref_reduce_host(a, out, n)) if R else (lambda: L.orc_sum_f32(n, a, out))
            what, units = "ReduceHost (reduce_fp32.cu:26-31)", 4 * n
        else:
            f = (lambda: R.ref_cux_nrm2_v0(n, a, out)) if R else (lambda: L.orc_nrm2_f32(n, a, out))
            what, units = "Nrm2HostV0 (nrm2.cpp:7-16)", 4 * n
        return f, units, "GB/s", 1, kind, what + " over the full N=2^28 vector(s); single thread, as in the reference (cux has no parallel host kernel for it)", extra
    if name == "scan":
        n = 1 << 27
        x = rng.integers(-2, 3, size=n).astype(np.int32)
        out = np.empty_like(x)
        extra = None
        if R is not None:
            xf, of = x.astype(np.float32), np.empty(n, np.float32)
            extra = ("reference_fp32_simd", lambda: R.ref_scan_f32(2, xf, n, of), 1,
                     "the reference's fp32 VectorScanSIMDv2 (vector_scan.cpp:66-79) on the same slice")
        return (lambda: L.orc_scan_i32(n, x, out, 0)), 8 * n, "GB/s", 1, "port", \
            "int32 restatement of VectorScanNormal (vector_scan.cpp:21-26) over N=2^27 (1/8 of the workload), single thread " \
            "(the reference has no int32 scan)", extra
    if name == "sgemm":
        n, rows = SGEMM_N, 96
        A = rng.integers(-2, 6, size=(rows, n)).astype(np.float32)
        B = rng.integers(-2, 6, size=(n, n)).astype(np.float32)
        C = np.zeros((rows, n), np.float32)
        units = 2.0 * rows * n * n
        if R is None:
            return (lambda: L.orc_gemm_f32_ikj(rows, n, n, 1.0, A, n, B, n, 0.0, C, n, ncores)), units, "TFLOP/s", ncores, "port", \
                "GemmHostV1 order, OpenMP over rows, %d-row slice of the 8192^3 problem (rows are independent)" % rows, None
        extra = ("GemmHostV4", lambda: R.ref_cux_gemm(4, rows, n, n, 1.0, A, n, B, n, 0.0, C, n), min(6, ncores),
                 "GemmHostV4 (gemm.cpp:161-218, AVX + OpenMP num_threads(6)) on the same slice")
        return (lambda: R.ref_omp_gemm(1, rows, n, n, 1.0, A, n, B, n, C, n)), units, "TFLOP/s", R.ref_num_threads(), kind, \
            "MatrixMulMP (openmp/alg_matrix_multiply.cpp:33-51, all cores) on a %d-row slice of the 8192^3 problem (rows are independent)" % rows, extra
    if name == "hgemm":
        n, rows = HGEMM_N, 1
        A16 = rng.integers(-10, 10, size=(rows, n)).astype(np.float16).view(np.uint16)
        B16 = rng.integers(-10, 10, size=(n, n)).astype(np.float16).view(np.uint16)
        C16 = np.zeros((rows, n), np.uint16)
        units = 2.0 * rows * n * n
        if R is None:
            return (lambda: L.orc_hgemm_f16(rows, n, n, 1.0, A16, n, B16, n, 0.0, C16.ctypes.data, n, 0, 1)), units, "TFLOP/s", 1, "port", \
                "%d-row slice of the 16384^3 problem" % rows, None
        extra = None
        if R.ref_cpu_has_avx512bf16():
            r2 = 64
            Ab = (rng.integers(-10, 10, size=(r2, n)).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
            Bb, Cb = B16.astype(np.uint16), np.zeros((r2, n), np.uint16)   # bit patterns only need to be valid bf16 for timing
            extra = ("LinearAvx512Bf16", lambda: R.ref_linear_bf16(1, r2, n, n, Ab, Bb, Cb), 1,
                     "fastest 16-bit CPU path in the repo, LinearAvx512Bf16 (linear.cpp:156-182, 1 thread) on a 64-row slice", 2.0 * r2 * n * n)
        return (lambda: R.ref_hgemm_host(rows, n, n, A16, n, B16, n, C16, n)), units, "TFLOP/s", 1, kind, \
            "GemmHostV1-half (gemm_fp16_wmma.cu:75-90) on a %d-row slice of the 16384^3 problem, single thread (rows are " \
            "independent; the full problem would take hours)" % rows, extra
    if name == "linear_s8":
        n, rows = LIN8_N, 128
        A = rng.integers(0, 17, size=(rows, n)).astype(np.int8)
        B = rng.integers(-8, 9, size=(n, n)).astype(np.int8)
        C = np.zeros((rows, n), np.int32)
        units = 2.0 * rows * n * n
        if R is not None and R.ref_cpu_has_avx512():
            return (lambda: R.ref_linear_s8(1, rows, n, n, A, B, C)), units, "TOP/s", 1, kind, \
                "LinearAvx512Int8 (linear.cpp:214-237, VNNI dpbusd; valid because A >= 0) on a %d-row slice of the 8192^3 problem, single thread as in the sample" % rows, None
        if R is not None:
            return (lambda: R.ref_linear_s8(0, rows, n, n, A, B, C)), units, "TOP/s", 1, kind, \
                "LinearNormalInt8 (linear.cpp:199-211) on a %d-row slice, single thread" % rows, None
        return (lambda: L.orc_linear_s8(rows, n, n, A, B, C, 1)), units, "TOP/s", 1, "port", "LinearNormalInt8 restatement, %d-row slice" % rows, None
    raise ValueError(name)


def _per_s(units, t, unit):
    return units / t / (1e9 if unit == "GB/s" else 1e12)


def time_cpu_case(case, warmup, steps, budget_s):
    """W untimed + up to K individually timed passes (fewer only if the budget runs out: the count is reported)."""
    f, units, unit, cores, kind, what, extra = case
    for _ in range(max(1, warmup)):
        f()
    ts, t_end = [], time.time() + budget_s
    while len(ts) < steps:
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
        if time.time() > t_end and len(ts) >= 3:
            break
    cb = {"value": _per_s(units, float(np.median(ts)), unit), "unit": unit, "cores": cores, "kind": kind,
          "sample": "%s; %d passes timed one by one (median), -O3 build of the reference's sources" % (what, len(ts))}
    if extra is not None:
        key, g, ecores, ewhat = extra[:4]
        eunits = extra[4] if len(extra) > 4 else units
        g()
        es = []
        for _ in range(min(5, steps)):
            t0 = time.perf_counter()
            g()
            es.append(time.perf_counter() - t0)
        cb[key] = {"value": _per_s(eunits, float(np.median(es)), unit), "unit": unit, "cores": ecores, "what": ewhat}
    return cb, ts


def cpu_baseline_for(name, budget_s=3.0):
    cb, _ = time_cpu_case(cpu_reference_case(name), 1, 64, budget_s)
    return cb


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the workload, same metric/unit/config.
    A step is ONE pass over the bounded sample; W warm-up passes, then K passes each timed on its own."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    d = describe(args.workload, world, args.scaling)
    cb, ts = time_cpu_case(cpu_reference_case(args.workload), max(1, args.warmup), args.steps, 240.0)
    line = {
        "impl": "reference", "metric": d["metric"], "value": cb["value"], "unit": d["unit"], "n_gpus": args.gpus,
        "steps": len(ts), "warmup": max(1, args.warmup), "ms_per_step": float(np.mean(ts)) * 1e3, "higher_is_better": True,
        "scaling": args.scaling if world > 1 else "weak", "vs_baseline": None, "dtype": d["dtype"], "data": "synthetic", "config": d["config"],
        "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "ms_per_step_median": float(np.median(ts)) * 1e3,
        "note": "host CPU only, rank 0 only; value = sample units / median pass time",
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- main (GPU arm)
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="dot", choices=ALL)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"],
                    help="N>1: exchange fused into the kernel over peer memory (default) or a NCCL collective after it")
    ap.add_argument("--bcast", default="chain", choices=["chain", "star", "mcast"],
                    help="N>1 GEMMs, how B reaches the non-root ranks: chain = rank r pulls from rank r-1's copy (every GPU "
                         "sends B once; default), star = everybody pulls from the root, mcast = every receiver pulls 1/(N-1) "
                         "of B from the root and stores it to an NVSwitch multicast address (measured slower: DESIGN.md 6)")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="N>1.  strong (default): the named size is the whole job, each GPU takes 1/N of the elements "
                         "(vectors) or of the rows of A and C (GEMMs); weak: the named size PER GPU")
    ap.add_argument("--also", default="all", help="comma-separated workloads to measure besides the primary one, 'all' or 'none'")
    ap.add_argument("--no-also", action="store_true", help="same as --also none")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs, the probes and the library context")
    ap.add_argument("--quick", action="store_true", help="skip e2e, sustained loops and second data sets (development)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    from hpc_b200 import capi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # keep stdout to the single JSON line: NCCL's version banner (NCCL_DEBUG=VERSION) goes there
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", ""):
            os.environ["NCCL_DEBUG"] = "WARN"
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = capi.Context(local)
    peers, peers_note = None, None
    if world > 1 and args.collective == "fused":
        try:
            from hpc_b200 import multi
            peers = multi.PeerExchange(dist, torch.device("cuda", local))
        except Exception as ex:  # reported in the JSON line, never silent
            peers_note = "peer-memory rendezvous failed (%s): NCCL collective used instead" % str(ex)[:200]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    env = {"torch": torch, "ctx": ctx, "dist": dist, "args": args, "world": world, "rank": rank, "peers": peers,
           "peaks": load_peaks(), "probes": {}, "sampler": sampler}
    if world == 1 and not args.no_cpu:
        env["probes"] = run_probes(env)

    primary = measure(args.workload, env, args.steps, primary=True)

    names = [] if (args.no_also or args.also == "none") else (ALL if args.also == "all" else args.also.split(","))
    also = {}
    for name in names:
        if name == args.workload:
            continue
        try:
            also[name] = measure(name, env, max(20, min(args.steps, 50)))
            if world == 1 and rank == 0 and not args.no_cpu:
                also[name]["cpu_baseline"] = cpu_baseline_for(name, 2.0)
                lc = library_context(name, env)
                if lc is not None:
                    also[name]["library_context"] = lc
        except Exception as ex:  # reported, not hidden -- and it fails the run
            also[name] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300]), "verified": False}
        torch.cuda.empty_cache()
    if world == 1 and rank == 0 and not args.no_cpu and names:
        try:
            also["gemm_1024_host_path"] = config0_leg(env)
        except Exception as ex:
            also["gemm_1024_host_path"] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300]), "verified": False}

    # ---- NVLink ceiling beside the broadcast numbers (N > 1; rank 0 measures the 0 <- 1 pair)
    nvlink_probe = None
    if world > 1 and not args.no_cpu:
        if rank == 0:
            try:
                nvlink_probe = {"pair": "GPU 1 -> GPU 0", "unit": "GB/s", "ceiling_per_direction": NVLINK_GBS_PER_DIRECTION}
                for key, mode, depth in (("bulk_depth2", "bulk", 2), ("bulk_depth8", "bulk", 8), ("kernel_pull", "pull", 0),
                                         ("kernel_push", "push", 0), ("copy_engine", "memcpy", 0)):
                    nvlink_probe[key] = ctx.probe_nvlink(1, mode, depth, 512 << 20, 3)
            except Exception as ex:
                nvlink_probe = {"error": str(ex)[:300]}
        dist.barrier()

    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu:
        cpu = cpu_baseline_for(args.workload, 10.0 if args.workload == "dot" else 4.0)

    all_ok = bool(primary.get("verified")) and all(v.get("verified", False) for v in also.values())
    if rank == 0:
        d = describe(args.workload, world, args.scaling)
        line = {
            "metric": d["metric"], "value": primary["value"], "unit": d["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": primary["warmup"], "ms_per_step": primary["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling if world > 1 else "weak",
            "vs_baseline": None, "dtype": d["dtype"], "data": "synthetic", "config": d["config"],
            "roofline": primary["roofline"], "cpu_baseline": cpu, "e2e": primary.get("e2e"),
            "gpu_launches": primary["gpu_launches"], "clocks": primary.get("clocks"),
            "collective": primary["collective"] if peers_note is None else peers_note,
            "verified": primary["verified"], "verification": primary["verification"], "all_verified": all_ok,
            "variants": primary.get("variants"), "weak": primary.get("weak"),
            "second_data_set": primary.get("second_data_set"), "device_result": primary["device_result"],
            "probes": env["probes"] or None, "nvlink_probe": nvlink_probe,
            "peer_plumbing": getattr(peers, "plumbing", None), "also": also,
        }
        print(json.dumps(line))
        sys.stdout.flush()
    sampler.stop()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if not all_ok:
        sys.exit(3)


if __name__ == "__main__":
    main()

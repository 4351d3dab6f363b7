#!/usr/bin/env python3
"""bench.py -- the measurement contract for the cux operator hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]
    torchrun ... bench.py --gpus N ...           (one rank per GPU, NCCL)

Prints ONE JSON line.  The default workload is BASELINE.json configs[1]: the cux `dot`
operator, fp32, N = 2^28 elements per GPU (2 x 1 GiB read per GPU, far larger than the 126 MB L2).
A step is one pass of the hot path: the dot kernel over this rank's vectors and, at N > 1, one NCCL
all-reduce of the per-GPU partial (north_star: "a per-GPU partial and then a single NCCL allreduce").
  value  = bytes all ranks read / max-over-ranks device time   [GB/s], inputs resident in HBM
  e2e    = the same metric through the host-buffer C-ABI call (cuxb_dot_f32_host): pinned host
           vectors are copied host->device every step and the scalar is read back
  roofline / cpu_baseline / clocks / gpu_launches: see the task contract and DESIGN.md.
Other workloads (--workload sum|nrm2|scan|sgemm|hgemm|linear_s8) time the remaining BASELINE
configs; at N=1 a short pass over each is reported under "also".

--impl reference times the reference's own CPU implementation of the same workload (oracle/_ref
when built, else the oracle port) on the host cores; that is the only product-side use of oracle/.
"""
import argparse
import ctypes
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
SHARE = 1   # --scaling strong: every size above is the WHOLE job and each of the N GPUs takes 1/N of it
FP32_FFMA_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4: 148 SMs x 128 lanes x 2 x clocks.max.sm
INT8_NOMINAL_TOPS = 4500.0


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tensor_tflops": p["bf16_tflops"],
                "tensor_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tensor_tflops": 1590.0, "tensor_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (recipe in B200_PROFILING.md)."""
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

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, power = [], [], set(), []
        for ts, line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[1]))
                    power.append(float(f[3]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            if t0 - 0.05 <= ts <= t1 + 0.15:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- workloads (GPU)
class Workload:
    """name, per-rank algorithmic work per step, the step itself (device-resident inputs)."""

    def __init__(self, torch, ctx, world):
        self.torch, self.ctx, self.world = torch, ctx, world

    def gen_int(self, shape, lo, hi, dtype, seed):
        g = self.torch.Generator(device="cuda")
        g.manual_seed(seed)
        return self.torch.randint(lo, hi + 1, shape, device="cuda", generator=g, dtype=self.torch.int32).to(dtype)


def make_workload(name, torch, ctx, world, rank, dist, peers=None, bcast="chain"):
    from hpc_b200 import capi, multi
    W = Workload(torch, ctx, world)
    W.collective = "none"
    stream = lambda: torch.cuda.current_stream().cuda_stream
    seed = 20240607 + 17 * rank
    if name in ("dot", "sum", "nrm2"):
        n = N_DOT // SHARE
        a = W.gen_int((n,), -2, 2, torch.float32, seed)        # cux fill, demo.cpp:30
        b = W.gen_int((n,), -2, 2, torch.float32, seed + 1) if name == "dot" else None
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
        step = step_nccl
        W.collective = "none (1 GPU)" if world == 1 else "nccl all-reduce of the scalar after the kernel"
        W.step_nccl = step_nccl
        if world > 1 and peers is not None:
            mg = {"dot": lambda: ctx.mg_dot_f32(n, a.data_ptr(), b.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream()),
                  "sum": lambda: ctx.mg_sum_f32(n, a.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream()),
                  "nrm2": lambda: ctx.mg_nrm2_f32(n, a.data_ptr(), res.data_ptr(), world, rank, peers.ptrs, stream())}[name]
            step = mg
            W.collective = "fused: partials exchanged over NVLink peer memory inside the reduction kernel (no NCCL launch)"
        W.kernel = fn
        W.step = step
        W.units = (8 if name == "dot" else 4) * n
        W.unit, W.bound, W.dtype = "GB/s", "hbm", "f32"
        W.desc = "cux %s fp32 N=2^28 per GPU (configs[1]); inputs %d MiB per GPU > L2, no flush needed" % (
            name, W.units >> 20)
        W.metric = "reduce HBM GB/s (fp32 %s, N=2^28 per GPU)" % ("dot-product" if name == "dot" else name)
        W.result = res
        W.launches_per_step = 1

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
        n = N_SCAN // SHARE
        x = W.gen_int((n,), -2, 2, torch.int32, seed)
        out = torch.empty_like(x)
        tot = torch.zeros(1, dtype=torch.int32, device="cuda")
        if world == 1:
            fn = lambda: ctx.scan_i32(n, x.data_ptr(), out.data_ptr(), 0, stream())
            step = fn
            W.launches_per_step = 1
        else:
            gathered = torch.zeros(world, dtype=torch.int32, device="cuda")
            fn = lambda: ctx.scan_i32(n, x.data_ptr(), out.data_ptr(), 0, stream())

            def step():
                # phase 1: local total; exchange; phase 2: scan seeded with the exclusive offset
                ctx.sum_i32(n, x.data_ptr(), tot.data_ptr(), stream())
                seed_val = multi.scan_seed(dist, tot, gathered)
                ctx.scan_i32(n, x.data_ptr(), out.data_ptr(), seed_val, stream())
            W.launches_per_step = 2
            W.collective = "nccl all-gather of per-GPU totals + host-computed seed"
            W.step_nccl = step
            if peers is not None:
                step = lambda: ctx.mg_scan_i32(n, x.data_ptr(), out.data_ptr(), world, rank, peers.ptrs, stream())
                W.collective = "fused: totals exchanged over NVLink peer memory inside the reduction kernel; seed read from device memory (no host sync, no NCCL launch)"
        W.kernel, W.step = fn, step
        W.units = 8 * n
        W.unit, W.bound, W.dtype = "GB/s", "hbm", "i32"
        W.desc = "int32 inclusive scan N=2^30 per GPU (configs[4]); 8 GiB moved per GPU > L2"
        W.metric = "scan HBM GB/s (int32 inclusive prefix scan, N=2^30 per GPU)"
        W.result = out[-1:]

        def make_e2e():
            hx = torch.empty(n, dtype=torch.int32, pin_memory=True).copy_(x)
            hout = torch.empty(n, dtype=torch.int32, pin_memory=True)
            return {"call": lambda: ctx.scan_i32_host(n, hx.data_ptr(), hout.data_ptr(), 0), "h2d": 4 * n, "d2h": 4 * n,
                    "api": "cuxb_scan_i32_host", "result": hout[-1:], "combine": None, "keep": (hx, hout)}
        W.make_e2e = make_e2e if world == 1 else None   # sharded e2e would need the seed exchange on the host
    elif name in ("sgemm", "hgemm", "linear_s8"):
        if name == "sgemm":
            n = SGEMM_N
            m = n // SHARE
            A = W.gen_int((m, n), -2, 5, torch.float32, seed)
            B = W.gen_int((n, n), -2, 5, torch.float32, 20240607 + 99)   # B identical on every rank
            C = torch.zeros((m, n), device="cuda")
            fn = lambda: ctx.sgemm_f32(m, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, C.data_ptr(), n, stream())
            W.dtype, W.peak_kind = "f32", "ffma"
            W.desc = "fp32 SGEMM %d^3 per GPU on CUDA cores (configs[2]); rows of C sharded, B replicated" % n
            W.metric = "GEMM TFLOP/s (fp32 SGEMM 8192^3 per GPU, CUDA cores)"
        elif name == "hgemm":
            n = HGEMM_N
            m = n // SHARE
            A = W.gen_int((m, n), -10, 9, torch.float16, seed)
            B = W.gen_int((n, n), -10, 9, torch.float16, 20240607 + 99)
            C = torch.zeros((m, n), dtype=torch.float16, device="cuda")
            fn = lambda: ctx.hgemm_f16(m, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, C.data_ptr(), n,
                                       capi.FLOAT16, stream())
            W.dtype, W.peak_kind = "f16 (fp32 accumulate)", "tensor"
            W.desc = "fp16 GEMM %d^3 per GPU, fp32 accumulate, tcgen05/TMEM (configs[3]); rows of C sharded, B replicated" % n
            W.metric = "GEMM TFLOP/s (fp16 GEMM 16384^3 per GPU, fp32 accumulate, tcgen05)"
        else:
            n = LIN8_N
            m = n // SHARE
            A = W.gen_int((m, n), 0, 16, torch.int8, seed)
            B = W.gen_int((n, n), -8, 8, torch.int8, 20240607 + 99)
            C = torch.zeros((m, n), dtype=torch.int32, device="cuda")
            fn = lambda: ctx.linear_s8(m, n, n, A.data_ptr(), B.data_ptr(), C.data_ptr(), stream())
            W.dtype, W.peak_kind = "s8 (int32 accumulate)", "tensor_int8"
            W.desc = "int8 linear %d^3 per GPU, int32 accumulate, tcgen05 kind::i8 (configs[4])" % n
            W.metric = "GEMM TOP/s (int8 linear 8192^3 per GPU, int32 accumulate)"
        W.kernel = W.step = fn
        if world > 1 and peers is not None:
            # B lives on rank 0 only; every other rank's GEMM kernel pulls it over NVLink while multiplying
            # chained: rank r pulls from rank r-1's copy (rank 1 from the root), so every GPU sends B once
            if bcast == "chain":
                bsym, b_up = peers.chain(ctx, B)
            else:                                  # star: everybody pulls from the root (comparison)
                bsym = peers.symmetric_like(B)
                b_up = bsym.ptrs[0]
            if rank == 0:
                bsym.tensor.copy_(B)
            torch.cuda.synchronize()
            dist.barrier()
            is_root = rank == 0
            W.step_nccl = fn  # comparison: B already resident on every GPU (no broadcast in the step)
            if name == "hgemm":
                W.step = lambda: ctx.mg_hgemm_f16(m, n, n, 1.0, A.data_ptr(), n, b_up, bsym.ptrs[rank], n, 0.0,
                                                  C.data_ptr(), n, is_root, capi.FLOAT16, stream())
                how = "by the otherwise idle warp of every CTA (TMA), column panel by column panel"
            elif name == "sgemm":
                W.step = lambda: ctx.mg_sgemm_f32(m, n, n, 1.0, A.data_ptr(), n, b_up, bsym.ptrs[rank], n, 0.0,
                                                  C.data_ptr(), n, is_root, stream())
                how = "by the first 48 CTAs of the launch (the slots its last wave of tiles leaves idle), 128 rows of B per readiness flag"
            else:
                W.step = lambda: ctx.mg_linear_s8(m, n, n, A.data_ptr(), b_up, bsym.ptrs[rank], C.data_ptr(),
                                                  is_root, stream())
                how = "by the otherwise idle warp of every CTA (TMA), n-block panel by n-block panel"
            route = ("from the rank before it (chain: each GPU sends B once)" if bcast == "chain"
                     else "straight from rank 0 (star)")
            W.collective = ("fused: B (%d MiB) lives on rank 0; every other rank's GEMM kernel pulls it over NVLink peer "
                            "memory %s, %s, overlapped with the mainloop (no NCCL launch); rows of C sharded"
                            % ((B.numel() * B.element_size()) >> 20, route, how))
            W.keep = bsym

        def make_e2e():
            pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)
            hA, hB = pin(A), pin(B)
            hC = torch.empty(C.shape, dtype=C.dtype, pin_memory=True)
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
        W.make_e2e = make_e2e
        W.units = 2 * m * n * n
        W.unit = "TOP/s" if name == "linear_s8" else "TFLOP/s"
        W.bound = "tensor" if name != "sgemm" else "ffma"
        W.result = C[-1, -1:]
        W.launches_per_step = 1
    else:
        raise SystemExit("unknown workload " + name)
    W.name = name
    return W


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


def peak_for(W, peaks):
    if W.bound == "hbm":
        return peaks["hbm_gbs"], "GB/s", "MEASURED_PEAKS.json hbm_gbs (%s)" % peaks["source"]
    if W.name == "sgemm":
        return FP32_FFMA_PEAK_TFLOPS, "TFLOP/s", "FP32 FFMA 148 SM x 128 x 2 x 1.965 GHz (nominal; no measured FFMA peak)"
    if W.name == "linear_s8":
        return 2 * peaks["tensor_tflops"], "TOP/s", "2 x MEASURED_PEAKS.json bf16_tflops (int8 = 2x bf16 rate; %s)" % peaks["source"]
    return peaks["tensor_tflops"], "TFLOP/s", "MEASURED_PEAKS.json bf16_tflops burst (%s)" % peaks["source"]


# --------------------------------------------------------------------------- CPU reference arm
def _timed(f, budget_s, max_reps=64):
    f()  # warm caches / page in
    times = []
    t_end = time.time() + budget_s
    while True:
        t0 = time.perf_counter()
        f()
        times.append(time.perf_counter() - t0)
        if time.time() > t_end or len(times) >= max_reps:
            break
    return float(np.median(times)), len(times)


def cpu_baseline_for(name, budget_s=3.0):
    """The reference's own CPU implementation of `name` (oracle/_ref, built from /root/reference; the
    oracle port where the reference has no such function) timed on this host on a BOUNDED sample of the
    same workload.  Returns the cpu_baseline dict; 'cores' is the number of threads actually used."""
    import oracle
    R, L = oracle.ref(), oracle.lib()
    rng = np.random.default_rng(20240607)
    ncores = os.cpu_count() or 1
    if name in ("dot", "sum", "nrm2"):
        n = N_DOT
        a = rng.integers(-2, 3, size=n).astype(np.float32)
        out = np.zeros(1, np.float32)
        if name == "dot":
            b = rng.integers(-2, 3, size=n).astype(np.float32)
            f = (lambda: R.ref_cux_dot_v0(n, a, b, out)) if R else (lambda: L.orc_dot_f32(n, a, b, out))
            fn, nbytes = "DotHostV0 (dot_product.cpp:7-13)", 8 * n
        elif name == "sum":
            f = (lambda: R.ref_reduce_host(a, out, n)) if R else (lambda: L.orc_sum_f32(n, a, out))
            fn, nbytes = "ReduceHost (reduce_fp32.cu:26-31)", 4 * n
        else:
            f = (lambda: R.ref_cux_nrm2_v0(n, a, out)) if R else (lambda: L.orc_nrm2_f32(n, a, out))
            fn, nbytes = "Nrm2HostV0 (nrm2.cpp:7-16)", 4 * n
        t, reps = _timed(f, budget_s * 0.6 if name == "dot" else budget_s)
        cb = {"value": nbytes / t / 1e9, "unit": "GB/s", "cores": 1, "kind": "reference" if R else "port",
              "sample": "%s over the full N=2^28 vector(s), %d passes (median); single thread, as in the reference" % (fn, reps)}
        if name == "dot" and R is not None and hasattr(R, "ref_std_dot_mt"):
            # not a cux op, but the repo's own all-cores dot product (one std::thread per slice): reported beside it
            tm, rm = _timed(lambda: R.ref_std_dot_mt(n, a, b, ncores, out), budget_s * 0.4)
            cb["all_cores"] = {"value": nbytes / tm / 1e9, "unit": "GB/s", "cores": ncores,
                               "what": "DotProductMTv1 (std/alg_vector_dot_product.cpp:19-56), %d passes (median)" % rm}
        return cb
    if name == "scan":
        n = 1 << 27
        x = rng.integers(-2, 3, size=n).astype(np.int32)
        out = np.empty_like(x)
        t, reps = _timed(lambda: L.orc_scan_i32(n, x, out, 0), budget_s)
        note = ""
        if R is not None:
            xf, of = x.astype(np.float32), np.empty(n, np.float32)
            t2, _ = _timed(lambda: R.ref_scan_f32(2, xf, n, of), 1.5)
            note = "; the reference's fp32 VectorScanSIMDv2 (vector_scan.cpp:66-79) on the same slice: %.2f GB/s" % (8 * n / t2 / 1e9)
        return {"value": 8 * n / t / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                "sample": "int32 restatement of VectorScanNormal (vector_scan.cpp:21-26) over N=2^27 (1/8 of the workload), "
                          "%d passes (median), single thread%s" % (reps, note)}
    if name == "sgemm":
        n, rows = SGEMM_N, 96
        A = rng.integers(-2, 6, size=(rows, n)).astype(np.float32)
        B = rng.integers(-2, 6, size=(n, n)).astype(np.float32)
        C = np.zeros((rows, n), np.float32)
        flop = 2.0 * rows * n * n
        if R is None:
            t, reps = _timed(lambda: L.orc_gemm_f32_ikj(rows, n, n, 1.0, A, n, B, n, 0.0, C, n, ncores), budget_s, 8)
            return {"value": flop / t / 1e12, "unit": "TFLOP/s", "cores": ncores, "kind": "port",
                    "sample": "GemmHostV1 order, OpenMP over rows, %d-row slice of the 8192^3 problem (rows are independent)" % rows}
        t4, _ = _timed(lambda: R.ref_cux_gemm(4, rows, n, n, 1.0, A, n, B, n, 0.0, C, n), budget_s / 2, 8)
        tm, _ = _timed(lambda: R.ref_omp_gemm(1, rows, n, n, 1.0, A, n, B, n, C, n), budget_s / 2, 8)
        best, who, cores = (t4, "GemmHostV4 (gemm.cpp:161-218, AVX + OpenMP num_threads(6))", min(6, ncores)) if t4 <= tm else \
                           (tm, "MatrixMulMP (openmp/alg_matrix_multiply.cpp:33-51, all cores)", R.ref_num_threads())
        return {"value": flop / best / 1e12, "unit": "TFLOP/s", "cores": cores, "kind": "reference",
                "sample": "%s on a %d-row slice of the 8192^3 problem (rows are independent); GemmHostV4 %.1f GFLOP/s, "
                          "MatrixMulMP %.1f GFLOP/s" % (who, rows, flop / t4 / 1e9, flop / tm / 1e9)}
    if name == "hgemm":
        n, rows = HGEMM_N, 1
        A16 = rng.integers(-10, 10, size=(rows, n)).astype(np.float16).view(np.uint16)
        B16 = rng.integers(-10, 10, size=(n, n)).astype(np.float16).view(np.uint16)
        C16 = np.zeros((rows, n), np.uint16)
        flop = 2.0 * rows * n * n
        if R is not None:
            t0 = time.perf_counter()
            R.ref_hgemm_host(rows, n, n, A16, n, B16, n, C16, n)
            t = time.perf_counter() - t0
            note = ""
            if R.ref_cpu_has_avx512bf16():
                r2 = 64
                Ab = rng.integers(-10, 10, size=(r2, n)).astype(np.float32).view(np.uint32) >> 16
                Ab = Ab.astype(np.uint16)
                Bb = (B16.astype(np.uint16))  # bit patterns only need to be valid bf16 for timing
                Cb = np.zeros((r2, n), np.uint16)
                t2, _ = _timed(lambda: R.ref_linear_bf16(1, r2, n, n, Ab, Bb, Cb), 1.5, 4)
                note = "; fastest 16-bit CPU path in the repo, LinearAvx512Bf16 (linear.cpp:156-182, 1 thread) on a 64-row slice: %.1f GFLOP/s" % (2.0 * r2 * n * n / t2 / 1e9)
            return {"value": flop / t / 1e12, "unit": "TFLOP/s", "cores": 1, "kind": "reference",
                    "sample": "GemmHostV1-half (gemm_fp16_wmma.cu:75-90) on a %d-row slice of the 16384^3 problem, one pass, "
                              "single thread (rows are independent; the full problem would take hours)%s" % (rows, note)}
        t, _ = _timed(lambda: L.orc_hgemm_f16(rows, n, n, 1.0, A16, n, B16, n, 0.0, C16.ctypes.data, n, 0, 1), budget_s, 2)
        return {"value": flop / t / 1e12, "unit": "TFLOP/s", "cores": 1, "kind": "port", "sample": "%d-row slice" % rows}
    if name == "linear_s8":
        n, rows = LIN8_N, 128
        A = rng.integers(0, 17, size=(rows, n)).astype(np.int8)
        B = rng.integers(-8, 9, size=(n, n)).astype(np.int8)
        C = np.zeros((rows, n), np.int32)
        ops = 2.0 * rows * n * n
        if R is not None and R.ref_cpu_has_avx512():
            t, reps = _timed(lambda: R.ref_linear_s8(1, rows, n, n, A, B, C), budget_s, 8)
            who, kind = "LinearAvx512Int8 (linear.cpp:214-237, VNNI dpbusd; valid because A >= 0)", "reference"
        elif R is not None:
            t, reps = _timed(lambda: R.ref_linear_s8(0, rows, n, n, A, B, C), budget_s, 4)
            who, kind = "LinearNormalInt8 (linear.cpp:199-211)", "reference"
        else:
            t, reps = _timed(lambda: L.orc_linear_s8(rows, n, n, A, B, C, 1), budget_s, 4)
            who, kind = "LinearNormalInt8 restatement", "port"
        return {"value": ops / t / 1e12, "unit": "TOP/s", "cores": 1, "kind": kind,
                "sample": "%s on a %d-row slice of the 8192^3 problem, %d passes (median), single thread as in the sample" % (who, rows, reps)}
    raise ValueError(name)


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the workload, same metric/unit/config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    d = make_desc(args.workload)
    # each "step" = one pass over the bounded sample; K steps timed after W warm-up passes
    t_start = time.perf_counter()
    cb = cpu_baseline_for(args.workload, budget_s=max(1.0, min(20.0, 0.25 * args.steps)))
    wall = time.perf_counter() - t_start
    v = cb["value"]
    per_pass_ms = wall * 1e3 / max(1, args.steps)  # wall time of the whole arm spread over the requested steps
    line = {
        "impl": "reference", "metric": d["metric"], "value": v, "unit": d["unit"], "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_pass_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": d["dtype"], "data": "synthetic",
        "config": {"workload": d["desc"] + " -- on the host CPU through the reference's own code", "sample": cb["sample"]},
        "cpu_baseline": cb, "e2e": {"value": v, "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- main (GPU arm)
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="dot", choices=["dot", "sum", "nrm2", "scan", "sgemm", "hgemm", "linear_s8"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"],
                    help="N>1: exchange fused into the kernel over peer memory (default) or a NCCL collective after it")
    ap.add_argument("--bcast", default="chain", choices=["chain", "star"],
                    help="N>1 GEMMs: pull B from the previous rank's copy (default) or every rank from the root")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the named size PER GPU (default, what the driver's 1-8 GPU run uses); strong: the named size is "
                         "the whole job, each GPU takes 1/N of the elements (vectors) or of the rows of A and C (GEMMs)")
    ap.add_argument("--no-also", action="store_true", help="skip the short passes over the other workloads")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
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
    if args.scaling == "strong":
        global SHARE
        SHARE = world
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # keep stdout to the single JSON line: NCCL's version banner (NCCL_DEBUG=VERSION) goes there
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", ""):
            os.environ["NCCL_DEBUG"] = "WARN"
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = capi.Context(local)
    peaks = load_peaks()
    peers, peers_note = None, None
    if world > 1 and args.collective == "fused":
        try:
            from hpc_b200 import multi
            peers = multi.PeerExchange(dist, torch.device("cuda", local))
        except Exception as ex:  # reported in the JSON line, never silent
            peers_note = "peer-memory rendezvous failed (%s): NCCL collective used instead" % str(ex)[:200]
    W = make_workload(args.workload, torch, ctx, world, rank, dist, peers, args.bcast)
    warmup = max(args.warmup, 3)

    # ---- device-resident timing (value)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = capi.launch_count()
    ms_total, t0, t1 = time_steps(torch, W.step, args.steps, warmup, dist)
    launches = (capi.launch_count() - l0) - warmup * W.launches_per_step
    ms_step = ms_total / args.steps
    value = rate(W.units * world, ms_step, W.unit)
    collective_desc = W.collective if peers_note is None else peers_note

    nccl_variant = None
    if world > 1 and peers is not None and hasattr(W, "step_nccl"):
        ms_n, _, _ = time_steps(torch, W.step_nccl, args.steps, 3, dist)
        nccl_variant = {"ms_per_step": ms_n / args.steps, "value": rate(W.units * world, ms_n / args.steps, W.unit),
                        "what": "same kernel followed by a NCCL collective on the scalar(s)" if W.bound == "hbm" else
                        "B already resident on every GPU: no broadcast inside the step"}

    # ---- the dominant kernel alone, on the launching stream (roofline.achieved)
    ms_k, _, t1k = time_steps(torch, W.kernel, args.steps, 3, None)
    ms_kernel = ms_k / args.steps
    # spread of single launches (cux's GpuTimer protocol, timer.h:16-38: one event pair per run), reported
    # next to the average the roofline uses
    n_single = max(5, min(args.steps, 50))
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_single)]
    for a_, b_ in evs:
        a_.record()
        W.kernel()
        b_.record()
    torch.cuda.synchronize()
    singles = sorted(a_.elapsed_time(b_) for a_, b_ in evs)
    clocks = None
    if rank == 0:
        # nvidia-smi cannot sample faster than ~10 Hz: keep the GPU under the same load a little longer
        # if the timed regions were too short to catch a few samples, and say which window was used
        t_load_end = t1k
        while time.time() - t0 < 0.6:
            _, _, t_load_end = time_steps(torch, W.kernel, max(args.steps, 50), 0, None)
        clocks = sampler.stop(t0, t_load_end)
        clocks["window"] = "timed steps + the kernel-only timing loop that follows them (same kernel, same inputs)"
    achieved = rate(W.units, ms_kernel, W.unit)
    peak, punit, pnote = peak_for(W, peaks)
    roofline = {"bound": "hbm" if W.bound == "hbm" else "tensor" if W.bound == "tensor" else "ffma",
                "achieved": achieved, "peak": peak, "unit": punit, "frac": achieved / peak,
                "traffic": None, "kernel_ms": ms_kernel, "kernel_ms_median": singles[len(singles) // 2],
                "kernel_ms_min": singles[0], "kernel_ms_max": singles[-1], "single_launches_timed": n_single,
                "peak_source": pnote,
                "algorithmic_units_per_launch": W.units}
    # the same ceilings measured here, by the library's own probes (N = 1 only: they take a second)
    if world == 1 and not args.no_cpu:
        try:
            if W.bound == "hbm":
                kind = "copy" if W.name == "scan" else "read"
                roofline["measured_here"] = {"value": ctx.probe_hbm(kind, 2 << 30, 5), "unit": "GB/s",
                                             "what": "cuxb_probe_hbm %s stream over 2 GiB, best of 5" % kind}
            elif W.name == "sgemm":
                roofline["measured_here"] = {"value": ctx.probe_fp32(True, 5), "unit": "TFLOP/s",
                                             "what": "cuxb_probe_fp32: packed FFMA2 chains, best of 5"}
            if "measured_here" in roofline:
                roofline["measured_here"]["frac"] = achieved / roofline["measured_here"]["value"]
        except Exception as ex:
            roofline["measured_here"] = {"error": str(ex)[:200]}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        roofline["traffic"] = json.load(open(tpath)).get(W.name)

    # ---- end to end through the host-buffer C ABI (pinned host memory, H2D + kernel + D2H per step)
    e2e = None
    if getattr(W, "make_e2e", None) is not None:
        E = W.make_e2e()
        part = torch.zeros(1, device="cuda")

        def e2e_step():
            E["call"]()                                  # H2D of the inputs, kernel, D2H of the result: blocking
            if world > 1 and E["combine"] == "sum":      # the scalar of every rank, combined like the device path
                part.copy_(E["result"], non_blocking=True)
                dist.all_reduce(part)
                E["result"].copy_(part)
        e2e_steps = max(3, min(args.steps, 10 if W.bound == "hbm" else 5))
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
        e2e = {"value": rate(W.units * world, e2e_ms, W.unit), "unit": W.unit, "h2d_bytes_per_step": int(E["h2d"]),
               "d2h_bytes_per_step": int(E["d2h"]), "ms_per_step": e2e_ms, "steps": e2e_steps,
               "api": "%s (pinned host buffers; host wall clock around the blocking call%s)" % (
                   E["api"], "; + one all-reduce of the scalar" if world > 1 and E["combine"] else ""),
               "result": float(E["result"][0])}
        del E

    # ---- short passes over the other configs (N=1 only)
    also = {}
    if world == 1 and not args.no_also:
        device_result = float(W.result[0])
        W = None
        torch.cuda.empty_cache()
        for name in ["sum", "nrm2", "scan", "sgemm", "hgemm", "linear_s8"]:
            if name == args.workload:
                continue
            try:
                w2 = make_workload(name, torch, ctx, 1, 0, None)
                ms2, _, _ = time_steps(torch, w2.kernel, 10 if w2.bound == "hbm" else 5, 3, None)
                ms2 /= (10 if w2.bound == "hbm" else 5)
                ach = rate(w2.units, ms2, w2.unit)
                pk, pu, pn = peak_for(w2, peaks)
                also[name] = {"value": ach, "unit": w2.unit, "ms": ms2, "peak": pk, "frac": ach / pk,
                              "peak_source": pn, "workload": w2.desc}
                del w2
                torch.cuda.empty_cache()
                if not args.no_cpu:
                    also[name]["cpu_baseline"] = cpu_baseline_for(name, 2.0)
            except Exception as ex:  # reported, not hidden
                also[name] = {"error": str(ex)[:300]}

    # ---- CPU baseline leg (rank 0, N=1): DotHostV0 on the host cores
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu:
        cpu = cpu_baseline_for(args.workload, 10.0 if args.workload == "dot" else 4.0)

    if rank == 0:
        wl = make_desc(args.workload)
        if args.scaling == "strong" and world > 1:   # the named size is the whole job
            for k in ("metric", "desc"):
                wl[k] = wl[k].replace(" per GPU", " in total, 1/%d of it per GPU" % world)
        line = {
            "metric": wl["metric"], "value": value, "unit": wl["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": {"workload": wl["desc"], "l2": "inputs larger than L2 (no flush needed)" if wl["hbm"] else
                       "GEMM operands exceed L2; no flush",
                       "parallelism": "1 GPU" if world == 1 else
                       ("one rank per GPU, contiguous slice each; " if wl["hbm"] else "one rank per GPU, rows of A and C sharded; ")
                       + collective_desc},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "collective": collective_desc, "nccl_variant": nccl_variant, "also": also,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def make_desc(name):
    d = {
        "dot": ("reduce HBM GB/s (fp32 dot-product, N=2^28 per GPU)", "GB/s", "f32",
                "cux dot fp32 N=2^28 per GPU (configs[1]), integer-valued fill [-2,2] as cux's demo", True),
        "sum": ("reduce HBM GB/s (fp32 sum, N=2^28 per GPU)", "GB/s", "f32", "fp32 sum-reduce N=2^28 per GPU (configs[1])", True),
        "nrm2": ("reduce HBM GB/s (fp32 nrm2, N=2^28 per GPU)", "GB/s", "f32", "cux nrm2 fp32 N=2^28 per GPU", True),
        "scan": ("scan HBM GB/s (int32 inclusive scan, N=2^30 per GPU)", "GB/s", "i32", "int32 prefix scan N=2^30 per GPU (configs[4])", True),
        "sgemm": ("GEMM TFLOP/s (fp32 SGEMM 8192^3 per GPU, CUDA cores)", "TFLOP/s", "f32", "fp32 SGEMM 8192^3 per GPU (configs[2])", False),
        "hgemm": ("GEMM TFLOP/s (fp16 GEMM 16384^3 per GPU, fp32 accumulate, tcgen05)", "TFLOP/s", "f16",
                  "fp16 GEMM 16384^3 per GPU, fp32 accumulate (configs[3])", False),
        "linear_s8": ("GEMM TOP/s (int8 linear 8192^3 per GPU)", "TOP/s", "s8", "int8 linear 8192^3 per GPU (configs[4])", False),
    }[name]
    return {"metric": d[0], "unit": d[1], "dtype": d[2], "desc": d[3], "hbm": d[4]}


if __name__ == "__main__":
    main()

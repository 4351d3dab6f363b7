#!/usr/bin/env python3
"""Summarise the ncu --set full reports of tools/ncu_capture.sh into profiles/ (text, tracked) and refresh
profiles/traffic.json (DRAM bytes per launch + the sha1 of each kernel's source at capture time, which bench.py
uses to flag a stale figure).      python tools/ncu_summary.py gpurun_out/ncu_r02"""
import csv, hashlib, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "lts__t_sector_hit_rate.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
        "lts__t_bytes.sum", "sm__cycles_active.avg", "smsp__cycles_active.avg"]
SRC = {"dot": "reduce.cu", "sum": "reduce.cu", "nrm2": "reduce.cu", "scan": "scan.cu", "sgemm": "sgemm.cu", "hgemm": "umma_gemm.cu",
       "linear_s8": "umma_gemm.cu"}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main(d):
    traffic = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) from the ncu --set full captures "
                           "summarised in profiles/r02_ncu_full_*.txt", "capture": "ncu --set full --clock-control none, round 2",
               "source_sha1": {}}
    for w in ("dot", "sum", "scan", "sgemm", "hgemm", "linear_s8"):
        rep = os.path.join(d, "r02_%s.ncu-rep" % w)
        if not os.path.exists(rep):
            print("missing", rep)
            continue
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        head, units, vals = rows[0], rows[1], rows[2]
        col = {h: i for i, h in enumerate(head)}
        out = ["# source: %s (tools/ncu_capture.sh; summarised by tools/ncu_summary.py)" % rep, "",
               "kernel: %s  grid %s block %s" % (vals[col["Kernel Name"]], vals[col.get("Grid Size", 0)], vals[col.get("Block Size", 0)])]
        tot = 0.0
        for m in KEEP:
            if m in col:
                out.append("  %-75s %s %s" % (m, vals[col[m]], units[col[m]]))
                if m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(vals[col[m]].replace(",", "")) * UNIT.get(units[col[m]], 1.0)
        open(os.path.join(ROOT, "profiles", "r02_ncu_full_%s.txt" % w), "w").write("\n".join(out) + "\n")
        traffic[w] = int(tot)
        src = SRC[w]
        traffic["source_sha1"][src] = hashlib.sha1(open(os.path.join(ROOT, "hpc_b200", "csrc", src), "rb").read()).hexdigest()[:12]
        print(w, "DRAM bytes per launch", int(tot))
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    ll = os.path.join(d, "r02_launches_dot.csv")
    if os.path.exists(ll):
        open(os.path.join(ROOT, "profiles", "r02_launches_dot.csv"), "w").write(open(ll).read())


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "ncu_r02"))

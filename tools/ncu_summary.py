"""Summarise gpurun_out/ ncu artefacts into profiles/ (tracked): launch-share table, per-kernel metric tables, and
profiles/ncu_traffic.json (average DRAM bytes per launch per C entry point, read by bench.py's roofline.traffic).

    python tools/ncu_summary.py <tag>        # e.g. r01_mxfp8
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
        "sm__cycles_elapsed.avg.per_second", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]

# first match wins (so "dswiglu" is tested before "swiglu")
ENTRY = {"fp8_gemm_kernel": None, "mx_dswiglu_quantize_kernel": "llmfp8_mx_dswiglu_quantize", "mx_swiglu_quantize_kernel": "llmfp8_mx_swiglu_quantize",
         "mx_quantize_kernel": "llmfp8_mx_quantize", "dswiglu_quant_kernel": "llmfp8_dswiglu_quant_bwd",
         "swiglu_quant_kernel": "llmfp8_swiglu_quant_fwd", "rmsnorm_fwd_kernel": "llmfp8_rmsnorm_quant_fwd", "rmsnorm_bwd_kernel": "llmfp8_rmsnorm_bwd",
         "cast_amax_bgrad_kernel": "llmfp8_quant_cast_amax_bgrad", "cast_amax_kernel": "llmfp8_quant_cast_amax"}


def to_bytes(v, unit):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(v.replace(",", "")) * m.get(unit, 1)


def launches(tag):
    p = os.path.join(OUT, "launches.csv")
    if not os.path.exists(p):
        return
    lines = [l for l in open(p) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        n = re.sub(r"<.*", "", row["Kernel Name"])[:80]
        agg[n][0] += 1
        agg[n][1] += v
    tot = sum(v for _, v in agg.values())
    with open(os.path.join(PROF, f"{tag}_launch_shares.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  (cold-cache, serialised: compare SHARES)\n")
        f.write(f"# total {tot/1e3:.2f} ms over {sum(c for c, _ in agg.values())} launches\n")
        f.write("# share   launches   avg_us   kernel\n")
        for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{v/tot*100:6.2f}%  {c:5d}  {v/c:10.1f}  {n}\n")
    os.replace(p, os.path.join(PROF, f"{tag}_launches.csv"))


def report(name, tag, traffic):
    rep = os.path.join(OUT, f"{name}.ncu-rep")
    csv_path = os.path.join(OUT, f"{name}_raw.csv")       # exported on the GPU box (the reports themselves exceed gpurun's 64 MiB return limit)
    if os.path.exists(csv_path):
        raw = open(csv_path).read()
    elif os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    else:
        return
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    per = collections.defaultdict(list)
    with open(os.path.join(PROF, f"{tag}_{name}.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on ; source report: gpurun_out/{name}.ncu-rep (scratch)\n")
        for r in rows[2:]:
            kn = r[idx["Kernel Name"]]
            f.write(f"--- {kn[:140]}\n")
            for k in KEYS:
                if k in idx:
                    f.write(f"    {k} = {r[idx[k]]} {units[idx[k]]}\n")
            b = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) + to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
            for short, entry in ENTRY.items():
                if short in kn:
                    if entry is None:
                        # fp8_gemm_kernel<CG, A_MN, B_MN, MX, RES>: the 4th template argument says block-scaled
                        entry = "llmfp8_fp8_gemm"
                        m = re.search(r"fp8_gemm_kernel<(.*?)>\(", kn)
                        if m:
                            targs = [re.sub(r"\(.*?\)", "", t).strip() for t in m.group(1).split(",")]
                            if len(targs) >= 4 and targs[3] in ("1", "true"):
                                entry = "llmfp8_mxfp8_gemm"
                    per[entry].append(b)
                    break
    for e, v in per.items():
        traffic[e] = {"dram_bytes_per_launch_avg": sum(v) / len(v), "launches_captured": len(v), "source": f"profiles/{tag}_{name}.txt"}


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    os.makedirs(PROF, exist_ok=True)
    launches(tag)
    tp = os.path.join(PROF, "ncu_traffic.json")
    traffic = json.load(open(tp)) if os.path.exists(tp) else {}
    for name in ("prof_gemm", "prof_quant", "prof_gemm_pt", "prof_quant_pt"):
        report(name, tag, traffic)
    json.dump(traffic, open(tp, "w"), indent=1)
    print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()

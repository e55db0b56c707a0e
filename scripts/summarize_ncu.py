#!/usr/bin/env python
"""Turn the two ncu outputs of scripts/gpu_validate_and_profile.sh into the summaries kept under profiles/.

    python scripts/summarize_ncu.py --tag r01e [--launches gpurun_out/launches_latest.csv]
                                    [--rep gpurun_out/lensll_latest.ncu-rep] [--rows 37888]
writes  profiles/<tag>_launches_bench_296prompts.csv   (the launch list as ncu wrote it)
        profiles/<tag>_launch_summary.txt              (time and share per kernel)
        profiles/<tag>_lensll_ncu_full_selected.csv    (the metrics DESIGN.md quotes, per captured launch)
        profiles/lens_traffic.json                     (DRAM bytes per row, read by bench.py)
Runs where ncu is installed (this container: reading a report needs no GPU).
"""
import argparse
import collections
import csv
import io
import json
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEEP = [
    "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__cycles_active.avg", "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma_realtime.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__cluster_dim_x", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__cycles_active.avg", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "gpc__cycles_elapsed.max",
    "sm__cycles_elapsed.avg.per_second", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
]


def launch_summary(path, tag):
    text = open(path).read()
    start = text.find('"ID"')
    rows = list(csv.DictReader(io.StringIO(text[start:])))
    per = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "s": 1e3,
                  "second": 1e3}.get(unit, 1e-6)
        k = per.setdefault(r["Kernel Name"], [0.0, 0])
        k[0] += ms
        k[1] += 1
    total = sum(v[0] for v in per.values())
    lines = []
    for name, (ms, n) in sorted(per.items(), key=lambda kv: -kv[1][0]):
        lines.append(f"{ms:10.2f} ms {n:5d} launches {100 * ms / total:5.1f}%  {name[:150]}")
    lines.append(f"{total:10.2f} ms total device time over {sum(v[1] for v in per.values())} launches")
    shutil.copy(path, os.path.join(ROOT, "profiles", f"{tag}_launches_bench_296prompts.csv"))
    out = os.path.join(ROOT, "profiles", f"{tag}_launch_summary.txt")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:8]))


def full_selected(rep, tag, captured_rows, d, V):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.find('"ID"'):])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = os.path.join(ROOT, "profiles", f"{tag}_lensll_ncu_full_selected.csv")
    picked = {}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
        for i, name in enumerate(hdr):
            short = name.split(".", 2)[-1] if name.count(".") >= 2 and name.split(".")[0].isupper() else name
            if any(name == k or name.endswith("." + k) or short == k for k in KEEP):
                w.writerow([name, units[i]] + [r[i] for r in data])
                picked[name] = [r[i] for r in data]

    def metric(suffix):
        for k, v in picked.items():
            if k.endswith(suffix):
                u = units[hdr.index(k)]
                scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}.get(u, 1.0)
                return [float(x.replace(",", "")) * scale for x in v]
        return None

    rd, wr = metric("dram__bytes_read.sum"), metric("dram__bytes_write.sum")
    if rd and wr:
        per_launch = [a + b for a, b in zip(rd, wr)]
        kname = picked.get("Kernel Name", ["?"])[0]
        j = {
            "kernel": kname.split("(")[0].replace("void ", "").replace("tb::", ""),
            "captured_rows": captured_rows,
            "dram_bytes_per_launch": per_launch,
            "dram_bytes_per_row": sum(per_launch) / len(per_launch) / captured_rows,
            "algorithmic_dram_bytes_per_launch": float(2 * d * (captured_rows + V)),
            "source": f"profiles/{tag}_lensll_ncu_full_selected.csv (ncu --set full, bench.py --prompts 296)",
            "note": "traffic scales with the row count at a fixed chunking (W_U is re-streamed once per group of "
                    "resident row blocks); bench.py scales dram_bytes_per_row to its launch size",
        }
        json.dump(j, open(os.path.join(ROOT, "profiles", "lens_traffic.json"), "w"), indent=1)
        print(json.dumps(j)[:400])
    for k, v in picked.items():
        print(k, v)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", required=True)
    ap.add_argument("--launches", default=os.path.join(ROOT, "gpurun_out", "launches_latest.csv"))
    ap.add_argument("--rep", default=os.path.join(ROOT, "gpurun_out", "lensll_latest.ncu-rep"))
    ap.add_argument("--rows", type=int, default=296 * 128)
    ap.add_argument("--d", type=int, default=3584)
    ap.add_argument("--vocab", type=int, default=256000)
    a = ap.parse_args()
    if os.path.exists(a.launches):
        launch_summary(a.launches, a.tag)
    if os.path.exists(a.rep):
        full_selected(a.rep, a.tag, a.rows, a.d, a.vocab)


if __name__ == "__main__":
    main()

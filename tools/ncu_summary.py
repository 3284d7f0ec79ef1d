"""Condenses `ncu --page raw --csv` exports (tools/ncu_capture.sh) into one markdown table per workload:
per kernel (aggregated over its launches in the step) duration, DRAM bytes, achieved DRAM bandwidth against the measured
copy bandwidth, shared-memory (LSU) wavefront utilisation, issue-slot and tensor-pipe utilisation.

    python tools/ncu_summary.py gpurun_out/mlp_full_raw.csv [...]  > profiles/r2_ncu_kernel_table.md
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HBM = 6552.0
try:
    HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def short(name):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"\(.*", "", name)
    return name.replace("void ", "")


def to_bytes(v, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)
    return float(v) * mult


def to_ms(v, unit):
    return float(v) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1.0)


def table(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    need = {"t": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
            "lsu": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
            "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "tensor": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "l2": "lts__t_sector_hit_rate.pct", "regs": "launch__registers_per_thread", "warps": "sm__warps_active.avg.pct_of_peak_sustained_active"}
    agg = {}
    for r in data:
        if len(r) < len(hdr):
            continue
        k = short(r[col["Kernel Name"]])
        a = agg.setdefault(k, {"n": 0, "ms": 0.0, "bytes": 0.0, "lsu": 0.0, "issue": 0.0, "tensor": 0.0, "l2": 0.0, "regs": r[col[need["regs"]]], "warps": 0.0})
        ms = to_ms(r[col[need["t"]]], units[col[need["t"]]])
        a["n"] += 1
        a["ms"] += ms
        a["bytes"] += to_bytes(r[col[need["rd"]]], units[col[need["rd"]]]) + to_bytes(r[col[need["wr"]]], units[col[need["wr"]]])
        for key in ("lsu", "issue", "tensor", "l2", "warps"):
            try:
                a[key] += float(r[col[need[key]]]) * ms       # time-weighted
            except ValueError:
                pass
    total = sum(a["ms"] for a in agg.values())
    out = [f"### {os.path.basename(path)} ({len(data)} launches, {total:.2f} ms under ncu: serialised, cold caches)", "",
           "| kernel | launches | ms | share | DRAM GB | DRAM GB/s | of HBM peak | LSU shared wavefronts % | issue active % | tensor active % | L2 hit % | regs | warps active % |",
           "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        w = max(a["ms"], 1e-12)
        gbs = a["bytes"] / (a["ms"] * 1e-3) / 1e9 if a["ms"] else 0.0
        out.append(f"| {k} | {a['n']} | {a['ms']:.3f} | {100 * a['ms'] / total:.1f}% | {a['bytes'] / 1e9:.3f} | {gbs:.0f} | {gbs / HBM:.2f} | "
                   f"{a['lsu'] / w:.1f} | {a['issue'] / w:.1f} | {a['tensor'] / w:.1f} | {a['l2'] / w:.1f} | {a['regs']} | {a['warps'] / w:.1f} |")
    return "\n".join(out)


if __name__ == "__main__":
    print(f"Per-kernel Nsight Compute summary (`ncu --set full --clock-control none`, one bench step per workload via "
          f"tools/profile_step.py; HBM peak = {HBM:.0f} GB/s measured copy bandwidth).  Durations under ncu are serialised and "
          f"cold-cache: shares, not absolutes, compare with bench.py's trace.\n")
    for p in sys.argv[1:]:
        print(table(p))
        print()

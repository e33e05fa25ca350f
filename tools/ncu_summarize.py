#!/usr/bin/env python
"""Turn ncu outputs into the small JSON summaries committed under profiles/.

  launches <csv> <out.json>            per-kernel launch count / total / average from a
                                       `--metrics gpu__time_duration.sum --csv` launch list
  full <report.ncu-rep> <out.json>     selected raw metrics of every captured launch of a
                                       `--set full` report (reads it with `ncu -i ... --page raw --csv`)
"""
import csv
import io
import json
import re
import subprocess
import sys
from collections import OrderedDict

OURS = ("attn_", "kv_write", "kv_indices", "gemm_tn", "fp8_", "w4a16", "awq_", "b200", "quant",
        "rmsnorm", "rope", "silu", "allreduce", "gptq")

FULL_KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__warps_eligible.avg.per_cycle_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_dense_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
]


def short_name(k: str) -> str:
    k = re.sub(r"^void\s+", "", k)
    k = re.sub(r"\(anonymous namespace\)::", "", k)
    m = re.match(r"(?:b200::)?([A-Za-z0-9_:]+)", k)
    name = m.group(1) if m else k
    if name.startswith("at::") or name.startswith("c10::"):
        name = "torch:" + name.split("<")[0]
    return name


def launches(path, out):
    text = open(path).read()
    start = text.find('"ID"')
    rows = list(csv.DictReader(io.StringIO(text[start:])))
    agg = OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
        name = short_name(r["Kernel Name"])
        a = agg.setdefault(name, [0, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1
        a[1] += us
    ours_total = sum(a[1] for n, a in agg.items() if any(n.startswith(p) or p in n for p in OURS))
    res = []
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        ours = any(n.startswith(p) or p in n for p in OURS)
        res.append({"kernel": n, "launches": a[0], "total_us": round(a[1], 1),
                    "avg_us": round(a[1] / a[0], 2), "grid": a[2], "block": a[3],
                    "share_of_b200_kernels": round(a[1] / ours_total, 4) if ours and ours_total else None})
    json.dump(res, open(out, "w"), indent=1)
    for r in res[:12]:
        print(r)


def full(rep, out):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt[txt.find('"ID"'):])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    res = []
    for d in data:
        rec = {}
        row = dict(zip(hdr, d))
        u = dict(zip(hdr, units))
        for k in FULL_KEYS:
            if k in row and row[k] != "":
                rec[k] = f"{row[k]} {u[k]}".strip()
        for k in ("Grid Size", "Block Size", "Kernel Name"):
            rec[k] = row.get(k)
        res.append(rec)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res[0], indent=1))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])

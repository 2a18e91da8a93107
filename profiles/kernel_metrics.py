#!/usr/bin/env python
"""Key metrics of every kernel instance in an .ncu-rep (from `ncu --set full`), as text + JSON.
usage: kernel_metrics.py report.ncu-rep [out.json]"""
import csv
import io
import json
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__waves_per_multiprocessor", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "lts__t_requests_srcunit_tex_op_red.sum",
    "smsp__inst_executed_op_global_red.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main(rep, out_json=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    res = []
    for r in rows[2:]:
        d = {"kernel": r[idx["Kernel Name"]].split("(")[0]}
        for w in WANT:
            if w in idx and r[idx[w]] not in ("", "n/a"):
                u = units[idx[w]]
                d[w] = to_bytes(r[idx[w]], u) if "byte" in u else float(r[idx[w]].replace(",", ""))
                if "byte" not in u and u:
                    d[w + "#unit"] = u
        st = [(float(r[i].replace(",", "")), h.replace("smsp__pcsamp_warps_issue_stalled_", ""))
              for h, i in idx.items() if h.startswith("smsp__pcsamp_warps_issue_stalled") and r[i] not in ("", "n/a")]
        tot = sum(v for v, _ in st) or 1
        d["stall_pct"] = {h: round(100 * v / tot, 1) for v, h in sorted(st, reverse=True)[:8]}
        d["dram_traffic_bytes"] = d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
        res.append(d)
    for d in res:
        print(f"== {d['kernel']}")
        for k, v in d.items():
            if k != "kernel" and not k.endswith("#unit"):
                print(f"   {k:62s} {v}")
    if out_json:
        json.dump(res, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)

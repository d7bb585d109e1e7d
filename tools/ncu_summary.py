#!/usr/bin/env python
"""Summarise an ncu report: tools/ncu_summary.py gpurun_out/x.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "inst"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%act"), ("sm__cycles_active.avg", "sm_act"),
        ("sm__cycles_elapsed.max", "sm_el"), ("lts__t_sector_hit_rate.pct", "l2hit"), ("l1tex__t_sector_hit_rate.pct", "l1hit"),
        ("lts__t_bytes.sum", "l2bytes"), ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wf")]
for r in data:
    name = r[idx["Kernel Name"]].split("(")[0][-28:]
    out = [name]
    for k, short in want:
        if k in idx:
            v = r[idx[k]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.3g}"
            except Exception:
                pass
            out.append(f"{short}={v}{units[idx[k]] if short in ('time','rd','wr','l2bytes') else ''}")
    print(" ".join(out))

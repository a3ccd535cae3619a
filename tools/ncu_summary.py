#!/usr/bin/env python
"""Summarise an `ncu --set full` report (+ optionally a gpu__time_duration launch list) into a small JSON/markdown pair that
is committed under profiles/:  ncu_summary.py <report.ncu-rep> <out.json> [launches.csv]
Per kernel: duration, DRAM bytes read+written (the `traffic` of bench.py's roofline), issue-slot and pipe utilisation,
occupancy and its limiter, L1/L2 hit rates, sectors per global load request (32 = uncoalesced, 4 = one 128-byte line)."""
import csv, io, json, subprocess, sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units = rows[0], rows[1]
M = {"ms": "gpu__time_duration.sum", "dram_read": "dram__bytes_read.sum", "dram_write": "dram__bytes_write.sum",
     "dram_pct_of_peak": "dram__throughput.avg.pct_of_peak_sustained_elapsed",
     "sm_throughput_pct": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
     "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "alu_pipe_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "fma_pipe_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
     "lsu_pipe_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
     "warp_instructions": "smsp__inst_executed.sum", "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "registers": "launch__registers_per_thread", "l1_hit_pct": "l1tex__t_sector_hit_rate.pct", "l2_hit_pct": "lts__t_sector_hit_rate.pct",
     "ld_sectors": "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "ld_requests": "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
     "stall_long_scoreboard": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
     "stall_wait": "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
     "stall_math_throttle": "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}
ks = []
for r in rows[2:]:
    d = dict(zip(h, r))
    k = {"kernel": d["Kernel Name"].split("(")[0].replace("void ", ""), "grid": d["Grid Size"], "block": d["Block Size"]}
    for name, m in M.items():
        if m in d and d[m] not in ("", "n/a"):
            v = float(d[m].replace(",", ""))
            u = units[h.index(m)]
            if name in ("ms", "dram_read", "dram_write"):
                v *= SCALE.get(u, 1.0)
            k[name] = v
    if "dram_read" in k:
        k["dram_bytes_per_launch"] = k["dram_read"] + k.get("dram_write", 0.0)
        k["dram_GBps"] = k["dram_bytes_per_launch"] / (k["ms"] * 1e-3) / 1e9
    if k.get("ld_requests"):
        k["sectors_per_ld_request"] = k["ld_sectors"] / k["ld_requests"]
    ks.append(k)
res = {"report": rep.split("/")[-1], "kernels": ks}
if len(sys.argv) > 3:
    tot, per = 0.0, {}
    rr = [r for r in csv.reader(open(sys.argv[3])) if len(r) > 10]
    hh = rr[0]
    for r in rr[1:]:
        if r[hh.index("Metric Name")] != "gpu__time_duration.sum":
            continue
        v = float(r[hh.index("Metric Value")].replace(",", "")) * SCALE.get(r[hh.index("Metric Unit")], 1.0)
        n = r[hh.index("Kernel Name")].split("(")[0].replace("void ", "")
        per.setdefault(n, [0, 0.0]); per[n][0] += 1; per[n][1] += v; tot += v
    res["launch_list"] = {"file": sys.argv[3].split("/")[-1], "total_ms": tot,
                          "kernels": [{"kernel": n, "launches": c, "ms": ms, "share": ms / tot} for n, (c, ms) in sorted(per.items(), key=lambda kv: -kv[1][1])]}
json.dump(res, open(out, "w"), indent=1)
for k in ks:
    print("%-22s %8.3f ms  dram %7.2f GB (%6.0f GB/s)  issue %5.1f%%  alu %5.1f%%  lsu %5.1f%%  warps %5.1f%%  regs %3d  L1 %5.1f%%  L2 %5.1f%%  sect/req %5.1f" % (
        k["kernel"][:22], k["ms"], k.get("dram_bytes_per_launch", 0) / 1e9, k.get("dram_GBps", 0), k.get("issue_active_pct", 0), k.get("alu_pipe_pct", 0),
        k.get("lsu_pipe_pct", 0), k.get("warps_active_pct", 0), int(k.get("registers", 0)), k.get("l1_hit_pct", 0), k.get("l2_hit_pct", 0), k.get("sectors_per_ld_request", 0)))
if "launch_list" in res:
    for k in res["launch_list"]["kernels"][:12]:
        print("   %-24s x%-3d %9.3f ms  %5.1f%%" % (k["kernel"][:24], k["launches"], k["ms"], 100 * k["share"]))

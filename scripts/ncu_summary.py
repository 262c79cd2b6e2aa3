#!/usr/bin/env python
"""Summarise one kernel of an ncu report (--set full) into a small JSON + markdown for profiles/.

    python scripts/ncu_summary.py REPORT.ncu-rep KERNEL_SUBSTRING OUT_PREFIX [LAUNCH_INDEX]

Reads the report with `ncu -i ... --page raw --csv`; needs no GPU.
"""
import csv
import json
import subprocess
import sys

WANT = {
    "duration_us": "gpu__time_duration.sum",
    "dram_bytes_read": "dram__bytes_read.sum",
    "dram_bytes_write": "dram__bytes_write.sum",
    "sm_cycles_elapsed_max": "sm__cycles_elapsed.max",
    "inst_executed": "smsp__inst_executed.sum",
    "ipc_active": "sm__inst_executed.avg.per_cycle_active",
    "pipe_fp64_pct_of_peak_active": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "pipe_fp64_cycles_active_pct": "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "pipe_alu_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "pipe_fma_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "pipe_lsu_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "pipe_xu_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "achieved_occupancy_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "theoretical_occupancy_pct": "sm__maximum_warps_per_active_cycle_pct",
    "registers_per_thread": "launch__registers_per_thread",
    "dynamic_smem_per_block": "launch__shared_mem_per_block_dynamic",
    "static_smem_per_block": "launch__shared_mem_per_block_static",
    "grid_size": "launch__grid_size",
    "block_size": "launch__block_size",
    "waves_per_sm": "launch__waves_per_multiprocessor",
    "l2_hit_rate_pct": "lts__t_sector_hit_rate.pct",
    "l1_hit_rate_pct": "l1tex__t_sector_hit_rate.pct",
    "dram_throughput_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm_throughput_pct": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smem_wavefronts": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "local_load_sectors": "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum",
    "local_store_sectors": "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def main():
    rep, kern, out = sys.argv[1], sys.argv[2], sys.argv[3]
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    head, units = rows[0], rows[1]
    ik = head.index("Kernel Name")
    sel = [r for r in rows[2:] if kern in r[ik]]
    if not sel:
        raise SystemExit("no launch of a kernel matching %r in %s" % (kern, rep))
    r = sel[min(which, len(sel) - 1)]
    res = {"report": rep.split("/")[-1], "kernel": r[ik], "launches_in_report": len(sel)}
    for key, name in WANT.items():
        if name in head:
            i = head.index(name)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            u = units[i]
            u0 = u.split("/")[0]
            if u0 in UNIT_SCALE and ("bytes" in key or "smem" in key or key == "duration_us"):
                v *= UNIT_SCALE[u0]
            res[key] = v
    # warp stall breakdown (pc sampling, all samples)
    stalls = {}
    for i, n in enumerate(head):
        if n.startswith("smsp__pcsamp_warps_issue_stalled_") and not n.endswith("_not_issued"):
            try:
                v = float(r[i])
            except ValueError:
                continue
            if v > 0:
                stalls[n[len("smsp__pcsamp_warps_issue_stalled_"):]] = v
    tot = sum(stalls.values()) or 1.0
    res["stall_share"] = {k: round(v / tot, 4) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]}
    if "dram_bytes_read" in res:
        res["dram_bytes_per_launch"] = res["dram_bytes_read"] + res.get("dram_bytes_write", 0.0)
    json.dump(res, open(out + ".json", "w"), indent=1)
    with open(out + ".md", "w") as f:
        f.write("# ncu --set full summary: %s\n\nreport `%s`, launch %d of %d in the capture (cold-cache, serialised: use shares, not absolutes)\n\n" %
                (res["kernel"].split("(")[0], res["report"], which, len(sel)))
        f.write("| metric | value |\n|---|---|\n")
        for k, v in res.items():
            if k in ("report", "kernel", "stall_share"):
                continue
            f.write("| %s | %s |\n" % (k, ("%.4g" % v) if isinstance(v, float) else v))
        f.write("\nwarp-stall shares (pc sampling): " + ", ".join("%s %.1f%%" % (k, 100 * v) for k, v in res["stall_share"].items()) + "\n")
    print(json.dumps(res))


if __name__ == "__main__":
    main()

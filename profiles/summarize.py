#!/usr/bin/env python3
"""Turns gpurun_out/ ncu artefacts into the small text summaries committed under profiles/.
  python profiles/summarize.py launches gpurun_out/launches.csv profiles/r01_launches.txt
  python profiles/summarize.py full gpurun_out/prof_trace.ncu-rep profiles/r01_k_trace_full.txt
  python profiles/summarize.py json gpurun_out/x.ncu-rep profiles/r02x_cfg3_ncu.json <spp> <segments> <config> <kernel variant>
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread", "launch__waves_per_multiprocessor",
    "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__sass_average_branch_targets_threads_uniform.pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_bytes.sum",
    "lts__t_bytes.sum.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sectors.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
]


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, mi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[mi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        name = re.sub(r"\(.*", "", r[ki])
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += v
    s = sum(v[1] for v in tot.values())
    with open(dst, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none ; source {src}\n")
        f.write(f"# total {s:.1f} us over {sum(v[0] for v in tot.values())} launches (cold-cache, serialised: compare shares)\n")
        for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{v[1]:12.1f} us {v[0]:5d} x {100 * v[1] / s:6.2f}%  {k}\n")
    print(open(dst).read())


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full --clock-control none ; source {src}\n")
        for r in rows[2:]:
            f.write(f"## kernel: {r[hdr.index('Kernel Name')][:120]}  grid {r[hdr.index('Grid Size')]} block {r[hdr.index('Block Size')]}\n")
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write(f"{k:75s} {r[i]:>20s} {units[i]}\n")
    print(open(dst).read())


def as_json(src, dst, spp="0", segments="0", config="3", kernel_variant="pixel"):
    """Machine-readable digest of one `ncu --set full` capture of k_trace; bench.py reads the
    newest profiles/*_k_trace_ncu.json for roofline.traffic (DRAM bytes of that launch)."""
    import json

    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, r = rows[0], rows[1], rows[2]

    def val(k):
        if k not in hdr:
            return None
        i = hdr.index(k)
        v = float(r[i].replace(",", ""))
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[i], 1.0)
        return v * scale

    d = {
        "source": src, "config": int(config), "kernel_variant": kernel_variant, "kernel": r[hdr.index("Kernel Name")], "spp_of_launch": int(spp), "segments_of_launch": int(segments),
        "duration_ms_under_ncu": val("gpu__time_duration.sum"),
        "dram_bytes": (val("dram__bytes_read.sum") or 0) + (val("dram__bytes_write.sum") or 0),
        "l2_bytes": (val("lts__t_sectors.sum") or 0) * 32.0,
        "l1_global_load_bytes": (val("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum") or 0) * 32.0,
        "l1_lsu_wavefronts": val("l1tex__data_pipe_lsu_wavefronts.sum") or (val("SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg") or 0) * 148.0,
        "l1_lsu_wavefronts_pct_of_peak": val("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
        "l2_throughput_pct": val("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        "dram_throughput_pct": val("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "fma_pipe_pct": val("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        "alu_pipe_pct": val("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "threads_per_instruction": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "registers_per_thread": val("launch__registers_per_thread"),
        "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
    }
    json.dump(d, open(dst, "w"), indent=1)
    print(json.dumps(d, indent=1))


if __name__ == "__main__":
    {"launches": launches, "full": full, "json": as_json}[sys.argv[1]](*sys.argv[2:])

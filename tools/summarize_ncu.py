#!/usr/bin/env python
"""Turn gpurun_out ncu artefacts into the small text summaries committed under profiles/.
  python tools/summarize_ncu.py launches gpurun_out/launches_r1.csv profiles/r1_launches.md
  python tools/summarize_ncu.py full gpurun_out/prof_r1_dr_v2.ncu-rep profiles/r1_ncu_dr_stages.md
"""
import collections
import csv
import io
import re
import subprocess
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    m = re.match(r"hmcb::k_transition<hmcb::(\w+)<(\w+)>, \w+, (?:\(int\))?(\d+), (?:\(int\))?(\d+), (?:\(bool\))?(\d), (?:\(int\))?(\d)", name)
    if m:
        mode = {"0": "HMC", "1": "DR stage 0", "2": "DR stage>=1"}[m.group(6)]
        return "hmcb::k_transition<%s,%s,EPL=%s,LPC=%s> [%s]" % (m.group(1), m.group(2), m.group(3), m.group(4), mode)
    m = re.match(r"hmcb::k_transition_ws<hmcb::(\w+)<(\w+)>, \w+, (?:\(int\))?(\d+), (?:\(int\))?(\d)", name)
    if m:
        mode = {"0": "HMC", "1": "DR stage 0"}[m.group(4)]
        return "hmcb::k_transition_ws<%s,%s,EPL=%s> [%s, warp-specialised]" % (m.group(1), m.group(2), m.group(3), mode)
    name = name.replace("hmcb::lr::", "hmcb::lr::").replace("(bool)", "").replace("(int)", "")
    return re.sub(r"\(.*", "", name)[:100]


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    idx = {h: i for i, h in enumerate(rows[0])}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = short(r[idx["Kernel Name"]])
        a = agg.setdefault(k, [0, 0.0, r[idx["Grid Size"]], r[idx["Block Size"]]])
        a[0] += 1
        a[1] += float(r[idx["Metric Value"]].replace(",", "")) / 1e3
    total = sum(a[1] for a in agg.values())
    mine = sum(a[1] for k, a in agg.items() if k.startswith("hmcb::"))
    with open(dst, "w") as f:
        f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none)\n\n")
        f.write("Source: `%s`.  Per-launch times under ncu are cold-cache and serialised: read the SHARES.\n\n" % src)
        f.write("All launches: %d, %.1f us total; libhmcb kernels: %.1f us = %.1f%% of GPU time.\n\n" %
                (sum(a[0] for a in agg.values()), total, mine, 100 * mine / total))
        f.write("| kernel | launches | total us | avg us | share | grid | block |\n|---|---:|---:|---:|---:|---|---|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %d | %.1f | %.1f | %.1f%% | %s | %s |\n" % (k, a[0], a[1], a[1] / a[0], 100 * a[1] / total, a[2], a[3]))


WANT = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "sm__sass_inst_executed_op_local_ld.sum", "sm__sass_inst_executed_op_local_st.sum",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "TPC.TriageCompute.sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg",
        "smsp__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_bytes.sum"]


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(dst, "w") as f:
        f.write("# ncu --set full summary\n\nSource: `%s` (ncu -i ... --page raw --csv).\n\n" % src)
        for r in rows[2:]:
            f.write("## %s\n\n| metric | value | unit |\n|---|---:|---|\n" % short(r[idx["Kernel Name"]]))
            for w in WANT:
                if w in idx and r[idx[w]] not in ("", "n/a"):
                    f.write("| %s | %s | %s |\n" % (w, r[idx[w]], units[idx[w]]))
            f.write("\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])

"""Turns an .ncu-rep (kept out of git) into the small text summary that is committed.

    python profiles/summarize.py gpurun_out/prof.ncu-rep profiles/r01/name.md [series_in_launch]
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.sum", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    n_series = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    d = dict(zip(hdr, vals))
    u = dict(zip(hdr, units))
    lines = [f"# ncu summary of `{rep.split('/')[-1]}`", "", f"kernel: `{d.get('Kernel Name', '?')}`", "",
             "| metric | value | unit |", "|---|---|---|"]
    for k in KEYS:
        if k in d:
            lines.append(f"| {k} | {d[k]} | {u.get(k, '')} |")
    stalls = sorted(((float(v), h) for h, v in d.items()
                     if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h and v),
                    reverse=True)[:8]
    lines += ["", "Warp stall reasons (warps per issue-active cycle):", ""]
    lines += [f"- {h.split('issue_stalled_')[1].split('_per_issue')[0]}: {v:.2f}" for v, h in stalls]
    try:
        dur = float(d["gpu__time_duration.sum"])
        unit = u.get("gpu__time_duration.sum", "ns")
        dur_s = dur * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(unit, 1e-9)
        tr = float(d["dram__bytes_read.sum"]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u.get("dram__bytes_read.sum"), 1)
        tw = float(d["dram__bytes_write.sum"]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u.get("dram__bytes_write.sum"), 1)
        lines += ["", f"DRAM traffic per launch: read {tr / 1e6:.1f} MB + write {tw / 1e6:.1f} MB = {(tr + tw) / 1e6:.1f} MB "
                      f"({(tr + tw) / dur_s / 1e9:.0f} GB/s over {dur_s * 1e3:.3f} ms under ncu)"]
        if n_series:
            inst = float(d.get("smsp__inst_executed.sum", 0))
            lines += [f"Per series: {(tr + tw) / n_series:.0f} B of DRAM traffic, {inst / n_series:.0f} warp instructions"]
    except Exception as e:  # keep the table even if a unit is unexpected
        lines += ["", f"(derived numbers unavailable: {e})"]
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()

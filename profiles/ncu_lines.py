"""Executed warp instructions per source line of one kernel: joins the SASS listing of an ncu
report (`ncu -i X.ncu-rep --page source --csv`, column "Instructions Executed") with the line table
of the same kernel in the built library (`cuobjdump -xelf` + `nvdisasm -g -c`), instruction by
instruction.  Needs the library the report was taken from (same build).

    python profiles/ncu_lines.py gpurun_out/X.ncu-rep foremast_1_b200/lib/libforemast_cuda.so <kernel substring> \
           <series in the launch> [--phases file:lo-hi=name ...]
"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_with_lines(so, pat):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
    out = []
    for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
        txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
        cur, where = None, ("?", 0)
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                cur = m.group(1)
                continue
            if cur is None or pat not in cur:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                where = (m.group(1).split("/")[-1], int(m.group(2)))
                continue
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m:
                out.append((m.group(1), where))
        if out:
            return out
    raise SystemExit("kernel not found in " + so)


def main():
    rep, so, pat, n_series = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[1]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("Source")
    stall_cols = {h: i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h}
    samp = hdr.index("# Samples")
    execd = [(r[cs].split()[0] if not r[cs].strip().startswith("@") else r[cs].split()[1], int(r[ci]), int(r[samp]),
              {h: int(r[i] or 0) for h, i in stall_cols.items()}) for r in rows[2:] if len(r) > ci]
    sass = sass_with_lines(so, pat)
    if len(sass) != len(execd):
        print(f"warning: {len(sass)} instructions in the library, {len(execd)} in the report", file=sys.stderr)
    per_line = collections.Counter()
    per_line_samples = collections.Counter()
    per_op = collections.Counter()
    per_line_op = collections.Counter()
    stalls = collections.Counter()
    for (op, where), (op2, n, smp, st) in zip(sass, execd):
        per_line[where] += n
        per_line_samples[where] += smp
        per_op[op.split(".")[0]] += n
        per_line_op[(where, op.split(".")[0])] += n
        for h, v in st.items():
            stalls[h] += v
    total = sum(per_line.values())
    tot_s = sum(per_line_samples.values()) or 1
    print(f"{total / n_series:.0f} warp instructions per series ({total} in the launch)")
    print("\nper source line (>= 0.4 % of the instructions or of the stall samples):")
    for (f, l), n in sorted(per_line.items()):
        if n >= 0.004 * total or per_line_samples[(f, l)] >= 0.004 * tot_s:
            print(f"  {n / n_series:8.1f}  {100.0 * per_line_samples[(f, l)] / tot_s:5.1f} % samples  {f}:{l}")
    print("\nper opcode:")
    for o, n in per_op.most_common(30):
        print(f"  {n / n_series:8.1f}  {o}")
    if "--line-ops" in sys.argv:
        print("\nper (line, opcode), >= 0.3 % of the instructions:")
        for ((f, l), o), n in sorted(per_line_op.items()):
            if n >= 0.003 * total:
                print(f"  {n / n_series:8.1f}  {f}:{l}  {o}")
    print("\nstall samples:")
    for h, v in stalls.most_common(10):
        print(f"  {100.0 * v / tot_s:5.1f} %  {h}")


if __name__ == "__main__":
    main()

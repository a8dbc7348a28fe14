"""Static SASS statistics of one kernel, read from `nvdisasm -g -c <cubin>` output.

    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -cubin -o k.cubin file.cu
    nvdisasm -g -c k.cubin > k.sass
    python profiles/sass_hist.py k.sass <kernel-name-substring> [--lines] [--ops] [--inlined]

Prints (i) instructions per source file, (ii) with --lines per source line (file:line of the
innermost frame the line table names), (iii) with --ops the opcode histogram.  The counts are
STATIC: they equal the executed counts only for straight-line code, which is what the hot
kernels here mostly are (the networks and the per-key loops are fully unrolled).
"""
import collections
import re
import sys


def sections(path):
    cur, name = None, None
    with open(path) as f:
        for ln in f:
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                name = m.group(1)
                cur = []
                yield (name, cur)      # filled while the generator runs on: consume with list()
                continue
            if cur is not None:
                if re.match(r"\s*\.section|\s*\.text\.", ln):
                    cur = None
                else:
                    cur.append(ln)


def main():
    path, pat = sys.argv[1], sys.argv[2]
    want_lines, want_ops = "--lines" in sys.argv, "--ops" in sys.argv
    found = False
    for name, body in list(sections(path)):
        if pat not in name:
            continue
        found = True
        per_file = collections.Counter()
        per_line = collections.Counter()
        ops = collections.Counter()
        where = ("?", 0)
        n = 0
        for ln in body:
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                where = (m.group(1).split("/")[-1], int(m.group(2)))
                continue
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if not m:
                continue
            op = m.group(1)
            n += 1
            per_file[where[0]] += 1
            per_line[where] += 1
            ops[op.split(".")[0]] += 1
        print(f"## {name}: {n} instructions")
        for f, c in per_file.most_common():
            print(f"  {c:7d}  {f}")
        if want_lines:
            print("  -- per line (>= 8 instructions)")
            for (f, l), c in sorted(per_line.items()):
                if c >= 8:
                    print(f"  {c:7d}  {f}:{l}")
        if want_ops:
            print("  -- opcodes")
            for o, c in ops.most_common(40):
                print(f"  {c:7d}  {o}")
    if not found:
        sys.exit(f"no kernel matching {pat!r}")


if __name__ == "__main__":
    main()

"""Opcode histogram (static SASS) of the main kernels of the built library, to keep the
Blackwell-native evidence in the tree: UBLKCP (cp.async.bulk through the TMA unit), SYNCS (mbarrier
arrive / try_wait with transaction counts), ATOMS / LDS / STS (shared-memory multiset and search),
FMNMX (register sorting network), no HMMA / UTC*MMA (nothing here is a contraction).

    python profiles/sass_report.py > profiles/r02/sass_opcodes.txt
"""
import collections
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "foremast_1_b200", "lib", "libforemast_cuda.so")
KERNELS = ["k_canary_count", "k_canary_team", "k_canary_sortILi64ELi48", "k_canary_elect", "k_ma_onlyILi4", "k_ma_onlyILi16",
           "k_groups_sort", "k_score_series", "k_score_groups", "k_score_pairs", "k_gather_rows"]
SHOW = ["UBLKCP", "SYNCS", "ATOMS", "LDS", "STS", "LDG", "STG", "SHFL", "FMNMX", "FSETP", "ISETP", "IMAD", "DFMA", "DADD",
        "F2F", "REDUX", "VOTE", "BAR", "CCTL", "HMMA", "UTCHMMA", "RED", "ATOMG", "BRA"]


def main():
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=tmp, check=True, capture_output=True)
    hist = {}
    for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
        txt = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
        cur = None
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                cur = m.group(1)
                hist.setdefault(cur, collections.Counter())
                continue
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
            if m and cur:
                hist[cur][m.group(1)] += 1
    print("static SASS opcode counts per kernel (cuobjdump -xelf + nvdisasm of foremast_1_b200/lib/libforemast_cuda.so, sm_100a)\n")
    for k in KERNELS:
        for name, h in sorted(hist.items()):
            if k in name and sum(h.values()):
                tot = sum(h.values())
                cells = "  ".join(f"{o} {h[o]}" for o in SHOW if h.get(o))
                print(f"{k}: {tot} instructions\n    {cells}\n")
                break


if __name__ == "__main__":
    main()

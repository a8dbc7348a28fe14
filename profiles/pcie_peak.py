"""Pinned host -> device copy rate of the box (what bounds the cold end-to-end number of bench.py).

    python profiles/pcie_peak.py          GPU 0 alone, both directions
    python profiles/pcie_peak.py --all    every GPU alone, then all of them at the same time (one
                                          stream and one pinned buffer per GPU): the ratio says how
                                          much of the cold end-to-end scaling loss is the host side
"""
import json
import sys

import torch

GB = 1 << 30


def rate(pairs, reps=5):
    """pairs: [(device index, src, dst)]; all copies of a repetition are in flight together."""
    streams = [torch.cuda.Stream(device=i) for i, _, _ in pairs]
    for (i, src, dst), st in zip(pairs, streams):
        with torch.cuda.stream(st):
            dst.copy_(src, non_blocking=True)
    for i, _, _ in pairs:
        torch.cuda.synchronize(i)
    ev = []
    for (i, src, dst), st in zip(pairs, streams):
        with torch.cuda.stream(st):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            for _ in range(reps):
                dst.copy_(src, non_blocking=True)
            b.record(st)
            ev.append((a, b))
    for i, _, _ in pairs:
        torch.cuda.synchronize(i)
    return [reps * src.numel() / (a.elapsed_time(b) / 1e3) / 1e9 for (a, b), (_, src, _) in zip(ev, pairs)]


def main():
    n = torch.cuda.device_count() if "--all" in sys.argv else 1
    host = [torch.empty(GB, dtype=torch.uint8).pin_memory() for _ in range(n)]
    dev = [torch.empty(GB, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
    out = {"gpus": n}
    out["h2d_GBps_alone"] = [round(rate([(i, host[i], dev[i])])[0], 1) for i in range(n)]
    out["d2h_GBps_alone"] = [round(rate([(i, dev[i], host[i])])[0], 1) for i in range(n)]
    if n > 1:
        tog = rate([(i, host[i], dev[i]) for i in range(n)])
        out["h2d_GBps_together"] = [round(x, 1) for x in tog]
        out["h2d_GBps_together_sum"] = round(sum(tog), 1)
        out["h2d_together_over_alone"] = round(sum(tog) / sum(out["h2d_GBps_alone"]), 3)
    else:
        out["h2d_GBps"], out["d2h_GBps"] = out["h2d_GBps_alone"][0], out["d2h_GBps_alone"][0]
    print(json.dumps(out))


if __name__ == "__main__":
    main()

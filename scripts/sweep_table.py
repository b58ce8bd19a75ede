#!/usr/bin/env python
"""Markdown table of a scripts/comm_check.py --sweep JSON: per size the time of every algorithm, the best bus bandwidth against
NCCL's on the same box, and the achieved fraction of the NVLS traffic model S(N+1)/N per direction at 900 GB/s."""
import json
import sys

d = json.load(open(sys.argv[1]))
N = d["world"]
print("%d x B200, fp32, in place, timed inside one CUDA graph (10 back-to-back all-reduces per replay), max over ranks.  "
      "NVLS multicast bound: %s.\n" % (N, d["nvls"]))
print("| bytes | one-shot us | two-shot us | NVLS us | NVLS 32 CTAs us | NCCL us | best bus GB/s | NCCL bus GB/s | NCCL time / best | of S(N+1)/N at 900 GB/s |")
print("|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
for r in d["sweep"]:
    b = r["bytes"]
    ours = min(v for k, v in r.items() if k.endswith("_us") and not k.startswith("nccl"))
    bus = b / (ours * 1e-6) * 2 * (N - 1) / N / 1e9
    ideal = b * (N + 1) / N / 900e9 * 1e6
    f = lambda k: ("%.1f" % r[k]) if k in r else "-"
    nccl = r.get("nccl_us")
    print("| %d | %s | %s | %s | %s | %s | %.1f | %s | %s | %.0f%% |" % (
        b, f("oneshot_us"), f("twoshot_us"), f("nvls_us"), f("nvls_c32_us"), f("nccl_us"), bus,
        ("%.1f" % r["nccl_busbw_GBs"]) if nccl else "-", ("%.2fx" % (nccl / ours)) if nccl else "-", 100 * ideal / ours))

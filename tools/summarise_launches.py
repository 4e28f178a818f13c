"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> per-kernel totals as a markdown table.
usage: python tools/summarise_launches.py gpurun_out/launches.csv > table.md"""
import csv
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        ms = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v
        rows.append((r["Kernel Name"], ms))
tot, cnt = defaultdict(float), defaultdict(int)
for k, ms in rows:
    tot[k] += ms
    cnt[k] += 1
total = sum(tot.values())
OURS = ("bgpu", "widen_kernel", "gather_kernel", "small_pack_kernel", "small_unpack_kernel", "check_offsets_kernel",
        "partition_offsets_noalloc_kernel", "concat_kernel", "ipc_", "lz4_")


def is_ours(k):
    return any(t in k for t in OURS)


ours = sum(v for k, v in tot.items() if is_ours(k))
print(f"{len(rows)} launches, {total:.2f} ms in all kernels, {ours:.2f} ms in this library's kernels\n")
print("| kernel | launches | total ms | share of all | share of ours |\n|---|---:|---:|---:|---:|")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:28]:
    mine = is_ours(k)
    print(f"| `{k[:100]}` | {cnt[k]} | {v:.3f} | {100 * v / total:.1f}% | {f'{100 * v / ours:.1f}%' if mine and ours else ''} |")

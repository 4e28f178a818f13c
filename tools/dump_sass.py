"""Summarise the SASS of one kernel family of libbgpu.so as markdown (mnemonic histogram, every memory instruction form,
an excerpt around the first LDGSTS / LDG). usage: python tools/dump_sass.py <name substring> "<title>" > profiles/sass/<x>.md"""
import collections
import re
import subprocess
import sys

name, title = sys.argv[1], sys.argv[2]
lib = "arrow-ballista_b200/libbgpu.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
funcs, cur, keep = [], None, False
lines = []
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        keep = name in m.group(1)
        if keep:
            funcs.append(m.group(1))
        continue
    if keep and re.match(r"\s*/\*[0-9a-f]{4,}\*/", line):
        lines.append(line)
mn = collections.Counter()
for l in lines:
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", l)
    if m:
        mn[m.group(1)] += 1
demangled = subprocess.run(["c++filt"] + funcs, capture_output=True, text=True).stdout.split("\n")
print(f"# SASS of `{name}` (sm_100a): {title}\n")
print(f"`cuobjdump -sass {lib}`, functions matching `{name}` (nvcc 12.9, `-gencode arch=compute_100a,code=sm_100a -lineinfo -O3`):\n")
for d in demangled:
    if d.strip():
        print(f"* `{d.strip()[:160]}`")
print(f"\n{sum(mn.values())} instructions. Mnemonic histogram (full mnemonics incl. modifiers, top 40):\n")
print("| mnemonic | count |\n|---|---:|")
for k, v in mn.most_common(40):
    print(f"| `{k}` | {v} |")
mem = sorted((k, v) for k, v in mn.items() if re.match(r"(LD|ST|ATOM|RED|LDGSTS|LDGDEPBAR|DEPBAR|MATCH|SHFL|VOTE|UBLKCP|SYNCS|BAR)", k))
print("\nMemory / warp-wide / barrier instructions (every distinct form):\n")
for k, v in mem:
    print(f"* `{k}` x {v}")
first = next((i for i, l in enumerate(lines) if "LDGSTS" in l), None)
if first is None:
    first = next((i for i, l in enumerate(lines) if "LDG" in l), 0)
print("\nExcerpt (the instructions around the first LDGSTS / LDG):\n\n```")
for l in lines[max(0, first - 24): first + 40]:
    print(l.rstrip()[:118])
print("```")

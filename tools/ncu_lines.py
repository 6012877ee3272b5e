#!/usr/bin/env python
"""Per-source-line warp-stall samples of one kernel: joins `ncu --page source` (SASS, sampled) with `nvdisasm --print-line-info`.

    python tools/ncu_lines.py <rep.ncu-rep> <kernel-substring> [lib.so] [top] [mangled-substring]
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

rep, kern = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else "physics_informed_reinforcement_learning_b200/libpirl_rcwa.so"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sect = sys.argv[5] if len(sys.argv) > 5 else kern      # substring of the MANGLED name when several instances exist

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
dis = ""          # the library holds one cubin per translation unit
for f in sorted(os.listdir(tmp)):
    if f.endswith(".cubin"):
        dis += subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, f)], capture_output=True, text=True).stdout
# offset -> line for the wanted function
off2line, cur, infn = {}, None, False
for ln in dis.splitlines():
    if ln.startswith("\t.section\t.text."):
        infn = sect in ln
    if not infn:
        continue
    m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        off2line[int(m.group(1), 16)] = (cur, m.group(2).strip())

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
agg = collections.Counter()
inst = collections.Counter()
total = 0
i = 0
seen_kernel = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Kernel Name":
        use = kern in r[1] and seen_kernel == 0
        if kern in r[1]:
            seen_kernel += 1
        hdr = rows[i + 1]
        ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
        j = i + 2
        base = None
        while j < len(rows) and rows[j] and rows[j][0] != "Kernel Name":
            if use:
                addr = int(rows[j][ia], 16)
                if base is None:
                    base = addr
                key = off2line.get(addr - base, ((None, -1), ""))[0]
                s = int(rows[j][isamp] or 0)
                agg[key] += s
                inst[key] += int(rows[j][iex] or 0)
                total += s
            j += 1
        i = j
    else:
        i += 1
print(f"total samples {total}")
src_cache = {}
for key, s in agg.most_common(top):
    f, l = key if key else (None, -1)
    text = ""
    if f:
        for root, _, files in os.walk("physics_informed_reinforcement_learning_b200/csrc"):
            if f in files:
                if f not in src_cache:
                    src_cache[f] = open(os.path.join(root, f)).read().splitlines()
                text = src_cache[f][l - 1].strip()[:100] if l - 1 < len(src_cache[f]) else ""
    print(f"{100.0 * s / max(total, 1):5.1f}%  inst {inst[key]:9d}  {f}:{l}  {text}")

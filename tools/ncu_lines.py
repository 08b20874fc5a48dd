"""Dev: stall samples / executed instructions of one kernel of an .ncu-rep, aggregated per SOURCE LINE.
The CLI's source page is per SASS instruction without line numbers; `nvdisasm -g` of the same cubin gives the line of every
instruction in the same order, so the two lists are joined by position.
usage: ncu_lines.py report.ncu-rep kernel_regex mangled_function_substring [top_n]   (library must be the profiled build)"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kre, mangled = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
lib = os.path.join(ROOT, "imagined-speech-decoding_b200", "libfastb200.so")

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
lines_of = None
for f in sorted(os.listdir(tmp)):
    if not f.endswith(".cubin"):
        continue
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur, fn, seq = None, None, collections.defaultdict(list)
    for ln in dis.splitlines():
        m = re.match(r"^\.text\.(\S+):", ln)
        if m:
            fn = m.group(1)
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if fn and re.match(r"^\s+/\*[0-9a-f]{4,}\*/", ln):
            seq[fn].append(cur)
    for k, v in seq.items():
        if mangled in k:
            lines_of = v
            break
    if lines_of:
        break
assert lines_of, "function not found in the library"

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
hx = {c: i for i, c in enumerate(hdr)}
data = []
for r in rows[2:]:                                   # the page repeats per matching launch: keep the first one
    if r and r[0] == "Kernel Name":
        break
    if len(r) == len(hdr) and r[0] != "Address":
        data.append(r)
assert len(data) == len(lines_of), (len(data), len(lines_of))
agg = collections.defaultdict(lambda: [0.0, 0.0, collections.Counter()])
stalls = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
for r, key in zip(data, lines_of):
    a = agg[key]
    a[0] += float(r[hx["# Samples"]] or 0)
    a[1] += float(r[hx["Instructions Executed"]] or 0)
    for s in stalls:
        v = float(r[hx[s]] or 0)
        if v:
            a[2][s] += v
tot_s = sum(a[0] for a in agg.values()) or 1
tot_i = sum(a[1] for a in agg.values()) or 1
src_cache = {}


def src(key):
    fn, ln = key
    if fn not in src_cache:
        p = os.path.join(ROOT, "imagined-speech-decoding_b200", "csrc", fn)
        src_cache[fn] = open(p).read().splitlines() if os.path.exists(p) else []
    t = src_cache[fn]
    return t[ln - 1].strip()[:90] if 0 < ln <= len(t) else ""


print(f"# {rows[0][1][:100]}: {tot_s:.0f} samples, {tot_i:.3g} warp instructions, {len(data)} SASS instructions")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    why = ", ".join(f"{k[6:]} {v / a[0] * 100:.0f}%" for k, v in a[2].most_common(3)) if a[0] else ""
    print(f"{a[0] / tot_s * 100:5.1f}% smp {a[1] / tot_i * 100:5.1f}% inst  {key[0]}:{key[1]:<4d} {src(key)}   [{why}]")

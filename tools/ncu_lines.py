"""Dynamic instruction counts and stall samples per SOURCE LINE of one kernel: joins the SASS page
of an .ncu-rep (executed counts, in program order) with nvdisasm's line info of the same cubin.
usage: python tools/ncu_lines.py report.ncu-rep <cubin> <mangled kernel name> [units] [top]
`units` divides the counts (e.g. the number of warp-tiles) so they read as "per unit"."""
import collections
import csv
import io
import re
import subprocess
import sys

rep, cubin, kern = sys.argv[1:4]
units = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text." + kern + ":"))
lines, cur = [], ("?", 0)
for l in dis[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section") and lines:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+\S", l):
        lines.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blks = out.split('"Kernel Name"')[1:]
base = re.sub(r"^_ZN4lynx\d+", "", kern).split("I")[0].split("E")[0]  # e.g. track_pipe_kernel
blk = next((b for b in blks if base in b.splitlines()[0]), blks[0])
rows = list(csv.reader(io.StringIO("\n".join(blk.splitlines()[1:]))))
hdr = rows[0]
E, S = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = [r for r in rows[1:] if len(r) > E and r[E].isdigit()]
print(f"# {rep}: {len(data)} SASS rows in the report, {len(lines)} in the cubin")
n = min(len(data), len(lines))
ex, sm = collections.Counter(), collections.Counter()
for (f, ln), r in zip(lines[:n], data[:n]):
    ex[(f, ln)] += int(r[E])
    sm[(f, ln)] += int(r[S]) if r[S].isdigit() else 0
tot, tots = sum(ex.values()), sum(sm.values())
print(f"# total warp instructions {tot} ({tot / units:.1f} per unit), stall samples {tots}")
src = {}
for (f, ln), v in ex.most_common(top):
    if f not in src:
        try:
            src[f] = open("lynx_b200/csrc/" + f).read().splitlines()
        except OSError:
            src[f] = []
    text = src[f][ln - 1].strip()[:90] if 0 < ln <= len(src[f]) else ""
    print(f"{v / units:8.1f} {100 * v / tot:5.1f}%  samples {100 * sm[(f, ln)] / max(tots, 1):5.1f}%  {f}:{ln}  {text}")

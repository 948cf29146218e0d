"""Summarise an ncu report's source page: top SASS instructions by stall samples and the
breakdown of stall reasons.  usage: python tools/ncu_stalls.py report.ncu-rep [top] [launch]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name"')
blk = '"Kernel Name"' + blocks[int(sys.argv[3]) if len(sys.argv) > 3 else 1]
lines = blk.splitlines()
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]
S = hdr.index("# Samples")
src = hdr.index("Source")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
tot = {hdr[i]: 0 for i in stall_cols}
total = 0
for r in rows[1:]:
    if len(r) <= S or not r[S].isdigit():
        continue
    n = int(r[S])
    total += n
    for i in stall_cols:
        try:
            tot[hdr[i]] += int(r[i])
        except ValueError:
            pass
    data.append((n, r[src], {hdr[i]: r[i] for i in stall_cols if r[i] not in ("0", "")}))
print("kernel:", lines[0][:120])
print("total samples", total)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]:
    print(f"  {k:28s} {v:8d} {100.0 * v / max(total, 1):5.1f}%")
data.sort(key=lambda t: -t[0])
for n, s, d in data[:top]:
    print(f"{n:6d} {100.0 * n / max(total, 1):5.1f}%  {s[:70]:70s} {d}")

"""Top stall sites of one kernel from an ncu report's source page (SASS level).
   python tools/ncu_hot.py report.ncu-rep <kernel-name-substring | #index> [topN]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]   # pat: substring of the full kernel name, or #<launch index in the report>
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
if pat.startswith("#"):
    k = int(pat[1:])
else:
    k = next(n for n, i in enumerate(start) if pat in lines[i])
start = start[k:k + 2]
seg = lines[start[0] + 1:(start[1] if len(start) > 1 else len(lines))]
rows = list(csv.reader(seg))
hdr, data = rows[0], rows[1:]
si = hdr.index("# Samples"); src = hdr.index("Source")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si] or 0) for r in data)
print(lines[start[0]][:120]); print("total samples", tot)
order = sorted(range(len(data)), key=lambda i: -int(data[i][si] or 0))[:top]
for i in sorted(order):
    r = data[i]
    st = sorted(((int(r[j] or 0), hdr[j][6:]) for j in stalls), reverse=True)[:3]
    print(f"{i:5d} {int(r[si]):7d} {100*int(r[si])/max(tot,1):5.1f}%  {r[src].strip()[:70]:70s} {' '.join(f'{n}:{c}' for c,n in st if c)}")

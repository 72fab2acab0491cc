"""Address-ordered SASS of a kernel from an ncu report with per-instruction executed counts and active threads:
    python profiles/sass_profile.py X.ncu-rep > out.txt     (share = % of the kernel's warp instructions)"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; data = []
for r in rows:
    if r and r[0] == "Address": hdr = r; continue
    if hdr and len(r) >= len(hdr) and r[0].startswith("0x"):
        try: data.append((int(r[0], 16), r[1].strip(), int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("Thread Instructions Executed")]), int(r[hdr.index("# Samples")] or 0)))
        except ValueError: pass
total = sum(d[2] for d in data) or 1
base = data[0][0]; cum = 0
print(f"# total warp instructions {total}")
for a, s, n, t, smp in data:
    cum += n
    print(f"{a - base:5x} {100 * n / total:6.3f}% cum {100 * cum / total:6.2f}%  thr {t / max(n, 1):5.2f}  smp {smp:6d}  {s}")

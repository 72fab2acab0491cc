"""Issue-slot split of a kernel by SOURCE LINE from an ncu report (needs -lineinfo + --import-source on):
    python profiles/region_split.py gpurun_out/X.ncu-rep [top_n]
Per source file:line: warp instructions executed, share of the kernel, average active threads. Run here (no GPU)."""
import csv
import io
import subprocess
import sys
from collections import defaultdict


def main(path, top=45):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    cur_file = None; hdr = None; cur_line = None
    agg = defaultdict(lambda: [0, 0, ""])
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]; continue
        if r[0] == "Line No":
            hdr = r; continue
        if hdr is None or len(r) < len(hdr):
            continue
        if r[0] != "":   # a source line row: its SASS rows follow
            cur_line = (cur_file, int(r[0]), r[1].strip()); continue
        try:
            inst = int(r[hdr.index("Instructions Executed")]); thr = int(r[hdr.index("Thread Instructions Executed")])
        except ValueError:
            continue
        if cur_line:
            a = agg[cur_line[:2]]; a[0] += inst; a[1] += thr; a[2] = cur_line[2]
    total = sum(a[0] for a in agg.values()) or 1
    print(f"total warp instructions {total}, mean threads {sum(a[1] for a in agg.values()) / total:.2f}")
    by_file = defaultdict(lambda: [0, 0])
    for (f, l), a in agg.items():
        by_file[f][0] += a[0]; by_file[f][1] += a[1]
    for f, a in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print(f"  {f:24s} {100 * a[0] / total:5.1f}%  threads {a[1] / max(a[0], 1):5.2f}")
    print("\nshare   threads  file:line  source")
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{100 * a[0] / total:5.2f}%  {a[1] / max(a[0], 1):5.2f}  {f}:{l}  {a[2][:110]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)

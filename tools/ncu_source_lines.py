"""Per-source-line instruction totals of an ncu report (kernel launch ID given): the hottest
lines by warp instructions executed, with lanes on and stall samples.
Usage: python tools/ncu_source_lines.py REPORT.ncu-rep [launch-index] [top-n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
idx = sys.argv[2] if len(sys.argv) > 2 else "0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass",
                      "--launch-skip", idx, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit():
        d = dict(zip(hdr, r))
        num = lambda k: int(d[k]) if d.get(k, "").isdigit() else 0
        lines.append((cur_file, int(r[0]), r[1].strip(), num("Instructions Executed"),
                      num("Thread Instructions Executed"), num("# Samples")))
tot = sum(x[3] for x in lines)
tots = sum(x[5] for x in lines)
print(f"total warp instructions {tot}, samples {tots}")
byfile = {}
for f, ln, src, ex, th, s in lines:
    a = byfile.setdefault(f, [0, 0, 0]); a[0] += ex; a[1] += th; a[2] += s
for f, (ex, th, s) in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print(f"  {f:28s} {100 * ex / tot:5.1f}% of instr, lanes {th / max(1, ex):4.1f}, {100 * s / max(1, tots):5.1f}% of samples")
for f, ln, src, ex, th, s in sorted(lines, key=lambda x: -x[3])[:top]:
    print(f"{100 * ex / tot:5.1f}% lanes {th / max(1, ex):4.1f} smp {100 * s / max(1, tots):4.1f}%  {f}:{ln}  {src[:100]}")

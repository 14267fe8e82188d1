"""Per-region instruction totals of an `ncu --page source --csv` dump (one kernel).
Usage: python tools/ncu_source_regions.py FILE.csv [markers]   -- prints every VOTE / LDG.128 /
CALL with its offset so regions can be read off, then totals between the given hex offsets."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = []
for r in rows[hi + 1:]:
    if r and r[0] in ("Kernel Name", "Address"):
        break  # next kernel of the report
    if len(r) > 10:
        data.append(r)
ia, isrc = hdr.index("Address"), hdr.index("Source")
iex, ith, ismp = (hdr.index(k) for k in ("Instructions Executed", "Thread Instructions Executed", "# Samples"))
base = int(data[0][ia], 16)
tot_ex = sum(int(r[iex]) for r in data)
tot_th = sum(int(r[ith]) for r in data)
tot_s = sum(int(r[ismp]) for r in data)
print("total warp instr", tot_ex, "thread instr", tot_th, "avg thr %.1f" % (tot_th / tot_ex), "samples", tot_s)
marks = [int(x, 16) for x in sys.argv[2:]]
if not marks:
    for r in data:
        off = int(r[ia], 16) - base
        if any(k in r[isrc] for k in ("VOTE", "LDG.E.128", "CALL", "BAR", "ATOM")):
            print(hex(off), r[isrc].strip()[:70], r[iex], round(int(r[ith]) / max(1, int(r[iex])), 1))
else:
    marks = [0] + marks + [1 << 30]
    for lo, hi_ in zip(marks[:-1], marks[1:]):
        ex = th = s = 0
        for r in data:
            off = int(r[ia], 16) - base
            if lo <= off < hi_:
                ex += int(r[iex]); th += int(r[ith]); s += int(r[ismp])
        print(f"[{lo:#7x},{hi_:#7x}) warp-instr {ex:>11d} ({100 * ex / tot_ex:5.1f}%) avg thr {th / max(1, ex):5.1f} samples {100 * s / max(1, tot_s):5.1f}%")

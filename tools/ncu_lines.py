"""Per-source-line instruction and stall-sample shares from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None; lines = {}
for r in rows:
    if r and r[0] == "Line No": hdr = r; iI = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit(): continue
    k = int(r[0]); a = lines.setdefault(k, [r[1], 0, 0]); a[1] += num(r[iI]); a[2] += num(r[iS])
tot = sum(v[1] for v in lines.values()) or 1; ts = sum(v[2] for v in lines.values()) or 1
print("total warp instructions", tot, "samples", ts)
for k, v in sorted(lines.items(), key=lambda x: -x[1][1])[:top]:
    print(f"{k:5d} {100*v[1]/tot:5.1f}% inst {100*v[2]/ts:5.1f}% samp | {v[0][:120]}")

"""Top SASS lines by stall samples from `ncu --page source --csv` output.  python tools/ncu_hot.py file.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, iall, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
body = [r for r in rows[hi + 1:] if len(r) > iall and r[iall].strip().isdigit()]
tot = sum(int(r[iall]) for r in body)
print("total samples", tot, "instructions", len(body))
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
order = sorted(range(len(body)), key=lambda k: -int(body[k][iall]))[:top]
for k in sorted(order):
    r = body[k]
    st = sorted(((int(r[i]) if r[i].isdigit() else 0, hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{k:6d} {int(r[iall]):7d} {100*int(r[iall])/tot:5.1f}%  ex={r[iex]:>9s}  {r[isrc][:70]:70s} {st}")

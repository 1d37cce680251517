"""Shared-memory instructions with excess wavefronts from `ncu --page source --csv`.  python tools/ncu_smem.py file.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
isrc, iex = hdr.index("Source"), hdr.index("Instructions Executed")
iw = hdr.index("L1 Wavefronts Shared"); ie = hdr.index("L1 Wavefronts Shared Excessive"); iid = hdr.index("L1 Wavefronts Shared Ideal")
body = [r for r in rows[hi + 1:] if len(r) > ie]
body = body[:len(body) // 2]
f = lambda s: int(float(s)) if s.strip() else 0
print("total wavefronts", sum(f(r[iw]) for r in body), "excessive", sum(f(r[ie]) for r in body))
for k in sorted(range(len(body)), key=lambda k: -f(body[k][ie]))[:top]:
    r = body[k]
    print(f"{k:6d} ex={r[iex]:>9s} wave={f(r[iw]):>10d} ideal={f(r[iid]):>10d} excess={f(r[ie]):>10d}  {r[isrc][:80]}")

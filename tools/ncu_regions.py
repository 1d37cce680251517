"""Stall-sample totals by reason and by opcode for an instruction range of `ncu --page source --csv` output.
python tools/ncu_regions.py file.csv [first last]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
isrc, iall, iex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
body = [r for r in rows[hi + 1:] if len(r) > iall and r[iall].strip().isdigit()]
body = body[:len(body) // 2] if len(body) > 2 and body[0][isrc] == body[len(body) // 2][isrc] else body
lo, hi_ = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (0, len(body))
sel = body[lo:hi_]
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iall]) for r in sel)
print("instructions", len(sel), "samples", tot, "of", sum(int(r[iall]) for r in body))
by = collections.Counter()
for r in sel:
    for i in stall_cols:
        if r[i].isdigit(): by[hdr[i][6:]] += int(r[i])
print("by reason:", [(k, v, f"{100*v/tot:.1f}%") for k, v in by.most_common(8)])
op = collections.Counter(); opn = collections.Counter()
for r in sel:
    s = r[isrc].split()
    o = s[1] if s and s[0].startswith("@") else (s[0] if s else "")
    o = o.split(".")[0] + ("." + o.split(".")[1] if o.startswith("MUFU") else "")
    op[o] += int(r[iall]); opn[o] += 1
print("by opcode:", [(k, v, f"{100*v/tot:.1f}%", opn[k]) for k, v in op.most_common(14)])

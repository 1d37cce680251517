"""ncu launch-list CSV (gpu__time_duration.sum) -> markdown table.  python tools/launch_list_md.py in.csv steps title > out.md"""
import csv, re, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr, body = rows[0], rows[1:]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in body:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("b200::", "").replace("(anonymous namespace)::", "")
    name = re.sub(r"\((int|bool)\)", "", name)
    v = float(r[iv].replace(",", "")) * ({"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[iu], 1e-6))
    tot[name] += v; cnt[name] += 1
steps = int(sys.argv[2]); total = sum(tot.values())
print(f"# {sys.argv[3]}\n")
print("| kernel | launches | total ms | share |\n|---|---|---|---|")
for k, v in tot.most_common():
    print(f"| `{k}` | {cnt[k]} | {v:.3f} | {100 * v / total:.1f}% |")
print(f"\nTotal {total:.1f} ms over {steps} steps = {total / steps:.1f} ms/step under ncu (cold-cache, serialised launches: compare shares, not absolutes).")

"""Print the hottest SASS/source lines (by stall samples) of an ncu --page source --csv dump."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
isamp = hdr.index("# Samples")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    if len(r) <= isamp:
        continue
    try:
        n = int(r[isamp] or 0)
    except ValueError:
        continue
    data.append((n, r))
tot = sum(n for n, _ in data) or 1
print("total samples", tot)
agg = {}
for n, r in data:
    for i in stalls:
        try:
            v = int(r[i] or 0)
        except ValueError:
            v = 0
        agg[hdr[i]] = agg.get(hdr[i], 0) + v
print({k: round(100 * v / tot, 1) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]})
for n, r in sorted(data, key=lambda x: -x[0])[:top]:
    st = sorted(((int(r[i] or 0), hdr[i]) for i in stalls), reverse=True)[:2]
    print(f"{100*n/tot:5.1f}%  {r[1][:110]:110s} {st}")

"""Summarise an `ncu --page source --csv` export: per kernel instance, the SASS lines with the most stall samples.
    python tools/ncu_src_top.py file.src.csv [instance] [topN]"""
import csv, sys
csv.field_size_limit(10**9)
rows = list(csv.reader(open(sys.argv[1])))
inst = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
blk = rows[starts[inst]:starts[inst + 1]]
h = blk[1]
ci = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
body = [r for r in blk[2:] if len(r) == len(h)]
tot = sum(int(r[ci["# Samples"]] or 0) for r in body)
print("instance", inst, "lines", len(body), "samples", tot)
agg = {s: sum(int(r[ci[s]] or 0) for r in body) for s in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
for r in sorted(body, key=lambda r: -int(r[ci["# Samples"]] or 0))[:top]:
    st = {s[6:]: int(r[ci[s]]) for s in stalls if int(r[ci[s]] or 0)}
    st = dict(sorted(st.items(), key=lambda x: -x[1])[:3])
    print(f'{r[ci["Address"]][-5:]} {int(r[ci["# Samples"]]):6d} {int(r[ci["Instructions Executed"]] or 0):9d}  {r[ci["Source"]][:70]:70s} {st}')

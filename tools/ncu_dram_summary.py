"""Summarise an ncu metrics pass (`--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`)
of one bench step: per-kernel launch count, time and DRAM bytes; writes the conv1d_tc per-launch average that
bench.py reports as roofline.traffic.

    python tools/ncu_dram_summary.py gpurun_out/r1_dram.csv <workload> profiles/conv1d_tc_dram.json profiles/r1_launches_summary.md
"""
import csv, json, sys
from collections import defaultdict

src, workload, out_json, out_md = sys.argv[1:5]
rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
h = rows[0]
ki, mi, vi, ui, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0,
         "msecond": 1e3, "second": 1e6}
per = defaultdict(lambda: defaultdict(float))
ids = defaultdict(set)
for r in rows[1:]:
    name = r[ki].split("(")[0].replace("void ", "").replace("b200svs::", "")
    val = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1.0)
    per[name][r[mi]] += val
    ids[name].add(r[ii])
tot_t = sum(v["gpu__time_duration.sum"] for v in per.values())
lines = ["| kernel | launches | time (ms) | share | DRAM read (GB) | DRAM write (GB) | DRAM GB/s |", "|---|---|---|---|---|---|---|"]
for name, v in sorted(per.items(), key=lambda x: -x[1]["gpu__time_duration.sum"]):
    t = v["gpu__time_duration.sum"]
    rd, wr = v["dram__bytes_read.sum"], v["dram__bytes_write.sum"]
    lines.append(f"| {name} | {len(ids[name])} | {t / 1e3:.2f} | {100 * t / tot_t:.1f} % | {rd / 1e9:.2f} | {wr / 1e9:.2f} | {(rd + wr) / max(t, 1e-9) / 1e3:.0f} |")
open(out_md, "w").write(f"# ncu launch summary - {workload}\n\nSource: `{src}` (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,"
                        f"dram__bytes_write.sum --clock-control none; per-launch times are serialised and cold-cache, compare shares).\n\n"
                        + "\n".join(lines) + f"\n\nTotal kernel time {tot_t / 1e3:.2f} ms.\n")
# both instantiations of the conv kernel (<0> single CTA, <1> CTA pairs) and the fused pair kernel (the same implicit-GEMM conv)
c = defaultdict(float)
n = 0
for name in per:
    if name.startswith("conv1d_tc_kernel") or name.startswith("resblock_pair_kernel"):
        n += len(ids[name])
        for m_, v_ in per[name].items():
            c[m_] += v_
json.dump({"workload": workload, "launches": n,
           "dram_bytes_per_launch": (c["dram__bytes_read.sum"] + c["dram__bytes_write.sum"]) / max(n, 1),
           "dram_read_bytes": c["dram__bytes_read.sum"], "dram_write_bytes": c["dram__bytes_write.sum"],
           "time_ms": c["gpu__time_duration.sum"] / 1e3, "share_of_step": c["gpu__time_duration.sum"] / tot_t,
           "source": f"ncu metrics pass over one step ({src.split('/')[-1]}), average over {n} launches"}, open(out_json, "w"), indent=1)
print(open(out_md).read())

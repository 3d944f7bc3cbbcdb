"""Print the headline and the per-stage table of bench.py JSON lines: python tools/bench_show.py file.json ..."""
import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1),
          "roofline", d["roofline"]["achieved"], d["roofline"]["frac"])
    st = d.get("stages_ms", {})
    flat = {}
    for k, v in st.items():
        if isinstance(v, dict):
            for k2, v2 in v.items():
                flat[f"{k}:{k2}"] = v2
        else:
            flat[k] = v
    print("   ", {k: round(v, 1) for k, v in sorted(flat.items(), key=lambda x: -x[1])[:14]})

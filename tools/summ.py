"""Print the measured fields of bench.py JSON lines (development helper)."""
import json
import sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    r, par = d["roofline"], d.get("parity") or {}
    print(f, "N", d.get("n_gpus"), "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 4), "frac", round(r["frac"], 3),
          "kms", round(r.get("kernel_ms_per_step") or 0, 4), "e2e", round(d["e2e"]["value"] or 0, 1), "parity", par.get("pass"),
          par.get("bit_exact"), par.get("exchange"))
    for n, c in d.get("configs", {}).items():
        if isinstance(c, dict) and "value" in c:
            r = c["roofline"]
            print("  ", n, "value", round(c["value"], 1), "ms", round(c["ms_per_step"], 4), "frac", round(r["frac"], 3), "unovl",
                  round(r.get("frac_of_unoverlapped_bound") or 0, 3), "e2e", round(c["e2e"]["value"], 1), "parity",
                  c["parity"]["pass"], c["parity"]["bit_exact"], c["parity"].get("exchange"), "wall", round(c["wall_s"], 1),
                  "launches/step", r.get("launches_per_step"))
        else:
            print("  ", n, c)

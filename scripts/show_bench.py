"""Print the phase times and rooflines of a bench.py JSON line (scratch helper)."""
import json, sys
for f in sys.argv[1:]:
    d = json.load(open(f))
    c = d["config"].get("counters", {})
    print(f, "ms/step %.3f value %.3gM e2e %.3gM" % (d["ms_per_step"], d["value"] / 1e6, d["e2e"]["value"] / 1e6))
    print("  ", {k: round(v, 3) for k, v in c.items() if k.startswith("ms_")})
    for r in d.get("rooflines", []):
        print("  ", r["kernel"], "ms %.3f frac %.3f" % (r["ms"], r["frac"]),
              {k: round(v, 3) for k, v in r.items() if k in ("frac_of_random_gather", "dpx_frac", "gcups")})

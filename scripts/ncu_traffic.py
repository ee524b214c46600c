#!/usr/bin/env python3
"""profiles/r2_traffic.json from an ncu CSV of bench.py's own command (GPU box):
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \\
      --log-file gpurun_out/traffic.csv python bench.py --chain-only --steps 1 --warmup 1 [--mode ...]
  python scripts/ncu_traffic.py gpurun_out/traffic.csv profiles/r2_traffic.json --batch 262144 --genome-mbp 3100 --nseq 24 --mode pairs
Per kernel: DRAM bytes (read + write) of its LAST launch(es) in the run -- the timed step -- in bench.py's kernel names
(the fill and the backtrace launch twice per paired step: seed-extension windows, then mate windows)."""
import argparse
import collections
import csv
import json

ap = argparse.ArgumentParser()
ap.add_argument("csv"); ap.add_argument("out")
ap.add_argument("--batch", type=int, default=262144); ap.add_argument("--genome-mbp", type=int, default=3100)
ap.add_argument("--nseq", type=int, default=24); ap.add_argument("--mode", default="pairs")
ap.add_argument("--command", default="")
a = ap.parse_args()
rows = list(csv.reader(l for l in open(a.csv) if l.startswith('"')))
h = rows[0]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = collections.OrderedDict()          # launch id -> (kernel, {metric: value})
for r in rows[1:]:
    d = per.setdefault(r[ii], [r[ki].split("(")[0].split("<")[0], {}])
    d[1][r[mi]] = float(r[vi].replace(",", ""))
launches = collections.defaultdict(list)
for _, (k, m) in per.items():
    launches[k].append(m)
unit = {r[mi]: r[h.index("Metric Unit")] for r in rows[1:]}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def dram(m):
    return sum(m.get(x, 0.0) * scale.get(unit.get(x, "byte"), 1.0) for x in ("dram__bytes_read.sum", "dram__bytes_write.sum"))


paired = a.mode == "pairs"
names = {"k_seed_exact": "k_seed_exact", "k_exact_sweep": "k_exact_sweep", "k_extend": "k_extend", "k_locate": "k_locate"}
out = {}
for k, ms in launches.items():
    base = k.split("::")[-1].replace("void ", "").strip()
    if base in names:
        out[names[base]] = dram(ms[-1])
    elif base.startswith("k_dp_fill") or base.startswith("k_dp_backtrace") or base.startswith("k_dp_gather"):
        key = "k_dp_fill" if "fill" in base else ("k_dp_backtrace" if "backtrace" in base else "k_dp_gather")
        if paired and len(ms) >= 2:
            out[key] = out.get(key, 0.0) + dram(ms[-2]); out[key + "[mate windows]"] = out.get(key + "[mate windows]", 0.0) + dram(ms[-1])
        else:
            out[key] = out.get(key, 0.0) + dram(ms[-1])
json.dump({"workload": {"batch": a.batch, "genome_mbp": a.genome_mbp, "nseq": a.nseq, "mode": a.mode},
           "source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum on `%s` (%s)" % (a.command or "python bench.py --chain-only --steps 1 --warmup 1", a.csv),
           "kernels": out}, open(a.out, "w"), indent=1)
print(json.dumps(out, indent=1))

"""Prints the interesting fields of a bench.py JSON line read from stdin (last line that parses)."""
import json
import sys

d = None
for ln in sys.stdin:
    ln = ln.strip()
    if ln.startswith("{"):
        try:
            d = json.loads(ln)
        except ValueError:
            pass
if d is None:
    sys.exit("no JSON line")
print(f"n_gpus {d['n_gpus']} value {d['value']:.4g} ms/step {d['ms_per_step']:.2f} e2e {d['e2e']['value']:.4g} ({d['e2e']['ms_per_step']:.2f} ms) "
      f"frac {d['roofline']['frac']:.3f}")
print("phases", {k: round(v, 2) for k, v in d["phases_ms_per_step"].items()})
for k in ("parity_vs_oracle", "parity_vs_1gpu", "forest_digest_same_on_all_ranks", "comm_bytes_per_step"):
    if k in d:
        print(k, d[k])
print("ms_each_step", d.get("ms_each_step"))
for k in ("configs2", "configs3", "predict"):
    if k in d:
        print(k, json.dumps(d[k]))

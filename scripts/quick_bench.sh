#!/bin/bash
# quick_bench.sh — short bench lines for several workloads (prints a compact summary per line)
fmt='import sys,json
for l in sys.stdin:
    l=l.strip()
    if not l.startswith("{"): print(l); continue
    d=json.loads(l); print({k:(round(d[k],3) if isinstance(d[k],float) else d[k]) for k in ("value","ms_per_step","gteps_min","gteps_max","levels_mean")}, "frac",round(d["roofline"]["frac"],3), "e2e",round(d["e2e"]["value"],2))'
for args in "$@"; do
  echo "== $args"
  timeout 150 python bench.py $args --no-cpu-baseline 2>&1 | tail -1 | python -c "$fmt"
done

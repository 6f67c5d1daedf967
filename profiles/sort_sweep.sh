#!/bin/bash
# Usage: profiles/sort_sweep.sh lib1.so lib2.so ...   one bench_sort.py summary line per library and table size
for lib in "$@"; do
  echo "== $lib"
  python bench_sort.py --no-cpu --lib "$lib" 2>&1 | python -c '
import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        print(d["config"]["rows"], "rows", round(d["ms_per_step"], 4), "ms", round(d["value"] / 1e9, 2), "G rows/s  frac", round(d["roofline"]["frac"], 3))
    else:
        print(l.rstrip())
'
done

#!/bin/bash
# rows x lanes sweep of the headline pass (bench.py --no-extras --no-cpu-baseline): usage tools/sweep_rows_lanes.sh OUT "rows lanes" ...
out=$1; shift
: > "$out"
for cfg in "$@"; do
  set -- $cfg
  line=$(timeout 200 python bench.py --no-extras --no-cpu-baseline --rows $1 --lanes $2 2>/dev/null | tail -n 1)
  echo "rows=$1 lanes=$2 $(echo "$line" | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("value %.0f e2e %.0f ms %.1f clocks %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]))')" >> "$out"
done

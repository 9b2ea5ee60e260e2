#!/bin/bash
# A/B timing of environment switches on the real decode step (tools/step_bench.py): usage tools/ab_env.sh OUT "VAR=a VAR2=b" ...
out=$1; shift
: > "$out"
for cfg in "$@"; do
  for shape in "15 256" "44 89" "1024 60"; do
    echo "== $cfg | B S = $shape" >> "$out"
    env $cfg timeout 120 python tools/step_bench.py $shape 1 fp16 2>&1 | tail -n 1 >> "$out"
  done
done

#!/bin/bash
# GPU check-in: parity tests, smoke, bench, launch list.  Usage: tools_gpu.sh [bench sentences]
N=${1:-4096}
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
run() { name=$1; shift; timeout ${TMO:-300} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" | tee -a gpurun_out/summary.txt; tail -n ${TAILN:-6} gpurun_out/$name.log; }
PT="python -m pytest -q -m gpu --timeout 200 -p no:cacheprovider -x"
run tests_kernels $PT tests/test_gpu_kernels.py
run tests_decode  $PT tests/test_gpu_decode.py -s
run smoke     python -c "import __graft_entry__ as g; g.smoke()"
run bench     python bench.py --sentences $N --steps 2 --warmup 3
CMD="python bench.py --sentences 64 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu launches exit=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/summary.txt

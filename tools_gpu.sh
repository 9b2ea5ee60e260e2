#!/bin/bash
# GPU check-in (run under gpurun): parity tests, smoke, bench, launch list of a 4096-row search step.
# Usage: tools_gpu.sh            logs land in gpurun_out/
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
run() { name=$1; shift; timeout ${TMO:-1800} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" | tee -a gpurun_out/summary.txt; tail -n ${TAILN:-4} gpurun_out/$name.log; }
run tests   python -m pytest tests -m gpu -q --timeout 1200 -p no:cacheprovider
run smoke   python -c "import __graft_entry__ as g; g.smoke()"
run bench   python bench.py
run kernels python tools/kernel_bench.py 1024 40 30
CMD="python tools/step_bench.py 128 40 8 fp16"
timeout 600 $CMD > gpurun_out/plain_step.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 4100 -c 140 --csv --log-file gpurun_out/launches_step_4096rows.csv $CMD > gpurun_out/ncu_step.log 2>&1
echo "ncu launch list exit=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/summary.txt

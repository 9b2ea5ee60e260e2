#!/bin/bash
# First bring-up on the GPU box: each group in its own process so one CUDA fault cannot poison the rest.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; timeout 200 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" | tee -a gpurun_out/summary.txt; tail -n 25 gpurun_out/$name.log; }
PT="python -m pytest -q -m gpu --timeout 120 -p no:cacheprovider"
run k_basic   $PT tests/test_gpu_kernels.py -k "f32_cuda_cores or embed or attention"
run k_beam    $PT tests/test_gpu_kernels.py -k "beam_step"
run k_simt16  $PT tests/test_gpu_kernels.py -k "linear_bf16 and simt"
run k_tc      $PT tests/test_gpu_kernels.py -k "tcgen05"
run d_fp32    $PT tests/test_gpu_decode.py -k "fp32 or list_wrapped"
run d_16s     $PT tests/test_gpu_decode.py -s -k "teacher_forced and simt"
run d_16t     $PT tests/test_gpu_decode.py -s -k "teacher_forced and tcgen05"
run d_16x     $PT tests/test_gpu_decode.py -s -k "six_layer or free_running"
run smoke     python -c "import __graft_entry__ as g; g.smoke()"
run bench     python bench.py --sentences 512 --steps 1 --warmup 1 --cpu-sentences 2
run bench16   python bench.py --sentences 512 --steps 1 --warmup 1 --no-cpu-baseline --dtype fp16
cat gpurun_out/summary.txt
CMD="python bench.py --sentences 64 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 700 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu launches exit=$?" | tee -a gpurun_out/summary.txt

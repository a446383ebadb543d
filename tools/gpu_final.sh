#!/bin/bash
# Round-end evidence: whole GPU suite, bench line, per-kernel roofline table, other BASELINE configs, ncu launch list.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
( time python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
python tools/fine_kernels.py > gpurun_out/fine_kernels.log 2>&1
python tools/measure.py c3 c4 planes > gpurun_out/measure.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ls -la gpurun_out

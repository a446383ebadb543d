#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the per-kernel roofline table, the other
# BASELINE configs, then (after each command exited without ncu) the ncu launch list and --set full captures.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
python tools/fine_kernels.py > gpurun_out/fine_kernels.log 2>&1
python tools/measure.py c3 c4 planes > gpurun_out/measure.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:^k -f -o gpurun_out/fine2d_full \
    python tools/fine_kernels.py --only 2d --reps 1 --warm 0 > gpurun_out/ncu_full2d.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:^k -f -o gpurun_out/fine3d_full \
    python tools/fine_kernels.py --only 3d --n3 257 --reps 1 --warm 0 > gpurun_out/ncu_full3d.log 2>&1
ls -la gpurun_out

#!/bin/bash
# quick GPU check: kernel + solver parity tests, the per-kernel table, other-config timings
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels2d.py tests/test_gpu_solver2d.py tests/test_gpu_kernels3d.py tests/test_gpu_solver3d.py -x -q 2>&1 | tail -15 > gpurun_out/pytest_q.log
python tools/fine_kernels.py $FK_ARGS > gpurun_out/fine_kernels_q.log 2>&1
python tools/measure.py $MEASURE_ARGS > gpurun_out/measure_q.log 2>&1
if [ -n "$NCU_K" ]; then
  ncu --set full --clock-control none --import-source on -k "regex:$NCU_K" -f -o gpurun_out/ncu_q \
      python tools/fine_kernels.py --reps 1 --warm 0 $NCU_FK_ARGS > gpurun_out/ncu_q.log 2>&1
fi

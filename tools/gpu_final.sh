#!/bin/bash
# Round-end evidence: whole GPU suite, smoke, bench line, per-kernel roofline table, ncu launch list.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
( time timeout 600 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1
( timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python tools/fine_kernels.py > gpurun_out/fine_kernels.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ls -la gpurun_out

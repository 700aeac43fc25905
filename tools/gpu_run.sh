#!/bin/bash
# usage: tools/gpu_run.sh TAG   (run on the GPU box via gpurun; outputs under gpurun_out/)
TAG=${1:-r1l}
set -x
nvidia-smi -L
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_$TAG.log
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_check_$TAG.json 2> gpurun_out/bench_check_$TAG.err
python bench.py --steps 20 --warmup 3 --mode solve_check --no-cpu-baseline > gpurun_out/bench_solve_$TAG.json 2> gpurun_out/bench_solve_$TAG.err
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_check_rs -s 3 -c 1 -o gpurun_out/prof_check_$TAG python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2_$TAG.log 2>&1
tail -3 gpurun_out/pytest_$TAG.log; cat gpurun_out/bench_check_$TAG.json; cat gpurun_out/bench_solve_$TAG.json; tail -3 gpurun_out/bench_check_$TAG.err

#!/bin/bash
# round 2, GPU run 1: parity suite, A/B of the round-2 solver changes, short bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_1.log 2>&1; tail -5 gpurun_out/r2_gpu_tests_1.log
{
echo "== default"; python tools/time_tile.py
echo "== minb64=3"; MGB200_LIB=multigrid_b200/variants/libmgb200_minb64_3.so python tools/time_tile.py
for env in "MGB200_SHORT_SEGMENTS=1 MGB200_FUSE_ZERO_GUESS=1" "MGB200_SHORT_SEGMENTS=0 MGB200_FUSE_ZERO_GUESS=1" "MGB200_SHORT_SEGMENTS=1 MGB200_FUSE_ZERO_GUESS=0" "MGB200_SHORT_SEGMENTS=0 MGB200_FUSE_ZERO_GUESS=0"; do
  echo "== solve $env"; env $env python tools/time_solve.py 2>&1 | tail -12
done
} > gpurun_out/r2_ab_1.log 2>&1
cat gpurun_out/r2_ab_1.log
python bench.py --steps 5 --warmup 3 --no-dg > gpurun_out/r2_bench_1.json 2> gpurun_out/r2_bench_1.err; tail -c 1500 gpurun_out/r2_bench_1.json; tail -5 gpurun_out/r2_bench_1.err

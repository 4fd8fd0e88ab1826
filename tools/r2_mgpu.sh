#!/bin/bash
# round 2, multi-GPU run: the 17 parity cases of tests/test_multi_gpu.py (2, 4 and 8 ranks against the global oracle)
# and the weak-scaling bench lines at N = 8, 4, 2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_multi_gpu_tests.log
python -m pytest tests/test_multi_gpu.py -m gpu -q -rA >> gpurun_out/r2_multi_gpu_tests.log 2>&1
tail -25 gpurun_out/r2_multi_gpu_tests.log
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $n --steps 6 --warmup 3 > gpurun_out/r2_bench_${n}gpu.json 2> gpurun_out/r2_bench_${n}gpu.err
  python -c "import json,sys; d=json.loads(open('gpurun_out/r2_bench_${n}gpu.json').read().strip().splitlines()[-1]); print($n, d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['cg_iterations'], d['l2_error'])" || tail -5 gpurun_out/r2_bench_${n}gpu.err
done
python bench.py --gpus 1 --steps 6 --warmup 3 --no-dg --no-cpu-baseline > gpurun_out/r2_bench_1gpu_scale.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/r2_bench_1gpu_scale.json').read().strip().splitlines()[-1]); print(1, d['ms_per_step'], d['value'], d['e2e']['ms_per_step'])"

#!/bin/bash
# 8-GPU box: multi-GPU tests on real NCCL, then STRONG scaling of the fixed jobs (bench.py under torchrun, as the driver runs it)
O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q > $O/pytest_multi_r02.log 2>&1; echo "pytest multi rc=$?" | tee -a $O/pytest_multi_r02.log; tail -3 $O/pytest_multi_r02.log
run() {  # name gpus extra-args...
  name=$1; n=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --gpus $n --no-cpu-baseline "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"
  python -c "
import json
d=json.load(open('$O/$name.json'))
print('$name', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],1), d['config']['parallelism'][:40], d.get('mgpu_check'))"
}
run bench_mg8_r02 8 --steps 3 --warmup 2
run bench_mg8_tiles1 8 --steps 3 --warmup 2 --tiles 1 --no-extras
run bench_mg8_tiles2 8 --steps 3 --warmup 2 --tiles 2 --no-extras
run bench_mg4_r02 4 --steps 3 --warmup 2 --no-extras
run bench_C3_mg8_r02 8 --config C3 --steps 3 --warmup 2 --no-extras

#!/bin/bash
# 2 GPUs: multi-GPU tests, bench at N=2 (strong) and N=1 of the same short job
O=gpurun_out; mkdir -p $O
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $O/c8_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c8_pytest.log
tail -15 $O/c8_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus 2 --steps 2 --warmup 1 > $O/c8_bench2.json 2> $O/c8_bench2_err.log; echo "bench2 rc=$?"
tail -5 $O/c8_bench2_err.log; python -c "
import json; d=json.load(open('$O/c8_bench2.json')); print(d['value'], d['e2e']['value'], d['config']['parallelism'], d['mgpu_check'], d['ms_per_step'], d['rank0_last_render_ms'])"

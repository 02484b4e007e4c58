#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/c7_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c7_pytest.log
tail -8 $O/c7_pytest.log
( time timeout 900 python bench.py --steps 2 --warmup 1 > $O/c7_bench.json 2> $O/c7_bench_err.log ) 2>&1 | grep real
tail -5 $O/c7_bench_err.log; head -c 6000 $O/c7_bench.json
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/c7_bench_ref.json 2>> $O/c7_bench_err.log ) 2>&1 | grep real
cat $O/c7_bench_ref.json

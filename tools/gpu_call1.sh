#!/bin/bash
# round-2 call 1: microbenchmark, GPU tests (all, not -x), A/B of the FHFMA slab test against the round-1 library
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/c1_smi.txt
./build/fhfma_mb > $O/c1_mb.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > $O/c1_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c1_pytest.log
for v in base new; do
  if [ $v = base ]; then export WGPT_LIB=$PWD/build/libwgpt_base.so; else unset WGPT_LIB; fi
  timeout 300 python tools/stage_bench.py C2,C3,C4 64 > $O/c1_stages_$v.txt 2>&1
  timeout 120 python tools/visit_stats.py C2,C4 2 > $O/c1_visits_$v.txt 2>&1
done
unset WGPT_LIB
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 20 -c 1 -f -o $O/extend_C4_c1 python tools/prof_run.py C4 64 > $O/c1_ncu.log 2>&1
tail -5 $O/c1_pytest.log; cat $O/c1_mb.txt $O/c1_stages_*.txt $O/c1_visits_*.txt

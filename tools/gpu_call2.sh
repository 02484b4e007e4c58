#!/bin/bash
# round-2 call 2: variants of the FHFMA extend kernel against the round-1 library; per-iteration dump; tests
O=gpurun_out; mkdir -p $O
for v in base b b80 bp; do
  export WGPT_LIB=$PWD/build/libwgpt_$v.so
  timeout 300 python tools/stage_bench.py C2,C4 64 > $O/c2_stages_$v.txt 2>&1
done
export WGPT_LIB=$PWD/build/libwgpt_b.so
rm -f $O/c2_iters_b.txt
WGPT_PROF_DUMP=$O/c2_iters_b.txt timeout 200 python tools/stage_bench.py C4 64 > /dev/null 2>&1
unset WGPT_LIB
timeout 1200 python -m pytest tests -m gpu -q > $O/c2_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c2_pytest.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 20 -c 1 -f -o $O/extend_C4_c2 python tools/prof_run.py C4 64 > $O/c2_ncu.log 2>&1
tail -3 $O/c2_pytest.log; cat $O/c2_stages_*.txt

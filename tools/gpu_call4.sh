#!/bin/bash
O=gpurun_out; mkdir -p $O
for v in base c c80; do
  export WGPT_LIB=$PWD/build/libwgpt_$v.so
  timeout 300 python tools/stage_bench.py C2,C3,C4 64 > $O/c4_stages_$v.txt 2>&1
done
unset WGPT_LIB
timeout 1200 python -m pytest tests -m gpu -q -x > $O/c4_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c4_pytest.log
rm -f $O/c4_iters.txt
WGPT_PROF_DUMP=$O/c4_iters.txt timeout 200 python tools/stage_bench.py C4 64 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 20 -c 1 -f -o $O/extend_C4_c4 python tools/prof_run.py C4 64 > $O/c4_ncu.log 2>&1
tail -3 $O/c4_pytest.log; cat $O/c4_stages_*.txt

#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/c6_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c6_pytest.log
tail -4 $O/c6_pytest.log
for v in base a12 a8 a8r64 a12r64 s12r64; do
  export WGPT_LIB=$PWD/build/libwgpt_$v.so
  timeout 300 python tools/stage_bench.py C2,C3,C4 64 > $O/c6_stages_$v.txt 2>&1
  echo $v; cat $O/c6_stages_$v.txt
done
export WGPT_LIB=$PWD/build/libwgpt_a12.so
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 20 -c 1 -f -o $O/extend_C4_c6 python tools/prof_run.py C4 64 > $O/c6_ncu.log 2>&1

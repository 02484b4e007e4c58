#!/bin/bash
O=gpurun_out; mkdir -p $O
for v in base s8 s12 s16; do
  export WGPT_LIB=$PWD/build/libwgpt_$v.so
  timeout 300 python tools/stage_bench.py C2,C3,C4 64 > $O/c5_stages_$v.txt 2>&1
done
export WGPT_LIB=$PWD/build/libwgpt_s12.so
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 20 -c 1 -f -o $O/extend_C4_c5 python tools/prof_run.py C4 64 > $O/c5_ncu.log 2>&1
for v in base s8 s12 s16; do echo $v; cat $O/c5_stages_$v.txt; done

#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/c3_pytest.log 2>&1; echo "pytest rc=$?" >> $O/c3_pytest.log
tail -15 $O/c3_pytest.log
export WGPT_LIB=$PWD/build/libwgpt_dbg.so
timeout 300 python tools/visit_stats.py C4 24 > $O/c3_longwalks.txt 2>&1
grep -c "long walk" $O/c3_longwalks.txt; grep "long walk" $O/c3_longwalks.txt | sort -t' ' -k3 -n -r | head -40; tail -2 $O/c3_longwalks.txt

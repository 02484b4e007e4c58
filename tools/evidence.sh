#!/bin/bash
# One-box evidence refresh (run under gpurun): GPU tests, both bench arms, the ncu launch list of the
# bench command, and one `--set full` capture per (stage, config) at a steady-state launch.
# Everything lands in gpurun_out/; tools/evidence_collect.sh turns it into profiles/*_rNN.*
R=${1:-r01}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$R.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$R.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_$R.json 2> $O/bench_err.log; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$R.json 2>> $O/bench_err.log; echo "bench ref rc=$?"
timeout 200 python bench.py --config C1 --steps 5 --warmup 3 > $O/bench_C1_$R.json 2>> $O/bench_err.log
timeout 300 python bench.py --config C3 --steps 3 --warmup 3 --cpu-baseline-frames 4 > $O/bench_C3_$R.json 2>> $O/bench_err.log
timeout 500 python bench.py --config C4 --steps 2 --warmup 3 --cpu-baseline-frames 4 > $O/bench_C4_$R.json 2>> $O/bench_err.log
timeout 120 python tools/visit_stats.py C1,C2,C3,C4,C5 2 > $O/visits_$R.txt 2>&1
timeout 120 python tools/k4_bench.py > $O/k4_$R.txt 2>&1
timeout 300 python tools/stage_bench.py C1,C2,C3,C4,C5 64 > $O/stages_$R.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches_$R.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_bench.log 2>&1; echo "ncu list rc=$?"
for cfg in C2 C4; do
  for st in extend shade; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_$st -s 20 -c 1 -f -o $O/${st}_${cfg}_$R \
        python tools/prof_run.py $cfg 64 > $O/ncu_${st}${cfg}.log 2>&1; echo "ncu $st $cfg rc=$?"
  done
done
ls -la $O | tail -20

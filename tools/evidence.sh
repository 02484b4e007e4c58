#!/bin/bash
# One-box evidence refresh (run under gpurun): GPU tests, one `--set full` ncu capture per (stage, config) at a
# steady-state launch, profiles/traffic.json from those captures (stamped with the kernel-source hash, written ON the
# box so that the bench line of the same run can quote it), both bench arms, the ncu launch list of the bench command.
# Everything lands in gpurun_out/; tools/evidence_collect.py turns it into profiles/*_rNN.*
R=${1:-r02}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu_$R.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$R.log
for cfg in C4 C2; do
  for st in extend shade; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_$st -s 20 -c 1 -f -o $O/${st}_${cfg}_$R \
        python tools/prof_run.py $cfg 64 > $O/ncu_${st}${cfg}.log 2>&1; echo "ncu $st $cfg rc=$?"
  done
done
python tools/evidence_collect.py $R --traffic-only > $O/traffic_collect.log 2>&1; echo "traffic rc=$?"; cp profiles/traffic.json $O/traffic.json
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench_$R.json 2> $O/bench_err.log; echo "bench rc=$?"
[ -n "$LITE" ] && { ls -la $O | tail -30; exit 0; }   # LITE=1: tests, captures, traffic stamp and the default bench only (a host-side change)
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$R.json 2>> $O/bench_err.log; echo "bench ref rc=$?"
timeout 200 python bench.py --config C1 --steps 5 --warmup 3 --no-extras --cpu-baseline-frames 8 > $O/bench_C1_$R.json 2>> $O/bench_err.log
timeout 200 python bench.py --config C2 --steps 5 --warmup 3 --no-extras --cpu-baseline-frames 8 > $O/bench_C2_$R.json 2>> $O/bench_err.log
timeout 300 python bench.py --config C3 --steps 3 --warmup 3 --no-extras --cpu-baseline-frames 4 > $O/bench_C3_$R.json 2>> $O/bench_err.log
timeout 600 python bench.py --config C5 --steps 1 --warmup 1 --no-extras --cpu-baseline-frames 1 > $O/bench_C5_$R.json 2>> $O/bench_err.log
timeout 120 python tools/visit_stats.py C1,C2,C3,C4,C5 2 > $O/visits_$R.txt 2>&1
timeout 120 python tools/k4_bench.py > $O/k4_$R.txt 2>&1
timeout 300 python tools/stage_bench.py C1,C2,C3,C4,C5 64 > $O/stages_$R.txt 2>&1
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/launches_$R.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $O/ncu_bench.log 2>&1; echo "ncu list rc=$?"
ls -la $O | tail -30

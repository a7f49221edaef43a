#!/bin/bash
# Run on the GPU box (gpurun): launch list + full ncu capture of the dominant kernel for the bench workload.
R=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_plain_$R.json 2> gpurun_out/bench_plain_$R.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c2_$R.csv \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launch_$R.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_score_fast -s 6 -c 1 -o gpurun_out/prof_c2_score_$R \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_full_c2_$R.log 2>&1
python bench.py --config C3 --steps 2 --warmup 3 > gpurun_out/bench_plain_c3_$R.json 2>/dev/null &&
ncu --set full --clock-control none --import-source on -k regex:k_score_fast -s 3 -c 1 -o gpurun_out/prof_c3_score_$R \
    python bench.py --config C3 --steps 2 --warmup 3 > gpurun_out/ncu_full_c3_$R.log 2>&1
ls -la gpurun_out

#!/bin/bash
# Run on the GPU box (gpurun): launch lists + full ncu captures of the kernels of the hypothesis stage (round tag $1).
#   C3 (default bench workload, k_score_fast ring kernel) and C2 (k_score_resident); each ncu command follows its own plain run.
R=${1:-r2}
mkdir -p gpurun_out
B3="python bench.py --steps 2 --warmup 3 --no-cpu --no-secondary"
B2="python bench.py --config C2 --steps 3 --warmup 3 --no-cpu"
$B3 > gpurun_out/plain_c3_$R.json 2> gpurun_out/plain_c3_$R.err || exit 1
$B2 > gpurun_out/plain_c2_$R.json 2> gpurun_out/plain_c2_$R.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3_$R.csv $B3 > gpurun_out/ncu_launch_c3_$R.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c2_$R.csv $B2 > gpurun_out/ncu_launch_c2_$R.log 2>&1
for k in k_score_fast k_sample_solve k_tail; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o gpurun_out/prof_c3_${k}_$R $B3 > gpurun_out/ncu_full_c3_${k}_$R.log 2>&1
done
for k in k_score_resident k_sample_solve k_tail; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -o gpurun_out/prof_c2_${k}_$R $B2 > gpurun_out/ncu_full_c2_${k}_$R.log 2>&1
done
# the non-minimal fit and the local-optimisation pass (gcr_fit_score_models, gcr_lo_round): launch list + one full capture of k_fit_rows
REPS=3 python tools/time_fit.py C1 > gpurun_out/plain_fit_$R.txt 2>&1 || exit 1
REPS=3 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_fit_$R.csv python tools/time_fit.py C1 > gpurun_out/ncu_launch_fit_$R.log 2>&1
REPS=3 ncu --set full --clock-control none --import-source on -k regex:k_fit_rows -s 10 -c 1 -o gpurun_out/prof_fit_k_fit_rows_$R python tools/time_fit.py C1 > gpurun_out/ncu_full_fit_$R.log 2>&1
tools/ubench --json gpurun_out/ubench_b200.json > gpurun_out/ubench_b200.txt 2>&1
ls -la gpurun_out | tail -30

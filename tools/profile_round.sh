#!/bin/bash
# Run on the GPU box (gpurun): launch list + full ncu captures of the three kernels of the hypothesis stage.
R=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_plain_$R.json 2> gpurun_out/bench_plain_$R.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c2_$R.csv \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launch_$R.log 2>&1
for k in k_score_resident k_sample_solve k_tail; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -o gpurun_out/prof_c2_${k}_$R \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_full_c2_${k}_$R.log 2>&1
done
python bench.py --config C3 --steps 3 --warmup 3 > gpurun_out/bench_plain_c3_$R.json 2>/dev/null &&
ncu --set full --clock-control none --import-source on -k regex:k_score_fast -s 3 -c 1 -o gpurun_out/prof_c3_k_score_fast_$R \
    python bench.py --config C3 --steps 2 --warmup 3 > gpurun_out/ncu_full_c3_$R.log 2>&1
ls -la gpurun_out

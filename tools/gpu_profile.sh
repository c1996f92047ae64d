#!/bin/bash
# One GPU visit: plain runs first (each must exit 0), then the ncu passes. Outputs under gpurun_out/r02_*.
set -u
O=gpurun_out
python tools/ncu_targets.py > $O/r02_targets.log 2>&1 || { tail -5 $O/r02_targets.log; exit 1; }
python bench.py --steps 2 --warmup 3 --no-microbench --no-cpu-baseline --pipeline-depth 1 > $O/r02_bench_depth1.json 2> $O/r02_bench_depth1.err || { tail -5 $O/r02_bench_depth1.err; exit 1; }
NCU="ncu --clock-control none"
$NCU --metrics gpu__time_duration.sum -k regex:'gmm_|eb_' -c 2000 --csv --log-file $O/r02_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-microbench --no-cpu-baseline --pipeline-depth 1 > $O/r02_ncu_bench.log 2>&1
full() { # name regex count which
  $NCU --set full --import-source on -k regex:"$2" -c $3 -f -o $O/r02_$1 python tools/ncu_targets.py $4 > $O/r02_ncu_$1.log 2>&1 || tail -3 $O/r02_ncu_$1.log
}
full sf 'gmm_startfreq_kmajor4' 3 sf
full rows 'gmm_cdf_window_batched' 6 dec
full pop 'gmm_pop_rows' 4 dec
full push 'gmm_push_kernel|gmm_startfreq_batched' 2 push
full lik 'gmm_likelihood' 4 lik
full il 'gmm_ileave' 2 il
full glue 'gmm_prologue|gmm_epilogue|eb_records' 4 glue
ls -la $O/r02_*.ncu-rep

#!/bin/bash
# One GPU box visit: tests, bench (both arms), ncu launch list, ncu --set full of the four kernels.
# usage: tools/gpu_round.sh TAG   (outputs under gpurun_out/TAG_*)
T=${1:-sXX}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; tail -1 $O/${T}_pytest.log
timeout 900 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -c 600 $O/${T}_bench.json
timeout 900 python bench.py --impl reference > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; tail -c 400 $O/${T}_bench_ref.json
B="python bench.py --no-microbench --no-cpu-baseline --steps 2 --warmup 3"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gmm_ -c 4000 --csv --log-file $O/${T}_launches.csv $B > $O/${T}_ncu_launch.log 2>&1
for K in "pop:gmm_pop_rows_kernel:10" "rows:gmm_cdf_rows_batched_kernel:10" "sfb:gmm_startfreq_batched_kernel:3" "push:gmm_push_kernel:3" "sf:gmm_startfreq_scalar:200"; do
  IFS=: read name re skip <<< "$K"
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:$re -s $skip -c 1 -o $O/${T}_$name -f $B > $O/${T}_ncu_$name.log 2>&1
done
ls -la $O/${T}_*

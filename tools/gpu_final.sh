#!/bin/bash
# Final visit of a round: tests, smoke, default bench (both arms), refreshed ncu captures of the kernels that changed.
set -u
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/r02_final_tests.log; cat $O/r02_final_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-200
timeout 600 python bench.py > $O/r02_final_bench.json 2> $O/r02_final_bench.err || tail -3 $O/r02_final_bench.err
timeout 400 python bench.py --impl reference > $O/r02_final_bench_ref.json 2> /dev/null
NCU="ncu --clock-control none"
full() { $NCU --set full --import-source on -k regex:"$2" -c $3 -f -o $O/r02b_$1 python tools/ncu_targets.py $4 > $O/r02b_ncu_$1.log 2>&1 || tail -3 $O/r02b_ncu_$1.log; }
if [ "${NCU:-0}" = "1" ]; then
full rows 'gmm_cdf_window_batched' 24 dec
full lik 'eb_likelihood|gmm_likelihood' 4 lik
fi
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_final_bench.json"))
for k in ("value","ms_per_step","value_one_batch_in_flight","e2e","latency_single_image_ms","gpu_launches","model_codec","model_codec_two_lanes","cpu_baseline","cpu_baseline_full_call"): print(k, json.dumps(d.get(k))[:400])
print(json.dumps(d["microbench"])[:900])
PY

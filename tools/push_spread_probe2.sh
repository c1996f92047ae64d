#!/bin/bash
O=gpurun_out
run() {
  tag=$1; shift; extra=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu-baseline --no-microbench $extra --steps 12 --warmup 3 > $O/push_$tag.json 2> $O/push_$tag.err || tail -3 $O/push_$tag.err
  python - "$tag" <<'PY'
import json, sys
d = json.load(open(f"gpurun_out/push_{sys.argv[1]}.json"))
print(sys.argv[1], {k: d.get(k) for k in ("value", "value_one_batch_in_flight", "encode_ms_per_step")}, "e2e", (d.get("e2e") or {}).get("value"), flush=True)
PY
}
run s28_w1  --no-e2e FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=28
run s40_w1  --no-e2e FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=40
run s75_w1  --no-e2e FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=75
run s56_w1_e2e  "" FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=56
run s56_dflt_e2e "" FLASHGMM_PUSH_SPREAD_KB=56
run s56_w3_e2e  "" FLASHGMM_WORKERS=3 FLASHGMM_PUSH_SPREAD_KB=56

#!/bin/bash
# Dev probe: device push with the wait kernel (default) vs spread by reserved shared memory (FLASHGMM_PUSH_SPREAD_KB).
# FLASHGMM_WORKERS=1 reproduces the 8-GPU box's condition (the cost model then sends the long streams to the push kernel).
O=gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu-baseline --no-microbench --no-e2e --steps 12 --warmup 3 > $O/push_$tag.json 2> $O/push_$tag.err || tail -3 $O/push_$tag.err
  python - "$tag" <<'PY'
import json, sys
d = json.load(open(f"gpurun_out/push_{sys.argv[1]}.json"))
print(sys.argv[1], {k: d.get(k) for k in ("value", "value_one_batch_in_flight", "encode_ms_per_step", "decode_ms_per_step")}, flush=True)
PY
}
run wait      FLASHGMM_WORKERS=1
run spread100 FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=100
run spread56  FLASHGMM_WORKERS=1 FLASHGMM_PUSH_SPREAD_KB=56
run spread100_forced FLASHGMM_PUSH=device FLASHGMM_PUSH_SPREAD_KB=100
run wait_forced      FLASHGMM_PUSH=device
FLASHGMM_PUSH_SPREAD_KB=100 timeout 300 python -m pytest tests/test_push_gpu.py -x -q -m gpu 2>&1 | tail -3

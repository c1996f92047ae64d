#!/bin/bash
O=gpurun_out
run() {
  tag=$1; shift; extra=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu-baseline --no-microbench --no-e2e $extra > $O/lane_$tag.json 2> $O/lane_$tag.err || tail -3 $O/lane_$tag.err
  python - "$tag" <<'PY'
import json, sys
d = json.load(open(f"gpurun_out/lane_{sys.argv[1]}.json"))
print(sys.argv[1], {k: d.get(k) for k in ("value", "value_one_batch_in_flight", "encode_ms_per_step", "gpu_launches", "steps", "warmup")}, flush=True)
PY
}
run off     "" FLASHGMM_PUSH_BUSY_LANES=0
run auto    ""
run forced  "" FLASHGMM_PUSH=device
run auto_w12 "--warmup 12"
run off_w12  "--warmup 12" FLASHGMM_PUSH_BUSY_LANES=0

#!/bin/bash
# short GPU check used while tuning kernels: parity tests of the render path, then the per-stage times of the bench
#   tools/quick_bench.sh <tag> [extra bench args]
TAG=${1:-qb}; shift
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py tests/test_gpu_grad_golden.py tests/test_gpu_guards.py -q 2>&1 | tail -3
python bench.py --no-cpu-baseline --no-extras "$@" > gpurun_out/${TAG}.json 2> gpurun_out/${TAG}.err || tail -5 gpurun_out/${TAG}.err
python - "$TAG" <<'PY'
import json, sys
d=json.load(open(f"gpurun_out/{sys.argv[1]}.json"))
print("fwd+bwd %.1f  fwd %.1f  e2e %s  regions %s" % (d["value"], d["forward_only"]["value"], d["e2e"] and round(d["e2e"]["value"],1), d["clocks"]["regions_ms_per_step"]))
print([(s["stage"], round(s["ms_per_frame"],4)) for s in d["stages"]])
PY

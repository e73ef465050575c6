#!/bin/bash
# short GPU check used while tuning kernels: parity tests of the render path, then the per-stage times of the bench
python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -q -x 2>&1 | tail -2
python bench.py --no-cpu-baseline ${QB_ARGS:-} > gpurun_out/qb.json 2> gpurun_out/qb.err || tail -5 gpurun_out/qb.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/qb.json"))
print("fwd+bwd %.1f  fwd %.1f  e2e %s" % (d["value"], d["forward_only"]["value"], d["e2e"] and round(d["e2e"]["value"],1)))
print([(s["stage"], round(s["ms_per_frame"],4)) for s in d["stages"]])
PY

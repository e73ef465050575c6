#!/bin/bash
# A/B of the priority lanes, the number of view streams and the CTA sizes of the stages before compositing:
#   tools/ab_lanes.sh "L S PRE SC" ...   -> one line per configuration
for cfg in "$@"; do
  set -- $cfg; L=$1; S=$2; PRE=$3; SC=$4
  GSR_PRIORITY_LANES=$L GSR_PRE_THREADS=$PRE GSR_SCATTER_THREADS=$SC python bench.py --no-cpu-baseline --no-extras --streams $S > gpurun_out/ab.json 2>/dev/null
  python - "$cfg" <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab.json"))
st = {s["stage"]: round(s["ms_per_frame"], 4) for s in d["stages"]}
print("lanes/streams/pre/scatter", sys.argv[1], "fwd+bwd %.1f fwd %.1f e2e %.1f" % (d["value"], d["forward_only"]["value"], d["e2e"]["value"]),
      "pre %.4f scatter %.4f" % (st["preprocess"], st["duplicate"]), "e2e host ms/step", d["e2e"].get("host_ms_per_step_before_loss_read"))
PY
done

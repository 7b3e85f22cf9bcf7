#!/bin/bash
# bench.py at several launch-group sizes of the device-pointer path (SPB_GROUP_ROWS: a sizing knob, results unchanged)
for G in "$@"; do
  SPB_GROUP_ROWS=$G timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null > /tmp/sweep_$G.json
  python - "$G" <<'PY'
import json, sys
g = sys.argv[1]
d = json.loads(open(f"/tmp/sweep_{g}.json").read().splitlines()[-1])
print("group_rows", g, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), {k: round(v, 2) for k, v in d["roofline"]["kernel_ms_per_step"].items()},
      "MHz", d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
PY
done

#!/bin/bash
# bench.py with the utterance starts of the synthetic arena aligned to A samples (SPB_BENCH_ALIGN)
for A in "$@"; do
  SPB_BENCH_ALIGN=$A timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null > /tmp/sweepa_$A.json
  python - "$A" <<'PY'
import json, sys
a = sys.argv[1]
d = json.loads(open(f"/tmp/sweepa_{a}.json").read().splitlines()[-1])
print("align", a, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), {k: round(v, 2) for k, v in d["roofline"]["kernel_ms_per_step"].items()}, "MHz", d["clocks"]["sm_mhz"])
PY
done

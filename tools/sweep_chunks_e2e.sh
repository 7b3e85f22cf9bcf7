#!/bin/bash
# bench.py's end-to-end (host buffer) leg at several pipeline-chunk sizes of the host-pointer calls (SPB_CHUNK_ROWS)
for G in "$@"; do
  SPB_CHUNK_ROWS=$G timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null > /tmp/sweepc_$G.json
  python - "$G" <<'PY'
import json, sys
g = sys.argv[1]
d = json.loads(open(f"/tmp/sweepc_{g}.json").read().splitlines()[-1])
print("chunk_rows", g, "e2e", round(d["e2e"]["value"], 1), "f32 arena", round(d["e2e"].get("float32_arena_value", 0), 1), "value", round(d["value"], 1), "match", d["e2e"].get("matches_device_path"))
PY
done

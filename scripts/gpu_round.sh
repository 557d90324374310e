#!/bin/bash
# One GPU visit: smoke, parity tests, the bench line, a few throughput points.
mkdir -p gpurun_out
python __graft_entry__.py smoke 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
( time timeout 1500 python bench.py --steps 5 --warmup 3 2> gpurun_out/bench.err | grep "^{" > gpurun_out/bench.json ) 2>&1 | tail -3
tail -5 gpurun_out/bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
for k in ("value", "ms_per_step", "verified", "roofline", "e2e", "cpu_baseline", "gpu_launches", "clocks"):
    print(k, d.get(k))
print("other_mode", {k: d["other_mode"][k] for k in ("value", "e2e", "parse_kernel_ms")})
print("next_rows", d.get("next_rows"))
cr = d.get("config_rows") or {}
for k, v in cr.items():
    if k != "size_sweep":
        print(k, v)
sw = cr.get("size_sweep") or {}
for k, v in sw.items():
    print(k, v if not isinstance(v, dict) else {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items()})
PY
for f in 3 4; do echo "== fpw=$f"; SNAPPY_B200_LANE_FPW=$f timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1; done

#!/bin/bash
mkdir -p gpurun_out
for v in 40 104 168 232; do echo "== var=$v"; SNAPPY_B200_LANE_VAR=$v timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_lane | tail -1; done
timeout 900 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 2> gpurun_out/bench_n2.err | grep "^{" > gpurun_out/bench_n2.json
tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n2.json"))
for k in ("value", "ms_per_step", "n_gpus", "verified", "e2e", "sharded_stream_2gib", "gpu_launches", "config"):
    print(k, d.get(k))
PY

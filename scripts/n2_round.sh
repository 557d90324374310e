#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 2> gpurun_out/bench_n2.err | grep "^{" > gpurun_out/bench_n2.json
tail -3 gpurun_out/bench_n2.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n2.json"))
for k in ("value", "ms_per_step", "n_gpus", "verified", "e2e", "sharded_stream_2gib", "gpu_launches"):
    print(k, d.get(k))
PY

#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 2> gpurun_out/bench_n$N.err | grep "^{" > gpurun_out/bench_n$N.json
tail -4 gpurun_out/bench_n$N.err
python - $N <<'PY'
import json, sys
d = json.load(open("gpurun_out/bench_n%s.json" % sys.argv[1]))
for k in ("value", "ms_per_step", "n_gpus", "verified", "e2e", "sharded_stream_2gib", "gpu_launches", "clocks"):
    print(k, d.get(k))
print("other_mode", {k: d["other_mode"][k] for k in ("value", "e2e")})
PY

#!/bin/bash
for g in 32 64 128; do echo "== L2 fetch $g"
  SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact | tail -1
  SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode exact_pipe | tail -1
  SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --mode fast | tail -1
  SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --mib 1024 --reps 3 --data random | tail -1
  SNAPPY_B200_L2_FETCH=$g timeout 200 python scripts/run_one.py --batch 4096 --mib 1024 --reps 2 --decode | tail -2
done
python - <<'PY'
import ctypes, torch
torch.cuda.init()
rt = ctypes.CDLL("libcudart.so.12")
v = ctypes.c_size_t(0)
print("default cudaLimitMaxL2FetchGranularity:", rt.cudaDeviceGetLimit(ctypes.byref(v), 5), v.value)
PY

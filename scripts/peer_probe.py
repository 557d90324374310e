#!/usr/bin/env python
"""Time the peer-memory exchange alone: every rank sends a fixed payload to the root's archive.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/peer_probe.py [--mib 442]
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=float, default=442.6)
    ap.add_argument("--steps", type=int, default=8)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pkg = g.load_package()
    from compression_accelerator_b200 import multi
    ctx = pkg.Context(local)
    n = int(a.mib * (1 << 20)) + 13 * rank           # odd sizes: unaligned offsets
    payload = torch.randint(0, 255, (n,), dtype=torch.uint8, device=dev)
    size = torch.tensor([n], dtype=torch.int64, device=dev)
    pa = multi.PeerArchive(ctx, dev, n + 4096)
    torch.cuda.synchronize()
    step = 0
    for _ in range(2):
        pa.gather(payload, size, step)
        step += 1
    pa.wait(step - 1)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        pa.gather(payload, size, step)
        step += 1
    total, offs = pa.wait(step - 1)
    dt = (time.perf_counter() - t0) / a.steps
    ok = bool(torch.equal(pa.archive((step - 1) & 1)[offs[rank]:offs[rank] + n], payload)) if rank == 0 else None
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        moved = total - n                             # the root's own payload does not cross NVLink
        print("world %d: %.3f ms per exchange, %.1f GB/s into the root over NVLink (%.1f GB/s incl. local copy), root slice ok: %s"
              % (world, t.item() * 1e3, moved / t.item() / 1e9, total / t.item() / 1e9, ok), flush=True)
    dist.barrier()
    if rank != 0:
        pa.close()
    dist.barrier()
    pa.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Sharded single stream over NCCL on >= 2 GPUs: every rank compresses its run of whole fragments
with the CUDA path (header-less), the exchange step assembles the stream on rank 0, and the
result must be byte-identical to the oracle's one-shot stream.  Skipped on a 1-GPU box."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import corpus

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, q):
    for p in (HERE, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    import __graft_entry__ as g
    import oracle_lib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        pkg = g.load_package()
        from compression_accelerator_b200 import multi
        ctx = pkg.Context(rank)
        host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n_total)
        sh = multi.ShardedStream(n_total, dev)
        shard = torch.from_numpy(host[sh.lo:sh.hi].copy()).to(dev)
        stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
        torch.cuda.synchronize()
        payload, size = ctx.compress_device(shard, header=False)
        dst = torch.zeros(pkg.max_compressed_length(n_total) + 16, dtype=torch.uint8, device=dev) if rank == 0 else None
        torch.cuda.synchronize()      # the context's stream is non-blocking: order the allocations before it
        with torch.cuda.stream(stream):
            total = sh.assemble(payload, size, dst)
        torch.cuda.synchronize()
        if rank == 0:
            want = oracle_lib.checker().compress(host.tobytes())
            q.put((total == len(want), bytes(dst[:total].cpu().numpy()) == want, total))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [65536 * 37 + 4321, 1 << 26])
def test_sharded_stream_nccl(n_total):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    len_ok, bytes_ok, total = q.get(timeout=5)
    assert len_ok and bytes_ok, (n_total, total)


def _peer_worker(rank, world, port, q):
    for p in (HERE, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    import __graft_entry__ as g
    import oracle_lib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        pkg = g.load_package()
        from compression_accelerator_b200 import multi
        ctx = pkg.Context(rank)
        text = corpus.fixture("all-mtg-cards.txt")
        # every rank its own stream, different lengths, three pipelined steps with different inputs
        n = 3 * 65536 + 1000 * (rank + 1)
        cap = pkg.max_compressed_length(n)
        pa = multi.PeerArchive(ctx, dev, cap)
        inputs = [text[k * 100000 + rank * 7: k * 100000 + rank * 7 + n] for k in range(3)]
        srcs = [torch.frombuffer(bytearray(b), dtype=torch.uint8).to(dev) for b in inputs]
        bufs = [(torch.empty(cap, dtype=torch.uint8, device=dev), torch.zeros(1, dtype=torch.int64, device=dev))
                for _ in range(2)]
        torch.cuda.synchronize()
        chk = oracle_lib.checker()
        ok = True
        for k in range(3):
            pa.reuse(k - 2)
            ctx.compress_device(srcs[k], bufs[k & 1][0], bufs[k & 1][1])
            pa.gather(bufs[k & 1][0], bufs[k & 1][1], k)
            if k >= 1:                                   # check step k-1 while step k is in flight
                total, offs = pa.wait(k - 1)
                want = chk.compress(inputs[k - 1])
                size = (offs[rank + 1] if rank + 1 < world else total) - offs[rank]
                got = bytes(pa.archive((k - 1) & 1)[offs[rank]:offs[rank] + size].cpu().numpy())
                ok = ok and got == want
        total, offs = pa.wait(2)
        want = chk.compress(inputs[2])
        got = bytes(pa.archive(0)[offs[rank]:offs[rank] + len(want)].cpu().numpy())
        ok = ok and got == want and (offs[rank + 1] if rank + 1 < world else total) - offs[rank] == len(want)
        q.put((rank, ok, total))
        dist.barrier()
        if rank != 0:                                # importers unmap before the root frees the archive
            pa.close()
        dist.barrier()
        pa.close()
    finally:
        dist.destroy_process_group()


def test_peer_archive_over_nvlink():
    """Exchange step over CUDA-IPC peer memory: every rank finds its own stream, byte-identical to the
    oracle's, at its offset of the root's archive slot, for three pipelined steps."""
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    res = [q.get(timeout=5) for _ in range(world)]
    assert all(ok for _, ok, _ in res), res
    assert len({t for _, _, t in res}) == 1          # every rank saw the same total


def _peer_stream_worker(rank, world, port, n_total, q):
    for p in (HERE, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    import __graft_entry__ as g
    import oracle_lib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        pkg = g.load_package()
        from compression_accelerator_b200 import multi
        ctx = pkg.Context(rank)
        host = corpus.repeat_to(corpus.fixture("all-mtg-cards.txt"), n_total)
        sh = multi.PeerShardedStream(ctx, n_total, dev)
        shard = torch.from_numpy(host[sh.lo:sh.hi].copy()).to(dev)
        torch.cuda.synchronize()
        want = oracle_lib.checker().compress(host.tobytes()) if rank == 0 else None
        ok = True

        def check(k):
            total, x = sh.finish(k)
            return rank != 0 or (total == len(want) and bytes(x.cpu().numpy()) == want)

        sh.submit(shard)                                  # the same stream three times, two in flight
        sh.submit(shard)
        ok = check(0) and ok
        sh.submit(shard)
        ok = check(1) and ok
        ok = check(2) and ok
        if rank == 0:
            q.put(ok)
        dist.barrier()
        if rank != 0:
            sh.close()
        dist.barrier()
        sh.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [65536 * 37 + 4321, 1 << 26])
def test_sharded_stream_over_peer_memory(n_total):
    """One stream, fragments sharded over the GPUs, assembled over NVLink peer memory: byte-identical
    to the oracle's one-shot stream."""
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_stream_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert q.get(timeout=5)

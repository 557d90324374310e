"""Multi-rank exchange step (compression-accelerator_b200/multi.py) over gloo on CPU, world_size 2
and 3.  The local compression of each rank's fragment run is played by the CPU oracle here (test
infrastructure); on the GPU box the same exchange carries the CUDA path's payloads
(test_multi_gpu_sharded_stream, bench.py --gpus N)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import corpus

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
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g.load_package()
        from compression_accelerator_b200 import multi
        chk = oracle_lib.checker()
        data = corpus.repeat_to(corpus.fixture("alice29.txt"), n_total).tobytes() if n_total else b""
        sh = multi.ShardedStream(n_total, torch.device("cpu"))
        shard = data[sh.lo:sh.hi]
        # header-less payload of a run of whole fragments = the shard's own stream minus its preamble
        payload = chk.compress(shard)[len(multi.varint32(len(shard))):] if shard else b""
        pt = torch.frombuffer(bytearray(payload) + bytearray(16), dtype=torch.uint8)
        dst = torch.zeros(chk.max_compressed_length(n_total) + 16, dtype=torch.uint8) if rank == 0 else None
        total = sh.assemble(pt, len(payload), dst)
        if rank == 0:
            want = chk.compress(data)
            q.put((total == len(want), bytes(dst[:total].numpy()) == want, total))
        # archive shape: every rank its own whole stream, packed back to back on the root
        ag = multi.ArchiveGather(None, torch.device("cpu"), chk.max_compressed_length(len(shard) + BLOCK))
        mine = chk.compress(shard)
        ag.exchange(torch.frombuffer(bytearray(mine), dtype=torch.uint8), torch.tensor([len(mine)]))
        if rank == 0:
            ok = True
            for r in range(world):
                lo, hi = multi.fragment_range(n_total, r, world)
                w = chk.compress(data[lo:hi])
                ok &= bytes(ag.archive[ag.offsets[r]:ag.offsets[r] + len(w)].numpy()) == w
            q.put((ok, ag.total))
        dist.barrier()
    finally:
        dist.destroy_process_group()


BLOCK = 65536


@pytest.mark.parametrize("world,n_total", [(2, 65536 * 5 + 1234), (3, 65536 * 2 + 7), (2, 100), (2, 0)])
def test_sharded_stream_gloo(world, n_total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    len_ok, bytes_ok, total = q.get(timeout=5)
    assert len_ok and bytes_ok, (n_total, total)
    arch_ok, arch_total = q.get(timeout=5)
    assert arch_ok


def test_fragment_ranges_cover_stream():
    import __graft_entry__ as g
    g.load_package()
    from compression_accelerator_b200 import multi
    for n in (0, 1, 65535, 65536, 65537, 65536 * 7 + 3, (1 << 32) - 1):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = multi.fragment_range(n, r, world)
                assert lo == prev and lo % 65536 == 0 or lo == n
                prev = hi
            assert prev == n
    assert multi.varint32(0) == b"\x00" and multi.varint32(4096) == b"\x80\x20"
    assert multi.varint32(1 << 30) == bytes([0x80, 0x80, 0x80, 0x80, 0x04])

"""Multi-GPU Snappy: fragments sharded over the GPUs of one box, one process per GPU.

Snappy's fragments are independent (fresh table and base pointer per 64 KiB block,
reference tests/snappy/snappy.cc:1013-1061); the only datum that crosses a fragment boundary is
the running output offset.  So a stream shards with exactly one exchange step:

    rank r compresses fragments [F*r/R, F*(r+1)/R) of the stream, header-less   (local kernels)
    all_gather of one int64 per rank: the payload sizes                         (NCCL, 8 B/rank)
    exclusive scan of the sizes -> payload offsets (+ the varint header length)
    payload of rank r lands at dst + hdr + offset[r] on the root                (NCCL send/recv)

The root writes varint32(N).  The result is byte-identical to the single-GPU stream.  The same
exchange with hdr = 0 and whole streams per rank builds an "archive" of streams (inputs of 4 GiB
and more cannot be ONE Snappy stream: the preamble is a varint32, snappy.cc:1006-1007).

torch.distributed is plumbing only; the exchange also runs over gloo with CPU tensors, which is
how tests/test_multi.py covers it without a GPU.
"""
import torch
import torch.distributed as dist

BLOCK = 65536


def fragment_range(n_total: int, rank: int, world: int):
    """Byte range [lo, hi) of rank's run of whole fragments: fragments [F*r/R, F*(r+1)/R)."""
    n_frag = (n_total + BLOCK - 1) // BLOCK
    f_lo = n_frag * rank // world
    f_hi = n_frag * (rank + 1) // world
    return min(f_lo * BLOCK, n_total), min(f_hi * BLOCK, n_total)


def varint32(n: int) -> bytes:
    """tests/snappy/snappy-stubs-internal.h:552-578 (host-side copy for the root's header)."""
    out = bytearray()
    while n >= 128:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out)


class PayloadExchange:
    """The exchange step: all_gather(sizes) -> scan -> gather payload on `root`.

    payload / dst are uint8 tensors on the process group's device type (CUDA for NCCL, CPU for
    gloo).  Every call is collective over `group`."""

    def __init__(self, device, group=None, root: int = 0):
        self.device, self.group, self.root = device, group, root
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._sizes = torch.zeros(self.world, dtype=torch.int64, device=device)
        self._mine = torch.zeros(1, dtype=torch.int64, device=device)
        self.launch_like_ops = 0          # collectives issued (not kernels of ours)

    def sizes(self, my_size) -> torch.Tensor:
        """all_gather of one int64 per rank.  my_size: python int or 1-element int64 tensor."""
        if isinstance(my_size, torch.Tensor):
            self._mine.copy_(my_size.reshape(1).to(torch.int64))
        else:
            self._mine.fill_(int(my_size))
        dist.all_gather_into_tensor(self._sizes, self._mine, group=self.group)
        self.launch_like_ops += 1
        return self._sizes

    def gather(self, payload: torch.Tensor, my_size, dst: torch.Tensor = None, base: int = 0):
        """Returns (total_len, offsets list).  On the root, dst[base + off[r] : +size[r]] holds rank r's
        payload afterwards (the root's own payload is copied locally)."""
        sizes = self.sizes(my_size).tolist()              # host sync: 8 bytes per rank
        offs, run = [], base
        for s in sizes:
            offs.append(run)
            run += s
        if self.rank == self.root:
            if dst is None or dst.numel() < run:
                raise ValueError("root dst too small: need %d bytes" % run)
            ops = []
            for r in range(self.world):
                if sizes[r] == 0:
                    continue
                view = dst[offs[r]:offs[r] + sizes[r]]
                if r == self.root:
                    view.copy_(payload[:sizes[r]])
                else:
                    ops.append(dist.P2POp(dist.irecv, view, r, group=self.group))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
                self.launch_like_ops += 1
        elif sizes[self.rank]:
            ops = [dist.P2POp(dist.isend, payload[:sizes[self.rank]], self.root, group=self.group)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            self.launch_like_ops += 1
        return run, offs


class ShardedStream:
    """ONE Snappy stream of n_total bytes (< 2^32) compressed by all ranks of the group.

    compress_fragments(src_shard) -> (payload tensor, size) is the local, header-less compression of
    a run of whole fragments: on the GPU it is Context.compress_device(..., header=False)."""

    def __init__(self, n_total: int, device, group=None, root: int = 0):
        if n_total >= 1 << 32:
            raise ValueError("a Snappy stream holds < 2^32 bytes; use an archive of streams")
        self.n_total = n_total
        self.ex = PayloadExchange(device, group, root)
        self.lo, self.hi = fragment_range(n_total, self.ex.rank, self.ex.world)
        self.header = varint32(n_total)

    def assemble(self, payload: torch.Tensor, my_size, dst: torch.Tensor = None) -> int:
        """Collective.  Returns the stream length; on the root dst[:length] is the stream."""
        hdr = len(self.header)
        if self.ex.rank == self.ex.root and dst is not None:
            dst[:hdr].copy_(torch.frombuffer(bytearray(self.header), dtype=torch.uint8))
        total, _ = self.ex.gather(payload, my_size, dst, base=hdr)
        return total


class ArchiveGather:
    """Weak-scaling shape of bench.py: every rank compresses its own whole stream; the streams are
    packed back to back on the root, offsets from the all-gathered sizes."""

    def __init__(self, ctx, device, cap_per_rank: int, group=None, root: int = 0):
        self.ctx = ctx
        self.ex = PayloadExchange(device, group, root)
        self.archive = None
        if self.ex.rank == root:
            self.archive = torch.empty(cap_per_rank * self.ex.world, dtype=torch.uint8, device=device)
        self.offsets = None
        self.total = 0

    def exchange(self, stream: torch.Tensor, out_n: torch.Tensor) -> int:
        self.total, self.offsets = self.ex.gather(stream, out_n, self.archive, base=0)
        return self.total


class _DevicePtr:
    """A raw device allocation of the library seen by torch (torch.as_tensor over
    __cuda_array_interface__)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerArchive:
    """The exchange step of bench.py's weak-scaling shape over NVLink peer memory: every rank
    compresses its own stream, the streams land back to back in an archive slot on the root.  Sizes,
    offsets and payload bytes move by the library's own kernels through CUDA-IPC peer pointers
    (csrc/peer.cuh); the process group is used once, to pass the IPC handles around.  Nothing in a
    step needs the host, so step k's payload moves while step k+1 is being compressed.

        pa = PeerArchive(ctx, device, cap_per_rank)
        for k in range(steps):
            pa.reuse(k - 2)                      # buffer k % 2 is free again
            ctx.compress_device(src, dst[k % 2], out_n[k % 2])
            pa.gather(dst[k % 2], out_n[k % 2], k)
        total, offsets = pa.wait(steps - 1)      # slot (steps - 1) % 2 on the root is complete
    """

    def __init__(self, ctx, device, cap_per_rank: int, group=None, root: int = 0):
        import ctypes as C
        from . import api
        self._api, self._C = api, C
        self.ctx, self.device, self.group, self.root = ctx, device, group, root
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.slot_bytes = cap_per_rank * self.world
        L = api.lib()
        hb = L.snappy_b200_peer_handle_bytes()
        mine = C.create_string_buffer(hb)
        h = C.c_void_p()
        self._h = None
        # every rank learns whether every rank got through a stage, so a failure raises everywhere
        # instead of leaving the others inside a collective
        st = L.snappy_b200_peer_create(ctx._h, device.index, self.rank, self.world, root, self.slot_bytes,
                                       mine, C.byref(h))
        self._agree(st, "peer_create")
        self._h = h
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine.raw, group=group)
        st = L.snappy_b200_peer_connect(self._h, b"".join(everyone))
        self._agree(st, "peer_connect")          # also: every mailbox is mapped before anyone stores into one

    def _agree(self, status, what):
        text = self._api.lib().snappy_b200_peer_last_error().decode() if status else ""
        flag = torch.tensor([1 if status == 0 else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            self.close()
            raise self._api.SnappyB200Error(status or 3, "%s failed on some rank%s" % (what, ": " + text if text else ""))

    def _check(self, status):
        if status != 0:
            raise self._api.SnappyB200Error(status, self._api.lib().snappy_b200_peer_last_error().decode())

    def gather(self, payload: torch.Tensor, out_n: torch.Tensor, step: int, base: int = 0):
        """Asynchronous; ordered behind the work queued on the context's stream.  base = bytes the root
        keeps at the front of the slot (the varint preamble of a stream sharded over the ranks)."""
        self._check(self._api.lib().snappy_b200_peer_gather(self._h, payload.data_ptr(), out_n.data_ptr(), step, base))

    def reuse(self, step: int):
        if step >= 0:
            self._check(self._api.lib().snappy_b200_peer_reuse(self._h, step))

    def wait(self, step: int, base: int = 0):
        """Blocks until every rank's payload of `step` is in the root's slot; (total, offsets)."""
        C, L = self._C, self._api.lib()
        self._check(L.snappy_b200_peer_wait(self._h, step))
        total = C.c_ulonglong()
        offs = (C.c_ulonglong * self.world)()
        self._check(L.snappy_b200_peer_result(self._h, step, base, C.byref(total), offs))
        return total.value, list(offs)

    def wait_async(self, step: int):
        """Queue the wait only (the context's stream then follows the exchange stream)."""
        self._check(self._api.lib().snappy_b200_peer_wait(self._h, step))

    def archive(self, slot: int) -> torch.Tensor:
        """The slot as a uint8 tensor (on non-root ranks: the root's memory as mapped here)."""
        ptr = self._api.lib().snappy_b200_peer_archive(self._h, slot)
        return torch.as_tensor(_DevicePtr(ptr, self.slot_bytes))

    @property
    def launches(self) -> int:
        return self._api.lib().snappy_b200_peer_launch_count(self._h)

    def close(self):
        if getattr(self, "_h", None):
            self._api.lib().snappy_b200_peer_destroy(self._h)
            self._h = None


class PeerShardedStream:
    """ONE Snappy stream of n_total bytes (< 2^32) compressed by all ranks, assembled on the root over
    NVLink peer memory: rank r compresses fragments [F*r/R, F*(r+1)/R) header-less, its payload lands
    behind the varint preamble and the payloads of the ranks below (PeerArchive with base = preamble).
    Byte-identical to the single-GPU stream (tests/test_multi_gpu.py)."""

    def __init__(self, ctx, n_total: int, device, group=None, root: int = 0):
        if n_total >= 1 << 32:
            raise ValueError("a Snappy stream holds < 2^32 bytes; use an archive of streams")
        from . import api
        self.ctx, self.n_total, self.device, self.root = ctx, n_total, device, root
        world = dist.get_world_size(group)
        self.lo, self.hi = fragment_range(n_total, dist.get_rank(group), world)
        self.header = varint32(n_total)
        cap = api.max_compressed_length(n_total) + 64
        self.pa = PeerArchive(ctx, device, -(-cap // world), group, root)
        shard_cap = api.max_compressed_length(self.hi - self.lo)
        self._bufs = [(torch.empty(shard_cap, dtype=torch.uint8, device=device),
                       torch.zeros(1, dtype=torch.int64, device=device)) for _ in range(2)]
        self.step = 0

    def submit(self, src_shard: torch.Tensor) -> int:
        """Queue this rank's part of one stream (src_shard = bytes [lo, hi) of the input); returns the step.
        Two streams may be in flight: finish(k) before submit of step k + 2 (its slot is reused)."""
        k = self.step
        self.pa.reuse(k - 2)
        d, o = self._bufs[k & 1]
        self.ctx.compress_device(src_shard, d, o, header=False)
        self.pa.gather(d, o, k, base=len(self.header))
        self.step = k + 1
        return k

    def finish(self, step: int):
        """Blocks until the stream of `step` is complete on the root; returns (length, tensor or None)."""
        total, _ = self.pa.wait(step, base=len(self.header))
        if self.pa.rank != self.root:
            return total, None
        slot = self.pa.archive(step & 1)
        slot[:len(self.header)].copy_(torch.frombuffer(bytearray(self.header), dtype=torch.uint8))
        torch.cuda.synchronize(self.device)
        return total, slot[:total]

    def close(self):
        self.pa.close()

"""Row-range sharding of the lookup over the GPUs of one box.

Rows are independent byte-aligned streams, so the path shards with no data-path
collective: the reference itself cuts a batch into `ceil(n / T)` contiguous
chunks, one per host thread (src/reader.cpp:65-80), and proves that the cut never
changes a bit (src/tests.cpp:90-113).  Here the same cut is made per GPU: every
rank (one process per GPU, torch.distributed) holds a replica of the compressed
table and decodes its own contiguous slice of the OUTPUT rows.

The only collective is optional: `all_gather_rows` assembles the dense matrix on
every rank (NCCL all-gather over NVLink on the GPU box; gloo in the CPU tests)
when one device must hold all of it.  torch is imported lazily and is plumbing
only (process group, device tensors); every row is decoded by the sm_100a
kernels behind `_memb`.
"""


def shard_bounds(n, world_size):
    """[begin_0, begin_1, ..., begin_G] with chunks of ceil(n / G) rows, the cut of
    src/reader.cpp:65-80 (jobSize = ceil(n / numThreads); the tail ranks may be
    short or empty)."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    job = (n + world_size - 1) // world_size
    return [min(n, g * job) for g in range(world_size + 1)]


def shard_range(n, rank, world_size):
    """(begin, end) of the rows rank `rank` owns."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    bounds = shard_bounds(n, world_size)
    return bounds[rank], bounds[rank + 1]


def _rank_world(group=None):
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def all_gather_rows(local, n_total, group=None):
    """Assemble the (n_total, dim) matrix from the per-rank slices of `shard_range`.

    `local` is this rank's (end - begin, dim) tensor.  The gathered buffer holds `world`
    slots of ceil(n / G) rows; the local slice is copied into its slot and ONE in-place
    all_gather_into_tensor moves everything (pad rows of short tail slices are cut off
    again).  Use `decode_range_gathered` to decode straight into the slot instead."""
    import torch
    import torch.distributed as dist

    rank, world = _rank_world(group)
    if world == 1:
        return local
    job = (n_total + world - 1) // world
    begin, end = shard_range(n_total, rank, world)
    if local.shape[0] != end - begin:
        raise ValueError("local slice has %d rows, rank %d owns %d" % (local.shape[0], rank, end - begin))
    gathered = torch.empty((world * job, local.shape[1]), dtype=local.dtype, device=local.device)
    slot = gathered[rank * job:(rank + 1) * job]
    slot[: local.shape[0]].copy_(local)
    dist.all_gather_into_tensor(gathered, slot, group=group)
    return gathered[:n_total]


def decode_range_gathered(reader, begin=0, end=None, *, group=None, out=None):
    """Full-vocabulary decode where EVERY rank ends up with the dense (end - begin, dim) matrix:
    this rank decodes its row range directly into its slot of the gathered buffer (no send
    copy, no padded temporary), then one in-place NCCL all-gather.  `out`: optional
    (world * ceil(n / world), dim) float32 buffer on the reader's device to reuse.
    Returns the (end - begin, dim) view of it."""
    import torch
    import torch.distributed as dist

    rank, world = _rank_world(group)
    end = reader.rows if end is None else end
    n = end - begin
    job = (n + world - 1) // world
    device = torch.device("cuda", reader._impl.device())
    if out is None:
        out = torch.empty((world * job, reader.dim), dtype=torch.float32, device=device)
    _check_out(out, world * job, reader.dim, device)
    lo, hi = shard_range(n, rank, world)
    slot = out[rank * job:(rank + 1) * job]
    stream = torch.cuda.current_stream(device).cuda_stream
    reader._impl.decode_range_into(begin + lo, begin + hi, slot.data_ptr(), True, stream)
    if world > 1:
        dist.all_gather_into_tensor(out, slot, group=group)
    return out[:n]


class GatheredMatrix:
    """The dense (rows, dim) matrix of a reader on EVERY rank, filled by a fused decode -> all-gather:
    the buffer is torch symmetric memory with an NVLS multicast mapping, each rank decodes its row
    range with stores to the multicast address (memb_b200_decode_range_multicast), and the NVSwitch
    replicates every store into all ranks' buffers while the decode is still running -- no
    separate collective, no staging copy.  `fill()` returns the (rows, dim) view once every
    rank's rows have arrived (a one-element NCCL all-reduce on the current stream serves as the
    barrier: the symmetric-memory handle's own device barrier, queued right before the decode
    kernel without a host synchronisation in between, hung on the B200 box -- r02).

    Raises RuntimeError when the process group has no multicast support (then use
    decode_range_gathered: decode in place + one NCCL all-gather)."""

    def __init__(self, reader, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        self.reader = reader
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.rows, self.dim = reader.rows, reader.dim
        self.job = (self.rows + self.world - 1) // self.world
        self.device = torch.device("cuda", reader._impl.device())
        self.buffer = symm.empty((self.world * self.job, self.dim), dtype=torch.float32, device=self.device)
        self.handle = symm.rendezvous(self.buffer, self.group.group_name)
        if not self.handle.multicast_ptr:
            raise RuntimeError("symmetric memory without a multicast mapping (no NVLS on this box)")
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)

    def fill(self):
        import torch

        lo, hi = shard_range(self.rows, self.rank, self.world)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        slot = self.handle.multicast_ptr + self.rank * self.job * self.dim * 4
        import torch.distributed as dist

        dist.all_reduce(self._flag, group=self.group)  # nobody still reads the previous contents
        self.reader._impl.decode_range_multicast(lo, hi, slot, stream)
        dist.all_reduce(self._flag, group=self.group)  # every rank's kernel is complete: its rows have landed everywhere
        return self.buffer[: self.rows]


def _check_out(out, rows, dim, device):
    import torch

    if (tuple(out.shape) != (rows, dim) or out.dtype != torch.float32 or not out.is_contiguous()
            or out.device != device):
        raise ValueError("out must be a contiguous float32 (%d, %d) tensor on %s" % (rows, dim, device))


def decode_range_sharded(reader, begin=0, end=None, *, gather=False, group=None, out=None):
    """Full-vocabulary decode (Reader.tokenizer_embedding / to_keyed_vectors path,
    python/memb/reader.py:19-30,94-111) sharded by row range: this rank decodes its
    slice of rows [begin, end) on its own GPU into a device tensor.

    Returns (tensor, (slice_begin, slice_end)); with gather=True the tensor is the
    whole (end - begin, dim) matrix on every rank (one NCCL all-gather)."""
    import torch

    rank, world = _rank_world(group)
    rows = reader.rows
    end = rows if end is None else end
    lo, hi = shard_range(end - begin, rank, world)
    device = torch.device("cuda", reader._impl.device())
    if out is None:
        out = torch.empty((hi - lo, reader.dim), dtype=torch.float32, device=device)
    _check_out(out, hi - lo, reader.dim, device)
    stream = torch.cuda.current_stream(device).cuda_stream
    reader._impl.decode_range_into(begin + lo, begin + hi, out.data_ptr(), True, stream)
    if gather:
        return all_gather_rows(out, end - begin, group), (begin, end)
    return out, (begin + lo, begin + hi)


def batch_embedding_ids_sharded(reader, ids, *, gather=False, group=None):
    """Batched lookup of row ids with the QUERIES sharded by rank: `ids` is the whole
    batch (identical on every rank, int32 numpy or torch); this rank decodes rows
    shard_range(len(ids), rank, world) of the output."""
    import torch

    rank, world = _rank_world(group)
    n = len(ids)
    lo, hi = shard_range(n, rank, world)
    device = torch.device("cuda", reader._impl.device())
    ids_t = torch.as_tensor(ids[lo:hi], dtype=torch.int32).to(device, non_blocking=True).contiguous()
    out = torch.empty((hi - lo, reader.dim), dtype=torch.float32, device=device)
    stream = torch.cuda.current_stream(device).cuda_stream
    reader._impl.decode_ids_into(ids_t.data_ptr(), True, hi - lo, out.data_ptr(), True, stream)
    if gather:
        return all_gather_rows(out, n, group), (0, n)
    return out, (lo, hi)

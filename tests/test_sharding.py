"""Row-range sharding (memb_b200/sharding.py): partition arithmetic and the all-gather
plumbing, on CPU with world_size-2 gloo.  The decode itself needs a GPU (tests marked gpu)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_bounds_follow_the_reference_chunking():
    from memb_b200.sharding import shard_bounds, shard_range

    # jobSize = ceil(n / T), src/reader.cpp:65-80: 1025 words over 4 workers -> 257,257,257,254
    assert shard_bounds(1025, 4) == [0, 257, 514, 771, 1025]
    assert shard_bounds(0, 3) == [0, 0, 0, 0]
    assert shard_bounds(5, 8) == [0, 1, 2, 3, 4, 5, 5, 5, 5]  # tail ranks are empty
    assert shard_bounds(2_200_000, 8)[-1] == 2_200_000
    for n in (0, 1, 31, 32, 33, 1023, 1024, 1025, 2_200_000):
        for world in (1, 2, 3, 4, 8):
            bounds = [shard_range(n, r, world) for r in range(world)]
            assert bounds[0][0] == 0 and bounds[-1][1] == n
            assert all(bounds[i][1] == bounds[i + 1][0] for i in range(world - 1))
            job = -(-n // world)
            assert all(0 <= e - b <= job for b, e in bounds)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    with pytest.raises(ValueError):
        shard_bounds(10, 0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gather_worker(rank, world, port, n, dim, result_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    from memb_b200.sharding import all_gather_rows, shard_range

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(n * dim, dtype=torch.float32).reshape(n, dim)  # stands in for decoded rows
        begin, end = shard_range(n, rank, world)
        gathered = all_gather_rows(full[begin:end].clone(), n)
        assert gathered.shape == (n, dim)
        assert torch.equal(gathered, full), "rank %d: gathered matrix differs" % rank
        with pytest.raises(ValueError):
            all_gather_rows(full[: max(end - begin - 1, 0) + 2].clone(), n + 7)
        np.save(os.path.join(result_dir, "ok%d.npy" % rank), np.array([begin, end]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1025, 7, 2])
def test_all_gather_rows_world_size_2_gloo(tmp_path, n):
    import torch.multiprocessing as mp

    world, dim = 2, 12
    port = _free_port()
    mp.spawn(_gather_worker, args=(world, port, n, dim, str(tmp_path)), nprocs=world, join=True)
    spans = [np.load(tmp_path / ("ok%d.npy" % r)) for r in range(world)]
    assert spans[0][0] == 0 and spans[0][1] == spans[1][0] and spans[1][1] == n


@pytest.mark.gpu
def test_sharded_decode_equals_whole_decode(oracle, synthetic):
    """threadedDecoderWorks (src/tests.cpp:90-113) at device granularity: decoding the
    shards of a 1-, 2-, 3- and 8-way partition one after the other on one GPU gives the same
    bits as one whole decode (the N-GPU run is the same calls on different devices)."""
    import memb_b200
    from memb_b200.sharding import shard_range

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=11)
    reader = memb_b200.Reader(syn.path)
    whole = reader.decode_range(0, reader.rows)
    ids = np.random.default_rng(5).integers(-1, reader.rows, 1025).astype(np.int32)
    whole_ids = reader.batch_embedding_ids(ids)
    for world in (1, 2, 3, 8):
        parts = [reader.decode_range(*shard_range(reader.rows, r, world)) for r in range(world)]
        assert np.array_equal(np.concatenate(parts).view(np.uint32), whole.view(np.uint32))
        parts = []
        for r in range(world):
            b, e = shard_range(len(ids), r, world)
            parts.append(reader.batch_embedding_ids(ids[b:e]))
        assert np.array_equal(np.concatenate(parts).view(np.uint32), whole_ids.view(np.uint32))


@pytest.mark.gpu
def test_decode_range_sharded_single_process(synthetic):
    import torch

    import memb_b200
    from memb_b200.sharding import batch_embedding_ids_sharded, decode_range_sharded

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=11)
    reader = memb_b200.Reader(syn.path)
    out, span = decode_range_sharded(reader, gather=True)
    torch.cuda.synchronize()
    assert span == (0, reader.rows)
    assert np.array_equal(out.cpu().numpy().view(np.uint32), reader.decode_range(0, reader.rows).view(np.uint32))
    ids = np.arange(0, 3000, 7, dtype=np.int32)
    rows, span = batch_embedding_ids_sharded(reader, ids)
    torch.cuda.synchronize()
    assert span == (0, len(ids))
    assert np.array_equal(rows.cpu().numpy().view(np.uint32), reader.batch_embedding_ids(ids).view(np.uint32))


@pytest.mark.gpu
def test_decode_range_gathered_and_out_validation(oracle, synthetic):
    """decode_range_gathered decodes straight into this rank's slot of the gathered buffer (no send copy);
    with one process the slot is the whole buffer.  `out=` arguments are validated instead of being written
    out of bounds."""
    import torch

    import memb_b200
    from memb_b200.sharding import decode_range_gathered, decode_range_sharded

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=11)
    reader = memb_b200.Reader(syn.path)
    want = oracle.open(syn.path).decode_range(0, 3000)
    got = decode_range_gathered(reader)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32))
    part = decode_range_gathered(reader, 100, 1133)
    torch.cuda.synchronize()
    assert np.array_equal(part.cpu().numpy().view(np.uint32), want[100:1133].view(np.uint32))
    reuse = torch.empty((3000, 300), dtype=torch.float32, device="cuda")
    assert decode_range_gathered(reader, out=reuse).data_ptr() == reuse.data_ptr()
    for bad in (torch.empty((2999, 300), dtype=torch.float32, device="cuda"),
                torch.empty((3000, 300), dtype=torch.float64, device="cuda"),
                torch.empty((3000, 600), dtype=torch.float32, device="cuda")[:, ::2],
                torch.empty((3000, 300), dtype=torch.float32)):
        with pytest.raises(ValueError):
            decode_range_sharded(reader, out=bad)
        with pytest.raises(ValueError):
            decode_range_gathered(reader, out=bad)

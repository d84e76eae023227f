"""DM sharding + the single candidate gather (peasoup_b200/shard.py) with world_size 2 over gloo
on CPU: the N>1 path of bench.py without GPUs."""
import os
import socket
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.multiprocessing as mp  # noqa: E402

from peasoup_b200.host import CAND_DTYPE  # noqa: E402
from peasoup_b200.shard import canonical_order, shard_range  # noqa: E402


def fake_records(rank, ndm_local, offset, rng):
    recs = []
    for d in range(ndm_local):
        for c in range(int(rng.integers(0, 4))):
            top = (10.0 * (offset + d), offset + d, 0.0, 0, 20.0 + c, 100.0 + c, 1)
            recs.append(top)
            for a in range(int(rng.integers(0, 3))):
                recs.append((10.0 * (offset + d), offset + d, 1.0 + a, 1, 10.0 + a, 100.0 + c, 0))
    return np.array(recs, CAND_DTYPE) if recs else np.zeros(0, CAND_DTYPE)


def worker(rank, world, port, ndm, out_dir):
    import torch.distributed as dist
    from peasoup_b200.shard import gather_candidates, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    start, count = shard_range(ndm, rank, world)
    recs = fake_records(rank, count, start, np.random.default_rng(100 + rank))
    np.save(os.path.join(out_dir, f"local{rank}.npy"), recs)
    merged = gather_candidates(recs)
    if rank == 0:
        np.save(os.path.join(out_dir, "merged.npy"), merged)
    else:
        assert merged is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions():
    for ndm in (0, 1, 7, 1024, 4096):
        for world in (1, 2, 3, 8):
            spans = [shard_range(ndm, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == ndm
            assert all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_canonical_order_keeps_associates_behind_parent():
    rng = np.random.default_rng(0)
    a = fake_records(0, 5, 5, rng)
    b = fake_records(1, 5, 0, rng)
    merged = canonical_order(np.concatenate([a, b]))
    assert merged["is_top"][0] == 1
    tops = np.maximum.accumulate(np.where(merged["is_top"] == 1, np.arange(len(merged)), 0))
    assert np.all(np.diff(merged["dm_idx"][tops]) >= 0)
    assert np.array_equal(np.sort(merged, order=list(CAND_DTYPE.names)), np.sort(np.concatenate([a, b]), order=list(CAND_DTYPE.names)))
    assert np.all(merged["dm_idx"] == merged["dm_idx"][tops])


def test_gather_candidates_world_size_2_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ndm = 11
    mp.spawn(worker, args=(2, port, ndm, str(tmp_path)), nprocs=2, join=True)
    merged = np.load(tmp_path / "merged.npy")
    want = canonical_order(np.concatenate([np.load(tmp_path / "local0.npy"), np.load(tmp_path / "local1.npy")]))
    assert np.array_equal(merged, want)
    assert len(merged) > 0 and np.all(np.diff(merged["dm_idx"][merged["is_top"] == 1]) >= 0)


def strong_worker(rank, world, port, ndm, out_dir):
    """bench.py --config 4 without GPUs: shard the DM range, gather ONCE, rank 0 runs the post-join
    distillations (ps_distill_merged, host code of the product library) on the merged list."""
    import torch.distributed as dist
    from peasoup_b200.host import distill_merged
    from peasoup_b200.shard import gather_candidates, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    whole = np.load(os.path.join(out_dir, "whole.npy"))
    start, count = shard_range(ndm, rank, world)
    tops = np.maximum.accumulate(np.where(whole["is_top"] == 1, np.arange(len(whole)), 0))
    mine = whole[(whole["dm_idx"][tops] >= start) & (whole["dm_idx"][tops] < start + count)]
    merged = gather_candidates(mine)
    if rank == 0:
        np.save(os.path.join(out_dir, "final.npy"), distill_merged(merged, 1e-4, 16))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_job_gives_the_unsharded_final_list_gloo(tmp_path, lib):
    """The fixed job of BASELINE configs[3] must not depend on how it is sharded: per-DM candidate
    lists split over two ranks, gathered and distilled on rank 0, equal the distillation of the
    unsharded list (distinct S/N values: the reference's sort is unstable on ties)."""
    from peasoup_b200.host import distill_merged
    rng = np.random.default_rng(5)
    ndm = 13
    parts = [fake_records(0, 1, d, rng) for d in range(ndm)]
    whole = np.concatenate(parts)
    whole["freq"] = (rng.choice([3.3, 6.6, 9.9, 50.0, 51.7], len(whole)) * (1 + rng.normal(0, 2e-5, len(whole)))).astype(np.float32)
    whole["snr"] = (9.0 + 0.173 * rng.permutation(len(whole))).astype(np.float32)
    np.save(tmp_path / "whole.npy", whole)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(strong_worker, args=(2, port, ndm, str(tmp_path)), nprocs=2, join=True)
    final = np.load(tmp_path / "final.npy")
    want = distill_merged(canonical_order(whole), 1e-4, 16)
    assert len(final) == len(want) >= len(whole) > 0
    for name in CAND_DTYPE.names:
        assert np.array_equal(final[name], want[name]), name
    assert (final["is_top"] == 1).sum() < (whole["is_top"] == 1).sum()   # the distillers did merge something

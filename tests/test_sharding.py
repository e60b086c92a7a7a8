"""Multi-GPU sharding logic on CPU: world-index ranges, shard-reproducible scenes and the
end-of-run gather of summary statistics, exercised with world_size 2 over the gloo backend.  The step path itself
has no collective; what is tested here is everything around it that bench.py and a multi-rank caller rely on."""
import os
import socket

import numpy as np
import pytest

import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes


def test_shard_ranges_partition_the_batch():
    for n in (0, 1, 7, 100_000, 2**20 + 3):
        for ws in (1, 2, 3, 8):
            ranges = [cs.shard_range(n, r, ws) for r in range(ws)]
            assert ranges[0][0] == 0 and sum(c for _, c in ranges) == n
            for (f0, c0), (f1, _) in zip(ranges, ranges[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in ranges) - min(c for _, c in ranges) <= 1
    with pytest.raises(ValueError):
        cs.shard_range(10, 2, 2)


@pytest.mark.parametrize("maker", [scenes.c2_perturbed, scenes.c3_stack])
def test_scene_shards_equal_slices_of_the_whole(maker):
    whole = maker(37)
    for ws in (2, 3):
        for r in range(ws):
            first, count = cs.shard_range(37, r, ws)
            part = maker(count, first)
            assert np.array_equal(part.view(np.uint8), whole[first:first + count].view(np.uint8))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world_size, port, n_total, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        first, count = cs.shard_range(n_total, rank, world_size)
        shard = scenes.c3_stack(count, first)
        # per-rank summary statistics as bench.py gathers them: [elapsed_ms, body_substeps, checksum of the shard]
        checksum = float(shard["position"].sum())
        stats = torch.tensor([10.0 + rank, float(count * 10 * 20), checksum], dtype=torch.float64)
        gathered = [torch.zeros(3, dtype=torch.float64) for _ in range(world_size)]
        dist.all_gather(gathered, stats)
        t = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max-over-ranks timing rule
        if rank == 0:
            np.save(os.path.join(out_dir, "gathered.npy"), torch.stack(gathered).numpy())
            np.save(os.path.join(out_dir, "tmax.npy"), t.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_gather_matches_unsharded(tmp_path):
    import torch.multiprocessing as mp
    n_total, ws = 101, 2
    mp.spawn(_rank_main, args=(ws, _free_port(), n_total, str(tmp_path)), nprocs=ws, join=True)
    g = np.load(tmp_path / "gathered.npy")
    whole = scenes.c3_stack(n_total)
    assert g[:, 1].sum() == n_total * 10 * 20  # body-substeps of one frame over all ranks
    assert g[:, 2].sum() == pytest.approx(float(whole["position"].sum()), rel=1e-12)
    assert float(np.load(tmp_path / "tmax.npy")[0]) == 11.0  # the slower rank's time is the job's time

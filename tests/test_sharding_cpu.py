"""world_size-2 `gloo` test of the multi-GPU host logic (crocoddyl_b200/sharding.py): contiguous sharding of a batch of
independent OCPs and the single all-gather of per-problem records. The per-shard solve is done by the CPU oracle
here (test infrastructure); on the GPUs the same functions run over NCCL in bench.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, out_dir):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import torch.distributed as dist
    import oracle_lib as ol
    from crocoddyl_b200 import sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    T = 20
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, n_total), rng.uniform(-2, 2, n_total), rng.uniform(-np.pi, np.pi, n_total)], 1)
    lo, hi = sharding.shard_range(n_total, rank, world)
    xs = np.repeat(x0[lo:hi, None, :], T + 1, 1)
    us = np.zeros((hi - lo, T, 2))
    ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0[lo:hi])
    ob.solve(xs, us, 100, False, 1e-9, nthreads=1)
    s = ob.scalars()
    scal = np.zeros((hi - lo, 17))
    scal[:, 0], scal[:, 9], scal[:, 1], scal[:, 11] = s["cost"], s["iter"], s["stop"], s["converged"]
    rec = sharding.gather_records(sharding.records_from_scalars(scal), n_total, dist)
    np.save(os.path.join(out_dir, "rec%d.npy" % rank), rec)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [11, 8])
def test_two_rank_sharded_solve_matches_single_process(tmp_path, n_total):
    sys.path.insert(0, HERE)
    import oracle_lib as ol
    from crocoddyl_b200 import sharding

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_total, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rec0.npy"), np.load(tmp_path / "rec1.npy")
    assert r0.shape == (n_total, 4)
    assert np.array_equal(r0, r1)  # every rank holds the same gathered table
    T = 20
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, n_total), rng.uniform(-2, 2, n_total), rng.uniform(-np.pi, np.pi, n_total)], 1)
    ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0)
    ob.solve(np.repeat(x0[:, None, :], T + 1, 1), np.zeros((n_total, T, 2)), 100, False, 1e-9, nthreads=1)
    s = ob.scalars()
    assert np.array_equal(r0[:, 0], s["cost"]) and np.array_equal(r0[:, 1], s["iter"])
    assert np.array_equal(r0[:, 2], s["stop"]) and np.array_equal(r0[:, 3], s["converged"])


def test_shard_ranges_partition_the_batch():
    from crocoddyl_b200 import sharding

    for n in (0, 1, 7, 8, 8192, 65537):
        for world in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sz = [b - a for a, b in r]
            assert max(sz) - min(sz) <= 1 and sz == sharding.shard_sizes(n, world)
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)

"""CPU test of the N>1 host logic: contiguous env-id sharding and the one collective of the path,
the end-of-run all-reduce of the tracking-error statistics, with world_size 2 over gloo."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from robobee3d_b200.batch import gather_metrics, shard_range
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nenv = 1001  # ragged on purpose
    lo, hi = shard_range(nenv, rank, world)
    # per-env metrics keyed by GLOBAL env id (what the device reduction of a shard would produce)
    ids = np.arange(lo, hi, dtype=np.float64)
    err, eff = ids * 0.5, ids * 0.25
    local = np.array([err.sum(), eff.sum(), float(hi - lo) * 3, float(hi - lo) * 10, float(rank == 1), err.max() if len(err) else 0, 0, 0])
    tot = gather_metrics(local)
    q.put((rank, lo, hi, tot.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_cover_and_do_not_overlap():
    from robobee3d_b200.batch import shard_range
    for n in (1, 7, 1000, 65536, 262144, 1001):
        for w in (1, 2, 3, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            assert all(hi >= lo for lo, hi in r)


def test_gather_metrics_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    ids = np.arange(1001, dtype=np.float64)
    exp = [ids.sum() * 0.5, ids.sum() * 0.25, 1001 * 3, 1001 * 10, 1.0, 1000 * 0.5, 0, 0]
    for rank, lo, hi, tot in res:
        assert np.allclose(tot, exp), (rank, tot, exp)
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == 1001


def test_gather_metrics_single_process_is_identity():
    from robobee3d_b200.batch import gather_metrics
    s = np.arange(8, dtype=np.float64)
    assert np.array_equal(gather_metrics(s), s)

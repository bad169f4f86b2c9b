"""Host-side logic of the N>1 path on CPU: two gloo ranks shard 20 patients by contiguous ranges with
no data-path exchange, and the only cross-rank step is the integer sum of per-rank success counts
(the analogue of `reduction(+:successCount)`, main_parallel.cpp:336)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import common as cm


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_patients, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nm03 = cm.load_product()
    lo, hi = nm03.shard_range(n_patients, world, rank)
    # every rank "processes" its own patients: here the oracle on one tiny slice per patient
    ok = 0
    digest = 0
    for pat in range(lo, hi):
        rc, r = cm.o_pipeline(cm.phantom(100, 100, pat, 0), want=("dilated",))
        ok += int(rc == 0)
        digest += int(r["dilated"].sum())
    t = torch.tensor([ok, hi - lo, digest], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)          # host-side join only
    tmax = torch.tensor([float(hi - lo)])
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)       # bench.py's max-over-ranks timing reduction
    q.put((rank, lo, hi, t.tolist(), tmax.item()))
    dist.destroy_process_group()


def test_two_rank_patient_sharding():
    world, n_patients = 2, 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_patients, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert [(r[1], r[2]) for r in res] == [(0, 2), (2, 5)]
    single = sum(int(cm.o_pipeline(cm.phantom(100, 100, pat, 0), want=("dilated",))[1]["dilated"].sum())
                 for pat in range(n_patients))
    for r in res:
        assert r[3] == [n_patients, n_patients, single]   # same totals on every rank, equal to one process
        assert r[4] == 3.0


@pytest.mark.parametrize("p,g", [(20, 1), (20, 2), (20, 4), (20, 8), (3, 8), (0, 4)])
def test_shard_range_partitions(lib, p, g):
    spans = [lib.shard_range(p, g, r) for r in range(g)]
    assert spans[0][0] == 0 and spans[-1][1] == p
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1
    if (p, g) == (20, 8):
        assert sorted(sizes) == [2, 2, 2, 2, 3, 3, 3, 3]   # SURVEY §7.4 load balance

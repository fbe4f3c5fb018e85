"""world_size-2 gloo test (CPU) of the host-side logic of the item-sharded path: the shard plan the C-ABI uses and the
in-place chunked all-gather every exchange of a sweep goes through (gpirt_b200/distributed.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpirt_b200.distributed import all_gather_chunks, shard_plan


def test_shard_plan_covers_items_and_respondents_once():
    for n, m, world in ((2000, 5000, 1), (2000, 5000, 8), (101, 37, 4), (10, 9, 8), (10000, 50000, 8)):
        p = shard_plan(n, m, world)
        items = [j for r in p["ranks"] for j in range(r["item0"], r["item0"] + r["m_local"])]
        resp = [i for r in p["ranks"] for i in range(r["resp0"], r["resp0"] + r["n_local"])]
        assert items == list(range(m)) and resp == list(range(n))
        assert all(r["m_local"] <= p["m_chunk"] and r["n_local"] <= p["n_chunk"] for r in p["ranks"])
        assert all(r["item0"] == r["rank"] * p["m_chunk"] for r in p["ranks"])  # global index = rank*chunk + local: streams keyed globally


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, m, ngrid):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = shard_plan(n, m, world)
        me = plan["ranks"][rank]
        ld = (plan["m_chunk"] + 15) // 16 * 16
        # the f* exchange buffer: world chunks of ngrid x ld (grid-row-major); each rank fills its own chunk
        table = torch.zeros(world * ngrid * ld, dtype=torch.float64)
        mine = table[rank * ngrid * ld:(rank + 1) * ngrid * ld].view(ngrid, ld)
        for jl in range(me["m_local"]):
            mine[:, jl] = torch.arange(ngrid, dtype=torch.float64) * 1000 + (me["item0"] + jl)
        all_gather_chunks(table, rank, world)
        full = table.view(world, ngrid, ld)
        for J in range(m):  # the theta kernel's addressing: chunk r = J // m_chunk, column jl = J % m_chunk
            r, jl = divmod(J, plan["m_chunk"])
            assert torch.equal(full[r, :, jl], torch.arange(ngrid, dtype=torch.float64) * 1000 + J)
        # theta exchange: chunk r holds respondents of rank r; global index = position
        theta = torch.full((world * plan["n_chunk"],), -1.0, dtype=torch.float64)
        theta[me["resp0"]:me["resp0"] + me["n_local"]] = torch.arange(me["resp0"], me["resp0"] + me["n_local"], dtype=torch.float64)
        all_gather_chunks(theta, rank, world)
        assert torch.equal(theta[:n], torch.arange(n, dtype=torch.float64))
        # per-item ll exchange + ordered total: identical on every rank
        ll = torch.zeros(world * plan["m_chunk"], dtype=torch.float64)
        ll[me["item0"]:me["item0"] + me["m_local"]] = -torch.arange(me["item0"], me["item0"] + me["m_local"], dtype=torch.float64)
        all_gather_chunks(ll, rank, world)
        tot = ll[:m].sum()
        both = [torch.zeros(()) .double() for _ in range(world)]
        dist.all_gather(both, tot)
        assert all(torch.equal(b, both[0]) for b in both) and float(tot) == -m * (m - 1) / 2
    finally:
        dist.destroy_process_group()


def test_chunked_all_gather_world2_gloo():
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 11, 9, 7), nprocs=2, join=True)

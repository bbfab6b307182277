"""CPU-only, world_size 2 over gloo: the N>1 path of the host logic -- every rank derives
the same row-strip plan, fills only its own rows, and the reduction of per-rank timings /
unit counts that bench.py performs.  Compute is done by the CPU oracle here (no GPU);
the GPU version of the same split is tests/test_gpu_parity.py::test_panel_size_and_sharding."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sequencealigner_b200 import AlignConfig, synth
from sequencealigner_b200.aligner import plan_strips


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        import bench
        store, c = synth.make("c1", n=700)
        cfg = AlignConfig("nw", "blosum62", gap_pen=4).resolve()
        plan = plan_strips(store, world)
        n = store.num
        mine = np.zeros(n * (n - 1) // 2, dtype=np.int32)
        cells = 0
        for row0, rows, owner in plan:
            if owner != rank:
                continue
            part, cl = oracle.align_all(cfg, store, first_j=max(row0, 1), last_j=row0 + rows)
            lo, hi = row0 * (row0 - 1) // 2, (row0 + rows) * (row0 + rows - 1) // 2
            mine[lo:hi] = part[lo:hi]
            cells += cl
        filled = torch.from_numpy((mine != 0).astype(np.int32))
        total = torch.from_numpy(mine.copy())
        dist.all_reduce(total)  # disjoint rows -> the sum is the merged matrix
        dist.all_reduce(filled)
        ms, units = bench.reduce_over_ranks(10.0 + rank, cells)
        q.put((rank, total.numpy(), int(filled.max()), ms, units, plan))
    finally:
        dist.destroy_process_group()


def test_two_ranks_fill_disjoint_rows_and_reduce():
    from oracle import oracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    store, _ = synth.make("c1", n=700)
    want, cells = oracle.align_all(AlignConfig("nw", "blosum62", gap_pen=4).resolve(), store)
    for rank, merged, overlap, ms, units, plan in res:
        assert np.array_equal(merged, want)
        assert overlap <= 1  # no cell filled by two ranks
        assert ms == pytest.approx(11.0) and units == cells  # max over ranks, sum of units
        assert plan == res[0][5]

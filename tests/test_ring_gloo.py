"""CPU tests (-m "not gpu") of the N>1 host path with two real processes over torch.distributed/gloo:
ring neighbours, blob exchange, the single av_vels reduction and the collate gather, checked against
the oracle's emulated ranks (reference decomposition, d2q9-bgk.c:341-447, 483-503)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, size, port, out):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(size))
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        lbm = importlib.import_module("lattice-boltzmann_b200")
        ring = importlib.import_module("lattice-boltzmann_b200.ring")
        import oracle_lib
        from helpers import load_grid
        # ring bookkeeping + blob exchange
        up, down = ring.ring_neighbours(rank, size)
        blob = bytes([rank]) * lbm.IPC_BLOB_BYTES
        b_up, b_down, allb = ring.exchange_blobs(dist, torch, blob, rank, size)
        assert b_up == bytes([up]) * lbm.IPC_BLOB_BYTES and b_down == bytes([down]) * lbm.IPC_BLOB_BYTES
        assert len(allb) == size
        # a short run of the whole domain by the oracle (every rank computes the same thing), then each
        # rank keeps only ITS slab and ITS per-step partial, as a GPU rank would hand them back
        p, mask, lines = load_grid("128x256")
        steps = 25
        ora = oracle_lib.Oracle()
        f, tot = ora.run(p["nx"], p["ny"], steps, p["density"], p["accel"], p["omega"], mask)
        rows = [lbm.allocate_rows(p["ny"], size, r) for r in range(size)]
        start, n = rows[rank]
        local = np.ascontiguousarray(f[:, start:start + n])
        # per-rank partial of the last step = sum of |u| over the rank's fluid cells; the velocities of the
        # pulled state equal those of the post-collision state (BGK conserves momentum), so the final
        # state's |u| plane gives them
        _, _, u, _ = ora.final_state(f, p["density"], mask)
        partial = np.array([u[start:start + n][mask[start:start + n] == 0].astype(np.float64).sum()])
        inv = np.float32(1.0) / np.float32(p["nx"] * p["ny"] - lines)
        av = ring.reduce_av_vels(dist, torch, partial, inv)
        expect = np.float32(tot[-1]) * inv
        assert abs(float(av[0]) - float(expect)) / float(expect) < 1e-4, (av, expect)
        g = ring.gather_rows(dist, torch, local, [r[1] for r in rows], rank)
        if rank == 0:
            assert np.array_equal(g, f)
        out.put((rank, "ok"))
    except Exception as e:   # report to the parent instead of hanging the group
        out.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_ring_over_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for pr in procs:
        pr.start()
    res = [out.get(timeout=240) for _ in procs]
    for pr in procs:
        pr.join(60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res

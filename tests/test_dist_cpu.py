"""CPU, world_size 2, gloo: the multi-GPU path of LinkSimulator.sweep -- contiguous frame
shards per rank, then a SUM all-reduce of the counter table -- with the per-rank kernel
chain replaced by a deterministic stand-in (no GPU here).  The union over ranks must be
identical to a single-rank run, for even and uneven splits."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from waveforms_b200 import linksim


def _fake_run(self, ebn0_db, n_frames, first_frame=0, active=None):
    """Counters that depend on every (point, frame) pair individually, like the real chain
    (same signature as LinkSimulator.run)."""
    eb = np.atleast_1d(np.asarray(ebn0_db, dtype=np.float64))
    out = np.zeros((eb.size, 3), dtype=np.int64)
    for p in range(eb.size):
        if active is not None and not active[p]:
            continue
        f = np.arange(first_frame, first_frame + n_frames, dtype=np.int64)
        out[p, 0] = int(np.sum((f * 2654435761 + p * 40503) % 97))
        out[p, 1] = int(np.sum((f * 40503 + p) % 89))
        out[p, 2] = n_frames * 4093
    return out


def _worker(rank, world, port, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    linksim.LinkSimulator.run = _fake_run
    sim = object.__new__(linksim.LinkSimulator)
    sim.device = "cpu"
    got = sim.sweep([0.0, 6.0, 12.0], n_frames, first_frame=5)
    q.put((rank, got))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_world(n_frames, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


def test_sweep_two_ranks_equals_one_rank():
    for n_frames in (64, 7, 1):
        want = _fake_run(None, [0.0, 6.0, 12.0], n_frames, 5)
        res = _run_world(n_frames)
        assert np.array_equal(res[0], want) and np.array_equal(res[1], want), n_frames


def test_sweep_without_process_group_is_run():
    linksim_run = linksim.LinkSimulator.run
    try:
        linksim.LinkSimulator.run = _fake_run
        sim = object.__new__(linksim.LinkSimulator)
        assert np.array_equal(sim.sweep([3.0], 10, 2), _fake_run(None, [3.0], 10, 2))
    finally:
        linksim.LinkSimulator.run = linksim_run


# ---------------------------------------------------------------- waterfall with a stop rule
def _fake_run_active(self, ebn0_db, n_frames, first_frame=0, active=None):
    """Like _fake_run, but honours the `active` mask of LinkSimulator.waterfall; errors thin out
    with the point index like a BER curve does."""
    eb = np.atleast_1d(np.asarray(ebn0_db, dtype=np.float64))
    out = np.zeros((eb.size, 3), dtype=np.int64)
    for p in range(eb.size):
        if active is not None and not active[p]:
            continue
        f = np.arange(first_frame, first_frame + n_frames, dtype=np.int64)
        out[p, 0] = int(np.sum(((f * 2654435761 + p * 40503) % 1000) < (300 >> (3 * p))))
        out[p, 2] = n_frames * 4093
    return out


def _waterfall_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    linksim.LinkSimulator.run = _fake_run_active
    sim = object.__new__(linksim.LinkSimulator)
    sim.device, sim.max_batch, sim.n_compared = "cpu", 8, 4093
    got = sim.waterfall([0.0, 4.0, 8.0, 12.0], min_errors=50, max_bits=4093 * 640, chunk_frames=32, first_frame=3)
    q.put((rank, got))
    dist.barrier()
    dist.destroy_process_group()


def _waterfall_single():
    run = linksim.LinkSimulator.run
    try:
        linksim.LinkSimulator.run = _fake_run_active
        sim = object.__new__(linksim.LinkSimulator)
        sim.device, sim.max_batch, sim.n_compared = "cpu", 8, 4093
        return sim.waterfall([0.0, 4.0, 8.0, 12.0], min_errors=50, max_bits=4093 * 640, chunk_frames=32, first_frame=3)
    finally:
        linksim.LinkSimulator.run = run


def test_waterfall_stop_rule():
    got = _waterfall_single()
    # every point stopped for one of the two reasons, in whole chunks of 32 frames
    assert np.all((got[:, 0] >= 50) | (got[:, 2] >= 4093 * 640))
    assert np.all(got[:, 2] % (32 * 4093) == 0)
    # the easy point stops as soon as it has its errors, the hard points run to the bit budget
    assert got[0, 0] >= 50 and got[0, 2] == 6 * 32 * 4093
    assert np.all(got[1:, 2] == 640 * 4093) and np.all(got[1:, 0] < 50)
    assert got[1, 0] > got[2, 0] > got[3, 0] == 0


def test_waterfall_two_ranks_equals_one_rank():
    want = _waterfall_single()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_waterfall_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(res[0], want) and np.array_equal(res[1], want)


# ---------------------------------------------------------------- work list of LinkSimulator.run
def test_run_work_list_covers_every_point_frame_pair_once():
    """run() cuts the point-major (point, frame) list into batches that ignore point boundaries:
    every pair appears exactly once, in order, for any batch size; inactive points get no rows."""
    pts = np.array([0, 2, 3])  # point 1 inactive
    n_frames, first = 37, 1000
    total = pts.size * n_frames
    for max_batch in (1, 5, 37, 40, 64, 500):
        seen = []
        for w0 in range(0, total, max_batch):
            pw, fw = linksim.work_slice(pts, n_frames, first, w0, min(max_batch, total - w0))
            seen += list(zip(pw.tolist(), fw.tolist()))
        assert seen == [(p, first + f) for p in pts for f in range(n_frames)], max_batch


def test_run_bit_offsets_are_exact_for_far_frames():
    sim = object.__new__(linksim.LinkSimulator)
    sim.cfg = linksim.LinkConfig(n_bits=4096)
    for degree in (15, 23, 33):
        sim.period = (1 << degree) - 1
        frames = np.array([0, 1, 7, 123456789, 2 ** 40 + 3], dtype=np.int64)
        want = [(int(f) * 4096) % sim.period for f in frames]
        assert sim._bit_offsets(frames).tolist() == want, degree

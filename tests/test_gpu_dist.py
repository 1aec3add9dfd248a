"""GPU, world_size 2, NCCL: the REAL kernel chain sharded over two B200s.  Each rank runs
LinkSimulator.sweep() on its contiguous share of the frames and the int64 counter table is
all-reduced over NVLink; the result must equal a single-GPU run of the same (point, frame) list
bit for bit (PRBS offsets and Philox counters are functions of the GLOBAL frame id), for even and
uneven splits and for the waterfall's stop rule.  Skipped on a box with one GPU
(tests/test_dist_cpu.py covers the host logic there); run with `gpurun --gpus 2`."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

EBN0 = [2.0, 6.0, 9.0]
CASES = [("soqpsk-tg", "PAM", {}, 257), ("soqpsk-tg", "PAM", {"fused": True}, 64), ("soqpsk-tg", "PT", {}, 33),
         ("pcmfm", "PAM", {}, 31), ("multih", "PAM", {}, 18)]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist

    from waveforms_b200.linksim import LinkConfig, LinkSimulator

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    out = {}
    for wf, det, extra, n_frames in CASES:
        sim = LinkSimulator(LinkConfig(waveform=wf, detector=det, n_bits=1024, extra=extra), device=dev, max_batch=48)
        out[(wf, det, bool(extra), n_frames)] = sim.sweep(EBN0, n_frames, first_frame=1000)
    sim = LinkSimulator(LinkConfig(n_bits=512), device=dev, max_batch=32)
    out["waterfall"] = sim.waterfall([0.0, 5.0, 9.0], min_errors=400, max_bits=512 * 64 * 40, chunk_frames=64, first_frame=7)
    q.put((rank, out))
    dist.barrier(device_ids=[rank])
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two CUDA devices")
def test_two_gpu_counters_equal_one_gpu_counters():
    import torch.multiprocessing as mp

    from waveforms_b200.linksim import LinkConfig, LinkSimulator

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=600) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    dev = torch.device("cuda", 0)
    for wf, det, extra, n_frames in CASES:
        sim = LinkSimulator(LinkConfig(waveform=wf, detector=det, n_bits=1024, extra=extra), device=dev, max_batch=48)
        want = sim.run(EBN0, n_frames, first_frame=1000)
        key = (wf, det, bool(extra), n_frames)
        assert want[:, 2].min() == n_frames * sim.n_compared and want[0, 0] > want[2, 0]
        assert np.array_equal(res[0][key], want) and np.array_equal(res[1][key], want), (key, res[0][key], want)
    sim = LinkSimulator(LinkConfig(n_bits=512), device=dev, max_batch=32)
    want = sim.waterfall([0.0, 5.0, 9.0], min_errors=400, max_bits=512 * 64 * 40, chunk_frames=64, first_frame=7)
    assert np.array_equal(res[0]["waterfall"], want) and np.array_equal(res[1]["waterfall"], want)
    assert want[0, 2] < want[2, 2]  # the easy point retired earlier than the hard one

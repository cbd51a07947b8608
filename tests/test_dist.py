"""Host logic of the data-parallel path (matcha_b200/dist.py): partitioning and the final gather, world_size 2 over gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from matcha_b200 import dist as mdist


def test_partition_balances_frames_and_covers_everything():
    lengths = [451, 120, 300, 299, 451, 80, 10, 400, 222, 37]
    for world in (1, 2, 3, 4, 8):
        parts = mdist.partition(lengths, world)
        assert len(parts) == world
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(len(lengths)))
        loads = [sum(lengths[i] for i in p) for p in parts]
        assert max(loads) - min(loads) <= max(lengths)          # LPT bound
        for p in parts:                                         # sorted by decreasing length inside a rank
            assert [lengths[i] for i in p] == sorted((lengths[i] for i in p), reverse=True)
    assert mdist.partition([], 2) == [[], []]
    assert mdist.partition([5], 4) == [[0], [], [], []]
    with pytest.raises(ValueError):
        mdist.partition([1], 0)


def test_partition_is_deterministic_on_ties():
    assert mdist.partition([7, 7, 7, 7], 2) == [[0, 2], [1, 3]]


def test_micro_batches_respect_the_frame_budget():
    lengths = [400, 390, 200, 190, 50, 40, 30]
    idx = list(range(len(lengths)))
    for budget in (400, 800, 1200, 10 ** 6):
        mbs = mdist.micro_batches(idx, lengths, budget)
        assert [i for mb in mbs for i in mb] == idx
        for mb in mbs:
            assert len(mb) == 1 or len(mb) * max(lengths[i] for i in mb) <= budget


def test_single_process_gather_reorders_rows():
    local = torch.arange(3 * 2 * 4, dtype=torch.float32).reshape(3, 2, 4)
    out = mdist.gather_padded(local, [2, 0, 1], 3)
    assert torch.equal(out[2], local[0]) and torch.equal(out[0], local[1]) and torch.equal(out[1], local[2])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_path(mu, lens):
    """Stand-in for the CUDA hot path with the same shape contract: (b, 80, T) -> (b, 1, 4 * T_local_max)."""
    t_max = int(lens.max())
    wav = mu[:, :1, :t_max].repeat_interleave(4, dim=2) * 0.5
    for i, n in enumerate(lens.tolist()):
        wav[i, :, 4 * n:] = 0
    return wav


def _worker(rank, world, port, lengths, n_empty_ok, result):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        n = len(lengths)
        mu = torch.randn(n, 80, max(lengths), generator=g)
        lens = torch.tensor(lengths)
        out = mdist.synthesise_sharded(_fake_path, lengths, mu, lens, dst=0)
        if rank == 0:
            ref = _fake_path(mu, lens)
            assert out.shape == ref.shape
            assert torch.equal(out, ref)
            result.put("ok")
        else:
            assert out is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("lengths", [[30, 12, 25, 7, 19], [9], [16, 16, 16, 16]])
def test_sharded_path_world2_matches_single_process(lengths):
    ctx = mp.get_context("spawn")
    result = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, True, result)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert result.get(timeout=5) == "ok"


def _grad_worker(rank, world, port, result):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(5)
        shapes = [(256, 160, 3), (256,), (1024, 256), (1024,), (80, 256, 1), (7,)]
        per_rank = [[torch.randn(*s, generator=g) for s in shapes] for _ in range(world)]
        frames = [torch.tensor(float(100 + 37 * r)) for r in range(world)]
        for weighted in (False, True):
            params = [torch.nn.Parameter(torch.zeros(*s)) for s in shapes]
            for p, gr in zip(params, per_rank[rank]):
                p.grad = gr.clone()
            params.append(torch.nn.Parameter(torch.zeros(3)))           # a parameter without a gradient is skipped
            mdist.allreduce_gradients(params, bucket_bytes=300_000, valid_frames=frames[rank] if weighted else None)
            tot = sum(float(f) for f in frames)
            for i, p in enumerate(params[:-1]):
                if weighted:
                    want = sum(per_rank[r][i] * (float(frames[r]) / tot) for r in range(world))
                else:
                    want = sum(per_rank[r][i] for r in range(world)) / world
                assert torch.allclose(p.grad, want, atol=1e-6), (weighted, i)
            assert params[-1].grad is None
        if rank == 0:
            result.put("ok")
    finally:
        dist.destroy_process_group()


def test_gradient_allreduce_world2_buckets_and_frame_weighting():
    ctx = mp.get_context("spawn")
    result = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, result)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert result.get(timeout=5) == "ok"

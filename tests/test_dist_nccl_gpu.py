"""The sharded synthesis path on real GPUs: `dist.synthesise_sharded` over NCCL with the real `Plan.synthesise`, world
size 2, against the single-process result on the same inputs.  Utterances are independent and every
kernel is run-to-run deterministic, but a rank's share is a smaller batch with its own T_pad: results depend on the batch's
padded length, so both runs pad every utterance to the same T_pad and the outputs must be
EQUAL.  Needs two GPUs; skipped on a one-GPU box."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

T_PAD, N_STEPS, MEAN, STD = 64, 3, -5.536622, 2.116101
LENS = [64, 21, 50, 64, 33, 8, 47]


def _inputs():
    from matcha_b200 import synthetic as weights
    g = torch.Generator().manual_seed(5)
    n = len(LENS)
    return (torch.randn(n, 80, T_PAD, generator=g), torch.randn(n, 80, T_PAD, generator=g) * 0.667,
            weights.prefix_mask(LENS, T_PAD))


def _path(plan):
    def fn(mu, z, mask):
        # the vocoder sees the global longest length, as the single-process call does
        _, _, wav = plan.synthesise(mu, mask, z, N_STEPS, max(LENS), MEAN, STD)
        return wav
    return fn


def _plan():
    from matcha_b200 import synthetic as weights
    from matcha_b200._lib import Plan
    plan = Plan("fp16", 160)
    plan.load_estimator(weights.estimator_state_dict(160, seed=3))
    plan.load_generator(weights.generator_state_dict(4))
    return plan


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from matcha_b200 import dist as mdist
        mu, z, mask = (t.cuda() for t in _inputs())
        wav = mdist.synthesise_sharded(_path(_plan()), LENS, mu, z, mask, dst=0)
        if rank == 0:
            torch.save(wav.cpu(), out_path)
        else:
            assert wav is None
    finally:
        dist.destroy_process_group()


def test_sharded_synthesis_over_nccl_equals_single_process(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out_path = str(tmp_path / "wav.pt")
    mp.spawn(_worker, args=(2, port, out_path), nprocs=2, join=True)
    sharded = torch.load(out_path)
    mu, z, mask = (t.cuda() for t in _inputs())
    single = _path(_plan())(mu, z, mask).cpu()
    assert sharded.shape == single.shape == (len(LENS), 1, 256 * max(LENS))
    assert torch.equal(sharded, single)

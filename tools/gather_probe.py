"""Times the bench's final exchange (NCCL gather of mel + wav to rank 0) alone, per transport, at world_size N.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/gather_probe.py
"""
import os
import torch
import torch.distributed as dist

B, T = 64, 451
local = int(os.environ["LOCAL_RANK"]); rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
mel = torch.randn(B, 80, T, device=dev)
wav = torch.randn(B, 1, 256 * T, device=dev)
gm = [torch.empty_like(mel) for _ in range(world)] if rank == 0 else None
gw = [torch.empty_like(wav) for _ in range(world)] if rank == 0 else None
am = torch.empty(world, *mel.shape, device=dev)
aw = torch.empty(world, *wav.shape, device=dev)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def f_gather():
    dist.gather(mel, gm, dst=0); dist.gather(wav, gw, dst=0)


def f_allgather():
    dist.all_gather_into_tensor(am, mel); dist.all_gather_into_tensor(aw, wav)


def f_sendrecv():
    if rank == 0:
        for r in range(1, world):
            dist.recv(am[r], src=r); dist.recv(aw[r], src=r)
    else:
        dist.send(mel, dst=0); dist.send(wav, dst=0)


for name, fn in (("gather", f_gather), ("all_gather_into_tensor", f_allgather), ("send/recv", f_sendrecv)):
    ms = timed(fn)
    if rank == 0:
        print(f"[gather_probe] world={world} {name}: {ms:.3f} ms per step ({(mel.numel()+wav.numel())*4/1e6:.1f} MB per rank)", flush=True)
can = torch.cuda.can_device_access_peer(0, 1) if torch.cuda.device_count() > 1 else None
if rank == 0:
    print("[gather_probe] can_device_access_peer(0,1) =", can, flush=True)
dist.destroy_process_group()

#!/usr/bin/env python
"""All-reduce of the view-parallel gradient buffer ([P,61] fp32 = 732 MB at C4) on N GPUs: one call vs chunks, for the
NCCL algorithm selected by the environment (NCCL_ALGO=Ring|Tree|NVLS|NVLSTree; unset = NCCL's choice).
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/allreduce_probe.py [P_gaussians]"""
import json
import os
import sys

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
buf = torch.ones(P * 61, dtype=torch.float32, device=dev)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def chunks(n):
    step = (buf.numel() + n - 1) // n
    ws = [dist.all_reduce(buf[i:i + step], async_op=True) for i in range(0, buf.numel(), step)]
    for w in ws:
        w.wait()


out = dict(world=world, bytes=buf.numel() * 4, algo=os.environ.get("NCCL_ALGO", "default"))
out["one_call_ms"] = timed(lambda: dist.all_reduce(buf))
for n in (4, 8, 16):
    out[f"chunks{n}_ms"] = timed(lambda: chunks(n))
out["busbw_gbs"] = out["bytes"] * 2 * (world - 1) / world / (out["one_call_ms"] * 1e-3) / 1e9
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()

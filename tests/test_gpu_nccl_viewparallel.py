"""NCCL parity of the view-parallel step on real hardware (needs >= 2 GPUs; skipped on a 1-GPU box): the gradient buffer
every rank holds after `ViewParallelRasterizer.step` on 2 ranks -- views sharded, rows all-reduced over NCCL underneath
the deferred SH backward -- equals the buffer ONE GPU produces for the whole batch (SURVEY 8(e) "Semantics note":
all-reduced gradient = sum over the single-view gradients computed on one GPU; the one-GPU batch step is itself checked
against the sum of the oracle's per-view gradients in test_gpu_viewparallel.py).  Also covers the parameter replicas
completed by the in-place all-gather."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from pdgs_b200 import scenes

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")]
N_VIEWS = 7   # 4 + 3 views: uneven shards


def _inputs(dev, split):
    sc = scenes.make_config("C4", P=6000, n_views=N_VIEWS, scale_image=0.1)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(21)
    targets = {i: torch.rand(3, H, W, generator=gen).to(dev) for i in range(N_VIEWS)}
    g = {k: getattr(sc, k).to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    if split:
        g = dict(means3D=g["means3D"], opacities=g["opacities"], scales=g["scales"], rotations=g["rotations"],
                 shs_dc=g["shs"][:, :1].contiguous(), shs_rest=g["shs"][:, 1:].contiguous())
    views = [helpers.cuda_settings(sc, v, dev) for v in sc.views]
    return sc, g, views, targets


def _worker(rank, world, port, split, chunks, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PDGS_REDUCE_CHUNKS=str(chunks))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc, g, views, targets = _inputs(dev, split)
    vp = ViewParallelRasterizer(sc.P, 16, dev)
    out = vp.step(g, views, targets, lambda_dssim=0.2)
    torch.cuda.synchronize(dev)
    # the staged-parameter path: every rank uploads 1/world, an all-gather completes the replicas
    host = {k: v.cpu().pin_memory() for k, v in g.items()}
    staged, ev = vp.stage_gaussians(host)
    if ev is not None:
        ev.synchronize()
    replicas_ok = all(torch.equal(staged[k].cpu(), host[k]) for k in host)
    q.put((rank, out["local_views"], vp.buf.flat.cpu().numpy(), vp.buf.max_radii2D.cpu().numpy(), vp.loss_acc.cpu().numpy(),
           replicas_ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("split,chunks", [(False, 1), (True, 3)])
def test_two_ranks_over_nccl_equal_one_gpu(split, chunks):
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    dev = torch.device("cuda", 0)
    sc, g, views, targets = _inputs(dev, split)
    single = ViewParallelRasterizer(sc.P, 16, dev, lanes=1)
    single.step(g, views, targets, lambda_dssim=0.2)
    torch.cuda.synchronize(dev)
    ref_flat, ref_radii, ref_loss = single.buf.flat.cpu(), single.buf.max_radii2D.cpu(), single.loss_acc.cpu()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, split, chunks, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=300) for _ in range(2)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][1] == [0, 1, 2, 3] and got[1][1] == [4, 5, 6]
    for rank, _, flat, radii, loss, replicas_ok in got:
        flat = torch.from_numpy(flat)
        assert replicas_ok, rank
        assert helpers.rel_err(flat, ref_flat) < 1e-5, rank          # fp32 sums in a different order
        assert torch.equal(torch.from_numpy(radii), ref_radii), rank
        assert torch.allclose(torch.from_numpy(loss), ref_loss, rtol=1e-5, atol=1e-7), rank
    assert torch.equal(torch.from_numpy(got[0][2]), torch.from_numpy(got[1][2]))   # both ranks hold the same bits

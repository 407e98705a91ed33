"""CPU tests of the view-parallel host logic: sharding, flat gradient buffer layout and the
all-reduce, with world_size 2 over gloo.  The per-view backend injected here is the float64 oracle (test
infrastructure); the product backend is the CUDA library and is covered by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from pdgs_b200 import scenes
from pdgs_b200.viewparallel import GradBuffer, ViewParallelRasterizer, shard_views


def test_shard_views_partitions_contiguously():
    for n in (0, 1, 7, 64, 65):
        for world in (1, 2, 3, 8):
            parts = [list(shard_views(n, r, world)) for r in range(world)]
            assert sum(parts, []) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_overlapped_reduce_slices_cover_the_buffer_once():
    """The chunked multi-GPU reduce all-reduces tail_slice() plus, per chunk of rows, row_slices(lo, hi): together they
    must cover every float of the flat gradient buffer exactly once (padding words aside)."""
    for P, split in ((1000, False), (1000, True), (777, False), (5, True)):
        b = GradBuffer(P, 16, "cpu")
        b.flat.zero_()
        b.tail_slice().add_(1.0)
        step = 256
        for lo in range(0, P, step):
            for t in b.row_slices(lo, min(P, lo + step), split):
                t.add_(1.0)
        for name, t in b.accumulate_dict(split).items():
            if name != "max_radii2D":
                assert bool((t == 1.0).all()), (P, split, name)
        assert float(b.flat.max()) == 1.0


def test_parameter_upload_shares_cover_the_tensor():
    from pdgs_b200.viewparallel import shard_span
    for numel in (0, 1, 7, 48, 9_000_001):
        for world in (1, 2, 3, 8):
            spans = [shard_span(numel, r, world) for r in range(world)]
            chunk = spans[0][0]
            assert all(s[0] == chunk for s in spans) and world * chunk >= numel
            covered = 0
            for r, (_, lo, hi) in enumerate(spans):
                assert lo == min(r * chunk, numel) and lo <= hi <= numel and hi - lo <= chunk
                assert lo == covered or hi == lo  # contiguous, in rank order
                covered = max(covered, hi)
            assert covered == numel


def test_grad_buffer_layout():
    for P in (0, 1, 5, 1001):
        b = GradBuffer(P, 16, "cpu")
        assert b.floats_per_gaussian == 61
        seen = torch.zeros_like(b.flat)
        for name, t in b.accumulate_dict().items():
            if name == "max_radii2D":
                assert t.dtype == torch.int32 and t.shape == (P,)
                continue
            assert t.data_ptr() % 16 == 0, name
            t += 1.0
        assert float(b.flat.sum()) == 61 * P  # segments are disjoint and cover 61 floats per Gaussian
        assert b.view("shs").shape == (P, 16, 3) and b.view("means3D").shape == (P, 3)
        assert b.view("opacities").shape == (P, 1) and b.view("densify_denom").shape == (P,)
        # split layout (features_dc / features_rest): same 61 floats, both tensors 16-byte aligned and disjoint
        b2 = GradBuffer(P, 16, "cpu")
        d2 = b2.accumulate_dict(split=True)
        assert "shs" not in d2 and d2["shs_dc"].shape == (P, 1, 3) and d2["shs_rest"].shape == (P, 15, 3)
        for name, t in d2.items():
            if name != "max_radii2D":
                assert t.data_ptr() % 16 == 0, name
                t += 1.0
        assert float(b2.flat.sum()) == 61 * P


def _oracle_backend(scene):
    def run(gaussians, view, target, prior, mask, w_depth, acc, loss_acc):
        o, res, st = helpers.run_oracle(scene, view)
        gc = torch.sign(res.color - target.double()) / target.numel()
        g = o.backward(gc, None)
        for k in ("means3D", "shs", "opacities", "scales", "rotations"):
            acc[k] += g[k].view_as(acc[k]).float()
        vis = torch.from_numpy(st["radii"] > 0)
        acc["densify_grad_accum"][vis] += g["means2D"][vis, :2].norm(dim=-1).float()
        acc["densify_denom"][vis] += 1
        acc["max_radii2D"][:] = torch.maximum(acc["max_radii2D"], torch.from_numpy(st["radii"]))
        loss_acc[0] += float((res.color - target.double()).abs().mean())
        return dict(num_rendered=st["num_rendered"], launches=0)
    return run


def _scene():
    return scenes.make_config("C4", P=300, n_views=5, scale_image=0.04)


def _targets(sc):
    out = {}
    for i, v in enumerate(sc.views):
        gen = torch.Generator().manual_seed(50 + i)
        out[i] = torch.rand(3, v.image_height, v.image_width, generator=gen)
    return out


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    sc = _scene()
    vp = ViewParallelRasterizer(sc.P, 16, "cpu", backend=_oracle_backend(sc))
    out = vp.step({}, sc.views, _targets(sc))
    q.put((rank, out["local_views"], vp.buf.flat.clone().numpy(), vp.buf.max_radii2D.clone().numpy(),
           float(vp.loss_acc[0])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_allreduce_equals_single_process_sum():
    sc = _scene()
    single = ViewParallelRasterizer(sc.P, 16, "cpu", backend=_oracle_backend(sc))
    single.step({}, sc.views, _targets(sc))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == [0, 1, 2] and res[1][1] == [3, 4]
    for r in res:
        np.testing.assert_allclose(r[2], single.buf.flat.numpy(), rtol=1e-5, atol=1e-9)   # fp32 reassociation only
        np.testing.assert_array_equal(r[3], single.buf.max_radii2D.numpy())
        assert abs(r[4] - float(single.loss_acc[0])) < 1e-5
    assert float(single.buf.view("densify_denom").max()) <= 5 and float(single.buf.view("densify_denom").sum()) > 0


def test_cuda_backend_is_mandatory_without_injection():
    with pytest.raises(RuntimeError, match="no CPU path"):
        ViewParallelRasterizer(10, 16, "cpu")


def test_prefetched_images_are_used_only_when_every_tensor_is_this_steps():
    """prefetch_images hand-over: the upload staged during the previous step replaces this step's upload of a lane's
    first view only if target, prior, mask and alpha mask are all the very tensors this step passes; otherwise the
    step's own arguments win (a stale or incomplete prefetch must not drop the depth-prior term or the alpha mask)."""
    seen = []

    def backend(gaussians, view, target, prior, mask, w_depth, acc, loss_acc):
        seen.append((target, prior, mask))
        return dict(num_rendered=0, launches=0)

    sc = scenes.make_config("C4", P=8, n_views=2, scale_image=0.02)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    tg = {i: torch.rand(3, H, W) for i in range(2)}
    pr = {i: torch.rand(1, H, W) for i in range(2)}
    vp = ViewParallelRasterizer(sc.P, 16, "cpu", backend=backend)
    # the prefetch names the targets only, the next step also passes priors: staged upload must be dropped
    vp.step({}, sc.views, tg, prefetch_images=dict(targets=tg))
    assert 0 in vp._pre_images
    seen.clear()
    vp.step({}, sc.views, tg, priors=pr)
    assert seen[0][0] is tg[0] and seen[0][1] is pr[0]
    assert not vp._pre_images
    # a complete prefetch is taken
    vp.step({}, sc.views, tg, priors=pr, prefetch_images=dict(targets=tg, priors=pr))
    staged = vp._pre_images[0][1]
    seen.clear()
    vp.step({}, sc.views, tg, priors=pr)
    assert seen[0][0] is staged[0][0] and seen[0][1] is staged[0][1] and seen[0][1] is pr[0]
    # priors replaced since the prefetch: this step's tensors are used
    vp.step({}, sc.views, tg, priors=pr, prefetch_images=dict(targets=tg, priors=pr))
    pr2 = {i: torch.rand(1, H, W) for i in range(2)}
    seen.clear()
    vp.step({}, sc.views, tg, priors=pr2)
    assert seen[0][1] is pr2[0]

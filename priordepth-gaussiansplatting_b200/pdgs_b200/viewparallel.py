"""View-parallel batch step over the rasterizer.

The reference renders one camera per optimiser step (/root/reference/train.py:129-143, 184) and keeps
the densification statistics with torch ops on the caller side (scene/gaussian_model.py:502-506,
train.py:313-314).  BASELINE.json config 4 instead renders a BATCH of cameras: the views are sharded
contiguously over the ranks (one process per GPU), the Gaussians are replicated, every rank
accumulates the gradients of its views IN PLACE inside the preprocess-backward kernel
(PdgsGrads.accumulate = 1) into one flat fp32 buffer, and a single all-reduce (NCCL over NVLink on the
GPUs, gloo in the CPU tests) sums that buffer over ranks.  The result equals what repeated
`loss.backward()` calls without `zero_grad` produce on one GPU.

The rasterizer backend is libpdgs.so through `pdgs_b200.rasterizer`; there is no CPU path.  The
backend argument exists so the host logic (sharding, buffer layout, collectives) can be exercised on
CPU tensors with `gloo` by the tests, which inject their own per-view gradient function.
"""
from __future__ import annotations

import os
import threading
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

# segments of the flat gradient buffer: (name, floats per Gaussian); shs_rest uses (M-1)*3.  The SH gradient occupies
# the adjacent shs_dc + shs_rest segments: as two tensors for a split scene (features_dc / features_rest), or as ONE
# packed [P,M,3] tensor ("shs") laid over the start of the same region.
_SEGMENTS = (("means3D", 3), ("shs_dc", 3), ("shs_rest", None), ("opacities", 1), ("scales", 3), ("rotations", 4),
             ("densify_grad_accum", 1), ("densify_denom", 1))


def shard_views(n_views: int, rank: int, world: int) -> range:
    """Contiguous split: rank r renders views [r*B/n, (r+1)*B/n); remainders go to the low ranks."""
    base, rem = divmod(int(n_views), int(world))
    lo = rank * base + min(rank, rem)
    return range(lo, lo + base + (1 if rank < rem else 0))


def shard_span(numel: int, rank: int, world: int):
    """(chunk, lo, hi): rank r uploads elements [lo, hi) of a flat parameter tensor; chunks are equal (all-gather needs
    equal counts), so the padded buffer holds world * chunk >= numel elements and the last shares may be short or empty."""
    chunk = (int(numel) + world - 1) // world
    return chunk, min(rank * chunk, numel), min((rank + 1) * chunk, numel)


class GradBuffer:
    """One flat fp32 buffer holding every per-Gaussian gradient plus the densification statistics
    (61 floats per Gaussian for SH degree 3), each segment starting on a 16-byte boundary so the kernels'
    float4 accesses stay aligned, and an int32 max_radii2D vector (reduced with MAX)."""

    def __init__(self, P: int, M: int, device):
        self.P, self.M = int(P), int(M)
        off = 0
        self._spans: Dict[str, tuple] = {}
        for name, width in _SEGMENTS:
            w = 3 * (self.M - 1) if width is None else width
            self._spans[name] = (off, self.P * w, w)
            off += (self.P * w + 3) // 4 * 4
        self._spans["shs"] = (self._spans["shs_dc"][0], self.P * 3 * self.M, 3 * self.M)
        self.flat = torch.zeros(max(off, 4), dtype=torch.float32, device=device)
        self.max_radii2D = torch.zeros(self.P, dtype=torch.int32, device=device)

    def view(self, name: str) -> torch.Tensor:
        off, n, w = self._spans[name]
        t = self.flat[off:off + n]
        if name == "shs":
            return t.view(self.P, self.M, 3)
        if name == "shs_dc":
            return t.view(self.P, 1, 3)
        if name == "shs_rest":
            return t.view(self.P, max(self.M - 1, 0), 3)
        return t.view(self.P, w) if w > 1 or name == "opacities" else t.view(self.P)

    def zero_(self, skip_sh: bool = False) -> "GradBuffer":
        """skip_sh: leave the SH gradient region alone (the deferred SH backward overwrites every row of it)."""
        if skip_sh and self.P > 0:
            a = self._spans["shs_dc"][0]
            b = self._spans["shs_rest"][0] + (self._spans["shs_rest"][1] + 3) // 4 * 4
            self.flat[:a].zero_()
            self.flat[b:].zero_()
        else:
            self.flat.zero_()
        self.max_radii2D.zero_()
        return self

    def accumulate_dict(self, split: bool = False) -> Dict[str, torch.Tensor]:
        """split: the SH gradient as shs_dc / shs_rest (scene given as features_dc / features_rest), else packed `shs`."""
        d = {name: self.view(name) for name, _ in _SEGMENTS if not name.startswith("shs")}
        if split:
            d["shs_dc"], d["shs_rest"] = self.view("shs_dc"), self.view("shs_rest")
        else:
            d["shs"] = self.view("shs")
        d["max_radii2D"] = self.max_radii2D
        return d

    def all_reduce(self, group=None) -> None:
        """The only data-path collective of the view-parallel step."""
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(self.max_radii2D, op=dist.ReduceOp.MAX, group=group)

    def tail_slice(self) -> torch.Tensor:
        """The contiguous part of the flat buffer behind the SH gradient: opacities | scales | rotations | densify
        statistics (10 floats per Gaussian).  Nothing writes it after the last view's preprocess backward."""
        return self.flat[self._spans["opacities"][0]:]

    def row_slices(self, lo: int, hi: int, split: bool) -> List[torch.Tensor]:
        """Flat slices holding rows [lo, hi) of the segments the deferred SH backward writes: means3D and the SH
        gradient (packed, or features_dc + features_rest)."""
        out = []
        for name in ("means3D",) + (("shs_dc", "shs_rest") if split else ("shs",)):
            off, n, w = self._spans[name]
            if w > 0 and hi > lo:
                out.append(self.flat[off + lo * w: off + hi * w])
        return out

    @property
    def floats_per_gaussian(self) -> int:
        return sum((3 * (self.M - 1) if w is None else w) for _, w in _SEGMENTS)


def _cuda_view_backend(gaussians: Dict[str, torch.Tensor], rs, target, prior, mask, w_depth, acc: Dict, loss_acc,
                       lambda_dssim: float = 0.0, scratch: Optional[Dict] = None, clamp: bool = True, alpha_mask=None):
    """forward -> image loss -> backward (accumulating) of ONE view on the CUDA library.  lambda_dssim > 0 selects
    the reference's training loss (1-l)*L1 + l*(1-SSIM) (/root/reference/train.py:153-159), else plain L1.  clamp /
    alpha_mask: the loss sees clamp(render, 0, 1) * alpha_mask like the reference's caller
    (gaussian_renderer/__init__.py:121, train.py:148-150); the clamp and its backward run inside the loss kernels."""
    from . import rasterizer as rz
    empty = torch.empty(0, device=gaussians["means3D"].device)
    # split SH (features_dc / features_rest) and raw parameters: SURVEY 8(f) rank 2, see rz.rasterize_gaussians
    dc = gaussians.get("shs_dc")
    shs = gaussians.get("shs_rest") if dc is not None else gaussians.get("shs")
    raw = bool(gaussians.get("raw_params", False))
    colors = gaussians.get("colors_precomp")
    num_rendered, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
        rs.bg, gaussians["means3D"], empty if colors is None else colors, gaussians["opacities"], gaussians["scales"],
        gaussians["rotations"], rs.scale_modifier, empty, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
        rs.image_height, rs.image_width, empty if shs is None else shs, rs.sh_degree, rs.campos, False, rs.antialiasing,
        rs.debug, dc=dc, raw_params=raw)
    launches = rz.LAUNCHES.forward
    use_depth = prior is not None and w_depth != 0.0
    if lambda_dssim > 0.0:
        buf = None if scratch is None else scratch.get("ssim")
        need = rz.ssim_scratch_bytes(rs.image_width, rs.image_height)
        if buf is None or buf.numel() < need:
            buf = torch.empty(need, dtype=torch.uint8, device=color.device)
            if scratch is not None:
                scratch["ssim"] = buf
        dcolor, ddepth, _ = rz.image_loss_ssim(color, target, invd if use_depth else None, prior if use_depth else None,
                                               mask if use_depth else None, lambda_dssim, w_depth, loss_acc, buf,
                                               clamp=clamp, alpha_mask=alpha_mask)
    else:
        dcolor, ddepth, _ = rz.image_loss(color, target, invd if use_depth else None, prior if use_depth else None,
                                          mask if use_depth else None, 1.0, w_depth, loss_acc, clamp=clamp,
                                          alpha_mask=alpha_mask)
    launches += rz.LAUNCHES.loss
    rz.rasterize_gaussians_backward(
        rs.bg, gaussians["means3D"], radii, empty if colors is None else colors, gaussians["opacities"],
        gaussians["scales"], gaussians["rotations"], rs.scale_modifier, empty, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
        rs.tanfovy, dcolor, ddepth, empty if shs is None else shs, rs.sh_degree, rs.campos, geom, num_rendered, binning,
        img, rs.antialiasing, rs.debug, accumulate_into=acc, dc=dc, raw_params=raw)
    launches += rz.LAUNCHES.backward
    return dict(num_rendered=num_rendered, launches=launches, radii=radii, color=color, invdepth=invd)


class ViewParallelRasterizer:
    """step(): fwd + loss + bwd of this rank's shard of a batch of views, then one all-reduce."""

    def __init__(self, P: int, M: int, device, group=None, backend: Optional[Callable] = None, lanes: Optional[int] = None):
        self.device = torch.device(device)
        self.group = group
        self.backend = backend or _cuda_view_backend
        if backend is None and self.device.type != "cuda":
            raise RuntimeError("ViewParallelRasterizer: the rasterizer needs a CUDA device (there is no CPU path)")
        self.buf = GradBuffer(P, M, self.device)
        self.loss_acc = torch.zeros(3, dtype=torch.float32, device=self.device)  # sums over views: L1, depth L1, SSIM
        self._scratch: Dict = {}
        self.last_launches = 0
        self.last_num_rendered: List[int] = []
        # views in flight at once (CUDA backend): each lane = one host thread + one stream; PDGS_LANES overrides
        self.lanes = max(1, int(lanes if lanes is not None else os.environ.get("PDGS_LANES", "3")))
        self._pool = None
        self._lane_streams = []
        self.h2d_bytes = 0  # bytes this object copied host -> device (parameters and per-view images)
        self._count_lock = threading.Lock()
        self._copy_stream = None
        self._color_grads = None
        self._pre_images: Dict[int, tuple] = {}  # view index -> ((host target, prior, mask, alpha mask), staged upload) for the NEXT step
        # chunks along P of the deferred SH backward when its rows are all-reduced behind it (world > 1); PDGS_REDUCE_CHUNKS.
        # Measured on 8 x B200 (profiles/README.md r02h): 1 chunk 17.99 ms/step, 4 chunks 18.31 -- NCCL loses more on the
        # smaller messages (732 MB in one call 1.80 ms, in 4 calls 2.07 ms) than the 0.5 ms SH kernel can hide, so the
        # default is one chunk: only the 10 non-SH floats per Gaussian are reduced underneath the SH kernel.
        self.reduce_chunks = max(1, int(os.environ.get("PDGS_REDUCE_CHUNKS", "1")))

    def _upload(self, i, targets, priors, masks, alphas=None):
        items = [targets[i], None if priors is None else priors.get(i), None if masks is None else masks.get(i),
                 None if alphas is None else alphas.get(i)]
        if self.device.type != "cuda" or all(t is None or t.device == self.device for t in items):
            return items, None
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            with self._count_lock:
                self.h2d_bytes += sum(t.numel() * t.element_size() for t in items if t is not None and t.device != self.device)
            items = [None if t is None else t.to(self.device, non_blocking=True) for t in items]
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        return items, ev

    def stage_gaussians(self, host: Dict[str, torch.Tensor]):
        """Copy a dict of (pinned) host tensors -- the Gaussian parameters of an upcoming step -- to the device and return a
        handle for step(gaussians=handle).  With several ranks each rank uploads its 1/world share and an all-gather over
        NVLink completes the replicas (so every rank must call this together).  To overlap the upload with a running step
        use step(prefetch_gaussians=...) instead, which slices the copy behind the per-view image uploads."""
        if self.device.type != "cuda":
            return dict(host), None
        plan = self._prefetch_begin(host)
        self._prefetch_slice(plan, 0, 1)
        return self._prefetch_end(plan)

    def _prefetch_begin(self, host: Dict[str, torch.Tensor]):
        """Device buffers for the next step's Gaussians.  With several ranks every rank uploads only its 1/world share of
        each tensor over PCIe; _prefetch_end completes the replicas with an all-gather over NVLink (the parameters are
        identical on all ranks, so sending them world times through the host's memory system is the slow way)."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        world, rank = self.world, self.rank
        plan = {}
        with torch.cuda.stream(self._copy_stream):  # allocated on the stream that fills them
            for name, src in host.items():
                if not torch.is_tensor(src):
                    plan[name] = src
                    continue
                n = src.numel()
                chunk, lo, hi = shard_span(n, rank, world)
                flat = torch.empty(max(world * chunk, 1), dtype=src.dtype, device=self.device)
                plan[name] = (flat, src.reshape(-1), n, chunk, lo, hi, tuple(src.shape))
        return plan

    def _prefetch_slice(self, plan, k: int, n_slices: int) -> None:
        with torch.cuda.stream(self._copy_stream):
            for name, p in plan.items():
                if not isinstance(p, tuple):
                    continue
                flat, src, n, chunk, lo, hi = p[:6]
                a, b = lo + (hi - lo) * k // n_slices, lo + (hi - lo) * (k + 1) // n_slices
                if b > a:
                    flat[a:b].copy_(src[a:b], non_blocking=True)
                    with self._count_lock:
                        self.h2d_bytes += (b - a) * src.element_size()

    def _prefetch_end(self, plan):
        world, rank = self.world, self.rank
        dev = {}
        with torch.cuda.stream(self._copy_stream):
            for name, p in plan.items():
                if not isinstance(p, tuple):
                    dev[name] = p
                    continue
                flat, src, n, chunk, lo, hi, shape = p
                if world > 1 and chunk > 0:  # in place: this rank's share already sits at its position in `flat`
                    dist.all_gather_into_tensor(flat[:world * chunk], flat[rank * chunk:(rank + 1) * chunk], group=self.group)
                dev[name] = flat[:n].view(shape)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        return dev, ev

    def _await(self, staged):
        items, ev = staged
        if ev is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for t in items:
                if t is not None:
                    t.record_stream(cur)
        return items

    @property
    def world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    @property
    def rank(self) -> int:
        return dist.get_rank(self.group) if dist.is_available() and dist.is_initialized() else 0

    def step(self, gaussians: Dict[str, torch.Tensor], views: Sequence, targets: Dict[int, torch.Tensor],
             priors: Optional[Dict[int, torch.Tensor]] = None, masks: Optional[Dict[int, torch.Tensor]] = None,
             w_depth: float = 0.0, reduce: bool = True, keep_outputs: bool = False, lambda_dssim: float = 0.0,
             prefetch_gaussians: Optional[Dict[str, torch.Tensor]] = None, clamp: bool = True,
             alpha_masks: Optional[Dict[int, torch.Tensor]] = None, prefetch_images: Optional[Dict] = None) -> Dict:
        """gaussians: means3D/shs/opacities/scales/rotations (post-activation, as render() passes them,
        /root/reference/gaussian_renderer/__init__.py:57-112); optionally `shs_dc` + `shs_rest` instead of `shs`
        (features_dc / features_rest, gradients then in buf.view("shs_dc") / ("shs_rest")) and `raw_params=True`
        (opacities / scales / rotations are the pre-activation parameters, gradients w.r.t. them); views: one GaussianRasterizationSettings per
        camera of the WHOLE batch (identical on every rank); targets/priors/masks: dict view index -> tensor
        for at least this rank's shard.  Returns summed-over-views gradients (all ranks) in self.buf.
        clamp (default True) / alpha_masks (view index -> [1,H,W]): the loss of every view is taken on
        clamp(render, 0, 1) * alpha_mask, what the reference's render() and training step do before l1_loss / ssim
        (gaussian_renderer/__init__.py:121, train.py:148-150).

        prefetch_gaussians: pinned HOST tensors of the NEXT step's Gaussians.  They are uploaded during this step, one
        slice per view, on the same copy stream and behind that view's image upload (host->device copies of different
        streams share one DMA queue, so a single 236 B/Gaussian copy issued up front would hold the images back and
        stall the kernels); with several ranks each rank uploads 1/world of every tensor and an all-gather over NVLink
        completes the replicas.  The handle comes back as out["prefetched"] and is passed as `gaussians` to the next
        step.  Every rank must pass prefetch_gaussians in the same steps (it contains a collective).

        prefetch_images: dict(targets=..., priors=..., masks=..., alpha_masks=...) of the NEXT step (pinned host tensors).
        Each lane uploads the images of the view it will start the next step with as soon as it has launched its last
        view of this step, so the next step's first kernels do not wait for a 20 MB image to cross PCIe.  It must carry
        EVERY per-view tensor the next step will pass (the same objects): a staged upload is used only when its target,
        prior, mask and alpha mask are identical to the next step's arguments for that view, otherwise it is dropped
        and the view is uploaded again."""
        if isinstance(gaussians, tuple):  # handle from stage_gaussians(): wait for the upload on the compute stream
            gaussians, ev = gaussians
            if ev is not None:
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                for t in gaussians.values():
                    if torch.is_tensor(t):
                        t.record_stream(cur)
        my = shard_views(len(views), self.rank, self.world)
        split = gaussians.get("shs_dc") is not None
        sh_scene = self.backend is _cuda_view_backend and (split or gaussians.get("shs") is not None)
        defer = sh_scene and len(my) > 0
        self.buf.zero_(skip_sh=defer)  # with the deferred SH backward every SH gradient row is written, not added to
        self.loss_acc.zero_()
        acc = self.buf.accumulate_dict(split)
        self.last_launches = 0
        self.last_num_rendered = []
        outs = []
        # Deferred SH backward (CUDA backend, SH colours): every view leaves only its masked dL/drgb [P,3]; the SH
        # coefficients are read and their gradient written once per batch by pdgs_sh_backward_views.
        if defer:
            need = (len(my), self.buf.P, 3)
            if self._color_grads is None or tuple(self._color_grads.shape) != need:
                self._color_grads = torch.empty(need, dtype=torch.float32, device=self.device)
        # Per-view images (target / prior / mask) may live in pinned HOST memory: they are uploaded on a side
        # stream one view ahead of the rasterizer, so the host->device copies overlap the kernels.
        nxt = self._prefetch_begin(prefetch_gaussians) if prefetch_gaussians is not None else None
        jobs = list(enumerate(my))
        lanes = self.lanes if (self.backend is _cuda_view_backend and len(jobs) >= 2 * self.lanes) else 1

        def run_lane(lane: int, sub):
            """The views `sub` = [(k, i), ...] of one lane, in order, on the calling thread's current stream."""
            res = []
            scratch = self._scratch.setdefault(lane, {})
            pending = None
            if sub:
                i0 = sub[0][1]
                pre = self._pre_images.pop(i0, None)   # uploaded during the previous step?
                # the staged upload is taken only if it was made from exactly the host tensors this step passes for
                # that view: target, prior, depth mask and alpha mask, each by identity (None matches only None)
                now = (targets[i0],) + tuple(d.get(i0) if d is not None else None for d in (priors, masks, alpha_masks))
                if pre is not None and all(a is b for a, b in zip(pre[0], now)):
                    pending = pre[1]
                else:
                    pending = self._upload(sub[0][1], targets, priors, masks, alpha_masks)
            for n, (k, i) in enumerate(sub):
                cur = pending
                pending = self._upload(sub[n + 1][1], targets, priors, masks, alpha_masks) if n + 1 < len(sub) else None
                if nxt is not None:
                    self._prefetch_slice(nxt, k, len(jobs))
                tgt, pri, msk, alp = self._await(cur)
                acc_k = dict(acc)  # per view: the slot that receives this view's masked colour gradient
                if defer:
                    acc_k["color_grad_view"] = self._color_grads[k]
                if self.backend is _cuda_view_backend:
                    r = self.backend(gaussians, views[i], tgt, pri, msk, w_depth, acc_k, self.loss_acc, lambda_dssim, scratch,
                                     clamp, alp)
                else:
                    r = self.backend(gaussians, views[i], tgt, pri, msk, w_depth, acc_k, self.loss_acc)
                res.append((k, r))
            if prefetch_images is not None and sub:
                i0, nt = sub[0][1], prefetch_images["targets"]
                if i0 in nt:
                    others = tuple(prefetch_images.get(k) for k in ("priors", "masks", "alpha_masks"))
                    src = (nt[i0],) + tuple(d.get(i0) if d is not None else None for d in others)
                    self._pre_images[i0] = (src, self._upload(i0, nt, *others))
            return res

        if lanes == 1:
            results = run_lane(0, jobs)
        else:
            # self.lanes (default 3) pipelines in flight: lane t renders views t, t + lanes, ... on its own stream from its own
            # host thread.  The blend kernels are issue / latency-bound and the preprocess and sort kernels HBM-bound, so
            # kernels of different views co-scheduled on the SMs use both (C4: 2.28 ms per view with one lane, 2.07 with two,
            # 2.01 with three).  All lanes add
            # into the same gradient buffers: K9 accumulates with red.global.add, the loss with atomics.
            main = torch.cuda.current_stream(self.device)
            start = torch.cuda.Event()
            start.record(main)

            def lane_main(t):
                st = self._lane_streams[t]
                with torch.cuda.device(self.device), torch.cuda.stream(st):
                    st.wait_event(start)
                    out = run_lane(t, jobs[t::lanes])
                    ev = torch.cuda.Event()
                    ev.record(st)
                return out, ev

            if self._pool is None:
                from concurrent.futures import ThreadPoolExecutor
                self._pool = ThreadPoolExecutor(max_workers=self.lanes, thread_name_prefix="pdgs-lane")
                self._lane_streams = [torch.cuda.Stream(self.device) for _ in range(self.lanes)]
            futs = [self._pool.submit(lane_main, t) for t in range(lanes)]
            results = []
            for f in futs:
                out, ev = f.result()
                main.wait_event(ev)
                results.extend(out)
            results.sort(key=lambda kr: kr[0])
        for k, r in results:
            self.last_launches += int(r.get("launches", 0))
            self.last_num_rendered.append(int(r.get("num_rendered", 0)))
            if keep_outputs:
                outs.append(r)
        # the sequence of collectives must not depend on how many views THIS rank has (a rank without views still takes
        # part in every all-reduce, with zero rows)
        overlap = sh_scene and reduce and self.world > 1
        works = []
        if overlap:
            # Everything behind the SH gradient (opacities, scales, rotations, densification statistics: 10 of the 61
            # floats per Gaussian) is final once the last view's preprocess backward has run: its all-reduce starts
            # now, on NCCL's stream, and runs while the deferred SH backward computes.
            works.append(dist.all_reduce(self.buf.tail_slice(), op=dist.ReduceOp.SUM, group=self.group, async_op=True))
            works.append(dist.all_reduce(self.buf.max_radii2D, op=dist.ReduceOp.MAX, group=self.group, async_op=True))
            works.append(dist.all_reduce(self.loss_acc, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        if defer or overlap:
            from . import rasterizer as rz
            campos = torch.stack([views[i].campos.reshape(3) for i in my]).to(self.device, torch.float32) if defer else None
            # The SH pass runs in chunks along P: rows [lo, hi) of means3D and of the SH gradient are final
            # when their chunk is done, so their all-reduce overlaps the next chunk's kernel.
            P = self.buf.P
            n_chunks = self.reduce_chunks if overlap else 1
            step_rows = max(256, (-(-P // n_chunks) + 255) // 256 * 256)
            for lo in range(0, max(P, 1), step_rows):
                hi = min(P, lo + step_rows)
                if not defer:
                    pass  # no local view: rows stay zero, the collectives below still run
                elif split:
                    rz.sh_backward_views_split(gaussians["means3D"], gaussians["shs_dc"], gaussians.get("shs_rest"),
                                               views[my[0]].sh_degree, campos, self._color_grads, acc["shs_dc"],
                                               acc["shs_rest"], acc["means3D"], accumulate=False, first=lo, count=hi - lo)
                else:
                    rz.sh_backward_views(gaussians["means3D"], gaussians["shs"], views[my[0]].sh_degree, campos,
                                         self._color_grads, acc["shs"], acc["means3D"], accumulate=False, first=lo,
                                         count=hi - lo)
                if defer:
                    self.last_launches += rz.LAUNCHES.sh_bwd
                if overlap:
                    for t in self.buf.row_slices(lo, hi, split):
                        works.append(dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        if overlap:
            for w in works:
                w.wait()  # the current stream waits for the collective; the host does not block
        elif reduce and self.world > 1:
            self.buf.all_reduce(self.group)
            dist.all_reduce(self.loss_acc, op=dist.ReduceOp.SUM, group=self.group)
        prefetched = None
        if nxt is not None:
            if len(my) == 0:
                self._prefetch_slice(nxt, 0, 1)
            prefetched = self._prefetch_end(nxt)
        return dict(loss_sums=self.loss_acc, grads=self.buf, local_views=list(my), outputs=outs, prefetched=prefetched)

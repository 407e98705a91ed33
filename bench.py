#!/usr/bin/env python
"""Benchmark of the rasterizer hot path on BASELINE.json's headline configuration.

Workload (config 4): a batch of 64 cameras (1600x1064) around ~3M Gaussians (SH degree 3), fwd + image
loss + bwd per view, views sharded contiguously over the ranks (one process per GPU), per-Gaussian
gradients summed with one all-reduce per step.  One "step" = the whole 64-view batch.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference ...                     CPU arm: the float64 oracle on the host cores

Prints ONE JSON line (rank 0).  `value` = views/s with all inputs resident in HBM; `e2e` = the same
batch step fed from pinned HOST buffers (Gaussians, cameras, target images copied host->device and
the loss read back inside the timed region).  `roofline` describes the kernel that takes the largest
share of the step, `stages` every stage; durations are CUDA-event pairs recorded on the launching
stream inside the timed region (pdgs_set_stage_timing).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "priordepth-gaussiansplatting_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

METRIC = "fwd+bwd raster views/s @1.6K px, ~3M Gaussians (batch of 64 cameras, view-parallel)"
UNIT = "views/s"
FP32_LANES = 128  # fp32 lanes per SM; the SM count is read from the device


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C4")
    ap.add_argument("--views", type=int, default=64)
    ap.add_argument("--P", type=int, default=None, help="override the Gaussian count (debugging only)")
    ap.add_argument("--scale-image", type=float, default=1.0, help="debugging only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-window", type=int, nargs=2, default=[8, 6], help="tile window of the CPU sample")
    ap.add_argument("--no-dropin", action="store_true", help="skip the drop-in API row")
    ap.add_argument("--no-library-baselines", action="store_true", help="skip the CUB / reference fused-ssim rows")
    ap.add_argument("--dropin-only", action="store_true", help=argparse.SUPPRESS)  # child process of the dropin row
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


# --------------------------------------------------------------------------------------------
# CPU arm: the float64 oracle on bounded samples (tile windows of one view), extrapolated by pairs
# --------------------------------------------------------------------------------------------
CPU_POSITIONS = ((0.5, 0.5), (0.25, 0.3), (0.75, 0.7), (0.3, 0.75), (0.7, 0.25))


def oracle_sample(scene, view, window_tiles, threads, return_oracle=False, position=(0.5, 0.5), st=None):
    """fwd+bwd of the oracle (oracle/window.py) on the Gaussians touching a window of tiles of `view`.
    Returns (seconds, pairs in the window, pairs in the whole view, Gaussians in the window)."""
    from oracle import window as ow
    torch.set_num_threads(threads)
    W, H = view.image_width, view.image_height
    win = ow.place_window(W, H, window_tiles, position)
    if st is None:
        st = ow.structure(scene, view)
    gen = torch.Generator().manual_seed(7)
    target = torch.rand(3, H, W, generator=gen, dtype=torch.float64)
    t0 = time.perf_counter()
    o, ids, info = ow.window_oracle(scene, view, win, st=st, need_grad=True)
    res = o.forward()
    img = res.color.clamp(0, 1)
    gc = torch.sign(img - target) / target.numel() * ((res.color >= 0) & (res.color <= 1))
    o.backward(gc, None)
    dt = time.perf_counter() - t0
    if return_oracle:
        return dt, info["pairs_window"], info["pairs_view"], int(ids.numel()), dict(result=res, window=win, ids=ids)
    return dt, info["pairs_window"], info["pairs_view"], int(ids.numel())


def cpu_arm(args, scene, positions, samples, warm):
    """`positions` window positions of view 0, per position `warm` untimed + `samples` timed fwd+bwd passes of the oracle.
    Every sample is extrapolated to a whole view by its own pair count; the value is the mean over all timed samples and
    the spread over positions is reported (the figure is an EXTRAPOLATED sample, not a measured whole-view time)."""
    from oracle import window as ow
    threads = os.cpu_count() or 1
    view = scene.views[0]
    st = ow.structure(scene, view)
    per_pos, all_t, all_v, frac = [], [], [], []
    for pos in CPU_POSITIONS[:positions]:
        vals = []
        for it in range(warm + samples):
            dt, pw, pv, n = oracle_sample(scene, view, args.cpu_window, threads, position=pos, st=st)
            if it >= warm:
                vals.append(1.0 / (dt * pv / max(pw, 1)))
                all_t.append(dt)
        frac.append(pw / max(pv, 1))
        per_pos.append(dict(position=list(pos), gaussians=n, pairs_window=pw, views_per_s=sum(vals) / len(vals),
                            min=min(vals), max=max(vals)))
        all_v.extend(vals)
    value = sum(all_v) / len(all_v)
    t_sample = sum(all_t) / len(all_t)
    sample = (f"float64 PyTorch-CPU oracle (clamp + L1 loss), fwd+bwd, {args.cpu_window[0]}x{args.cpu_window[1]}-tile windows of "
              f"view 0 at {len(per_pos)} position(s), {warm} warm + {samples} timed passes each; a pass covers "
              f"{100 * sum(frac) / len(frac):.2f} % of the view's {pv} (tile,Gaussian) pairs and takes {t_sample:.2f} s; "
              f"EXTRAPOLATED to a whole view by pair count")
    return dict(value=value, unit=UNIT, cores=threads, kind="port", sample=sample, extrapolated=True,
                sample_fraction=sum(frac) / len(frac), spread=dict(min=min(all_v), max=max(all_v)),
                per_position=per_pos), t_sample


# --------------------------------------------------------------------------------------------
class _NvmlSampler:
    """SM clock + throttle reasons every 5 ms from a host thread (NVML): the timed region of a multi-GPU run lasts tens
    of milliseconds, far less than one `nvidia-smi -lms` period."""

    def __init__(self, gpu_index):
        import threading
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        h = None
        try:
            pr = torch.cuda.get_device_properties(gpu_index)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        self.h = h
        self.sm, self.reasons, self.stop = [], 0, False
        self.max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        nv = self.nv
        while not self.stop:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.reasons |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.005)

    def finish(self):
        self.stop = True
        self.t.join(timeout=2)
        nv = self.nv
        names = [("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                 ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                 ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap)]
        sm = sorted(self.sm)
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=self.max,
                    reasons=sorted(n for n, bit in names if self.reasons & bit), samples=len(sm), source="nvml, 5 ms period")


def start_clock_sampler(gpu_index):
    try:
        return _NvmlSampler(gpu_index), None
    except Exception:
        pass
    f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                 "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
    except OSError:
        return None, f.name
    return proc, f.name


def stop_clock_sampler(proc, path):
    if isinstance(proc, _NvmlSampler):
        return proc.finish()
    out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
    if proc is not None:
        proc.terminate()
        try:
            proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            proc.kill()
    try:
        rows = [r.strip().split(",") for r in open(path) if r.strip()]
        os.unlink(path)
    except OSError:
        rows = []
    sm, reasons = [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in rows:
        if len(r) < 7:
            continue
        try:
            sm.append(float(r[0]))
            out["sm_max_mhz"] = float(r[1])
        except ValueError:
            continue
        for k, name in enumerate(names):
            if r[3 + k].strip().lower() == "active":
                reasons.add(name)
    if sm:
        sm.sort()
        out["sm_mhz"] = sm[len(sm) // 2]
        out["samples"] = len(sm)
    out["reasons"] = sorted(reasons)
    return out


def _jsonable(o, path="line"):
    """Plain python for json.dumps; a tensor / numpy scalar that slipped into the line is converted (and reported on
    stderr with its key path, so that it gets fixed at the source)."""
    if isinstance(o, dict):
        return {str(k): _jsonable(v, f"{path}.{k}") for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return [_jsonable(v, f"{path}[{i}]") for i, v in enumerate(o)]
    if torch.is_tensor(o):
        print(f"bench.py: tensor at {path} converted for JSON", file=sys.stderr)
        return o.item() if o.numel() == 1 else o.tolist()
    if hasattr(o, "item") and not isinstance(o, (int, float, bool, str)):
        return o.item()
    return o


def _pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))]


def dropin_row(args, scene, dev, iters=50, warmup=10, vary_r=False):
    """SURVEY 8(d) timing protocol on the DROP-IN API: `GaussianRasterizer.forward` + `loss.backward()` of the caller's
    loss `|clamp(render, 0, 1) - gt|.mean()` (torch ops, as /root/reference/gaussian_renderer/__init__.py:121 and
    train.py:153 do), one view per iteration cycling through the cameras, CUDA events, 10 warm-up + 50 timed,
    median / p10 / p90.  Honours CUDA_LAUNCH_BLOCKING from the environment (the reference sets it, train.py:65).
    vary_r: every other camera renders with scale_modifier 2.5 (every splat 2.5x larger: about 3x the pairs) and the
    cameras are visited in a seeded random order, as a training loop pops them (train.py:130-135), so the speculative
    binning capacity is exercised with a pair count that jumps by more than 2x between calls; misses are counted."""
    import random
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from pdgs_b200 import rasterizer as rz
    from pdgs_b200 import scenes
    torch.cuda.set_device(dev)
    lv = {k: getattr(scene, k).to(dev).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    means2D = torch.zeros(scene.P, 3, device=dev, requires_grad=True)
    n = min(len(scene.views), 8)
    rast, tgts = [], []
    for i in range(n):
        v = scene.views[i]
        rs = GaussianRasterizationSettings(
            image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy, bg=scene.bg.to(dev),
            scale_modifier=2.5 if (vary_r and i % 2 == 1) else 1.0, viewmatrix=v.world_view_transform.to(dev), projmatrix=v.full_proj_transform.to(dev),
            sh_degree=scene.sh_degree, campos=v.camera_center.to(dev), prefiltered=False, debug=False, antialiasing=False)
        rast.append(GaussianRasterizer(raster_settings=rs))
        gen = torch.Generator().manual_seed(scenes.SEED0 + 1000 + i)
        tgts.append(torch.rand(3, v.image_height, v.image_width, generator=gen).to(dev))
    fwd, tot, pairs = [], [], []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    rng = random.Random(7)
    rz.binning_stats(reset=True)
    for it in range(warmup + iters):
        for t in list(lv.values()) + [means2D]:
            t.grad = None
        k = rng.randrange(n) if vary_r else it % n
        ev[0].record()
        color, radii, invd = rast[k](means3D=lv["means3D"], means2D=means2D, opacities=lv["opacities"], shs=lv["shs"],
                                     scales=lv["scales"], rotations=lv["rotations"])
        ev[1].record()
        loss = (color.clamp(0, 1) - tgts[k]).abs().mean()
        loss.backward()
        ev[2].record()
        torch.cuda.synchronize(dev)
        if it >= warmup:
            fwd.append(ev[0].elapsed_time(ev[1]))
            tot.append(ev[0].elapsed_time(ev[2]))
            pairs.append(int(color.grad_fn.num_rendered))
        elif it == warmup - 1:
            rz.binning_stats(reset=True)
    return dict(binning=rz.binning_stats(), pairs=dict(min=min(pairs), max=max(pairs)), api="diff_gaussian_rasterization.GaussianRasterizer.forward + loss.backward() (torch clamp + L1 loss), one view "
                    "per call", iters=iters, warmup=warmup, cuda_launch_blocking=os.environ.get("CUDA_LAUNCH_BLOCKING", "") == "1",
                fwd_ms=dict(median=_pct(fwd, 0.5), p10=_pct(fwd, 0.1), p90=_pct(fwd, 0.9)),
                fwd_bwd_ms=dict(median=_pct(tot, 0.5), p10=_pct(tot, 0.1), p90=_pct(tot, 0.9)),
                views_per_s=1e3 / _pct(tot, 0.5))


def dropin_row_blocking(args):
    """The same row in a child process with CUDA_LAUNCH_BLOCKING=1 (it must be set before CUDA initialises)."""
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING="1")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    cmd = [sys.executable, os.path.abspath(__file__), "--dropin-only", "--config", args.config, "--views", str(min(args.views, 8)),
           "--scale-image", str(args.scale_image)] + (["--P", str(args.P)] if args.P else [])
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        for ln in reversed(r.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)
        return dict(error=(r.stderr or r.stdout)[-400:])
    except Exception as e:  # noqa: BLE001
        return dict(error=f"{type(e).__name__}: {e}")


def library_baselines(scene, view_settings, g_dev, dev, H, W, P):
    """GPU denominators on the same box (tools/gpu_baselines.py): the reference rasterizer's K2 / K4 library calls
    (cub::DeviceScan::InclusiveSum, cub::DeviceRadixSort::SortPairs on the reference-format 64-bit keys) on the arrays
    of one real view, and the reference's fused-ssim CUDA extension next to pdgs_image_loss_ssim."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gpu_baselines as gb
    from pdgs_b200 import rasterizer as rz
    out = {}
    if gb.cub() is not None:
        rs = view_settings
        empty = torch.empty(0, device=dev)
        with torch.no_grad():
            R, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
                rs.bg, g_dev["means3D"], empty, g_dev["opacities"], g_dev["scales"], g_dev["rotations"], 1.0, empty,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, g_dev["shs"], rs.sh_degree, rs.campos, False,
                False, False)
        sv = rz.state_view(geom, binning, img, P, W, H, R, to_cpu=False)
        chain_t = dict(warmup=5, iters=30)
        ku, vu = gb.reference_binning_chain(sv, radii, W, H, timing=chain_t)
        T = ((W + 15) // 16) * ((H + 15) // 16)
        c = gb.cub_binning(sv["tiles_touched"], ku, vu, T, warmup=5, iters=30)
        out["cub"] = dict(
            what="reference-shaped K2 + K3 + K4 on this view's K1 outputs (xy, radii, depth bits, tiles_touched; nothing of the "
                 "new pipeline's binning is read): cub::DeviceScan::InclusiveSum(tiles_touched[P]), duplicateWithKeys with one "
                 "thread per Gaussian in index order (tools/cub_baseline.cu, SURVEY A.6) and "
                 "cub::DeviceRadixSort::SortPairs(u64 tile<<32|depth, u32 id) over bits [0, 32+bits(T)); CUDA 12.9 CUB",
            P=P, R=int(R), end_bit=c["end_bit"], scan_ms=c["scan_ms"], duplicate_ref_shaped_ms=chain_t["duplicate_ms"],
            sort_pairs_u64_ms=c["sort_ms"],
            reference_k2_k3_k4_ms=c["scan_ms"]["median"] + chain_t["duplicate_ms"]["median"] + c["sort_ms"]["median"],
            sorted_list_equals_ours=bool(torch.equal(c["sorted_vals"], sv["point_list"]) and
                                         torch.equal(c["sorted_keys"], sv["sorted_keys"])))
        del sv, ku, vu, c
    else:
        out["cub"] = dict(unavailable="tools/libcub_baseline.so not built")
    try:
        out["fused_ssim"] = gb.ssim_loss_baseline(H, W, dev)
        if out["fused_ssim"].get("reference_ms") is None:
            out["fused_ssim"]["unavailable"] = "oracle/_ref/fused_ssim_cuda.so not built"
    except Exception as e:  # noqa: BLE001
        out["fused_ssim"] = dict(error=f"{type(e).__name__}: {e}"[:300])
    torch.cuda.empty_cache()
    return out


def grad_checksum(buf):
    """Per-segment (sum g, sum |g|) of the all-reduced flat gradient buffer in float64, and the sum of max_radii2D: equal
    (to fp32 reassociation) for every N if N GPUs compute what one GPU computes (SURVEY 8(e) semantics note)."""
    out = {}
    for name in ("means3D", "shs", "opacities", "scales", "rotations", "densify_grad_accum", "densify_denom"):
        t = buf.view(name).double()
        out[name] = [float(t.sum()), float(t.abs().sum())]
    out["max_radii2D"] = int(buf.max_radii2D.sum(dtype=torch.int64))
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from pdgs_b200 import scenes

    if args.dropin_only:
        print(json.dumps(dropin_row(args, scenes.make_config(args.config, P=args.P, n_views=args.views,
                                                             scale_image=args.scale_image), torch.device("cuda", 0))))
        return 0

    if args.impl == "reference":
        if rank != 0:
            return 0
        scene = scenes.make_config(args.config, P=args.P, n_views=args.views, scale_image=args.scale_image)
        t0 = time.perf_counter()
        # K timed steps = K bounded samples of the workload (fwd+bwd of a tile window of a view), spread over three window
        # positions, at most 12 in total so the arm ends within a few minutes; one warm pass per position
        per_pos = max(1, min(4, -(-max(1, args.steps) // 3)))
        base, t_sample = cpu_arm(args, scene, 3, per_pos, 1 if args.warmup > 0 else 0)
        # `steps` is what was timed (window samples), not what was asked for; `ms_per_step` describes the same quantity as
        # `value` (one step = all views of the batch, extrapolated from the samples by pair count); the duration of one
        # measured sample is kept apart as `ms_per_sample`
        n_views = len(scene.views)
        line = dict(impl="reference", metric=METRIC, value=base["value"], unit=UNIT, n_gpus=args.gpus, steps=3 * per_pos,
                    steps_requested=args.steps, warmup=args.warmup,
                    ms_per_step=(1e3 * n_views / base["value"]) if base["value"] > 0 else None,
                    ms_per_sample=1e3 * t_sample, steps_timed=3 * per_pos, higher_is_better=True,
                    scaling="strong", vs_baseline=None, dtype="f64", data="synthetic", extrapolated=True,
                    config=workload_config(args, scene, world=max(1, args.gpus)), cpu_baseline=base,
                    e2e=dict(value=base["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    gpu_launches=0, wall_s=time.perf_counter() - t0,
                    note="EXTRAPOLATED CPU-ORACLE SAMPLE, not a speed-up denominator: the reference rasterizer is CUDA-only and "
                         "its source is absent from the checkout (empty submodule), so this arm times the float64 CPU "
                         "transcription of its maths (oracle/) on tile windows of one view (`ms_per_sample` = one measured "
                         "sample, `steps` = samples timed) and scales by pair count to views/s (`ms_per_step` = the "
                         "extrapolated time of one batch of views).  GPU-side denominators (CUB, the reference's "
                         "fused-ssim kernel) are in the b200 arm's `library_baselines`.")
        print(json.dumps(line))
        return 0

    assert torch.cuda.is_available(), "bench.py needs a B200: the rasterizer has no CPU path"
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    from pdgs_b200 import rasterizer as rz
    from pdgs_b200.viewparallel import ViewParallelRasterizer, shard_views

    scene = scenes.make_config(args.config, P=args.P, n_views=args.views, scale_image=args.scale_image)
    P, M = scene.P, scene.shs.shape[1]
    host = dict(means3D=scene.means3D, shs=scene.shs, opacities=scene.opacities, scales=scene.scales,
                rotations=scene.rotations)
    host = {k: v.contiguous().pin_memory() for k, v in host.items()}
    g_dev = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    my = shard_views(len(scene.views), rank, world)
    H, W = scene.views[0].image_height, scene.views[0].image_width

    def settings(view, mats):
        return GaussianRasterizationSettings(
            image_height=view.image_height, image_width=view.image_width, tanfovx=view.tanfovx, tanfovy=view.tanfovy,
            bg=mats["bg"], scale_modifier=1.0, viewmatrix=mats["view"], projmatrix=mats["proj"],
            sh_degree=scene.sh_degree, campos=mats["campos"], prefiltered=False, debug=False, antialiasing=False)

    # host-side per-view inputs (pinned): camera block [view 16 | proj 16 | campos 3 | bg 3] and the target image
    cam_host, tgt_host = {}, {}
    for i in my:
        v = scene.views[i]
        cam_host[i] = torch.cat([v.world_view_transform.reshape(-1), v.full_proj_transform.reshape(-1),
                                 v.camera_center.reshape(-1), scene.bg.reshape(-1)]).float().pin_memory()
        gen = torch.Generator().manual_seed(scenes.SEED0 + 1000 + i)
        tgt_host[i] = torch.rand(3, H, W, generator=gen).pin_memory()

    def cam_views(block):
        return dict(view=block[0:16].view(4, 4), proj=block[16:32].view(4, 4), campos=block[32:35], bg=block[35:38])

    cam_dev = {i: cam_host[i].to(dev) for i in my}
    tgt_dev = {i: tgt_host[i].to(dev) for i in my}
    views_dev = [settings(scene.views[i], cam_views(cam_dev[i])) if i in cam_dev else None for i in range(len(scene.views))]
    vp = ViewParallelRasterizer(P, M, dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---------------- resident-input throughput (`value`) ----------------
    for _ in range(args.warmup):
        vp.step(g_dev, views_dev, tgt_dev)
    barrier()
    sampler, sampler_path = start_clock_sampler(local_rank) if rank == 0 else (None, None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches, pairs = 0, []
    barrier()
    e0.record()
    for _ in range(args.steps):
        vp.step(g_dev, views_dev, tgt_dev)
        launches += vp.last_launches
        pairs.extend(vp.last_num_rendered)
    e1.record()
    barrier()
    clocks = stop_clock_sampler(sampler, sampler_path) if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    binning_stats = rz.binning_stats()        # warm-up + timed region of the batch path (this rank)
    checksum = grad_checksum(vp.buf)          # of the last timed step's all-reduced gradients (identical on all ranks)
    loss_sums_resident = [float(v) for v in vp.loss_acc.cpu()]
    # Per-stage durations (roofline / stages): a second measured region of the same steps with ONE view in flight.  In
    # the region above `lanes` views run concurrently on separate streams, so a stage's event pair would also cover
    # kernels of other views; one view at a time gives each kernel's own duration (the figure ncu's launch list shows).
    lanes = vp.lanes
    vp.lanes = 1
    vp.step(g_dev, views_dev, tgt_dev)
    barrier()
    rz.set_stage_timing(dev, True)
    rz.stage_times(dev)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(args.steps):
        vp.step(g_dev, views_dev, tgt_dev)
    s1.record()
    barrier()
    stage = rz.stage_times(dev)
    rz.set_stage_timing(dev, False)
    ms_one_lane = s0.elapsed_time(s1) / args.steps
    vp.lanes = lanes
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())
    ms_total = float(ms.item())
    n_views = len(scene.views)
    value = n_views * args.steps / (ms_total * 1e-3)
    loss_mean = loss_sums_resident[0] / n_views

    # ---------------- algorithmic work of this rank's views (untimed pass) ----------------
    work = dict(P=P, V=0, R=0, n_test=0, n_acc=0, n_test_bwd=0, views=len(my), tiles=((W + 15) // 16) * ((H + 15) // 16),
                HW=H * W)
    empty = torch.empty(0, device=dev)
    rz.set_work_counters(dev, True)   # N_test / N_acc counters of the forward blend: only this untimed pass needs them
    with torch.no_grad():
        for i in my:
            rs = views_dev[i]
            R, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
                rs.bg, g_dev["means3D"], empty, g_dev["opacities"], g_dev["scales"], g_dev["rotations"], 1.0, empty,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, g_dev["shs"], rs.sh_degree, rs.campos, False,
                False, False)
            c = rz.image_counters(geom, binning, img, P, W, H)
            work["V"] += int((radii > 0).sum().item())
            work["R"] += int(R)
            work["n_test"] += c["n_tested"]
            work["n_acc"] += c["n_accepted"]
            work["n_test_bwd"] += c["n_contrib"]
    rz.set_work_counters(dev, False)
    torch.cuda.synchronize(dev)

    # ---------------- end-to-end with host buffers (`e2e`) ----------------
    e2e = None
    if not args.no_e2e:
        d2h = 12  # the three loss sums, read back every step

        staged = [None]
        # The step's result (three loss sums) is read back EVERY step: an asynchronous device->host copy into pinned
        # memory behind the step, consumed by the host one step later (after its event), so that the host keeps
        # enqueueing the next step while the copy lands -- the loss.item() of a training loop that logs one step late.
        loss_pinned = [torch.empty(3, dtype=torch.float32).pin_memory() for _ in range(2)]
        loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
        loss_seen = []

        def e2e_step(k, prefetch_next=True):
            # this step's Gaussians: uploaded by the previous step's prefetch, or now (first step of a run)
            g = staged[0] if staged[0] is not None else vp.stage_gaussians(host)
            vs = [None] * n_views
            tg = {}
            for i in my:
                vs[i] = settings(scene.views[i], cam_views(cam_host[i].to(dev, non_blocking=True)))
                tg[i] = tgt_host[i]  # pinned host tensor: step() uploads it one view ahead on a copy stream
            # the NEXT step's Gaussians are uploaded by step() itself, slice by slice behind the image uploads, so the copy
            # overlaps this step's kernels; every step's host->device copy still happens inside the timed region
            # (`steps` parameter uploads for `steps` steps)
            out = vp.step(g, vs, tg, prefetch_gaussians=host if prefetch_next else None,
                          prefetch_images=dict(targets=tg) if prefetch_next else None)
            staged[0] = out["prefetched"]
            if os.environ.get("PDGS_E2E_SYNC_READBACK", "0") == "1":   # A/B switch: blocking read of this step's result
                loss_seen.append(out["loss_sums"].cpu())
                loss_ready[k & 1].record()
                loss_pinned[k & 1].copy_(loss_seen[-1])
                return
            loss_pinned[k & 1].copy_(out["loss_sums"], non_blocking=True)
            loss_ready[k & 1].record()
            if k > 0:
                loss_ready[(k - 1) & 1].synchronize()
                loss_seen.append(loss_pinned[(k - 1) & 1].clone())

        def e2e_drain(k_last):
            loss_ready[k_last & 1].synchronize()
            if os.environ.get("PDGS_E2E_SYNC_READBACK", "0") != "1":
                loss_seen.append(loss_pinned[k_last & 1].clone())
            return loss_seen[-1]

        nw = max(1, min(args.warmup, 2))
        for k in range(nw):
            e2e_step(k, prefetch_next=False)
        e2e_drain(nw - 1)
        barrier()
        # N > 1: every rank uploaded 1/world of each tensor and an in-place all-gather over NVLink completed the
        # replicas; they must equal the host tensors bit for bit on every rank
        replicas_ok = None
        if world > 1:
            staged_dev, ev = vp.stage_gaussians(host)
            if ev is not None:
                torch.cuda.current_stream(dev).wait_event(ev)
            ok = all(torch.equal(staged_dev[k].cpu(), host[k]) for k in host)
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            replicas_ok = bool(flag.item())
            assert replicas_ok, "gathered parameter replicas differ from the host tensors"
            del staged_dev
            barrier()
        h2d0 = vp.h2d_bytes
        t0 = time.perf_counter()
        e0.record()
        loss_seen.clear()
        for k in range(args.steps):
            e2e_step(k, prefetch_next=(k + 1 < args.steps))
        last = e2e_drain(args.steps - 1)   # every step's result has reached the host inside the timed region
        e1.record()
        barrier()
        assert len(loss_seen) == args.steps
        wall = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([max(wall, e0.elapsed_time(e1))], dtype=torch.float64, device=dev)
        # bytes actually copied host -> device per step by this rank: Gaussians (first step: the whole replica; later steps:
        # this rank's 1/world share, the rest arrives by all-gather over NVLink), target images, camera blocks
        h2d = (vp.h2d_bytes - h2d0) // args.steps + sum(cam_host[i].numel() * 4 for i in my)
        hb = torch.tensor([h2d, d2h], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(hb)
        e2e = dict(value=n_views * args.steps / (float(t.item()) * 1e-3), unit=UNIT, h2d_bytes_per_step=int(hb[0].item()),
                   d2h_bytes_per_step=int(hb[1].item()), ms_per_step=float(t.item()) / args.steps,
                   api="pdgs_b200.viewparallel.ViewParallelRasterizer.step (C ABI: pdgs_forward / pdgs_image_loss / "
                       "pdgs_backward) fed from pinned host tensors; step k+1's Gaussians are uploaded during step k (each rank its "
                       "1/world share over PCIe, replicas completed by an all-gather over NVLink), target images one view ahead; the "
                       "loss sums of every step are copied device->host into pinned memory behind the step and read by the host "
                       "one step later",
                   loss_mean=float(last[0].item()) / n_views, replicas_equal_host=replicas_ok,
                   grad_checksum=grad_checksum(vp.buf))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---------------- roofline per stage (rank 0's launches) ----------------
    peaks = measured_peaks()
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count   # read from the device; B200: 148
    fp32_peak = sm_count * FP32_LANES * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12  # TFLOP/s at the max SM clock
    nv = max(work["views"], 1)
    Pn, V, R = work["P"], work["V"] / nv, work["R"] / nv
    T, HW = work["tiles"], work["HW"]
    npass = (max(1, T.bit_length()) + 7) // 8  # tile-id radix passes (8-bit digits); the depth sort always takes 4
    # Algorithmic bytes / flops per launch (= per view) of THIS pipeline.  Binning is
    # "depth sort of P Gaussians (u32 key, u32 index) -> scan/duplicate in depth order -> tile-id sort of R pairs
    # (u32, u32)", which yields the reference's sorted list from 8-byte instead of 12-byte pairs and 2 instead of 6
    # passes over R; SURVEY.md 8(d)'s figure for the reference's single 64-bit sort is 152 R.
    alg = {
        # SURVEY 8(d) / DESIGN.md section 4: read 12 P + 224 V, write 8 P (radii, tiles_touched) + 40 V = 20 P + 264 V
        "preprocess": ("hbm", 12 * Pn + 224 * V + 8 * Pn + 40 * V),
        "depth_sort": ("hbm", 4 * Pn + 12 * Pn + 3 * 16 * Pn),
        "scan": ("hbm", 12 * Pn),
        # fused scan + duplicate: sorted order 4P, tile rects 8P gathered, point_offsets 4P written, 8-byte pairs written
        "duplicate": ("hbm", 16 * Pn + 8 * R),
        "sort": ("hbm", 4 * R + npass * 16 * R),
        "ranges": ("hbm", 4 * R + 8 * T),
        "blend_fwd": ("fp32_simt", 15 * work["n_test"] / nv + 11 * work["n_acc"] / nv),
        "bwd_zero": ("hbm", 48 * Pn),
        "blend_bwd": ("fp32_simt", 15 * work["n_test_bwd"] / nv + 59 * work["n_acc"] / nv),
        # batch path: SH backward deferred -> per view K9 reads radii 4P, params 44V, moments 48V, records 48V, writes the
        # masked colour gradient 12P and read-modify-writes means/opacity/scales/rotations/densify (52V each way)
        "preprocess_bwd": ("hbm", 16 * Pn + 244 * V),
        # once per step: means 12, SH 192, V_local colour gradients 12 each, SH gradient 192 written, means gradient 24
        "sh_bwd": ("hbm", Pn * (12 + 192 + 12 * max(work["views"], 1) + 192 + 24)),
        "loss": ("hbm", 36 * HW),
    }
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    stages, total_stage_ms = [], sum(v[0] for v in stage.values())
    for name, (ms_sum, cnt) in stage.items():
        if cnt == 0:
            continue
        bound, amount = alg[name]
        avg_ms = ms_sum / cnt
        if bound == "hbm":
            achieved, peak, unit = amount / (avg_ms * 1e-3) / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            achieved, peak, unit = amount / (avg_ms * 1e-3) / 1e12, fp32_peak, "TFLOP/s"
        stages.append(dict(stage=name, bound=bound, avg_ms=avg_ms, launches=cnt, share=ms_sum / max(total_stage_ms, 1e-9),
                           achieved=achieved, peak=peak, unit=unit, frac=achieved / peak,
                           traffic=traffic.get(name)))
    stages.sort(key=lambda s: -s["share"])
    top = stages[0] if stages else None
    roofline = None
    if top:
        roofline = dict(kernel=top["stage"], bound=top["bound"], achieved=top["achieved"], peak=top["peak"], unit=top["unit"],
                        frac=top["frac"], traffic=top["traffic"],
                        traffic_source=dict(file=traffic.get("_file") or str(traffic.get("_source", "")).split(":")[0] or None,
                                            commit=traffic.get("_commit")),
                        avg_ms=top["avg_ms"], share_of_step=top["share"],
                        peak_source=("MEASURED_PEAKS.json hbm_gbs (%s)" % peaks["source"]) if top["bound"] == "hbm" else
                        "derived: %d SMs (device attribute) x 128 fp32 lanes x 2 x sm_max_mhz (MEASURED_PEAKS.json has no fp32 SIMT figure)" % sm_count)

    lib = None
    if not args.no_library_baselines:
        lib = library_baselines(scene, views_dev[my[0]], g_dev, dev, H, W, P)
        cubr = lib.get("cub", {})
        for st_ in stages:
            if st_["stage"] in ("sort", "depth_sort") and "sort_pairs_u64_ms" in cubr:
                st_["library_baseline_ms"] = cubr["sort_pairs_u64_ms"]["median"]
                st_["library_baseline"] = ("cub::DeviceRadixSort::SortPairs over the reference's 64-bit keys: the ONE sort "
                                           "the reference runs where this pipeline runs depth_sort + sort")
            if st_["stage"] == "duplicate" and "scan_ms" in cubr:
                st_["library_baseline_ms"] = cubr["scan_ms"]["median"]
                st_["library_baseline"] = "cub::DeviceScan::InclusiveSum (the scan fused into this kernel)"
        t = {s_["stage"]: s_["avg_ms"] for s_ in stages}
        if "sort_pairs_u64_ms" in cubr:
            cubr["ours_ms"] = dict(depth_sort=t.get("depth_sort"), scan_duplicate=t.get("duplicate"), tile_sort=t.get("sort"))
            cubr["ours_sorts_ms"] = (t.get("depth_sort") or 0.0) + (t.get("sort") or 0.0)
            cubr["reference_sort_plus_scan_ms"] = cubr["sort_pairs_u64_ms"]["median"] + cubr["scan_ms"]["median"]
            cubr["ours_k2_k3_k4_ms"] = cubr["ours_sorts_ms"] + (t.get("duplicate") or 0.0) + (t.get("scan") or 0.0)
        fs = lib.get("fused_ssim", {})
        if "loss" in t and fs.get("reference_ms"):
            fs["note"] = ("pdgs_image_loss_ssim = value + cotangent of (1-l)*L1 + l*(1-SSIM) on the raw render (clamp inside), "
                          "3 launches; reference = the same through torch ops + the reference's fused-ssim extension")

    dropin = None
    if world == 1 and not args.no_dropin:
        del vp
        torch.cuda.empty_cache()
        dropin = dict(default=dropin_row(args, scene, dev), varied_r=dropin_row(args, scene, dev, vary_r=True),
                      launch_blocking=dropin_row_blocking(args))

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        # the SAME protocol and process conditions as the reference arm: run it as a child process (fresh interpreter, no
        # CUDA context) on a bounded number of samples and take its cpu_baseline block
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "6", "--warmup", "1", "--config",
               args.config, "--views", str(args.views), "--scale-image", str(args.scale_image), "--cpu-window",
               str(args.cpu_window[0]), str(args.cpu_window[1])] + (["--P", str(args.P)] if args.P else [])
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
            cpu_baseline = json.loads([ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")][-1])["cpu_baseline"]
        except Exception as e:  # noqa: BLE001
            cpu_baseline = dict(error=f"{type(e).__name__}: {e}"[:300])

    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                ms_per_step=ms_total / args.steps, ms_per_view_per_gpu=ms_total / args.steps / max(len(my), 1),
                higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                config=workload_config(args, scene, world), clocks=clocks, e2e=e2e, gpu_launches=launches,
                roofline=roofline, stages=stages, cpu_baseline=cpu_baseline, library_baselines=lib, dropin=dropin,
                grad_checksum=checksum, loss_sums=loss_sums_resident, binning=binning_stats,
                stage_timing=dict(lanes=1, ms_per_step=ms_one_lane,
                                  note="stages / roofline come from a second measured region of the same steps with one view "
                                       "in flight; `value` is measured with config.lanes views in flight on separate streams"),
                work_per_view=dict(P=Pn, visible=V, pairs=R, n_test=work["n_test"] / nv, n_acc=work["n_acc"] / nv,
                                   n_test_bwd=work["n_test_bwd"] / nv),
                loss_mean=loss_mean)
    print(json.dumps(_jsonable(line)))
    if world > 1:
        dist.destroy_process_group()
    return 0


def workload_config(args, scene, world):
    v = scene.views[0]
    return dict(workload=f"{args.config}: {scene.P} Gaussians SH3, {len(scene.views)} cameras {v.image_width}x{v.image_height}, "
                         f"fwd + L1 loss on clamp(render,0,1) + bwd per view, grads summed over views",
                gaussians=scene.P, views_per_step=len(scene.views), image=[v.image_width, v.image_height],
                parallelism=f"view-parallel x{world} (Gaussians replicated, one all-reduce of [P,61] fp32 per step)",
                lanes=int(os.environ.get("PDGS_LANES", "3")),
                l2_policy="inputs larger than L2 (708 MB of Gaussian parameters streamed per view vs 126 MB L2); no flush")


if __name__ == "__main__":
    sys.exit(main())

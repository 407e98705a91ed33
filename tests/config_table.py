#!/usr/bin/env python
"""Per-config timing AND parity table through the reference-facing Python boundary (SURVEY.md 8(d) timing protocol).
Lives under tests/ because it uses the oracle as a checker (only tests/, smoke() and bench.py's CPU legs may touch oracle/).

For each BASELINE.json config (C1, C2, C3, C4 single view, C5 forward-only) this runs
`diff_gaussian_rasterization.GaussianRasterizer.forward` + `.backward()` of an L1 image loss (C3 adds the masked
inverse-depth L1 of /root/reference/train.py:174), CUDA events around forward and around backward separately,
W warm-up + K timed iterations, median / p10 / p90.  It also checks, at FULL size, radii and num_rendered against
the fp32 bit-model (oracle/, checker only) and times the float64 CPU oracle on a tile window (bench.py's CPU arm).

  python tests/config_table.py [--configs C1 C2 C3 C4 C5] [--iters 50] [--warmup 10] [--json out.json]

Prints one JSON object per config and a markdown table row for BASELINE.md section 4.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "priordepth-gaussiansplatting_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))]


def run_config(name, iters, warmup, cpu_window, check_ints=True, cpu=True):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from pdgs_b200 import scenes
    dev = torch.device("cuda", 0)
    t0 = time.perf_counter()
    sc = scenes.make_config(name, n_views=1 if name != "C4" else 64)
    view = sc.views[0]
    H, W = view.image_height, view.image_width
    gen_s = time.perf_counter() - t0
    fwd_only = name == "C5"
    leaves = dict(means3D=sc.means3D, shs=sc.shs, opacities=sc.opacities, scales=sc.scales, rotations=sc.rotations)
    leaves = {k: v.to(dev).requires_grad_(not fwd_only) for k, v in leaves.items()}
    means2D = torch.zeros(sc.P, 3, device=dev, requires_grad=not fwd_only)
    rs = GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=view.tanfovx, tanfovy=view.tanfovy, bg=sc.bg.to(dev), scale_modifier=1.0,
        viewmatrix=view.world_view_transform.to(dev), projmatrix=view.full_proj_transform.to(dev), sh_degree=sc.sh_degree,
        campos=view.camera_center.to(dev), prefiltered=False, debug=False, antialiasing=False)
    rast = GaussianRasterizer(raster_settings=rs)
    gen = torch.Generator().manual_seed(scenes.SEED0 + 77)
    target = torch.rand(3, H, W, generator=gen).to(dev)
    prior = torch.rand(1, H, W, generator=gen).to(dev) if name == "C3" else None
    mask = (torch.rand(1, H, W, generator=gen) > 0.2).float().to(dev) if name == "C3" else None
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    fwd, bwd, R = [], [], 0
    torch.cuda.reset_peak_memory_stats(dev)
    for it in range(warmup + iters):
        for t in list(leaves.values()) + [means2D]:
            t.grad = None
        ev[0].record()
        if fwd_only:
            with torch.no_grad():
                color, radii, invd = rast(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                          shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
            ev[1].record()
            ev[2].record()
        else:
            color, radii, invd = rast(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                      shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
            ev[1].record()
            loss = (color - target).abs().mean()
            if prior is not None:
                loss = loss + ((invd - prior) * mask).abs().mean()  # /root/reference/train.py:174
            loss.backward()
            ev[2].record()
        torch.cuda.synchronize(dev)
        if it >= warmup:
            fwd.append(ev[0].elapsed_time(ev[1]))
            bwd.append(ev[1].elapsed_time(ev[2]))
    out = dict(config=name, P=sc.P, W=W, H=H, visible=int((radii > 0).sum().item()), scene_gen_s=round(gen_s, 1),
               fwd_ms=dict(median=pct(fwd, 0.5), p10=pct(fwd, 0.1), p90=pct(fwd, 0.9)),
               bwd_ms=None if fwd_only else dict(median=pct(bwd, 0.5), p10=pct(bwd, 0.1), p90=pct(bwd, 0.9)),
               note="bwd includes the torch L1 loss kernels and autograd glue" if not fwd_only else "forward only",
               iters=iters, warmup=warmup, peak_mem_gb=torch.cuda.max_memory_allocated(dev) / 2**30)
    if hasattr(color, "grad_fn") and color.grad_fn is not None:
        out["num_rendered"] = int(color.grad_fn.num_rendered)
    if check_ints:
        import helpers
        st = helpers.bit_structure(sc, view)
        out["radii_bit_exact"] = bool((radii.cpu().numpy() == st["radii"]).all())
        out["num_rendered_model"] = int(st["num_rendered"])
        if "num_rendered" in out:
            out["num_rendered_equal"] = out["num_rendered"] == out["num_rendered_model"]
        else:
            out["num_rendered"] = out["num_rendered_model"]
    if cpu:
        import bench

        class A:
            pass
        a = A()
        a.cpu_window = cpu_window
        dt, pw, pv, n, orc = bench.oracle_sample(sc, view, cpu_window, os.cpu_count() or 1, return_oracle=True)
        # image / inverse-depth parity inside the window (the oracle blends only the window's tiles)
        x0, y0, x1, y1 = orc["window"]
        ys, xs = slice(16 * y0, min(16 * y1, H)), slice(16 * x0, min(16 * x1, W))
        res = orc["result"]
        ok = ~res.fragile[ys, xs]
        dc = (color.detach().double().cpu()[:, ys, xs] - res.color[:, ys, xs]).abs().max(dim=0).values
        dd = (invd.detach().double().cpu()[0, ys, xs] - res.invdepth[0, ys, xs]).abs()
        out["parity_window"] = dict(max_abs_color=float(dc[ok].max()), max_abs_invdepth=float(dd[ok].max()),
                                    invdepth_max=float(res.invdepth.max()), fragile_pixels=int((~ok).sum()),
                                    pixels=int(ok.numel()))
        out["cpu_oracle"] = dict(window_tiles=cpu_window, sample_s=dt, pairs_window=pw, pairs_view=pv,
                                 est_view_s=dt * pv / max(pw, 1), cores=os.cpu_count(),
                                 note="float64 CPU oracle fwd+bwd on a centred tile window, extrapolated by pairs")
    del leaves, means2D, target
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", nargs="+", default=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--cpu-window", type=int, nargs=2, default=[12, 8])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ints", action="store_true")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    assert torch.cuda.is_available(), "needs a B200"
    rows = []
    for name in args.configs:
        it = args.iters if name != "C5" else max(5, args.iters // 5)
        r = run_config(name, it, args.warmup if name != "C5" else 3, args.cpu_window, not args.no_ints, not args.no_cpu)
        rows.append(r)
        print(json.dumps(r), flush=True)
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    print("\n| Config | P / visible / pairs | fwd ms (p10-p90) | bwd ms (p10-p90) | fwd+bwd ms | CPU oracle est. s/view (cores) | max-abs img / invdepth (window) | ints bit-exact |")
    print("|---|---|---|---|---|---|---|---|")
    for r in rows:
        f, b = r["fwd_ms"], r["bwd_ms"]
        cpu = r.get("cpu_oracle")
        print(f"| {r['config']} {r['W']}x{r['H']} | {r['P']:,} / {r['visible']:,} / {r.get('num_rendered', 0):,} | "
              f"{f['median']:.3f} ({f['p10']:.3f}-{f['p90']:.3f}) | " +
              (f"{b['median']:.3f} ({b['p10']:.3f}-{b['p90']:.3f}) | {f['median'] + b['median']:.3f} | " if b else "- | - | ") +
              (f"{cpu['est_view_s']:.0f} ({cpu['cores']}) | " if cpu else "- | ") +
              (f"{r['parity_window']['max_abs_color']:.1e} / {r['parity_window']['max_abs_invdepth']:.1e} | "
               if "parity_window" in r else "- | ") +
              f"radii {r.get('radii_bit_exact')}, R {r.get('num_rendered_equal', '-')} |")


if __name__ == "__main__":
    main()

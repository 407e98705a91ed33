"""float64 PyTorch-CPU transcription of the reference rasterizer's math.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this file;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs do, as the checker / reported CPU baseline.

PARITY UNPINNED: the rasterizer source (graphdeco-inria/diff-gaussian-rasterization
@ 9c5c2028f6fbee2be239bc4c9421ff894fe4fbe0, branch dr_aa; /root/reference/.SUBMODULES.json
status "missing") is not in /root/reference, the reference ships no test or golden
vector for this path, and there is no reference binary.  The algorithm restated here
is the published one recalled in SURVEY.md Appendix A; the sub-steps the reference
DOES carry in-tree are pinned against it through tests/golden/:
  * SH -> RGB            /root/reference/utils/sh_utils.py:57-112, gaussian_renderer/__init__.py:76-80
  * cov3D = (RS)(RS)^T   /root/reference/utils/general_utils.py:64-110, scene/gaussian_model.py:34-38
  * view / projection    /root/reference/utils/graphics_utils.py:38-71, scene/cameras.py:80-89
Call site of the whole path: /root/reference/gaussian_renderer/__init__.py:36-52, 104-112.

Conventions:
  * backward = autograd of the float64 forward with a straight-through min(0.99, alpha)
  * radius / rect / alpha>=1/255 / T>=1e-4 tests are non-differentiable masks
  * grad_means2D is the gradient w.r.t. an NDC-scaled pixel offset (x 0.5*W, 0.5*H)
  * the 1.3*tanfov clamp detaches the clamped lateral coordinate (reference K8 behaviour)
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np
import torch

TILE = 16


def _f32(v: float) -> float:
    """The value the fp32 literal takes in the CUDA reference, as a python float."""
    return float(np.float32(v))


NEAR_Z = _f32(0.2)
DILATION = _f32(0.3)
AA_FLOOR = _f32(0.000025)
ALPHA_MIN = float(np.float32(1.0) / np.float32(255.0))
ALPHA_MAX = _f32(0.99)
T_EPS = _f32(0.0001)
W_EPS = _f32(0.0000001)
LIM = _f32(1.3)
RADIUS_FLOOR = _f32(0.1)

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435]


def eval_sh_basis(deg: int, d: torch.Tensor) -> torch.Tensor:
    """[N,3] unit dirs -> [N,(deg+1)^2] basis; polynomial of utils/sh_utils.py:74-100."""
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    out = [torch.full_like(x, SH_C0)]
    if deg > 0:
        out += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        out += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2.0 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if deg > 2:
        out += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
                SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
                SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    return torch.stack(out, dim=1)


def build_rotation_unnormalised(q: torch.Tensor) -> torch.Tensor:
    """utils/general_utils.py:78-99 without the renormalisation (A.3: quaternion used as given)."""
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


def cov3d_from_scale_rot(scales: torch.Tensor, mod: float, rots: torch.Tensor) -> torch.Tensor:
    """[N,6] packed [xx,xy,xz,yy,yz,zz] (utils/general_utils.py:64-75 ordering)."""
    R = build_rotation_unnormalised(rots)
    L = R * (mod * scales)[:, None, :]
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=1)


@dataclass
class Settings:
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool = False
    debug: bool = False
    antialiasing: bool = False


@dataclass
class ForwardResult:
    color: torch.Tensor            # [3,H,W] float64
    invdepth: torch.Tensor         # [1,H,W]
    radii: torch.Tensor            # [P] int32
    final_T: torch.Tensor          # [H,W]
    n_contrib: torch.Tensor        # [H,W] int32
    fragile: torch.Tensor          # [H,W] bool: a threshold test sat within rounding distance
    point_list: torch.Tensor       # [R] int64 sorted Gaussian ids
    point_tiles: torch.Tensor      # [R] int64 tile id per entry
    ranges: torch.Tensor           # [T,2] int64
    tiles_touched: torch.Tensor    # [P] int64
    rect: torch.Tensor             # [P,4] int64 (x0,y0,x1,y1)
    depth: torch.Tensor            # [P] float64 view-space z
    xy: torch.Tensor               # [P,2] pixel centre
    conic: torch.Tensor            # [P,3]
    opacity_eff: torch.Tensor      # [P]
    rgb: torch.Tensor              # [P,3]
    clamped: torch.Tensor          # [P,3] bool
    n_test: int = 0
    n_acc: int = 0
    n_test_bwd: int = 0
    aux: Dict = field(default_factory=dict)


def _leaf64(t: Optional[torch.Tensor], need_grad: bool) -> Optional[torch.Tensor]:
    if t is None or t.numel() == 0:
        return None
    out = t.detach().to(torch.float64).cpu().clone()
    out.requires_grad_(need_grad)
    return out


class Oracle:
    """One forward (+ optional backward) of the rasterizer in float64 on the CPU."""

    def __init__(self, settings: Settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                 rotations=None, cov3D_precomp=None, need_grad: bool = True, structure: Optional[dict] = None,
                 tile_window: Optional[tuple] = None, round_pixel_centre: bool = True):
        s = settings
        self.s = s
        self.W, self.H = int(s.image_width), int(s.image_height)
        self.gx, self.gy = (self.W + TILE - 1) // TILE, (self.H + TILE - 1) // TILE
        self.need_grad = need_grad
        self.structure = structure
        self.tile_window = tile_window  # (tx0, ty0, tx1, ty1) restrict blending to a tile window
        # False only for finite-difference self-checks (the fp32 rounding of the pixel centre is a step function)
        self.round_pixel_centre = round_pixel_centre
        self.means3D = _leaf64(means3D, need_grad)
        self.opacities = _leaf64(opacities, need_grad)
        self.shs = _leaf64(shs, need_grad)
        self.colors_precomp = _leaf64(colors_precomp, need_grad)
        self.scales = _leaf64(scales, need_grad)
        self.rotations = _leaf64(rotations, need_grad)
        self.cov3D_precomp = _leaf64(cov3D_precomp, need_grad)
        P = self.means3D.shape[0]
        self.P = P
        self.means2D = torch.zeros(P, 3, dtype=torch.float64, requires_grad=need_grad)
        if (self.shs is None) == (self.colors_precomp is None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((self.scales is None or self.rotations is None) and self.cov3D_precomp is None) or (
                (self.scales is not None or self.rotations is not None) and self.cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        self.V = s.viewmatrix.detach().to(torch.float64).cpu()
        self.Pm = s.projmatrix.detach().to(torch.float64).cpu()
        self.campos = s.campos.detach().to(torch.float64).cpu()
        self.bg = s.bg.detach().to(torch.float64).cpu()
        self.result: Optional[ForwardResult] = None

    # ------------------------------------------------------------------ K1
    def _preprocess(self):
        s, W, H = self.s, self.W, self.H
        P = self.P
        m = self.means3D
        ones = torch.ones(P, 1, dtype=torch.float64)
        hom = torch.cat([m, ones], dim=1)
        p_view = (hom @ self.V)[:, :3]
        p_hom = hom @ self.Pm
        p_w = 1.0 / (p_hom[:, 3] + W_EPS)
        ndc = p_hom[:, :2] * p_w[:, None]
        tanfovx, tanfovy = _f32(s.tanfovx), _f32(s.tanfovy)
        focal_x = float(np.float32(W) / (np.float32(2.0) * np.float32(s.tanfovx)))
        focal_y = float(np.float32(H) / (np.float32(2.0) * np.float32(s.tanfovy)))
        if self.cov3D_precomp is not None:
            cov3 = self.cov3D_precomp
        else:
            cov3 = cov3d_from_scale_rot(self.scales, _f32(s.scale_modifier), self.rotations)
        in_front = p_view[:, 2].detach() > NEAR_Z
        tz = torch.where(in_front, p_view[:, 2], torch.ones_like(p_view[:, 2]))  # keep culled rows finite
        limx, limy = LIM * tanfovx, LIM * tanfovy
        txtz, tytz = p_view[:, 0] / tz, p_view[:, 1] / tz
        cx = (txtz.detach() < -limx) | (txtz.detach() > limx)
        cy = (tytz.detach() < -limy) | (tytz.detach() > limy)
        tx = torch.where(cx, (txtz.clamp(-limx, limx) * tz).detach(), p_view[:, 0])
        ty = torch.where(cy, (tytz.clamp(-limy, limy) * tz).detach(), p_view[:, 1])
        J00, J02 = focal_x / tz, -(focal_x * tx) / (tz * tz)
        J11, J12 = focal_y / tz, -(focal_y * ty) / (tz * tz)
        Rw = self.V[:3, :3].T  # true world->view rotation, Rw[r][c] = vm[4c+r]
        M0 = J00[:, None] * Rw[0][None, :] + J02[:, None] * Rw[2][None, :]
        M1 = J11[:, None] * Rw[1][None, :] + J12[:, None] * Rw[2][None, :]
        S = torch.stack([cov3[:, 0], cov3[:, 1], cov3[:, 2], cov3[:, 1], cov3[:, 3], cov3[:, 4], cov3[:, 2],
                         cov3[:, 4], cov3[:, 5]], dim=1).view(P, 3, 3)
        v0 = torch.einsum('pij,pj->pi', S, M0)
        v1 = torch.einsum('pij,pj->pi', S, M1)
        a0, b, c0 = (M0 * v0).sum(1), (M0 * v1).sum(1), (M1 * v1).sum(1)
        det0 = a0 * c0 - b * b
        a, c = a0 + DILATION, c0 + DILATION
        det = a * c - b * b
        det_ok = det.detach() != 0
        det_s = torch.where(det_ok, det, torch.ones_like(det))
        if s.antialiasing:
            ratio = det0 / det_s
            floor_active = ratio.detach() <= AA_FLOOR
            h = torch.sqrt(torch.where(floor_active, torch.full_like(ratio, AA_FLOOR), ratio))
        else:
            h = torch.ones_like(det)
        conic = torch.stack([c / det_s, -b / det_s, a / det_s], dim=1)
        mid = 0.5 * (a + c)
        disc = torch.sqrt(torch.clamp(mid * mid - det, min=RADIUS_FLOOR))
        lam = torch.maximum(mid + disc, mid - disc)
        radius = torch.ceil(3.0 * torch.sqrt(lam.detach())).to(torch.int64)
        pix = torch.stack([((ndc[:, 0] + 1.0) * W - 1.0) * 0.5, ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5], dim=1)
        pix32 = pix.detach().to(torch.float32).to(torch.float64)  # reference rounds the pixel centre to fp32
        if not self.round_pixel_centre:
            pix32 = pix.detach()
        depth = p_view[:, 2]
        if self.structure is not None:  # integer side taken from the fp32 bit-model
            st = self.structure
            radius = torch.as_tensor(st["radius_raw"], dtype=torch.int64)
            rect = torch.as_tensor(st["rect"], dtype=torch.int64)
            depth_key = torch.as_tensor(st["depth_bits"], dtype=torch.int64)
            pix32 = torch.as_tensor(st["xy"], dtype=torch.float64)
        else:
            rf = radius.to(torch.float64)
            lo = torch.trunc((pix32 - rf[:, None]) / TILE)
            hi = torch.trunc((pix32 + rf[:, None] + (TILE - 1)) / TILE)
            g = torch.tensor([self.gx, self.gy], dtype=torch.float64)
            lo = torch.minimum(torch.clamp(lo, min=0), g)
            hi = torch.minimum(torch.clamp(hi, min=0), g)
            rect = torch.cat([lo, hi], dim=1).to(torch.int64)
            d32 = depth.detach().to(torch.float32)
            depth_key = torch.from_numpy(d32.numpy().view(np.int32).astype(np.int64))
        area = (rect[:, 2] - rect[:, 0]) * (rect[:, 3] - rect[:, 1])
        visible = in_front & det_ok & (area > 0)
        if self.structure is not None:
            visible = torch.as_tensor(self.structure["radii"], dtype=torch.int64) > 0
            area = torch.where(visible, area, torch.zeros_like(area))
        # pixel centre used by the blend: the fp32-rounded value, differentiable w.r.t. the fp64 one,
        # plus the means2D leaf in NDC units
        half = torch.tensor([0.5 * W, 0.5 * H], dtype=torch.float64)
        xy = pix + (pix32 - pix.detach()) + self.means2D[:, :2] * half
        # colour
        if self.colors_precomp is not None:
            rgb = self.colors_precomp
            clamped = torch.zeros(P, 3, dtype=torch.bool)
        else:
            d = m - self.campos[None, :]
            d = d / d.norm(dim=1, keepdim=True)
            basis = eval_sh_basis(int(s.sh_degree), d)
            n = basis.shape[1]
            raw = torch.einsum('pk,pkc->pc', basis, self.shs[:, :n, :]) + 0.5
            clamped = raw.detach() < 0
            rgb = torch.where(clamped, torch.zeros_like(raw), raw)
        opacity_eff = self.opacities.view(-1) * h
        self._pre = dict(depth=depth, xy=xy, conic=conic, opacity_eff=opacity_eff, rgb=rgb, clamped=clamped,
                         radius=radius, rect=rect, area=area, visible=visible, depth_key=depth_key,
                         invz=1.0 / tz)

    # ------------------------------------------------------------ K2-K5
    def _bin(self):
        pre = self._pre
        vis = pre["visible"]
        ids = torch.nonzero(vis).view(-1)
        rect = pre["rect"][ids]
        w = rect[:, 2] - rect[:, 0]
        cnt = pre["area"][ids]
        R = int(cnt.sum())
        rep = torch.repeat_interleave(torch.arange(ids.numel()), cnt)
        start = torch.cumsum(cnt, 0) - cnt
        k = torch.arange(R) - start[rep]
        ty = rect[rep, 1] + k // w[rep]
        tx = rect[rep, 0] + k % w[rep]
        tile = ty * self.gx + tx
        gid = ids[rep]
        key = (tile << 32) | pre["depth_key"][gid]
        order = torch.argsort(key, stable=True)
        self.point_list = gid[order]
        self.point_tiles = tile[order]
        self.sorted_keys = key[order]
        T = self.gx * self.gy
        counts = torch.bincount(self.point_tiles, minlength=T)
        ends = torch.cumsum(counts, 0)
        starts = ends - counts
        ranges = torch.stack([starts, ends], dim=1)
        ranges[counts == 0] = 0
        self.ranges = ranges
        self.R = R

    # --------------------------------------------------------------- K6
    def _blend_tile(self, t: int, g2d: dict, want_counts: bool = True):
        W, H = self.W, self.H
        tx, ty = t % self.gx, t // self.gx
        s, e = int(self.ranges[t, 0]), int(self.ranges[t, 1])
        ids = self.point_list[s:e]
        L = ids.numel()
        px = (tx * TILE + torch.arange(TILE)).repeat(TILE)
        py = (ty * TILE + torch.arange(TILE)).repeat_interleave(TILE)
        inside = (px < W) & (py < H)
        npx = TILE * TILE
        pxf, pyf = px.to(torch.float64), py.to(torch.float64)
        out = {}
        if L == 0:
            zc = torch.zeros(npx, 3, dtype=torch.float64)
            out.update(C=zc, D=torch.zeros(npx, dtype=torch.float64), T=torch.ones(npx, dtype=torch.float64),
                       n_contrib=torch.zeros(npx, dtype=torch.int64), fragile=torch.zeros(npx, dtype=torch.bool),
                       inside=inside, px=px, py=py, n_test=0, n_acc=0, n_test_bwd=0)
            return out
        xy = g2d["xy"][ids]
        con = g2d["conic"][ids]
        op = g2d["opacity_eff"][ids]
        rgb = g2d["rgb"][ids]
        invz = g2d["invz"][ids]
        dx = xy[:, 0:1] - pxf[None, :]
        dy = xy[:, 1:2] - pyf[None, :]
        power = -0.5 * (con[:, 0:1] * dx * dx + con[:, 2:3] * dy * dy) - con[:, 1:2] * dx * dy
        G = torch.exp(power)
        a_raw = op[:, None] * G
        alpha = a_raw + (torch.clamp(a_raw, max=ALPHA_MAX) - a_raw).detach()  # straight-through clamp
        pw = power.detach()
        skip = (pw > 0) | (alpha.detach() < ALPHA_MIN)
        a_eff = torch.where(skip, torch.zeros_like(alpha), alpha)
        T_incl = torch.cumprod(1.0 - a_eff, dim=0)
        T_excl = torch.cat([torch.ones(1, npx, dtype=torch.float64), T_incl[:-1]], dim=0)
        stop_cond = (~skip) & (T_incl.detach() < T_EPS)
        idx = torch.arange(L)[:, None]
        big = torch.full((1, npx), L, dtype=torch.int64)
        stop_idx = torch.where(stop_cond, idx.expand(L, npx), big.expand(L, npx)).min(dim=0).values  # [npx]
        active = (~skip) & (idx < stop_idx[None, :]) & inside[None, :]
        wgt = a_eff * T_excl * active
        C = torch.einsum('lp,lc->pc', wgt, rgb)
        D = (wgt * invz[:, None]).sum(0)
        # final T = T after the last applied Gaussian
        gather_idx = torch.clamp(stop_idx, max=L - 1)[None, :]
        T_stop = torch.gather(T_excl, 0, gather_idx)[0]
        T_end = T_incl[L - 1]
        T_fin = torch.where(stop_idx < L, T_stop, T_end)
        T_fin = torch.where(inside, T_fin, torch.ones_like(T_fin))
        pos = (idx + 1).expand(L, npx)
        n_contrib = torch.where(active, pos, torch.zeros_like(pos)).max(dim=0).values
        out.update(C=C, D=D, T=T_fin, n_contrib=n_contrib, inside=inside, px=px, py=py)
        if want_counts:
            tested = (idx <= stop_idx[None, :]) & inside[None, :]
            rel_a = (a_raw.detach() / ALPHA_MIN - 1.0).abs() < 2e-5
            rel_t = ((T_incl.detach() / T_EPS - 1.0).abs() < 2e-4) & (~skip)
            rel_p = (pw.abs() < 1e-6) & (op[:, None].detach() >= ALPHA_MIN)
            rel_c = ((a_raw.detach() / ALPHA_MAX - 1.0).abs() < 2e-6)
            frag = ((rel_a | rel_t | rel_p | rel_c) & tested).any(dim=0)
            out["fragile"] = frag
            out["n_test"] = int(tested.sum())
            out["n_acc"] = int(active.sum())
            out["n_test_bwd"] = int(((idx < n_contrib[None, :]) & inside[None, :]).sum())
        return out

    def forward(self) -> ForwardResult:
        self._preprocess()
        self._bin()
        pre = self._pre
        W, H = self.W, self.H
        if self.need_grad:
            self._g2d = {k: pre[k].detach().clone().requires_grad_(True) for k in ("xy", "conic", "opacity_eff", "rgb", "invz")}
        else:
            self._g2d = {k: pre[k].detach() for k in ("xy", "conic", "opacity_eff", "rgb", "invz")}
        color = torch.zeros(3, H, W, dtype=torch.float64)
        invd = torch.zeros(1, H, W, dtype=torch.float64)
        final_T = torch.ones(H, W, dtype=torch.float64)
        n_contrib = torch.zeros(H, W, dtype=torch.int32)
        fragile = torch.zeros(H, W, dtype=torch.bool)
        n_test = n_acc = n_test_bwd = 0
        for t in self._tiles():
            with torch.no_grad():
                o = self._blend_tile(t, self._g2d)
            ins = o["inside"]
            px, py = o["px"][ins], o["py"][ins]
            color[:, py, px] = (o["C"][ins] + o["T"][ins][:, None] * self.bg[None, :]).T
            invd[0, py, px] = o["D"][ins]
            final_T[py, px] = o["T"][ins]
            n_contrib[py, px] = o["n_contrib"][ins].to(torch.int32)
            fragile[py, px] = o["fragile"][ins]
            n_test += o["n_test"]
            n_acc += o["n_acc"]
            n_test_bwd += o["n_test_bwd"]
        if self.tile_window is not None:  # outside the window: background only (declared sample)
            pass
        radii = torch.where(pre["visible"], pre["radius"], torch.zeros_like(pre["radius"])).to(torch.int32)
        self.result = ForwardResult(
            color=color, invdepth=invd, radii=radii, final_T=final_T, n_contrib=n_contrib, fragile=fragile,
            point_list=self.point_list, point_tiles=self.point_tiles, ranges=self.ranges,
            tiles_touched=pre["area"], rect=pre["rect"], depth=pre["depth"].detach(), xy=pre["xy"].detach(),
            conic=pre["conic"].detach(), opacity_eff=pre["opacity_eff"].detach(), rgb=pre["rgb"].detach(),
            clamped=pre["clamped"], n_test=n_test, n_acc=n_acc, n_test_bwd=n_test_bwd,
            aux=dict(R=self.R, sorted_keys=self.sorted_keys))
        return self.result

    def _tiles(self):
        if self.tile_window is None:
            return range(self.gx * self.gy)
        x0, y0, x1, y1 = self.tile_window
        return [ty * self.gx + tx for ty in range(y0, min(y1, self.gy)) for tx in range(x0, min(x1, self.gx))]

    # ----------------------------------------------------------- K7-K9
    def backward(self, grad_color: torch.Tensor, grad_invdepth: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        assert self.need_grad and self.result is not None
        gc = grad_color.detach().to(torch.float64).cpu()
        gd = None if grad_invdepth is None else grad_invdepth.detach().to(torch.float64).cpu()
        g2d = self._g2d
        for v in g2d.values():
            v.grad = None
        for t in self._tiles():
            if int(self.ranges[t, 1]) == int(self.ranges[t, 0]):
                continue
            o = self._blend_tile(t, g2d, want_counts=False)
            ins = o["inside"]
            px, py = o["px"], o["py"]
            pxc, pyc = torch.clamp(px, max=self.W - 1), torch.clamp(py, max=self.H - 1)
            g_c = gc[:, pyc, pxc].T * ins[:, None]
            out_c = o["C"] + o["T"][:, None] * self.bg[None, :]
            loss = (out_c * g_c).sum()
            if gd is not None:
                loss = loss + (o["D"] * (gd[0, pyc, pxc] * ins)).sum()
            loss.backward()
        pre = self._pre
        outs, cots = [], []
        for k in ("xy", "conic", "opacity_eff", "rgb", "invz"):
            if g2d[k].grad is not None and pre[k].requires_grad:
                outs.append(pre[k])
                cots.append(g2d[k].grad)
        if outs:
            torch.autograd.backward(outs, cots)
        z = lambda t: None if t is None else (t.grad if t.grad is not None else torch.zeros_like(t))
        self.inter_grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in g2d.items()}
        return dict(means3D=z(self.means3D), means2D=z(self.means2D), shs=z(self.shs),
                    colors_precomp=z(self.colors_precomp), opacities=z(self.opacities), scales=z(self.scales),
                    rotations=z(self.rotations), cov3D_precomp=z(self.cov3D_precomp))


def mark_visible(means3D: torch.Tensor, viewmatrix: torch.Tensor) -> torch.Tensor:
    """K10 checkFrustum: only the near-plane test survives upstream (A.2)."""
    m = means3D.detach().to(torch.float64).cpu()
    V = viewmatrix.detach().to(torch.float64).cpu()
    z = m @ V[:3, 2] + V[3, 2]
    return z > NEAR_Z

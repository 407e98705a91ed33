"""`SparseGaussianAdam`: the optimiser class the reference imports from the rasterizer package when it is available
(/root/reference/train.py:47-52, scene/gaussian_model.py:25-28, 196-198) and steps with
`optimizer.step(visible, N)` (/root/reference/train.py:332-334).  Host side only: the arithmetic is
pdgs_sparse_adam in libpdgs.so (one launch for all parameter groups); there is no CPU path."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


def sparse_adam_update(groups, visible, N, beta1=0.9, beta2=0.999, eps=1e-15, step=0):
    """groups: list of (param, grad, exp_avg, exp_avg_sq, lr) fp32 CUDA tensors with numel == N * M each.
    visible: bool/uint8 [N] tensor, or an int32 radii tensor (visible = radii > 0)."""
    if not groups:
        return
    lib = L.lib()
    dev = groups[0][0].device
    if not groups[0][0].is_cuda:
        raise L.PdgsError("SparseGaussianAdam needs CUDA tensors (no CPU path)")
    N = int(N)
    vis_ptr = rad_ptr = None
    if visible.dtype == torch.int32:
        rad_ptr = C.c_void_p(visible.contiguous().data_ptr())
    elif visible.dtype in (torch.bool, torch.uint8):
        visible = visible.contiguous()
        vis_ptr = C.c_void_p(visible.data_ptr())
    else:
        raise L.PdgsError("visibility must be a bool / uint8 mask or int32 radii")
    if visible.numel() != N or visible.device != dev:
        raise L.PdgsError("visibility must have N elements on the parameters' device")
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for i0 in range(0, len(groups), 8):
            chunk = groups[i0:i0 + 8]
            arr = (L.PdgsAdamGroup * len(chunk))()
            for a, (p, g, m, v, lr) in zip(arr, chunk):
                for t in (p, g, m, v):
                    if t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev or t.numel() != p.numel():
                        raise L.PdgsError("SparseGaussianAdam: tensors must be contiguous fp32 on one device")
                if N == 0 or p.numel() % N:
                    raise L.PdgsError("SparseGaussianAdam: parameter numel is not a multiple of N")
                a.param, a.grad, a.exp_avg, a.exp_avg_sq = p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr()
                a.lr, a.M = float(lr), p.numel() // N
            L.check(lib.pdgs_sparse_adam(h, arr, len(chunk), vis_ptr, rad_ptr, N, float(beta1), float(beta2), float(eps),
                                         int(step), stream), "pdgs_sparse_adam")


class SparseGaussianAdam(torch.optim.Adam):
    """Adam that only touches the rows of visible Gaussians.  Same constructor and `step(visibility, N)` as the class
    the reference expects; state keys (`step`, `exp_avg`, `exp_avg_sq`) are torch.optim.Adam's, so the reference's
    densification code that edits optimiser state (scene/gaussian_model.py:330-420) keeps working.

    bias_correction=False (default) is the upstream sparse rule; True applies torch.optim.Adam's bias correction with
    a global step count (an extension)."""

    def __init__(self, params, lr, eps, bias_correction=False):
        super().__init__(params=params, lr=lr, eps=eps)
        self.bias_correction = bool(bias_correction)
        self._global_step = 0

    @torch.no_grad()
    def step(self, visibility, N):
        self._global_step += 1
        batch = {}
        for group in self.param_groups:
            assert len(group["params"]) == 1, "more than one tensor in group"
            param = group["params"][0]
            if param.grad is None:
                continue
            state = self.state[param]
            if len(state) == 0:
                state["step"] = torch.tensor(0.0, dtype=torch.float32)
                state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
            b1, b2 = group["betas"]
            key = (float(b1), float(b2), float(group["eps"]))
            batch.setdefault(key, []).append((param.data, param.grad.contiguous(), state["exp_avg"], state["exp_avg_sq"],
                                              group["lr"]))
        for (b1, b2, eps), groups in batch.items():
            sparse_adam_update(groups, visibility, N, b1, b2, eps, self._global_step if self.bias_correction else 0)

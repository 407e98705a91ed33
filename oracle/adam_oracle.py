"""TEST INFRASTRUCTURE (oracle): CPU restatement of the visibility-sparse Adam step.  Imported only by tests/,
__graft_entry__.smoke() and bench.py's CPU legs -- never by the product path.

What it restates
  * the reference's optimiser call sites: `gaussians.optimizer.step(visible, radii.shape[0])`
    (/root/reference/train.py:332-334) on the six parameter groups built at
    /root/reference/scene/gaussian_model.py:186-199 (`SparseGaussianAdam(l, lr=0.0, eps=1e-15)`), betas (0.9, 0.999);
  * the update rule of `SparseGaussianAdam` / `adamUpdate`, which lives in the accelerated branch of
    graphdeco-inria/diff-gaussian-rasterization (NOT in /root/reference: the submodule directory is empty), restated
    from its published form [UPSTREAM-RECALL]: for rows of visible Gaussians only
        m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;  p += -lr * m / (sqrt(v) + eps)        (no bias correction)
  * `step > 0`: torch.optim.Adam's bias-corrected rule (the reference's default optimiser,
    scene/gaussian_model.py:194-195).

Parity status: the bias-corrected mode is PINNED against torch.optim.Adam itself (tests/test_adam_oracle.py, all rows
visible); the sparse rule has no reference artefact in /root/reference to pin against -> "parity unpinned" for it.

All arithmetic is numpy float32 with every product / sum / quotient rounded separately, in the order the CUDA
kernel uses (csrc/pdgs_optim.cu adam_elem), so the kernel is compared BIT-EXACTLY."""
import numpy as np

f32 = np.float32


def sparse_adam_step(param, grad, exp_avg, exp_avg_sq, visible, lr, b1=0.9, b2=0.999, eps=1e-15, step=0):
    """In-place update of float32 arrays [N, ...]; `visible` bool [N].  Returns (param, exp_avg, exp_avg_sq)."""
    assert param.dtype == np.float32 and grad.dtype == np.float32
    N = param.shape[0]
    vis = np.asarray(visible).astype(bool).reshape(N)
    lr_d, b1_d, b2_d = float(lr), float(b1), float(b2)
    lr, b1, b2, eps = f32(lr), f32(b1), f32(b2), f32(eps)
    if step > 0:   # torch: host doubles rounded to fp32 once
        omb1, omb2 = f32(1.0 - b1_d), f32(1.0 - b2_d)
    else:          # upstream sparse kernel: (1.0f - b) in fp32
        omb1, omb2 = f32(1.0) - b1, f32(1.0) - b2
    g = grad[vis]
    m = (b1 * exp_avg[vis]).astype(f32) + (omb1 * g).astype(f32)
    v = (b2 * exp_avg_sq[vis]).astype(f32) + ((omb2 * g).astype(f32) * g).astype(f32)
    m, v = m.astype(f32), v.astype(f32)
    with np.errstate(divide="ignore", invalid="ignore"):
        if step > 0:
            step_size = f32(lr_d / (1.0 - b1_d ** step))
            bc2s = f32(np.sqrt(1.0 - b2_d ** step))
            denom = ((np.sqrt(v).astype(f32) / bc2s).astype(f32) + eps).astype(f32)
            upd = ((step_size * m).astype(f32) / denom).astype(f32)
        else:
            upd = ((lr * m).astype(f32) / (np.sqrt(v).astype(f32) + eps).astype(f32)).astype(f32)
    param[vis] = (param[vis] - upd).astype(f32)
    exp_avg[vis] = m
    exp_avg_sq[vis] = v
    return param, exp_avg, exp_avg_sq

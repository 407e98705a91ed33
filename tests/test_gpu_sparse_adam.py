"""GPU tests of the visibility-sparse Adam kernel (pdgs_sparse_adam via pdgs_b200.optim) against oracle/adam_oracle.py:
BIT-EXACT parameters and moments on the reference's six parameter groups (scene/gaussian_model.py:186-193), several
steps, ragged N, mask given as bool or as int32 radii; the class through `step(visible, N)` as train.py:332-334 calls
it; and the bias-corrected mode against torch.optim.Adam on the GPU."""
import numpy as np
import pytest
import torch

from oracle.adam_oracle import sparse_adam_step
from pdgs_b200.optim import SparseGaussianAdam, sparse_adam_update

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.025, "scaling": 5e-3, "rotation": 1e-3}


def make(N, seed):
    rng = np.random.default_rng(seed)
    return {k: rng.standard_normal((N,) + s).astype(np.float32) for k, s in SHAPES.items()}


@pytest.mark.parametrize("N,step_mode,use_radii", [(1000, 0, False), (4099, 0, True), (777, 1, False), (1, 0, False)])
def test_kernel_bit_exact_vs_oracle(N, step_mode, use_radii):
    P = make(N, 0)
    m = {k: np.zeros_like(v) for k, v in P.items()}
    v2 = {k: np.zeros_like(v) for k, v in P.items()}
    tp = {k: torch.from_numpy(v.copy()).to(DEV) for k, v in P.items()}
    tm = {k: torch.zeros_like(t) for k, t in tp.items()}
    tv = {k: torch.zeros_like(t) for k, t in tp.items()}
    rng = np.random.default_rng(7)
    for it in range(1, 5):
        G = {k: (g * 10.0 ** rng.uniform(-6, 0)).astype(np.float32) for k, g in make(N, 100 + it).items()}
        radii = (rng.random(N) < 0.6) * rng.integers(1, 50, N)
        vis = radii > 0
        step = it if step_mode else 0
        for k in P:
            sparse_adam_step(P[k], G[k], m[k], v2[k], vis, LRS[k], eps=1e-15, step=step)
        groups = [(tp[k], torch.from_numpy(G[k]).to(DEV), tm[k], tv[k], LRS[k]) for k in P]
        mask = torch.from_numpy(radii.astype(np.int32)).to(DEV) if use_radii else torch.from_numpy(vis).to(DEV)
        sparse_adam_update(groups, mask, N, 0.9, 0.999, 1e-15, step)
        torch.cuda.synchronize()
        for k in P:
            np.testing.assert_array_equal(tp[k].cpu().numpy().view(np.uint32), P[k].view(np.uint32), err_msg=f"param {k} it {it}")
            np.testing.assert_array_equal(tm[k].cpu().numpy().view(np.uint32), m[k].view(np.uint32), err_msg=f"m {k}")
            np.testing.assert_array_equal(tv[k].cpu().numpy().view(np.uint32), v2[k].view(np.uint32), err_msg=f"v {k}")


def test_class_step_like_train_py():
    N = 2500
    P = make(N, 3)
    params = {k: torch.nn.Parameter(torch.from_numpy(v.copy()).to(DEV)) for k, v in P.items()}
    l = [{"params": [params[k]], "lr": LRS[k], "name": k} for k in P]
    opt = SparseGaussianAdam(l, lr=0.0, eps=1e-15)
    m = {k: np.zeros_like(v) for k, v in P.items()}
    v2 = {k: np.zeros_like(v) for k, v in P.items()}
    rng = np.random.default_rng(5)
    for it in range(3):
        radii = torch.from_numpy(((rng.random(N) < 0.5) * 7).astype(np.int32)).to(DEV)
        G = make(N, 50 + it)
        for k in P:
            params[k].grad = torch.from_numpy(G[k]).to(DEV)
        visible = radii > 0
        opt.step(visible, radii.shape[0])
        opt.zero_grad(set_to_none=True)
        for k in P:
            sparse_adam_step(P[k], G[k], m[k], v2[k], visible.cpu().numpy(), LRS[k], eps=1e-15)
    for k in P:
        np.testing.assert_array_equal(params[k].detach().cpu().numpy().view(np.uint32), P[k].view(np.uint32))
        st = opt.state[params[k]]
        assert set(st) >= {"step", "exp_avg", "exp_avg_sq"}  # the keys densification edits (gaussian_model.py:330-420)
        np.testing.assert_array_equal(st["exp_avg"].cpu().numpy().view(np.uint32), m[k].view(np.uint32))


def test_bias_corrected_matches_torch_adam_on_gpu():
    N = 3001
    P = make(N, 9)
    a = {k: torch.nn.Parameter(torch.from_numpy(v.copy()).to(DEV)) for k, v in P.items()}
    b = {k: torch.nn.Parameter(torch.from_numpy(v.copy()).to(DEV)) for k, v in P.items()}
    oa = SparseGaussianAdam([{"params": [a[k]], "lr": LRS[k]} for k in P], lr=0.0, eps=1e-15, bias_correction=True)
    ob = torch.optim.Adam([{"params": [b[k]], "lr": LRS[k]} for k in P], lr=0.0, eps=1e-15)
    allvis = torch.ones(N, dtype=torch.bool, device=DEV)
    for it in range(4):
        G = make(N, 70 + it)
        for k in P:
            a[k].grad = torch.from_numpy(G[k] * 1e-3).to(DEV)
            b[k].grad = a[k].grad.clone()
        oa.step(allvis, N)
        ob.step()
    for k in P:
        d = float((a[k] - b[k]).abs().max())
        assert d <= 4e-6 * LRS[k] / 1e-4 + 4e-7 * float(b[k].abs().max()), (k, d)


def test_edge_shapes_unaligned_and_generic_row_length():
    """Row lengths without a compiled specialisation (M = 2, 7), tensors that start 4 bytes off a 16-byte boundary (scalar
    path), N = 0, and more than 8 groups (chunked launches): still bit-exact."""
    rng = np.random.default_rng(1)
    N = 1237
    vis = rng.random(N) < 0.5
    mask = torch.from_numpy(vis).to(DEV)
    groups, refs = [], []
    for gi, M in enumerate([2, 7, 1, 3, 4, 45, 48, 5, 9, 16]):
        p, g = rng.standard_normal((N, M)).astype(np.float32), rng.standard_normal((N, M)).astype(np.float32)
        m, v = np.abs(rng.standard_normal((N, M))).astype(np.float32) * 1e-3, np.abs(rng.standard_normal((N, M))).astype(np.float32) * 1e-6
        def dev(a, off):
            buf = torch.empty(a.size + 4, dtype=torch.float32, device=DEV)
            t = buf[off:off + a.size].view(a.shape)   # off = 1: 4 bytes past the 16-byte boundary
            t.copy_(torch.from_numpy(a))
            return t
        off = gi % 2
        tp, tg, tm, tv = dev(p, off), dev(g, off), dev(m, off), dev(v, off)
        lr = 10.0 ** rng.uniform(-4, -2)
        groups.append((tp, tg, tm, tv, lr))
        sparse_adam_step(p, g, m, v, vis, lr, eps=1e-15)
        refs.append((p, m, v))
    sparse_adam_update(groups, mask, N, 0.9, 0.999, 1e-15, 0)
    torch.cuda.synchronize()
    for (tp, tg, tm, tv, lr), (p, m, v) in zip(groups, refs):
        np.testing.assert_array_equal(tp.cpu().numpy().view(np.uint32), p.view(np.uint32))
        np.testing.assert_array_equal(tm.cpu().numpy().view(np.uint32), m.view(np.uint32))
        np.testing.assert_array_equal(tv.cpu().numpy().view(np.uint32), v.view(np.uint32))
    with pytest.raises(Exception):
        sparse_adam_update([(groups[0][0], groups[0][1], groups[0][2], groups[0][3], 1e-3)], mask.float(), N)  # bad mask dtype
    with pytest.raises(Exception):
        sparse_adam_update([(groups[0][0], groups[0][1], groups[0][2], groups[0][3], 1e-3)], mask[:-1], N)   # wrong length

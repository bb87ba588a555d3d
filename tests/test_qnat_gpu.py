"""GPU: QuantumNAT post-measurement kernels against oracle/qnat_oracle.py."""
import numpy as np
import pytest
import torch

import qris_b200
from helpers import assert_close
from oracle import qnat_oracle as QO

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,n", [(2, 1), (7, 3), (256, 6), (4099, 10), (65536, 10), (1000, 16)])
@pytest.mark.parametrize("normalise,quantise,sigma", [(True, True, 0.1), (True, False, 0.0), (False, True, 0.5)])
def test_qnat_forward_backward(B, n, normalise, quantise, sigma):
    g = torch.Generator().manual_seed(B + n)
    y = (torch.rand(B, n, generator=g) * 2 - 1) * 0.7
    noise = torch.randn(B, n, generator=g)
    gout = torch.randn(B, n, generator=g)
    lo, hi, levels = -2.0, 2.0, 5
    head = qris_b200.QuantumNATHead(sigma=sigma, lo=lo, hi=hi, levels=levels, normalise=normalise, quantise=quantise).cuda().train()
    yd = y.cuda().requires_grad_(True)
    out = head(yd, noise.cuda() if sigma else None)
    out.backward(gout.cuda())
    o_out, o_yn, o_mean, o_rstd = QO.qnat_forward(y, noise if sigma else None, sigma=sigma, lo=lo, hi=hi, levels=levels,
                                                   normalise=normalise, quantise=quantise)
    o_gy = QO.qnat_backward(gout, y, o_yn, o_mean, o_rstd, lo=lo, hi=hi, normalise=normalise, quantise=quantise)
    res = out.detach().cpu().double()
    if quantise:
        # H4: compare levels exactly, except where the oracle's pre-quantisation value is
        # within fp32 tolerance of a rounding edge or of the clamp bounds.
        step = (hi - lo) / (levels - 1)
        t = (o_yn.clamp(lo, hi) - lo) / step
        near_edge = (t - t.floor() - 0.5).abs() < 1e-4
        same = (res - o_out).abs() < 1e-6
        assert bool((same | near_edge).all()), f"{(~(same | near_edge)).sum()} quantised outputs differ away from edges"
        assert near_edge.float().mean() < 1e-3
        near_clamp = ((o_yn - lo).abs() < 1e-4) | ((o_yn - hi).abs() < 1e-4)
    else:
        assert_close(res.numpy(), o_out.numpy(), rel=2e-5, what="out")
        near_clamp = torch.zeros_like(o_yn, dtype=torch.bool)
    gy = yd.grad.cpu().double()
    scale = float(o_gy.abs().max())
    err = (gy - o_gy).abs()
    tol = 3e-5 * max(1.0, scale)
    if normalise and quantise:
        # a flipped STE mask near a clamp bound changes the batch sums by one term; allow for it
        tol += float(near_clamp.sum()) * scale / B
    bad = (err > tol) & ~near_clamp
    assert not bool(bad.any()), f"gy worst err {err.max():.3e} tol {tol:.3e}"


def test_eval_mode_has_no_noise_and_is_deterministic():
    head = qris_b200.QuantumNATHead(sigma=0.3).cuda().eval()
    y = torch.rand(128, 6, device="cuda")
    assert torch.equal(head(y), head(y))
    head.train()
    torch.manual_seed(5); a = head(y)
    torch.manual_seed(5); b = head(y)
    assert torch.equal(a, b), "noise must come from torch's generator so seeds reproduce"

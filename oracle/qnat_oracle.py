"""CPU oracle for the QuantumNAT post-measurement pipeline.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference has no post-measurement normalisation, noise or quantisation
(its "QuantumNAT" is the weight-space noise at Estimators_QuantumNAT_onchipQNN.py:176-196,
restated in ``weight_noise`` below).  BASELINE.json's north_star asks for the pipeline of the
QuantumNAT paper (Wang et al., DAC 2022); this file fixes its definition for this build
(SURVEY.md section 5.9):

1. normalise, per qubit, over the batch:  ``yh = (y - mean_b y) * rsqrt(var_b y + eps)`` with
   the unbiased variance (torch's default ``std``);
2. inject noise (training only):          ``yn = yh + sigma * noise`` -- ``noise`` is a tensor
   drawn by the caller from torch's generator, so seeds reproduce;
3. quantise to ``levels`` uniform levels on ``[lo, hi]``:
   ``q = lo + step * rint((clamp(yn, lo, hi) - lo) / step)``, ``step = (hi-lo)/(levels-1)``,
   round-half-to-even; straight-through gradient where ``lo <= yn <= hi``, zero outside.

Everything is evaluated in float64 unless ``dtype`` says otherwise.
"""
from __future__ import annotations

import torch


def qnat_forward(y, noise=None, sigma=0.0, lo=-2.0, hi=2.0, levels=5, eps=1e-10,
                 normalise=True, quantise=True, dtype=torch.float64):
    """Returns (out, yn, mean, rstd).  ``yn`` is the pre-quantisation value."""
    y = y.to(dtype)
    B = y.shape[0]
    if normalise:
        mean = y.mean(dim=0)
        var = ((y - mean) ** 2).sum(dim=0) / max(B - 1, 1)
        rstd = torch.rsqrt(var + eps)
        yh = (y - mean) * rstd
    else:
        mean = torch.zeros(y.shape[1], dtype=dtype)
        rstd = torch.ones(y.shape[1], dtype=dtype)
        yh = y
    yn = yh + sigma * noise.to(dtype) if (noise is not None and sigma != 0.0) else yh
    if quantise:
        step = (hi - lo) / (levels - 1)
        out = lo + step * torch.round((torch.clamp(yn, lo, hi) - lo) / step)
    else:
        out = yn
    return out, yn, mean, rstd


def qnat_backward(gout, y, yn, mean, rstd, lo=-2.0, hi=2.0, normalise=True, quantise=True,
                  dtype=torch.float64):
    """d loss / d y given d loss / d out (straight-through estimator + batch-norm backward)."""
    g = gout.to(dtype)
    if quantise:
        g = torch.where((yn >= lo) & (yn <= hi), g, torch.zeros_like(g))
    if not normalise:
        return g
    y = y.to(dtype)
    B = y.shape[0]
    yh = (y - mean) * rstd
    s1 = g.sum(dim=0)
    s2 = (g * yh).sum(dim=0)
    return rstd * (g - s1 / B - yh * s2 / max(B - 1, 1))


def qnat_autograd_check(y, gout, noise, sigma, lo, hi, levels, eps=1e-10):
    """Same backward obtained from torch autograd with an explicit STE, for cross-checking."""
    yr = y.to(torch.float64).clone().requires_grad_(True)
    B = yr.shape[0]
    mean = yr.mean(dim=0)
    var = ((yr - mean) ** 2).sum(dim=0) / max(B - 1, 1)
    yh = (yr - mean) * torch.rsqrt(var + eps)
    yn = yh + sigma * noise.to(torch.float64)
    step = (hi - lo) / (levels - 1)
    q = lo + step * torch.round((torch.clamp(yn, lo, hi) - lo) / step)
    inside = ((yn >= lo) & (yn <= hi)).to(torch.float64)
    out = yn * inside + (q - yn * inside).detach()      # STE: d out / d yn = inside
    out.backward(gout.to(torch.float64))
    return yr.grad


def weight_noise(weights: torch.Tensor, noise_level: float, generator=None) -> torch.Tensor:
    """Estimators_QuantumNAT_onchipQNN.py:186-188: ``w + noise_level * randn_like(w)``."""
    n = torch.randn(weights.shape, dtype=weights.dtype, generator=generator)
    return weights + noise_level * n

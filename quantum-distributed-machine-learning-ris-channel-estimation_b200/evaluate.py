"""Evaluation protocol of Test.py:64-275 on synthetic data.

Per SNR: fresh samples from every (scenario, user) pair; the scenario classifier's argmax routes each
sample to one of the three scenario-specific extractors (Test.py:167-214), the shared mapper produces
the channel estimate, NMSE is taken against the perfect channel (Test.py:221-243) next to the LS
estimate's NMSE, and classifier accuracy is recorded (Test.py:246-256).  The reference routes sample by
sample through Python lists; here routing is three masked batched calls.  Estimators are stock PyTorch.
"""
from __future__ import annotations

import math
from typing import Dict, Sequence

import torch

from . import data as D
from .estimators import NMSE_cuda


@torch.no_grad()
def routed_estimate(classifier, convs, fc, yp_img: torch.Tensor):
    """Predict the scenario, run each sample through the extractor of ITS predicted scenario, map."""
    pred = classifier(yp_img).argmax(dim=1)
    feats = None
    for s in range(3):
        idx = (pred == s).nonzero(as_tuple=True)[0]
        if idx.numel() == 0:
            continue
        f = convs[s](yp_img[idx])
        if feats is None:
            feats = torch.empty(yp_img.shape[0], f.shape[1], device=f.device, dtype=f.dtype)
        feats[idx] = f
    return fc(feats), pred


@torch.no_grad()
def snr_sweep(gen: D.SyntheticRIS, classifiers: Dict[str, torch.nn.Module], convs, fc,
              snrs: Sequence[float] = (5, 7, 9, 11, 13, 15), n_per_pair: int = 1000, batch: int = 200, seed: int = 60000):
    """Returns {"snr": [...], "LS": [dB...], "<name>": [dB...], "acc_<name>": [...]} (Test.py:259-336)."""
    for m in list(classifiers.values()) + list(convs) + [fc]:
        m.eval()
    out = {"snr": list(snrs), "LS": []}
    for name in classifiers:
        out[name], out["acc_" + name] = [], []
    for snr in snrs:
        num = {k: torch.zeros((), device=gen.device, dtype=torch.float64) for k in ["LS"] + list(classifiers)}
        den = torch.zeros((), device=gen.device, dtype=torch.float64)
        correct = {k: 0 for k in classifiers}
        total = 0
        for s in range(3):
            for u in range(3):
                yp, hls, hperf, ind = gen.pair_set(s, u, n_per_pair, snr, seed)
                for b0 in range(0, n_per_pair, batch):
                    sl = slice(b0, b0 + batch)
                    img = D.pilot_image(yp[sl], gen.P)
                    den += (hperf[sl].double() ** 2).sum()
                    num["LS"] += ((hls[sl].double() - hperf[sl].double()) ** 2).sum()
                    for name, clf in classifiers.items():
                        est, pred = routed_estimate(clf, convs, fc, img)
                        num[name] += ((est.double() - hperf[sl].double()) ** 2).sum()
                        correct[name] += int((pred == ind[sl]).sum())
                    total += img.shape[0]
        out["LS"].append(10 * math.log10(float(num["LS"] / den)))
        for name in classifiers:
            out[name].append(10 * math.log10(float(num[name] / den)))
            out["acc_" + name].append(correct[name] / total)
    return out

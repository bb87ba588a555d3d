"""Synthetic DeepMIMO-shaped data, resident on the device.

The reference reads ``available_data/{Yp,Hlabel,Hperf}{scenario}_{P}_1024_{SNR}dB_{uid}_datalen_{N}.npy``
through a ``generate_data`` module (Runner_P128_...py:16,48-95); neither the files nor the module are in
the repository.  This stands in with the same shapes and roles, inferred from the call sites
(SURVEY.md 3.4): per (scenario s in 0..2, user u in 0..2) a channel ``H`` of 64 antennas x 16
sub-carriers = 1024 complex entries built from a few steering-vector paths whose number, angular spread
and delay spread depend on the scenario (so scenarios are separable, as in the reference's figures);
``Yp = A H + noise`` where the P pilots observe a regular h x w sub-grid of the 64 x 16
(antenna, sub-carrier) plane (so the (2, h, w) pilot image the CNN front-ends see keeps the array
structure) scaled by fixed seeded unit-modulus pilot symbols; ``Hlabel`` the least-squares
(minimum-norm) estimate; ``Hperf = H``; label = scenario id (Run:58-61).
Everything stays on the GPU and batches are index selections: the loader never touches the host.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import torch

SCENARIOS = (  # (paths, angular spread [rad], max delay [fraction of symbol], centre of user 0 [rad])
    (2, 0.03, 0.02, -0.9),
    (4, 0.15, 0.10, 0.0),
    (8, 0.45, 0.30, 0.8),
)
N_ANT, N_SUB = 64, 16
N_H = N_ANT * N_SUB   # 1024


def pilot_shape(pilot_num: int) -> Tuple[int, int]:
    """(h, w) of the 2-channel pilot image; 16 x 8 for P = 128 is the reference's (Run:345)."""
    return {32: (8, 4), 64: (8, 8), 128: (16, 8), 256: (16, 16)}[pilot_num]


class SyntheticRIS:
    def __init__(self, pilot_num: int = 128, seed: int = 0, device="cuda"):
        self.P, self.device = pilot_num, torch.device(device)
        g = torch.Generator(device="cpu").manual_seed(seed)
        h, w = pilot_shape(pilot_num)
        ant = torch.arange(h) * (N_ANT // h)
        sub = torch.arange(w) * (N_SUB // w)
        cols = (ant[:, None] * N_SUB + sub[None, :]).reshape(-1)                 # observed entries of H, row-major (h, w)
        phase = torch.polar(torch.ones(pilot_num), 2 * math.pi * torch.rand(pilot_num, generator=g))
        A = torch.zeros(pilot_num, N_H, dtype=torch.complex64)
        A[torch.arange(pilot_num), cols] = phase                                 # unit-modulus pilot symbols
        self.A = A.to(self.device)
        self.A_pinv = A.conj().T.contiguous().to(self.device)                    # rows orthonormal: pinv = A^H, (1024, P)
        self.pilot_phase = phase.to(self.device)

    def channels(self, scenario: int, user: int, n: int, generator: torch.Generator) -> torch.Tensor:
        paths, spread, tau_max, centre = SCENARIOS[scenario]
        centre = centre + 0.25 * user
        dev = self.device
        theta = centre + spread * torch.randn(n, paths, generator=generator, device=dev)
        tau = tau_max * torch.rand(n, paths, generator=generator, device=dev)
        gain = torch.complex(torch.randn(n, paths, generator=generator, device=dev),
                             torch.randn(n, paths, generator=generator, device=dev)) / math.sqrt(2 * paths)
        k = torch.arange(N_ANT, device=dev, dtype=torch.float32)
        m = torch.arange(N_SUB, device=dev, dtype=torch.float32)
        steer = torch.polar(torch.ones(1, device=dev), math.pi * k[None, None, :] * torch.sin(theta)[..., None])   # (n, paths, 64)
        delay = torch.polar(torch.ones(1, device=dev), -2 * math.pi * m[None, None, :] * tau[..., None])           # (n, paths, 16)
        H = torch.einsum("np,npa,nps->nas", gain, steer, delay)
        return H.reshape(n, N_H).to(torch.complex64)

    def observe(self, H: torch.Tensor, snr_db: float, generator: torch.Generator):
        """Yp = A H + CN(0, sigma^2) at the given SNR; LS estimate = pinv(A) Yp."""
        y = H @ self.A.T
        sig_pow = (y.real ** 2 + y.imag ** 2).mean()
        sigma = torch.sqrt(sig_pow / (10 ** (snr_db / 10)) / 2)
        noise = torch.complex(torch.randn(y.shape, generator=generator, device=self.device),
                              torch.randn(y.shape, generator=generator, device=self.device)) * sigma
        yp = y + noise
        return yp * self.pilot_phase.conj(), yp @ self.A_pinv.T     # receiver removes the known pilot symbols

    def pair_set(self, scenario: int, user: int, n: int, snr_db: float, seed: int):
        """(Yp (n,P) complex, Hlabel (n,2048), Hperf (n,2048), indicator (n,)) for one (scenario, user) file."""
        g = torch.Generator(device=self.device).manual_seed(seed * 1000 + scenario * 10 + user)
        H = self.channels(scenario, user, n, g)
        yp, hls = self.observe(H, snr_db, g)
        as_real = lambda z: torch.cat([z.real, z.imag], dim=1).float()
        ind = torch.full((n,), scenario, dtype=torch.long, device=self.device)
        return yp, as_real(hls), as_real(H), ind


def pilot_image(yp: torch.Tensor, pilot_num: int) -> torch.Tensor:
    """cat([real, imag], dim=1) -> (B, 2, h, w) float32, as Run:344-345."""
    h, w = pilot_shape(pilot_num)
    return torch.cat([yp.real, yp.imag], dim=1).float().reshape(yp.shape[0], 2, h, w)


class NineWayLoader:
    """Device-resident stand-in for ``DataLoader(DatasetFolder_DML(...), batch_size, shuffle=True)``
    (Run:87-93): each iteration yields ``sutd[sid][uid] = (Yp, Hlabel, Hperf, indicator)`` batches."""

    def __init__(self, sets: Dict[Tuple[int, int], tuple], batch_size: int, shuffle: bool, seed: int = 0,
                 rank: int = 0, world: int = 1):
        self.sets, self.bs, self.shuffle = sets, batch_size, shuffle
        self.n = min(v[0].shape[0] for v in sets.values())
        self.rank, self.world = rank, world
        dev = next(iter(sets.values()))[0].device
        self.gen = torch.Generator(device=dev).manual_seed(seed)
        self.dev = dev

    def __len__(self):
        return (self.n + self.bs - 1) // self.bs

    def __iter__(self):
        perm = torch.randperm(self.n, generator=self.gen, device=self.dev) if self.shuffle else torch.arange(self.n, device=self.dev)
        for it in range(len(self)):
            idx = perm[it * self.bs:(it + 1) * self.bs]
            if self.world > 1:                               # data-parallel: this rank's slice of every batch
                per = -(-idx.numel() // self.world)
                idx = idx[self.rank * per:(self.rank + 1) * per]
            yield [[tuple(t[idx] for t in self.sets[(s, u)]) for u in range(3)] for s in range(3)]


def make_loaders(pilot_num=128, data_len=20000, snr_db=10.0, batch_size=256, seed=0, device="cuda", rank=0, world=1):
    """90/10 train/validation split of nine (scenario, user) sets (Run:67-84)."""
    gen = SyntheticRIS(pilot_num, seed=seed, device=device)
    n_train = int(0.9 * data_len)
    train, val = {}, {}
    for s in range(3):
        for u in range(3):
            full = gen.pair_set(s, u, data_len, snr_db, seed)
            train[(s, u)] = tuple(t[:n_train] for t in full)
            val[(s, u)] = tuple(t[n_train:] for t in full)
    return (NineWayLoader(train, batch_size, True, seed, rank, world),
            NineWayLoader(val, batch_size, False, seed, rank, world), gen)

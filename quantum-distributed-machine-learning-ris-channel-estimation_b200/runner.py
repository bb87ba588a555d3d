"""Training loops: the host-side mirror of ``Y2HRunner`` (Runner_P128_QuantumNAT_onchipQNN.py:19-426)
for the path this repository accelerates, on synthetic device-resident data (``data.py``).

``train_QSC_P128`` follows Run:307-426 step for step: nine forward calls (3 scenarios x 3 users) of one
batch each, ``loss = sum(nll_loss) / 9``, ONE backward, gradient pruning, AdamW(lr, weight_decay=0.01),
validation accuracy after every epoch, best / last checkpoints under the reference's file names.  What
differs is plumbing only: data never leaves the GPU, the optimiser tail is one kernel
(``optim.FlatPrunedAdamW``), the loss is read back once per epoch instead of once per iteration
(Run:370), and under ``torchrun`` every rank takes its slice of each batch with one gradient all-reduce
per step.  ``train_Conv_Linear_of_HDCE`` (Run:134-283) is stock PyTorch and kept short.
"""
from __future__ import annotations

import os
import time
from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import data as D
from .estimators import Conv_P128, FC_P128, NMSELoss, QSC_P128, SC_P128
from .optim import FlatPrunedAdamW


class GraphedQSCStep:
    """One optimiser step of the scenario classifier (``Y2HRunner.qsc_step``) as a CUDA graph.

    At the reference's sizes (nine batches of 256, 6 qubits) a step is about a hundred small kernels and is
    bound by launch and Python overhead, not by the GPU.  The first ``warmup`` calls run eagerly (they are
    real training steps); the next call records the step - batch copied into fixed input buffers, nine
    forwards, backward into the flat gradient bucket, pruning + AdamW with its step count on the device -
    and every later call copies its batch into the same buffers and replays the recording.  A batch of a
    different shape (the last, shorter one of an epoch) runs eagerly.  Single process only: the gradient
    all-reduce of a multi-GPU run is not recorded.
    """

    def __init__(self, runner, model, opt, concat_users: bool = False, warmup: int = 3):
        if opt.steps_done is None:
            raise ValueError("GraphedQSCStep needs FlatPrunedAdamW(device_step_count=True)")
        self.runner, self.model, self.opt, self.concat = runner, model, opt, concat_users
        self.warmup, self.calls = warmup, 0
        self.static = None          # fixed input buffers [sid][uid] -> tuple of tensors
        self.graph = None
        self.loss = None
        self.stream = torch.cuda.Stream(device=runner.device)

    def _shapes(self, sutd):
        return [tuple((t.shape, t.dtype) for t in sutd[s][u]) for s in range(3) for u in range(3)]

    def _load(self, sutd):
        for s in range(3):
            for u in range(3):
                for dst, src in zip(self.static[s][u], sutd[s][u]):
                    dst.copy_(src)

    def __call__(self, sutd):
        self.calls += 1
        if self.static is not None and self._shapes(sutd) != self._shapes(self.static):
            return self.runner.qsc_step(self.model, self.opt, sutd, self.concat)       # ragged batch: eager
        if self.static is None:
            self.static = [[tuple(t.clone() for t in sutd[s][u]) for u in range(3)] for s in range(3)]
        self._load(sutd)
        if self.graph is not None:
            self.graph.replay()
            return self.loss
        if self.calls <= self.warmup:
            # eager steps on the side stream the recording will use (allocator and library warm-up)
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                loss = self.runner.qsc_step(self.model, self.opt, self.static, self.concat)
            torch.cuda.current_stream().wait_stream(self.stream)
            return loss
        self.graph = torch.cuda.CUDAGraph()
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.loss = self.runner.qsc_step(self.model, self.opt, self.static, self.concat)
        self.graph.replay()          # recording does not execute: run the step it recorded
        return self.loss


def can_concat_users(model) -> bool:
    """Whether the nine user batches of one step may run as ONE model call with the same result: the model must
    treat samples independently (no batch statistics: BatchNorm in training mode, QuantumNATHead) and must not
    draw per-call randomness (the weight noise of Est:176-196 is drawn once per call, so nine calls see nine
    draws)."""
    from .qnat import QuantumNATHead
    for m in model.modules():
        if isinstance(m, (torch.nn.modules.batchnorm._BatchNorm, QuantumNATHead)):
            return False
    return not (getattr(model, "use_quantumnat", False) and getattr(model, "noise_level", 0.0) > 0)


class Y2HRunner:
    def __init__(self, Pilot_num=128, data_len=20000, SNRdb=10, batch_size_DML=256, lr=1e-3, n_epochs=100,
                 device=None, seed=0, workdir="./workspace", n_qubits=6, n_layers=3, verbose=True):
        self.Pilot_num, self.data_len, self.SNRdb = Pilot_num, data_len, SNRdb        # Run:20-34
        self.batch_size_DML, self.lr, self.n_epochs = batch_size_DML, lr, n_epochs
        self.seed, self.workdir, self.verbose = seed, workdir, verbose
        self.n_qubits, self.n_layers = n_qubits, n_layers
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        if device is None:
            device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        self.device = torch.device(device)
        self.train_QSC_losses, self.val_QSC_losses, self.val_QSC_accuracies = [], [], []   # Run:36-38

    def log(self, *a):
        if self.verbose and self.rank == 0:
            print(*a)

    def get_dataloader_DML(self, data_len: Optional[int] = None):
        return D.make_loaders(self.Pilot_num, data_len or self.data_len, self.SNRdb, self.batch_size_DML, self.seed,
                              self.device, self.rank, self.world)[:2]

    def _save_dir(self):
        d = os.path.join(self.workdir, f"Pn_{self.Pilot_num}", "HDCE")
        os.makedirs(d, exist_ok=True)       # the reference assumes it exists (Run:418)
        return d

    # ------------------------------------------------------------------------------------------
    def qsc_step(self, model, opt, sutd, concat_users: bool = False):
        """One optimiser step of Run:338-369.  Returns the (device) loss.

        ``concat_users``: run the nine equal-sized user batches as ONE model call.  The loss is the same
        number (a sum of nine equal-size means / 9 is the mean over the concatenation) with
        nine times fewer kernel launches; not valid for models that normalise over the batch."""
        opt.zero_grad()
        if concat_users:
            yp = torch.cat([sutd[s][u][0] for s in range(3) for u in range(3)])
            ind = torch.cat([sutd[s][u][3] for s in range(3) for u in range(3)])
            outputs = model(D.pilot_image(yp, self.Pilot_num))
            loss = F.nll_loss(outputs, torch.clamp(ind.long().reshape(-1), 0, outputs.size(-1) - 1))
            loss.backward()
            opt.step(local_rows=ind.numel() if self.world > 1 else None)
            return loss.detach()
        loss = 0
        for sid in range(3):
            for uid in range(3):
                yp, _, _, ind = sutd[sid][uid]
                outputs = model(D.pilot_image(yp, self.Pilot_num))                    # Run:344-352
                labels = torch.clamp(ind.long().reshape(-1), 0, outputs.size(-1) - 1)  # Run:346-360
                loss = loss + F.nll_loss(outputs, labels) / 9                          # Run:362
        loss.backward()                                                               # Run:364
        rows = sum(sutd[s][u][3].numel() for s in range(3) for u in range(3))
        opt.step(local_rows=rows if self.world > 1 else None)                         # Run:367-369 (prune + AdamW)
        return loss.detach()

    def train_QSC_P128(self, n_epochs: Optional[int] = None, use_quantumnat=False, use_gradient_pruning=False,
                       max_iters_per_epoch: Optional[int] = None, save=True, concat_users: Optional[bool] = None,
                       cuda_graph=False) -> Dict[str, list]:
        """``concat_users``: run the nine user batches of a step as one model call (``qsc_step``); the default
        ``None`` does so whenever that gives the same result (``can_concat_users``: no batch statistics, no
        per-call weight noise) -- 6.5x fewer kernel launches per step at the reference's sizes.
        ``cuda_graph``: record one optimiser step (nine forwards, backward, prune + AdamW) into a CUDA graph
        after three eager steps and replay it for every later batch of the same shape (single process only;
        see ``GraphedQSCStep``)."""
        torch.manual_seed(self.seed)
        model = QSC_P128(n_qubits=self.n_qubits, n_layers=self.n_layers, use_quantumnat=use_quantumnat,
                         use_gradient_pruning=use_gradient_pruning,
                         pilot_hw=D.pilot_shape(self.Pilot_num)).to(self.device)       # Run:313-316
        opt = FlatPrunedAdamW(model.parameters(), lr=self.lr, weight_decay=0.01,        # Run:320
                              prune_threshold=model.gradient_threshold if use_gradient_pruning else None,
                              device_step_count=bool(cuda_graph))
        if concat_users is None:
            concat_users = can_concat_users(model)
        graphed = GraphedQSCStep(self, model, opt, concat_users) if cuda_graph and self.world == 1 else None
        train_loader, val_loader = self.get_dataloader_DML()
        self.log("Data Loaded for QSC with QuantumNAT + On-chip QNN optimization!")
        best_acc, t0 = 0.0, time.time()
        self.train_QSC_losses = []
        for epoch in range(n_epochs or self.n_epochs):
            model.train()
            epoch_loss = torch.zeros((), device=self.device)
            n_batches = 0
            for it, sutd in enumerate(train_loader):
                if max_iters_per_epoch is not None and it >= max_iters_per_epoch:
                    break
                epoch_loss += graphed(sutd) if graphed is not None else self.qsc_step(model, opt, sutd, concat_users)
                n_batches += 1
            avg = float(epoch_loss.item()) / max(n_batches, 1)          # one device->host read per epoch
            self.train_QSC_losses.append(avg)
            val_loss, acc = self.validate_QSC(model, val_loader)
            self.val_QSC_losses.append(val_loss)
            self.val_QSC_accuracies.append(acc)
            self.log(f"Epoch {epoch + 1}, Average Loss: {avg:.4f}, Validation Loss: {val_loss:.4f}, "
                     f"Validation Accuracy: {acc:.2%}")
            if save and self.rank == 0:
                d = self._save_dir()
                if acc > best_acc:                                                     # Run:417-421
                    torch.save(model.state_dict(), os.path.join(d, f"QSC_OPT_{self.batch_size_DML}_{self.SNRdb}dB_best_DML.pth"))
                if epoch == (n_epochs or self.n_epochs) - 1:                           # Run:423-426
                    torch.save(model.state_dict(), os.path.join(d, f"QSC_OPT_{self.batch_size_DML}_{self.SNRdb}dB_epoch{epoch}_DML.pth"))
            best_acc = max(best_acc, acc)
        self.log(f"Quantum Scenario Classifier training time: {(time.time() - t0) / 60:.2f} minutes")
        self.model_QSC = model
        return {"train_loss": self.train_QSC_losses, "val_loss": self.val_QSC_losses, "val_acc": self.val_QSC_accuracies}

    @torch.no_grad()
    def validate_QSC(self, model, val_loader):
        """Run:378-409: nine-way validation loss and argmax accuracy (summed over ranks)."""
        model.eval()
        stats = torch.zeros(4, device=self.device, dtype=torch.float64)   # loss sum, batches, correct, total
        for suvd in val_loader:
            batch_loss = 0
            for sid in range(3):
                for uid in range(3):
                    yp, _, _, ind = suvd[sid][uid]
                    out = model(D.pilot_image(yp, self.Pilot_num))
                    labels = ind.long().reshape(-1)
                    batch_loss = batch_loss + F.nll_loss(out, labels) / 9
                    stats[2] += (out.argmax(dim=1) == labels).sum()
                    stats[3] += labels.numel()
            stats[0] += batch_loss
            stats[1] += 1
        if self.world > 1:
            dist.all_reduce(stats)
        s = stats.tolist()
        return s[0] / max(s[1], 1), s[2] / max(s[3], 1)

    # ------------------------------------------------------------------------------------------
    def train_Conv_Linear_of_HDCE(self, n_epochs: int = 1, max_iters_per_epoch: Optional[int] = None, lr=1e-3):
        """Run:134-283 in brief: three scenario-specific extractors + one shared mapper, nine NMSE losses
        per iteration, Adam, learning rate halved every 30 epochs with a 1e-6 floor.  Stock PyTorch."""
        torch.manual_seed(self.seed + 1)
        hw = D.pilot_shape(self.Pilot_num)
        convs = [Conv_P128(hw).to(self.device) for _ in range(3)]
        fc = FC_P128(hw).to(self.device)
        params = [p for m in convs + [fc] for p in m.parameters()]
        opt = torch.optim.Adam(params, lr=lr)
        crit = NMSELoss()
        train_loader, _ = self.get_dataloader_DML()
        hist = []
        for epoch in range(n_epochs):
            if epoch > 0 and epoch % 30 == 0:
                for gp in opt.param_groups:
                    gp["lr"] = max(gp["lr"] / 2, 1e-6)
            tot, nb = torch.zeros((), device=self.device), 0
            for it, sutd in enumerate(train_loader):
                if max_iters_per_epoch is not None and it >= max_iters_per_epoch:
                    break
                opt.zero_grad()
                loss = 0
                for sid in range(3):
                    for uid in range(3):
                        yp, _, hperf, _ = sutd[sid][uid]
                        loss = loss + crit(fc(convs[sid](D.pilot_image(yp, self.Pilot_num))), hperf)
                loss.backward()
                if self.world > 1:
                    for p in params:
                        dist.all_reduce(p.grad)
                        p.grad /= self.world
                opt.step()
                tot += loss.detach() / 9
                nb += 1
            hist.append(float(tot.item()) / max(nb, 1))
            self.log(f"HDCE epoch {epoch + 1}: mean NMSE loss {hist[-1]:.4f}")
        self.convs, self.fc = convs, fc
        return hist

    def train_SC_P128(self, n_epochs: int = 1, max_iters_per_epoch: Optional[int] = None):
        """Classical scenario classifier trained like the QSC (for the comparison in Test.py:158)."""
        torch.manual_seed(self.seed + 2)
        model = SC_P128(D.pilot_shape(self.Pilot_num)).to(self.device)
        opt = torch.optim.AdamW(model.parameters(), lr=self.lr, weight_decay=0.01)
        train_loader, _ = self.get_dataloader_DML()
        for _ in range(n_epochs):
            for it, sutd in enumerate(train_loader):
                if max_iters_per_epoch is not None and it >= max_iters_per_epoch:
                    break
                opt.zero_grad()
                loss = 0
                for sid in range(3):
                    for uid in range(3):
                        yp, _, _, ind = sutd[sid][uid]
                        loss = loss + F.nll_loss(model(D.pilot_image(yp, self.Pilot_num)), ind) / 9
                loss.backward()
                opt.step()
        self.model_SC = model
        return model

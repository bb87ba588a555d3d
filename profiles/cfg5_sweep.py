#!/usr/bin/env python
"""BASELINE configs[4]: pilot-length sweep P = 32 / 64 / 128 / 256, end-to-end training + NMSE-vs-SNR parity.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 profiles/cfg5_sweep.py \
        [--data-len 20000] [--epochs 3] [--hdce-epochs 2] [--n-per-pair 1000] [--out profiles/r02_cfg5_sweep.json]

For every pilot length: train the quantum scenario classifier (runner.Y2HRunner.train_QSC_P128: the loop of
Runner_P128_...py:307-426, batch sharded over the N ranks, one gradient all-reduce per step) and the HDCE
estimators (Run:134-283, stock PyTorch) on data.SyntheticRIS, then run the evaluation protocol of
Test.py:64-275 (evaluate.snr_sweep, SNR 5...15 dB) twice: with the kernel-backed classifier and with the same
classifier driven by the CPU oracle (oracle/ is used here as the checker only).  The SNR points are spread over
the ranks.  Records per P: NMSE pairs in dB, |delta|, classifier accuracy of both, argmax agreement, epoch times.
"""
import argparse
import copy
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from helpers import oracle_ops  # noqa: E402
from oracle import qnn_oracle as O  # noqa: E402
from qris_b200 import data as D  # noqa: E402
from qris_b200 import evaluate, runner  # noqa: E402


class OracleQLayer(torch.nn.Module):          # checker: the circuit evaluated by the CPU oracle (complex128)
    def __init__(self, circuit, weights):
        super().__init__()
        self.circuit, self.weights = circuit, torch.nn.Parameter(weights.detach().double().cpu().clone())

    def forward(self, x):
        return O.simulate(oracle_ops(self.circuit), self.circuit.n_qubits, x.double(), self.weights.reshape(-1))


class OracleClassifier(torch.nn.Module):      # evaluates on the CPU, answers on the GPU
    def __init__(self, model, dev):
        super().__init__()
        self.twin = copy.deepcopy(model).cpu().double()
        self.twin.qlayer = OracleQLayer(model.qlayer.circuit, model.qlayer.weights)
        self.twin.eval()
        self.dev = dev

    def forward(self, img):
        with torch.no_grad():
            return self.twin(img.cpu().double()).float().to(self.dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--data-len", type=int, default=20000)       # Run:22
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--hdce-epochs", type=int, default=2)
    ap.add_argument("--n-per-pair", type=int, default=1000)
    ap.add_argument("--pilots", default="32,64,128,256")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_cfg5_sweep.json"))
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    snrs = (5, 7, 9, 11, 13, 15)                                  # Test.py:66
    record = {"config": "BASELINE configs[4]: pilot-length sweep, end-to-end training + NMSE-vs-SNR parity",
              "n_gpus": world, "data_len_per_pair": args.data_len, "batch_size_DML": 256, "qsc_epochs": args.epochs,
              "hdce_epochs": args.hdce_epochs, "n_per_pair_eval": args.n_per_pair, "snr_db": list(snrs),
              "gpu": torch.cuda.get_device_name(dev), "pilots": {}}
    for P in [int(p) for p in args.pilots.split(",")]:
        r = runner.Y2HRunner(Pilot_num=P, data_len=args.data_len, batch_size_DML=256, n_epochs=args.epochs, lr=5e-3,
                             device=dev, verbose=(rank == 0), workdir=os.path.join(ROOT, "gpurun_out", "cfg5_workspace"))
        torch.cuda.synchronize(); t0 = time.time()
        hist = r.train_QSC_P128(save=False)
        torch.cuda.synchronize(); t_qsc = time.time() - t0
        t0 = time.time()
        hd = r.train_Conv_Linear_of_HDCE(n_epochs=args.hdce_epochs)
        torch.cuda.synchronize(); t_hdce = time.time() - t0
        gen = D.SyntheticRIS(P, seed=0, device=dev)
        clfs = {"quantum": r.model_QSC, "quantum_oracle": OracleClassifier(r.model_QSC, dev)}
        mine = [s for i, s in enumerate(snrs) if i % world == rank]
        t0 = time.time()
        res = evaluate.snr_sweep(gen, clfs, r.convs, r.fc, snrs=mine, n_per_pair=args.n_per_pair, batch=200) if mine else None
        t_eval = time.time() - t0
        parts = [None] * world
        if world > 1:
            dist.all_gather_object(parts, res)
        else:
            parts = [res]
        if rank == 0:
            merged = {}
            for part in parts:
                if part:
                    for i, s in enumerate(part["snr"]):
                        merged[s] = {k: part[k][i] for k in part if k != "snr"}
            rows = [dict(snr=s, **merged[s]) for s in sorted(merged)]
            deltas = [abs(x["quantum"] - x["quantum_oracle"]) for x in rows]
            record["pilots"][str(P)] = {
                "pilot_image_hw": list(D.pilot_shape(P)), "qsc_train_loss": hist["train_loss"], "qsc_val_acc": hist["val_acc"],
                "qsc_seconds_per_epoch": t_qsc / args.epochs, "hdce_loss": hd, "hdce_seconds_per_epoch": t_hdce / args.hdce_epochs,
                "eval_seconds_this_rank": t_eval, "rows": rows, "max_abs_delta_db": max(deltas),
                "nmse_within_0p05_db": max(deltas) <= 0.05,
                "identical_accuracy": all(x["acc_quantum"] == x["acc_quantum_oracle"] for x in rows)}
            print(f"P={P}: max |dNMSE| = {max(deltas):.4f} dB, identical accuracy: {record['pilots'][str(P)]['identical_accuracy']}, "
                  f"QSC {t_qsc / args.epochs:.2f} s/epoch, HDCE {t_hdce / args.hdce_epochs:.2f} s/epoch", flush=True)
    if rank == 0:
        with open(args.out, "w") as f:
            json.dump(record, f, indent=1)
        print("wrote", args.out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

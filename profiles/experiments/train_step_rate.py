"""Experiment: optimiser steps per second of the reference's as-shipped training configuration
(n = 6, L = 3, nine user batches of 256; Runner_P128...py:20-34) through runner.qsc_step."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from qris_b200 import estimators as E, optim, runner
r = runner.Y2HRunner(data_len=2560, batch_size_DML=256, device="cuda", verbose=False)
loader, _ = r.get_dataloader_DML()
batches = list(loader)
for concat, graph in ((False, False), (True, False), (False, True), (True, True)):
    torch.manual_seed(0)
    model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).cuda().train()
    opt = optim.FlatPrunedAdamW(model.parameters(), lr=1e-3, weight_decay=0.01, device_step_count=graph)
    step = runner.GraphedQSCStep(r, model, opt, concat) if graph else (lambda b: r.qsc_step(model, opt, b, concat))
    for b in batches[:5]:
        step(b)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    n = 0
    for rep in range(5):
        for b in batches:
            step(b); n += 1
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"concat_users={concat} cuda_graph={graph}: {n / dt:.1f} optimiser steps/s = {n * 9 * 256 / dt / 1e3:.1f} k circuit-samples/s "
          f"(fwd+bwd+optimiser, whole QSC model); one reference epoch (71 steps) = {71 * dt / n:.2f} s")

import sys; sys.path.insert(0,'/root/repo')
import torch, qris_b200
from qris_b200 import runner
for lr in (1e-3, 5e-3):
    r = runner.Y2HRunner(data_len=1500, batch_size_DML=64, n_epochs=4, lr=lr, device="cuda", verbose=True)
    h = r.train_QSC_P128(save=False)

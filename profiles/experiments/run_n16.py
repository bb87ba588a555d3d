import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, qris_b200
n, L, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
c = qris_b200.u3cu3_circuit(n, L)
x = 2 * torch.rand(B, n, device="cuda") - 1; w = 6.28 * torch.rand(c.n_weights, device="cuda"); gy = torch.randn(B, n, device="cuda")
for _ in range(2):
    qris_b200.qnn_forward_raw(c, x, w); qris_b200.qnn_backward_raw(c, x, w, gy)
torch.cuda.synchronize()
d = c.describe()
import collections
names={0:'END',1:'ROT',2:'ROTX',4:'CUL',5:'CUP',6:'DIAG',7:'DIAGR',8:'XCHG',9:'ROT5',10:'ROTX5',11:'CHAIN'}
print(collections.Counter(names[o[0]] for o in d['ops']))

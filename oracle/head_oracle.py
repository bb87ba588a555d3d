"""CPU restatement of the classifier's angle head - TEST INFRASTRUCTURE ONLY.

Only ``tests/`` may import this module; the product (``quantum-distributed-machine-learning-ris-channel-
estimation_b200/head.py`` over ``csrc/head.cu``) never does.

What it restates: the last two stages of the reference's ``preprocess`` stack,
``nn.Linear(32*4*2, n_qubits)`` and ``nn.Tanh()`` (Estimators_QuantumNAT_onchipQNN.py:161-162), applied at
``Est:173`` (``x = self.preprocess(x)``).  This part of the reference IS runnable here (plain torch), so the
restatement is pinned against torch itself in ``tests/test_oracle.py::test_head_oracle_matches_torch``
(float64, forward and autograd gradients).
"""
import numpy as np


def head_forward(f, W, b):
    """x = tanh(f W^T + b); f (B, K), W (n, K), b (n,); float64."""
    f, W, b = (np.asarray(a, np.float64) for a in (f, W, b))
    return np.tanh(f @ W.T + b)


def head_backward(f, W, x, gx):
    """Gradients of sum(gx * x) w.r.t. f, W, b given the forward output x."""
    f, W, x, gx = (np.asarray(a, np.float64) for a in (f, W, x, gx))
    gz = gx * (1.0 - x * x)
    return gz @ W, gz.T @ f, gz.sum(axis=0)

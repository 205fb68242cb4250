"""TEST INFRASTRUCTURE ONLY (imported by tests/): CPU restatement of the optimizer the reference's train step uses,
src/solver.py:101-106 `torch.optim.Adadelta(params, lr, eps=1e-8)` stepped at solver.py:182.  The algorithm lives in
PyTorch (torch/optim/adadelta.py, `_single_tensor_adadelta`); `adadelta_step` restates it in numpy float32 and
tests/test_oracle_golden.py pins it against torch.optim.Adadelta itself on the CPU."""
import numpy as np


def adadelta_step(param, grad, square_avg, acc_delta, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0.0):
    """One in-place update of float32 arrays; returns nothing."""
    f = np.float32
    g = grad.astype(np.float32)
    if weight_decay != 0.0:
        g = g + f(weight_decay) * param
    square_avg *= f(rho)
    square_avg += f(1.0 - rho) * g * g
    std = np.sqrt(square_avg + f(eps))
    delta = np.sqrt(acc_delta + f(eps)) / std * g
    acc_delta *= f(rho)
    acc_delta += f(1.0 - rho) * delta * delta
    param -= f(lr) * delta

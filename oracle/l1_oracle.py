"""ORACLE (test infrastructure) — L1 pixel loss value and gradient.

Restates reference utils/loss.py:32-39: `sum(loss_weight * nn.L1Loss()(inputs, gt))` with
loss_weight = [w]  =>  w * mean|x - y|;  d/dx = w * sign(x - y) / numel, sign(0) = 0
(aten::l1_loss backward).  The reference's default ctor argument `lossweight=1` is an int and
crashes on current torch; the restatement takes the float the reference means.
Pinned by tests/golden/l1_golden.npz (reference utils.loss.L1Loss(1.0) run in the build container).
"""
import numpy as np


def l1_loss(pred: np.ndarray, tgt: np.ndarray, w: float = 1.0) -> np.float32:
    d = pred.astype(np.float32) - tgt.astype(np.float32)
    # torch mean on CPU accumulates pairwise in fp32; fp64 sum then cast is within 1 ulp of it
    return np.float32(w * np.abs(d).astype(np.float64).sum() / d.size)


def l1_grad(pred: np.ndarray, tgt: np.ndarray, w: float = 1.0, grad_scale: float = 1.0) -> np.ndarray:
    d = pred.astype(np.float32) - tgt.astype(np.float32)
    return (np.sign(d) * np.float32(w * grad_scale / d.size)).astype(np.float32)


# nn.MSELoss() — train.py:13, :233, :264 (the loss of a run without --enchant), default mean reduction:
# mean((x - y)^2);  d/dx = 2 (x - y) / numel.  Pinned against torch.nn.MSELoss itself in tests/test_oracle.py.
def mse_loss(pred: np.ndarray, tgt: np.ndarray, w: float = 1.0) -> np.float32:
    d = pred.astype(np.float32) - tgt.astype(np.float32)
    return np.float32(w * (d.astype(np.float64) ** 2).sum() / d.size)


def mse_grad(pred: np.ndarray, tgt: np.ndarray, w: float = 1.0, grad_scale: float = 1.0) -> np.ndarray:
    d = pred.astype(np.float32) - tgt.astype(np.float32)
    return (np.float32(2.0) * d * np.float32(w * grad_scale / d.size)).astype(np.float32)

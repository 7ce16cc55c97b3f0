"""Oracle: multi-bandwidth Gaussian-kernel MMD (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/Code/cgnn/utils/Loss.py:10-32.

Two forms are given:

* ``mmd_loss_literal``  - the reference's op graph, line by line, in torch
  (any dtype, autograd-capable): concat, X X^T, row norms, the signed weight
  matrix S = s s^T, seven exp/multiply/sum passes.  This is the truth the CUDA
  kernels are compared with (fp64) and the reference's own noise floor (fp32).
* ``mmd_blocks`` / ``mmd_grad_analytic`` - the block form the CUDA kernels use
  (PP + TT - 2 PT, direct differences, closed-form gradient);
  checked against autograd of the literal form in tests/test_oracle.py.
"""
import numpy as np
import torch

from . import GAMMAS


def mmd_loss_literal(xy_true: torch.Tensor, xy_pred: torch.Tensor) -> torch.Tensor:
    """Loss.py:12-32, literally.  Returns a 0-d tensor of xy_pred's dtype."""
    N = xy_pred.shape[0]                                     # Loss.py:14
    X = torch.cat([xy_pred, xy_true], 0)                     # Loss.py:16
    XX = X @ X.t()                                           # Loss.py:17
    X2 = (X * X).sum(1, keepdim=True)                        # Loss.py:18
    exponent = -2 * XX + X2 + X2.t()                         # Loss.py:19
    s1 = torch.full((N, 1), 1.0 / N, dtype=X.dtype)          # Loss.py:21
    s2 = -torch.full((N, 1), 1.0 / N, dtype=X.dtype)         # Loss.py:22
    s = torch.cat([s1, s2], 0)                               # Loss.py:23
    S = s @ s.t()                                            # Loss.py:24
    loss = 0                                                 # Loss.py:26
    for g in GAMMAS:                                         # Loss.py:28
        kernel_val = torch.exp(-g * exponent)                # Loss.py:29
        loss = loss + (S * kernel_val).sum()                 # Loss.py:30
    return loss


def mmd_loss_and_grad_autograd(xy_true: np.ndarray, xy_pred: np.ndarray, dtype=torch.float64):
    """Literal loss and d loss / d xy_pred by autograd."""
    t = torch.as_tensor(np.asarray(xy_true), dtype=dtype)
    p = torch.as_tensor(np.asarray(xy_pred), dtype=dtype).clone().requires_grad_(True)
    loss = mmd_loss_literal(t, p)
    (g,) = torch.autograd.grad(loss, p)
    return float(loss.detach()), g.numpy()


def _sqdist(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    d = a[:, None, :] - b[None, :, :]
    return (d * d).sum(-1)


def mmd_blocks(xy_true: np.ndarray, xy_pred: np.ndarray):
    """(PP, TT, PT) = full ordered N x N kernel sums over the 7 bandwidths (fp64)."""
    t = np.asarray(xy_true, np.float64)
    p = np.asarray(xy_pred, np.float64)
    out = []
    for a, b in ((p, p), (t, t), (p, t)):
        D = _sqdist(a, b)
        out.append(sum(np.exp(-g * D).sum() for g in GAMMAS))
    return tuple(out)


def mmd_loss_blocks(xy_true, xy_pred) -> float:
    N = np.asarray(xy_pred).shape[0]
    PP, TT, PT = mmd_blocks(xy_true, xy_pred)
    return (PP + TT - 2.0 * PT) / (N * N)


def mmd_grad_analytic(xy_true: np.ndarray, xy_pred: np.ndarray) -> np.ndarray:
    """d loss / d xy_pred, closed form:
    (4/N^2) sum_g g [ sum_j k(P_i,T_j)(P_i-T_j) - sum_j k(P_i,P_j)(P_i-P_j) ]."""
    t = np.asarray(xy_true, np.float64)
    p = np.asarray(xy_pred, np.float64)
    N = p.shape[0]
    Dpt = _sqdist(p, t)
    Dpp = _sqdist(p, p)
    Wpt = sum(g * np.exp(-g * Dpt) for g in GAMMAS)
    Wpp = sum(g * np.exp(-g * Dpp) for g in GAMMAS)
    grad = (Wpt.sum(1)[:, None] * p - Wpt @ t) - (Wpp.sum(1)[:, None] * p - Wpp @ p)
    return 4.0 / (N * N) * grad


# ---- random-feature ("Fourier") MMD: the use_Fast_MMD=True alternative (Loss.py:35-56) ---------------
def rp_literal(k, d, rng):
    """Loss.py:35-37: W = [2*gamma_i * N(0,1)[k,d] for the 7 gammas ; U(0, 2 pi)[7k,1]]^T -> [d+1, 7k].
    rng: numpy RandomState-like with standard_normal / uniform (draw order as the TF graph creates the ops)."""
    blocks = [2.0 * g * rng.standard_normal((k, d)) for g in GAMMAS]
    phase = rng.uniform(0.0, 2.0 * np.pi, (k * len(GAMMAS), 1))
    return np.concatenate([np.concatenate(blocks, 0), phase], 1).T


def fourier_mmd_literal(xy_true: torch.Tensor, xy_pred: torch.Tensor, wz: torch.Tensor, k: int) -> torch.Tensor:
    """Loss.py:39-56 with the random matrix wz [d+1, 7k] supplied by the caller."""
    N = xy_pred.shape[0]

    def f1(x):                                                   # Loss.py:39-45
        ones = torch.ones((N, 1), dtype=x.dtype)
        return torch.cos(torch.cat([x, ones], 1) @ wz)

    e1 = float(np.sqrt(2.0 / k)) * f1(xy_true).mean(0)           # Loss.py:53
    e2 = float(np.sqrt(2.0 / k)) * f1(xy_pred).mean(0)           # Loss.py:54
    return ((e1 - e2) ** 2).sum()                                # Loss.py:56

"""Oracle: pairwise generator GNN (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/Code/cgnn/GNN.py:
  :21-29   init          N(0, init_std^2) for every tensor, biases included
  :55-67   generator     hid = relu([X, e] W_in + b_in);  out_y = hid W_out + b_out
  :70-72   loss          MMD([X, Y], [X, out_y])   (cause column copied)
  :74-75   Adam.minimize over [W_in, b_in, W_out, b_out]
  :92-96   train loop    fresh noise each sess.run; fetched loss is pre-update
  :115-126 evaluate      plain mean of test_epochs forward losses
  :129-166 both directions; second call on m[:, [1, 0]]
  :196-207 score         (mean BA - mean AB) / (mean BA + mean AB)

Flat parameter layout shared with the CUDA engine (documented in DESIGN.md):
  theta = [ W_in row 0 (weights of X) (h) | W_in row 1 (weights of e) (h) | b_in (h) | W_out (h) | b_out (1) ]
Noise and initial parameters are always supplied by the caller (the reference
is unseeded), so a kernel and the oracle can be driven identically.
"""
import numpy as np
import torch

from .mmd import mmd_loss_literal
from .adam import TFAdam


def n_params(h: int) -> int:
    return 4 * h + 1


def unpack(theta, h):
    W_in = theta[: 2 * h].reshape(2, h)
    b_in = theta[2 * h: 3 * h]
    W_out = theta[3 * h: 4 * h].reshape(h, 1)
    b_out = theta[4 * h: 4 * h + 1]
    return W_in, b_in, W_out, b_out


def forward_torch(theta: torch.Tensor, x: torch.Tensor, e: torch.Tensor, h: int) -> torch.Tensor:
    """GNN.py:64-67.  x, e: [N,1]; returns out_y [N,1]."""
    W_in, b_in, W_out, b_out = unpack(theta, h)
    hid = torch.relu(torch.cat([x, e], 1) @ W_in + b_in)
    return hid @ W_out + b_out


def loss_torch(theta, data, e, h):
    """GNN.py:70-72: MMD(concat[X,Y], concat[X,out_y]); data [N,2] with col 0 the cause."""
    x, y = data[:, :1], data[:, 1:2]
    out_y = forward_torch(theta, x, e, h)
    return mmd_loss_literal(torch.cat([x, y], 1), torch.cat([x, out_y], 1)), out_y


def step_debug(theta0, data, e, h, dtype=torch.float64, wz=None, k=None):
    """One training evaluation: returns dict(loss, out_y, dloss_dout, grad) (no update).
    wz [3, 7k]: use the random-feature loss of GNN.py:69-70 (use_Fast_MMD) with that matrix."""
    th = torch.as_tensor(np.asarray(theta0), dtype=dtype).clone().requires_grad_(True)
    d = torch.as_tensor(np.asarray(data), dtype=dtype)
    ee = torch.as_tensor(np.asarray(e), dtype=dtype).reshape(-1, 1)
    if wz is None:
        loss, out_y = loss_torch(th, d, ee, h)
    else:
        from .mmd import fourier_mmd_literal
        x, y = d[:, :1], d[:, 1:2]
        out_y = forward_torch(th, x, ee, h)
        loss = fourier_mmd_literal(torch.cat([x, y], 1), torch.cat([x, out_y], 1),
                                   torch.as_tensor(np.asarray(wz), dtype=dtype), k)
    out_y.retain_grad()
    loss.backward()
    return dict(loss=float(loss.detach()), out_y=out_y.detach().numpy()[:, 0].copy(),
                dloss_dout=out_y.grad.numpy()[:, 0].copy(), grad=th.grad.numpy().copy())


def train(theta0, data, noise, h, lr=0.01, dtype=torch.float64, wz=None, k=None):
    """GNN.py:92-96.  noise: [train_epochs, N]; wz: None or [train_epochs, 3, 7k] (use_Fast_MMD).
    Returns (theta, losses[train_epochs], thetas per step)."""
    npdt = np.float64 if dtype == torch.float64 else np.float32
    theta = np.asarray(theta0, npdt).copy()
    opt = TFAdam(theta.size, lr, npdt)
    losses, trail = [], []
    for it in range(noise.shape[0]):
        r = step_debug(theta, data, noise[it], h, dtype, None if wz is None else wz[it], k)
        losses.append(r["loss"])
        theta = opt.step(theta, r["grad"].astype(npdt)).astype(npdt)
        trail.append(theta.copy())
    return theta, np.array(losses), trail


def evaluate(theta, data, noise, h, dtype=torch.float64, wz=None, k=None):
    """GNN.py:115-126.  noise: [test_epochs, N].  Returns (mean, per-epoch losses)."""
    th = torch.as_tensor(np.asarray(theta), dtype=dtype)
    d = torch.as_tensor(np.asarray(data), dtype=dtype)
    out = []
    with torch.no_grad():
        for it in range(noise.shape[0]):
            ee = torch.as_tensor(np.asarray(noise[it]), dtype=dtype).reshape(-1, 1)
            if wz is None:
                out.append(float(loss_torch(th, d, ee, h)[0]))
            else:
                from .mmd import fourier_mmd_literal
                x, y = d[:, :1], d[:, 1:2]
                out.append(float(fourier_mmd_literal(torch.cat([x, y], 1), torch.cat([x, forward_torch(th, x, ee, h)], 1),
                                                     torch.as_tensor(np.asarray(wz[it]), dtype=dtype), k)))
    avg = 0.0
    for s in out:                       # GNN.py:113,120: running sum then / test_epochs
        avg += s
    return avg / len(out), np.array(out)


def run_instance(theta0, data, noise_train, noise_eval, h, lr=0.01, dtype=torch.float64):
    """GNN.py:129-132 for one direction (column 0 of ``data`` is the cause)."""
    theta, _, _ = train(theta0, data, noise_train, h, lr, dtype)
    return evaluate(theta, data, noise_eval, h, dtype)[0]


def causal_score(scores_ab, scores_ba) -> float:
    """GNN.py:196-207."""
    ab, ba = np.mean(scores_ab), np.mean(scores_ba)
    return (ba - ab) / (ba + ab)

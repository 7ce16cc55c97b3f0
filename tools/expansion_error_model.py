"""CPU model (numpy, emulated fp32) of the graph kernel's distance arithmetic, to bound its error before any GPU run.

The kernel forms the squared distance of a generated row a and a column b as
    D = K_a + R_b - 2 a'.b,   a' = a - m,  K_a = |a'|^2 + 2 a'.m,  R_b = |b - m|^2
(m = column means of the generated matrix of the step; one FMA per coordinate instead of a subtract and an
FMA), falls back to direct differences for pairs with D < TH, and accumulates the gradient as a.S - A with
S = sum w, A = sum w b in fp32 runs of one observed + one generated tile.  This script replays that in emulated fp32
on the bundled 22-variable example (initial, collapsed state and a trained state) and prints the error of
dLoss/dPred against the fp64 oracle.  Usage: python tools/expansion_error_model.py [train_steps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pandas as pd

from oracle import GAMMAS, cgnn as oc, mmd
from cgnn_b200 import DirectedGraph

f32 = np.float32
TH = 2.0


def fl(x):
    return np.asarray(x, np.float64).astype(f32)


def dot_chains(a, B, n_chain=4):
    """a [d] fp32, B [n,d] fp32 -> sequential FMA chains in fp32 (products exact in fp64, one rounding per FMA)."""
    d = a.shape[0]
    ch = [np.zeros(B.shape[0], f32) for _ in range(n_chain)]
    for c in range(d):
        k = (c // 2) % (n_chain // 2) * 2 + (c & 1)          # packed pairs: lanes x/y of two FFMA2 chains
        ch[k] = fl(ch[k].astype(np.float64) + np.float64(a[c]) * B[:, c].astype(np.float64))
    s0, s1 = fl(ch[0].astype(np.float64) + ch[2]), fl(ch[1].astype(np.float64) + ch[3])
    return fl(s0.astype(np.float64) + s1)


def wsum64(D):
    D = D.astype(np.float64)
    return fl(sum(g * np.exp(-g * D) for g in GAMMAS))


def model_grad(P, T):
    N, d = P.shape
    m = fl(P.astype(np.float64).mean(0))
    ap = fl(P.astype(np.float64) - m)                       # a' rows (registers)
    K64 = (ap.astype(np.float64) ** 2).sum(1) + 2.0 * ap.astype(np.float64) @ m.astype(np.float64)
    K_hi = fl(K64); K_lo = fl(K64 - K_hi)
    R_P = fl(((P.astype(np.float64) - m) ** 2).sum(1))
    R_T = fl(((T.astype(np.float64) - m) ** 2).sum(1))
    G = np.zeros((N, d), f32)
    slow = 0
    for i in range(N):
        a = fl(ap[i].astype(np.float64) + m)
        tot = np.zeros(d, f32)
        for j0 in range(0, N, 32):
            S = f32(0); A = np.zeros(d, f32)
            for sign, X, R in ((1.0, T, R_T), (-1.0, P, R_P)):
                B = X[j0:j0 + 32]
                dot = dot_chains(ap[i], B)
                D = fl(fl(K_hi[i] - 2.0 * dot.astype(np.float64)).astype(np.float64) + R[j0:j0 + 32])
                D = fl(D.astype(np.float64) + K_lo[i])
                near = D < TH
                if near.any():
                    slow += int(near.sum())
                    bp = fl(B.astype(np.float64) - m)
                    df = fl(ap[i].astype(np.float64) - bp)
                    Dd = np.zeros(B.shape[0], f32)
                    for c in range(d):
                        Dd = fl(Dd.astype(np.float64) + df[:, c].astype(np.float64) ** 2)
                    D = np.where(near, Dd, D)               # (the kernel recomputes the whole pair of columns)
                w = wsum64(D) * f32(sign)
                for jj in range(B.shape[0]):
                    S = fl(np.float64(S) + w[jj])
                    A = fl(A.astype(np.float64) + np.float64(w[jj]) * B[jj].astype(np.float64))
            tot = fl(tot.astype(np.float64) + fl(a.astype(np.float64) * np.float64(S) - A.astype(np.float64)))
        G[i] = tot
    return G * f32(4.0 / (N * N)), slow


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 150
    ex = os.path.join(ROOT, "examples")
    df = pd.read_csv(os.path.join(ex, "Example_graph_numdata.csv"))
    dag = DirectedGraph(pd.read_csv(os.path.join(ex, "Example_graph_target.csv")))
    plan = oc.build_plan(dag)
    X = df[plan["list_nodes"]].values
    T = ((X - X.mean(0)) / X.std(0)).astype(f32)
    N, d = T.shape
    h = 20
    rng = np.random.default_rng(0)
    theta = (0.05 * rng.normal(size=oc.n_params(plan, h)))
    noise = rng.normal(size=(steps + 1, plan["n_noise"], N))
    from oracle.adam import TFAdam
    opt = TFAdam(theta.size, 0.01)
    for it in range(steps + 1):
        r = oc.step_debug(theta, T.astype(np.float64), noise[it], plan, h)
        if it in (0, 10, 30, steps):
            P = r["pred"].astype(f32)
            ref = mmd.mmd_grad_analytic(T, P)
            got, slow = model_grad(P, T)
            err = np.max(np.abs(got - ref)) / np.max(np.abs(ref))
            rowerr = np.max(np.abs(got - ref), 1) / np.max(np.abs(ref), 1)
            print("step %4d  loss %.5f  dL/dP rel err (max-norm) %.2e   worst row %.2e   direct-form pairs %d of %d"
                  % (it, r["loss"], err, rowerr.max(), slow, 2 * N * N), flush=True)
        theta = opt.step(theta, r["grad"])


if __name__ == "__main__":
    main()

"""Measures, on the GPU, the error of every exp_policy against the fp64 oracle (loss, dLoss/dyhat,
parameter gradient) at initialisation and at a trained state.  Run under gpurun; prints a table."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from util import relerr
from cgnn_b200 import Engine, synth
from oracle import gnn as og

eng = Engine(0)
for (n, h, steps) in [(1500, 30, 300), (500, 20, 400)]:
    xy = synth.make_pair(3, n)[0]
    rng = np.random.default_rng(0)
    theta0 = (0.05 * rng.normal(size=4 * h + 1)).astype(np.float32)
    noise = rng.normal(size=(steps, n)).astype(np.float32)
    for pol in range(4):
        got = eng.pairwise_step_debug(xy, noise, theta0, h_layer_dim=h, exp_policy=pol)
        rows = []
        for s in (0, steps // 2, steps - 1):
            th = theta0 if s == 0 else got["params"][s - 1]
            r = og.step_debug(th.astype(np.float64), xy.astype(np.float64), noise[s].astype(np.float64), h)
            rows.append((s, r["loss"], abs(got["loss"][s] - r["loss"]) / abs(r["loss"]),
                         relerr(got["dldy"][s], r["dloss_dout"]), relerr(got["grad"][s], r["grad"])))
        for s, l, el, eg, ep in rows:
            print("N=%d h=%d pol=%d step=%3d loss=%.5f  err: loss %.2e  dldy %.2e  grad %.2e" % (n, h, pol, s, l, el, eg, ep))
    # fp32 literal graph (the reference's own arithmetic) for comparison
    import torch
    for s in (0, steps - 1):
        th = theta0 if s == 0 else got["params"][s - 1]
        r64 = og.step_debug(th.astype(np.float64), xy.astype(np.float64), noise[s].astype(np.float64), h)
        r32 = og.step_debug(th, xy, noise[s], h, dtype=torch.float32)
        print("N=%d fp32-literal step=%3d  err: loss %.2e dldy %.2e grad %.2e" % (
            n, s, abs(r32["loss"] - r64["loss"]) / abs(r64["loss"]), relerr(r32["dloss_dout"], r64["dloss_dout"]),
            relerr(r32["grad"], r64["grad"])))

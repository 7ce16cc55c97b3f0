"""Oracle: graph generators CGNN and CGNN_confounders (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/Code/cgnn/CGNN.py:
  :55-57   node list / n_var from graph.get_list_nodes()
  :63-84   build order: repeated scan of list_nodes, emit a node once all parents are
           generated; per node W_in[|pa|+1,h], b_in[h], W_out[h,1], b_out[1];
           input = [generated parents (get_parents order) ..., own noise]
  :86-90   prediction matrix = generated columns in list_nodes order
  :95      loss = MMD(data, prediction)
  :97-99   Adam over all parameters (in creation order)
  :116-121 train, :140-153 evaluate (as GNN.py)
and /root/reference/Code/cgnn/CGNN_confounders.py:
  :57      nodes from graph.skeleton.get_list_nodes()
  :64-72   one shared N(0,1) column per skeleton edge (get_list_edges_without_duplicate)
  :84-96   input = [generated parents ..., own noise, shared noise of every skeleton
           neighbour (get_neighbors order) ...]; W_in[|pa|+|nb|+1, h]

A *plan* is a plain-data description of that build, shared in spirit with the
engine's graph descriptor: ``nodes`` in build order, each with
  col      output column (index in list_nodes order)
  parents  list of input columns (generated values)
  noise    list of noise-column ids: own noise first (id = col), then shared ones
           (id = n_var + skeleton-edge index)
Flat parameters: per node in build order  [W_in rows (n_in x h) | b_in (h) | W_out (h) | b_out (1)].
External noise layout: [steps, n_noise_cols, N].
"""
import numpy as np
import torch

from .mmd import mmd_loss_literal
from .adam import TFAdam


def build_plan(graph, confounders: bool = False):
    """graph: any object with the reference Graph API (get_list_nodes, get_parents[, skeleton])."""
    if confounders:
        skel = graph.skeleton
        list_nodes = skel.get_list_nodes()                         # CGNN_confounders.py:57
        skel_edges = skel.get_list_edges_without_duplicate()       # :64
        edge_id = {}
        for k, e in enumerate(skel_edges):                         # :67-72
            edge_id[(e[0], e[1])] = k
            edge_id[(e[1], e[0])] = k
    else:
        list_nodes = graph.get_list_nodes()                        # CGNN.py:55
        skel_edges = []
    n_var = len(list_nodes)
    col = {v: i for i, v in enumerate(list_nodes)}
    done, nodes = set(), []
    while len(done) < n_var:                                       # CGNN.py:63
        progressed = False
        for var in list_nodes:                                     # :65
            par = graph.get_parents(var)                           # :67
            if var not in done and set(par).issubset(done):        # :68-69
                noise = [col[var]]
                if confounders:
                    for nb in graph.skeleton.get_neighbors(var):   # CGNN_confounders.py:82,94-95
                        noise.append(n_var + edge_id[(nb, var)])
                nodes.append(dict(col=col[var], parents=[col[p] for p in par], noise=noise))
                done.add(var)
                progressed = True
        if not progressed:
            raise ValueError("graph is cyclic: the reference's build loop would never terminate")
    return dict(n_var=n_var, n_noise=n_var + len(skel_edges), nodes=nodes, list_nodes=list_nodes)


def n_params(plan, h):
    return sum((len(nd["parents"]) + len(nd["noise"])) * h + 2 * h + 1 for nd in plan["nodes"])


def forward_torch(theta, noise, plan, h):
    """noise: [n_noise_cols, N] tensor.  Returns prediction [N, n_var]."""
    cols = [None] * plan["n_var"]
    o = 0
    for nd in plan["nodes"]:
        n_in = len(nd["parents"]) + len(nd["noise"])
        W_in = theta[o:o + n_in * h].reshape(n_in, h); o += n_in * h
        b_in = theta[o:o + h]; o += h
        W_out = theta[o:o + h].reshape(h, 1); o += h
        b_out = theta[o:o + 1]; o += 1
        inp = [cols[p] for p in nd["parents"]] + [noise[c].reshape(-1, 1) for c in nd["noise"]]
        inp = torch.cat(inp, 1)                                    # CGNN.py:76-78
        out = torch.relu(inp @ W_in + b_in) @ W_out + b_out        # :80-81
        cols[nd["col"]] = out
    return torch.cat(cols, 1)                                      # :86-90


def step_debug(theta0, data, noise, plan, h, dtype=torch.float64, wz=None, k=None):
    th = torch.as_tensor(np.asarray(theta0), dtype=dtype).clone().requires_grad_(True)
    d = torch.as_tensor(np.asarray(data), dtype=dtype)
    nz = torch.as_tensor(np.asarray(noise), dtype=dtype)
    pred = forward_torch(th, nz, plan, h)
    pred.retain_grad()
    if wz is None:
        loss = mmd_loss_literal(d, pred)                           # CGNN.py:95
    else:
        from .mmd import fourier_mmd_literal
        loss = fourier_mmd_literal(d, pred, torch.as_tensor(np.asarray(wz), dtype=dtype), k)      # CGNN.py:93
    loss.backward()
    return dict(loss=float(loss.detach()), pred=pred.detach().numpy().copy(),
                dloss_dpred=pred.grad.numpy().copy(), grad=th.grad.numpy().copy())


def train(theta0, data, noise, plan, h, lr=0.01, dtype=torch.float64):
    """noise: [train_epochs, n_noise_cols, N]."""
    npdt = np.float64 if dtype == torch.float64 else np.float32
    theta = np.asarray(theta0, npdt).copy()
    opt = TFAdam(theta.size, lr, npdt)
    losses, trail = [], []
    for it in range(noise.shape[0]):
        r = step_debug(theta, data, noise[it], plan, h, dtype)
        losses.append(r["loss"])
        theta = opt.step(theta, r["grad"].astype(npdt)).astype(npdt)
        trail.append(theta.copy())
    return theta, np.array(losses), trail


def evaluate(theta, data, noise, plan, h, dtype=torch.float64):
    th = torch.as_tensor(np.asarray(theta), dtype=dtype)
    d = torch.as_tensor(np.asarray(data), dtype=dtype)
    out = []
    with torch.no_grad():
        for it in range(noise.shape[0]):
            nz = torch.as_tensor(np.asarray(noise[it]), dtype=dtype)
            out.append(float(mmd_loss_literal(d, forward_torch(th, nz, plan, h))))
    tot = 0.0
    for s in out:                                                  # CGNN.py:138,146
        tot += s
    return tot / len(out), np.array(out)


def generate(theta, noise, plan, h, dtype=torch.float64):
    """CGNN.py:155-160: one sampled [N, n_var] matrix."""
    with torch.no_grad():
        return forward_torch(torch.as_tensor(np.asarray(theta), dtype=dtype),
                             torch.as_tensor(np.asarray(noise), dtype=dtype), plan, h).numpy()


def graph_score(run_scores, n_edges=None, complexity=5e-5):
    """CGNN.py:217: mean over finite run scores; confounders add complexity*|E| (CGNN_confounders.py:237)."""
    s = np.mean([i for i in run_scores if np.isfinite(i)])
    if n_edges is not None:
        s += complexity * n_edges
    return s

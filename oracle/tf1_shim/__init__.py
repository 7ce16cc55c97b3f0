"""A tiny graph-mode TensorFlow-1 API emulator on torch (TEST INFRASTRUCTURE ONLY).

Purpose: execute the reference's OWN, UNMODIFIED model code (Code/cgnn/utils/Loss.py, GNN.py,
CGNN.py, CGNN_confounders.py) in a container where TensorFlow cannot be installed, so that the
oracle restatement is pinned against the reference's source rather than against a reading of it.
Only the TF-1 surface those four files use is provided:

  placeholder, Variable, random_normal, random_uniform, constant, ones, concat, matmul, transpose,
  reduce_sum(keep_dims=), reduce_mean, exp, cos, sqrt, nn.relu, float32, device, ConfigProto,
  Session(.run), global_variables_initializer, reset_default_graph, train.AdamOptimizer(.minimize)

Semantics kept from TF-1 graph mode: ops build a lazy graph; every `Session.run` evaluates each
node at most once (so a random_normal op yields ONE fresh draw per run, shared by all fetches of
that run, and a loss fetched together with the train op is the pre-update loss); Variables hold
state across runs; static shapes are inferred well enough for `get_shape().as_list()`.
What is NOT TensorFlow here: the numeric kernels (torch, in `DTYPE`, fp64 by default) and Adam,
which follows TF-1's documented update (oracle/adam.py).  Random draws come from `RNG` (a numpy
RandomState) and are logged in `DRAWS` so a test can replay them elsewhere.
"""
import contextlib
import sys
import types

import numpy as np
import torch

DTYPE = torch.float64
RNG = np.random.RandomState(0)
DRAWS = []          # (node id, phase, ndarray) in evaluation order; phase = "init" | "run"
RUN_LOG = []        # per Session.run: list of fetched values (python floats / ndarrays)
float32 = "float32"
_NEXT_ID = [0]


def reset(seed=0, dtype=torch.float64):
    global DTYPE, RNG
    DTYPE = dtype
    RNG = np.random.RandomState(seed)
    del DRAWS[:]
    del RUN_LOG[:]
    _NEXT_ID[0] = 0


class _Shape(object):
    def __init__(self, dims):
        self.dims = list(dims)

    def as_list(self):
        return list(self.dims)


class Node(object):
    def __init__(self, shape):
        self.shape = list(shape) if shape is not None else None
        self.id = _NEXT_ID[0]
        _NEXT_ID[0] += 1

    def get_shape(self):
        return _Shape(self.shape)

    def value(self, ctx):
        if self.id not in ctx.cache:
            ctx.cache[self.id] = self._eval(ctx)
        return ctx.cache[self.id]

    # python operators used by the reference
    def __add__(self, o): return _Bin(torch.add, self, o)
    def __radd__(self, o): return _Bin(torch.add, o, self)
    def __sub__(self, o): return _Bin(torch.sub, self, o)
    def __rsub__(self, o): return _Bin(torch.sub, o, self)
    def __mul__(self, o): return _Bin(torch.mul, self, o)
    def __rmul__(self, o): return _Bin(torch.mul, o, self)
    def __truediv__(self, o): return _Bin(torch.div, self, o)
    def __pow__(self, o): return _Bin(torch.pow, self, o)
    def __neg__(self): return _Un(torch.neg, self)


def _as_node(x):
    return x if isinstance(x, Node) else _Const(x)


def _bshape(a, b):
    if a is None or b is None:
        return a if b is None else b
    out = []
    for k in range(max(len(a), len(b))):
        da = a[len(a) - 1 - k] if k < len(a) else 1
        db = b[len(b) - 1 - k] if k < len(b) else 1
        out.append(db if da == 1 else da if db in (1, da) else (da if da is not None else db))
    return out[::-1]


class _Const(Node):
    def __init__(self, v, shape=None):
        arr = np.asarray(v, dtype=np.float64)
        if shape is not None:
            arr = np.broadcast_to(arr, shape).copy()
        super(_Const, self).__init__(list(arr.shape))
        self.arr = arr

    def _eval(self, ctx):
        return torch.as_tensor(self.arr, dtype=DTYPE)


class _Bin(Node):
    def __init__(self, fn, a, b):
        self.fn, self.a, self.b = fn, _as_node(a), _as_node(b)
        super(_Bin, self).__init__(_bshape(self.a.shape, self.b.shape))

    def _eval(self, ctx):
        return self.fn(self.a.value(ctx), self.b.value(ctx))


class _Un(Node):
    def __init__(self, fn, a, shape=None):
        self.fn, self.a = fn, _as_node(a)
        super(_Un, self).__init__(self.a.shape if shape is None else shape)

    def _eval(self, ctx):
        return self.fn(self.a.value(ctx))


class _Placeholder(Node):
    def _eval(self, ctx):
        if self not in ctx.feed:
            raise KeyError("placeholder was not fed")
        return torch.as_tensor(np.asarray(ctx.feed[self]), dtype=DTYPE)


class _Random(Node):
    def __init__(self, shape, draw):
        super(_Random, self).__init__(list(shape))
        self.draw = draw

    def _eval(self, ctx):
        arr = self.draw(tuple(self.shape))
        DRAWS.append((self.id, ctx.phase, arr.copy()))
        return torch.as_tensor(arr, dtype=DTYPE)


_VARIABLES = []


class _Variable(Node):
    def __init__(self, init):
        self.init = _as_node(init)
        super(_Variable, self).__init__(self.init.shape)
        self.tensor = None
        _VARIABLES.append(self)

    def _eval(self, ctx):
        if self.tensor is None:
            raise RuntimeError("variable used before global_variables_initializer")
        return self.tensor


class _Concat(Node):
    def __init__(self, nodes, axis):
        self.nodes, self.axis = [_as_node(n) for n in nodes], axis
        rank = len(self.nodes[0].shape)
        shape = []
        for d in range(rank):
            dims = [n.shape[d] for n in self.nodes]
            if d == axis:
                shape.append(None if any(x is None for x in dims) else sum(dims))
            else:
                known = [x for x in dims if x is not None]
                shape.append(known[0] if known else None)
        super(_Concat, self).__init__(shape)

    def _eval(self, ctx):
        return torch.cat([n.value(ctx) for n in self.nodes], self.axis)


class _MatMul(Node):
    def __init__(self, a, b):
        self.a, self.b = _as_node(a), _as_node(b)
        super(_MatMul, self).__init__([self.a.shape[0], self.b.shape[1]])

    def _eval(self, ctx):
        return self.a.value(ctx) @ self.b.value(ctx)


class _Reduce(Node):
    def __init__(self, fn, x, axis, keep):
        self.fn, self.x, self.axis, self.keep = fn, _as_node(x), axis, keep
        if axis is None:
            shape = []
        else:
            shape = list(self.x.shape)
            if keep:
                shape[axis] = 1
            else:
                del shape[axis]
        super(_Reduce, self).__init__(shape)

    def _eval(self, ctx):
        v = self.x.value(ctx)
        return self.fn(v) if self.axis is None else self.fn(v, self.axis, keepdim=self.keep)


class _Ctx(object):
    def __init__(self, feed, phase):
        self.feed, self.cache, self.phase = feed or {}, {}, phase


# ---- the TF-1 surface -----------------------------------------------------------------------------
def placeholder(dtype, shape=None):
    return _Placeholder(shape)


def Variable(init):
    return _Variable(init)


def random_normal(shape, mean=0.0, stddev=1.0):
    return _Random(shape, lambda s: mean + stddev * RNG.standard_normal(s))


def random_uniform(shape, minval=0.0, maxval=1.0):
    return _Random(shape, lambda s: RNG.uniform(minval, maxval, s))


def constant(value, shape=None):
    return _Const(value, shape)


def ones(shape):
    return _Const(1.0, tuple(shape))


def concat(values, axis):
    return _Concat(values, axis)


def matmul(a, b):
    return _MatMul(a, b)


def transpose(x):
    x = _as_node(x)
    return _Un(lambda t: t.t(), x, shape=list(x.shape[::-1]))


def reduce_sum(x, axis=None, keep_dims=False):
    return _Reduce(torch.sum, x, axis, keep_dims)


def reduce_mean(x, axis=None, keep_dims=False):
    return _Reduce(torch.mean, x, axis, keep_dims)


def exp(x): return _Un(torch.exp, x)
def cos(x): return _Un(torch.cos, x)
def sqrt(x): return _Un(torch.sqrt, x) if isinstance(x, Node) else float(np.sqrt(x))


nn = types.SimpleNamespace(relu=lambda x: _Un(torch.relu, x))


@contextlib.contextmanager
def device(name):
    yield


class ConfigProto(object):
    def __init__(self):
        self.gpu_options = types.SimpleNamespace(allow_growth=False)


class _InitOp(object):
    pass


def global_variables_initializer():
    return _InitOp()


def reset_default_graph():
    del _VARIABLES[:]


class _TrainOp(object):
    def __init__(self, loss, var_list, lr):
        self.loss, self.vars, self.lr = loss, list(var_list), lr
        self.m = self.v = None
        self.b1p = self.b2p = 1.0


class _Adam(object):
    """TF-1 AdamOptimizer with default betas/epsilon ("epsilon hat" form), see oracle/adam.py."""

    def __init__(self, learning_rate=0.001):
        self.lr = learning_rate

    def minimize(self, loss, var_list=None):
        return _TrainOp(loss, var_list, self.lr)


train = types.SimpleNamespace(AdamOptimizer=_Adam)


class Session(object):
    def __init__(self, config=None):
        pass

    def _train(self, op, ctx):
        loss = op.loss.value(ctx)
        grads = torch.autograd.grad(loss, [v.tensor for v in op.vars])
        if op.m is None:
            op.m = [torch.zeros_like(v.tensor) for v in op.vars]
            op.v = [torch.zeros_like(v.tensor) for v in op.vars]
        op.b1p *= 0.9
        op.b2p *= 0.999
        lr_t = op.lr * np.sqrt(1.0 - op.b2p) / (1.0 - op.b1p)
        new = []
        with torch.no_grad():
            for k, (var, g) in enumerate(zip(op.vars, grads)):
                op.m[k] = 0.9 * op.m[k] + 0.1 * g
                op.v[k] = 0.999 * op.v[k] + 0.001 * g * g
                new.append(var.tensor - lr_t * op.m[k] / (torch.sqrt(op.v[k]) + 1e-8))
        return new

    def run(self, fetches, feed_dict=None):
        if isinstance(fetches, _InitOp):
            ctx = _Ctx(None, "init")
            for v in _VARIABLES:
                if v.tensor is None:
                    v.tensor = v.init.value(ctx).detach().clone().requires_grad_(True)
            return None
        single = not isinstance(fetches, (list, tuple))
        items = [fetches] if single else list(fetches)
        ctx = _Ctx(feed_dict, "run")
        out, updates = [], []
        for f in items:                       # values first: everything sees the pre-update variables
            if isinstance(f, _TrainOp):
                updates.append((f, self._train(f, ctx)))
                out.append(None)
            else:
                v = f.value(ctx).detach()
                out.append(np.float64(v.item()) if v.dim() == 0 else v.numpy().copy())
        for op, new in updates:
            for var, t in zip(op.vars, new):
                var.tensor = t.detach().clone().requires_grad_(True)
        RUN_LOG.append([o for o in out if o is not None])
        return out[0] if single else out


def install():
    """Make `import tensorflow` resolve to this module (for loading the reference's files)."""
    sys.modules["tensorflow"] = sys.modules[__name__]
    return sys.modules[__name__]

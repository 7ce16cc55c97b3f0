"""Host-side handle on the CUDA engine: one `Engine` = one cgnn_ctx = one GPU.

This is the batched replacement of the reference's per-instance scorers
(`tf_run_instance`, Code/cgnn/GNN.py:135-166; `run_CGNN_tf`, CGNN.py:163-195;
`run_CGNN_confounders_tf`, CGNN_confounders.py:182-216).  It only marshals buffers
across the C ABI (include/cgnn_b200.h); all arithmetic happens in the CUDA kernels.
There is no CPU path: constructing an Engine without the built library or without an
sm_100 GPU raises.
"""
import ctypes as C

import numpy as np

from . import _lib
from .utils.Settings import SETTINGS


class EngineError(RuntimeError):
    pass


def _ptr(a):
    """Raw address of a numpy array (host) or torch tensor (host or device); None -> NULL."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    return C.c_void_p(a.data_ptr())          # torch.Tensor


def make_hparams(**kw):
    """cgnn_hparams from SETTINGS with the reference's kwargs-override rule
    (GNN.py:45-48,90,112; CGNN.py:48-51,115,137): the same names, per call."""
    hp = _lib.Hparams()
    hp.h_layer_dim = int(kw.get("h_layer_dim", SETTINGS.h_layer_dim))
    hp.train_epochs = int(kw.get("train_epochs", SETTINGS.train_epochs))
    hp.test_epochs = int(kw.get("test_epochs", SETTINGS.test_epochs))
    hp.learning_rate = float(kw.get("learning_rate", SETTINGS.learning_rate))
    hp.init_std = float(kw.get("init_std", SETTINGS.init_weights))
    hp.exp_policy = int(kw.get("exp_policy", SETTINGS.exp_policy))
    hp.trace_every = int(kw.get("trace_every", 0))
    fast = kw.get("use_Fast_MMD", SETTINGS.use_Fast_MMD)            # GNN.py:47-48, CGNN.py:50-51
    hp.fast_mmd_vectors = int(kw.get("nb_vectors_approx_MMD", SETTINGS.nb_vectors_approx_MMD)) if fast else 0
    hp.seed = int(kw.get("seed", SETTINGS.seed)) & 0xFFFFFFFFFFFFFFFF
    hp.graph_cluster = int(kw.get("graph_cluster", 0))
    hp.graph_noise_regen = int(kw.get("graph_noise_regen", 0))
    return hp


class Engine:
    def __init__(self, device=0):
        self._lib = _lib.load()
        h = C.c_void_p()
        rc = self._lib.cgnn_b200_create(C.byref(h), int(device))
        if rc != 0:
            raise EngineError("cgnn_b200_create(device=%d) failed with status %d "
                              "(no sm_100 GPU? this engine has no CPU fallback)" % (device, rc))
        self._h = h
        self.device = int(device)

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.cgnn_b200_destroy(h)
            except Exception:          # interpreter shutdown: the library may already be unloaded
                pass

    def __del__(self):
        self.close()

    def _check(self, rc, what):
        if rc != 0:
            msg = self._lib.cgnn_b200_last_error(self._h)
            raise EngineError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else ""))

    # -- info -------------------------------------------------------------------------------
    @property
    def last_kernel_ms(self):
        return float(self._lib.cgnn_b200_last_kernel_ms(self._h))

    @property
    def last_launch_count(self):
        return int(self._lib.cgnn_b200_last_launch_count(self._h))

    @property
    def last_graph_cluster(self):
        return int(self._lib.cgnn_b200_last_graph_cluster(self._h))

    @property
    def last_graph_mode(self):
        """Shared-memory plan of the last graph call: (observed matrix: 0 smem / 1 ring / 2 global) + 4 if the noise
        was regenerated at every use instead of buffered."""
        return int(self._lib.cgnn_b200_last_graph_mode(self._h))

    def device_info(self):
        a, b = C.c_int(), C.c_int()
        self._check(self._lib.cgnn_b200_device_info(self._h, C.byref(a), C.byref(b)), "device_info")
        return dict(n_sm=a.value, ctas_per_sm_pairwise=b.value)

    def set_stream(self, cuda_stream_ptr):
        self._check(self._lib.cgnn_b200_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "set_stream")

    def probe_peak(self, kind, reps=5):
        """kind 'mufu' -> exp/s, 'ffma' / 'ffma2' -> flop/s of scalar / packed (FP32x2) FMA chains (micro-benchmarks)."""
        out = C.c_double()
        k = {"mufu": 0, "ffma": 1, "ffma2": 2, "fadd2_slots": 3, "graph_mix_slots": 4}[kind]
        self._check(self._lib.cgnn_b200_probe_peak(self._h, k, reps, C.byref(out)), "probe_peak")
        return out.value

    # -- pairwise ---------------------------------------------------------------------------
    def run_pairwise(self, xy, n_rows=None, n_runs=None, uid_base=0, noise=None, init_params=None,
                     want_trace=False, want_params=False, scores_out=None, **kw):
        """Score every (pair, run, direction) instance.

        xy: [n_pairs, ld_rows, 2] float32, numpy (host) or torch (host/cuda) - column 0 = A.
        Returns dict(scores [n_pairs, n_runs, 2], trace?, params?) as numpy arrays, or writes the
        scores into `scores_out` (e.g. a cuda tensor) when given.
        """
        hp = make_hparams(**kw)
        n_runs = int(n_runs if n_runs is not None else kw.get("nb_runs", SETTINGS.NB_RUNS))
        if isinstance(xy, np.ndarray):
            xy = np.ascontiguousarray(xy, dtype=np.float32)
        n_pairs, ld_rows = int(xy.shape[0]), int(xy.shape[1])
        if tuple(xy.shape[2:]) != (2,):
            raise ValueError("xy must be [n_pairs, rows, 2]")
        if n_rows is None:
            n_rows = np.full(n_pairs, ld_rows, np.int32)
        n_rows = np.ascontiguousarray(n_rows, dtype=np.int32)
        n_inst = n_pairs * n_runs * 2
        P = 4 * hp.h_layer_dim + 1
        epochs = hp.train_epochs + hp.test_epochs
        if noise is not None and isinstance(noise, np.ndarray):
            noise = np.ascontiguousarray(noise, np.float32)
            if noise.size != n_inst * epochs * ld_rows:
                raise ValueError("noise must be [n_inst, train+test epochs, ld_rows]")
        if init_params is not None and isinstance(init_params, np.ndarray):
            init_params = np.ascontiguousarray(init_params, np.float32)
            if init_params.size != n_inst * P:
                raise ValueError("init_params must be [n_inst, 4h+1]")
        scores = scores_out if scores_out is not None else np.empty((n_pairs, n_runs, 2), np.float32)
        trace = np.empty((n_inst, epochs), np.float32) if want_trace else None
        pout = np.empty((n_inst, P), np.float32) if want_params else None
        rc = self._lib.cgnn_b200_run_pairwise(self._h, C.byref(hp), _ptr(xy), _ptr(n_rows), n_pairs, ld_rows,
                                              n_runs, C.c_ulonglong(int(uid_base)), _ptr(noise), _ptr(init_params),
                                              _ptr(scores), _ptr(trace), _ptr(pout))
        self._check(rc, "cgnn_b200_run_pairwise")
        return dict(scores=scores, trace=trace, params=pout)

    def pairwise_step_debug(self, xy, noise, init_params, fourier_wz=None, n_steps=None, **kw):
        """Parity probe: consecutive train steps of ONE instance (column 0 = cause).
        fourier_wz [n_steps, 3, 7k]: with use_Fast_MMD=True, the random matrix of Loss.py:35-37 per step."""
        hp = make_hparams(**kw)
        if fourier_wz is not None:
            fourier_wz = np.ascontiguousarray(fourier_wz, np.float32)
            assert fourier_wz.shape[1:] == (3, 7 * hp.fast_mmd_vectors)
        xy = np.ascontiguousarray(xy, np.float32)
        n_rows = xy.shape[0]
        P = 4 * hp.h_layer_dim + 1
        if noise is not None:                    # None: the kernel's own Philox noise (stream uid 0)
            noise = np.ascontiguousarray(noise, np.float32)
            n_steps = noise.shape[0]
            assert noise.shape == (n_steps, n_rows)
        if init_params is not None:
            init_params = np.ascontiguousarray(init_params, np.float32)
            assert init_params.shape == (P,)
        out = dict(loss=np.empty(n_steps, np.float32), yhat=np.empty((n_steps, n_rows), np.float32),
                   dldy=np.empty((n_steps, n_rows), np.float32), grad=np.empty((n_steps, P), np.float32),
                   params=np.empty((n_steps, P), np.float32))
        rc = self._lib.cgnn_b200_pairwise_step_debug(self._h, C.byref(hp), _ptr(xy), n_rows, n_steps, _ptr(noise),
                                                     _ptr(init_params), _ptr(fourier_wz), _ptr(out["loss"]), _ptr(out["yhat"]),
                                                     _ptr(out["dldy"]), _ptr(out["grad"]), _ptr(out["params"]))
        self._check(rc, "cgnn_b200_pairwise_step_debug")
        return out

    # -- graph ------------------------------------------------------------------------------
    def run_graph(self, data, plans, n_runs=None, row_subsets=None, uid_base=0, graph_uids=None, noise=None,
                  init_params=None, want_trace=False, want_params=False, **kw):
        """Score every (candidate graph, run) instance.

        data: [n_rows_total, n_var] float32 (numpy or torch), columns as the plans' `col` indices.
        plans: list of plans (cgnn_b200.plan.build_plan).  row_subsets: None or int32
        [n_graphs*n_runs, n_rows] (the per-call subsample of CGNN.py:183-185).
        Returns dict(scores [n_graphs, n_runs], trace?, params?).
        """
        from .plan import GraphBatch, n_params
        hp = make_hparams(**kw)
        n_runs = int(n_runs if n_runs is not None else kw.get("nb_runs", SETTINGS.NB_RUNS))
        if isinstance(data, np.ndarray):
            data = np.ascontiguousarray(data, dtype=np.float32)
        n_total = int(data.shape[0])
        gb = GraphBatch(plans)
        n_graphs = len(plans)
        n_inst = n_graphs * n_runs
        if row_subsets is not None:
            row_subsets = np.ascontiguousarray(row_subsets, np.int32).reshape(n_inst, -1)
            n_rows = row_subsets.shape[1]
        else:
            n_rows = n_total
        ld = max(n_params(p, hp.h_layer_dim) for p in plans)
        epochs = hp.train_epochs + hp.test_epochs
        if noise is not None:
            noise = np.ascontiguousarray(noise, np.float32)
            if noise.size != n_inst * epochs * gb.struct.n_noise_cols * n_rows:
                raise ValueError("noise must be [n_inst, epochs, n_noise_cols, n_rows]")
        if init_params is not None:
            init_params = np.ascontiguousarray(init_params, np.float32)
            if init_params.size != n_inst * ld:
                raise ValueError("init_params must be [n_inst, %d]" % ld)
        guids = None if graph_uids is None else np.ascontiguousarray(graph_uids, np.uint64)
        if guids is not None and guids.size != n_graphs:
            raise ValueError("graph_uids must have one entry per graph")
        scores = np.empty((n_graphs, n_runs), np.float32)
        trace = np.empty((n_inst, epochs), np.float32) if want_trace else None
        pout = np.empty((n_inst, ld), np.float32) if want_params else None
        rc = self._lib.cgnn_b200_run_graph(self._h, C.byref(hp), _ptr(data), n_total, C.byref(gb.struct), n_runs,
                                           _ptr(row_subsets), n_rows, C.c_ulonglong(int(uid_base)), _ptr(guids), _ptr(noise),
                                           _ptr(init_params), ld, _ptr(scores), _ptr(trace), _ptr(pout))
        self._check(rc, "cgnn_b200_run_graph")
        return dict(scores=scores, trace=trace, params=pout)

    def graph_step_debug(self, data, plan, noise, init_params, fourier_wz=None, **kw):
        """Parity probe: consecutive train steps of ONE graph instance.
        fourier_wz [n_steps, n_var+1, 7k]: with use_Fast_MMD=True, the random matrix of Loss.py:35-37 per step."""
        from .plan import GraphBatch, n_params
        hp = make_hparams(**kw)
        data = np.ascontiguousarray(data, np.float32)
        noise = np.ascontiguousarray(noise, np.float32)
        init_params = np.ascontiguousarray(init_params, np.float32)
        gb = GraphBatch([plan])
        n_rows, n_var = data.shape
        n_steps = noise.shape[0]
        P = n_params(plan, hp.h_layer_dim)
        assert noise.shape == (n_steps, plan["n_noise"], n_rows) and init_params.shape == (P,)
        out = dict(loss=np.empty(n_steps, np.float32), pred=np.empty((n_steps, n_rows, n_var), np.float32),
                   dldp=np.empty((n_steps, n_rows, n_var), np.float32), grad=np.empty((n_steps, P), np.float32),
                   params=np.empty((n_steps, P), np.float32))
        if fourier_wz is not None:
            fourier_wz = np.ascontiguousarray(fourier_wz, np.float32)
            assert fourier_wz.shape == (n_steps, n_var + 1, 7 * hp.fast_mmd_vectors)
        rc = self._lib.cgnn_b200_graph_step_debug(self._h, C.byref(hp), _ptr(data), n_rows, C.byref(gb.struct), n_steps,
                                                  _ptr(noise), _ptr(init_params), P, _ptr(fourier_wz), _ptr(out["loss"]), _ptr(out["pred"]),
                                                  _ptr(out["dldp"]), _ptr(out["grad"]), _ptr(out["params"]))
        self._check(rc, "cgnn_b200_graph_step_debug")
        return out

    def graph_generate(self, plan, params, noise, **kw):
        """One sample of the generative model (CGNN_tf.generate, CGNN.py:155-160): the generated matrix
        [N, n_var] for given parameters (flat, caller-visible layout) and noise [n_noise, N]."""
        noise = np.ascontiguousarray(noise, np.float32)
        n_rows = noise.shape[1]
        dummy = np.zeros((n_rows, plan["n_var"]), np.float32)
        out = self.graph_step_debug(dummy, plan, noise[None], np.asarray(params, np.float32), **kw)
        return out["pred"][0]

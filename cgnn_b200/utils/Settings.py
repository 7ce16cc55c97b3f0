"""Global configuration singleton, API-compatible with the reference's `cgnn.SETTINGS`
(Code/cgnn/utils/Settings.py:9-45): the same 14 knob names with the same defaults, read at call
time and overridable per call through **kwargs of the same (lower-case) names, as the reference
does (GNN.py:45-48; CGNN.py:48-51,208-209).

Engine-only additions: `seed` (the reference is unseeded), `exp_policy` (how the CUDA kernel
produces the 7 Gaussian kernel values, see include/cgnn_b200.h; -1 = the fastest per path), `verbose` and `work_stealing`
(chunked dynamic schedule over GPUs / ranks, cgnn_b200/runtime.py).
`NB_JOBS` is accepted and ignored (no process pool: one launch scores the whole batch);
`NB_GPU`/`GPU_OFFSET` select which GPUs a single process drives.
"""

_REFERENCE_DEFAULTS = (
    # ensemble / placement                      Settings.py:27-34
    ("NB_RUNS", 32), ("NB_JOBS", 1), ("GPU", True), ("NB_GPU", 1), ("GPU_OFFSET", 0),
    ("learning_rate", 0.01), ("init_weights", 0.05), ("max_nb_points", 1500),
    # generators                                Settings.py:36-41
    ("h_layer_dim", 20), ("train_epochs", 1000), ("test_epochs", 500),
    ("use_Fast_MMD", False), ("nb_vectors_approx_MMD", 100), ("complexity_graph_param", 0.00005),
)
_ENGINE_DEFAULTS = (("seed", 0), ("exp_policy", -1), ("verbose", True), ("work_stealing", False))


class EngineSettings(object):
    __slots__ = tuple(k for k, _ in _REFERENCE_DEFAULTS + _ENGINE_DEFAULTS)

    def __init__(self):
        self.reset()

    def reset(self):
        for k, v in _REFERENCE_DEFAULTS + _ENGINE_DEFAULTS:
            setattr(self, k, v)

    def as_dict(self):
        return {k: getattr(self, k) for k in self.__slots__}


SETTINGS = EngineSettings()

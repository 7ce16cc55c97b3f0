"""Oracle: TensorFlow-1.x Adam (TEST INFRASTRUCTURE ONLY).

The reference calls ``tf.train.AdamOptimizer(learning_rate).minimize`` with
default betas/epsilon (Code/cgnn/GNN.py:74-75, CGNN.py:97-99,
CGNN_confounders.py:116-118).  TensorFlow is a third-party, un-vendored and
unpinned dependency (Code/requirements.txt:4); its published TF-1 update is

    t <- t+1;  lr_t = lr * sqrt(1-b2^t) / (1-b1^t)
    m <- b1 m + (1-b1) g;  v <- b2 v + (1-b2) g^2
    theta <- theta - lr_t * m / (sqrt(v) + eps)        ("epsilon hat" form)

with b1=0.9, b2=0.999, eps=1e-8.  (PyTorch's Adam puts eps inside the
bias-corrected denominator and is NOT equivalent.)
"""
import numpy as np

BETA1, BETA2, EPS = 0.9, 0.999, 1e-8


class TFAdam:
    def __init__(self, n_params: int, lr: float = 0.01, dtype=np.float64):
        self.lr = lr
        self.m = np.zeros(n_params, dtype)
        self.v = np.zeros(n_params, dtype)
        self.b1p = 1.0   # beta1^t, beta2^t accumulators, as TF keeps them
        self.b2p = 1.0

    def step(self, theta: np.ndarray, grad: np.ndarray) -> np.ndarray:
        self.b1p *= BETA1
        self.b2p *= BETA2
        lr_t = self.lr * np.sqrt(1.0 - self.b2p) / (1.0 - self.b1p)
        self.m = BETA1 * self.m + (1.0 - BETA1) * grad
        self.v = BETA2 * self.v + (1.0 - BETA2) * grad * grad
        return theta - lr_t * self.m / (np.sqrt(self.v) + EPS)

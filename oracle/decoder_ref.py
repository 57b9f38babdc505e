"""TEST INFRASTRUCTURE ONLY -- CPU restatement of InnerProductDecoder's edge-wise decode
(reference STEM-GNN/model/encoder.py:353-366) and of its gradient, in plain torch / numpy.
Pinned against the reference's own class in tests/golden/dec_*.npz (tests/test_oracle_golden.py).
"""
import numpy as np
import torch


def edge_dot_ref(z: torch.Tensor, edge_index: torch.Tensor, sigmoid: bool = True) -> torch.Tensor:
    """encoder.py:364-366: value = (z[edge_index[0]] * z[edge_index[1]]).sum(dim=1); sigmoid."""
    value = (z[edge_index[0]] * z[edge_index[1]]).sum(dim=1)
    return torch.sigmoid(value) if sigmoid else value


def edge_dot_grad_ref(z: np.ndarray, edge_index: np.ndarray, g_out: np.ndarray, sigmoid: bool) -> np.ndarray:
    """Closed-form gradient in float64, edge by edge in edge order:
    g_z[src] += w_e z[dst], g_z[dst] += w_e z[src], w_e = g_e (* s(1-s) after the sigmoid)."""
    z64 = z.astype(np.float64)
    src, dst = edge_index[0], edge_index[1]
    v = (z64[src] * z64[dst]).sum(1)
    w = g_out.astype(np.float64)
    if sigmoid:
        s = 1.0 / (1.0 + np.exp(-v))
        w = w * s * (1.0 - s)
    g = np.zeros_like(z64)
    for e in range(src.shape[0]):
        g[src[e]] += w[e] * z64[dst[e]]
        g[dst[e]] += w[e] * z64[src[e]]
    return g

"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/c/agg_ref.c (built by `make -C oracle/c`
or `__graft_entry__.build()` into oracle/_build/liboracle_agg.so)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_agg.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "c")])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _prep(edge_index, x, xe, w):
    ei = np.ascontiguousarray(np.asarray(edge_index, dtype=np.int64))
    x = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
    xe = None if xe is None else np.ascontiguousarray(np.asarray(xe, dtype=np.float32))
    w = None if w is None else np.ascontiguousarray(np.asarray(w, dtype=np.float32))
    return ei, x, xe, w


def aggregate_fwd(edge_index, x, num_nodes, xe=None, w=None, act=1, reduce=1):
    """out[dst] = reduce_e act(x[src] + xe[e]) * w[e]; sequential in edge order."""
    ei, x, xe, w = _prep(edge_index, x, xe, w)
    E, F = ei.shape[1], x.shape[1]
    out = np.empty((num_nodes, F), dtype=np.float32)
    src, dst = np.ascontiguousarray(ei[0]), np.ascontiguousarray(ei[1])
    lib().oracle_aggregate_fwd(_p(src), _p(dst), ctypes.c_int64(E), ctypes.c_int64(num_nodes),
                               ctypes.c_int64(F), _p(x), _p(xe), _p(w), act, reduce, _p(out))
    return out


def aggregate_bwd(edge_index, x, g_out, xe=None, w=None, act=1, reduce=1):
    ei, x, xe, w = _prep(edge_index, x, xe, w)
    g_out = np.ascontiguousarray(np.asarray(g_out, dtype=np.float32))
    E, F, N = ei.shape[1], x.shape[1], g_out.shape[0]
    assert x.shape[0] == N, "square aggregation only (source set == destination set)"
    g_x = np.empty((N, F), dtype=np.float32)
    src, dst = np.ascontiguousarray(ei[0]), np.ascontiguousarray(ei[1])
    lib().oracle_aggregate_bwd(_p(src), _p(dst), ctypes.c_int64(E), ctypes.c_int64(N),
                               ctypes.c_int64(F), _p(x), _p(xe), _p(w), act, reduce,
                               _p(g_out), _p(g_x))
    return g_x

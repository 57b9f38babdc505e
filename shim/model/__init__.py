"""Import shim: put `<repo>/shim` (and `<repo>`) in front of the reference checkout on sys.path and the
reference's `from model.encoder import Encoder` / `from model.vq import VectorQuantize`
(STEM-GNN/pretrain.py:16-18, finetune.py:14-16, model/pt_model.py, model/ft_model.py) resolve to the
B200 modules; everything else under the reference's own `model/` package (pt_model, ft_model, ...) is
still found there, because this package extends its search path with the other `model/` directories
on sys.path."""
import os as _os
import sys as _sys

_here = _os.path.dirname(_os.path.abspath(__file__))
for _p in list(_sys.path):
    _cand = _os.path.join(_p or ".", "model")
    if _os.path.isdir(_cand) and _os.path.abspath(_cand) != _here and _cand not in __path__:
        __path__.append(_cand)          # after ours: encoder.py / vq.py come from this directory

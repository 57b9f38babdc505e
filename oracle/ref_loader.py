"""TEST INFRASTRUCTURE ONLY -- imports the reference's own `model` package (vq.py, encoder.py,
pt_model.py, ft_model.py; /root/reference/STEM-GNN/model) either from the checkout, when this machine
has one, or from the archive oracle/make_ref.py packed (oracle/_ref/, git-ignored, travels to the GPU
box).  encoder.py / pt_model.py / ft_model.py import torch_geometric and torch_scatter, which are not
installable here: unless the real packages are importable, oracle/pyg_shim stands in for the handful
of symbols they use (see oracle/__init__.py: that arithmetic is "parity unpinned by the reference").

The modules are loaded under private names, so the `model` package of the import shim (shim/model,
which serves the B200 drop-ins under the reference's import names) is left alone."""
import importlib
import os
import sys
import types
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
CHECKOUT = "/root/reference/STEM-GNN"
ARCHIVE = os.path.join(HERE, "_ref", "stemgnn_reference_model.zip")
PYG_SHIM = os.path.join(HERE, "pyg_shim")
_cache = {}


def source():
    """('checkout' | 'archive' | None, sys.path entry)"""
    if os.path.isdir(os.path.join(CHECKOUT, "model")):
        return "checkout", CHECKOUT
    if os.path.exists(ARCHIVE):
        return "archive", ARCHIVE
    return None, None


def pyg_kind():
    """'torch_geometric <version>' when the real package is importable, else 'oracle/pyg_shim'."""
    if PYG_SHIM in sys.path:
        return "oracle/pyg_shim"
    try:
        import torch_geometric
        if "oracle_shim" not in getattr(torch_geometric, "__version__", ""):
            return f"torch_geometric {torch_geometric.__version__}"
    except ImportError:
        pass
    return "oracle/pyg_shim"


def _ensure_pyg():
    try:
        import torch_geometric  # noqa: F401
        import torch_scatter  # noqa: F401
    except ImportError:
        if PYG_SHIM not in sys.path:
            sys.path.insert(0, PYG_SHIM)


def load(names=("vq", "encoder", "pt_model", "ft_model")):
    """SimpleNamespace with the requested reference modules (ns.vq.VectorQuantize, ns.encoder.Encoder,
    ns.pt_model.PretrainModel, ns.ft_model.TaskModel), plus ns.kind / ns.path / ns.pyg.  Raises
    FileNotFoundError when neither the checkout nor the archive is present."""
    kind, path = source()
    if kind is None:
        raise FileNotFoundError("the reference's model/ package is neither at /root/reference nor packed in "
                                "oracle/_ref (run `python oracle/make_ref.py` where the checkout exists)")
    if any(n != "vq" for n in names):
        _ensure_pyg()
    held = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "model" or k.startswith("model.")}
    sys.path.insert(0, path)
    importlib.invalidate_caches()
    out = types.SimpleNamespace(kind=kind, path=path, pyg=None)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")           # torch.cuda.amp.autocast FutureWarning (vq.py:438,623)
            for n in names:
                key = (path, n)
                if key not in _cache:
                    _cache[key] = importlib.import_module("model." + n)
                setattr(out, n, _cache[key])
    finally:
        sys.path.remove(path)
        for k in list(sys.modules):
            if k == "model" or k.startswith("model."):
                del sys.modules[k]
        sys.modules.update(held)
    out.pyg = pyg_kind() if any(n != "vq" for n in names) else None
    return out

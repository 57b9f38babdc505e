"""Importable name of the package whose sources live in `stem-gnn_b200/` (a directory name
Python cannot import directly).  All modules are resolved from that directory."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                 "stem-gnn_b200"))

from ._api import *  # noqa: E402,F401,F403
from ._api import __all__  # noqa: E402,F401

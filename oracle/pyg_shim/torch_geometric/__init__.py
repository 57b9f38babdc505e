"""TEST INFRASTRUCTURE ONLY -- stand-in for torch_geometric 2.3.0 (see nn/__init__.py)."""
__version__ = "2.3.0+oracle_shim"

"""TEST INFRASTRUCTURE ONLY."""


def spmm(*a, **k):
    raise NotImplementedError("pyg_shim: spmm path (encoder.py:99-102) needs a SparseTensor; unused")

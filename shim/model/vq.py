"""`model.vq` of the reference (STEM-GNN/model/vq.py) served by stemgnn_b200."""
from stemgnn_b200 import CosineSimCodebook, EuclideanCodebook, VectorQuantize  # noqa: F401

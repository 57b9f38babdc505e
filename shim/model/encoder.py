"""`model.encoder` of the reference (STEM-GNN/model/encoder.py) served by stemgnn_b200."""
from stemgnn_b200 import (Encoder, GCNConv, GINConv, InnerProductDecoder, MixtureSageLayer,  # noqa: F401
                          MySAGEConv)

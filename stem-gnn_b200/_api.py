"""Public surface: the drop-in modules plus the functional operators underneath them."""
from . import _lib
from ._lib import StemGnnError
from .conv import GCNConv, GINConv, MixtureSageLayer, MySAGEConv
from .encoder import Encoder, InnerProductDecoder
from .graph import CSR, GraphStructure, clear_structure_cache, structure_of
from .partition import PartitionedGraph, block_size, local_edges, partition_1d
from .ops import (aggregate, l2norm_rows, vq_assign_raw, vq_code_stats, vq_ema_apply, vq_quantize,
                  METRIC_DOT, METRIC_EUCLID, PATH_AUTO, PATH_EXACT, PATH_TENSOR)
from .dense import linear_raw, linear_sum, linear_wgrad_raw
from .decoder import edge_dot
from .augment import dropout_adj, edge_concat_linear, gather_rows, mask_feature, negative_sampling
from .sampler import NeighborLoader, NeighborSampler, neighbor_sample
from .norm_act import bn_act_drop, batch_stats
from . import norm_act
from .vq import CosineSimCodebook, EuclideanCodebook, VectorQuantize

__all__ = ["StemGnnError", "GCNConv", "GINConv", "MixtureSageLayer", "MySAGEConv", "Encoder",
           "InnerProductDecoder", "CSR", "GraphStructure", "clear_structure_cache", "structure_of",
           "PartitionedGraph", "block_size", "local_edges", "partition_1d",
           "aggregate", "l2norm_rows", "edge_dot", "bn_act_drop", "batch_stats", "norm_act", "linear_raw", "linear_sum", "linear_wgrad_raw", "vq_assign_raw", "vq_code_stats", "vq_ema_apply", "vq_quantize",
           "METRIC_DOT", "METRIC_EUCLID", "PATH_AUTO", "PATH_EXACT", "PATH_TENSOR",
           "CosineSimCodebook", "EuclideanCodebook", "VectorQuantize",
           "dropout_adj", "gather_rows", "mask_feature", "negative_sampling", "edge_concat_linear", "NeighborLoader", "NeighborSampler", "neighbor_sample"]

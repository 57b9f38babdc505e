"""1-D partition of a graph by destination node across the GPUs of one box.

The reference is single-GPU (pretrain.py:84); north_star asks for the propagate step to scale by
splitting the nodes into `world` contiguous, equal-sized blocks: rank r owns the rows
[r*per, min(N, (r+1)*per)), per = ceil(N / world), of every node tensor (x, z, gradients, VQ rows).

* forward  (aggregate AT edge_index[1]): rank r keeps the edges whose destination it owns, in their
  original order, and needs the features of ALL sources -> one all-gather of the row blocks per
  layer.  Because the blocks are equal-sized and contiguous, a node's row in the gathered matrix is
  its global id: no index remap.
* backward (autograd of the above): g_x of an owned source sums over the edges whose SOURCE the
  rank owns, gathering g_out / count of all destinations; each rank divides its own g_out rows by
  its own counts first (K2a), so one all-gather of the scaled g_out blocks is all that moves.

A row's edges all live on one rank and keep their relative order, so the sums are the same fp32
sums, in the same order, as on one GPU: partitioned results are bit-identical to single-GPU ones.
The integer bookkeeping below is plain torch and runs on any device (tests/test_partition_cpu.py
checks it on CPU with gloo); the kernels are the same C-ABI calls as the single-GPU path.
"""
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib
from .graph import CSR


def partition_1d(num_nodes: int, world: int) -> List[Tuple[int, int]]:
    """[(lo, hi)] per rank: contiguous blocks of ceil(N / world) nodes (the last ones may be short
    or empty)."""
    per = -(-int(num_nodes) // int(world)) if num_nodes > 0 else 0
    return [(min(num_nodes, r * per), min(num_nodes, (r + 1) * per)) for r in range(world)]


def block_size(num_nodes: int, world: int) -> int:
    return -(-int(num_nodes) // int(world)) if num_nodes > 0 else 0


def local_edges(edge_index: torch.Tensor, lo: int, hi: int, owner_row: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Edges whose endpoint `edge_index[owner_row]` lies in [lo, hi), in original order.  Returns
    (local edge list [2, E_loc] with the owner row shifted to local ids and the other row left
    global, original edge ids [E_loc])."""
    key = edge_index[owner_row]
    keep = (key >= lo) & (key < hi)
    eid = keep.nonzero(as_tuple=False).view(-1)
    loc = edge_index[:, eid].clone()
    loc[owner_row] -= lo
    return loc.contiguous(), eid


class PartitionedGraph:
    """This rank's share of a 1-D destination-partitioned graph.  Accepted wherever the drop-in
    layers take `edge_index` (like GraphStructure).

    edge_index: the GLOBAL [2, E] int64 edge list (every rank passes the same tensor), or, with
    `pre_split=(fwd_edges, bwd_edges)`, the two local lists produced by `local_edges` elsewhere.
    `gather` maps this rank's padded row block [per, F] to the gathered [world*per, F] matrix;
    the default is `dist.all_gather_into_tensor` on `group`.
    """

    def __init__(self, edge_index: Optional[torch.Tensor], num_nodes: int, rank: Optional[int] = None,
                 world: Optional[int] = None, group=None,
                 gather: Optional[Callable[[torch.Tensor], torch.Tensor]] = None, pre_split=None):
        if world is None:
            world = dist.get_world_size(group) if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.num_nodes_global, self.rank, self.world, self.group = int(num_nodes), int(rank), int(world), group
        self.per = block_size(num_nodes, world)
        self.lo, self.hi = partition_1d(num_nodes, world)[rank]
        self.num_nodes = self.hi - self.lo                      # rows this rank owns
        if pre_split is None:
            fwd, _ = local_edges(edge_index, self.lo, self.hi, 1)
            bwd, _ = local_edges(edge_index, self.lo, self.hi, 0)
        else:
            fwd, bwd = pre_split
        self.fwd_edges, self.bwd_edges = fwd, bwd
        self._gather = gather
        self._csr = {}

    # ---- structure (K0 on the local edge lists) ------------------------------------------------
    def fwd_csr(self) -> CSR:
        """rows = owned destinations (local ids), col = global source ids."""
        if "f" not in self._csr:
            self._csr["f"] = CSR(self.fwd_edges, self.num_nodes, 1)
        return self._csr["f"]

    def bwd_csr(self) -> CSR:
        """rows = owned sources (local ids), col = global destination ids."""
        if "b" not in self._csr:
            self._csr["b"] = CSR(self.bwd_edges, self.num_nodes, 0)
        return self._csr["b"]

    def gather_rows(self, local: torch.Tensor) -> torch.Tensor:
        """[n_local, F] -> [world*per, F]: every rank's block, padded to `per` rows, in rank order
        (so row g of the result is global node g).  This is the halo exchange."""
        n, f = local.shape
        if self.world == 1 and self._gather is None:
            return local
        blk = local
        if n != self.per:
            blk = local.new_zeros((self.per, f))
            blk[:n] = local
        if self._gather is not None:
            return self._gather(blk)
        out = local.new_empty((self.world * self.per, f))
        with _lib.traced("halo_all_gather", (self.world - 1) * self.per * f * local.element_size(), 0, 0):
            dist.all_gather_into_tensor(out, blk.contiguous(), group=self.group)
        return out

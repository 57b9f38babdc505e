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
import os
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


def block_edges(edges: torch.Tensor, per: int, world: int, other_row: int) -> List[Tuple[torch.Tensor, torch.Tensor]]:
    """Cut a rank's local edge list by the OWNER BLOCK of the other endpoint (row `other_row`, global ids):
    [(edges of block b with that row shifted to ids inside the block, positions in `edges`)] for b in
    range(world).  Inside a block the edges keep their relative order."""
    out = []
    owner = torch.div(edges[other_row], max(per, 1), rounding_mode="floor")
    for b in range(world):
        pos = (owner == b).nonzero(as_tuple=False).view(-1)
        sub = edges[:, pos].clone()
        sub[other_row] -= b * per
        out.append((sub.contiguous(), pos))
    return out


def chunk_edges(edges: torch.Tensor, per: int, world: int, chunks: int, other_row: int):
    """Cut a rank's local edge list by ROW CHUNK of the other endpoint inside its owner's block: chunk c holds the
    edges whose gathered row (global id g, owner b = g // per, local row l = g - b * per) has l // cr == c, with
    cr = ceil(per / chunks); the endpoint is renumbered to b * cr + (l - c * cr), its row in the all-gather of the
    ranks' c-th row chunks.  Returns [(edges of chunk c, positions in `edges`)]; order inside a chunk is kept."""
    cr = -(-max(per, 1) // chunks)
    g = edges[other_row]
    b = torch.div(g, max(per, 1), rounding_mode="floor")
    loc = g - b * per
    c_of = torch.div(loc, cr, rounding_mode="floor")
    out = []
    for c in range(chunks):
        pos = (c_of == c).nonzero(as_tuple=False).view(-1)
        sub = edges[:, pos].clone()
        sub[other_row] = b[pos] * cr + (loc[pos] - c * cr)
        out.append((sub.contiguous(), pos))
    return out


def exchange_order(rank: int, world: int, transport: str) -> List[int]:
    """Owner blocks in the order a rank consumes them in the overlapped exchange: its own block first (no wire),
    then in order of arrival -- round k of the pairwise exchange brings the block of rank - k; the broadcasts
    arrive in rank order."""
    if transport == "p2p":
        return [(rank - k) % world for k in range(world)]
    return [rank] + [b for b in range(world) if b != rank]


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
                 gather: Optional[Callable[[torch.Tensor], torch.Tensor]] = None, pre_split=None,
                 halo_slices: Optional[int] = None, halo_dtype: Optional[torch.dtype] = None,
                 stat_gather=None, stat_reduce=None, halo_mode: str = "gather", halo_transport: str = "p2p",
                 blocks: Optional[Callable[[torch.Tensor], List[torch.Tensor]]] = None, halo_chunks: int = 4):
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
        self._stat_gather, self._stat_reduce = stat_gather, stat_reduce     # BatchNorm collectives (tests inject them)
        self._csr = {}
        # halo pipeline: number of feature slices per exchange (None: by width) and the wire dtype
        # (None = fp32, bit-identical to one GPU; torch.bfloat16 halves the bytes at a stated 2^-9
        # relative error per gathered feature)
        self.halo_slices, self.halo_dtype = halo_slices, halo_dtype
        if halo_dtype not in (None, torch.float32, torch.bfloat16):
            raise ValueError("halo_dtype must be None, torch.float32 or torch.bfloat16")
        # halo_mode "gather": one blocking all-gather per exchange, then ONE aggregation over the rank's CSR --
        #   every row's edges in original order, bit-identical to one GPU (the default, and what parity tests use).
        # halo_mode "blocks": the overlapped exchange.  The rank's CSR is cut by source-owner block (K0, once per
        #   graph); the row blocks travel one by one (pairwise send/recv rounds or broadcasts, all issued up front
        #   on the collective's own stream) and block b is aggregated into `out` (stemgnn_aggregate, accumulate = 1)
        #   as soon as it has landed, the rank's own block first -- while the later blocks are still on the wire.
        #   A row's sum then runs block by block instead of in edge order: same terms, different order,
        #   |difference| <= a few fp32 roundings of the row's partial sums (tests hold 1e-5 of the output scale).
        # halo_mode "chunks": the all-gather itself pipelined over ROW chunks.  Every rank's block is cut into
        #   `halo_chunks` row ranges; chunk c of all ranks is one all-gather (the collective NCCL is best at, all of
        #   them issued up front), and the edges whose gathered row lies in chunk c (a sub-CSR per chunk) are
        #   aggregated as soon as that all-gather has completed, while the next chunks are on the wire.  The row
        #   sums are cut C-fold instead of P-fold (at P = 8 the owner-block calls see 3 edges per row and re-read
        #   `out` 8 times: aggregation 5.6 -> 13.3 ms per step, more than the overlap wins).  Same 1e-5 bar.
        if halo_mode not in ("gather", "blocks", "chunks") or halo_transport not in ("p2p", "bcast"):
            raise ValueError("halo_mode is 'gather', 'blocks' or 'chunks'; halo_transport is 'p2p' or 'bcast'")
        self.halo_mode, self.halo_transport, self._blocks = halo_mode, halo_transport, blocks
        self.halo_chunks = max(1, int(halo_chunks))
        # SMs the dense kernels leave to NCCL while an exchange is in flight (sage.py); STEMGNN_SM_RESERVE for the A/B
        self.sm_reserve = int(os.environ.get("STEMGNN_SM_RESERVE", "32"))

    def batch_sync(self):
        """Collectives for full-batch BatchNorm over the partition (norm_act.BatchSync), or None when
        this process holds the whole graph."""
        if self.world == 1 and self._stat_gather is None:
            return None
        from .norm_act import BatchSync
        return BatchSync(self.num_nodes_global, self.world, self.group, self._stat_gather, self._stat_reduce)

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

    def block_csrs(self, backward: bool) -> List[CSR]:
        """The rank's CSR cut by owner block of the gathered endpoint: entry b holds the edges whose source
        (forward) / destination (backward) rank b owns, col = row inside b's block."""
        key = "bb" if backward else "fb"
        if key not in self._csr:
            edges, own_row = (self.bwd_edges, 0) if backward else (self.fwd_edges, 1)
            self._csr[key] = [CSR(sub, self.num_nodes, own_row)
                              for sub, _ in block_edges(edges, self.per, self.world, 1 - own_row)]
        return self._csr[key]

    @property
    def n_groups(self) -> int:
        """Aggregation calls one exchange is cut into (overlapped modes)."""
        return self.world if self.halo_mode == "blocks" else self.halo_chunks

    def group_csrs(self, backward: bool) -> List[CSR]:
        """Sub-CSRs of the overlapped modes: by owner block ("blocks") or by row chunk ("chunks")."""
        if self.halo_mode == "blocks":
            return self.block_csrs(backward)
        key = "bc" if backward else "fc"
        if key not in self._csr:
            edges, own_row = (self.bwd_edges, 0) if backward else (self.fwd_edges, 1)
            self._csr[key] = [CSR(sub, self.num_nodes, own_row)
                              for sub, _ in chunk_edges(edges, self.per, self.world, self.halo_chunks, 1 - own_row)]
        return self._csr[key]

    def exchange_groups(self, local: torch.Tensor):
        """Start the exchange and yield (group id, gathered rows of that group [rows, F] fp32) as they complete."""
        if self.halo_mode == "blocks":
            yield from self.exchange_blocks(local)
            return
        C = self.halo_chunks
        cr = -(-max(self.per, 1) // C)
        n, f = local.shape
        blk = local.new_zeros((C * cr, f))
        blk[:n] = local
        if self._blocks is not None:                       # tests: the other ranks' padded blocks are known
            got = self._blocks(self._padded_block(local))
            for c in range(C):
                rows = []
                for b in range(self.world):
                    src = blk if b == self.rank else torch.nn.functional.pad(got[b], (0, 0, 0, C * cr - got[b].size(0)))
                    rows.append(src[c * cr:(c + 1) * cr])
                yield c, torch.cat(rows, dim=0)
            return
        if self.world == 1:
            for c in range(C):
                yield c, blk[c * cr:(c + 1) * cr]
            return
        wire = blk if self.halo_dtype in (None, torch.float32) else blk.to(self.halo_dtype)
        pending = []
        for c in range(C):
            out = wire.new_empty((self.world * cr, f))
            work = dist.all_gather_into_tensor(out, wire[c * cr:(c + 1) * cr], group=self.group, async_op=True)
            pending.append((c, out, work))
        nbytes = (self.world - 1) * cr * f * wire.element_size()
        for c, out, work in pending:
            with _lib.traced("halo_wait", nbytes, 0, 0):
                work.wait()
            yield c, (out if out.dtype == torch.float32 else out.float())

    def exchange_blocks(self, local: torch.Tensor):
        """Start the exchange of this rank's row block and yield (owner b, block [per, F] fp32) in
        `exchange_order`: the own block at once, every other one after its transfer has completed (the wait is
        stream-ordered for NCCL: the host does not block)."""
        blk = self._padded_block(local)
        order = exchange_order(self.rank, self.world, self.halo_transport)
        if self._blocks is not None:                       # tests: the other ranks' blocks are known
            got = self._blocks(blk)
            for b in order:
                yield b, (blk if b == self.rank else got[b])
            return
        if self.world == 1:
            yield self.rank, blk
            return
        wire = blk if self.halo_dtype in (None, torch.float32) else blk.to(self.halo_dtype)
        wire = wire.contiguous()
        nbytes = wire.numel() * wire.element_size()
        ranks = dist.get_process_group_ranks(self.group) if self.group is not None else list(range(self.world))
        pending = []
        if self.halo_transport == "p2p":
            for k in range(1, self.world):
                src, dst = (self.rank - k) % self.world, (self.rank + k) % self.world
                buf = torch.empty_like(wire)
                ops = [dist.P2POp(dist.isend, wire, ranks[dst], self.group),
                       dist.P2POp(dist.irecv, buf, ranks[src], self.group)]
                pending.append((src, buf, dist.batch_isend_irecv(ops)))
        else:
            for b in range(self.world):
                buf = wire if b == self.rank else torch.empty_like(wire)
                work = dist.broadcast(buf, ranks[b], group=self.group, async_op=True)
                if b != self.rank:
                    pending.append((b, buf, [work]))
                else:
                    pending.append((None, None, [work]))
        yield self.rank, blk
        for b, buf, works in pending:
            with _lib.traced("halo_wait", nbytes if b is not None else 0, 0, 0):
                for w in works:
                    w.wait()
            if b is not None:
                yield b, (buf if buf.dtype == torch.float32 else buf.float())

    def feature_slices(self, f: int) -> List[Tuple[int, int]]:
        """Column ranges the halo exchange of an [n, f] matrix is pipelined over (16-byte aligned)."""
        if (self.world == 1 and self._gather is None) or f % 4:
            return [(0, f)]
        s = self.halo_slices
        if s is None:
            # measured on 2 B200s (F=256, 1.2M nodes): 4 slices of 64 columns 7.60 ms vs 7.54 ms
            # unsliced -- what the overlap hides, the narrower aggregation launches give back; slices
            # stay >= 256 columns wide by default
            s = max(1, f // 256)
        s = max(1, min(int(s), f // 4))
        w = -(-(f // 4) // s) * 4                       # slice width: multiple of 4 columns
        return [(lo, min(f, lo + w)) for lo in range(0, f, w)]

    def _padded_block(self, local: torch.Tensor) -> torch.Tensor:
        n, f = local.shape
        if n == self.per and local.is_contiguous():
            return local
        blk = local.new_zeros((self.per, f))
        blk[:n] = local
        return blk

    def gather_cols_async(self, local: torch.Tensor, lo: int, hi: int):
        """Start the exchange of columns [lo, hi) of this rank's block; returns a callable that waits
        for it and hands back the gathered fp32 [world*per, hi-lo] matrix."""
        blk = self._padded_block(local[:, lo:hi])
        if self._gather is not None:
            full = self._gather(blk)
            return lambda: full
        wire = blk if self.halo_dtype in (None, torch.float32) else blk.to(self.halo_dtype)
        out = wire.new_empty((self.world * self.per, hi - lo))
        nbytes = (self.world - 1) * self.per * (hi - lo) * wire.element_size()
        work = dist.all_gather_into_tensor(out, wire.contiguous(), group=self.group, async_op=True)

        def finish():
            with _lib.traced("halo_wait", nbytes, 0, 0):
                work.wait()
            return out if out.dtype == torch.float32 else out.float()
        return finish

    def gather_rows(self, local: torch.Tensor) -> torch.Tensor:
        """[n_local, F] -> [world*per, F]: every rank's block, padded to `per` rows, in rank order
        (so row g of the result is global node g).  This is the halo exchange."""
        if self.world == 1 and self._gather is None:
            return local
        return self.gather_cols_async(local, 0, local.size(1))()

"""Fan-out neighbour sampling on the GPU (K10) behind the names the reference uses: a
`NeighborLoader(data, num_neighbors, input_nodes, batch_size, shuffle)` iterator like
torch_geometric.loader.NeighborLoader 2.3.0 (STEM-GNN/pretrain.py:151-153, utils/loader.py:11-24)
and the functional `neighbor_sample` underneath it.  Sampling, relabelling and the attribute
slices all run on the device; one batch needs no host synchronisation until its sizes are read."""
import copy
import ctypes as C
import math

import torch

from . import _lib
from .augment import gather_rows
from .graph import CSR


class NeighborSampler:
    """Holds the in-edge CSR of one graph (K0, built once), the persistent node map and the Philox
    seed; `sample(seeds)` returns (n_id, edge_index_local, e_id) trimmed to their true sizes."""

    def __init__(self, edge_index, num_nodes, num_neighbors, seed=None):
        if not edge_index.is_cuda:
            raise _lib.StemGnnError("NeighborSampler: edge_index must be a CUDA tensor (no CPU fallback)")
        self.num_nodes, self.num_edges = int(num_nodes), edge_index.size(1)
        self.fanout = [int(k) for k in num_neighbors]
        self.device = edge_index.device
        self.csr = CSR(edge_index.contiguous(), self.num_nodes, 1)
        self._fan = (C.c_int32 * len(self.fanout))(*self.fanout)
        self.node_map = torch.empty(self.num_nodes, dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib().stemgnn_node_map_init(_lib.ptr(self.node_map), self.num_nodes, _lib.stream()))
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        self.seed = torch.tensor([seed], dtype=torch.int64, device=self.device)

    def bounds(self, num_seeds):
        nc, ec, ws = C.c_int64(0), C.c_int64(0), C.c_size_t(0)
        _lib.check(_lib.lib().stemgnn_neighbor_sample_bounds(
            num_seeds, self.num_nodes, self.num_edges, len(self.fanout), self._fan, C.byref(nc),
            C.byref(ec), C.byref(ws)))
        return nc.value, ec.value, ws.value

    def sample_padded(self, seeds):
        """Sync-free form: outputs at their capacity plus the device-resident `counts` vector
        ([0] nodes, [1] edges, then cumulative per hop, last = error bits).  The seed advances by one
        on the device so a captured graph draws fresh subsets on every replay."""
        L = _lib.lib()
        seeds = seeds.to(self.device, torch.int64).contiguous()
        B, H = seeds.numel(), len(self.fanout)
        nc, ec, wsb = self.bounds(B)
        n_id = torch.empty(max(nc, 1), dtype=torch.int64, device=self.device)
        ei = torch.empty(2, max(ec, 1), dtype=torch.int64, device=self.device)
        e_id = torch.empty(max(ec, 1), dtype=torch.int64, device=self.device)
        counts = torch.zeros(3 + 2 * H, dtype=torch.int64, device=self.device)
        ws = _lib.workspace(wsb, self.device)
        a = _lib.NeighborSampleArgs(
            rowptr=_lib.ptr(self.csr.rowptr), col=_lib.ptr(self.csr.col), perm=_lib.ptr(self.csr.perm),
            num_nodes=self.num_nodes, num_edges=self.num_edges, seeds=_lib.ptr(seeds), num_seeds=B,
            num_hops=H, fanout_host=self._fan, seed=_lib.ptr(self.seed), node_map=_lib.ptr(self.node_map),
            node_cap=n_id.numel(), edge_cap=ei.size(1), n_id=_lib.ptr(n_id), edge_index_out=_lib.ptr(ei),
            e_id=_lib.ptr(e_id), counts=_lib.ptr(counts))
        with _lib.traced("neighbor_sample", 0, 0, 4 + 9 * H):
            _lib.check(L.stemgnn_neighbor_sample(C.byref(a), _lib.ptr(ws), wsb, _lib.stream()))
        self.seed.add_(1)
        return n_id, ei, e_id, counts

    def sample(self, seeds):
        n_id, ei, e_id, counts = self.sample_padded(seeds)
        c = counts.tolist()                       # the one synchronisation of a batch
        if c[-1]:
            raise _lib.StemGnnError(f"neighbor_sample: device error bits {c[-1]} "
                                    "(1 = seed out of range, 2 / 4 = capacity exceeded)")
        n, e = c[0], c[1]
        return n_id[:n], ei[:, :e], e_id[:e]


def neighbor_sample(edge_index, num_nodes, seeds, num_neighbors, seed=None):
    """One-shot functional form (builds the CSR every call; keep a NeighborSampler for loops)."""
    return NeighborSampler(edge_index, num_nodes, num_neighbors, seed).sample(seeds)


class _Batch:
    """Attribute bag with the fields PyG's sampled `Data` carries (`n_id`, `e_id`, `edge_index`,
    `batch_size`, `input_id`, `num_nodes`, plus the sliced node / edge attributes)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def keys(self):
        return list(self.__dict__)

    def to(self, device):
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                self.__dict__[k] = v.to(device)
        return self


class NeighborLoader:
    """`for data in NeighborLoader(data, num_neighbors=[10, 10], input_nodes=idx, batch_size=1024,
    shuffle=True)`.  `data` is any object with `.edge_index` ([2, E] int64) and tensor attributes;
    like PyG, attributes whose first dimension equals the node count are sliced by `n_id`, those
    matching the edge count by `e_id`, everything else is passed through (the reference's
    `node_text_feat` / `edge_text_feat` tables, pretrain.py:33-38)."""

    def __init__(self, data, num_neighbors, input_nodes=None, batch_size=1, shuffle=False,
                 num_nodes=None, device=None, drop_last=False, seed=None, **unused):
        attrs = dict(data) if isinstance(data, dict) else dict(vars(data))
        ei = attrs["edge_index"]
        device = torch.device(device) if device is not None else (ei.device if ei.is_cuda else torch.device("cuda"))
        self.device = device
        if num_nodes is None:
            num_nodes = attrs.get("num_nodes")
        if num_nodes is None:
            x = attrs.get("x")
            num_nodes = x.size(0) if torch.is_tensor(x) else int(ei.max()) + 1
        self.num_nodes, self.num_edges = int(num_nodes), ei.size(1)
        self.attrs = {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in attrs.items()
                      if k not in ("edge_index", "num_nodes")}
        self.sampler = NeighborSampler(ei.to(device), self.num_nodes, num_neighbors, seed)
        if input_nodes is None:
            input_nodes = torch.arange(self.num_nodes)
        if input_nodes.dtype == torch.bool:
            input_nodes = input_nodes.nonzero().view(-1)
        self.input_nodes = input_nodes.to(device, torch.int64)
        self.batch_size, self.shuffle, self.drop_last = int(batch_size), shuffle, drop_last

    def __len__(self):
        n = self.input_nodes.numel()
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def _slice(self, v, n_id, e_id):
        if not torch.is_tensor(v) or v.dim() == 0:
            return copy.copy(v)
        ids = n_id if v.size(0) == self.num_nodes else e_id if v.size(0) == self.num_edges else None
        if ids is None:
            return v
        if v.dtype == torch.float32 and v.dim() == 2 and v.size(1) % 4 == 0:
            return gather_rows(v, ids)
        return v[ids]

    def __iter__(self):
        n = self.input_nodes.numel()
        order = torch.randperm(n, device=self.device) if self.shuffle else torch.arange(n, device=self.device)
        for b in range(len(self)):
            input_id = order[b * self.batch_size:(b + 1) * self.batch_size]
            seeds = self.input_nodes[input_id]
            n_id, ei, e_id = self.sampler.sample(seeds)
            out = {k: self._slice(v, n_id, e_id) for k, v in self.attrs.items()}
            out.update(edge_index=ei, n_id=n_id, e_id=e_id, input_id=input_id, batch_size=seeds.numel(),
                       num_nodes=n_id.numel())
            yield _Batch(**out)

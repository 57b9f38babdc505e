"""Drop-in `Encoder` / `InnerProductDecoder` (reference model/encoder.py:132-380): same
constructor, forward/encode signature, attributes and state_dict keys (`layers.{i}.*`,
`norms.{i}.*`, `env_encoders.{j}.*`); the conv layers are the B200 ones from `conv.py`."""
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from .conv import GCNConv, GINConv, MixtureSageLayer, MySAGEConv
from . import dense, norm_act
from .decoder import edge_dot
from .graph import structure_of


class Encoder(nn.Module):
    def __init__(self, input_dim, hidden_dim, activation, num_layers, backbone="sage",
                 normalize="none", dropout=0.0, moe=False, num_experts=3, tau=1.0,
                 moe_layers="all"):
        super().__init__()
        self.input_dim, self.hidden_dim, self.num_layers = input_dim, hidden_dim, num_layers
        self.backbone, self.normalize = backbone, normalize
        self.moe = bool(moe) and num_experts > 1
        self.num_experts, self.tau, self.moe_layers = num_experts, tau, moe_layers
        self.activation = activation()
        self.dropout = nn.Dropout(dropout)
        self.layers, self.norms, self.env_encoders = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        self._last_env_reg: Optional[Tensor] = None
        self._moe_usage: Optional[List[dict]] = None
        self._router_cache: Optional[list] = None
        self._cache_router = False
        self.moe_layer_flags = self._build_moe_layer_flags()
        widths = [input_dim] + [hidden_dim] * num_layers
        for i in range(num_layers):
            fan_in, fan_out = widths[i], widths[i + 1]
            if self.moe_layer_flags[i] and backbone == "sage":
                self.layers.append(MixtureSageLayer(fan_in, fan_out, num_experts, residual=True))
                self.env_encoders.append(nn.Linear(fan_in, num_experts))
            else:
                self.layers.append(self._build_conv(fan_in, fan_out))
            self.norms.append(nn.BatchNorm1d(fan_out))   # always built (encoder.py:173)
        self.reset_parameters()

    def _build_moe_layer_flags(self):
        n = self.num_layers
        if not self.moe or self.moe_layers == "none":
            return [False] * n
        if self.moe_layers == "all":
            return [True] * n
        if self.moe_layers == "last":
            return [False] * max(n - 1, 0) + [True] * min(n, 1)
        raise ValueError(f"Unsupported moe_layers setting: {self.moe_layers}")

    def _build_conv(self, in_dim, out_dim):
        if self.backbone == "sage":
            return MySAGEConv(in_dim, out_dim, aggr="mean", normalize=False, root_weight=True)
        if self.backbone == "gcn":
            return GCNConv(in_dim, out_dim)
        if self.backbone == "gin":
            return GINConv(nn.Linear(in_dim, out_dim))
        if self.backbone == "gat":
            raise NotImplementedError("backbone='gat' (edge-softmax attention) is outside the "
                                      "aggregation hot path this package implements")
        raise ValueError(f"Unsupported backbone: {self.backbone}")

    def reset_parameters(self):
        for m in list(self.layers) + list(self.norms) + list(self.env_encoders):
            m.reset_parameters()
        self._last_env_reg = None
        self._moe_usage = None

    # ---- MoE bookkeeping (encoder.py:202-277) -------------------------------------------------
    @staticmethod
    def _reg_loss(weights, logits):
        return (weights * F.log_softmax(logits, dim=-1)).sum(-1).mean()

    def enable_router_cache(self, flag: bool = True):
        self._cache_router = flag
        self._router_cache = [] if flag else None

    def get_router_cache(self, reset: bool = True):
        out = self._router_cache or []
        if reset:
            self._router_cache = [] if self._cache_router else None
        return out

    def _update_moe_usage(self, slot, weights):
        if not self.moe:
            return
        w = weights.detach()
        if self._moe_usage is None:
            self._moe_usage = [dict(sum_prob=w.new_zeros(self.num_experts),
                                    sum_top1=w.new_zeros(self.num_experts), count=0)
                               for f in self.moe_layer_flags if f]
        st = self._moe_usage[slot]
        st["sum_prob"] += w.sum(0)
        st["sum_top1"] += torch.bincount(w.argmax(-1), minlength=self.num_experts).to(w.dtype)
        st["count"] += w.size(0)

    def get_moe_usage(self, reset=True):
        if not self.moe or self._moe_usage is None:
            return []
        layer_ids = [i for i, f in enumerate(self.moe_layer_flags) if f]
        usage = []
        for layer, st in zip(layer_ids, self._moe_usage):
            d = max(st["count"], 1)
            usage.append({"layer": layer, "avg_prob": (st["sum_prob"] / d).cpu().tolist(),
                          "top1_frac": (st["sum_top1"] / d).cpu().tolist()})
        if reset:
            self._moe_usage = None
        return usage

    # ---- forward --------------------------------------------------------------------------------
    def forward(self, x, edge_index, edge_attr=None):
        return self.encode(x, edge_index, edge_attr)

    def encode(self, x, edge_index, edge_attr=None):
        """conv -> BatchNorm (normalize != 'none') -> activation + dropout on all but the last
        layer (encoder.py:283-323).  The CSR/CSC structure of `edge_index` is built once and
        shared by every layer and by the backward pass."""
        graph = structure_of(edge_index, x.size(0)) if isinstance(edge_index, Tensor) else edge_index
        sync = graph.batch_sync() if hasattr(graph, "batch_sync") else None    # partitioned graph: full-batch BN
        z, slot, reg_sum, reg_n = x, 0, None, 0
        self._last_env_reg = None
        for i, layer in enumerate(self.layers):
            if isinstance(layer, MixtureSageLayer):
                logits = self.env_encoders[slot](z)
                if self.training:
                    gate = F.gumbel_softmax(logits, tau=self.tau, dim=-1)
                    r = self._reg_loss(gate, logits)
                    reg_sum, reg_n = (r if reg_sum is None else reg_sum + r), reg_n + 1
                else:
                    gate = F.softmax(logits, dim=-1)
                if self._cache_router:
                    self._router_cache = (self._router_cache or []) + [gate.detach()]
                self._update_moe_usage(slot, gate)
                z = (gate.unsqueeze(-1) * layer(z, graph, edge_attr)).sum(dim=1)
                slot += 1
            else:
                z = layer(z, graph, edge_attr)
            act_here = i < self.num_layers - 1
            if (self.normalize != "none" and self.dropout.p < 1.0
                    and norm_act.supported(z, self.norms[i], self.activation if act_here else None)):
                # BatchNorm (+ ReLU + dropout) in three passes over z (K7) on the same norms[i] module
                z = norm_act.bn_act_drop(z, self.norms[i], relu=act_here,
                                         p=self.dropout.p if act_here else 0.0, training=self.dropout.training,
                                         sync=sync)
            else:
                if self.normalize != "none":
                    if sync is not None and self.norms[i].training and not isinstance(self.norms[i], nn.SyncBatchNorm):
                        raise NotImplementedError("BatchNorm over a partitioned graph needs the fused K7 path "
                                                  "(float32, width % 4 == 0, ReLU or no activation)")
                    z = self.norms[i](z)
                if act_here:
                    z = self.dropout(self.activation(z))
        self._last_env_reg = reg_sum / reg_n if (reg_sum is not None and reg_n > 0) else z.new_zeros(1)
        return z

    def get_env_reg(self, reset=True):
        reg = self._last_env_reg
        if reg is None:
            reg = torch.zeros(1, device=next(self.parameters()).device)
        if reset:
            self._last_env_reg = None
        return reg


class InnerProductDecoder(nn.Module):
    """sigma(<z_i, z_j>) over given node pairs (encoder.py:336-380).  The edge-wise decode runs
    through K8 (`decoder.edge_dot`: no E x F gathered matrices, deterministic backward through
    K1); `forward_all` is a plain dense product."""

    def __init__(self, hidden_dim=None, output_dim=None):
        super().__init__()
        self.proj_z = hidden_dim is not None
        if self.proj_z:
            self.lin = nn.Linear(hidden_dim, output_dim)

    def _project(self, z):
        if not self.proj_z:
            return z
        w = self.lin.weight
        if z.dim() == 2:
            return dense.linear_sum([z.float()], [w], self.lin.bias)  # K5/K6 instead of torch's fp32 SGEMM
        return self.lin(z)

    def forward(self, z: Tensor, edge_index: Tensor, sigmoid: bool = True) -> Tensor:
        return edge_dot(self._project(z), edge_index, sigmoid)

    def forward_all(self, z: Tensor, sigmoid: bool = True) -> Tensor:
        z = self.lin(z) if self.proj_z else z
        adj = z @ z.t()
        return torch.sigmoid(adj) if sigmoid else adj

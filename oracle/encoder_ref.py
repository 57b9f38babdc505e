"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference encoder modules
(/root/reference/STEM-GNN/model/encoder.py) on top of oracle.propagate_ref.  Parameter
names match the reference's state_dict so weights can be copied across.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import propagate_ref as P


class SAGEConvRef(nn.Module):
    """MySAGEConv (encoder.py:17-106) with the arguments Encoder uses (encoder.py:193):
    out = lin_l(mean_i relu(x_j [+ xe])) + lin_r(x)."""

    def __init__(self, in_channels, out_channels, aggr="mean", normalize=False,
                 root_weight=True, bias=True):
        super().__init__()
        self.aggr, self.normalize, self.root_weight = aggr, normalize, root_weight
        self.lin_l = nn.Linear(in_channels, out_channels, bias=bias)      # encoder.py:58
        if root_weight:
            self.lin_r = nn.Linear(in_channels, out_channels, bias=False)  # encoder.py:60

    def forward(self, x, edge_index, edge_attr=None, size=None):
        out = P.sage_propagate_ref(x, edge_index, edge_attr, x.size(0), self.aggr)  # :82
        out = self.lin_l(out)                                                        # :83
        if self.root_weight:
            out = out + self.lin_r(x)                                                # :85-87
        if self.normalize:
            out = F.normalize(out, p=2.0, dim=-1)                                    # :89-90
        return out


class GCNConvRef(nn.Module):
    """PyG GCNConv as built at encoder.py:197: lin (no bias) -> gcn_norm propagate -> + bias."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.lin = nn.Linear(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels))

    def forward(self, x, edge_index, edge_weight=None, size=None):
        assert edge_weight is None
        return P.gcn_propagate_ref(self.lin(x), edge_index, x.size(0)) + self.bias


class GINConvRef(nn.Module):
    """PyG GINConv(nn.Linear) as built at encoder.py:199: nn((1 + eps) * x + sum_j x_j)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.nn = nn.Linear(in_channels, out_channels)
        self.register_buffer("eps", torch.zeros(1))

    def forward(self, x, edge_index, size=None):
        out = P.gin_aggregate_ref(x, edge_index)
        return self.nn(out + (1 + self.eps) * x)


class MixtureSageRef(nn.Module):
    """MixtureSageLayer (encoder.py:109-129)."""

    def __init__(self, in_dim, out_dim, num_experts, residual=True):
        super().__init__()
        self.residual = residual and in_dim == out_dim
        self.weights = nn.Parameter(torch.empty(num_experts, in_dim * 2, out_dim))
        nn.init.xavier_uniform_(self.weights)

    def forward(self, x, edge_index, edge_attr=None):
        agg = P.mixture_aggregate_ref(x, edge_index)                      # :123-124
        out = torch.einsum("nd,kdo->nko", torch.cat([agg, x], -1), self.weights)  # :125-126
        return out + x.unsqueeze(1) if self.residual else out             # :127-129


class EncoderRef(nn.Module):
    """Encoder.encode layer loop (encoder.py:283-323): conv -> BatchNorm (normalize != 'none')
    -> activation + dropout on all but the last layer.  MoE gating uses softmax (eval branch,
    encoder.py:300) when `deterministic_gate`, else gumbel_softmax like the reference."""

    def __init__(self, input_dim, hidden_dim, activation, num_layers, backbone="sage",
                 normalize="none", dropout=0.0, moe=False, num_experts=3, tau=1.0,
                 moe_layers="all"):
        super().__init__()
        self.num_layers, self.normalize, self.tau = num_layers, normalize, tau
        self.moe = moe and num_experts > 1
        self.activation = activation()
        self.dropout = nn.Dropout(dropout)
        self.layers, self.norms, self.env_encoders = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        if not self.moe or moe_layers == "none":
            flags = [False] * num_layers
        elif moe_layers == "all":
            flags = [True] * num_layers
        else:
            flags = [False] * (num_layers - 1) + [True]
        self.moe_layer_flags = flags
        dims = [input_dim] + [hidden_dim] * num_layers
        for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            if flags[i] and backbone == "sage":
                self.layers.append(MixtureSageRef(a, b, num_experts))
                self.env_encoders.append(nn.Linear(a, num_experts))
            elif backbone == "sage":
                self.layers.append(SAGEConvRef(a, b))
            elif backbone == "gcn":
                self.layers.append(GCNConvRef(a, b))
            elif backbone == "gin":
                self.layers.append(GINConvRef(a, b))
            else:
                raise ValueError(backbone)
            self.norms.append(nn.BatchNorm1d(b))

    def forward(self, x, edge_index, edge_attr=None):
        z, env = x, 0
        for i, layer in enumerate(self.layers):
            if isinstance(layer, MixtureSageRef):
                logits = self.env_encoders[env](z)
                w = F.gumbel_softmax(logits, tau=self.tau, dim=-1) if self.training \
                    else F.softmax(logits, dim=-1)
                z = torch.sum(w.unsqueeze(-1) * layer(z, edge_index, edge_attr), dim=1)
                env += 1
            else:
                z = layer(z, edge_index, edge_attr)
            if self.normalize != "none":
                z = self.norms[i](z)
            if i < self.num_layers - 1:
                z = self.dropout(self.activation(z))
        return z

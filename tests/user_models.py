"""Models the way USERS of the reference write them: plain ``nn.Module`` classes that know nothing about
this framework (no ``fused_spec``), used by the lowering / engine tests and by ``bench.py``.

``NYC_Model`` is the class a user finds in the reference's flagship example
(``/root/reference/examples/pytorch_nyctaxi.py:69-92``) and pastes into their own script; ``Dlrm`` follows
the module structure of the reference's DLRM notebook (``examples/pytorch_dlrm.ipynb`` cell 9: a bottom
model holding ``MultiTableEmbeddings`` + an MLP, a top model holding the interaction + an MLP + ``out``,
``forward(*inputs)`` taking one ``[B, 1]`` tensor per feature column).  ``tests/test_lowering.py`` also
runs the lowering on the reference's own source when the checkout is present.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class NYC_Model(nn.Module):
    def __init__(self, cols):
        super().__init__()
        self.fc1 = nn.Linear(cols, 256)
        self.fc2 = nn.Linear(256, 128)
        self.fc3 = nn.Linear(128, 64)
        self.fc4 = nn.Linear(64, 16)
        self.fc5 = nn.Linear(16, 1)
        self.bn1 = nn.BatchNorm1d(256)
        self.bn2 = nn.BatchNorm1d(128)
        self.bn3 = nn.BatchNorm1d(64)
        self.bn4 = nn.BatchNorm1d(16)

    def forward(self, x):
        x = F.relu(self.fc1(x))
        x = self.bn1(x)
        x = F.relu(self.fc2(x))
        x = self.bn2(x)
        x = F.relu(self.fc3(x))
        x = self.bn3(x)
        x = F.relu(self.fc4(x))
        x = self.bn4(x)
        x = self.fc5(x)
        return x


class GeluNet(nn.Module):
    """A model no fused whole-step engine covers: GELU, LayerNorm, a residual branch, dropout."""

    def __init__(self, cols, width=64, p=0.0):
        super().__init__()
        self.inp = nn.Linear(cols, width)
        self.norm = nn.LayerNorm(width)
        self.body = nn.Sequential(nn.Linear(width, 2 * width), nn.ReLU(), nn.Linear(2 * width, width))
        self.drop = nn.Dropout(p)
        self.head = nn.Linear(width, 1)

    def forward(self, x):
        h = F.gelu(self.inp(x))
        h = self.norm(h + self.drop(self.body(h)))
        return self.head(h)


# ---- DLRM, laid out like the reference notebook's classes ------------------------------------------
class Interaction:
    """Pairwise dot products of the (1 + tables) feature vectors, lower triangle, plus the dense vector."""

    def __init__(self, n_tables, dim):
        self.n, self.dim = n_tables + 1, dim
        rows = [i for i in range(self.n) for _ in range(i)]
        cols = [j for i in range(self.n) for j in range(i)]
        self.tril = torch.tensor([rows, cols])

    @property
    def width(self):
        return self.n * (self.n - 1) // 2 + self.dim + 1          # +1: zero pad to a multiple of 8

    def __call__(self, stacked, dense_vec):
        dots = torch.bmm(stacked, stacked.transpose(1, 2))[:, self.tril[0], self.tril[1]]
        pad = torch.zeros(stacked.shape[0], 1, dtype=stacked.dtype, device=stacked.device)
        return torch.cat((dense_vec, dots, pad), dim=1)


class Mlp(nn.Module):
    def __init__(self, in_dim, sizes):
        super().__init__()
        layers = []
        for out_dim in sizes:
            layers += [nn.Linear(in_dim, out_dim), nn.ReLU(inplace=True)]
            in_dim = out_dim
        self.layers = nn.Sequential(*layers)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / (m.in_features + m.out_features)))
                nn.init.normal_(m.bias, 0.0, math.sqrt(1.0 / m.out_features))

    def forward(self, x):
        return self.layers(x)


class Tables(nn.Module):
    def __init__(self, sizes, dim):
        super().__init__()
        self.embeddings = nn.ModuleList(
            nn.Embedding.from_pretrained(torch.empty(n, dim).uniform_(-math.sqrt(1 / n), math.sqrt(1 / n)),
                                         freeze=False, sparse=True) for n in sizes)

    def forward(self, cat):
        return [e(cat[:, i]).unsqueeze(1) for i, e in enumerate(self.embeddings)]


class Bottom(nn.Module):
    def __init__(self, n_dense, sizes, mlp_sizes, dim):
        super().__init__()
        self.embeddings = Tables(sizes, dim)
        self.mlp = Mlp(n_dense, mlp_sizes)

    def forward(self, dense, cat):
        vec = self.mlp(dense)
        return torch.cat([vec.view(len(dense), 1, -1)] + self.embeddings(cat), dim=1), vec


class Top(nn.Module):
    def __init__(self, sizes, interaction):
        super().__init__()
        self.interaction = interaction
        self.mlp = Mlp(interaction.width, sizes[:-1])
        self.out = nn.Linear(sizes[-2], sizes[-1])
        nn.init.zeros_(self.mlp.layers[0].weight[:, -1].data)

    def forward(self, stacked, vec):
        return self.out(self.mlp(self.interaction(stacked, vec)))


class Dlrm(nn.Module):
    def __init__(self, num_numerical_features, categorical_feature_sizes, bottom_mlp_sizes, top_mlp_sizes,
                 embedding_dim=32):
        super().__init__()
        self.num_numerical_features = num_numerical_features
        self.bottom_model = Bottom(num_numerical_features, categorical_feature_sizes, bottom_mlp_sizes, embedding_dim)
        self.top_model = Top(top_mlp_sizes, Interaction(len(categorical_feature_sizes), embedding_dim))

    def forward(self, *inputs):
        inputs = list(inputs)
        dense = torch.cat(inputs[:self.num_numerical_features], dim=1)
        cat = torch.cat(inputs[self.num_numerical_features:], dim=1)
        return self.top_model(*self.bottom_model(dense, cat))

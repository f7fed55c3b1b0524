"""Fast stand-in for the reference index build, used only to obtain *some* trained weights
and bucket assignment for benchmarks and tests.

Index build (K-means + MLP training, ``search/li/LearnedIndex.py:197-282``,
``search/li/model.py:149-199``) is OUT OF SCOPE of this project (BASELINE.json north_star:
"K-means/MLP index build stays on the reference path").  Parity only needs the same weights
on both sides; the reference build costs 8.2 h at 10M (README.md:18), so benchmarks train
the same architectures (``model.py:21-78``) with an ordinary mini-batch trainer in torch.
Nothing here is on the timed query path.
"""
from __future__ import annotations

import numpy as np
import torch

# hidden widths of the reference's layer stacks (search/li/model.py:21-70).
# 'MLP-9' (model.py:71-78) is shape-inconsistent in the reference and cannot run.
HIDDEN_WIDTHS = {
    "MLP": [128], "MLP-2": [64], "MLP-3": [256], "MLP-4": [512],
    "MLP-5": [256, 128], "MLP-6": [32], "MLP-7": [16], "MLP-8": [8],
}


def make_mlp(d_in: int, n_categories: int, model_type: str = "MLP-5") -> torch.nn.Sequential:
    widths = [d_in] + HIDDEN_WIDTHS[model_type] + [n_categories]
    layers = []
    for i in range(len(widths) - 1):
        layers.append(torch.nn.Linear(widths[i], widths[i + 1]))
        if i + 2 < len(widths):
            layers.append(torch.nn.ReLU())
    return torch.nn.Sequential(*layers)


def mlp_params(model: torch.nn.Sequential):
    """-> (weights [out,in] f32 ndarray list, biases list), the ``nn.Linear`` parameters in order."""
    lin = [m for m in model if isinstance(m, torch.nn.Linear)]
    return ([m.weight.detach().cpu().numpy().astype(np.float32).copy() for m in lin],
            [m.bias.detach().cpu().numpy().astype(np.float32).copy() for m in lin])


@torch.no_grad()
def kmeans(x: torch.Tensor, k: int, iters: int = 20, seed: int = 2023) -> torch.Tensor:
    g = torch.Generator(device="cpu").manual_seed(seed)
    c = x[torch.randperm(x.shape[0], generator=g)[:k].to(x.device)].clone()
    for _ in range(iters):
        lab = torch.cdist(x, c).argmin(1)
        sums = torch.zeros_like(c).index_add_(0, lab, x)
        cnt = torch.bincount(lab, minlength=k).clamp_min(1).unsqueeze(1)
        c = sums / cnt
    return c


def train_index(data_nav: torch.Tensor, n_categories: int, model_type: str = "MLP-5",
                sample: int = 400_000, epochs: int = 6, batch: int = 2048, lr: float = 2e-3,
                seed: int = 2023):
    """Returns (model on data_nav.device, weights, biases).  Deterministic given the seed."""
    dev = data_nav.device
    torch.manual_seed(seed)
    if dev.type == "cuda":
        torch.cuda.manual_seed(seed)
    n = data_nav.shape[0]
    g = torch.Generator(device="cpu").manual_seed(seed)
    idx = torch.randperm(n, generator=g)[: min(n, sample)].to(dev)
    xs = data_nav[idx]
    cent = kmeans(xs, n_categories, seed=seed)
    ys = torch.cdist(xs, cent).argmin(1)
    model = make_mlp(data_nav.shape[1], n_categories, model_type).to(dev)
    opt = torch.optim.Adam(model.parameters(), lr=lr)
    loss_fn = torch.nn.CrossEntropyLoss()
    for ep in range(epochs):
        perm = torch.randperm(xs.shape[0], generator=g).to(dev)
        for lo in range(0, xs.shape[0], batch):
            sel = perm[lo:lo + batch]
            opt.zero_grad(set_to_none=True)
            loss_fn(model(xs[sel]), ys[sel]).backward()
            opt.step()
    model.eval()
    w, b = mlp_params(model)
    return model, w, b

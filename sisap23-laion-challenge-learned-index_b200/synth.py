"""Synthetic LAION-shaped inputs (there is no network: the real ``clip768v2`` / ``pca96v2``
files that ``search/li/utils.py:71-82`` downloads cannot be fetched).

Recipe: a latent Gaussian mixture in R^32 pushed through a fixed random
linear map into R^768 plus isotropic noise, L2-normalised and **rounded to float16** (the
float16 array is the dataset, exactly like ``clip768v2``'s ``emb``, so norms are 1 +- 5e-4
and inverse norms still matter).  ``pca96`` navigation vectors are a PCA of the first 50k
rows followed by the caller-side ``sklearn.preprocessing.normalize`` (``search.py:50-52``).

Everything is generated with torch so the same code runs on the host (tests) and directly
in HBM (bench at 10M / 100M rows, where a host copy would cost more than the search).
"""
from __future__ import annotations

from dataclasses import dataclass, asdict

import torch


@dataclass
class SynthSpec:
    n: int = 300_000
    nq: int = 10_000
    d: int = 768
    d_nav: int = 96
    latent: int = 32
    n_blobs: int = 2000          # G: number of mixture components
    sigma: float = 0.7           # within-blob spread in latent space
    noise: float = 0.05          # isotropic noise added in R^d
    seed_data: int = 2023        # the reference's seed (search.py:14)
    seed_queries: int = 2024
    pca_rows: int = 50_000

    def as_dict(self):
        return asdict(self)


class SynthWorld:
    """Holds the fixed parts of the generator (centres, lift matrix, PCA) on ``device``."""

    def __init__(self, spec: SynthSpec, device="cpu"):
        self.spec = spec
        self.device = torch.device(device)
        g = torch.Generator(device="cpu").manual_seed(spec.seed_data)
        self.centres = torch.randn(spec.n_blobs, spec.latent, generator=g).to(self.device)
        self.lift = (torch.randn(spec.latent, spec.d, generator=g) / spec.latent ** 0.5).to(self.device)
        self.mean = None
        self.proj = None

    def _draw(self, rows: int, gen: torch.Generator) -> torch.Tensor:
        """rows x d float16 unit-ish vectors."""
        s = self.spec
        which = torch.randint(0, s.n_blobs, (rows,), generator=gen, device=self.device)
        z = self.centres[which] + s.sigma * torch.randn(rows, s.latent, generator=gen, device=self.device)
        x = z @ self.lift + s.noise * torch.randn(rows, s.d, generator=gen, device=self.device)
        x = x / x.norm(dim=1, keepdim=True).clamp_min(1e-12)
        return x.to(torch.float16)

    def fit_pca(self, sample_f16: torch.Tensor) -> None:
        x = sample_f16.float()
        self.mean = x.mean(0)
        xc = (x - self.mean).double()
        cov = (xc.T @ xc) / max(1, xc.shape[0] - 1)
        evals, evecs = torch.linalg.eigh(cov.cpu())           # ascending
        self.proj = evecs[:, -self.spec.d_nav:].flip(1).float().to(self.device).contiguous()

    def project(self, x_f16: torch.Tensor) -> torch.Tensor:
        """pca96 + the caller's L2 normalisation (search.py:50-52) -> float32 [rows, d_nav]."""
        y = (x_f16.float() - self.mean) @ self.proj
        return (y / y.norm(dim=1, keepdim=True).clamp_min(1e-12)).contiguous()

    def generate(self, chunk: int = 262_144):
        """Returns dict(data_f16 [n,d], data_nav [n,d_nav], queries_f16, queries_nav)."""
        s = self.spec
        gen = torch.Generator(device=self.device).manual_seed(s.seed_data)
        data = torch.empty(s.n, s.d, dtype=torch.float16, device=self.device)
        for lo in range(0, s.n, chunk):
            hi = min(s.n, lo + chunk)
            data[lo:hi] = self._draw(hi - lo, gen)
        self.fit_pca(data[: min(s.n, s.pca_rows)])
        nav = torch.empty(s.n, s.d_nav, dtype=torch.float32, device=self.device)
        for lo in range(0, s.n, chunk):
            hi = min(s.n, lo + chunk)
            nav[lo:hi] = self.project(data[lo:hi])
        genq = torch.Generator(device=self.device).manual_seed(s.seed_queries)
        queries = self._draw(s.nq, genq)
        return dict(data_f16=data, data_nav=nav, queries_f16=queries,
                    queries_nav=self.project(queries))


    # ---- chunk-seeded stream: any process can produce any row range (datasets that never exist in one piece) ----
    def draw_range(self, lo: int, hi: int, chunk: int = 262_144) -> torch.Tensor:
        """Rows [lo, hi) of the chunk-seeded dataset: block c (rows c*chunk ...) comes from its own
        generator seeded by (seed_data, c), so the rows do not depend on who draws them or in what order.
        (A different stream from ``generate()``'s sequential one.)"""
        s = self.spec
        out = torch.empty(hi - lo, s.d, dtype=torch.float16, device=self.device)
        for c in range(lo // chunk, (hi + chunk - 1) // chunk):
            gen = torch.Generator(device=self.device).manual_seed(s.seed_data * 1_000_003 + 1 + c)
            block = self._draw(chunk, gen)                      # always a full block: the tail must not change the stream
            b_lo, b_hi = max(lo, c * chunk), min(hi, (c + 1) * chunk)
            out[b_lo - lo:b_hi - lo] = block[b_lo - c * chunk:b_hi - c * chunk]
        return out

    def generate_range(self, lo: int, hi: int, chunk: int = 262_144):
        """dict(data_f16 [hi-lo,d], data_nav, queries_f16, queries_nav) for rows [lo, hi) of the chunk-seeded
        dataset of ``spec.n`` rows; PCA is fitted on rows [0, pca_rows) by every caller identically."""
        s = self.spec
        self.fit_pca(self.draw_range(0, min(s.n, s.pca_rows), chunk))
        data = self.draw_range(lo, hi, chunk)
        nav = torch.empty(hi - lo, s.d_nav, dtype=torch.float32, device=self.device)
        for a in range(0, hi - lo, chunk):
            nav[a:a + chunk] = self.project(data[a:a + chunk])
        genq = torch.Generator(device=self.device).manual_seed(s.seed_queries)
        queries = self._draw(s.nq, genq)
        return dict(data_f16=data, data_nav=nav, queries_f16=queries, queries_nav=self.project(queries))


def make_synthetic(spec: SynthSpec, device="cpu"):
    return SynthWorld(spec, device).generate()

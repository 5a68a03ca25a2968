"""Synthetic point clouds for the five BASELINE.json configs (numpy on the host, torch on the GPU).

uniform            config 1, 2, 4: uniform-random float3 in the unit cube
gaussian_blobs     config 3: one Gaussian blob per rank, centres on a grid, ~10 % spatial overlap
clustered_surface  config 5: strongly clustered + surface-sampled points (load imbalance)

The reference ships no generator for its programs (inputs are raw float3 files, /root/reference/README.md:25-39); its
library tests draw uniformly distributed coordinates (integers in [0, N] stored as floats,
cudaKDTree/testing/floatN-knn-and-fcp.cu:73-94 `generatePoints`); `uniform` is the continuous version of that in the
unit cube.
"""
from __future__ import annotations

import math
from typing import List

import numpy as np


def uniform(n: int, seed: int = 0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return rng.random((n, 3), dtype=np.float32)


def blob_centres(parts: int) -> np.ndarray:
    side = max(1, round(parts ** (1.0 / 3.0)))
    while side ** 3 < parts:
        side += 1
    c = np.array([[i, j, l] for i in range(side) for j in range(side) for l in range(side)], np.float32)
    return c[:parts]


def gaussian_blobs(n_per_part: int, parts: int, overlap: float = 0.10, seed: int = 0) -> List[np.ndarray]:
    """One blob per part on a unit-spaced grid. ``overlap`` ~ fraction of a blob's points that fall
    closer to another centre: sigma is chosen so P(|x - c| > 0.5 along one axis) ~ overlap/3."""
    from scipy.stats import norm
    if overlap <= 0.0:
        sigma = 0.05
    elif overlap >= 1.0:
        sigma = 10.0
    else:
        sigma = 0.5 / norm.isf(min(0.49, overlap / 6.0))
    out = []
    for p, c in enumerate(blob_centres(parts)):
        rng = np.random.default_rng(seed * 1000 + p)
        out.append((c + rng.normal(0.0, sigma, (n_per_part, 3))).astype(np.float32))
    return out


def clustered_surface(n: int, seed: int = 0, clusters: int = 64) -> np.ndarray:
    """Half the points in tight power-law-sized clusters, half sampled on a sphere surface."""
    rng = np.random.default_rng(seed)
    n_cl = n // 2
    w = rng.pareto(1.2, clusters) + 0.05
    sizes = np.floor(w / w.sum() * n_cl).astype(np.int64)
    sizes[0] += n_cl - sizes.sum()
    centres = rng.random((clusters, 3))
    parts = []
    for c, s in zip(centres, sizes):
        sig = 10 ** rng.uniform(-3.5, -1.5)
        parts.append(c + rng.normal(0, sig, (int(s), 3)))
    m = n - n_cl
    v = rng.normal(0, 1, (m, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    parts.append(0.5 + 0.45 * v + rng.normal(0, 1e-4, (m, 3)))
    pts = np.concatenate(parts).astype(np.float32)
    rng.shuffle(pts)
    return pts


def torch_uniform(n: int, device, seed: int = 0):
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return torch.rand((n, 3), generator=g, device=device, dtype=torch.float32)

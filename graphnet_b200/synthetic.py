"""Synthetic IceCube-shaped pulse maps for benchmarks and tests (host side).

Nothing here reads the reference tree: the detector geometry is generated
procedurally (a hexagonal string grid with the published IceCube-86 /
IceCube-Upgrade dimensions), and pulses follow the recipe of SURVEY.md
section 8(d), C2/C4: pulses per event ~ max(2, Poisson(mean)), sensors drawn
WITH replacement around a random vertex with probability ~ exp(-r / 40 m)
(the scale is set so that ~30 % of pulses share a sensor, the ratio measured on
the reference's bundled oscNext events: 38 unique sensors / 56 pulses; they
therefore share xyz exactly - the tie structure the kNN must reproduce), time = 1e4 + r/0.3 +
Exp(300) ns, charge ~ LogNormal(0, 1), rde in {1.0, 1.35}, pmt_area 0.0444.
Features are standardised with the reference's detector formulas
(`graphnet_b200.detector`).
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np
import torch

from .detector import (
    FEATURES_ICECUBE86,
    FEATURES_UPGRADE,
    IceCube86,
    IceCubeUpgrade,
)


def _hex_grid(n_rings: int, spacing: float) -> np.ndarray:
    pts = []
    for q in range(-n_rings, n_rings + 1):
        for r in range(max(-n_rings, -q - n_rings), min(n_rings, -q + n_rings) + 1):
            pts.append((spacing * (q + r / 2.0), spacing * (np.sqrt(3) / 2.0) * r))
    return np.asarray(pts, dtype=np.float64)


def icecube86_geometry() -> Dict[str, np.ndarray]:
    """78 strings on a 125 m hexagonal grid x 60 DOMs (17 m) + 8 DeepCore strings."""
    grid = _hex_grid(5, 125.0)  # 91 points
    d = np.hypot(grid[:, 0], grid[:, 1])
    grid = grid[np.argsort(d, kind="stable")][:78]
    xs, ys, zs, st, dn, rde = [], [], [], [], [], []
    for s, (gx, gy) in enumerate(grid):
        z = 500.0 - 17.0 * np.arange(60)
        xs.append(np.full(60, gx)); ys.append(np.full(60, gy)); zs.append(z)
        st.append(np.full(60, s + 1)); dn.append(np.arange(1, 61)); rde.append(np.full(60, 1.0))
    ang = np.arange(8) * (2 * np.pi / 8)
    for s in range(8):
        gx, gy = 60.0 * np.cos(ang[s]) + 20.0, 60.0 * np.sin(ang[s]) - 30.0
        z = np.concatenate([300.0 - 10.0 * np.arange(10), -150.0 - 7.0 * np.arange(50)])
        xs.append(np.full(60, gx)); ys.append(np.full(60, gy)); zs.append(z)
        st.append(np.full(60, 79 + s)); dn.append(np.arange(1, 61)); rde.append(np.full(60, 1.35))
    n = 86 * 60
    return {
        "x": np.concatenate(xs), "y": np.concatenate(ys), "z": np.concatenate(zs),
        "string": np.concatenate(st).astype(np.float64),
        "dom_number": np.concatenate(dn).astype(np.float64),
        "rde": np.concatenate(rde),
        "pmt_area": np.full(n, 0.0444),
        "pmt_number": np.zeros(n), "dom_type": np.full(n, 20.0),
        "pmt_dir": np.tile(np.array([0.0, 0.0, -1.0]), (n, 1)),
    }


def upgrade_geometry() -> Dict[str, np.ndarray]:
    """IceCube-86 plus 7 dense strings of multi-PMT modules (PMTs share xyz)."""
    base = icecube86_geometry()
    rng = np.random.default_rng(7)
    xs, ys, zs, st, dn, pn, dt, pd_ = [], [], [], [], [], [], [], []
    ang = np.arange(7) * (2 * np.pi / 7)
    for s in range(7):
        gx, gy = 35.0 * np.cos(ang[s]) + 20.0, 35.0 * np.sin(ang[s]) - 30.0
        for m in range(95):
            z = -150.0 - 3.0 * m
            npmt = 24 if m % 2 == 0 else 16
            v = rng.normal(size=(npmt, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True)
            xs.append(np.full(npmt, gx)); ys.append(np.full(npmt, gy)); zs.append(np.full(npmt, z))
            st.append(np.full(npmt, 87 + s)); dn.append(np.full(npmt, m + 1))
            pn.append(np.arange(npmt)); dt.append(np.full(npmt, 130.0 if npmt == 24 else 120.0))
            pd_.append(v)
    n_new = sum(len(a) for a in xs)
    out = {
        "x": np.concatenate([base["x"]] + xs), "y": np.concatenate([base["y"]] + ys),
        "z": np.concatenate([base["z"]] + zs),
        "string": np.concatenate([base["string"]] + st).astype(np.float64),
        "dom_number": np.concatenate([base["dom_number"]] + dn).astype(np.float64),
        "rde": np.concatenate([base["rde"], np.full(n_new, 1.0)]),
        "pmt_area": np.concatenate([base["pmt_area"], np.full(n_new, 0.005)]),
        "pmt_number": np.concatenate([base["pmt_number"]] + pn).astype(np.float64),
        "dom_type": np.concatenate([base["dom_type"]] + dt),
        "pmt_dir": np.concatenate([base["pmt_dir"]] + pd_),
    }
    return out


@dataclass
class PulseBatch:
    """A collated batch in the layout DynEdge consumes."""

    x: torch.Tensor  # [N, F] float32, standardised
    batch: torch.Tensor  # [N] int64, sorted
    n_pulses: torch.Tensor  # [B] int32
    ptr: torch.Tensor  # [B+1] int64

    @property
    def num_events(self) -> int:
        return int(self.n_pulses.shape[0])


def make_pulse_batch(
    num_events: int,
    mean_pulses: float = 60.0,
    detector: str = "icecube86",
    seed: int = 1234,
    fixed_pulses: Optional[int] = None,
    scale_m: float = 40.0,
    counts: Optional[np.ndarray] = None,
) -> PulseBatch:
    """Generate `num_events` standardised events (CPU tensors).

    `counts` (optional, [num_events]) fixes the pulses of every event (ragged batches)."""
    rng = np.random.default_rng(seed)
    geo = icecube86_geometry() if detector == "icecube86" else upgrade_geometry()
    pos = np.stack([geo["x"], geo["y"], geo["z"]], axis=1)
    S = pos.shape[0]
    if counts is not None:
        counts = np.maximum(2, np.asarray(counts, dtype=np.int64))
        assert counts.shape == (num_events,)
    elif fixed_pulses is not None:
        counts = np.full(num_events, int(fixed_pulses), dtype=np.int64)
    else:
        counts = np.maximum(2, rng.poisson(mean_pulses, size=num_events)).astype(np.int64)
    ptr = np.zeros(num_events + 1, dtype=np.int64)
    ptr[1:] = np.cumsum(counts)
    N = int(ptr[-1])
    sensor = np.empty(N, dtype=np.int64)
    rdist = np.empty(N, dtype=np.float64)
    lo = pos.min(0) * 0.8
    hi = pos.max(0) * 0.8
    for e in range(num_events):
        vtx = rng.uniform(lo, hi)
        r = np.linalg.norm(pos - vtx, axis=1)
        p = np.exp(-(r - r.min()) / scale_m)
        cdf = np.cumsum(p)
        u = rng.uniform(0.0, cdf[-1], size=counts[e])
        idx = np.minimum(np.searchsorted(cdf, u), S - 1)
        sensor[ptr[e]:ptr[e + 1]] = idx
        rdist[ptr[e]:ptr[e + 1]] = r[idx]
    time = 1.0e4 + rdist / 0.3 + rng.exponential(300.0, size=N)
    charge = rng.lognormal(0.0, 1.0, size=N)
    cols = {
        "dom_x": geo["x"][sensor], "dom_y": geo["y"][sensor], "dom_z": geo["z"][sensor],
        "dom_time": time, "charge": charge, "rde": geo["rde"][sensor],
        "pmt_area": geo["pmt_area"][sensor], "string": geo["string"][sensor],
        "pmt_number": geo["pmt_number"][sensor], "dom_number": geo["dom_number"][sensor],
        "pmt_dir_x": geo["pmt_dir"][sensor, 0], "pmt_dir_y": geo["pmt_dir"][sensor, 1],
        "pmt_dir_z": geo["pmt_dir"][sensor, 2], "dom_type": geo["dom_type"][sensor],
    }
    if detector == "icecube86":
        names, det = FEATURES_ICECUBE86, IceCube86()
    else:
        names, det = FEATURES_UPGRADE, IceCubeUpgrade()
    raw = torch.tensor(np.stack([cols[n] for n in names], axis=1), dtype=torch.float32)
    x = det(raw, names)
    batch = torch.repeat_interleave(torch.arange(num_events), torch.from_numpy(counts))
    return PulseBatch(
        x=x.contiguous(), batch=batch,
        n_pulses=torch.from_numpy(counts).to(torch.int32),
        ptr=torch.from_numpy(ptr),
    )

"""Shared fixtures.  `-m gpu` tests need a B200; everything else runs on CPU."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["oscnext5", "upgrade5", "prometheus50", "adversarial", "adversarial_k16_after"]
# options off the default path (LayerNorm, GELU, 1- and 3-Linear EdgeConv MLPs): fp32 only
GOLDEN_EXTRA = ["oscnext5_norm_gelu"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)
    out = {}
    for k in z.files:
        v = z[k]
        if v.dtype.kind in "US":
            out[k] = str(v)
        else:
            out[k] = torch.from_numpy(v) if v.ndim > 0 else v.item()
    out["dyn_kwargs"] = eval(out["dyn_kwargs"], {"__builtins__": {}}, {})  # repr of a plain dict
    return out


@pytest.fixture(params=GOLDEN_CASES + GOLDEN_EXTRA)
def golden(request):
    return load_golden(request.param)


def random_events(num_events, lo, hi, F=7, seed=0, ties=True):
    """Random events with duplicated xyz rows (same-sensor pulses)."""
    g = torch.Generator().manual_seed(seed)
    counts = torch.randint(lo, hi + 1, (num_events,), generator=g)
    N = int(counts.sum())
    x = torch.randn(N, F, generator=g)
    if ties and N > 4:
        # every ~4th pulse copies the xyz of an earlier pulse of the same event
        batch = torch.repeat_interleave(torch.arange(num_events), counts)
        ptr = torch.zeros(num_events + 1, dtype=torch.long)
        ptr[1:] = counts.cumsum(0)
        src = torch.randint(0, 1 << 30, (N,), generator=g)
        for i in range(N):
            if i % 4 == 3:
                e = int(batch[i])
                j = int(ptr[e]) + int(src[i]) % max(1, i - int(ptr[e]))
                if j < i:
                    x[i, :3] = x[j, :3]
    batch = torch.repeat_interleave(torch.arange(num_events), counts)
    return x.contiguous(), batch, counts.to(torch.int32)


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp(min=1e-30))


def assert_same_graph_up_to_near_ties(ei, want, coords, what, coord_noise=2e-5, max_frac=0.02):
    """The dynamic graphs are built from latents that carry fp32 rounding noise (a different
    summation order than the CPU reference), so neighbours at (nearly) equal distance may swap.
    Same targets and degrees are required; where a source differs, its squared distance to the
    target - measured in the REFERENCE latents - must equal the reference neighbour's up to what
    a relative coordinate noise `coord_noise` can move it: |delta d^2| <= 2 d eps + eps^2 with
    eps = coord_noise * |c|max."""
    assert ei.shape == want.shape, what
    assert torch.equal(ei[1], want[1]), what
    diff = (ei[0] != want[0]).nonzero().flatten()
    if diff.numel() == 0:
        return
    assert diff.numel() <= max(2, int(max_frac * ei.shape[1])), f"{what}: {diff.numel()} edges differ"
    c = coords.double()
    tgt = want[1][diff]
    d_mine = ((c[ei[0][diff]] - c[tgt]) ** 2).sum(1)
    d_want = ((c[want[0][diff]] - c[tgt]) ** 2).sum(1)
    eps = coord_noise * c.abs().max()
    bound = 2 * torch.sqrt(torch.maximum(d_mine, d_want)) * eps * 3 ** 0.5 + 3 * eps ** 2
    assert bool(((d_mine - d_want).abs() <= bound).all()), \
        f"{what}: a differing neighbour is not a near-tie"

"""Generate the golden vectors under tests/golden/ (run in the build container).

    python tests/golden/make_golden.py

Needs /root/reference: it executes the reference's OWN DynEdge / task / loss
code (via oracle/refstub.py, third-party kernels replaced by oracle/ops.py) on
the reference's bundled test databases and on small adversarial inputs, and
stores inputs + outputs as .npz.  The GPU box has no /root/reference; the
tests there read only the .npz files written here.

Cases
-----
oscnext5    data/tests/sqlite/oscNext_genie_level7_v02 (5 events, 9-15 pulses,
            FEATURES.ICECUBE86, IceCube86 standardisation), k=8,
            DynEdge(7, pools=[min,max,mean,sum]) + EnergyReconstruction.
upgrade5    data/tests/sqlite/upgrade_genie_step4... (5 events, 85-167 pulses,
            up to 30 pulses on one sensor -> in-degree 9 rows), FEATURES.UPGRADE,
            DynEdge(14, pools=[min,max,mean]) (QUESO shape), k=8.
prometheus  data/examples/sqlite/prometheus (50 events, 3-99 pulses, F=4),
            DynEdge(4) ctor defaults except pools=[min,max,mean,sum]; direction head.
oscnext5_norm_gelu  the oscNext events through DynEdge(add_norm_layer=True,
            activation_layer="gelu") with 3-, 1- and 2-Linear EdgeConv MLPs.
c2_1024_out / c4_256_out  BASELINE configs 2 and 4 at the benchmarked batch size (synthetic
            events regenerated from the seed): outputs + per-event digests of the dynamic graphs.
adversarial synthetic: events with n<k+1, all points identical, 2-pulse events,
            ties straddling the k+1 cut; DynEdge(7) node-level (no pooling).
"""

from __future__ import annotations

import os
import sqlite3
import sys

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import refstub  # noqa: E402
from graphnet_b200.detector import (  # noqa: E402
    FEATURES_ICECUBE86, FEATURES_PROMETHEUS, FEATURES_UPGRADE,
    IceCube86, IceCubeUpgrade, Prometheus,
)

REF_DATA = "/root/reference/data"
OUT = os.path.dirname(os.path.abspath(__file__))


def load_db(path, table, features, detector, index="event_no"):
    con = sqlite3.connect(path)
    df = pd.read_sql(f"select * from {table}", con)
    xs, counts = [], []
    for _, g in df.groupby(index, sort=True):
        raw = torch.tensor(g[features].to_numpy(dtype=np.float64), dtype=torch.float32)
        xs.append(detector(raw, features))
        counts.append(len(g))
    x = torch.cat(xs)
    counts = torch.tensor(counts)
    batch = torch.repeat_interleave(torch.arange(len(counts)), counts)
    return x, batch, counts.to(torch.int32)


def initial_graph(x, batch, k, columns=(0, 1, 2)):
    """KNNEdges on each event separately (edges.py:76-84 via dataset), then
    Batch.from_data_list offsets."""
    B = int(batch.max()) + 1
    parts, off = [], 0
    for e in range(B):
        xe = x[batch == e]
        parts.append(oracle.knn_graph(xe[:, list(columns)], k, None) + off)
        off += xe.shape[0]
    return torch.cat(parts, dim=1)


def run_case(name, ref, x, batch, n_pulses, k, dyn_kwargs, head=None, seed=0):
    ei = initial_graph(x, batch, k)
    torch.manual_seed(seed)
    model = ref.DynEdge(x.shape[1], **dyn_kwargs)
    tasks = {}
    if head == "energy":
        t = ref.reconstruction.EnergyReconstruction(
            hidden_size=model.nb_outputs, loss_function=ref.loss_functions.LogCoshLoss(),
            target_labels="energy")
        tasks["EnergyReconstruction"] = t
    elif head == "direction":
        t = ref.reconstruction.DirectionReconstructionWithKappa(
            hidden_size=model.nb_outputs,
            loss_function=ref.loss_functions.VonMisesFisher3DLoss(), target_labels="direction")
        tasks["DirectionReconstructionWithKappa"] = t
    data = ref.Data(x=x, edge_index=ei, batch=batch, n_pulses=n_pulses)
    with torch.no_grad():
        y = model(data)
        # dynamic graphs, recomputed layer by layer exactly as DynEdgeConv does
        spec = oracle.DynEdgeSpec(nb_inputs=x.shape[1], **dyn_kwargs)
        sd = {k_: v.detach().clone() for k_, v in model.state_dict().items()}
        y2, inter = oracle.dynedge_forward(spec, sd, x, ei, batch, n_pulses,
                                           return_intermediates=True)
        assert torch.equal(y, y2), "oracle restatement differs from the reference code"
    out = {
        "x": x.numpy(), "batch": batch.numpy(), "n_pulses": n_pulses.numpy(),
        "edge_index": ei.numpy(), "k": np.int64(k), "seed": np.int64(seed),
        "out": y.numpy(), "global_variables": inter["global_variables"].numpy(),
        "post": inter["post"].numpy(),
    }
    for i in range(len(spec.dynedge_layer_sizes)):
        out[f"conv{i}"] = inter[f"conv{i}"].numpy()
        out[f"graph{i + 1}"] = inter["graphs"][i + 1].numpy()
    if "pooled" in inter:
        out["pooled"] = inter["pooled"].numpy()
    # parameters are NOT stored (5.5 MB each): tests rebuild them with
    # oracle.init_dynedge_state_dict(spec, seed) and check this digest.
    out["param_digest"] = np.array(
        [float(sum(v.double().sum() for v in sd.values())),
         float(sum(v.double().abs().sum() for v in sd.values()))])
    sd2 = oracle.init_dynedge_state_dict(spec, seed)
    assert all(torch.equal(sd[k_], sd2[k_]) for k_ in sd) and set(sd) == set(sd2)
    out["dyn_kwargs"] = np.array(repr(dyn_kwargs))
    for kind, t in tasks.items():
        with torch.no_grad():
            p = t(y)
        out["head_kind"] = np.array(kind)
        out["head_weight"] = t._affine.weight.detach().numpy()
        out["head_bias"] = t._affine.bias.detach().numpy()
        out["head_out"] = p.numpy()
        g = torch.Generator().manual_seed(99)
        if kind == "EnergyReconstruction":
            tgt = torch.rand(p.shape[0], 1, generator=g) * 4
            loss = ref.loss_functions.LogCoshLoss()(p, tgt)
        else:
            tgt = torch.nn.functional.normalize(torch.randn(p.shape[0], 3, generator=g), dim=1)
            loss = ref.loss_functions.VonMisesFisher3DLoss()(p, tgt)
        out["target"] = tgt.numpy()
        out["loss"] = loss.numpy()
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **out)
    deg = np.bincount(ei[1].numpy(), minlength=x.shape[0])
    print(f"{name}: N={x.shape[0]} B={len(n_pulses)} E={ei.shape[1]} "
          f"deg[min,max]=[{deg.min()},{deg.max()}] -> {os.path.getsize(path) / 1e3:.0f} kB")


def adversarial_inputs():
    g = torch.Generator().manual_seed(5)
    ev = []
    ev.append(torch.randn(2, 7, generator=g))                      # 2-pulse event
    ev.append(torch.randn(5, 7, generator=g))                      # n < k+1
    e = torch.randn(9, 7, generator=g); ev.append(e)               # n == k+1
    e = torch.randn(12, 7, generator=g); e[:, :3] = 0.25; ev.append(e)   # all xyz identical
    e = torch.randn(30, 7, generator=g); e[3:18, :3] = e[3, :3]; ev.append(e)  # 15 on one sensor
    e = torch.randn(20, 7, generator=g)                            # ties straddling the cut:
    e[:, 1:3] = 0.0; e[:, 0] = torch.arange(20) // 2 * 0.5         # pairs on a line, equal spacing
    ev.append(e)
    e = torch.randn(40, 7, generator=g); e[:, :3] = torch.round(e[:, :3] * 2) / 2; ev.append(e)  # lattice
    e = torch.randn(3, 7, generator=g); e[:, :4] = 1.0; ev.append(e)  # homophily = 1 on xyzt
    x = torch.cat(ev)
    counts = torch.tensor([t.shape[0] for t in ev])
    batch = torch.repeat_interleave(torch.arange(len(ev)), counts)
    return x.contiguous(), batch, counts.to(torch.int32)


def run_extra_cases(ref):
    """Options off the default path: LayerNorm + GELU (dynedge.py:161-165, :202-203, :224-226)
    and EdgeConv MLPs with one / three Linear layers, on the oscNext events."""
    x, b, n = load_db(
        f"{REF_DATA}/tests/sqlite/oscNext_genie_level7_v02/oscNext_genie_level7_v02_first_5_frames.db",
        "SRTInIcePulses", FEATURES_ICECUBE86, IceCube86())
    run_case("oscnext5_norm_gelu", ref, x, b, n, 8,
             dict(global_pooling_schemes=["min", "max", "mean", "sum"], add_norm_layer=True,
                  activation_layer="gelu",
                  dynedge_layer_sizes=[(48, 64, 96), (80,), (64, 64)],
                  post_processing_layer_sizes=[72, 48], readout_layer_sizes=[40]),
             None, seed=5)


def run_jinst_case(ref):
    """DynEdgeJINST (gnn/dynedge_jinst.py) on the oscNext events, from the reference's own file."""
    from oracle.dynedge_jinst import init_jinst_state_dict, jinst_forward
    x, b, n = load_db(
        f"{REF_DATA}/tests/sqlite/oscNext_genie_level7_v02/oscNext_genie_level7_v02_first_5_frames.db",
        "SRTInIcePulses", FEATURES_ICECUBE86, IceCube86())
    ei = initial_graph(x, b, 8)
    torch.manual_seed(6)
    model = ref.DynEdgeJINST(x.shape[1])
    sd = {k_: v.detach().clone() for k_, v in model.state_dict().items()}
    sd2 = init_jinst_state_dict(x.shape[1], 4, seed=6)
    assert set(sd) == set(sd2) and all(torch.equal(sd[k_], sd2[k_]) for k_ in sd)
    with torch.no_grad():
        y = model(ref.Data(x=x, edge_index=ei, batch=b, n_pulses=n))
        y2, inter = jinst_forward(sd, x, ei, b, n, return_intermediates=True)
    assert torch.equal(y, y2), "oracle restatement differs from the reference code"
    path = os.path.join(OUT, "oscnext5_jinst.npz")
    np.savez_compressed(path, x=x.numpy(), batch=b.numpy(), n_pulses=n.numpy(),
                        edge_index=ei.numpy(), seed=np.int64(6), out=y.numpy(),
                        conv0=inter["convs"][0].numpy(), pooled=inter["pooled"].numpy())
    print(f"oscnext5_jinst: N={x.shape[0]} -> {os.path.getsize(path) / 1e3:.0f} kB")


def run_particlenet_case(ref):
    """ParticleNeT (gnn/particlenet.py) in eval mode on the oscNext events, from the reference's own
    file: default architecture (3-Linear MLPs with BatchNorm1d, aggr="mean", k=16, dynamic graphs)
    and a static-graph, two-pooling, GELU variant without BatchNorm."""
    from oracle.particlenet import init_particlenet_state_dict, particlenet_forward
    x, b, n = load_db(
        f"{REF_DATA}/tests/sqlite/oscNext_genie_level7_v02/oscNext_genie_level7_v02_first_5_frames.db",
        "SRTInIcePulses", FEATURES_ICECUBE86, IceCube86())
    ei = initial_graph(x, b, 16)
    out = {}
    variants = {
        "a": dict(),
        "b": dict(dynamic=False, dynedge_layer_sizes=[(48, 40), (72,)], readout_layer_sizes=[64, 32],
                  global_pooling_schemes=["max", "sum"], activation_layer="gelu",
                  add_batchnorm_layer=False, nb_neighbours=8),
    }
    for tag, kw in variants.items():
        torch.manual_seed(7)
        model = ref.ParticleNeT(x.shape[1], **kw)
        sizes = kw.get("dynedge_layer_sizes", [(64, 64, 64), (128, 128, 128), (256, 256, 256)])
        ro = kw.get("readout_layer_sizes", [256])
        pools = kw.get("global_pooling_schemes", ["mean"])
        bn = kw.get("add_batchnorm_layer", True)
        sd = init_particlenet_state_dict(x.shape[1], sizes, ro, pools, bn, seed=7)
        assert set(sd) == set(model.state_dict())
        model.load_state_dict(sd)
        model.eval()
        with torch.no_grad():
            y = model(ref.Data(x=x, edge_index=ei, batch=b, n_pulses=n))
            y2, inter = particlenet_forward(
                sd, x, ei, b, sizes, ro, pools, kw.get("nb_neighbours", 16), slice(0, 3),
                kw.get("dynamic", True), bn, kw.get("activation_layer", "relu"),
                return_intermediates=True)
        assert torch.equal(y, y2), "oracle restatement differs from the reference code"
        out[f"out_{tag}"] = y.numpy()
        out[f"conv0_{tag}"] = inter["convs"][0].numpy()
        out[f"graph1_{tag}"] = inter["graphs"][1].numpy()
    path = os.path.join(OUT, "oscnext5_particlenet.npz")
    np.savez_compressed(path, x=x.numpy(), batch=b.numpy(), n_pulses=n.numpy(),
                        edge_index=ei.numpy(), seed=np.int64(7), **out)
    print(f"oscnext5_particlenet: N={x.shape[0]} -> {os.path.getsize(path) / 1e3:.0f} kB")


def graph_digest(ei, batch, ptr, num_events):
    """Order-independent per-event digest of an edge list (int64 arithmetic wraps the same way
    on CPU and GPU): events whose dynamic graphs equal the golden run's have equal digests."""
    src, dst = ei[0], ei[1]
    ev = batch[dst]
    key = (src - ptr[ev]) * 4096 + (dst - ptr[ev]) + 1
    h = key * 0x9E3779B97F4A7C1 + (key * key) * 0x85EBCA6B
    out = torch.zeros(num_events, dtype=torch.int64)
    out.index_add_(0, ev, h)
    return out


def run_fullsize_case(name, ref, events, pulses, detector, k, pools, seed=0, data_seed=5):
    """BASELINE configs at the BENCHMARKED batch size (C2: 1024 x ~60 pulses, C4: 256 x ~500):
    inputs are regenerated from the seed by the tests (graphnet_b200.synthetic), only the outputs
    and per-event digests of the three dynamic graphs are stored."""
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(events, pulses, detector=detector, seed=data_seed)
    x, batch, n_pulses = pb.x, pb.batch, pb.n_pulses
    ei = oracle.knn_graph(x[:, :3], k, batch)
    dyn_kwargs = dict(nb_neighbours=k, global_pooling_schemes=pools)
    torch.manual_seed(seed)
    model = ref.DynEdge(x.shape[1], **dyn_kwargs)
    spec = oracle.DynEdgeSpec(nb_inputs=x.shape[1], **dyn_kwargs)
    sd = {k_: v.detach().clone() for k_, v in model.state_dict().items()}
    sd2 = oracle.init_dynedge_state_dict(spec, seed)
    assert all(torch.equal(sd[k_], sd2[k_]) for k_ in sd) and set(sd) == set(sd2)
    with torch.no_grad():
        y = model(ref.Data(x=x, edge_index=ei, batch=batch, n_pulses=n_pulses))
        y2, inter = oracle.dynedge_forward(spec, sd, x, ei, batch, n_pulses,
                                           return_intermediates=True)
    assert torch.equal(y, y2), "oracle restatement differs from the reference code"
    ptr = pb.ptr
    dig = torch.stack([graph_digest(inter["graphs"][i], batch, ptr, events) for i in (1, 2, 3)])
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(
        path, out=y.numpy(), graph_digest=dig.numpy(), k=np.int64(k), seed=np.int64(seed),
        data_seed=np.int64(data_seed), events=np.int64(events), pulses=np.float64(pulses),
        detector=np.array(detector), dyn_kwargs=np.array(repr(dyn_kwargs)),
        input_digest=np.array([float(x.double().sum()), float(x.shape[0]), float(ei.shape[1]),
                               float(ei.double().sum())]))
    print(f"{name}: N={x.shape[0]} B={events} E={ei.shape[1]} -> {os.path.getsize(path) / 1e3:.0f} kB")


def run_fullsize_cases(ref):
    run_fullsize_case("c2_1024_out", ref, 1024, 60.0, "icecube86", 8, ["min", "max", "mean", "sum"])
    run_fullsize_case("c4_256_out", ref, 256, 500.0, "upgrade", 16, ["min", "max", "mean"], seed=1)


def main():
    ref = refstub.load_reference()
    pools4 = ["min", "max", "mean", "sum"]
    x, b, n = load_db(
        f"{REF_DATA}/tests/sqlite/oscNext_genie_level7_v02/oscNext_genie_level7_v02_first_5_frames.db",
        "SRTInIcePulses", FEATURES_ICECUBE86, IceCube86())
    run_case("oscnext5", ref, x, b, n, 8, dict(global_pooling_schemes=pools4), "energy")
    x, b, n = load_db(
        f"{REF_DATA}/tests/sqlite/upgrade_genie_step4_140028_000998_first_5_frames/"
        "upgrade_genie_step4_140028_000998_first_5_frames.db",
        "SplitInIcePulses", FEATURES_UPGRADE, IceCubeUpgrade())
    run_case("upgrade5", ref, x, b, n, 8,
             dict(global_pooling_schemes=["min", "max", "mean"]), "energy", seed=1)
    x, b, n = load_db(f"{REF_DATA}/examples/sqlite/prometheus/prometheus-events.db",
                      "total", FEATURES_PROMETHEUS, Prometheus())
    keep = n > 1  # collate_fn drops events with <= 1 pulse (data/dataloader.py:11-17)
    assert bool(keep.all())
    run_case("prometheus50", ref, x, b, n, 8, dict(global_pooling_schemes=pools4),
             "direction", seed=2)
    run_extra_cases(ref)
    run_jinst_case(ref)
    run_particlenet_case(ref)
    x, b, n = adversarial_inputs()
    run_case("adversarial", ref, x, b, n, 8, dict(), None, seed=3)
    run_fullsize_cases(ref)
    run_case("adversarial_k16_after", ref, x, b, n, 8,
             dict(nb_neighbours=16, global_pooling_schemes=["mean"],
                  add_global_variables_after_pooling=True,
                  dynedge_layer_sizes=[(64, 96), (80, 64)],
                  post_processing_layer_sizes=[72, 48], readout_layer_sizes=[40, 24]),
             None, seed=4)


if __name__ == "__main__":
    if "--extra-only" in sys.argv:
        _ref = refstub.load_reference()
        run_extra_cases(_ref)
        run_jinst_case(_ref)
    elif "--particlenet-only" in sys.argv:
        run_particlenet_case(refstub.load_reference())
    elif "--fullsize-only" in sys.argv:
        run_fullsize_cases(refstub.load_reference())
    else:
        main()

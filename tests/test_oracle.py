"""CPU: the oracle against the golden vectors recorded from the reference's own code."""
import torch

import oracle
from oracle import tasks as otasks
from oracle.ops import knn_graph_padded
from conftest import load_golden, random_events, rel_l2, assert_same_graph_up_to_near_ties


def _spec_and_params(g):
    spec = oracle.DynEdgeSpec(nb_inputs=g["x"].shape[1], **g["dyn_kwargs"])
    sd = oracle.init_dynedge_state_dict(spec, seed=int(g["seed"]))
    digest = torch.tensor([
        float(sum(v.double().sum() for v in sd.values())),
        float(sum(v.double().abs().sum() for v in sd.values()))], dtype=torch.float64)
    assert torch.allclose(digest, g["param_digest"].double(), rtol=1e-12), \
        "seeded parameters differ from the ones the golden outputs were recorded with"
    return spec, sd


def test_oracle_reproduces_golden_layers(golden):
    """Layer by layer on the GOLDEN inputs and the GOLDEN graph of each layer (no near-tie chaos):
    summation order inside the host's sgemm differs between CPUs, so values are compared to 1e-5
    relative; the graph rebuilt from the golden latents is integer work and must be bit-exact."""
    from oracle import ops as oops
    from oracle.dynedge import _mlp, global_variables
    g = golden
    spec, sd = _spec_and_params(g)
    gv = global_variables(g["x"], g["edge_index"], g["batch"], g["n_pulses"])
    assert torch.allclose(gv, g["global_variables"], rtol=1e-6, atol=1e-7)
    x = g["x"] if spec.add_global_variables_after_pooling else torch.cat([g["x"], gv[g["batch"]]], 1)
    graph = g["edge_index"]
    skip = [x]
    for i, sizes in enumerate(spec.dynedge_layer_sizes):
        nn_ = lambda t, i=i, n=len(sizes): _mlp(spec, sd, f"_conv_layers.{i}.nn", n, t,
                                                spec.add_norm_layer)
        out = oops.edge_conv(x, graph, nn_, aggr="add")
        assert rel_l2(out, g[f"conv{i}"]) < 1e-5, f"conv{i}"
        x = g[f"conv{i}"]
        graph = oops.knn_graph(x[:, spec.features_subset], spec.nb_neighbours, g["batch"])
        assert torch.equal(graph, g[f"graph{i + 1}"]), f"graph{i + 1}"
        skip.append(x)
    post = _mlp(spec, sd, "_post_processing", len(spec.post_processing_layer_sizes),
                torch.cat(skip, 1), spec.add_norm_layer)
    assert rel_l2(post, g["post"]) < 1e-5


def test_oracle_reproduces_golden_forward(golden):
    """Free-running forward.  The graphs of layers 2-4 are rebuilt from latents that carry this
    host's sgemm rounding (~1e-6), so a neighbour may differ from the recorded one only where it is
    a near-tie; outputs agree to 1e-5 when no neighbour flipped, to 2e-3 otherwise."""
    g = golden
    spec, sd = _spec_and_params(g)
    out, inter = oracle.dynedge_forward(spec, sd, g["x"], g["edge_index"], g["batch"],
                                        g["n_pulses"], return_intermediates=True)
    assert torch.allclose(inter["global_variables"], g["global_variables"], rtol=1e-6, atol=1e-7)
    flips = 0
    for i in range(len(spec.dynedge_layer_sizes)):
        tol = 1e-5 if flips == 0 else 2e-3
        assert rel_l2(inter[f"conv{i}"], g[f"conv{i}"]) < tol, f"conv{i}"
        if flips == 0:  # later graphs are compared in the reference latents only while those agree
            assert_same_graph_up_to_near_ties(inter["graphs"][i + 1], g[f"graph{i + 1}"],
                                              g[f"conv{i}"][:, :3], f"graph{i + 1}")
        flips += int((inter["graphs"][i + 1] != g[f"graph{i + 1}"]).sum()) \
            if inter["graphs"][i + 1].shape == g[f"graph{i + 1}"].shape else 1
    assert rel_l2(out, g["out"]) < (1e-5 if flips == 0 else 2e-3)


def test_oracle_initial_graph_matches_golden(golden):
    g = golden
    ei = oracle.knn_graph(g["x"][:, :3], int(g["k"]), g["batch"])
    assert torch.equal(ei, g["edge_index"])
    assert torch.equal(knn_graph_padded(g["x"][:, :3], int(g["k"]), g["batch"]), ei)


def test_dense_distribute_is_gather():
    g = load_golden("oscnext5")
    spec, sd = _spec_and_params(g)
    a = oracle.dynedge_forward(spec, sd, g["x"], g["edge_index"], g["batch"], g["n_pulses"])
    b = oracle.dynedge_forward(spec, sd, g["x"], g["edge_index"], g["batch"], g["n_pulses"],
                               dense_distribute=True)
    assert torch.equal(a, b)


def test_knn_semantics_small():
    # 4 points on a line, k=2: ties broken by index, self dropped, degree rules
    x = torch.tensor([[0.0], [1.0], [2.0], [3.0]])
    ei = oracle.knn_graph(x, 2)
    # query 0: (0,self) 1 2 -> 1,2 ; query 1: self, 0(d=1), 2(d=1) -> 0,2 ; query 2: self,1,3 ; q3: self,2,1
    assert ei.tolist() == [[1, 2, 0, 2, 1, 3, 2, 1], [0, 0, 1, 1, 2, 2, 3, 3]]
    # all identical: self falls out of the first k+1 for index >= k+1 -> degree k+1
    x = torch.zeros(6, 3)
    ei = oracle.knn_graph(x, 2)
    deg = torch.bincount(ei[1], minlength=6)
    assert deg.tolist() == [2, 2, 2, 3, 3, 3]
    # n < k+1
    ei = oracle.knn_graph(torch.randn(3, 3), 8)
    assert torch.bincount(ei[1], minlength=3).tolist() == [2, 2, 2]
    # loop=True keeps self
    ei = oracle.knn_graph(torch.randn(5, 3), 2, loop=True)
    assert (ei[0] == ei[1]).sum() == 5


def test_knn_padded_equals_loop_random():
    for seed in range(3):
        x, batch, _ = random_events(12, 2, 70, seed=seed)
        for k in (1, 8, 16):
            assert torch.equal(oracle.knn_graph(x[:, :3], k, batch),
                               knn_graph_padded(x[:, :3], k, batch))


def test_scatter_and_homophily_semantics():
    src = torch.tensor([[1.0, -2.0], [3.0, 4.0], [5.0, -6.0]])
    idx = torch.tensor([0, 0, 2])
    assert oracle.scatter_sum(src, idx).tolist() == [[4.0, 2.0], [0.0, 0.0], [5.0, -6.0]]
    assert oracle.scatter_mean(src, idx).tolist() == [[2.0, 1.0], [0.0, 0.0], [5.0, -6.0]]
    mn, amn = oracle.scatter_min(src, idx)
    mx, amx = oracle.scatter_max(src, idx)
    assert mn.tolist() == [[1.0, -2.0], [0.0, 0.0], [5.0, -6.0]]
    assert mx.tolist() == [[3.0, 4.0], [0.0, 0.0], [5.0, -6.0]]
    assert amn.tolist() == [[0, 0], [3, 3], [2, 2]] and amx.tolist() == [[1, 1], [3, 3], [2, 2]]
    ei = torch.tensor([[1, 0, 3], [0, 1, 2]])
    y = torch.tensor([1.0, 1.0, 2.0, 3.0])
    batch = torch.tensor([0, 0, 1, 1])
    assert oracle.homophily(ei, y, batch).tolist() == [1.0, 0.0]


def test_heads_and_losses_match_golden():
    for name in ("oscnext5", "upgrade5", "prometheus50"):
        g = load_golden(name)
        p = otasks.task_forward(g["head_kind"], g["head_weight"], g["head_bias"], g["out"])
        assert torch.allclose(p, g["head_out"], rtol=1e-5, atol=1e-6)  # host sgemm order varies
        if g["head_kind"] == "EnergyReconstruction":
            loss = otasks.logcosh_loss(g["head_out"], g["target"])
        else:
            loss = otasks.vmf3d_loss(g["head_out"], g["target"])
        assert torch.allclose(loss, torch.as_tensor(g["loss"]), rtol=1e-5)


def test_vmf_m3_closed_form_is_the_reference_test_formula():
    # reference tests/training/test_loss_functions.py:71-101 checks LogCMK(m=3) against
    # log k - k - log(2 pi (1 - exp(-2k))) ; scipy is available, so check the Bessel form too
    import numpy as np
    import scipy.special
    k = torch.tensor(np.logspace(-3, 2, 20))
    bessel = 0.5 * torch.log(k) - torch.log(torch.from_numpy(scipy.special.iv(0.5, k.numpy()))) \
        - 1.5 * np.log(2 * np.pi)
    assert torch.allclose(otasks.log_cmk_m3_exact(k), bessel, rtol=1e-9, atol=1e-9)


def test_jinst_oracle_reproduces_reference_golden():
    """oracle/dynedge_jinst.py against the output of the reference's own dynedge_jinst.py."""
    import numpy as np, os
    from oracle.dynedge_jinst import init_jinst_state_dict, jinst_forward
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "oscnext5_jinst.npz"))
    x, b, n, ei = (torch.from_numpy(z[k]) for k in ("x", "batch", "n_pulses", "edge_index"))
    sd = init_jinst_state_dict(x.shape[1], 4, seed=int(z["seed"]))
    y = jinst_forward(sd, x, ei, b, n)
    # bit-equal on the host that recorded it; another CPU's sgemm sums in a different order
    assert rel_l2(y, torch.from_numpy(z["out"])) < 1e-5


def test_particlenet_oracle_reproduces_reference_golden():
    """oracle/particlenet.py against the output of the reference's own particlenet.py (eval mode):
    the default architecture and a static-graph / GELU / no-BatchNorm variant."""
    import numpy as np, os
    from oracle.particlenet import init_particlenet_state_dict, particlenet_forward
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "oscnext5_particlenet.npz"))
    x, b, ei = (torch.from_numpy(z[k]) for k in ("x", "batch", "edge_index"))
    sd = init_particlenet_state_dict(x.shape[1], seed=int(z["seed"]))
    y, inter = particlenet_forward(sd, x, ei, b, return_intermediates=True)
    assert rel_l2(inter["convs"][0], torch.from_numpy(z["conv0_a"])) < 1e-5
    assert rel_l2(y, torch.from_numpy(z["out_a"])) < 1e-4   # dynamic graphs: near-ties may flip
    sizes, ro, pools = [(48, 40), (72,)], [64, 32], ["max", "sum"]
    sd = init_particlenet_state_dict(x.shape[1], sizes, ro, pools, False, seed=int(z["seed"]))
    y = particlenet_forward(sd, x, ei, b, sizes, ro, pools, 8, slice(0, 3), False, False, "gelu")
    assert rel_l2(y, torch.from_numpy(z["out_b"])) < 1e-5

"""Gradients of the 16-bit training path against torch autograd (fp32) on the same graphs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _ell_to_pairs(graph, N):
    """(target i, source j) of every filled slot of an ELL / CSR graph."""
    if graph.ell_stride > 0:
        nbr = graph.col.reshape(N, graph.ell_stride).long()
        tgt = torch.arange(N, device=nbr.device)[:, None].expand_as(nbr)
        m = nbr >= 0
        return tgt[m], nbr[m]
    rp = graph.rowptr.long()
    deg = rp[1:] - rp[:-1]
    tgt = torch.repeat_interleave(torch.arange(N, device=rp.device), deg)
    return tgt, graph.col.long()


def _first_arg_pool(x, ptr, kind):
    outs = []
    p = ptr.tolist()
    for e in range(len(p) - 1):
        xs = x[p[e]:p[e + 1]]
        v = xs.min(0).values if kind == "min" else xs.max(0).values
        idx = (xs == v).float().argmax(0)            # first row attaining the value
        outs.append(xs.gather(0, idx[None])[0])
    return torch.stack(outs)


def _reference_forward(model, x0, layers, ptr, schemes, gv_after=None):
    """fp32 torch restatement of dynedge.py:296-350 on fixed graphs (autograd enabled)."""
    N = x0.shape[0]
    feats = [x0]
    x = x0
    for conv, L in zip(model._conv_layers, layers):
        tgt, src = _ell_to_pairs(L["graph"], N)
        msg = conv.nn(torch.cat([x[tgt], x[src] - x[tgt]], 1))
        x = torch.zeros((N, msg.shape[1]), device=x.device).index_add_(0, tgt, msg)
        feats.append(x)
    h = model._post_processing(torch.cat(feats, 1))
    pooled = []
    p = ptr.long()
    cnt = (p[1:] - p[:-1]).float()[:, None]
    seg = torch.repeat_interleave(torch.arange(len(p) - 1, device=h.device), p[1:] - p[:-1])
    for s in schemes:
        if s in ("min", "max"):
            pooled.append(_first_arg_pool(h, p, s))
        else:
            tot = torch.zeros((len(p) - 1, h.shape[1]), device=h.device).index_add_(0, seg, h)
            pooled.append(tot / cnt if s == "mean" else tot)
    if gv_after is not None:      # add_global_variables_after_pooling (dynedge.py:339-345)
        pooled.append(gv_after)
    return model._readout(torch.cat(pooled, 1))


@pytest.mark.parametrize("events,pulses,after", [(24, 18.0, False), (64, 40.0, False),
                                                 (24, 18.0, True)])
def test_gradients_match_autograd(dev, events, pulses, after):
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    from graphnet_b200.training import _DynEdgeTrain, _param_list

    schemes = ["min", "max", "mean", "sum"]
    pb = make_pulse_batch(events, pulses, seed=11)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=events)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev))
    torch.manual_seed(3)
    model = g.DynEdge(7, global_pooling_schemes=schemes, precision="bf16",
                      add_global_variables_after_pooling=after).to(dev)
    params = _param_list(model)

    out = g.training.dynedge_train_forward(model, data)
    w = torch.randn(out.shape, generator=torch.Generator(device=dev).manual_seed(5), device=dev)
    loss = (out * w).sum() / events
    loss.backward()
    got = [p.grad.detach().clone() for p in params]
    saved = out.grad_fn.saved if hasattr(out.grad_fn, "saved") else None
    assert saved is not None

    # reference: fp32 autograd on the graphs the forward actually used
    for p in params:
        p.grad = None
    gv = saved_gv(model, data)
    x0 = x if after else torch.cat([x, gv[batch]], 1)
    ref_out = _reference_forward(model, x0, saved["layers"], saved["ptr"], schemes,
                                 gv_after=gv if after else None)
    ((ref_out * w).sum() / events).backward()
    ref = [p.grad.detach().clone() for p in params]

    rel_out = float((out.detach() - ref_out.detach()).norm() / ref_out.detach().norm())
    assert rel_out < 2e-2, rel_out
    rels = [float((a - b).norm() / b.norm().clamp_min(1e-12)) for a, b in zip(got, ref)]
    print("relative gradient errors", [(tuple(p.shape), round(r, 4)) for p, r in zip(params, rels)])
    # The 16-bit forward differs from the fp32 one by ~1 % and flips the ReLU mask of
    # near-zero activations; with a random linear loss every flipped unit changes the gradient by
    # O(1) of its share.  The error is the same at every depth - including the readout bias,
    # whose gradient is a plain masked sum in fp32 - so it measures the forward's rounding, not
    # the backward kernels (those are compared exactly in test_gpu_backward_ops.py).
    for p, a, b, rel in zip(params, got, ref, rels):
        cos = float((a * b).sum() / (a.norm() * b.norm()).clamp_min(1e-30))
        assert rel < 0.12 and cos > 0.99, (tuple(p.shape), rel, cos)


def saved_gv(model, data):
    import graphnet_b200.ops as ops
    x, n_pulses, ptr, graph, N, B = model._graph_inputs(data, None)
    return ops.global_variables(x, graph, ptr, n_pulses)


def test_optimizer_step_reduces_loss(dev):
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(64, 30.0, seed=21)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=64)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev))
    torch.manual_seed(0)
    model = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16").to(dev)
    head = torch.nn.Linear(128, 1).to(dev)
    target = torch.rand(64, 1, device=dev) * 4
    opt = torch.optim.Adam(list(model.parameters()) + list(head.parameters()), lr=1e-3, eps=1e-3)
    losses = []
    for _ in range(12):
        opt.zero_grad(set_to_none=True)
        pred = head(g.training.dynedge_train_forward(model, data))
        d = pred - target
        loss = (d + torch.nn.functional.softplus(-2.0 * d) - 0.6931471805599453).mean()  # log-cosh
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0], losses


def test_standard_model_training_step(dev):
    """StandardModel.training_step -> backward -> Adam, through the public API."""
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(48, 25.0, seed=31)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=48)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev),
                  energy=torch.rand(48, device=dev) * 3 + 0.5)
    torch.manual_seed(1)
    backbone = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16")
    task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                  loss_function=g.LogCoshLoss())
    model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone,
                            optimizer_kwargs={"lr": 1e-3, "eps": 1e-3}).to(dev)
    model.train()
    opt = model.configure_optimizers()["optimizer"]
    losses = []
    for _ in range(10):
        opt.zero_grad(set_to_none=True)
        loss = model.training_step(data, 0)
        loss.backward()
        assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0], losses
    # inference afterwards still runs the kernel heads and carries no graph
    model.train(False)
    model.inference()
    with torch.no_grad():
        pred = model(data)[0]
    assert pred.shape == (48, 1) and not pred.requires_grad


def test_training_step_c4_shapes(dev):
    """Upgrade geometry, F = 14, k = 16 (17 slots per node), pools min/max/mean: the backward path
    on the shapes of BASELINE config 4."""
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(6, 120.0, detector="upgrade", seed=33)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 16, batch, num_events=6)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev))
    torch.manual_seed(2)
    model = g.DynEdge(14, nb_neighbours=16, global_pooling_schemes=["min", "max", "mean"],
                      precision="bf16").to(dev)
    head = torch.nn.Linear(128, 1).to(dev)
    target = torch.rand(6, 1, device=dev) * 4
    opt = torch.optim.Adam(list(model.parameters()) + list(head.parameters()), lr=1e-3, eps=1e-3)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        d = head(model(data)) - target
        loss = (d + torch.nn.functional.softplus(-2.0 * d) - 0.6931471805599453).mean()
        loss.backward()
        assert all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in model.parameters())
        opt.step()
        losses.append(float(loss.detach()))
    assert losses[-1] < losses[0], losses


@pytest.mark.parametrize("kwargs,F", [
    (dict(global_pooling_schemes=["max"], readout_layer_sizes=[64, 32]), 4),
    (dict(global_pooling_schemes=None), 7),
    (dict(global_pooling_schemes=["mean", "sum"], dynedge_layer_sizes=[(64, 96), (80, 64)],
          post_processing_layer_sizes=[72, 48], readout_layer_sizes=[40, 24]), 7),
])
def test_training_other_architectures(dev, kwargs, F):
    """Architectures off the default: one pooling scheme, node-level output, other layer sizes."""
    import graphnet_b200 as g
    pb_x = torch.randn(300, F, generator=torch.Generator().manual_seed(1))
    counts = torch.tensor([40, 25, 60, 35, 50, 90])
    batch = torch.repeat_interleave(torch.arange(6), counts)
    x = pb_x.to(dev)
    b = batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, b, num_events=6)
    data = g.Data(x=x, edge_index=ei, batch=b, n_pulses=counts.to(torch.int32).to(dev))
    torch.manual_seed(3)
    model = g.DynEdge(F, precision="bf16", **kwargs).to(dev)
    out = model(data)
    rows = 300 if kwargs["global_pooling_schemes"] is None else 6
    assert out.shape[0] == rows and out.requires_grad
    head = torch.nn.Linear(out.shape[1], 1).to(dev)
    target = torch.rand(rows, 1, device=dev)
    opt = torch.optim.Adam(list(model.parameters()) + list(head.parameters()), lr=1e-3, eps=1e-3)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        loss = (head(model(data)) - target).square().mean()
        loss.backward()
        assert all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in model.parameters())
        opt.step()
        losses.append(float(loss.detach()))
    assert losses[-1] < losses[0], losses

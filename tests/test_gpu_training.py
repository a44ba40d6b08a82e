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


def _relu_bits(z_pos, S):
    """[N, S, b] bool -> the [8, ld] int32 ReLU-bit buffer of gn_tc_edge_mlp_aggregate_mask: bit c
    of word (g, r) = z(r, 32 g + c) > 0, slot e = i*S + t sits in row r = (e / used) * 64 + e % used,
    used = (64 / S) * S (include/graphnet_b200.h)."""
    import graphnet_b200.ops as ops
    N, _, b = z_pos.shape
    mask = ops.edge_relu_mask(N, S, z_pos.device)
    mask.zero_()
    used = (64 // S) * S
    e = torch.arange(N * S, device=z_pos.device)
    r = (e // used) * 64 + e % used
    zb = z_pos.reshape(N * S, b)
    w = (1 << torch.arange(32, device=z_pos.device, dtype=torch.int64))
    for gidx in range((b + 31) // 32):
        blk = zb[:, 32 * gidx:32 * gidx + 32].to(torch.int64)
        word = (blk * w[:blk.shape[1]]).sum(1)
        word = torch.where(word >= 2 ** 31, word - 2 ** 32, word).to(torch.int32)
        mask[gidx, r] = word
    return mask


def test_backward_chain_teacher_forced(dev):
    """The backward kernels alone, without the 16-bit forward's rounding in the loop: every tensor
    the backward pass reads (layer inputs, P, Q, ReLU bits of both Linear layers, post-processing
    and readout activations, pooled values) is replaced by its value from an fp32 forward on the
    same graphs, rounded to the storage type.  What is left is the bf16 arithmetic of the backward
    GEMMs themselves: gradients must agree with fp32 autograd to 1 % - against 3-8 % for the free
    running step (test_gradients_match_autograd), which shows that gap to be the ReLU-mask flips
    of the 16-bit forward and not the backward chain."""
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    from graphnet_b200.training import _backward, _param_list

    schemes = ["min", "max", "mean", "sum"]
    events = 48
    pb = make_pulse_batch(events, 30.0, seed=13)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=events)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev))
    torch.manual_seed(4)
    model = g.DynEdge(7, global_pooling_schemes=schemes, precision="bf16").to(dev)
    params = _param_list(model)
    out = g.training.dynedge_train_forward(model, data)
    sv = out.grad_fn.saved
    N = sv["N"]
    gv = saved_gv(model, data)
    x0 = torch.cat([x, gv[batch]], 1)

    # ---- fp32 reference forward on the forward's graphs, with autograd, keeping intermediates
    feats = [x0]
    xin = x0
    skip = sv["skip"]
    for li, (conv, L) in enumerate(zip(model._conv_layers, sv["layers"])):
        graph, pk = L["graph"], L["pack"]
        S = graph.ell_stride
        assert S > 0
        l1, l2 = conv.nn[0], conv.nn[2]
        d = l1.weight.shape[1] // 2
        Wa, Wb = l1.weight[:, :d], l1.weight[:, d:]
        P = xin @ (Wa - Wb).t() + l1.bias
        Q = xin @ Wb.t()
        nbr = graph.col.reshape(N, S).long()
        valid = nbr >= 0
        h = torch.relu(P[:, None, :] + Q[nbr.clamp(min=0)])          # [N, S, a]
        z = h @ l2.weight.t() + l2.bias                                # [N, S, b]
        msg = torch.relu(z) * valid[:, :, None]
        xout = msg.sum(1)
        # teacher forcing: what the backward reads for this layer
        with torch.no_grad():
            a = l1.weight.shape[0]
            L["P"].zero_()
            L["Q"].zero_()
            L["P"][:, :a] = P.to(torch.float16)
            L["P"][:, a:a + 2] = 1.0
            L["Q"][:, :a] = Q.to(torch.float16)
            L["relu_mask"].copy_(_relu_bits((z > 0) & valid[:, :, None], S))
            skip[:, L["off_out"]:L["off_out"] + xout.shape[1]] = xout.to(skip.dtype)
        feats.append(xout)
        xin = xout
    hcat = torch.cat(feats, 1)
    post_lin = [m for m in model._post_processing if isinstance(m, torch.nn.Linear)]
    hpost = hcat
    for i, lin in enumerate(post_lin):
        hpost = torch.relu(lin(hpost))
        with torch.no_grad():
            dst = sv["post"][i]
            dst[:, :hpost.shape[1]] = hpost.to(dst.dtype)
    p = sv["ptr"].long()
    cnt = (p[1:] - p[:-1]).float()[:, None]
    seg = torch.repeat_interleave(torch.arange(len(p) - 1, device=dev), p[1:] - p[:-1])
    pooled = []
    for s in schemes:
        if s in ("min", "max"):
            pooled.append(_first_arg_pool(hpost, p, s))
        else:
            tot = torch.zeros((len(p) - 1, hpost.shape[1]), device=dev).index_add_(0, seg, hpost)
            pooled.append(tot / cnt if s == "mean" else tot)
    pooled = torch.cat(pooled, 1)
    ro_lin = [m for m in model._readout if isinstance(m, torch.nn.Linear)]
    hro = pooled
    new_ro = []
    for (lin, hin, hout, act), rl in zip(sv["readout"], ro_lin):
        hnext = torch.relu(rl(hro)) if act == "relu" else rl(hro)
        new_ro.append((lin, hro.detach().clone(), hnext.detach().clone(), act))
        hro = hnext
    with torch.no_grad():
        sv["pooled"].copy_(pooled)
        sv["readout"] = new_ro
    ref_out = hro
    w = torch.randn(ref_out.shape, generator=torch.Generator(device=dev).manual_seed(5), device=dev)
    for q in params:
        q.grad = None
    ((ref_out * w).sum() / events).backward()
    ref = [q.grad.detach().clone() for q in params]
    with torch.no_grad():
        got = _backward(model, sv, (w / events).contiguous())
    rels = [float((a_ - b_).norm() / b_.norm().clamp_min(1e-12)) for a_, b_ in zip(got, ref)]
    print("teacher-forced relative gradient errors",
          [(tuple(q.shape), round(r, 5)) for q, r in zip(params, rels)])
    for q, a_, b_, rel in zip(params, got, ref, rels):
        cos = float((a_ * b_).sum() / (a_.norm() * b_.norm()).clamp_min(1e-30))
        assert rel < 1e-2 and cos > 0.9999, (tuple(q.shape), rel, cos)


def test_loss_curve_follows_fp32_autograd(dev):
    """Twenty Adam steps on one batch: the loss of the 16-bit training path must follow the loss of
    an fp32 torch-autograd replica (same initial weights, same optimiser, graphs rebuilt by each
    from its own latents) within a few per cent at every step."""
    import copy
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch

    schemes = ["min", "max", "mean", "sum"]
    events = 64
    pb = make_pulse_batch(events, 30.0, seed=17)
    x = pb.x.to(dev)
    batch = pb.batch.to(dev)
    ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=events)
    data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=pb.n_pulses.to(dev))
    torch.manual_seed(6)
    model = g.DynEdge(7, global_pooling_schemes=schemes, precision="bf16").to(dev)
    head = torch.nn.Linear(128, 1).to(dev)
    ref_model, ref_head = copy.deepcopy(model), copy.deepcopy(head)
    target = torch.rand(events, 1, generator=torch.Generator().manual_seed(1)).to(dev) * 4
    gv = saved_gv(model, data)
    x0 = torch.cat([x, gv[batch]], 1)

    def logcosh(pred):
        d = pred - target
        return (d + torch.nn.functional.softplus(-2.0 * d) - 0.6931471805599453).mean()

    def ref_forward():
        # fp32 autograd replica of dynedge.py:296-350 with the library's own kNN for the graphs
        feats, xin = [x0], x0
        edges = ei
        for li, conv in enumerate(ref_model._conv_layers):
            src, dst = edges[0], edges[1]
            msg = conv.nn(torch.cat([xin[dst], xin[src] - xin[dst]], 1))
            xout = torch.zeros((x0.shape[0], msg.shape[1]), device=dev).index_add_(0, dst, msg)
            feats.append(xout)
            edges = g.knn_graph(xout[:, :3].detach().contiguous(), 8, batch, num_events=events)
            xin = xout
        h = ref_model._post_processing(torch.cat(feats, 1))
        p = torch.zeros(events + 1, dtype=torch.long, device=dev)
        p[1:] = torch.bincount(batch, minlength=events).cumsum(0)
        seg = batch
        cnt = (p[1:] - p[:-1]).float()[:, None]
        pooled = []
        for s in schemes:
            if s in ("min", "max"):
                pooled.append(_first_arg_pool(h, p, s))
            else:
                tot = torch.zeros((events, h.shape[1]), device=dev).index_add_(0, seg, h)
                pooled.append(tot / cnt if s == "mean" else tot)
        return ref_head(ref_model._readout(torch.cat(pooled, 1)))

    opt = torch.optim.Adam(list(model.parameters()) + list(head.parameters()), lr=1e-3, eps=1e-3)
    ropt = torch.optim.Adam(list(ref_model.parameters()) + list(ref_head.parameters()), lr=1e-3, eps=1e-3)
    ours, theirs = [], []
    for _ in range(20):
        opt.zero_grad(set_to_none=True)
        loss = logcosh(head(g.training.dynedge_train_forward(model, data)))
        loss.backward()
        opt.step()
        ours.append(float(loss.detach()))
        ropt.zero_grad(set_to_none=True)
        rloss = logcosh(ref_forward())
        rloss.backward()
        ropt.step()
        theirs.append(float(rloss.detach()))
    print("loss curves", [round(v, 4) for v in ours], [round(v, 4) for v in theirs])
    assert theirs[-1] < theirs[0] and ours[-1] < ours[0]
    # the two runs take different rounding paths (16-bit forward, atomically accumulated split-K
    # gradients: not bit-reproducible run to run), and Adam amplifies the difference step by step
    # on a loss that moves by factors per step: 5 % over the first five steps, 25 % at any later
    # one, 8 % on average (observed run to run: up to 13 % at single steps beyond the tenth)
    rel = [abs(a_ - b_) / (abs(b_) + 1e-3) for a_, b_ in zip(ours, theirs)]
    assert max(rel[:5]) <= 0.05 and max(rel) <= 0.25, (rel, ours, theirs)
    assert sum(rel) / len(rel) <= 0.08, (rel, ours, theirs)

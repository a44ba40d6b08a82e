"""CPU: host-side mirror of the reference interface (no compute)."""
import numpy as np
import pytest
import torch

import graphnet_b200 as g
import oracle
from graphnet_b200.synthetic import make_pulse_batch
from graphnet_b200.detector import FEATURES_ICECUBE86


@pytest.mark.parametrize("kwargs", [
    dict(),
    dict(global_pooling_schemes=["min", "max", "mean", "sum"]),
    dict(global_pooling_schemes="mean", add_global_variables_after_pooling=True,
         dynedge_layer_sizes=[(64, 96), (80, 64)], post_processing_layer_sizes=[72, 48],
         readout_layer_sizes=[40, 24]),
    dict(add_norm_layer=True, activation_layer="gelu"),
])
def test_state_dict_matches_reference_layout(kwargs):
    spec = oracle.DynEdgeSpec(nb_inputs=7, **kwargs)
    sd = oracle.init_dynedge_state_dict(spec, seed=3)
    torch.manual_seed(3)
    m = g.DynEdge(7, **kwargs)
    msd = m.state_dict()
    assert set(msd) == set(sd)
    for k in sd:
        assert msd[k].shape == sd[k].shape
        assert torch.equal(msd[k], sd[k]), k  # same construction order -> same seeded init
    m.load_state_dict(sd)
    assert m.nb_inputs == 7 and m.nb_outputs == spec.nb_outputs


def test_constructor_checks_follow_reference():
    with pytest.raises(AssertionError):
        g.DynEdge(7, global_pooling_schemes=["median"])
    with pytest.raises(AssertionError):
        g.DynEdge(7, add_global_variables_after_pooling=True)
    with pytest.raises(ValueError):
        g.DynEdge(7, activation_layer="tanh")
    with pytest.raises(AssertionError):
        g.DynEdge(7, dynedge_layer_sizes=[[128, 256]])


def test_batch_from_data_list():
    a = g.Data(x=torch.randn(3, 4), edge_index=torch.tensor([[1, 2], [0, 0]]),
               n_pulses=torch.tensor(3, dtype=torch.int32), energy=torch.tensor(1.5))
    b = g.Data(x=torch.randn(2, 4), edge_index=torch.tensor([[1], [0]]),
               n_pulses=torch.tensor(2, dtype=torch.int32), energy=torch.tensor(2.5))
    bt = g.Batch.from_data_list([a, b])
    assert bt.x.shape == (5, 4)
    assert bt.edge_index.tolist() == [[1, 2, 4], [0, 0, 3]]
    assert bt.batch.tolist() == [0, 0, 0, 1, 1]
    assert bt.n_pulses.tolist() == [3, 2] and bt.n_pulses.dtype == torch.int32
    assert bt.energy.tolist() == [1.5, 2.5] and bt.ptr.tolist() == [0, 3, 5]


def test_detector_standardisation():
    raw = torch.tensor([[250.0, -500.0, 100.0, 1.0e4, 10.0, 1.35, 0.0444]])
    x = g.IceCube86()(raw.clone(), FEATURES_ICECUBE86)
    assert torch.allclose(x, torch.tensor([[0.5, -1.0, 0.2, 0.0, 1.0, 0.4, 0.888]]), atol=1e-6)
    with pytest.raises(KeyError):
        g.IceCube86()(raw.clone(), ["bogus"] * 7)


def test_synthetic_batch_is_deterministic_and_has_ties():
    a = make_pulse_batch(32, 60.0, seed=5)
    b = make_pulse_batch(32, 60.0, seed=5)
    assert torch.equal(a.x, b.x) and torch.equal(a.batch, b.batch)
    assert a.x.shape[1] == 7 and a.num_events == 32
    assert int(a.n_pulses.sum()) == a.x.shape[0] and int(a.n_pulses.min()) >= 2
    # same-sensor pulses share xyz exactly
    xyz = a.x[:, :3]
    uniq = torch.unique(torch.cat([xyz, a.batch.unsqueeze(1).float()], 1), dim=0).shape[0]
    assert uniq < 0.95 * xyz.shape[0]
    u = make_pulse_batch(4, 300.0, detector="upgrade", seed=1)
    assert u.x.shape[1] == 14


def test_task_label_plumbing():
    t = g.EnergyReconstruction(hidden_size=128, target_labels="energy",
                               loss_function=g.LogCoshLoss(),
                               transform_prediction_and_target=lambda x: torch.log10(x))
    assert t._target_labels == ["energy"] and t._prediction_labels == ["energy_pred"]
    assert t._affine.weight.shape == (1, 128)
    d = g.DirectionReconstructionWithKappa(hidden_size=128, loss_function=g.VonMisesFisher3DLoss())
    assert d._affine.weight.shape == (3, 128) and len(d._prediction_labels) == 4
    with pytest.raises(AssertionError):
        g.EnergyReconstruction(hidden_size=8, transform_target=torch.log10,
                               transform_inference=lambda x: torch.pow(9.0, x))


def test_losses_match_oracle_on_cpu():
    from oracle import tasks as ot
    torch.manual_seed(0)
    pred = torch.randn(50, 1) * 3
    tgt = torch.randn(50, 1)
    assert torch.allclose(g.LogCoshLoss()(pred, tgt), ot.logcosh_loss(pred, tgt))
    v = torch.nn.functional.normalize(torch.randn(50, 3), dim=1)
    kappa = torch.rand(50, 1) * 150 + 1e-3
    p = torch.cat([v, kappa], 1)
    t = torch.nn.functional.normalize(torch.randn(50, 3), dim=1)
    assert torch.allclose(g.VonMisesFisher3DLoss()(p, t), ot.vmf3d_loss(p, t), rtol=1e-5)


def test_train_mode_forward_without_backward_raises():
    """precision='fp32' / 'fp16' (and DynEdgeJINST) have no backward pass: a train-mode forward with
    autograd on must fail loudly instead of returning an output without grad_fn (an optimizer
    would then update the task heads over a silently frozen backbone)."""
    import graphnet_b200 as g
    d = g.Data(x=torch.randn(6, 7), edge_index=torch.zeros(2, 0, dtype=torch.long),
               batch=torch.zeros(6, dtype=torch.long), n_pulses=torch.tensor([6], dtype=torch.int32))
    for prec in ("fp32", "fp16"):
        m = g.DynEdge(7, precision=prec)
        with pytest.raises(NotImplementedError, match="no backward"):
            m(d)
        m.eval()
        with pytest.raises(RuntimeError, match="CUDA"):   # past the mode check: no CPU fallback
            m(d)
    j = g.DynEdgeJINST(7)
    with pytest.raises(NotImplementedError, match="no backward"):
        j(d)

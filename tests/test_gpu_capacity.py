"""One captured training step for every batch that fits (training.BatchCapacity / pad_arrays_to_capacity): the
model.fit loop of the reference's force_*.py scripts sees a different pair / triple count in every batch."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _painn(depth=3):
    from gcnn_keras_b200.literature.PAiNN import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import PAINN_C3_CONFIG
    cfg = dict(PAINN_C3_CONFIG, depth=depth)
    torch.manual_seed(7)
    return EnergyForceModel(model_energy=make_model(**cfg), energy_output=0, output_to_tensor=True,
                            output_as_dict=False, output_squeeze_states=True, is_physical_force=False)


def _to_device(arrays, dev):
    return {k: torch.from_numpy(v).to(dev) for k, v in arrays.items()}


def _flat(step):
    return step.optimizer.flat.detach().clone()


def test_twenty_different_c3_batches_through_one_captured_graph_match_eager_bitwise():
    from gcnn_keras_b200.plan import clear_plan_cache
    from gcnn_keras_b200.synthetic import painn_c3_batch
    from gcnn_keras_b200.training import (BatchCapacity, CapturedForceTrainStep, ForceTrainStep, default_model_inputs,
                                          molecule_batch_to_arrays, pad_arrays_to_capacity)
    dev = torch.device("cuda:0")
    batches = [painn_c3_batch(12 + (s % 5), seed=s) for s in range(20)]     # 12..16 molecules, different pair counts
    assert len({int(b["edge_len"].sum()) for b in batches}) > 10
    cap = BatchCapacity.covering(batches, slack=0.02)
    max_nodes = max(int(b["node_len"].max()) for b in batches)

    model_a = _painn().to(dev)
    # the layers build lazily: one forward before copying so that the copies start from the same weights
    warm = _to_device(pad_arrays_to_capacity(molecule_batch_to_arrays(batches[0]), cap), dev)
    model_a(default_model_inputs(warm, max_nodes)[0], training=True)
    model_b = copy.deepcopy(model_a)
    model_c = copy.deepcopy(model_a)
    mk = lambda m: ForceTrainStep(m, lr=1e-3, force_loss_factor=95.0, clipnorm=1.0)
    step_a, step_b, step_c = mk(model_a), mk(model_b), mk(model_c)

    captured = CapturedForceTrainStep(step_a, batches[0], dev, capacity=cap, max_nodes=max_nodes, warmup=3)
    for _ in range(3):                                   # the capture's warm-up steps, replayed eagerly on B and C
        clear_plan_cache()
        step_b.step(*default_model_inputs(warm, max_nodes))
        plain = _to_device(molecule_batch_to_arrays(batches[0]), dev)
        step_c.step(*default_model_inputs(plain, max_nodes))
    assert torch.equal(_flat(step_a), _flat(step_b))

    for b in batches:
        loss_a = captured.run(b)
        torch.cuda.synchronize()
        assert captured.check()
        padded = _to_device(pad_arrays_to_capacity(molecule_batch_to_arrays(b), cap), dev)
        clear_plan_cache()
        loss_b = step_b.step(*default_model_inputs(padded, max_nodes))
        plain = _to_device(molecule_batch_to_arrays(b), dev)
        loss_c = step_c.step(*default_model_inputs(plain, max_nodes))
        assert float(loss_a) == float(loss_b)                                 # same kernels, same order: bitwise
        assert abs(float(loss_a) - float(loss_c)) <= 2e-5 * abs(float(loss_c))   # padding changes summation order only
    assert torch.equal(_flat(step_a), _flat(step_b))
    wa, wc = _flat(step_a), _flat(step_c)
    assert float((wa - wc).abs().max()) <= 1e-4 * float(wc.abs().max())


def test_batch_that_does_not_fit_raises_and_filler_is_inert():
    from gcnn_keras_b200.synthetic import painn_c3_batch
    from gcnn_keras_b200.training import BatchCapacity, molecule_batch_to_arrays, pad_arrays_to_capacity
    small, big = painn_c3_batch(4, seed=1), painn_c3_batch(9, seed=2)
    cap = BatchCapacity.covering([small])
    with pytest.raises(ValueError):
        pad_arrays_to_capacity(molecule_batch_to_arrays(big), cap)
    cap = BatchCapacity.covering([small, big], slack=0.1)
    p = pad_arrays_to_capacity(molecule_batch_to_arrays(small), cap)
    n_real = int(small["node_len"].sum())
    assert p["xyz"].shape[0] == cap.nodes and p["index"].shape[0] == cap.edges
    assert p["node_splits"].shape[0] == cap.graphs + 2 and p["node_splits"][-1] == cap.nodes
    assert np.all(p["sample_weight"][:4] == 1) and np.all(p["sample_weight"][4:] == 0)
    ghost = p["index"][int(small["edge_len"].sum()):]
    assert np.all(np.diff(ghost[:, 0]) >= 0) and np.all(ghost[:, 0] != ghost[:, 1])
    assert ghost.max() < cap.nodes - n_real                                   # local ids of the filler molecule


def test_fit_loop_over_shuffled_batches_follows_the_schedule_and_matches_eager():
    """examples/force_schnet.py in small: the captured fit equals the eager fit (same shuffling), the learning-rate
    schedule reaches the captured update kernel, and the loss goes down."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location(
        "force_schnet_example", os.path.join(os.path.dirname(os.path.dirname(__file__)), "examples", "force_schnet.py"))
    ex = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ex)
    from gcnn_keras_b200.training import fit_force_model, linear_learning_rate
    dev = torch.device("cuda:0")
    cfg = dict(ex.DEFAULTS, model_depth=2, energy_epochs=4)
    mols = ex.synthetic_molecules(200, 3, 5.0)
    sched = lambda ep: linear_learning_rate(ep, 1e-3, 1e-4, 0, 4)
    hists, finals = [], []
    for capture in (True, False):
        torch.manual_seed(0)
        model, step = ex.create_model(cfg, ex.create_model_config(cfg))
        hists.append(fit_force_model(step, mols[32:], 48, 4, dev, validation=mols[:32], learning_rate=sched,
                                     verbose=0, capture=capture, seed=5))
        finals.append(step.optimizer.flat.detach().clone())
    cap, eag = hists
    assert cap["lr"] == [sched(e) for e in range(4)]
    assert cap["loss"][-1] < cap["loss"][0]
    np.testing.assert_allclose(cap["loss"], eag["loss"], rtol=2e-4)
    np.testing.assert_allclose(cap["val_force_mae"], eag["val_force_mae"], rtol=2e-3)
    assert float((finals[0] - finals[1]).abs().max()) <= 2e-3 * float(finals[1].abs().max())

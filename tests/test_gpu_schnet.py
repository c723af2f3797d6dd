"""SchNet energy + force + force-training gradients on the GPU against the fp64 oracle (BASELINE configs[1])."""
import numpy as np
import pytest
import torch

from helpers import RTOL, device_ragged, oracle_ragged, rel_err
from oracle import models as OM

pytestmark = pytest.mark.gpu


def _setup(batch=6, depth=2, seed=0):
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    torch.manual_seed(seed)
    torch.backends.cuda.matmul.allow_tf32 = False
    cfg = dict(SCHNET_C2_CONFIG, depth=depth)
    b = schnet_c2_batch(batch, seed)
    energy_model = make_model(**cfg)
    model = EnergyForceModel(model_energy=energy_model, energy_output=0, output_to_tensor=False, output_as_dict=False,
                             output_squeeze_states=True, is_physical_force=False)
    inputs = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
              device_ragged(b["index"], b["edge_len"])]
    return cfg, b, model, inputs


def _oracle(cfg, b, params, dtype=torch.float64):
    params = {k: v.detach().cpu().to(dtype) for k, v in params.items()}
    for v in params.values():
        v.requires_grad_(True)
    z = oracle_ragged(b["z"], b["node_len"])
    xyz = oracle_ragged(b["xyz"], b["node_len"], dtype)
    edi = oracle_ragged(b["index"], b["edge_len"])

    def energy_fn(x):
        return OM.schnet_energy(params, z, x, edi, cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                cfg["last_mlp"]["activation"], "sum", None)

    eng, force = OM.energy_force(energy_fn, xyz)
    return params, eng, force.squeeze(-1)


def test_schnet_energy_and_forces_match_fp64_oracle():
    cfg, b, model, inputs = _setup()
    eng, force = model(inputs, training=False)
    params, o_eng, o_force = _oracle(cfg, b, model.energy_model.state_dict())
    assert eng.shape == (6, 1) and force.values.shape == (54, 3)
    assert rel_err(eng.cpu().numpy(), o_eng.detach().numpy()) < RTOL
    assert rel_err(force.values.cpu().numpy(), o_force.detach().numpy()) < RTOL


def test_schnet_force_training_gradients_match_fp64_oracle():
    """d loss / d weights where loss contains the forces (double backward), every weight tensor."""
    cfg, b, model, inputs = _setup()
    eng, force = model(inputs, training=True)
    e_t = torch.tensor(b["energy"]).cuda()
    f_t = torch.tensor(b["force"]).cuda()
    loss = (1 / 200) * torch.mean((eng - e_t) ** 2) + (199 / 200) * torch.mean((force.values - f_t) ** 2)
    names = [n for n, _ in model.energy_model.named_parameters()]
    grads = torch.autograd.grad(loss, list(model.energy_model.parameters()), allow_unused=True)
    params, o_eng, o_force = _oracle(cfg, b, model.energy_model.state_dict())
    o_loss = OM.force_training_loss(o_eng, o_force, torch.tensor(b["energy"], dtype=torch.float64),
                                    torch.tensor(b["force"], dtype=torch.float64))
    assert rel_err(loss.item(), o_loss.item()) < RTOL
    o_grads = torch.autograd.grad(o_loss, [params[n] for n in names], allow_unused=True)
    worst = 0.0
    for n, g, og in zip(names, grads, o_grads):
        if og is None:
            assert g is None or float(g.abs().max()) == 0.0, n
            continue
        err = rel_err(g.cpu().numpy(), og.numpy())
        worst = max(worst, err)
        assert err < RTOL, (n, err)
    assert worst > 0.0


def test_schnet_forces_match_finite_differences_of_oracle_energy():
    cfg, b, model, inputs = _setup(batch=2, depth=1)
    _, force = model(inputs, training=False)
    params, _, _ = _oracle(cfg, b, model.energy_model.state_dict())
    z = oracle_ragged(b["z"], b["node_len"])
    edi = oracle_ragged(b["index"], b["edge_len"])
    x0 = torch.tensor(b["xyz"], dtype=torch.float64)

    def energy_sum(xv):
        with torch.no_grad():
            return OM.schnet_energy(params, z, oracle_ragged(xv.numpy(), b["node_len"], torch.float64), edi,
                                    cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                    cfg["last_mlp"]["activation"], "sum", None).sum().item()

    h = 1e-5
    for (a, c) in [(0, 0), (4, 1), (11, 2), (17, 0)]:
        xp, xm = x0.clone(), x0.clone()
        xp[a, c] += h
        xm[a, c] -= h
        fd = (energy_sum(xp) - energy_sum(xm)) / (2 * h)
        assert abs(force.values[a, c].item() - fd) < 1e-4 * max(1.0, abs(fd))


def test_schnet_train_step_runs_and_is_deterministic():
    from gcnn_keras_b200.training import ForceTrainStep
    losses = []
    for _ in range(2):
        cfg, b, model, inputs = _setup(batch=8, depth=2, seed=3)
        step = ForceTrainStep(model)
        e_t, f_t = torch.tensor(b["energy"]).cuda(), torch.tensor(b["force"]).cuda()
        run = []
        for _ in range(3):
            model_out_loss = step.step(inputs, e_t, device_ragged(b["force"], b["node_len"]).values)
            run.append(model_out_loss.item())
        losses.append(run)
    assert losses[0] == losses[1]            # bitwise run-to-run reproducibility of the whole step
    assert losses[0][2] < losses[0][0]       # and it trains

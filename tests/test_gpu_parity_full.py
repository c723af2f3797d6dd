"""fp64-oracle parity AT the BASELINE configurations (not reduced copies of them):

  C1  GatherNodes -> Dense -> PoolingLocalMessages(sum), 32 molecules x 20 atoms x 400 edges, F = 64
  C2  SchNet energy + force training exactly as benchmarked: batch 256, depth 6, units 128
  C3  PaiNN depth 6, 64 molecules x 30 atoms
  C4  HDNNP2nd, 16 systems x 64 atoms, 25 neighbours, 'kj' triples

Energy, forces, the training loss and EVERY weight gradient through the force are compared with the fp64 oracle.
Two error measures are recorded per tensor:
    max-norm  relative error   max|a - b| / max|b|                        (the north-star's "1e-5 relative")
    element-wise relative error max_i |a_i - b_i| / max(|b_i|, FLOOR * max|b|),  FLOOR = 1e-2
together with err32 = the max-norm error of the fp32 evaluation of the oracle itself (the reference's op order in
fp32, the "TF-CPU-like" comparand).  Bar: max-norm < 1e-5 on every tensor, no allowance; element-wise (floored) error < 30 x that.

Full benchmarked batch sizes (C3 at 512 molecules, C4 at 256 systems): graphs do not interact, so the fp64 oracle is
evaluated in chunks of 64 / 16 molecules; energies and forces are concatenated, and the weight gradients of the
batch-mean loss are the sums of the chunk gradients of the same loss written with the batch-wide denominators.
C2-captured: the path bench.py times (CapturedForceTrainStep: weight-gradient sink -> grouped TN launch -> flat gradient
buffer -> clipnorm + Adam kernel) after ONE replay from known weights.  C5: the partitioned GCN layer against
oracle.conv.gcn.  Everything goes to gpurun_out/parity_r3.json (copied to profiles/parity_r3.json)."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import RTOL, c1_batch, device_ragged, oracle_ragged
from oracle import R
from oracle import layers as OL
from oracle import models as OM

pytestmark = pytest.mark.gpu

FLOOR = 1e-2
_RECORD = {}


def _errs(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    scale = max(float(np.max(np.abs(b))), 1e-300)
    diff = np.abs(a - b)
    return float(diff.max() / scale), float(np.max(diff / np.maximum(np.abs(b), FLOOR * scale)))


def _record(config, name, got, want, want32=None):
    e_max, e_elem = _errs(got, want)
    row = {"shape": list(np.shape(want)), "err_maxnorm": e_max, "err_elementwise": e_elem}
    if want32 is not None:
        row["err32_maxnorm"], row["err32_elementwise"] = _errs(want32, want)
    _RECORD.setdefault(config, {})[name] = row
    return row


@pytest.fixture(scope="module", autouse=True)
def _dump_record():
    yield
    out = {"floor": FLOOR, "bar_maxnorm": RTOL,
           "definition": {"err_maxnorm": "max|a-b| / max|b|", "err_elementwise": "max_i |a_i-b_i| / max(|b_i|, floor*max|b|)",
                          "err32": "the same measures for the fp32 evaluation of the oracle (reference op order) vs fp64"},
           "configs": _RECORD}
    for cfg, rows in _RECORD.items():
        over = {n: r for n, r in rows.items() if r["err_maxnorm"] >= RTOL}
        out.setdefault("summary", {})[cfg] = {
            "tensors": len(rows), "worst_maxnorm": max(r["err_maxnorm"] for r in rows.values()),
            "worst_elementwise": max(r["err_elementwise"] for r in rows.values()),
            "above_1e-5": {n: {"err": r["err_maxnorm"], "err32": r.get("err32_maxnorm")} for n, r in over.items()}}
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "parity_r3.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)


def _check(config):
    """max-norm bar (1e-5) on every recorded tensor of `config`, no allowance; element-wise error below 30 x the bar."""
    bad = [(name, r["err_maxnorm"]) for name, r in _RECORD[config].items() if not r["err_maxnorm"] < RTOL]
    assert not bad, bad
    for name, r in _RECORD[config].items():
        assert r["err_elementwise"] < 30 * RTOL, (name, r)


# --------------------------------------------------------------------------------------------------------
# C1: GatherNodes -> Dense -> PoolingLocalMessages (BASELINE configs[0])
# --------------------------------------------------------------------------------------------------------
def test_c1_gather_dense_pool_chain_at_full_size():
    from gcnn_keras_b200.layers.gather import GatherNodesOutgoing
    from gcnn_keras_b200.layers.modules import Dense
    from gcnn_keras_b200.layers.pooling import PoolingLocalMessages
    rng = np.random.default_rng(0)
    node_len, edge_len, index = c1_batch(rng, n_graphs=32, n_nodes=20, deg=20)
    assert int(edge_len[0]) == 400
    feat = 64
    x = rng.standard_normal((int(node_len.sum()), feat)).astype(np.float32)
    w = (rng.standard_normal((feat, feat)) / np.sqrt(feat)).astype(np.float32)
    bias = rng.standard_normal(feat).astype(np.float32) * 0.1
    cot = rng.standard_normal(x.shape).astype(np.float32)

    def oracle(dtype):
        ox = torch.tensor(x, dtype=dtype, requires_grad=True)
        ow = torch.tensor(w, dtype=dtype, requires_grad=True)
        ob = torch.tensor(bias, dtype=dtype, requires_grad=True)
        nodes, idx = R.from_row_lengths(ox, node_len), oracle_ragged(index, edge_len)
        msg = OL.gather_nodes_outgoing(nodes, idx)
        msg = msg.with_values(OL.dense(msg.values, ow, ob, "kgcnn>shifted_softplus"))
        y = OL.pooling_local_edges(nodes, msg, idx, "sum").values
        gx, gw, gb = torch.autograd.grad((y * torch.tensor(cot, dtype=dtype)).sum(), [ox, ow, ob], create_graph=True)
        ggw = torch.autograd.grad((gx ** 2).sum(), ow)[0]
        return [t.detach().numpy() for t in (y, gx, gw, gb, ggw)]

    want, want32 = oracle(torch.float64), oracle(torch.float32)
    dx = torch.tensor(x, device="cuda", requires_grad=True)
    nodes = device_ragged(x, node_len).with_values(dx)
    idx = device_ragged(index, edge_len)
    layer = Dense(feat, activation="kgcnn>shifted_softplus")
    layer.build((None, feat))
    layer.cuda()
    with torch.no_grad():
        layer.kernel.copy_(torch.tensor(w))
        layer.bias.copy_(torch.tensor(bias))
    msg = layer(GatherNodesOutgoing()([nodes, idx]))
    y = PoolingLocalMessages(pooling_method="sum")([nodes, msg, idx]).values
    gx, gw, gb = torch.autograd.grad((y * torch.tensor(cot, device="cuda")).sum(), [dx, layer.kernel, layer.bias],
                                     create_graph=True)
    ggw = torch.autograd.grad((gx ** 2).sum(), layer.kernel)[0]
    for name, g, o, o32 in zip(("y", "dx", "dW", "db", "ddW_through_dx"), (y, gx, gw, gb, ggw), want, want32):
        _record("C1", name, g.detach().cpu().numpy(), o, o32)
    _check("C1")


# --------------------------------------------------------------------------------------------------------
# shared: energy / force / loss / every weight gradient through the force
# --------------------------------------------------------------------------------------------------------
def _slice_batch(b, lo, hi):
    """Molecules [lo, hi) of a ragged batch dict (index columns are per-molecule local ids: slicing rows is enough)."""
    out = {}
    for values, lengths in (("z", "node_len"), ("xyz", "node_len"), ("force", "node_len"), ("index", "edge_len"),
                            ("angle_index", "angle_len")):
        if values in b:
            off = np.concatenate([[0], np.cumsum(b[lengths])])
            out[values] = b[values][off[lo]:off[hi]]
    for k in ("node_len", "edge_len", "angle_len", "energy"):
        if k in b:
            out[k] = b[k][lo:hi]
    return out


def _chunked_oracle(b, named, chunk_fn, force_loss_factor, chunk):
    """fp64 oracle of the full batch in chunks of `chunk` molecules.  chunk_fn(dtype, sub_batch) -> (params, energy,
    force).  loss = w_e mean_B (E - E^)^2 + w_f mean_{N,3} (F - F^)^2 = sum over chunks of the squared-error SUMS over
    the batch-wide element counts, so the gradients add."""
    n_mol = len(b["node_len"])
    w_e = 1.0 / force_loss_factor
    eng, force, total, grads = [], [], 0.0, None
    for lo in range(0, n_mol, chunk):
        sub = _slice_batch(b, lo, min(n_mol, lo + chunk))
        prm, oe, of = chunk_fn(torch.float64, sub)
        part = (w_e * ((oe - torch.tensor(sub["energy"], dtype=torch.float64)) ** 2).sum() / b["energy"].size
                + (1 - w_e) * ((of - torch.tensor(sub["force"], dtype=torch.float64)) ** 2).sum() / b["force"].size)
        og = torch.autograd.grad(part, [prm[n] for n, _ in named], allow_unused=True)
        grads = [None if g is None else g.numpy() for g in og] if grads is None else \
            [a if g is None else (g.numpy() if a is None else a + g.numpy()) for a, g in zip(grads, og)]
        eng.append(oe.detach().numpy())
        force.append(of.detach().numpy())
        total += float(part)
    return np.concatenate(eng), np.concatenate(force), total, grads


def _force_training_parity(config, model, inputs, b, oracle_fn, force_loss_factor, min_grads, chunk=None):
    """oracle_fn(dtype) -> (params dict, energy, force[N,3]) of the oracle in `dtype`; with `chunk` it is
    oracle_fn(dtype, sub_batch) and the fp64 oracle runs chunk molecules at a time (no fp32 column then)."""
    from gcnn_keras_b200.functional import no_coordinate_hessian
    eng, force = model(inputs, training=True)
    e_t, f_t = torch.tensor(b["energy"]).cuda(), torch.tensor(b["force"]).cuda()
    w_e = 1.0 / force_loss_factor
    loss = w_e * torch.mean((eng - e_t) ** 2) + (1 - w_e) * torch.mean((force.values - f_t) ** 2)
    named = [(n, p) for n, p in model.energy_model.named_parameters() if p.requires_grad]
    with no_coordinate_hessian():
        grads = torch.autograd.grad(loss, [p for _, p in named], allow_unused=True)

    def oracle(dtype):
        prm, oe, of = oracle_fn(dtype)
        ol = OM.force_training_loss(oe, of, torch.tensor(b["energy"], dtype=dtype),
                                    torch.tensor(b["force"], dtype=dtype), force_loss_factor=force_loss_factor)
        og = torch.autograd.grad(ol, [prm[n] for n, _ in named], allow_unused=True)
        return oe.detach().numpy(), of.detach().numpy(), float(ol), [None if g is None else g.numpy() for g in og]

    if chunk is None:
        e64, f64, l64, g64 = oracle(torch.float64)
        e32, f32, l32, g32 = oracle(torch.float32)
    else:
        e64, f64, l64, g64 = _chunked_oracle(b, named, oracle_fn, force_loss_factor, chunk)
        e32 = f32 = None
        l32, g32 = None, [None] * len(g64)
    _record(config, "energy", eng.detach().cpu().numpy(), e64, e32)
    _record(config, "force", force.values.detach().cpu().numpy(), f64, f32)
    _record(config, "loss", np.array([loss.item()]), np.array([l64]), None if l32 is None else np.array([l32]))
    n_checked = 0
    for (n, _), g, og, og32 in zip(named, grads, g64, g32):
        if og is None or float(np.abs(og).max()) == 0.0:
            assert g is None or float(g.abs().max()) < 1e-10, n
            continue
        _record(config, "grad:" + n, g.cpu().numpy(), og, og32)
        n_checked += 1
    assert n_checked >= min_grads, n_checked
    _check(config)


def _inputs(b):
    out = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
           device_ragged(b["index"], b["edge_len"])]
    if "angle_index" in b:
        out.append(device_ragged(b["angle_index"], b["angle_len"]))
    return out


def _force_model(energy_model):
    from gcnn_keras_b200.model.force import EnergyForceModel
    return EnergyForceModel(model_energy=energy_model, energy_output=0, output_to_tensor=False, output_as_dict=False,
                            output_squeeze_states=True, is_physical_force=False)


def _oracle_params(model, dtype):
    return {k: v.detach().cpu().to(dtype).requires_grad_(True)
            for k, v in model.energy_model.state_dict().items() if v.is_floating_point()}


# --------------------------------------------------------------------------------------------------------
# C2: SchNet exactly as benchmarked (BASELINE configs[1])
# --------------------------------------------------------------------------------------------------------
def test_c2_schnet_batch256_depth6_against_fp64_oracle():
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    torch.manual_seed(1234)
    cfg = SCHNET_C2_CONFIG
    assert cfg["depth"] == 6 and cfg["interaction_args"]["units"] == 128
    b = schnet_c2_batch(256, 0)
    model = _force_model(make_model(**cfg))
    inputs = _inputs(b)
    model(inputs, training=False)                      # lazy build of the layers

    def oracle_fn(dtype):
        prm = _oracle_params(model, dtype)
        z, edi = oracle_ragged(b["z"], b["node_len"]), oracle_ragged(b["index"], b["edge_len"])
        xyz = oracle_ragged(b["xyz"], b["node_len"], dtype)
        eng, force = OM.energy_force(
            lambda x: OM.schnet_energy(prm, z, x, edi, cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                       cfg["last_mlp"]["activation"], "sum", None), xyz)
        return prm, eng, force.squeeze(-1)

    _force_training_parity("C2", model, inputs, b, oracle_fn, 200.0, min_grads=60)


# --------------------------------------------------------------------------------------------------------
# C3: PaiNN depth 6, 64 molecules of 30 atoms (BASELINE configs[2])
# --------------------------------------------------------------------------------------------------------
def test_c3_painn_depth6_64_molecules_against_fp64_oracle():
    from gcnn_keras_b200.literature.PAiNN import make_model
    from gcnn_keras_b200.synthetic import PAINN_C3_CONFIG, painn_c3_batch
    torch.manual_seed(1234)
    cfg = PAINN_C3_CONFIG
    assert cfg["depth"] == 6
    b = painn_c3_batch(64, 0, n_atoms=30)
    model = _force_model(make_model(**cfg))
    inputs = _inputs(b)
    model(inputs, training=False)

    def oracle_fn(dtype):
        prm = _oracle_params(model, dtype)
        z, edi = oracle_ragged(b["z"], b["node_len"]), oracle_ragged(b["index"], b["edge_len"])
        xyz = oracle_ragged(b["xyz"], b["node_len"], dtype)
        eng, force = OM.energy_force(
            lambda x: OM.painn_energy(prm, z, x, edi, cfg["depth"], cfg["bessel_basis"], cfg["conv_args"],
                                      cfg["equiv_initialize_kwargs"]["method"], cfg["output_mlp"]["activation"]), xyz)
        return prm, eng, force.squeeze(-1)

    _force_training_parity("C3", model, inputs, b, oracle_fn, 95.0, min_grads=60)
    assert "grad:layers_by_name.bessel.frequencies" in _RECORD["C3"]      # trainable f_n (geom.py:766-770)


# --------------------------------------------------------------------------------------------------------
# C4: HDNNP2nd, 16 systems of 64 atoms (BASELINE configs[3])
# --------------------------------------------------------------------------------------------------------
def test_c4_hdnnp2nd_16_systems_of_64_atoms_against_fp64_oracle():
    from gcnn_keras_b200.literature.HDNNP2nd import make_model_behler
    from gcnn_keras_b200.synthetic import HDNNP_C4_CONFIG, hdnnp_c4_batch
    torch.manual_seed(1234)
    cfg = HDNNP_C4_CONFIG
    b = hdnnp_c4_batch(16, 0, n_atoms=64, max_neighbours=25)
    model = _force_model(make_model_behler(**cfg))
    inputs = _inputs(b)
    model(inputs, training=False)

    def oracle_fn(dtype):
        prm = _oracle_params(model, dtype)
        z, eij = oracle_ragged(b["z"], b["node_len"]), oracle_ragged(b["index"], b["edge_len"])
        ijk = oracle_ragged(b["angle_index"], b["angle_len"])
        xyz = oracle_ragged(b["xyz"], b["node_len"], dtype)
        eng, force = OM.energy_force(
            lambda x: OM.hdnnp2nd_energy(prm, z, x, eij, ijk, cfg["g2_kwargs"], cfg["g4_kwargs"],
                                         cfg["mlp_kwargs"]["activation"]), xyz)
        return prm, eng, force.squeeze(-1)

    _force_training_parity("C4", model, inputs, b, oracle_fn, 200.0, min_grads=6)


# --------------------------------------------------------------------------------------------------------
# C3 and C4 at the BENCHMARKED batch sizes (512 molecules / 256 systems): chunked fp64 oracle
# --------------------------------------------------------------------------------------------------------
def test_c3_painn_at_the_benchmarked_batch_of_512_against_chunked_fp64_oracle():
    from gcnn_keras_b200.literature.PAiNN import make_model
    from gcnn_keras_b200.synthetic import PAINN_C3_CONFIG, painn_c3_batch
    torch.manual_seed(1234)
    cfg = PAINN_C3_CONFIG
    b = painn_c3_batch(512, 0, n_atoms=30)
    model = _force_model(make_model(**cfg))
    inputs = _inputs(b)
    model(inputs, training=False)

    def oracle_fn(dtype, sub):
        prm = _oracle_params(model, dtype)
        z, edi = oracle_ragged(sub["z"], sub["node_len"]), oracle_ragged(sub["index"], sub["edge_len"])
        xyz = oracle_ragged(sub["xyz"], sub["node_len"], dtype)
        eng, force = OM.energy_force(
            lambda x: OM.painn_energy(prm, z, x, edi, cfg["depth"], cfg["bessel_basis"], cfg["conv_args"],
                                      cfg["equiv_initialize_kwargs"]["method"], cfg["output_mlp"]["activation"]), xyz)
        return prm, eng, force.squeeze(-1)

    _force_training_parity("C3@512", model, inputs, b, oracle_fn, 95.0, min_grads=60, chunk=64)


def test_c4_hdnnp2nd_at_the_benchmarked_batch_of_256_against_chunked_fp64_oracle():
    from gcnn_keras_b200.literature.HDNNP2nd import make_model_behler
    from gcnn_keras_b200.synthetic import HDNNP_C4_CONFIG, hdnnp_c4_batch
    torch.manual_seed(1234)
    cfg = HDNNP_C4_CONFIG
    b = hdnnp_c4_batch(256, 0, n_atoms=64, max_neighbours=25)
    model = _force_model(make_model_behler(**cfg))
    inputs = _inputs(b)
    model(inputs, training=False)

    def oracle_fn(dtype, sub):
        prm = _oracle_params(model, dtype)
        z, eij = oracle_ragged(sub["z"], sub["node_len"]), oracle_ragged(sub["index"], sub["edge_len"])
        ijk = oracle_ragged(sub["angle_index"], sub["angle_len"])
        xyz = oracle_ragged(sub["xyz"], sub["node_len"], dtype)
        eng, force = OM.energy_force(
            lambda x: OM.hdnnp2nd_energy(prm, z, x, eij, ijk, cfg["g2_kwargs"], cfg["g4_kwargs"],
                                         cfg["mlp_kwargs"]["activation"]), xyz)
        return prm, eng, force.squeeze(-1)

    _force_training_parity("C4@256", model, inputs, b, oracle_fn, 200.0, min_grads=6, chunk=16)


# --------------------------------------------------------------------------------------------------------
# C2 through the CAPTURED training step (what bench.py times): flat gradient buffer and one Adam update
# --------------------------------------------------------------------------------------------------------
def test_c2_captured_training_step_gradients_and_first_adam_update_against_fp64_oracle():
    """One replay of CapturedForceTrainStep at batch 256 from known weights (keep_weights: the capture's warm-up
    updates are undone).  (1) every variable's slice of the flat gradient buffer -- filled by the weight-gradient sink's
    grouped TN launch plus the directly accumulated gradients -- equals the fp64 oracle's gradient to 1e-5 max-norm;
    (2) the updated flat parameters equal Keras' Adam(1e-3) + clipnorm 1.0 rule (force_schnet.py:196-203) evaluated in
    fp64 ON the device's gradient to 1e-6 of the update size.  (2) is anchored on the device's gradient because the first
    Adam step is lr * g / (|g| + eps / sqrt(1 - beta2)): for the few elements with |g| below ~3e-6 it amplifies a
    1e-6-relative gradient difference to a finite fraction of lr, which says nothing about either side."""
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    from gcnn_keras_b200.training import CapturedForceTrainStep, ForceTrainStep
    torch.manual_seed(1234)
    cfg = SCHNET_C2_CONFIG
    b = schnet_c2_batch(256, 0)
    dev = torch.device("cuda:0")
    model = EnergyForceModel(model_energy=make_model(**cfg), energy_output=0, output_to_tensor=True,
                             output_as_dict=False, output_squeeze_states=True, is_physical_force=False)
    step = ForceTrainStep(model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0)
    captured = CapturedForceTrainStep(step, b, dev, keep_weights=True)
    opt = step.optimizer
    assert int(opt.step_count) == 0
    p0 = opt.flat.detach().clone()
    name_of = {id(p): n for n, p in model.energy_model.named_parameters()}
    names = [name_of[id(p)] for p in step.params]

    prm = _oracle_params(model, torch.float64)                   # the weights the replay starts from
    z, edi = oracle_ragged(b["z"], b["node_len"]), oracle_ragged(b["index"], b["edge_len"])
    xyz = oracle_ragged(b["xyz"], b["node_len"], torch.float64)
    oe, of = OM.energy_force(
        lambda x: OM.schnet_energy(prm, z, x, edi, cfg["depth"], cfg["gauss_args"], cfg["interaction_args"],
                                   cfg["last_mlp"]["activation"], "sum", None), xyz)
    ol = OM.force_training_loss(oe, of.squeeze(-1), torch.tensor(b["energy"], dtype=torch.float64),
                                torch.tensor(b["force"], dtype=torch.float64), force_loss_factor=200.0)
    og = torch.autograd.grad(ol, [prm[n] for n in names], allow_unused=True)

    loss = captured.run(b)
    torch.cuda.synchronize()
    assert captured.check() and int(opt.step_count) == 1
    _record("C2-captured", "loss", np.array([float(loss)]), np.array([float(ol)]))
    n_checked = 0
    for n, view, g in zip(names, opt.grad_views, og):
        if g is None or float(g.abs().max()) == 0.0:
            assert float(view.abs().max()) < 1e-10, n
            continue
        _record("C2-captured", "gflat:" + n, view.detach().cpu().numpy(), g.numpy())
        n_checked += 1
    assert n_checked >= 60
    # Keras Adam, first step, per-variable clipnorm -- fp64 on the device's gradient
    lr, b1, b2, eps, clip = opt.hyper
    want = p0.double().cpu().clone()
    gdev = opt.gflat.detach().double().cpu()
    for off, n_real in zip(opt.offsets, opt._real):
        g = gdev[off:off + n_real]
        g = g * clip / max(float(g.norm()), clip)
        m, v = (1 - b1) * g, (1 - b2) * g * g
        want[off:off + n_real] -= lr * np.sqrt(1 - b2) / (1 - b1) * m / (v.sqrt() + eps)
    delta_want = (want - p0.double().cpu()).numpy()
    delta_got = (opt.flat.detach().double().cpu() - p0.double().cpu()).numpy()
    # fp32 parameters: the update is rounded to the parameter's ulp (|p| <~ 1 -> 6e-8), i.e. 6e-5 of lr at worst
    assert float(np.abs(delta_got - delta_want).max()) <= 1e-4 * lr
    # recorded relative to the parameters (what the next step reads); the bound on the UPDATE is the assertion above
    _record("C2-captured", "updated_parameters_adam_on_device_gradient", opt.flat.detach().cpu().numpy(), want.numpy())
    bad = [(k, r["err_maxnorm"]) for k, r in _RECORD["C2-captured"].items() if not r["err_maxnorm"] < RTOL]
    assert not bad, bad


# --------------------------------------------------------------------------------------------------------
# C5: the partitioned GCN layer against oracle.conv.gcn (BASELINE configs[4] at reduced size, one rank)
# --------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_nodes,with_grad", [(200_000, False), (25_000, True)])
def test_c5_partitioned_gcn_layer_against_the_oracle(n_nodes, with_grad):
    """gcn_layer_partitioned (Dense -> halo plan -> fused weighted aggregation -> relu) with the C5 generator of
    bench.py at 1/10 size (forward; the fp64 oracle materialises [E, 128] like the reference) and at 1/80 size with the
    gradients of x, kernel and bias."""
    from bench import gcn_c5_shard
    from gcnn_keras_b200.partition import PartitionedGraph, gcn_layer_partitioned
    from oracle import conv as OC
    feat = 128
    recv, send, w = gcn_c5_shard(n_nodes, 0, n_nodes)
    rng = np.random.default_rng(5)
    x = rng.standard_normal((n_nodes, feat)).astype(np.float32)
    kernel = (rng.standard_normal((feat, feat)) / np.sqrt(feat)).astype(np.float32)
    bias = (rng.standard_normal(feat) * 0.1).astype(np.float32)
    cot = rng.standard_normal((n_nodes, feat)).astype(np.float32)

    graph = PartitionedGraph(n_nodes, recv, send, w, 0, 1, torch.device("cuda:0"))
    dx = torch.tensor(x, device="cuda", requires_grad=with_grad)
    dk = torch.tensor(kernel, device="cuda", requires_grad=with_grad)
    db = torch.tensor(bias, device="cuda", requires_grad=with_grad)
    y = gcn_layer_partitioned(graph, dx, dk, db, "relu")

    ox = torch.tensor(x, dtype=torch.float64, requires_grad=with_grad)
    ok = torch.tensor(kernel, dtype=torch.float64, requires_grad=with_grad)
    ob = torch.tensor(bias, dtype=torch.float64, requires_grad=with_grad)
    lengths = np.array([n_nodes]), np.array([len(recv)])
    with torch.set_grad_enabled(with_grad):
        want = OC.gcn(OC.P({"lay_dense.kernel": ok, "lay_dense.bias": ob}), R.from_row_lengths(ox, lengths[0]),
                      oracle_ragged(w.reshape(-1, 1), lengths[1], torch.float64),
                      oracle_ragged(np.stack([recv, send], 1), lengths[1]), "relu", "sum", False).values
    cfg = "C5/%d" % n_nodes
    _record(cfg, "y", y.detach().cpu().numpy(), want.detach().numpy())
    if with_grad:
        got = torch.autograd.grad(y, [dx, dk, db], torch.tensor(cot, device="cuda"))
        ref = torch.autograd.grad(want, [ox, ok, ob], torch.tensor(cot, dtype=torch.float64))
        for name, g, o in zip(("dx", "dkernel", "dbias"), got, ref):
            _record(cfg, name, g.cpu().numpy(), o.numpy())
    _check(cfg)

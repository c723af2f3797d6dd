"""BASELINE.json configurations at their FULL sizes (C2: SchNet 256 x 9 atoms, C3: PaiNN 512 x 30, C4: HDNNP2nd
256 x 64 with 11.3 M triples) through size-independent properties: run-to-run bitwise identity, zero net force per
molecule (translation invariance: tests that every VJP kernel scatters exactly the opposite of what it gathers),
rotation invariance of the energy / equivariance of the forces, and invariance under a permutation of the molecules
of the batch.  The fp64 oracle covers the same models at sizes it finishes in seconds (test_gpu_{schnet,painn,hdnnp})."""
import numpy as np
import pytest
import torch

import bench
from helpers import device_ragged, rel_err

pytestmark = pytest.mark.gpu


def _inputs(b, xyz=None):
    out = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"] if xyz is None else xyz, b["node_len"]),
           device_ragged(b["index"], b["edge_len"])]
    if "angle_index" in b:
        out.append(device_ragged(b["angle_index"], b["angle_len"]))
    return out


def _flat(force):
    """Ragged forces -> values; dense [B, Nmax, 3] (output_to_tensor=True, equal-sized molecules here) -> [N, 3]."""
    from gcnn_keras_b200 import RaggedTensor
    return (force.values if isinstance(force, RaggedTensor) else force.reshape(-1, 3)).detach()


def _net_force(force, b):
    seg = torch.as_tensor(np.repeat(np.arange(len(b["node_len"])), b["node_len"]), device="cuda")
    net = torch.zeros((len(b["node_len"]), 3), device="cuda").index_add_(0, seg, force)
    return float(net.abs().max()), float(force.abs().max())


@pytest.mark.parametrize("kind", ["schnet", "painn", "hdnnp"])
def test_full_size_model_properties(kind):
    build, batch = bench.FORCE_WORKLOADS[kind][0], bench.FORCE_WORKLOADS[kind][1]
    model, _, b = build(torch.device("cuda", 0), batch, 0)
    assert len(b["node_len"]) == batch
    e0, f0 = model(_inputs(b), training=False)
    f0 = _flat(f0)
    e0 = e0.detach()
    assert torch.isfinite(e0).all() and torch.isfinite(f0).all() and float(f0.abs().max()) > 0
    # 1. determinism: no float atomics anywhere on the path
    e1, f1 = model(_inputs(b), training=False)
    f1 = _flat(f1)
    assert torch.equal(e0, e1.detach()) and torch.equal(f0, f1)
    # 2. translation invariance -> the forces of every molecule sum to zero
    net, scale = _net_force(f0, b)
    assert net < 2e-4 * max(scale, 1.0), (net, scale)
    # 3. rotation: energies invariant, forces co-rotate
    q, _ = np.linalg.qr(np.random.default_rng(5).standard_normal((3, 3)))
    q = q.astype(np.float32)
    er, fr = model(_inputs(b, b["xyz"] @ q), training=False)
    fr = _flat(fr)
    e_scale = float(e0.abs().max())
    assert float((er.detach() - e0).abs().max()) < 2e-4 * max(e_scale, 1.0)
    assert rel_err(fr.cpu().numpy(), f0.cpu().numpy() @ q) < 2e-3
    # 4. reversing the order of the molecules reverses the outputs (graphs do not interact)
    n = len(b["node_len"])
    rev = {k: v for k, v in b.items()}
    node_off = np.concatenate([[0], np.cumsum(b["node_len"])])
    edge_off = np.concatenate([[0], np.cumsum(b["edge_len"])])
    order = np.arange(n)[::-1]
    take = lambda arr, off: np.concatenate([arr[off[g]:off[g + 1]] for g in order])
    rev["z"], rev["xyz"] = take(b["z"], node_off), take(b["xyz"], node_off)
    rev["index"] = take(b["index"], edge_off)
    rev["node_len"], rev["edge_len"] = b["node_len"][order], b["edge_len"][order]
    if "angle_index" in b:
        ang_off = np.concatenate([[0], np.cumsum(b["angle_len"])])
        rev["angle_index"], rev["angle_len"] = take(b["angle_index"], ang_off), b["angle_len"][order]
    ev, fv = model(_inputs(rev), training=False)
    fv = _flat(fv)
    assert float((ev.detach().flip(0) - e0).abs().max()) < 1e-5 * max(e_scale, 1.0)
    f0_rev = torch.as_tensor(take(f0.cpu().numpy(), node_off), device="cuda")
    assert float((fv - f0_rev).abs().max()) < 1e-5 * max(float(f0.abs().max()), 1.0)

"""Pin the oracle against every golden vector the reference's own tests hold for the hot path."""
import os

import numpy as np
import torch

from conftest import GOLDEN
from oracle import R
from oracle import conv as OC
from oracle import layers as OL


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


def test_gather_nodes_matches_reference_test():
    # test/test_gather.py:28-44
    d = _load("gather.npz")
    nodes = R.from_row_lengths(torch.tensor(d["nodes"]), d["node_lengths"])
    index = R.from_row_lengths(torch.tensor(d["index"]), d["index_lengths"])
    out_concat = OL.gather_embedding(nodes, index)
    assert np.sum(np.abs(out_concat[1].numpy() - d["expected_graph1_concat"])) < 1e-6
    out = OL.gather_embedding(nodes, index, concat_axis=None)
    assert np.sum(np.abs(out[1].numpy() - d["expected_graph1"])) < 1e-6


def _bessel(dtype):
    d = _load("bessel.npz")
    pos = R.from_row_lengths(torch.tensor(d["positions"], dtype=dtype), d["position_lengths"])
    idx = R.from_row_lengths(torch.tensor(d["index"]), d["index_lengths"])
    a, b = OL.node_position(pos, idx)
    dist = OL.node_distance_euclidean(a, b)
    return OL.bessel_basis(dist, int(d["num_radial"]), float(d["cutoff"])), d


def test_bessel_basis_golden_fp32_and_fp64():
    # test/test_geom.py:79-128 (tolerance 1e-5 there)
    for dtype in (torch.float32, torch.float64):
        out, d = _bessel(dtype)
        assert np.max(np.abs(out[0].numpy() - d["bessel_basis_0"])) < 1e-5
        assert np.max(np.abs(out[1].numpy() - d["bessel_basis_1"])) < 1e-5


def _acsf_inputs(dtype, angle_key="angle_index"):
    d = _load("acsf.npz")
    z = R.from_row_lengths(torch.tensor(d["atomic_number"]), d["atomic_number_lengths"])
    xyz = R.from_row_lengths(torch.tensor(d["positions"], dtype=dtype), d["position_lengths"])
    eij = R.from_row_lengths(torch.tensor(d["edge_index"]), d["edge_index_lengths"])
    ijk = R.from_row_lengths(torch.tensor(d[angle_key]), d[angle_key + "_lengths"])
    return d, z, xyz, eij, ijk


def test_acsf_g2_golden():
    # test/test_hdnnp.py:103-111 (tolerance 1e-4 there)
    for dtype in (torch.float32, torch.float64):
        d, z, xyz, eij, _ = _acsf_inputs(dtype)
        table, elements = OC.acsf_g2_param_table(d["g2_eta"], d["g2_rs"], float(d["g2_rc"]), d["g2_elements"])
        out = OC.acsf_g2(z, xyz, eij, table, elements)
        assert out.values.shape[1] == 12
        assert np.all(np.abs(out[2][0].numpy() - d["expected_g2"]) < 1e-4)


def test_acsf_g4_golden():
    # test/test_hdnnp.py:113-134 (tolerance 1e-4 there)
    for dtype in (torch.float32, torch.float64):
        d, z, xyz, _, ijk = _acsf_inputs(dtype)
        table, elements = OC.acsf_g4_param_table(d["g4_eta"], d["g4_zeta"], d["g4_lamda"], float(d["g4_rc"]),
                                                 d["g4_elements"])
        out = OC.acsf_g4(z, xyz, ijk, table, elements, multiplicity=float(d["g4_multiplicity"]))
        assert out.values.shape[1] == 48
        assert np.all(np.abs(out[2][0].numpy() - d["expected_g4"]) < 1e-4)


def test_gauss_basis_matches_reference_numpy_code():
    # kgcnn/graph/adj.py:507-533 is the reference's own CPU code for the same expansion
    d = _load("gauss.npz")
    dist = R(torch.tensor(d["distance"]), [0, len(d["distance"])])
    out = OL.gauss_basis(dist, bins=int(d["bins"]), dist=float(d["max_distance"]), sigma=float(d["sigma"]),
                         offset=float(d["offset"]))
    assert np.max(np.abs(out.values.numpy() - d["expected"][:, 0, :])) < 1e-12

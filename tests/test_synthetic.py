"""Host-side synthetic generators against the reference's own NumPy graph code (via the golden fixtures)."""
import os

import numpy as np

from conftest import GOLDEN


def test_angle_triples_match_reference_get_angle_indices():
    # golden arrays were produced by kgcnn.graph.adj.get_angle_indices (tests/golden/make_golden.py)
    from gcnn_keras_b200.synthetic import angle_triples
    d = np.load(os.path.join(GOLDEN, "acsf.npz"))
    e_splits = np.concatenate([[0], np.cumsum(d["edge_index_lengths"])])
    for key, pairing in (("angle_index", "ik"), ("angle_index_kj", "kj")):
        a_splits = np.concatenate([[0], np.cumsum(d[key + "_lengths"])])
        for g in range(3):
            idx = d["edge_index"][e_splits[g]:e_splits[g + 1]]
            want = d[key][a_splits[g]:a_splits[g + 1]]
            got = angle_triples(idx, pairing)
            assert got.shape == want.shape and np.array_equal(got, want), (key, g)


def test_c2_c3_c4_shapes():
    from gcnn_keras_b200.synthetic import hdnnp_c4_batch, painn_c3_batch, schnet_c2_batch
    b = schnet_c2_batch(3)
    assert b["edge_len"].tolist() == [72, 72, 72] and (np.diff(b["index"][:72, 0]) >= 0).all()
    b = painn_c3_batch(2)
    assert b["node_len"].tolist() == [30, 30] and b["edge_len"].max() <= 870
    b = hdnnp_c4_batch(1, n_atoms=27)
    assert b["edge_len"][0] == 27 * 25 and b["angle_index"].shape[1] == 3

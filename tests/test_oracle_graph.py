"""Oracle (NumPy restatement of the host-side graph construction) against fixtures produced by running the
reference's own functions (tests/golden/make_golden_graph.py): bit-exact, integer work."""
import ast
import os

import numpy as np

from conftest import GOLDEN
from gcnn_keras_b200 import synthetic
from oracle import graph as OG

G = np.load(os.path.join(GOLDEN, "graph.npz"))


def _kw(name):
    return ast.literal_eval(str(G[name]))


def test_range_indices_match_reference():
    for c in range(int(G["n_range"])):
        kw = _kw("range%d_kw" % c)
        idx, lens, dist = OG.batch_range_indices(G["range%d_xyz" % c], G["range%d_node_len" % c], **kw)
        assert np.array_equal(idx, G["range%d_index" % c]), (c, kw)
        assert np.array_equal(lens, G["range%d_edge_len" % c])
        assert dist.dtype == G["range%d_dist" % c].dtype and np.array_equal(dist, G["range%d_dist" % c])


def test_angle_indices_match_reference():
    for c in range(int(G["n_angle"])):
        kw = _kw("angle%d_kw" % c)
        ijk, ij_k, lens = OG.batch_angle_indices(G["angle%d_index" % c], G["angle%d_edge_len" % c], **kw)
        assert np.array_equal(ijk, G["angle%d_ijk" % c]), (c, kw)
        assert np.array_equal(ij_k, G["angle%d_ij_k" % c]), (c, kw)
        assert np.array_equal(lens, G["angle%d_angle_len" % c])


def test_pipeline_fixture():
    idx, _ = OG.range_indices(G["pipe_xyz"], max_distance=6.0, max_neighbours=12)
    assert np.array_equal(idx, G["pipe_index"])
    assert np.array_equal(OG.angle_indices(idx, edge_pairing="kj")[0], G["pipe_ijk"])
    # the vectorised generator of the synthetic C4 batches agrees with the reference loop
    assert np.array_equal(synthetic.angle_triples(idx, "kj"), G["pipe_ijk"])

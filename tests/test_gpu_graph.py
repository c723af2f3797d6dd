"""Device graph construction (kgcnn_range_*, kgcnn_angle_*) through the C-ABI: bit-exact against the fixtures
the reference's own NumPy code produced and against the oracle on larger seeded batches."""
import ast
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gcnn_keras_b200 import RaggedTensor, synthetic
from gcnn_keras_b200 import graph as KG
from oracle import graph as OG

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(GOLDEN, "graph.npz"))
DEV = "cuda"


def _kw(name):
    return ast.literal_eval(str(G[name]))


def _ragged(vals, lens):
    return RaggedTensor.from_row_lengths(torch.as_tensor(vals, device=DEV), torch.as_tensor(lens, device=DEV))


def _np(t):
    return t.cpu().numpy()


def test_range_neighbours_golden():
    for c in range(int(G["n_range"])):
        kw = _kw("range%d_kw" % c)
        xyz = _ragged(G["range%d_xyz" % c], G["range%d_node_len" % c])
        idx, dist = KG.range_neighbours(xyz, return_distance=True, **kw)
        assert idx.values.dtype == torch.int64
        assert np.array_equal(_np(idx.values), G["range%d_index" % c]), (c, kw)
        assert np.array_equal(_np(idx.row_lengths()), G["range%d_edge_len" % c])
        assert np.array_equal(_np(dist.values), G["range%d_dist" % c])       # same dtype, same bits


def test_angle_indices_golden():
    for c in range(int(G["n_angle"])):
        kw = _kw("angle%d_kw" % c)
        nodes = _ragged(np.zeros((int(G["angle%d_node_len" % c].sum()), 1), np.float32), G["angle%d_node_len" % c])
        index = _ragged(G["angle%d_index" % c], G["angle%d_edge_len" % c])
        ijk, ij_k = KG.get_angle_indices(index, nodes, return_edge_pairs=True, **kw)
        assert np.array_equal(_np(ijk.values), G["angle%d_ijk" % c]), (c, kw)
        assert np.array_equal(_np(ij_k.values), G["angle%d_ij_k" % c]), (c, kw)
        assert np.array_equal(_np(ijk.row_lengths()), G["angle%d_angle_len" % c])


def test_angle_errors():
    nodes = _ragged(np.zeros((3, 1), np.float32), [3])
    index = _ragged(np.array([[0, 1], [1, 2]]), [2])
    with pytest.raises(ValueError):
        KG.get_angle_indices(index, nodes, edge_pairing="ij")
    with pytest.raises(ValueError):
        KG.get_angle_indices(index, nodes, edge_pairing="kk")
    with pytest.raises(NotImplementedError):
        KG.get_angle_indices(index, nodes, edge_pairing="jk", allow_self_edges=True)


def test_empty_batch():
    xyz = _ragged(np.zeros((0, 3), np.float32), [0, 0])
    idx = KG.range_neighbours(xyz)
    assert tuple(idx.values.shape) == (0, 2) and _np(idx.row_splits).tolist() == [0, 0, 0]
    ang = KG.get_angle_indices(idx, xyz)
    assert tuple(ang.values.shape) == (0, 3) and _np(ang.row_splits).tolist() == [0, 0, 0]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_pipeline_vs_oracle_c4_shape(dtype):
    """C4-shaped molecules (64 atoms, 25 nearest, 'kj' triples) at a batch the oracle finishes in seconds."""
    rng = np.random.default_rng(5)
    B, n = 6, 64
    xyz = np.concatenate([synthetic.grid_molecule(n, (4, 4, 4), 2.8, 0.4, rng)[0] for _ in range(B)]).astype(dtype)
    lens = np.full(B, n)
    kw = dict(max_distance=11 * synthetic.BOHR, max_neighbours=25)
    o_idx, o_len, o_dist = OG.batch_range_indices(xyz, lens, **kw)
    o_ijk, _, o_alen = OG.batch_angle_indices(o_idx, o_len, edge_pairing="kj")
    r_xyz = _ragged(xyz, lens)
    idx, dist = KG.range_neighbours(r_xyz, return_distance=True, **kw)
    assert np.array_equal(_np(idx.values), o_idx) and np.array_equal(_np(idx.row_lengths()), o_len)
    assert np.array_equal(_np(dist.values), o_dist)
    ijk = KG.get_angle_indices(idx, r_xyz, edge_pairing="kj")
    assert np.array_equal(_np(ijk.values), o_ijk) and np.array_equal(_np(ijk.row_lengths()), o_alen)


def test_full_c4_batch_properties():
    """BASELINE config 4 at full size (256 x 64 atoms): size-independent properties + equality with the
    vectorised host generator of the synthetic batch."""
    b = synthetic.hdnnp_c4_batch(batch=256, seed=0)
    nodes = _ragged(b["xyz"], b["node_len"])
    index = _ragged(b["index"], b["edge_len"])
    ijk = KG.get_angle_indices(index, nodes, edge_pairing="kj")
    assert np.array_equal(_np(ijk.row_lengths()), b["angle_len"])
    assert torch.equal(ijk.values, torch.as_tensor(b["angle_index"], device=DEV))
    t = ijk.values
    assert bool((t[:, 0] != t[:, 2]).all()) and bool((t[:, 0] != t[:, 1]).all())   # 'kj': k is a third atom
    again = KG.get_angle_indices(index, nodes, edge_pairing="kj")
    assert torch.equal(again.values, t)                                      # run-to-run identical
    # neighbour lists: each atom has exactly min(25, n-1) neighbours inside the generous cutoff, sorted rows
    rn = KG.range_neighbours(_ragged(b["xyz"].astype(np.float64), b["node_len"]),
                             max_distance=11 * synthetic.BOHR, max_neighbours=25)
    e = rn.values
    assert int(e.shape[0]) == 256 * 64 * 25
    key = e[:, 0] * 64 + e[:, 1]
    seg = rn.value_rowids()
    assert bool(((key[1:] > key[:-1]) | (seg[1:] != seg[:-1])).all())

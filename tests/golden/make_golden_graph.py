"""Generate tests/golden/graph.npz by RUNNING the reference's own host-side graph construction (pure NumPy,
imports without TensorFlow) on seeded inputs -- build container only, /root/reference does not travel.

    python tests/golden/make_golden_graph.py [/root/reference]

Cases: SetRange = coordinates_to_distancematrix + define_adjacency_from_distance (kgcnn/graph/adj.py:466-481,
537-593; called as kgcnn/graph/preprocessor.py:302-304) and get_angle_indices (adj.py:300-385) for every
edge_pairing and flag combination the kernels support, on molecules with distinct distances (no ties)."""
import itertools
import os
import sys

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
from kgcnn.graph.adj import (coordinates_to_distancematrix, define_adjacency_from_distance,  # noqa: E402
                             get_angle_indices)

rng = np.random.default_rng(20241104)
store = {}

# ---- range neighbours: ragged batches of molecules, fp64 and fp32 coordinates
range_cases = []
for c, (dtype, sizes, kw) in enumerate([
        ("float64", [9, 1, 14, 30, 2], dict(max_distance=3.0, max_neighbours=8, exclusive=True, self_loops=False)),
        ("float32", [9, 12, 40, 3], dict(max_distance=2.5, max_neighbours=10000, exclusive=True, self_loops=False)),
        ("float64", [20, 33, 64], dict(max_distance=20.787, max_neighbours=25, exclusive=True, self_loops=False)),
        ("float32", [17, 8, 5], dict(max_distance=2.0, max_neighbours=3, exclusive=False, self_loops=False)),
        ("float64", [11, 6], dict(max_distance=2.5, max_neighbours=4, exclusive=True, self_loops=True)),
        ("float32", [13, 21], dict(max_distance=None, max_neighbours=5, exclusive=True, self_loops=False)),
        ("float64", [13, 21], dict(max_distance=3.5, max_neighbours=None, exclusive=True, self_loops=False)),
]):
    xyz = [rng.uniform(0, 1.6 * n ** (1 / 3), size=(n, 3)).astype(dtype) for n in sizes]
    idx, dist = [], []
    for x in xyz:
        d = coordinates_to_distancematrix(x)
        adj, ind = define_adjacency_from_distance(d, **kw)
        idx.append(ind.reshape(-1, 2).astype(np.int64))
        dist.append(d[np.array(adj, dtype=bool)].reshape(-1, 1))
    store["range%d_xyz" % c] = np.concatenate(xyz)
    store["range%d_node_len" % c] = np.array(sizes, dtype=np.int64)
    store["range%d_index" % c] = np.concatenate(idx)
    store["range%d_edge_len" % c] = np.array([len(i) for i in idx], dtype=np.int64)
    store["range%d_dist" % c] = np.concatenate(dist)
    store["range%d_kw" % c] = np.array(repr(kw))
store["n_range"] = np.array(c + 1)

# ---- angle triples: random directed multigraphs (duplicates, reverse edges, self loops), sorted and unsorted
c = 0
for pairing, (multi, rev, self_), sort in itertools.product(
        ["jk", "ik", "kj", "ki"], [(False, False, False), (True, False, False), (False, True, False),
                                   (True, True, False), (False, False, True), (True, True, True)], [True, False]):
    pos_fix = 0 if pairing[0] != "k" else 1
    pos_ij = 0 if "i" in pairing else 1
    if self_ and pos_fix != pos_ij:
        continue
    sizes = [(6, 14), (1, 0), (9, 30), (4, 12), (3, 1)]
    idx, ijk, ij_k = [], [], []
    for n, m in sizes:
        e = rng.integers(0, n, size=(m, 2)).astype(np.int64)
        if sort and m:
            e = e[np.lexsort((e[:, 1], e[:, 0]))]
        idx.append(e)
        if m == 0:
            ijk.append(np.zeros((0, 3), np.int64)), ij_k.append(np.zeros((0, 2), np.int64))
            continue
        _, t, l = get_angle_indices(e, check_sorted=True, allow_multi_edges=multi, allow_self_edges=self_,
                                    allow_reverse_edges=rev, edge_pairing=pairing)
        ijk.append(np.asarray(t, dtype=np.int64).reshape(-1, 3)), ij_k.append(np.asarray(l, dtype=np.int64).reshape(-1, 2))
    store["angle%d_index" % c] = np.concatenate(idx)
    store["angle%d_node_len" % c] = np.array([s[0] for s in sizes], dtype=np.int64)
    store["angle%d_edge_len" % c] = np.array([s[1] for s in sizes], dtype=np.int64)
    store["angle%d_ijk" % c] = np.concatenate(ijk)
    store["angle%d_ij_k" % c] = np.concatenate(ij_k)
    store["angle%d_angle_len" % c] = np.array([len(t) for t in ijk], dtype=np.int64)
    store["angle%d_kw" % c] = np.array(repr(dict(edge_pairing=pairing, allow_multi_edges=multi,
                                                 allow_reverse_edges=rev, allow_self_edges=self_)))
    c += 1
store["n_angle"] = np.array(c)

# ---- the C4 pipeline on one molecule: 25 nearest within 11 Bohr-ish, then 'kj' triples (prepare_data.py:34-35)
x = rng.uniform(0, 8.0, size=(40, 3))
d = coordinates_to_distancematrix(x)
adj, ind = define_adjacency_from_distance(d, max_distance=6.0, max_neighbours=12)
_, t, l = get_angle_indices(ind, edge_pairing="kj")
store["pipe_xyz"], store["pipe_index"], store["pipe_ijk"] = x, ind.astype(np.int64), t.astype(np.int64)

np.savez_compressed(os.path.join(OUT, "graph.npz"), **store)
print("wrote graph.npz: %d range cases, %d angle cases" % (int(store["n_range"]), int(store["n_angle"])))

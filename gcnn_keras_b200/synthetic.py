"""Seeded synthetic inputs of SURVEY.md 8(d) (no dataset files): host-side NumPy only."""
import numpy as np

ETHANOL_Z = np.array([6, 6, 8, 1, 1, 1, 1, 1, 1], dtype=np.int64)
ETHANOL_XYZ = np.array([[0.00, 0.00, 0.00], [1.52, 0.00, 0.00], [2.02, 1.34, 0.00], [-0.39, 1.02, 0.00],
                        [-0.39, -0.51, 0.89], [-0.39, -0.51, -0.89], [1.90, -0.52, 0.89], [1.90, -0.52, -0.89],
                        [2.98, 1.30, 0.00]], dtype=np.float64)


def pairs_within(xyz, cutoff, max_neighbours=None):
    """All ordered pairs (i, j), i != j, with |x_i - x_j| <= cutoff, row-major (receiver-sorted) like the
    reference's define_adjacency_from_distance (kgcnn/graph/adj.py:537-593); optional k nearest per receiver."""
    d = np.linalg.norm(xyz[:, None, :] - xyz[None, :, :], axis=-1)
    adj = d <= cutoff
    np.fill_diagonal(adj, False)
    if max_neighbours is not None and max_neighbours < len(xyz) - 1:
        dd = np.where(adj, d, np.inf)
        order = np.argsort(dd, axis=1, kind="stable")[:, :max_neighbours]
        keep = np.zeros_like(adj)
        np.put_along_axis(keep, order, True, axis=1)
        adj = adj & keep
    i, j = np.nonzero(adj)
    return np.stack([i, j], axis=1).astype(np.int64)


def molecule_batch(batch, template_xyz, template_z, cutoff, jitter, rng, max_neighbours=None):
    """`batch` jittered copies of one molecule -> dict of flat arrays + row lengths (ragged layout)."""
    n = len(template_z)
    xyz = template_xyz[None] + rng.normal(0.0, jitter, size=(batch, n, 3))
    idx, edge_len = [], []
    for b in range(batch):
        p = pairs_within(xyz[b], cutoff, max_neighbours)
        idx.append(p)
        edge_len.append(len(p))
    return {"z": np.tile(template_z, batch), "xyz": xyz.reshape(-1, 3).astype(np.float32),
            "node_len": np.full(batch, n, dtype=np.int64), "index": np.concatenate(idx, axis=0),
            "edge_len": np.array(edge_len, dtype=np.int64)}


def schnet_c2_batch(batch=256, seed=0):
    """C2: ethanol-shaped molecules (9 atoms, all pairs < 5 A -> 72 edges), labels ~ N(0,1)."""
    rng = np.random.default_rng(seed)
    b = molecule_batch(batch, ETHANOL_XYZ, ETHANOL_Z, 5.0, 0.05, rng)
    b["energy"] = rng.normal(size=(batch, 1)).astype(np.float32)
    b["force"] = rng.normal(size=(batch * len(ETHANOL_Z), 3)).astype(np.float32)
    return b


SCHNET_C2_CONFIG = {   # force_schnet.py:49-57, 127-160
    "input_embedding": {"node": {"input_dim": 95, "output_dim": 128}},
    "interaction_args": {"units": 128, "use_bias": True, "activation": "shifted_softplus", "cfconv_pool": "sum"},
    "node_pooling_args": {"pooling_method": "sum"},
    "depth": 6,
    "gauss_args": {"bins": 25, "distance": 5, "offset": 0.0, "sigma": 0.4},
    "verbose": 10,
    "last_mlp": {"use_bias": [True, True, True], "units": [128, 64, 1],
                 "activation": ["shifted_softplus", "shifted_softplus", "linear"]},
    "output_embedding": "graph", "output_to_tensor": True, "use_output_mlp": False, "output_mlp": None,
}

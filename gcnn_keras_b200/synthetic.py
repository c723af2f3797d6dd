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


def grid_molecule(n_atoms, grid, spacing, jitter, rng, z_choices=(1, 6, 7, 8), z_probs=(0.5, 0.3, 0.1, 0.1)):
    """First `n_atoms` sites of an (a x b x c) grid with uniform jitter; Z drawn from z_choices."""
    sites = np.array([[i, j, k] for i in range(grid[0]) for j in range(grid[1]) for k in range(grid[2])],
                     dtype=np.float64)[:n_atoms]
    xyz = sites * spacing + rng.uniform(-jitter, jitter, size=(n_atoms, 3))
    z = rng.choice(np.array(z_choices, dtype=np.int64), size=n_atoms, p=np.array(z_probs))
    return xyz, z


def painn_c3_batch(batch=512, seed=0, n_atoms=30):
    """C3: 30 atoms on a 4x4x2 grid (1.5 A spacing, +-0.2 jitter), all pairs < 5 A, Z in {1,6,7,8}."""
    rng = np.random.default_rng(seed)
    zs, xs, idx, edge_len = [], [], [], []
    for _ in range(batch):
        xyz, z = grid_molecule(n_atoms, (4, 4, 2), 1.5, 0.2, rng)
        p = pairs_within(xyz, 5.0)
        zs.append(z)
        xs.append(xyz)
        idx.append(p)
        edge_len.append(len(p))
    b = {"z": np.concatenate(zs), "xyz": np.concatenate(xs).astype(np.float32),
         "node_len": np.full(batch, n_atoms, dtype=np.int64), "index": np.concatenate(idx),
         "edge_len": np.array(edge_len, dtype=np.int64)}
    b["energy"] = rng.normal(size=(batch, 1)).astype(np.float32)
    b["force"] = rng.normal(size=(batch * n_atoms, 3)).astype(np.float32)
    return b


PAINN_C3_CONFIG = {   # force_painn.py:48-57, 127-148
    "input_embedding": {"node": {"input_dim": 95, "output_dim": 128}},
    "equiv_initialize_kwargs": {"dim": 3, "method": "eps"},
    "bessel_basis": {"num_radial": 20, "cutoff": 5.0, "envelope_exponent": 5},
    "pooling_args": {"pooling_method": "sum"},
    "conv_args": {"units": 128, "cutoff": None},
    "update_args": {"units": 128},
    "depth": 6, "verbose": 10,
    "output_embedding": "graph",
    "output_mlp": {"use_bias": [True, True], "units": [128, 1], "activation": ["swish", "linear"]},
}


def angle_triples(idx, edge_pairing="kj"):
    """Vectorised equivalent of the reference's get_angle_indices (kgcnn/graph/adj.py:300-385) with its default
    flags (no multi / reverse / self edges, check_sorted=True): node triples (i, j, k), ordered by the first
    edge and then by the partner edge.  'kj' is what the HDNNP data preparation uses
    (kgcnn/graph/preprocessor.py:341), 'ik' what test/test_hdnnp.py:114 uses."""
    idx = np.asarray(idx, dtype=np.int64)
    if len(idx) == 0:
        return np.zeros((0, 3), dtype=np.int64)
    pos_k = 0 if edge_pairing[0] == "k" else 1
    pos_fix = 1 - pos_k
    pos_ij = 0 if "i" in edge_pairing else 1
    order = np.argsort(idx[:, pos_fix], kind="stable")             # partner edges grouped by their fixed index
    keys_sorted = idx[order, pos_fix]
    want = idx[:, pos_ij]
    lo = np.searchsorted(keys_sorted, want, side="left")
    hi = np.searchsorted(keys_sorted, want, side="right")
    cnt = hi - lo
    n_rep = np.repeat(np.arange(len(idx)), cnt)                    # first edge n of every candidate pair
    start = np.repeat(lo, cnt)
    offs = np.arange(cnt.sum()) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    m = order[start + offs]                                        # partner edge label (ascending within n)
    e_n, e_m = idx[n_rep], idx[m]
    keep = (m != n_rep)
    keep &= (e_m[:, 0] != e_n[:, 0]) | (e_m[:, 1] != e_n[:, 1])    # no multi edges
    keep &= (e_m[:, 0] != e_n[:, 1]) | (e_m[:, 1] != e_n[:, 0])    # no reverse edges
    return np.concatenate([e_n[keep], e_m[keep][:, pos_k:pos_k + 1]], axis=1)


BOHR = 1.8897


def hdnnp_c4_batch(batch=256, seed=0, n_atoms=64, max_neighbours=25, edge_pairing="kj"):
    """C4: 64 atoms on a 4x4x4 grid (2.8 Bohr spacing, +-0.4 jitter), Z in {1,6,7,8}; range_indices = 25 nearest
    within 11*1.8897 Bohr (prepare_data.py:34-35,373); angle triples with the 'kj' pairing."""
    rng = np.random.default_rng(seed)
    side = int(round(n_atoms ** (1 / 3)))
    while side ** 3 < n_atoms:
        side += 1
    zs, xs, idx, ang, edge_len, ang_len = [], [], [], [], [], []
    for _ in range(batch):
        xyz, z = grid_molecule(n_atoms, (side, side, side), 2.8, 0.4, rng)
        p = pairs_within(xyz, 11 * BOHR, max_neighbours=max_neighbours)
        t = angle_triples(p, edge_pairing)
        zs.append(z), xs.append(xyz), idx.append(p), ang.append(t)
        edge_len.append(len(p)), ang_len.append(len(t))
    b = {"z": np.concatenate(zs), "xyz": np.concatenate(xs).astype(np.float32),
         "node_len": np.full(batch, n_atoms, dtype=np.int64), "index": np.concatenate(idx),
         "edge_len": np.array(edge_len, dtype=np.int64), "angle_index": np.concatenate(ang),
         "angle_len": np.array(ang_len, dtype=np.int64)}
    b["energy"] = rng.normal(size=(batch, 1)).astype(np.float32)
    b["force"] = rng.normal(size=(batch * n_atoms, 3)).astype(np.float32)
    return b


HDNNP_C4_CONFIG = {   # force_hdnnp2nd.py:43-61, 139-168
    "g2_kwargs": {"eta": [0.03, 0.08, 0.16, 0.3, 0.5], "rs": [0.0, 3.0, 4.0, 5.0, 6.0, 7.0, 8.0], "rc": 20,
                  "elements": [1, 6, 7, 8]},
    "g4_kwargs": {"eta": [0.03, 0.08, 0.16, 0.3, 0.5], "zeta": [1, 2, 4, 8, 16], "lamda": [-1, 1], "rc": 12,
                  "elements": [1, 6, 7, 8], "multiplicity": 2.0},
    "normalize_kwargs": {},
    "mlp_kwargs": {"units": [35, 35, 1], "num_relations": 30, "activation": ["tanh", "tanh", "linear"]},
    "node_pooling_args": {"pooling_method": "sum"},
    "verbose": 10, "output_embedding": "graph", "output_to_tensor": True, "use_output_mlp": False,
    "output_mlp": None,
}

"""Shared synthetic-input generators and comparison helpers for the tests (seeded, no dataset files)."""
import numpy as np
import torch

from oracle import R


def random_graph_batch(rng, n_graphs=6, max_nodes=12, max_deg=5, sorted_edges=True, allow_empty=True, k=2):
    """Ragged batch: node counts, edge index [E, k] in SAMPLE indexing (per-graph local ids)."""
    node_len, edge_len, idx = [], [], []
    for g in range(n_graphs):
        n = int(rng.integers(0 if allow_empty else 1, max_nodes + 1))
        if g == 1 and allow_empty:
            n = 0                                   # an empty graph in the middle of the batch
        e = int(rng.integers(0, n * max_deg + 1)) if n > 0 else 0
        if g == 2 and allow_empty:
            e = 0                                   # a graph without edges (all nodes unconnected)
        ii = rng.integers(0, max(n, 1), size=(e, k)) if n > 0 else np.zeros((0, k), dtype=np.int64)
        if n > 2 and e > 0:
            ii[:, 0] = np.minimum(ii[:, 0], n - 2)  # last node never receives: unconnected tail node
        if sorted_edges and e > 0:
            ii = ii[np.argsort(ii[:, 0], kind="stable")]
        node_len.append(n)
        edge_len.append(e)
        idx.append(ii.astype(np.int64))
    index = np.concatenate(idx, axis=0) if idx else np.zeros((0, k), dtype=np.int64)
    return np.array(node_len, dtype=np.int64), np.array(edge_len, dtype=np.int64), index


def c1_batch(rng, n_graphs=32, n_nodes=20, deg=20, shuffled=False):
    """20 distinct random senders per receiver, no self loops, receiver-sorted."""
    idx = []
    for _ in range(n_graphs):
        rows = []
        for i in range(n_nodes):
            others = np.array([j for j in range(n_nodes) if j != i])
            send = rng.permutation(others)[:min(deg, n_nodes - 1)]
            if deg > n_nodes - 1:
                send = np.concatenate([send, rng.choice(others, deg - (n_nodes - 1))])
            rows.append(np.stack([np.full(len(send), i), send], axis=1))
        ii = np.concatenate(rows, axis=0)
        if shuffled:
            ii = ii[rng.permutation(len(ii))]
        idx.append(ii)
    node_len = np.full(n_graphs, n_nodes, dtype=np.int64)
    edge_len = np.array([len(i) for i in idx], dtype=np.int64)
    return node_len, edge_len, np.concatenate(idx, axis=0).astype(np.int64)


def oracle_ragged(values, lengths, dtype=None):
    t = torch.as_tensor(np.asarray(values))
    if dtype is not None and t.is_floating_point():
        t = t.to(dtype)
    return R.from_row_lengths(t, lengths)


def device_ragged(values, lengths, device="cuda", dtype=None):
    from gcnn_keras_b200 import RaggedTensor
    t = torch.as_tensor(np.asarray(values))
    if dtype is not None and t.is_floating_point():
        t = t.to(dtype)
    return RaggedTensor.from_row_lengths(t.to(device), torch.as_tensor(np.asarray(lengths)).to(device))


def rel_err(a, b):
    """max |a-b| / max(|b|_max, tiny): the '1e-5 relative' bar of BASELINE.json is applied to this."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30))


RTOL = 1e-5   # north_star: fp32 outputs, gradients and forces within 1e-5 relative of the fp64 oracle


def schnet_params(cfg, rng, dtype=torch.float64):
    """Random SchNet weights with gcnn_keras_b200's state_dict names (for CPU-only oracle tests)."""
    units = cfg["interaction_args"]["units"]
    emb = cfg["input_embedding"]["node"]
    bins = cfg["gauss_args"]["bins"]
    p = {}

    def dense(name, fi, fo, bias=True):
        lim = np.sqrt(6.0 / (fi + fo))
        p[name + ".kernel"] = torch.tensor(rng.uniform(-lim, lim, (fi, fo)), dtype=dtype)
        if bias:
            p[name + ".bias"] = torch.tensor(rng.normal(0, 0.1, (fo,)), dtype=dtype)

    pre = "layers_by_name."
    p[pre + "embedding.embeddings"] = torch.tensor(rng.uniform(-0.05, 0.05, (emb["input_dim"], emb["output_dim"])),
                                                   dtype=dtype)
    dense(pre + "dense_in", emb["output_dim"], units)
    for i in range(cfg["depth"]):
        b = pre + "interaction_%d." % i
        dense(b + "lay_dense1", units, units, bias=False)
        dense(b + "lay_cfconv.lay_dense1", bins, units)
        dense(b + "lay_cfconv.lay_dense2", units, units)
        dense(b + "lay_dense2", units, units)
        dense(b + "lay_dense3", units, units)
    fi = units
    for k, u in enumerate(cfg["last_mlp"]["units"]):
        dense(pre + "last_mlp.mlp_dense_layer_list.%d" % k, fi, u)
        fi = u
    return p

"""Generate tests/golden/*.npz from the reference tree (run in the build container only).

    python tests/golden/make_golden.py [/root/reference]

The reference's layer tests need TensorFlow, which is not installable here, so this script does NOT run
them.  It (1) extracts the literal input fixtures and expected values from the reference's own test files
with `ast` (data only), (2) copies the golden Bessel arrays out of test/assets/bessel_basis_reference.npz,
and (3) calls the reference's pure-NumPy graph code (kgcnn.graph.adj.get_angle_indices) for the angle
triples exactly as test/test_hdnnp.py:114 does.  /root/reference does not exist on the GPU box, hence the
committed .npz files.
"""
import ast
import os
import sys

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _literal(node):
    """ast.literal_eval that also accepts unary minus with spaces ('- 2.5')."""
    return ast.literal_eval(ast.unparse(node))


def _class_assignments(path, class_name):
    tree = ast.parse(open(path).read())
    out = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == class_name:
            for item in node.body:
                if isinstance(item, ast.Assign) and len(item.targets) == 1:
                    out[item.targets[0].id] = item.value
    return out


def _function_assignments(path, class_name, func_name):
    tree = ast.parse(open(path).read())
    out = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == class_name:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name == func_name:
                    for stmt in ast.walk(item):
                        if isinstance(stmt, ast.Assign) and len(stmt.targets) == 1 and \
                                isinstance(stmt.targets[0], ast.Name):
                            out[stmt.targets[0].id] = stmt.value
    return out


def _first_arg(call_node):
    assert isinstance(call_node, ast.Call)
    return _literal(call_node.args[0])


def _ragged(rows, dtype):
    lengths = np.array([len(r) for r in rows], dtype=np.int64)
    flat = np.concatenate([np.asarray(r, dtype=dtype) for r in rows], axis=0)
    return flat, lengths


def gather_fixture():
    # test/test_gather.py:11-44
    a = _class_assignments(os.path.join(REF, "test/test_gather.py"), "TestGather")
    n1, ei1 = _literal(a["n1"]), _literal(a["ei1"])
    nodes, node_len = _ragged(n1, np.float32)
    idx, idx_len = _ragged(ei1, np.int64)
    # expected as computed by the reference test itself (NumPy fancy indexing of graph 1)
    np_gather = np.array(n1[1])[np.array(ei1[1])]          # [28, 2, 1]
    np.savez(os.path.join(OUT, "gather.npz"), nodes=nodes, node_lengths=node_len, index=idx,
             index_lengths=idx_len, expected_graph1=np_gather.astype(np.float32),
             expected_graph1_concat=np_gather.reshape(28, 2).astype(np.float32))


def bessel_fixture():
    # test/test_geom.py:79-128 + test/assets/bessel_basis_reference.npz
    a = _function_assignments(os.path.join(REF, "test/test_geom.py"), "TestBesselBasisLayer", "test_result_original")
    ei, ei1 = np.array(_first_arg(a["ei"])), np.array(_first_arg(a["ei1"]))
    x, x1 = np.array(_first_arg(a["x"])), np.array(_first_arg(a["x1"]))
    ref = np.load(os.path.join(REF, "test/assets/bessel_basis_reference.npz"), allow_pickle=True)
    np.savez(os.path.join(OUT, "bessel.npz"),
             positions=np.concatenate([x, x1]).astype(np.float64), position_lengths=np.array([len(x), len(x1)]),
             index=np.concatenate([ei, ei1]).astype(np.int64), index_lengths=np.array([len(ei), len(ei1)]),
             bessel_basis_0=np.array(ref["bessel_basis_0"]), bessel_basis_1=np.array(ref["bessel_basis_1"]),
             num_radial=10, cutoff=5.0)


def acsf_fixture():
    # test/test_hdnnp.py:9-134
    sys.path.insert(0, REF)
    from kgcnn.graph.adj import get_angle_indices  # pure NumPy, imports without TensorFlow
    path = os.path.join(REF, "test/test_hdnnp.py")
    a = _class_assignments(path, "ACSFTest")
    positions = _first_arg(a["positions"])
    atomic_number = _first_arg(a["atomic_number"])
    edge_index = _first_arg(a["edge_index"])
    g2 = _function_assignments(path, "ACSFTest", "test_acsf_g2")
    g4 = _function_assignments(path, "ACSFTest", "test_acsf_g4")
    exp_g2 = np.array(_first_arg(g2["expected_result"]))
    exp_g4 = np.array(_first_arg(g4["expected_result"]))
    g2_kwargs, g4_kwargs = _literal(g2["g2_kwargs"]), _literal(g4["g4_kwargs"])
    angle_ik = [get_angle_indices(np.array(x), edge_pairing="ik")[1] for x in edge_index]
    angle_kj = [get_angle_indices(np.array(x))[1] for x in edge_index]        # default pairing (drivers)
    pos, pos_len = _ragged(positions, np.float64)
    z, z_len = _ragged(atomic_number, np.int64)
    ei, ei_len = _ragged(edge_index, np.int64)
    ang, ang_len = _ragged([np.asarray(t).reshape(-1, 3) for t in angle_ik], np.int64)
    angkj, angkj_len = _ragged([np.asarray(t).reshape(-1, 3) for t in angle_kj], np.int64)
    np.savez(os.path.join(OUT, "acsf.npz"), positions=pos, position_lengths=pos_len, atomic_number=z,
             atomic_number_lengths=z_len, edge_index=ei, edge_index_lengths=ei_len, angle_index=ang,
             angle_index_lengths=ang_len, angle_index_kj=angkj, angle_index_kj_lengths=angkj_len,
             expected_g2=exp_g2, expected_g4=exp_g4,
             g2_eta=np.array(g2_kwargs["eta"]), g2_rs=np.array(g2_kwargs["rs"]), g2_rc=g2_kwargs["rc"],
             g2_elements=np.array(g2_kwargs["elements"]),
             g4_eta=np.array(g4_kwargs["eta"]), g4_zeta=np.array(g4_kwargs["zeta"]),
             g4_lamda=np.array(g4_kwargs["lamda"]), g4_rc=g4_kwargs["rc"],
             g4_elements=np.array(g4_kwargs["elements"]), g4_multiplicity=g4_kwargs["multiplicity"])


def gauss_fixture():
    """Reference's own NumPy Gaussian expansion (kgcnn/graph/adj.py:507-533) on the bessel positions."""
    sys.path.insert(0, REF)
    from kgcnn.graph.adj import distance_to_gauss_basis
    d = np.linspace(0.0, 6.0, 61).reshape(-1, 1)
    out = distance_to_gauss_basis(d, bins=25, distance=5.0, sigma=0.4, offset=0.0)
    np.savez(os.path.join(OUT, "gauss.npz"), distance=d, expected=out, bins=25, max_distance=5.0, sigma=0.4,
             offset=0.0)


if __name__ == "__main__":
    gather_fixture()
    bessel_fixture()
    acsf_fixture()
    gauss_fixture()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            d = np.load(os.path.join(OUT, f))
            print(f, {k: d[k].shape for k in d.files})

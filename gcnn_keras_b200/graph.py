"""Batch graph construction on the device: the step in front of the message-passing path.

The reference builds range (neighbour) and angle indices per molecule on the host:
``SetRange`` = ``coordinates_to_distancematrix`` + ``define_adjacency_from_distance``
(kgcnn/graph/preprocessor.py:288-313, kgcnn/graph/adj.py:466-481, 537-593: dense N x N matrix + argsort) and
``SetAngle`` = ``get_angle_indices`` (kgcnn/graph/preprocessor.py:341, kgcnn/graph/adj.py:300-385: a Python
loop over every edge).  Here both run for the whole ragged batch in a count -> offsets -> emit pass of
``libkgcnn_b200.so``; outputs are ragged int64 tensors in SAMPLE indexing, bit-identical to what the
reference functions return graph by graph (ties of equal distances in the neighbour ranking are broken by
index; NumPy leaves them unspecified).  There is no CPU fallback.
"""
import torch

from . import _lib
from ._lib import check
from .functional import _need_cuda, _ptr, _stream
from .plan import get_plan
from .ragged import RaggedTensor, RowPartition


def _offsets(count):
    n = int(count.shape[0])
    off = torch.empty((n + 1,), dtype=torch.int64, device=count.device)
    nbytes = int(_lib.lib().kgcnn_offsets_workspace_bytes(n))
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=count.device)
    check(_lib.lib().kgcnn_offsets_from_counts(_ptr(count) if n else None, n, _ptr(off), _ptr(ws), nbytes, _stream()),
          "kgcnn_offsets_from_counts")
    return off


def _pairing_positions(edge_pairing):
    """adj.py:329-335."""
    if "k" not in edge_pairing:
        raise ValueError("Edge pairing must have index 'k'.")
    if "i" not in edge_pairing and "j" not in edge_pairing:
        raise ValueError("Edge pairing must have at least one fix index 'i' or 'j'.")
    pos_fix = 0 if edge_pairing[0] != "k" else 1
    pos_ij = 0 if "i" in edge_pairing else 1
    return pos_fix, pos_ij


def get_angle_indices(index, nodes, check_sorted=True, allow_multi_edges=False, allow_self_edges=False,
                      allow_reverse_edges=False, edge_pairing="jk", return_edge_pairs=False):
    """Batched ``get_angle_indices`` (kgcnn/graph/adj.py:300-385).

    index: RaggedTensor int64 [M, 2] in sample indexing; nodes: any RaggedTensor with the node partition
    (only its row_splits are used).  Returns the ragged node triples ``idx_ijk`` [A, 3] (and the ragged edge
    labels ``idx_ij_k`` [A, 2] with return_edge_pairs), ordered by first edge, then by partner edge -- the
    order ``check_sorted=True`` produces.  check_sorted=False would return the same rows (the reference loop
    already emits them in that order), so the flag is accepted and ignored."""
    del check_sorted
    pos_fix, pos_ij = _pairing_positions(edge_pairing)
    if allow_self_edges and pos_fix != pos_ij:
        raise NotImplementedError("allow_self_edges with edge_pairing %r (fixed column != key column)" % edge_pairing)
    vals = index.values
    _need_cuda(vals)
    if vals.dim() != 2 or vals.shape[1] != 2:
        raise ValueError("edge index must have shape [M, 2]")
    plan = get_plan(index, nodes)
    rel = plan.relation(pos_fix)
    M, N, B = plan.E, plan.N, int(index.partition.nrows)
    dev = vals.device
    idx64 = vals.to(torch.int64).contiguous()
    flags = (int(bool(allow_multi_edges)), int(bool(allow_reverse_edges)), int(bool(allow_self_edges)))
    count = torch.empty((M,), dtype=torch.int32, device=dev)
    perm = rel.perm
    rowptr = rel.rowptr
    check(_lib.lib().kgcnn_angle_count(_ptr(plan.idx32), M, N, _ptr(rowptr), _ptr(perm) if perm is not None else None,
                                       pos_fix, pos_ij, *flags, _ptr(count), _stream()), "kgcnn_angle_count")
    off = _offsets(count)
    A = int(off[-1].item())                      # one scalar read sizes the output
    ijk = torch.empty((A, 3), dtype=torch.int64, device=dev)
    ij_k = torch.empty((A, 2), dtype=torch.int64, device=dev) if return_edge_pairs else None
    splits = torch.empty((B + 1,), dtype=torch.int64, device=dev)
    check(_lib.lib().kgcnn_angle_emit(_ptr(idx64), _ptr(plan.idx32), M, N, _ptr(rowptr),
                                      _ptr(perm) if perm is not None else None, pos_fix, pos_ij, *flags, _ptr(off),
                                      _ptr(index.partition.row_splits), B, _ptr(ijk),
                                      _ptr(ij_k) if ij_k is not None else None, _ptr(splits), _stream()),
          "kgcnn_angle_emit")
    part = RowPartition(splits, nvals=A)
    out = RaggedTensor(ijk, part)
    if return_edge_pairs:
        return out, RaggedTensor(ij_k, part)
    return out


def range_neighbours(coordinates, max_distance=4.0, max_neighbours=15, exclusive=True, self_loops=False,
                     return_distance=False):
    """Batched ``SetRange`` (kgcnn/graph/preprocessor.py:288-313): range indices [M, 2] (ragged, sample
    indexing, row-major order of the adjacency matrix) from ragged coordinates [N, 3] (fp32 or fp64; the
    distance arithmetic runs in that dtype like NumPy's).  ``max_distance`` / ``max_neighbours`` may be None
    (adj.py:563,572).  return_distance adds the ragged ``range_attributes`` [M, 1]."""
    xyz = coordinates.values
    _need_cuda(xyz)
    if xyz.dim() != 2 or xyz.shape[1] != 3:
        raise ValueError("coordinates must have shape [N, 3]")
    if xyz.dtype not in (torch.float32, torch.float64):
        raise TypeError("coordinates must be float32 or float64")
    xyz = xyz.detach().contiguous()
    part = coordinates.partition
    N, B = int(xyz.shape[0]), int(part.nrows)
    dev = xyz.device
    lens = part.row_lengths()
    max_nodes = int(lens.max().item()) if B > 0 else 0
    is64 = int(xyz.dtype == torch.float64)
    args = (max_nodes, int(max_distance is not None), float(max_distance if max_distance is not None else 0.0),
            int(max_neighbours) if max_neighbours is not None else -1, int(bool(exclusive)), int(bool(self_loops)))
    count = torch.empty((N,), dtype=torch.int32, device=dev)
    check(_lib.lib().kgcnn_range_count(_ptr(xyz) if N else None, is64, _ptr(part.row_splits), B, N, *args,
                                       _ptr(count) if N else None, _stream()), "kgcnn_range_count")
    off = _offsets(count)
    M = int(off[-1].item())
    edges = torch.empty((M, 2), dtype=torch.int64, device=dev)
    dist = torch.empty((M, 1), dtype=xyz.dtype, device=dev) if return_distance else None
    splits = torch.empty((B + 1,), dtype=torch.int64, device=dev)
    check(_lib.lib().kgcnn_range_emit(_ptr(xyz) if N else None, is64, _ptr(part.row_splits), B, N, *args, _ptr(off),
                                      _ptr(edges), _ptr(dist) if dist is not None else None, _ptr(splits), _stream()),
          "kgcnn_range_emit")
    epart = RowPartition(splits, nvals=M)
    out = RaggedTensor(edges, epart)
    if return_distance:
        return out, RaggedTensor(dist, epart)
    return out

"""UNCOMPILED / never run (no TensorFlow on either machine).  TensorFlow-side binding of the shim (source only; needs TensorFlow + tf_shim/build.sh).

Registers the gradients so that the ops are closed under differentiation exactly like the PyTorch binding
(gcnn_keras_b200/functional.py): gather <-> segment-sum, fused op <-> itself on the transposed relation + the
two-sided gather product.  A reference maintainer swaps the bodies of GatherEmbeddingSelection.call
(kgcnn/layers/gather.py:218-230) and PoolingLocalEdges.call (kgcnn/layers/pooling.py:37-79) for these calls.
"""
import os

import tensorflow as tf  # noqa: F401  (not installed in this repository's environment)

_ops = tf.load_op_library(os.path.join(os.path.dirname(__file__), "libkgcnn_tf_ops.so"))


class Relation:
    """ids[E] int32 (batch indexing) + CSR (rowptr, perm) built once per batch."""

    def __init__(self, ids, num_targets, assume_sorted=False):
        self.ids, self.num_targets = ids, num_targets
        self.rowptr, self.perm = _ops.kgcnn_plan_build(ids, num_targets=num_targets, assume_sorted=assume_sorted)
        self.sorted = assume_sorted


def batch_index(index_ragged, nodes_ragged):
    """partition_row_indexing once per batch: int32 [K, E]."""
    return _ops.kgcnn_shift_index(index_ragged.values, nodes_ragged.row_splits, index_ragged.row_splits)


def gather_mul_segment_sum(x, w, rel_out, rel_src):
    @tf.custom_gradient
    def fn(x, w):
        y = _ops.kgcnn_gather_mul_segment_reduce(x, w, rel_out.rowptr, rel_out.perm, rel_src.ids, op=0,
                                                 has_perm=not rel_out.sorted)

        def grad(g):
            gx = gather_mul_segment_sum(g, w, rel_src, rel_out)
            gw = _ops.kgcnn_gather2_mul(x, rel_src.ids, g, rel_out.ids, reduce_f=(w.shape[1] == 1))
            return gx, gw

        return y, grad

    return fn(x, w)


def gather_rows(x, rel):
    @tf.custom_gradient
    def fn(x):
        y = _ops.kgcnn_gather_rows(x, rel.ids)

        def grad(g):
            return _ops.kgcnn_gather_mul_segment_reduce(g, tf.zeros([0, 1]), rel.rowptr, rel.perm, rel.ids, op=0,
                                                        has_weight=False, has_perm=not rel.sorted, has_gather=False)

        return y, grad

    return fn(x)

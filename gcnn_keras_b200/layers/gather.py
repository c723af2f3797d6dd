"""Gather layers of kgcnn/layers/gather.py backed by the sm_100a row-gather kernel.

Same class names, aliases, constructor kwargs, call signature ``call([embeddings, tensor_index])`` and
``get_config`` keys as the reference; the index bookkeeping (partition_row_indexing, gather.py:78-82,
223-227) comes from the cached per-batch plan instead of being recomputed per call.
"""
import torch

from .. import functional as F
from ..plan import get_plan
from ..ragged import RaggedTensor
from .base import GraphBaseLayer, register_keras_serializable


@register_keras_serializable(package="kgcnn", name="GatherEmbedding")
class GatherEmbedding(GraphBaseLayer):
    """kgcnn/layers/gather.py:9-149.  Default: gather all K index columns and concatenate (concat_axis=2)."""

    def __init__(self, axis=1, concat_axis=2, split_axis=None, split_indices=None, concat_indices=None,
                 node_indexing="sample", **kwargs):
        super().__init__(node_indexing=node_indexing, **kwargs)
        self.concat_axis = concat_axis
        self.axis = axis
        self.split_axis = split_axis
        self.split_indices = split_indices
        self.concat_indices = concat_indices
        if split_axis is not None and concat_axis is not None:   # gather.py:60-61
            raise ValueError("Can not both split and concatenate new index axis. At least one must be `None`.")

    def build(self, input_shape):
        super().build(input_shape)
        if len(input_shape) != 2:
            print("Number of inputs for layer '%s' must be 2: `[nodes, indices]` ." % self.name)

    def call(self, inputs, **kwargs):
        nodes, index = self.assert_ragged_input_rank(inputs)
        if self.axis != 1 or self.concat_axis not in (None, 2) or self.split_axis not in (None, 2):
            raise NotImplementedError("GatherEmbedding: only axis=1 and concat/split axis 2 are on the hot path")
        plan = get_plan(index, nodes)
        x = nodes.values
        if self.concat_axis == 2:                                 # gather.py:85-90
            cols = list(self.concat_indices) if self.concat_indices else list(range(plan.K))
            return index.with_values(F.gather_concat(x, [plan.relation(c) for c in cols]))   # strided gathers, no cat
        if self.split_axis == 2:                                  # gather.py:92-97
            cols = list(self.split_indices) if self.split_indices else list(range(plan.K))
            return [index.with_values(F.gather(x, plan.relation(c))) for c in cols]
        parts = [F.gather(x, plan.relation(c)) for c in range(plan.K)]
        return index.with_values(torch.stack(parts, dim=1))       # [E, K, F]

    def get_config(self):
        config = super().get_config()
        config.update({"concat_axis": self.concat_axis, "axis": self.axis, "split_axis": self.split_axis,
                       "concat_indices": self.concat_indices, "split_indices": self.split_indices,
                       "node_indexing": self.node_indexing})
        return config


GatherNodes = GatherEmbedding


@register_keras_serializable(package="kgcnn", name="GatherEmbeddingSelection")
class GatherEmbeddingSelection(GraphBaseLayer):
    """kgcnn/layers/gather.py:153-245: list of gathered embeddings, one per selection index."""

    def __init__(self, selection_index, axis=1, axis_indices=2, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis
        self.axis_indices = axis_indices
        if not isinstance(selection_index, (list, tuple, int)):   # gather.py:195-196
            raise ValueError("Indices for selection must be list or tuple for layer `GatherEmbeddingSelection`.")
        self.selection_index = [selection_index] if isinstance(selection_index, int) else list(selection_index)

    def call(self, inputs, **kwargs):
        nodes, index = self.assert_ragged_input_rank(inputs)
        if self.axis != 1 or self.axis_indices != 2:
            raise NotImplementedError("GatherEmbeddingSelection: only axis=1, axis_indices=2 are on the hot path")
        plan = get_plan(index, nodes)
        outs = []
        for i in self.selection_index:
            out = index.with_values(F.gather(nodes.values, plan.relation(i)))
            out.provenance = ("gathered", nodes, plan.relation(i))      # lets the geometry layers fuse (geom.py)
            outs.append(out)
        return outs

    def get_config(self):
        config = super().get_config()
        config.update({"axis": self.axis, "axis_indices": self.axis_indices,
                       "selection_index": self.selection_index})
        return config


GatherNodesSelection = GatherEmbeddingSelection


@register_keras_serializable(package="kgcnn", name="GatherNodesIngoing")
class GatherNodesIngoing(GatherEmbeddingSelection):
    """kgcnn/layers/gather.py:249-282: receiving node (index column 0)."""

    def __init__(self, selection_index=0, **kwargs):
        super().__init__(selection_index=selection_index, **kwargs)

    def call(self, inputs, **kwargs):
        return super().call(inputs, **kwargs)[0]


@register_keras_serializable(package="kgcnn", name="GatherNodesOutgoing")
class GatherNodesOutgoing(GatherEmbeddingSelection):
    """kgcnn/layers/gather.py:286-319: sending node (index column 1)."""

    def __init__(self, selection_index=1, **kwargs):
        super().__init__(selection_index=selection_index, **kwargs)

    def call(self, inputs, **kwargs):
        return super().call(inputs, **kwargs)[0]


@register_keras_serializable(package="kgcnn", name="GatherState")
class GatherState(GraphBaseLayer):
    """kgcnn/layers/gather.py:323-375: repeat a per-graph state for every node/edge of the graph."""

    def call(self, inputs, **kwargs):
        env, target = inputs
        if not isinstance(target, RaggedTensor):
            target = self.assert_ragged_input_rank(target)
        rel = target.partition.relation()                         # value -> graph
        return target.with_values(F.gather(env, rel))

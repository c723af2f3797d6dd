"""PAiNN make_model of kgcnn/literature/PAiNN.py:18-155 wired with the B200-backed layers."""
from ..layers.casting import ChangeTensorType
from ..layers.conv.painn_conv import EquivariantInitialize, PAiNNconv, PAiNNUpdate
from ..layers.geom import (BesselBasisLayer, CosCutOffEnvelope, EdgeDirectionNormalized, NodeDistanceEuclidean,
                           NodePosition)
from ..layers.mlp import MLP, GraphMLP
from ..layers.modules import LazyAdd, OptionalInputEmbedding
from ..layers.pooling import PoolingNodes
from ..model.utils import Model, update_model_kwargs

__model_version__ = "2022.11.25"

model_default = {
    "name": "PAiNN",
    "inputs": [
        {"shape": (None,), "name": "node_attributes", "dtype": "float32", "ragged": True},
        {"shape": (None, 3), "name": "node_coordinates", "dtype": "float32", "ragged": True},
        {"shape": (None, 2), "name": "edge_indices", "dtype": "int64", "ragged": True}
    ],
    "input_embedding": {"node": {"input_dim": 95, "output_dim": 128}},
    "equiv_initialize_kwargs": {"dim": 3, "method": "zeros"},
    "bessel_basis": {"num_radial": 20, "cutoff": 5.0, "envelope_exponent": 5},
    "pooling_args": {"pooling_method": "sum"},
    "conv_args": {"units": 128, "cutoff": None, "conv_pool": "sum"},
    "update_args": {"units": 128},
    "equiv_normalization": False, "node_normalization": False,
    "depth": 3,
    "verbose": 10,
    "output_embedding": "graph", "output_to_tensor": True,
    "output_mlp": {"use_bias": [True, True], "units": [128, 1], "activation": ["swish", "linear"]}
}


@update_model_kwargs(model_default)
def make_model(inputs=None, input_embedding=None, equiv_initialize_kwargs=None, bessel_basis=None, depth=None,
               pooling_args=None, conv_args=None, update_args=None, equiv_normalization=None,
               node_normalization=None, name=None, verbose=None, output_embedding=None, output_to_tensor=None,
               output_mlp=None):
    """kgcnn/literature/PAiNN.py:45-155.  Inputs [node_attributes, node_coordinates, bond_indices(, equiv)]."""
    if equiv_normalization or node_normalization:
        raise NotImplementedError("GraphLayerNormalization / GraphBatchNormalization are outside the hot path")
    if output_embedding not in ("graph", "node"):
        raise ValueError("Unsupported output embedding for mode `PAiNN`")
    layers = {
        "embedding": OptionalInputEmbedding(**input_embedding["node"], use_embedding=len(inputs[0]["shape"]) < 2),
        "equiv_init": EquivariantInitialize(**equiv_initialize_kwargs),
        "node_position": NodePosition(),
        "direction": EdgeDirectionNormalized(),
        "distance": NodeDistanceEuclidean(),
        "envelope": CosCutOffEnvelope(conv_args["cutoff"]),
        "bessel": BesselBasisLayer(**bessel_basis),
        "add": LazyAdd(),
        "pool": PoolingNodes(**pooling_args),
        "output_mlp": MLP(**output_mlp) if output_embedding == "graph" else GraphMLP(**output_mlp),
        "cast": ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor"),
    }
    for i in range(depth):
        layers["conv_%d" % i] = PAiNNconv(**conv_args)
        layers["update_%d" % i] = PAiNNUpdate(**update_args)

    def wiring(L, x_in):
        node_input, xyz_input, edi = x_in[:3]
        z = L["embedding"](node_input)                                   # PAiNN.py:103-104
        v = x_in[3] if len(x_in) > 3 else L["equiv_init"](z)             # :106-109
        pos1, pos2 = L["node_position"]([xyz_input, edi])                # :115
        rij = L["direction"]([pos1, pos2])                               # :116
        d = L["distance"]([pos1, pos2])                                  # :117
        env = L["envelope"](d)                                           # :118
        rbf = L["bessel"](d)                                             # :119
        for i in range(depth):                                           # :121-133
            ds, dv = L["conv_%d" % i]([z, v, rbf, env, rij, edi])
            z = L["add"]([z, ds])
            v = L["add"]([v, dv])
            ds, dv = L["update_%d" % i]([z, v])
            z = L["add"]([z, ds])
            v = L["add"]([v, dv])
        if output_embedding == "graph":                                  # :136-138
            out = L["output_mlp"](L["pool"](z))
        else:                                                            # :139-142
            out = L["output_mlp"](z)
            if output_to_tensor:
                out = L["cast"](out)
        return out

    model = Model(layers, wiring, input_config=inputs, name=name, config=dict(
        inputs=inputs, input_embedding=input_embedding, equiv_initialize_kwargs=equiv_initialize_kwargs,
        bessel_basis=bessel_basis, depth=depth, pooling_args=pooling_args, conv_args=conv_args,
        update_args=update_args, equiv_normalization=equiv_normalization, node_normalization=node_normalization,
        name=name, verbose=verbose, output_embedding=output_embedding, output_to_tensor=output_to_tensor,
        output_mlp=output_mlp))
    model.__kgcnn_model_version__ = __model_version__
    return model

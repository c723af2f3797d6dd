"""SchNet make_model of kgcnn/literature/Schnet.py:24-148 wired with the B200-backed layers."""
from ..layers.casting import ChangeTensorType
from ..layers.conv.schnet_conv import SchNetInteraction
from ..layers.geom import GaussBasisLayer, NodeDistanceEuclidean, NodePosition
from ..layers.mlp import MLP, GraphMLP
from ..layers.modules import Dense, OptionalInputEmbedding
from ..layers.pooling import PoolingNodes
from ..model.utils import Model, update_model_kwargs

__model_version__ = "2022.11.25"

model_default = {
    "name": "Schnet",
    "inputs": [{"shape": (None,), "name": "node_attributes", "dtype": "float32", "ragged": True},
               {"shape": (None, 3), "name": "node_coordinates", "dtype": "float32", "ragged": True},
               {"shape": (None, 2), "name": "edge_indices", "dtype": "int64", "ragged": True}],
    "input_embedding": {"node": {"input_dim": 95, "output_dim": 64}},
    "make_distance": True, "expand_distance": True,
    "interaction_args": {"units": 128, "use_bias": True,
                         "activation": "kgcnn>shifted_softplus", "cfconv_pool": "sum"},
    "node_pooling_args": {"pooling_method": "sum"},
    "depth": 4,
    "gauss_args": {"bins": 20, "distance": 4, "offset": 0.0, "sigma": 0.4},
    "verbose": 10,
    "last_mlp": {"use_bias": [True, True], "units": [128, 64],
                 "activation": ["kgcnn>shifted_softplus", "kgcnn>shifted_softplus"]},
    "output_embedding": "graph", "output_to_tensor": True,
    "use_output_mlp": True,
    "output_mlp": {"use_bias": [True, True], "units": [64, 1],
                   "activation": ["kgcnn>shifted_softplus", "linear"]}
}


@update_model_kwargs(model_default)
def make_model(inputs=None, input_embedding=None, make_distance=None, expand_distance=None, gauss_args=None,
               interaction_args=None, node_pooling_args=None, depth=None, name=None, verbose=None, last_mlp=None,
               output_embedding=None, use_output_mlp=None, output_to_tensor=None, output_mlp=None):
    """kgcnn/literature/Schnet.py:46-148.  Inputs: [node_attributes, node_coordinates | edge_distance,
    edge_indices] as RaggedTensors; output: graph embeddings (batch, L) or node embeddings."""
    if output_embedding not in ("graph", "node"):
        raise ValueError("Unsupported output embedding for mode `SchNet`")
    layers = {
        "embedding": OptionalInputEmbedding(**input_embedding["node"], use_embedding=len(inputs[0]["shape"]) < 2),
        "node_position": NodePosition(),
        "distance": NodeDistanceEuclidean(),
        "gauss": GaussBasisLayer(**gauss_args),
        "dense_in": Dense(interaction_args["units"], activation="linear"),
        "last_mlp": GraphMLP(**last_mlp),
        "pool": PoolingNodes(**node_pooling_args),
        "cast": ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor"),
    }
    for i in range(depth):
        layers["interaction_%d" % i] = SchNetInteraction(**interaction_args)
    if use_output_mlp:
        layers["output_mlp"] = MLP(**output_mlp) if output_embedding == "graph" else GraphMLP(**output_mlp)

    def wiring(L, x_in):
        node_input, xyz_input, edi = x_in
        n = L["embedding"](node_input)                                   # Schnet.py:110-111
        if make_distance:                                                # :114-119
            pos1, pos2 = L["node_position"]([xyz_input, edi])
            ed = L["distance"]([pos1, pos2])
        else:
            ed = xyz_input
        if expand_distance:                                              # :121-122
            ed = L["gauss"](ed)
        n = L["dense_in"](n)                                             # :125
        for i in range(depth):                                           # :126-127
            n = L["interaction_%d" % i]([n, ed, edi])
        n = L["last_mlp"](n)                                             # :129
        if output_embedding == "graph":                                  # :132-135
            out = L["pool"](n)
            if use_output_mlp:
                out = L["output_mlp"](out)
        else:                                                            # :136-141
            out = n
            if use_output_mlp:
                out = L["output_mlp"](out)
            if output_to_tensor:
                out = L["cast"](out)
        return out

    model = Model(layers, wiring, input_config=inputs, name=name, config=dict(
        inputs=inputs, input_embedding=input_embedding, make_distance=make_distance,
        expand_distance=expand_distance, gauss_args=gauss_args, interaction_args=interaction_args,
        node_pooling_args=node_pooling_args, depth=depth, name=name, verbose=verbose, last_mlp=last_mlp,
        output_embedding=output_embedding, use_output_mlp=use_output_mlp, output_to_tensor=output_to_tensor,
        output_mlp=output_mlp))
    model.__kgcnn_model_version__ = __model_version__
    return model

"""GCN make_model of kgcnn/literature/GCN.py:18-112 wired with the B200-backed layers."""
from ..layers.casting import ChangeTensorType
from ..layers.conv.gcn_conv import GCN
from ..layers.mlp import MLP, GraphMLP
from ..layers.modules import Dense, OptionalInputEmbedding
from ..layers.pooling import PoolingNodes
from ..model.utils import Model, update_model_kwargs

__model_version__ = "2022.11.25"

model_default = {
    "name": "GCN",
    "inputs": [{"shape": (None,), "name": "node_attributes", "dtype": "float32", "ragged": True},
               {"shape": (None, 1), "name": "edge_weights", "dtype": "float32", "ragged": True},
               {"shape": (None, 2), "name": "edge_indices", "dtype": "int64", "ragged": True}],
    "input_embedding": {"node": {"input_dim": 95, "output_dim": 64},
                        "edge": {"input_dim": 25, "output_dim": 1}},
    "gcn_args": {"units": 100, "use_bias": True, "activation": "relu", "pooling_method": "sum"},
    "depth": 3, "verbose": 10,
    "output_embedding": "graph", "output_to_tensor": True,
    "output_mlp": {"use_bias": [True, True, False], "units": [25, 10, 1],
                   "activation": ["relu", "relu", "sigmoid"]}
}


@update_model_kwargs(model_default)
def make_model(inputs=None, input_embedding=None, depth=None, gcn_args=None, name=None, verbose=None,
               output_embedding=None, output_to_tensor=None, output_mlp=None):
    """kgcnn/literature/GCN.py:37-112.  Inputs [node_attributes, edge_weights (batch,[M],1), edge_indices]."""
    if inputs[1]["shape"][-1] != 1:                                      # GCN.py:82-84
        raise ValueError("No edge features available for GCN, only edge weights of pre-scaled adjacency matrix, "
                         "must be shape (batch, None, 1), but got (without batch-dimension): ", inputs[1]["shape"])
    if output_embedding not in ("graph", "node"):
        raise ValueError("Unsupported output embedding for `GCN`")
    layers = {
        "embedding_node": OptionalInputEmbedding(**input_embedding["node"],
                                                 use_embedding=len(inputs[0]["shape"]) < 2),
        "embedding_edge": OptionalInputEmbedding(**input_embedding["edge"],
                                                 use_embedding=len(inputs[1]["shape"]) < 2),
        "dense_in": Dense(gcn_args["units"], use_bias=True, activation="linear"),
        "pool": PoolingNodes(),
        "output_mlp": MLP(**output_mlp) if output_embedding == "graph" else GraphMLP(**output_mlp),
        "cast": ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor"),
    }
    for i in range(depth):
        layers["gcn_%d" % i] = GCN(**gcn_args)

    def wiring(L, x_in):
        node_input, edge_input, edi = x_in
        n = L["embedding_node"](node_input)                              # GCN.py:91-92
        ed = L["embedding_edge"](edge_input)                             # :93-94
        n = L["dense_in"](n)                                             # :98
        for i in range(depth):                                           # :99-100
            n = L["gcn_%d" % i]([n, ed, edi])
        if output_embedding == "graph":                                  # :103-105
            return L["output_mlp"](L["pool"](n))
        out = L["output_mlp"](n)                                         # :106-109
        return L["cast"](out) if output_to_tensor else out

    model = Model(layers, wiring, input_config=inputs, name=name, config=dict(
        inputs=inputs, input_embedding=input_embedding, depth=depth, gcn_args=gcn_args, name=name, verbose=verbose,
        output_embedding=output_embedding, output_to_tensor=output_to_tensor, output_mlp=output_mlp))
    model.__kgcnn_model_version__ = __model_version__
    return model

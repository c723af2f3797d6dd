"""HDNNP2nd make_model_behler (kgcnn/literature/HDNNP2nd.py:132-242) and make_model_weighted (:25-129) wired
with the B200-backed layers."""
from ..layers.casting import ChangeTensorType
from ..layers.conv.acsf_conv import ACSFG2, ACSFG4
from ..layers.conv.wacsf_conv import wACSFAng, wACSFRad
from ..layers.mlp import MLP, GraphMLP
from ..layers.modules import LazyConcatenate
from ..layers.pooling import PoolingNodes
from ..layers.relational import RelationalMLP
from ..model.utils import Model, update_model_kwargs

__model_version__ = "2023.01.11"

model_default_behler = {
    "name": "HDNNP2nd",
    "inputs": [{"shape": (None,), "name": "node_number", "dtype": "int64", "ragged": True},
               {"shape": (None, 3), "name": "node_coordinates", "dtype": "float32", "ragged": True},
               {"shape": (None, 2), "name": "range_indices", "dtype": "int64", "ragged": True},
               {"shape": (None, 3), "name": "angle_indices_nodes", "dtype": "int64", "ragged": True}],
    "g2_kwargs": {"eta": [0.0, 0.3], "rs": [0.0, 3.0], "rc": 10.0, "elements": [1, 6, 16]},
    "g4_kwargs": {"eta": [0.0, 0.3], "lamda": [-1.0, 1.0], "rc": 6.0,
                  "zeta": [1.0, 8.0], "elements": [1, 6, 16], "multiplicity": 2.0},
    "normalize_kwargs": {},
    "mlp_kwargs": {"units": [64, 64, 1], "num_relations": 96, "activation": ["swish", "swish", "linear"]},
    "node_pooling_args": {"pooling_method": "sum"},
    "verbose": 10,
    "output_embedding": "graph", "output_to_tensor": True,
    "use_output_mlp": False,
    "output_mlp": {"use_bias": [True, True], "units": [64, 1], "activation": ["swish", "linear"]}
}


@update_model_kwargs(model_default_behler)
def make_model_behler(inputs=None, node_pooling_args=None, name=None, verbose=None, normalize_kwargs=None,
                      g2_kwargs=None, g4_kwargs=None, mlp_kwargs=None, output_embedding=None, use_output_mlp=None,
                      output_to_tensor=None, output_mlp=None):
    """kgcnn/literature/HDNNP2nd.py:154-242.  Inputs [node_number, node_coordinates, edge_indices,
    angle_indices_nodes].  (The reference asserts on g2_kwargs['elemental_mapping'], a key the drivers never
    pass -- it only runs under `python -O`; that assert is not reproduced.)"""
    if normalize_kwargs:
        raise NotImplementedError("GraphBatchNormalization is skipped by the drivers (normalize_kwargs={})")
    if output_embedding not in ("graph", "node"):
        raise ValueError("Unsupported output embedding for mode `HDNNP2nd`")
    layers = {
        "g2": ACSFG2(**ACSFG2.make_param_table(**g2_kwargs)),
        "g4": ACSFG4(**ACSFG4.make_param_table(**g4_kwargs)),
        "concat": LazyConcatenate(),
        "mlp": RelationalMLP(**mlp_kwargs),
        "pool": PoolingNodes(**node_pooling_args),
        "cast": ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor"),
    }
    if use_output_mlp:
        layers["output_mlp"] = MLP(**output_mlp) if output_embedding == "graph" else GraphMLP(**output_mlp)

    def wiring(L, x_in):
        node_input, xyz_input, edge_index_input, angle_index_input = x_in
        rep_g2 = L["g2"]([node_input, xyz_input, edge_index_input])          # HDNNP2nd.py:213
        rep_g4 = L["g4"]([node_input, xyz_input, angle_index_input])         # :214
        rep = L["concat"]([rep_g2, rep_g4])                                  # :215
        n = L["mlp"]([rep, node_input])                                      # :222
        if output_embedding == "graph":                                      # :225-228
            out = L["pool"](n)
            if use_output_mlp:
                out = L["output_mlp"](out)
        else:
            out = n
            if use_output_mlp:
                out = L["output_mlp"](out)
            if output_to_tensor:
                out = L["cast"](out)
        return out

    model = Model(layers, wiring, input_config=inputs, name=name, config=dict(
        inputs=inputs, node_pooling_args=node_pooling_args, name=name, verbose=verbose,
        normalize_kwargs=normalize_kwargs, g2_kwargs=g2_kwargs, g4_kwargs=g4_kwargs, mlp_kwargs=mlp_kwargs,
        output_embedding=output_embedding, use_output_mlp=use_output_mlp, output_to_tensor=output_to_tensor,
        output_mlp=output_mlp))
    model.__kgcnn_model_version__ = __model_version__
    return model


model_default_weighted = {
    "name": "HDNNP2nd",
    "inputs": [{"shape": (None,), "name": "node_number", "dtype": "int64", "ragged": True},
               {"shape": (None, 3), "name": "node_coordinates", "dtype": "float32", "ragged": True},
               {"shape": (None, 2), "name": "range_indices", "dtype": "int64", "ragged": True},
               {"shape": (None, 3), "name": "angle_indices_nodes", "dtype": "int64", "ragged": True}],
    "w_acsf_ang_kwargs": {},
    "w_acsf_rad_kwargs": {},
    "normalize_kwargs": None,
    "mlp_kwargs": {"units": [64, 64, 1], "num_relations": 96, "activation": ["swish", "swish", "linear"]},
    "node_pooling_args": {"pooling_method": "sum"},
    "verbose": 10,
    "output_embedding": "graph", "output_to_tensor": True,
    "use_output_mlp": False,
    "output_mlp": {"use_bias": [True, True], "units": [64, 1], "activation": ["swish", "linear"]}
}


@update_model_kwargs(model_default_weighted)
def make_model_weighted(inputs=None, node_pooling_args=None, name=None, verbose=None, w_acsf_ang_kwargs=None,
                        w_acsf_rad_kwargs=None, normalize_kwargs=None, mlp_kwargs=None, output_embedding=None,
                        use_output_mlp=None, output_to_tensor=None, output_mlp=None):
    """kgcnn/literature/HDNNP2nd.py:46-129: wACSFRad + wACSFAng -> RelationalMLP -> PoolingNodes."""
    if normalize_kwargs:
        raise NotImplementedError("GraphBatchNormalization is out of scope (SURVEY.md 2); pass normalize_kwargs=None")
    if output_embedding not in ("graph", "node"):
        raise ValueError("Unsupported output embedding for mode `HDNNP2nd`")
    layers = {
        "rad": wACSFRad(**w_acsf_rad_kwargs),
        "ang": wACSFAng(**w_acsf_ang_kwargs),
        "concat": LazyConcatenate(),
        "mlp": RelationalMLP(**mlp_kwargs),
        "pool": PoolingNodes(**node_pooling_args),
        "cast": ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor"),
    }
    if use_output_mlp:
        layers["output_mlp"] = MLP(**output_mlp) if output_embedding == "graph" else GraphMLP(**output_mlp)

    def wiring(L, x_in):
        node_input, xyz_input, edge_index_input, angle_index_input = x_in
        rep_rad = L["rad"]([node_input, xyz_input, edge_index_input])        # HDNNP2nd.py:101
        rep_ang = L["ang"]([node_input, xyz_input, angle_index_input])       # :102
        rep = L["concat"]([rep_rad, rep_ang])                                # :103
        n = L["mlp"]([rep, node_input])                                      # :109
        if output_embedding == "graph":                                      # :112-115
            out = L["pool"](n)
            if use_output_mlp:
                out = L["output_mlp"](out)
        else:
            out = n
            if use_output_mlp:
                out = L["output_mlp"](out)
            if output_to_tensor:
                out = L["cast"](out)
        return out

    model = Model(layers, wiring, input_config=inputs, name=name, config=dict(
        inputs=inputs, node_pooling_args=node_pooling_args, name=name, verbose=verbose,
        w_acsf_ang_kwargs=w_acsf_ang_kwargs, w_acsf_rad_kwargs=w_acsf_rad_kwargs,
        normalize_kwargs=normalize_kwargs, mlp_kwargs=mlp_kwargs, output_embedding=output_embedding,
        use_output_mlp=use_output_mlp, output_to_tensor=output_to_tensor, output_mlp=output_mlp))
    model.__kgcnn_model_version__ = __model_version__
    return model


make_model = make_model_behler

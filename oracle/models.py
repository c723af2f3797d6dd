"""Oracle restatement of the named literature models and of EnergyForceModel (torch CPU autograd).
Parameter names follow gcnn_keras_b200's state_dict ("layers_by_name.<layer>.<sublayer>.kernel").
Test infrastructure only."""
import torch

from . import conv as C
from . import layers as L
from . import ops
from .conv import P
from .ragged import R


def _mlp(p, x, activations):
    """MLP / GraphMLP, kgcnn/layers/mlp.py:246-416: stack of Dense."""
    for i, act in enumerate(activations):
        sub = p.sub("mlp_dense_layer_list.%d" % i)
        x = L.dense(x, sub["kernel"], sub.get("bias"), act)
    return x


def schnet_energy(params, node_attr, xyz, edi, depth, gauss_args, interaction_args=None, last_mlp_act=None,
                  node_pooling="sum", output_mlp_act=None):
    """kgcnn/literature/Schnet.py:104-148 (make_distance=True, expand_distance=True, graph embedding)."""
    p = P(params, "layers_by_name.")
    ia = {"activation": "kgcnn>shifted_softplus", "cfconv_pool": "sum"}
    ia.update(interaction_args or {})
    table = p["embedding.embeddings"]
    n = node_attr.with_values(table[node_attr.values.to(torch.int64)])          # :110-111
    pos1, pos2 = L.node_position(xyz, edi)                                       # :116
    ed = L.node_distance_euclidean(pos1, pos2)                                   # :117
    ed = L.gauss_basis(ed, bins=gauss_args["bins"], dist=gauss_args["distance"], sigma=gauss_args["sigma"],
                       offset=gauss_args["offset"])                              # :122
    n = L.dense(n, p["dense_in.kernel"], p.get("dense_in.bias"), "linear")      # :125
    for i in range(depth):                                                       # :126-127
        n = C.schnet_interaction(p.sub("interaction_%d" % i), n, ed, edi, ia["activation"], ia["cfconv_pool"])
    n = _mlp(p.sub("last_mlp"), n, last_mlp_act)                                 # :129
    out = L.pooling_embedding(n, node_pooling)                                   # :133
    if output_mlp_act is not None:
        out = _mlp(p.sub("output_mlp"), out, output_mlp_act)                     # :135
    return out


def energy_force(energy_fn, xyz, is_physical_force=False, create_graph=True):
    """EnergyForceModel.call, kgcnn/model/force.py:136-201.  The reference pads, takes batch_jacobian of the
    energy (batch, states) wrt the padded coordinates and transposes to (batch, N, 3, states); because graphs
    do not interact this equals d(sum_b E[b, s])/dx on the ragged values, which is what autograd gives here."""
    xv = xyz.values.detach().clone().requires_grad_(True)
    eng = energy_fn(xyz.with_values(xv))
    cols = []
    for s in range(eng.shape[1]):
        (g,) = torch.autograd.grad(eng[:, s].sum(), xv, create_graph=create_graph, retain_graph=True)
        cols.append(g)
    de_dr = torch.stack(cols, dim=-1)
    if is_physical_force:
        de_dr = -de_dr
    return eng, de_dr


def force_training_loss(eng, force, e_target, f_target, force_loss_factor=200.0):
    """compile(loss=[mse, mse], loss_weights=[1/f, 1-1/f]) of force_schnet.py:194-203 (forces squeezed)."""
    w_e, w_f = 1.0 / force_loss_factor, 1.0 - 1.0 / force_loss_factor
    return w_e * torch.mean((eng - e_target) ** 2) + w_f * torch.mean((force - f_target) ** 2)


def painn_energy(params, node_attr, xyz, edi, depth, bessel_basis, conv_args=None, equiv_method="zeros",
                 output_mlp_act=("swish", "linear"), pooling="sum"):
    """kgcnn/literature/PAiNN.py:100-155 (graph embedding, no normalization)."""
    p = P(params, "layers_by_name.")
    ca = {"cutoff": None, "conv_pool": "sum"}
    ca.update(conv_args or {})
    table = p["embedding.embeddings"]
    z = node_attr.with_values(table[node_attr.values.to(torch.int64)])          # :103-104
    v = C.equivariant_initialize(z, dim=3, method=equiv_method)                  # :109
    pos1, pos2 = L.node_position(xyz, edi)                                       # :115
    rij = L.edge_direction_normalized(pos1, pos2)                                # :116
    d = L.node_distance_euclidean(pos1, pos2)                                    # :117
    env = L.cos_cutoff_envelope(d, ca["cutoff"])                                 # :118
    rbf = L.bessel_basis(d, bessel_basis["num_radial"], bessel_basis["cutoff"],
                         bessel_basis.get("envelope_exponent", 5), frequencies=p["bessel.frequencies"])   # :119
    for i in range(depth):                                                       # :121-133
        ds, dv = C.painn_conv(p.sub("conv_%d" % i), z, v, rbf, env, rij, edi, "swish", ca["cutoff"],
                              ca["conv_pool"])
        z = z.with_values(z.values + ds.values)
        v = v.with_values(v.values + dv.values)
        ds, dv = C.painn_update(p.sub("update_%d" % i), z, v, "swish")
        z = z.with_values(z.values + ds.values)
        v = v.with_values(v.values + dv.values)
    out = L.pooling_embedding(z, pooling)                                        # :137
    return _mlp(p.sub("output_mlp"), out, list(output_mlp_act))                  # :138


def relational_dense(x, relations, kernel, bias, activation):
    """RelationalDense.call, kgcnn/layers/relational.py:219-238."""
    kernel_per_feature = kernel[relations.values.to(torch.int64)]           # :233
    new = torch.einsum("nf,nfu->nu", x.values, kernel_per_feature)          # :234 batch_dot
    if bias is not None:
        new = new + bias                                                     # :235-236
    return x.with_values(ops.activation(activation)(new))                    # :237


def hdnnp2nd_energy(params, z, xyz, eij, ijk, g2_kwargs, g4_kwargs, mlp_activation, pooling="sum"):
    """make_model_behler, kgcnn/literature/HDNNP2nd.py:206-229 (graph embedding, no normalization)."""
    p = P(params, "layers_by_name.")
    g2 = dict(g2_kwargs)
    g4 = dict(g4_kwargs)
    mult = g4.pop("multiplicity", None)
    t2, e2 = C.acsf_g2_param_table(g2["eta"], g2["rs"], g2["rc"], g2["elements"])
    t4, e4 = C.acsf_g4_param_table(g4["eta"], g4["zeta"], g4["lamda"], g4["rc"], g4["elements"])
    rep_g2 = C.acsf_g2(z, xyz, eij, t2, e2)                                  # :213
    rep_g4 = C.acsf_g4(z, xyz, ijk, t4, e4, multiplicity=mult)               # :214
    n = xyz.with_values(torch.cat([rep_g2.values, rep_g4.values], dim=-1))   # :215
    for i, act in enumerate(mlp_activation):                                 # :222 RelationalMLP
        sub = p.sub("mlp.mlp_dense_layer_list.%d" % i)
        n = relational_dense(n, z, sub["kernel"], sub.get("bias"), act)
    return L.pooling_embedding(n, pooling)                                   # :226

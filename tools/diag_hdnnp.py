"""Diagnostics: HDNNP energy / force / gradient errors against the fp64 oracle, tcgen05 Dense vs library GEMM."""
import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
from helpers import device_ragged, oracle_ragged, rel_err
from oracle import models as OM
import gcnn_keras_b200.functional as F
from test_gpu_hdnnp import _setup

for lib in (False, True):
    F.DENSE_LIBRARY = lib
    cfg, b, model, inputs = _setup()
    eng, force = model(inputs, training=True)
    params = {k: v.detach().cpu().double().requires_grad_(True)
              for k, v in model.energy_model.state_dict().items() if v.is_floating_point()}
    z = oracle_ragged(b["z"], b["node_len"]); eij = oracle_ragged(b["index"], b["edge_len"])
    ijk = oracle_ragged(b["angle_index"], b["angle_len"]); xyz = oracle_ragged(b["xyz"], b["node_len"], torch.float64)
    fn = lambda x: OM.hdnnp2nd_energy(params, z, x, eij, ijk, cfg["g2_kwargs"], cfg["g4_kwargs"], cfg["mlp_kwargs"]["activation"])
    o_eng, o_force = OM.energy_force(fn, xyz); o_force = o_force.squeeze(-1)
    print("library" if lib else "tcgen05", "E", eng.detach().cpu().numpy().ravel(), o_eng.detach().numpy().ravel(), b["energy"].ravel())
    print("  eng err %.2e force err %.2e" % (rel_err(eng.detach().cpu().numpy(), o_eng.detach().numpy()),
                                            rel_err(force.values.detach().cpu().numpy(), o_force.detach().numpy())))
    e_t, f_t = torch.tensor(b["energy"]).cuda(), torch.tensor(b["force"]).cuda()
    loss = (1 / 200) * torch.mean((eng - e_t) ** 2) + (199 / 200) * torch.mean((force.values - f_t) ** 2)
    named = [(n, p) for n, p in model.energy_model.named_parameters() if p.requires_grad]
    with F.no_coordinate_hessian():
        grads = torch.autograd.grad(loss, [p for _, p in named], allow_unused=True)
    o_loss = OM.force_training_loss(o_eng, o_force, torch.tensor(b["energy"], dtype=torch.float64), torch.tensor(b["force"], dtype=torch.float64))
    o_grads = torch.autograd.grad(o_loss, [params[n] for n, _ in named], allow_unused=True)
    for (n, _), g, og in zip(named, grads, o_grads):
        print("  %-50s %.2e  |g| %.3e" % (n, rel_err(g.cpu().numpy(), og.numpy()), float(og.abs().max())))
    if 'bias' in named[-1][0]:
        print("  last bias", grads[-1].cpu().numpy().ravel(), o_grads[-1].numpy().ravel())

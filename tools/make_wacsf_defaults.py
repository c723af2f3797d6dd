"""Extracts the published wACSF default parameter tables (Gastegger et al. 2018, SI; numbers only) from the
reference's module-level assignments (kgcnn/layers/conv/wacsf_conv.py:12-95) without importing TensorFlow, and
writes them as data to gcnn_keras_b200/layers/conv/wacsf_defaults.npz.  Run in the build container only."""
import ast
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open("/root/reference/kgcnn/layers/conv/wacsf_conv.py").read()
ns = {"np": np}
for node in ast.parse(src).body:
    if isinstance(node, ast.Assign):
        names = [ast.unparse(t) for t in node.targets]
        if any(n.startswith(("radial_eta_mu_defaults", "angular_eta_mu_lambda_zeta_defaults")) for n in names):
            exec(compile(ast.Module([node], []), "wacsf_defaults", "exec"), ns)
np.savez_compressed(os.path.join(ROOT, "gcnn_keras_b200/layers/conv/wacsf_defaults.npz"),
                    radial_eta_mu_rc=ns["radial_eta_mu_defaults"],
                    angular_eta_mu_lambda_zeta_rc=ns["angular_eta_mu_lambda_zeta_defaults"])

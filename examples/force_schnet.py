"""SchNet energy + force training, the shape of the reference's force_schnet.py (create_model_config :127-160,
create_model :162-204, train_single_fold / model.fit :206-269) on the B200 path.

The reference reads a pickled MD data set; there is no data set on this machine, so the molecules are synthetic
(ethanol-like fragments of 6..12 atoms with a smooth pair potential as labels) -- the point of the script is the loop:
model built from the same JSON-style config, batches of DIFFERENT atom / pair counts drawn by shuffling, one captured
CUDA graph for all of them, linear learning-rate schedule, validation MAE per epoch.

    python examples/force_schnet.py --epochs 5 --molecules 2048 --batch-size 128
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from gcnn_keras_b200.literature.Schnet import make_model  # noqa: E402
from gcnn_keras_b200.model.force import EnergyForceModel  # noqa: E402
from gcnn_keras_b200.synthetic import pairs_within  # noqa: E402
from gcnn_keras_b200.training import ForceTrainStep, fit_force_model, linear_learning_rate, predict_energy_force  # noqa: E402

DEFAULTS = {   # force_schnet.py:40-57
    "energy_epochs": 500, "energy_initial_learning_rate": 1e-3, "energy_final_learning_rate": 1e-8,
    "energy_batch_size": 128, "force_loss_factor": 200,
    "input_embedding_dim": 128, "interaction_units": 128, "activation": "shifted_softplus", "model_depth": 6,
    "gauss_bins": 25, "gauss_distance": 5, "gauss_offset": 0.0, "gauss_sigma": 0.4, "last_mlp_units": [128, 64, 1],
}


def create_model_config(config):
    """force_schnet.py:127-160."""
    return {
        "input_embedding": {"node": {"input_dim": 95, "output_dim": config["input_embedding_dim"]}},
        "interaction_args": {"units": config["interaction_units"], "use_bias": True,
                             "activation": config["activation"], "cfconv_pool": "sum"},
        "node_pooling_args": {"pooling_method": "sum"},
        "depth": config["model_depth"],
        "gauss_args": {"bins": config["gauss_bins"], "distance": config["gauss_distance"],
                       "offset": config["gauss_offset"], "sigma": config["gauss_sigma"]},
        "verbose": 10,
        "last_mlp": {"use_bias": [True] * len(config["last_mlp_units"]), "units": config["last_mlp_units"],
                     "activation": [config["activation"]] * (len(config["last_mlp_units"]) - 1) + ["linear"]},
        "output_embedding": "graph", "output_to_tensor": True, "use_output_mlp": False, "output_mlp": None,
    }


def create_model(train_config, model_config):
    """force_schnet.py:162-204: EnergyForceModel around the SchNet energy model, Adam + clipnorm 1, loss weights
    [1/f, 1 - 1/f]."""
    model = EnergyForceModel(model_energy=make_model(**model_config), energy_output=0, esp_input=None,
                             esp_grad_input=None, output_to_tensor=True, output_as_dict=False,
                             output_squeeze_states=True, is_physical_force=False)
    step = ForceTrainStep(model, lr=train_config["energy_initial_learning_rate"],
                          force_loss_factor=train_config["force_loss_factor"], clipnorm=1.0)
    return model, step


def synthetic_molecules(n, seed, cutoff):
    """Fragments of 6..12 atoms; labels from a Morse-like pair potential, force column = dE/dx (is_physical_force=False)."""
    rng = np.random.default_rng(seed)
    mols = []
    for _ in range(n):
        k = int(rng.integers(6, 13))
        grid = np.array([[a, b, c] for a in range(3) for b in range(2) for c in range(2)], dtype=np.float64)[:k]
        xyz = grid * 1.4 + rng.normal(0.0, 0.12, size=(k, 3))
        z = rng.choice(np.array([1, 6, 7, 8]), size=k, p=[0.5, 0.3, 0.1, 0.1]).astype(np.int64)
        d = xyz[:, None, :] - xyz[None, :, :]
        r = np.linalg.norm(d, axis=-1) + np.eye(k)
        q = np.exp(-1.2 * (r - 1.4))
        pair = (q * q - 2.0 * q) * (1 - np.eye(k))
        dpair = (-2.4 * q * q + 2.4 * q) * (1 - np.eye(k))             # d pair / d r
        energy = 0.5 * pair.sum()
        grad = (dpair[:, :, None] * d / r[:, :, None]).sum(axis=1)      # dE/dx_i
        mols.append({"z": z, "xyz": xyz.astype(np.float32), "index": pairs_within(xyz, cutoff),
                     "energy": np.array([energy], dtype=np.float32), "force": grad.astype(np.float32)})
    return mols


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--molecules", type=int, default=2048)
    ap.add_argument("--batch-size", type=int, default=DEFAULTS["energy_batch_size"])
    ap.add_argument("--depth", type=int, default=DEFAULTS["model_depth"])
    ap.add_argument("--eager", action="store_true", help="no CUDA graph (comparison)")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("examples/force_schnet.py needs a CUDA device (there is no CPU path)")
    device = torch.device("cuda", 0)
    config = dict(DEFAULTS, energy_epochs=args.epochs, energy_batch_size=args.batch_size, model_depth=args.depth)
    model_config = create_model_config(config)
    print("Model config:")
    print(json.dumps(model_config))
    torch.manual_seed(0)
    model, step = create_model(config, model_config)
    mols = synthetic_molecules(args.molecules, 0, float(config["gauss_distance"]))
    n_val = max(args.batch_size, len(mols) // 10)
    train, val = mols[n_val:], mols[:n_val]
    # extensive label scaling (EnergyForceExtensiveLabelScaler): energies and forces in units of the force spread
    scale = float(np.std(np.concatenate([m["force"].reshape(-1) for m in train])))
    shift = float(np.mean([m["energy"][0] / len(m["z"]) for m in train]))
    for m in mols:
        m["energy"] = (m["energy"] - shift * len(m["z"])) / scale
        m["force"] = m["force"] / scale
    schedule = lambda epoch: linear_learning_rate(epoch, config["energy_initial_learning_rate"],
                                                  config["energy_final_learning_rate"], 0, config["energy_epochs"])
    hist = fit_force_model(step, train, args.batch_size, args.epochs, device, validation=val, shuffle=True,
                           learning_rate=schedule, capture=not args.eager)
    e, f = predict_energy_force(model, val, args.batch_size, device)
    e_true = np.stack([m["energy"] for m in val])
    print("validation MAE: energy %.5f  force %.5f  (scaled units)" % (
        float(np.mean(np.abs(e - e_true))),
        float(np.mean(np.concatenate([np.abs(fp - m["force"]).reshape(-1) for fp, m in zip(f, val)])))))
    print(json.dumps({"loss": hist["loss"], "val_force_mae": hist["val_force_mae"], "epoch_seconds": hist["time"]}))


if __name__ == "__main__":
    main()

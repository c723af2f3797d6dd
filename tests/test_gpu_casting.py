"""a20: ragged <-> padded (+ mask) through kgcnn_pad / kgcnn_unpad against the oracle restatement of
ChangeTensorType (kgcnn/layers/casting.py:78-84) and of the pad -> ragged cast of EnergyForceModel
(kgcnn/model/force.py:204-211).  Data movement only: bit-exact."""
import numpy as np
import pytest
import torch

from helpers import device_ragged, oracle_ragged
from oracle import layers as OL

pytestmark = pytest.mark.gpu

LENGTHS = [np.array([3, 0, 5, 1, 0, 4]), np.array([9] * 7), np.array([0, 0, 2]), np.array([1])]


@pytest.mark.parametrize("lengths", LENGTHS)
@pytest.mark.parametrize("tail", [(), (1,), (3,), (4,), (128,), (3, 5)])
@pytest.mark.parametrize("boolean_mask", [False, True])
def test_change_tensor_type_mask_bit_exact(lengths, tail, boolean_mask):
    from gcnn_keras_b200.layers.casting import ChangeTensorType
    rng = np.random.default_rng(0)
    vals = rng.standard_normal((int(lengths.sum()),) + tail).astype(np.float32)
    want_p, want_m = OL.change_tensor_type_mask(oracle_ragged(vals, lengths), default_value=-7.5,
                                                boolean_mask=boolean_mask)
    layer = ChangeTensorType(input_tensor_type="ragged", output_tensor_type="mask", default_value=-7.5,
                             boolean_mask=boolean_mask)
    got_p, got_m = layer(device_ragged(vals, lengths))
    assert got_p.dtype == torch.float32 and got_m.dtype == (torch.bool if boolean_mask else torch.float32)
    assert torch.equal(got_p.cpu(), want_p) and torch.equal(got_m.cpu(), want_m)
    # 'tensor' output: padded only, default 0
    got_t = ChangeTensorType(input_tensor_type="ragged", output_tensor_type="tensor")(device_ragged(vals, lengths))
    assert torch.equal(got_t.cpu(), OL.ragged_to_tensor(oracle_ragged(vals, lengths)))


@pytest.mark.parametrize("dtype", [np.int64, np.int32])
def test_pad_integer_values_and_fixed_shape(dtype):
    lengths = np.array([2, 0, 6, 3])
    vals = np.random.default_rng(1).integers(-2 ** 40 if dtype == np.int64 else -2 ** 20, 2 ** 20,
                                             size=(11, 2)).astype(dtype)
    r = device_ragged(vals, lengths)
    want = OL.ragged_to_tensor(oracle_ragged(vals, lengths), default_value=-1)
    assert torch.equal(r.to_tensor(default_value=-1).cpu(), want)
    # to_tensor(shape=(B, 4, ...)): longer rows are truncated, shorter ones padded
    want4 = OL.ragged_to_tensor(oracle_ragged(vals, lengths), default_value=0, max_len=4)
    assert torch.equal(r.to_tensor(shape=(None, 4, 2)).cpu(), want4)


@pytest.mark.parametrize("lengths", LENGTHS)
def test_unpad_inverts_pad_like_the_force_model_cast(lengths):
    from gcnn_keras_b200 import RaggedTensor
    vals = np.random.default_rng(2).standard_normal((int(lengths.sum()), 3)).astype(np.float32)
    r = device_ragged(vals, lengths)
    padded, mask = r.to_tensor_and_mask()
    back = RaggedTensor.from_padded(padded, r.partition)
    want = OL.padded_to_ragged(padded.cpu(), mask.cpu())
    assert torch.equal(back.values.cpu(), want.values) and torch.equal(back.row_splits.cpu(), want.row_splits)
    assert torch.equal(back.values.cpu(), torch.from_numpy(vals))


def test_pad_unpad_are_adjoint_and_differentiable_twice():
    from gcnn_keras_b200 import functional as F
    lengths = np.array([3, 0, 5, 1])
    rng = np.random.default_rng(3)
    r = device_ragged(rng.standard_normal((9, 3)).astype(np.float32), lengths)
    v = r.values.clone().requires_grad_(True)
    w = torch.randn(4, 5, 3, device="cuda")
    y = F.pad(v, r.row_splits, 5, 0.0)
    (g,) = torch.autograd.grad((y * w).sum(), v, create_graph=True)
    assert torch.equal(g, F.raw_unpad(w, r.row_splits, 9))                 # <pad v, w> = <v, unpad w>
    p = torch.randn(4, 5, 3, device="cuda", requires_grad=True)
    u = F.unpad(p, r.row_splits, 9)
    (gp,) = torch.autograd.grad((u * v.detach()).sum(), p)
    assert torch.equal(gp, F.raw_pad(v.detach(), r.row_splits, 5, 0.0)[0])


def test_energy_force_model_padded_outputs_match_ragged_values():
    """output_to_tensor=True (what force_schnet.py passes, kgcnn/model/force.py:189-194) pads the force with
    kgcnn_pad; the padded tensor holds exactly the ragged force values and zeros elsewhere."""
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    torch.manual_seed(0)
    b = schnet_c2_batch(4, 0)
    b["node_len"] = np.array([9, 9, 9, 9])
    em = make_model(**dict(SCHNET_C2_CONFIG, depth=1))
    kw = dict(energy_output=0, output_as_dict=False, output_squeeze_states=True, is_physical_force=False)
    inputs = [device_ragged(b["z"], b["node_len"]), device_ragged(b["xyz"], b["node_len"]),
              device_ragged(b["index"], b["edge_len"])]
    _, f_pad = EnergyForceModel(model_energy=em, output_to_tensor=True, **kw)(inputs, training=False)
    _, f_rag = EnergyForceModel(model_energy=em, output_to_tensor=False, **kw)(inputs, training=False)
    assert f_pad.shape == (4, 9, 3)
    assert torch.equal(f_pad.reshape(-1, 3), f_rag.values)

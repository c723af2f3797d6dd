"""The captured training step's plan-validation guard: a replayed batch that breaks the sortedness the graph was
captured with must leave parameters, Adam moments and the step counter untouched, and be reported on the host."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def small_schnet_step(device, batch=8, seed=0, depth=2):
    from gcnn_keras_b200.literature.Schnet import make_model
    from gcnn_keras_b200.model.force import EnergyForceModel
    from gcnn_keras_b200.synthetic import SCHNET_C2_CONFIG, schnet_c2_batch
    from gcnn_keras_b200.training import ForceTrainStep
    cfg = dict(SCHNET_C2_CONFIG, depth=depth)
    torch.manual_seed(1234)
    model = EnergyForceModel(model_energy=make_model(**cfg), energy_output=0, output_to_tensor=True,
                             output_as_dict=False, output_squeeze_states=True, is_physical_force=False).to(device)
    return model, ForceTrainStep(model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0), schnet_c2_batch(batch, seed)


def unsorted_copy(batch):
    """Same sizes, the edges of every molecule in reverse order: receivers descend where the capture saw them ascend."""
    bad = {k: np.array(v, copy=True) for k, v in batch.items()}
    off = 0
    for n in batch["edge_len"]:
        bad["index"][off:off + n] = batch["index"][off:off + n][::-1]
        off += int(n)
    return bad


def test_a_batch_that_breaks_the_captured_plan_skips_the_update_and_is_reported():
    from gcnn_keras_b200.training import CapturedForceTrainStep
    dev = torch.device("cuda:0")
    model, step, good = small_schnet_step(dev)
    captured = CapturedForceTrainStep(step, good, dev, warmup=2)
    opt = step.optimizer
    captured.run(good)
    torch.cuda.synchronize()
    assert captured.check()
    before = [t.clone() for t in (opt.flat, opt.m, opt.v, opt.step_count)]
    captured.run(unsorted_copy(good))
    torch.cuda.synchronize()
    assert not captured.check()
    for was, now in zip(before, (opt.flat, opt.m, opt.v, opt.step_count)):
        assert torch.equal(was, now)                              # nothing of the optimizer state moved
    with pytest.raises(RuntimeError, match="violated"):           # the next call reports the skipped step
        captured.run(good)
    captured.run(good)                                            # and a good batch updates again
    torch.cuda.synchronize()
    assert captured.check()
    assert int(opt.step_count) == int(before[3]) + 1
    assert not torch.equal(before[0], opt.flat)

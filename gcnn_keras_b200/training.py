"""Energy+force training step (what ``model.fit`` runs per batch in force_schnet.py:194-269).

loss = w_E * MSE(E) + w_F * MSE(dE/dx); gradients with respect to the weights go THROUGH the force (double
backward); Adam with Keras' per-variable ``clipnorm``.  The whole step can be captured in a CUDA graph
(config C2 is launch-bound: ~2 k kernels of a few microseconds each).
"""
import torch


def keras_mse(pred, target):
    return torch.mean((pred - target) ** 2)


class ForceTrainStep:
    def __init__(self, model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0, eps=1e-7):
        self.model = model
        self.w_e = 1.0 / force_loss_factor
        self.w_f = 1.0 - 1.0 / force_loss_factor
        self.clipnorm = clipnorm
        self.lr, self.eps = lr, eps
        self.params = None
        self.optimizer = None

    def _ensure_optimizer(self):
        if self.optimizer is None:
            self.params = [p for p in self.model.parameters() if p.requires_grad]
            self.optimizer = torch.optim.Adam(self.params, lr=self.lr, betas=(0.9, 0.999), eps=self.eps,
                                              capturable=True, foreach=True)

    def loss(self, inputs, e_target, f_target):
        out = self.model(inputs, training=True)
        eng, force = out[0], out[-1]
        return self.w_e * keras_mse(eng, e_target) + self.w_f * keras_mse(force, f_target)

    def step(self, inputs, e_target, f_target):
        loss = self.loss(inputs, e_target, f_target)
        self._ensure_optimizer()
        grads = torch.autograd.grad(loss, self.params, allow_unused=True)
        grads = [torch.zeros_like(p) if g is None else g for g, p in zip(grads, self.params)]
        if self.clipnorm is not None:                     # tf.clip_by_norm per variable (Keras clipnorm)
            norms = torch._foreach_norm(grads)
            scales = [self.clipnorm / torch.clamp(n, min=self.clipnorm) for n in norms]
            grads = torch._foreach_mul(grads, scales)
        for p, g in zip(self.params, grads):
            p.grad = g
        self.optimizer.step()
        return loss

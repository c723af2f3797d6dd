"""Energy+force training step (what ``model.fit`` runs per batch in force_schnet.py:194-269).

loss = w_E * MSE(E) + w_F * MSE(dE/dx); gradients with respect to the weights go THROUGH the force (double
backward); Adam with Keras' per-variable ``clipnorm``.  The whole step can be captured in a CUDA graph
(config C2 is launch-bound: ~2 k kernels of a few microseconds each).
"""
import torch


def keras_mse(pred, target):
    if isinstance(pred, torch.Tensor) and hasattr(target, "to_tensor"):
        target = target.to_tensor()          # padded prediction (output_to_tensor=True) vs ragged label
    pred = pred if isinstance(pred, torch.Tensor) else pred.values
    target = target if isinstance(target, torch.Tensor) else target.values
    return torch.mean((pred - target) ** 2)


class FlatAdam:
    """Keras Adam + per-variable clipnorm over ONE flat parameter buffer: cat(grads) -> [all-reduce] -> per-variable
    squared norms (CSR segment kernel) -> fused clip + Adam kernel.  4 launches instead of ~200."""

    def __init__(self, params, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, clipnorm=1.0):
        from . import functional as F
        from .plan import Relation
        self.params = list(params)
        dev = self.params[0].device
        real = [p.numel() for p in self.params]
        sizes = [(n + 63) // 64 * 64 for n in real]              # every variable starts 256 B aligned
        self.total = sum(sizes)
        self.flat = torch.zeros(self.total, dtype=torch.float32, device=dev)
        off = 0
        for p, n, padded in zip(self.params, real, sizes):
            self.flat[off:off + n].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + n].view_as(p)          # parameters become views of the flat buffer
            off += padded
        self._real, self._padded = real, sizes
        self.offsets = [int(o) for o in ([0] + list(torch.tensor(sizes).cumsum(0)[:-1]))]
        self.gflat = torch.zeros_like(self.flat)                 # flat gradient buffer, same layout (padding stays 0)
        self.grad_views = [self.gflat[o:o + n].view_as(p) for o, n, p in zip(self.offsets, real, self.params)]
        self.m = torch.zeros_like(self.flat)
        self.v = torch.zeros_like(self.flat)
        offsets = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int32, device=dev)
        var_of = torch.repeat_interleave(torch.arange(len(sizes), device=dev), torch.tensor(sizes, device=dev))
        self.var_of = var_of.to(torch.int32)
        self.rel = Relation(self.var_of, len(sizes), rowptr=offsets, perm=None, is_sorted=True)
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.hyper = (float(lr), float(beta1), float(beta2), float(eps), float(clipnorm or 0.0))
        self._F = F

    def step(self, grads, accumulate=False):
        """grads: per-variable gradients (None allowed).  With `accumulate` they are ADDED to the flat gradient
        buffer, which the weight-gradient sink has already accumulated into (after `zero_grad()`)."""
        if not accumulate:
            self.zero_grad()
        pairs = [(v, gr) for v, gr in zip(self.grad_views, grads) if gr is not None]
        if pairs:
            torch._foreach_add_([v for v, _ in pairs], [gr.reshape(v.shape) for v, gr in pairs])
        self.step_flat()

    def zero_grad(self):
        self.gflat.zero_()

    def step_flat(self):
        from . import _lib
        from .distributed import allreduce_flat
        F = self._F
        g = allreduce_flat(self.gflat)                           # data parallel by graph: the step's only collective
        lr, b1, b2, eps, clip = self.hyper
        sumsq = F.raw_gms(g.unsqueeze(1), g.unsqueeze(1), self.rel, None) if clip > 0 else None
        # inside a captured step the plan-validation flags guard the update: a replayed batch that breaks the
        # recorded sortedness / index-range assumptions leaves parameters, moments and step counter untouched
        from .plan import current_mode
        mode = current_mode()
        guard = mode.flag_buf if (mode is not None and mode.kind == "capture") else None
        _lib.check(_lib.lib().kgcnn_adam_clipnorm(
            self.flat.data_ptr(), g.data_ptr(), self.m.data_ptr(), self.v.data_ptr(), self.var_of.data_ptr(),
            None if sumsq is None else sumsq.data_ptr(), self.total, lr, b1, b2, eps, clip,
            self.step_count.data_ptr(), None if guard is None else guard.data_ptr(),
            None if guard is None else mode.mask_buf.data_ptr(), 0 if guard is None else guard.numel(),
            torch.cuda.current_stream().cuda_stream), "kgcnn_adam_clipnorm")


class ForceTrainStep:
    def __init__(self, model, lr=1e-3, force_loss_factor=200.0, clipnorm=1.0, eps=1e-7):
        self.model = model
        self.w_e = 1.0 / force_loss_factor
        self.w_f = 1.0 - 1.0 / force_loss_factor
        self.clipnorm = clipnorm
        self.lr, self.eps = lr, eps
        self.params = None
        self.optimizer = None
        self.sink = None

    def _ensure_optimizer(self):
        if self.optimizer is None:
            self.params = [p for p in self.model.parameters() if p.requires_grad]
            self.optimizer = FlatAdam(self.params, lr=self.lr, eps=self.eps, clipnorm=self.clipnorm)
            # data parallel by graph: every replica starts from rank 0's weights (layers build lazily, so this
            # is the first point at which all variables exist); one broadcast of the flat buffer
            from .distributed import broadcast_flat
            broadcast_flat(self.optimizer.flat)

    def loss(self, inputs, e_target, f_target):
        out = self.model(inputs, training=True)
        eng, force = out[0], out[-1]
        return self.w_e * keras_mse(eng, e_target) + self.w_f * keras_mse(force, f_target)

    def step(self, inputs, e_target, f_target):
        loss = self.loss(inputs, e_target, f_target)
        self._ensure_optimizer()
        from .functional import WeightGradSink, no_coordinate_hessian, weight_grad_sink
        opt = self.optimizer
        if self.sink is None:
            self.sink = WeightGradSink(self.params, opt.gflat, opt.offsets)
        opt.zero_grad()
        # weights only: no d2/dx2 through the fused geometry kernels; the Dense weight gradients (leaves of this
        # sweep) are deferred into the sink and computed by ONE grouped tensor-core launch at its exit
        with no_coordinate_hessian(), weight_grad_sink(self.sink):
            grads = torch.autograd.grad(loss, self.params, allow_unused=True)
        opt.step(grads, accumulate=True)
        return loss


class StaticBatch:
    """Pinned host staging + fixed device buffers for one batch signature (the shapes a CUDA graph was
    captured for).  `load()` is asynchronous on the current stream: host arrays -> pinned -> device."""

    def __init__(self, example, device, device_angles=None):
        """device_angles: None, or the `edge_pairing` with which the angle triples are rebuilt ON THE DEVICE from
        the pair list after every load (gcnn_keras_b200.graph.get_angle_indices) instead of being copied from
        the host -- at C4 the triples are 97 % of the bytes of a batch."""
        self.device = device
        self.device_angles = device_angles
        self.shapes = {k: (tuple(v.shape), v.dtype) for k, v in example.items()}
        derived = ("angle_index", "angle_splits") if device_angles else ()
        self.pinned = {k: torch.from_numpy(v.copy()).pin_memory() for k, v in example.items() if k not in derived}
        self.dev = {k: t.to(device) for k, t in self.pinned.items()}
        for k in derived:
            self.dev[k] = torch.zeros(example[k].shape, dtype=torch.from_numpy(example[k][:0]).dtype, device=device)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.pinned.values())

    def signature_matches(self, batch):
        return all(k in batch and tuple(batch[k].shape) == s for k, (s, _) in self.shapes.items())

    def load(self, batch=None):
        if batch is not None:
            for k, t in self.pinned.items():
                t.copy_(torch.from_numpy(batch[k]))
        for k, t in self.pinned.items():
            self.dev[k].copy_(t, non_blocking=True)
        if self.device_angles:
            self._derive_angles()

    def _derive_angles(self):
        from .graph import get_angle_indices
        from .ragged import RaggedTensor, RowPartition
        dev = self.dev
        nodes = RaggedTensor(dev["xyz"], RowPartition(dev["node_splits"], nvals=int(dev["xyz"].shape[0])))
        index = RaggedTensor(dev["index"], RowPartition(dev["edge_splits"], nvals=int(dev["index"].shape[0])))
        ang = get_angle_indices(index, nodes, edge_pairing=self.device_angles)
        if tuple(ang.values.shape) != tuple(dev["angle_index"].shape):
            raise ValueError("batch has %d angle triples, the captured signature has %d"
                             % (int(ang.values.shape[0]), int(dev["angle_index"].shape[0])))
        dev["angle_index"].copy_(ang.values)
        dev["angle_splits"].copy_(ang.row_splits)


def molecule_batch_to_arrays(b):
    """synthetic / dataset batch dict -> flat arrays incl. int64 row_splits (host side, NumPy)."""
    import numpy as np
    out = {"z": b["z"].astype(np.int64), "xyz": b["xyz"].astype(np.float32), "index": b["index"].astype(np.int64),
           "node_splits": np.concatenate([[0], np.cumsum(b["node_len"])]).astype(np.int64),
           "edge_splits": np.concatenate([[0], np.cumsum(b["edge_len"])]).astype(np.int64),
           "energy": b["energy"].astype(np.float32), "force": b["force"].astype(np.float32)}
    if "angle_index" in b:
        out["angle_index"] = b["angle_index"].astype(np.int64)
        out["angle_splits"] = np.concatenate([[0], np.cumsum(b["angle_len"])]).astype(np.int64)
    return out


def default_model_inputs(dev, max_nodes=None):
    """[z, xyz, edge_index(, angle_index)] RaggedTensors over the device buffers, sharing ONE node partition."""
    from .ragged import RaggedTensor, RowPartition
    node_part = RowPartition(dev["node_splits"], nvals=int(dev["xyz"].shape[0]), max_len=max_nodes)
    edge_part = RowPartition(dev["edge_splits"], nvals=int(dev["index"].shape[0]))
    inputs = [RaggedTensor(dev["z"], node_part), RaggedTensor(dev["xyz"], node_part),
              RaggedTensor(dev["index"], edge_part)]
    if "angle_index" in dev:
        inputs.append(RaggedTensor(dev["angle_index"], RowPartition(dev["angle_splits"],
                                                                    nvals=int(dev["angle_index"].shape[0]))))
    return inputs, dev["energy"], RaggedTensor(dev["force"], node_part)


class CapturedForceTrainStep:
    """The whole training step (index shift + plan, forward, dE/dx, loss, weight gradients through the force,
    clipnorm, Adam) captured ONCE in a CUDA graph and replayed per batch of the same signature.

    Sortedness of every relation is detected during an eager warm-up (one sync each), trusted during capture,
    and re-validated after every replay from device-side flags (`check()`), so no host sync sits inside a step."""

    def __init__(self, train_step, example_batch, device, make_inputs=default_model_inputs, warmup=3,
                 device_angles=None):
        from . import _lib
        from .plan import PlanMode, clear_plan_cache, plan_mode
        self.train_step = train_step
        self.static = StaticBatch(molecule_batch_to_arrays(example_batch), device, device_angles=device_angles)
        self.make_inputs = make_inputs
        self.max_nodes = int(example_batch["node_len"].max())
        self.static.load()
        record = PlanMode("record")
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):
            for i in range(warmup):
                clear_plan_cache()
                mode = PlanMode("record")
                with plan_mode(mode):
                    inputs, e_t, f_t = self.make_inputs(self.static.dev, self.max_nodes)
                    train_step.step(inputs, e_t, f_t)
                record = mode
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        clear_plan_cache()
        self.mode = PlanMode("capture", hints=record.hints, device=device)
        self.graph = torch.cuda.CUDAGraph()
        _lib.launch_count_reset()
        with plan_mode(self.mode):
            with torch.cuda.graph(self.graph):
                self.mode.flag_buf.zero_()                   # guard words are OR-accumulated per replay
                inputs, e_t, f_t = self.make_inputs(self.static.dev, self.max_nodes)
                self.loss = train_step.step(inputs, e_t, f_t)
        self.mode.finalize_mask()
        self.flags = self.mode.flag_buf[:len(self.mode.flags)] if self.mode.flags else None
        self._flags_event, self._flags_pending = torch.cuda.Event(), False
        self.kernels_per_step = _lib.launch_count()      # our kernels recorded in the graph
        self.loss_host = torch.zeros((), dtype=torch.float32).pin_memory()
        self.flags_host = None if self.flags is None else torch.zeros_like(self.flags, device="cpu").pin_memory()

    def run(self, batch=None, load=True):
        """H2D (if load) -> replay -> async D2H of the loss (+ validation flags).  Returns the pinned loss."""
        if batch is not None:
            if not self.static.signature_matches(molecule_batch_to_arrays(batch)):
                raise ValueError("batch signature differs from the captured one; capture a new "
                                 "CapturedForceTrainStep")
            if int(batch["node_len"].max()) > self.max_nodes:
                raise ValueError("batch has a molecule of %d atoms, the captured padded outputs hold %d"
                                 % (int(batch["node_len"].max()), self.max_nodes))
        # the previous replay's validation flags, if they have landed: a violated plan assumption raises here
        # (the device-side guard already kept that step from touching the weights)
        if self._flags_pending and self._flags_event.query():
            self._flags_pending = False
            if not self.check():
                raise RuntimeError("a replayed batch violated the sortedness / index-range assumptions of the "
                                   "captured plan; the optimizer update of that step was skipped")
        if load:
            self.static.load(None if batch is None else molecule_batch_to_arrays(batch))
        self.graph.replay()
        self.loss_host.copy_(self.loss.detach(), non_blocking=True)
        if self.flags is not None:
            self.flags_host.copy_(self.flags, non_blocking=True)
            self._flags_event.record()
            self._flags_pending = True
        return self.loss_host

    def check(self):
        """After a synchronize: were the sortedness / index-range assumptions of the capture still true?"""
        if self.flags is None:
            return True
        return not self.mode.violated(list(self.flags_host))

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


def weighted_mse(pred, target, sample_weight, inv_graphs):
    """Keras mean-squared error over the molecules with weight 1: sum_b w_b |pred_b - target_b|^2 / (n_real * elements
    per molecule).  With all weights 1 and inv_graphs = 1 / B this is keras_mse."""
    if isinstance(pred, torch.Tensor) and hasattr(target, "to_tensor"):
        target = target.to_tensor()
    pred = pred if isinstance(pred, torch.Tensor) else pred.values
    target = target if isinstance(target, torch.Tensor) else target.values
    err = (pred - target) ** 2
    per_graph = err.reshape(err.shape[0], -1)
    return (per_graph * sample_weight.unsqueeze(1)).sum() * (inv_graphs.reshape(()) / per_graph.shape[1])


GHOST_ATOMS = 3     # filler atoms a capacity-padded batch always carries (a triangle: pairs and triples stay regular)


class BatchCapacity:
    """Static sizes a captured step is built for: `graphs` real molecules at most, and TOTAL atoms / pairs / triples
    including the filler molecule (see pad_arrays_to_capacity)."""

    def __init__(self, graphs, nodes, edges, angles=None):
        self.graphs, self.nodes, self.edges = int(graphs), int(nodes), int(edges)
        self.angles = None if angles is None else int(angles)

    @classmethod
    def covering(cls, batches, slack=0.0):
        """Smallest capacity that holds every batch of `batches` (dicts of synthetic.* / dataset batches), plus
        `slack` (fraction) of head room."""
        up = lambda v: int(v * (1.0 + slack)) + 1
        g = max(len(b["node_len"]) for b in batches)
        n = max(int(b["node_len"].sum()) for b in batches) + GHOST_ATOMS
        e = max(int(b["edge_len"].sum()) for b in batches)
        a = max(int(b["angle_len"].sum()) for b in batches) if "angle_len" in batches[0] else None
        return cls(g, up(n), up(e), None if a is None else up(a))

    def fits(self, arrays):
        ok = (len(arrays["node_splits"]) - 1 <= self.graphs and arrays["xyz"].shape[0] + GHOST_ATOMS <= self.nodes
              and arrays["index"].shape[0] <= self.edges)
        if self.angles is not None:
            ok = ok and arrays["angle_index"].shape[0] <= self.angles
        return ok


def pad_arrays_to_capacity(a, cap):
    """Bring a batch (molecule_batch_to_arrays) to the static shapes of `cap` without changing its result.

    The batch becomes cap.graphs + 1 molecules: the real ones, empty ones up to cap.graphs, and ONE filler molecule
    that owns every unused atom / pair / triple slot.  The filler atoms sit on a zig-zag line 1.5 A apart (finite
    distances and derivatives everywhere), its pairs and triples connect filler atoms only and keep the receiver
    column ascending, so it never touches a real atom; its loss weight is 0, so it contributes exact zeros to every
    gradient.  Adds `sample_weight` [graphs + 1] and `inv_graphs` [1] (= 1 / number of real molecules)."""
    import numpy as np
    if not cap.fits(a):
        raise ValueError("batch does not fit the captured capacity (graphs %d, atoms %d + %d filler, pairs %d%s)"
                         % (cap.graphs, cap.nodes - GHOST_ATOMS, GHOST_ATOMS, cap.edges,
                            "" if cap.angles is None else ", triples %d" % cap.angles))
    B, N, E = len(a["node_splits"]) - 1, a["xyz"].shape[0], a["index"].shape[0]
    G = cap.nodes - N
    k = np.arange(G)
    out = {}
    out["z"] = np.concatenate([a["z"], np.full((G,) + a["z"].shape[1:], a["z"][0] if N else 1, dtype=a["z"].dtype)])
    ghost_xyz = np.stack([1.5 * k, 0.8 * (k % 2), 0.3 * (k % 3)], axis=1).astype(np.float32)
    out["xyz"] = np.concatenate([a["xyz"], ghost_xyz])
    out["force"] = np.concatenate([a["force"], np.zeros((G, 3), dtype=np.float32)])
    out["energy"] = np.concatenate([a["energy"], np.zeros((cap.graphs + 1 - B,) + a["energy"].shape[1:], np.float32)])

    def splits(s, total):
        return np.concatenate([s, np.full((cap.graphs - B,), s[-1], dtype=np.int64), [total]]).astype(np.int64)

    def filler(count, width):          # rows (r, r+1, r+2, ...) mod G with r ascending: local ids of the filler molecule
        r = (np.arange(count) * G) // max(count, 1)
        return np.stack([(r + c) % G for c in range(width)], axis=1).astype(np.int64)

    out["node_splits"] = splits(a["node_splits"], cap.nodes)
    out["index"] = np.concatenate([a["index"], filler(cap.edges - E, 2)])
    out["edge_splits"] = splits(a["edge_splits"], cap.edges)
    if cap.angles is not None:
        A = a["angle_index"].shape[0]
        out["angle_index"] = np.concatenate([a["angle_index"], filler(cap.angles - A, 3)])
        out["angle_splits"] = splits(a["angle_splits"], cap.angles)
    w = np.zeros((cap.graphs + 1,), dtype=np.float32)
    w[:B] = 1.0
    out["sample_weight"] = w
    out["inv_graphs"] = np.array([1.0 / max(B, 1)], dtype=np.float32)
    return out


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
        # flat gradient buffer, same layout (padding stays 0).  Data parallel over NCCL ranks: the buffer lives in
        # peer-mapped memory and the all-reduce is our own kernel over NVLink (distributed.PeerAllReduce)
        from .distributed import PeerAllReduce, world_info
        self.world = world_info()[1]
        # with more than one rank the gradient buffer carries one more (256 B aligned) slot behind the last variable:
        # word [total] is the captured step's plan-validation verdict, reduced with the gradients (step_flat)
        self.gtotal = self.total + (64 if self.world > 1 else 0)
        self.peer = PeerAllReduce.create(self.gtotal, dev)
        self.gflat = self.peer.buffer if self.peer is not None else torch.zeros(self.gtotal, dtype=torch.float32,
                                                                                device=dev)
        self.all_bits = torch.full((1,), -1, dtype=torch.int32, device=dev)
        self.reduced_guard = None          # after a captured step at world > 1: view of the reduced verdict word
        self.grad_views = [self.gflat[o:o + n].view_as(p) for o, n, p in zip(self.offsets, real, self.params)]
        self.m = torch.zeros_like(self.flat)
        self.v = torch.zeros_like(self.flat)
        offsets = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int32, device=dev)
        var_of = torch.repeat_interleave(torch.arange(len(sizes), device=dev), torch.tensor(sizes, device=dev))
        self.var_of = var_of.to(torch.int32)
        self.rel = Relation(self.var_of, len(sizes), rowptr=offsets, perm=None, is_sorted=True)
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.hyper = (float(lr), float(beta1), float(beta2), float(eps), float(clipnorm or 0.0))
        self.lr_device = torch.full((1,), float(lr), dtype=torch.float32, device=dev)   # read by the update kernel
        self._F = F

    def set_lr(self, lr):
        """Learning-rate schedules (kgcnn/training/scheduler.py): the kernel reads the rate from device memory, so
        a captured step follows the schedule without being re-captured."""
        self.hyper = (float(lr),) + self.hyper[1:]
        self.lr_device.fill_(float(lr))

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
        # inside a captured step the plan-validation flags guard the update: a replayed batch that breaks the
        # recorded sortedness / index-range assumptions leaves parameters, moments and step counter untouched
        from .plan import current_mode
        mode = current_mode()
        guard = mode.flag_buf if (mode is not None and mode.kind == "capture") else None
        guard_ptr = None if guard is None else guard.data_ptr()
        mask_ptr = None if guard is None else mode.mask_buf.data_ptr()
        n_guard = 0 if guard is None else guard.numel()
        stream = torch.cuda.current_stream().cuda_stream
        if guard is not None and self.world > 1:
            # the verdict must be the GROUP's: this rank's word rides behind the gradients through the all-reduce
            _lib.check(_lib.lib().kgcnn_guard_fold(guard_ptr, mask_ptr, n_guard, self.gflat[self.total:].data_ptr(),
                                                   stream), "kgcnn_guard_fold")
        # data parallel by graph: the step's only collective
        g = self.peer.reduce() if self.peer is not None else allreduce_flat(self.gflat)
        if guard is not None and self.world > 1:
            # mean of the 0/1 words: non-zero bits on every rank iff any rank tripped -> all replicas skip together
            self.reduced_guard = g[self.total:self.total + 1]
            guard_ptr, mask_ptr, n_guard = self.reduced_guard.data_ptr(), self.all_bits.data_ptr(), 1
        g = g[:self.total]
        lr, b1, b2, eps, clip = self.hyper
        sumsq = F.raw_gms(g.unsqueeze(1), g.unsqueeze(1), self.rel, None) if clip > 0 else None
        _lib.check(_lib.lib().kgcnn_adam_clipnorm(
            self.flat.data_ptr(), g.data_ptr(), self.m.data_ptr(), self.v.data_ptr(), self.var_of.data_ptr(),
            None if sumsq is None else sumsq.data_ptr(), self.total, lr, self.lr_device.data_ptr(), b1, b2, eps, clip,
            self.step_count.data_ptr(), guard_ptr, mask_ptr, n_guard, stream), "kgcnn_adam_clipnorm")


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

    def loss(self, inputs, e_target, f_target, weights=None):
        out = self.model(inputs, training=True)
        eng, force = out[0], out[-1]
        if weights is None:
            return self.w_e * keras_mse(eng, e_target) + self.w_f * keras_mse(force, f_target)
        # capacity-padded batch (pad_arrays_to_capacity): per-molecule weights (0 for the filler molecules) and
        # 1 / (number of real molecules) as DEVICE tensors, so one captured graph serves every batch that fits
        sample_weight, inv_graphs = weights
        return (self.w_e * weighted_mse(eng, e_target, sample_weight, inv_graphs)
                + self.w_f * weighted_mse(force, f_target, sample_weight, inv_graphs))

    def step(self, inputs, e_target, f_target, weights=None):
        from .functional import WeightGradSink, no_coordinate_hessian, static_weights, weight_grad_sink
        # the variables live in FlatAdam's buffer and its update kernel (followed by the step-counter kernel) is their
        # only writer between two steps: this step vouches for that, so the fused kernels may stage them early
        with static_weights(None if self.optimizer is None else self.optimizer.flat):
            loss = self.loss(inputs, e_target, f_target, weights)
        self._ensure_optimizer()
        opt = self.optimizer
        if self.sink is None:
            self.sink = WeightGradSink(self.params, opt.gflat, opt.offsets)
        opt.zero_grad()
        # weights only: no d2/dx2 through the fused geometry kernels; the Dense weight gradients (leaves of this
        # sweep) are deferred into the sink and computed by ONE grouped tensor-core launch at its exit
        with no_coordinate_hessian(), weight_grad_sink(self.sink), static_weights(opt.flat):
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
        # two pinned staging sets: the host may fill the next batch while the previous step's copy is still queued
        self._pinned_sets = [{k: torch.from_numpy(v.copy()).pin_memory() for k, v in example.items()
                              if k not in derived} for _ in range(2)]
        self._copied = [None, None]              # event after the H2D copies out of each set
        self._slot = 0
        self.pinned = self._pinned_sets[0]
        self.dev = {k: t.to(device) for k, t in self.pinned.items()}
        for k in derived:
            self.dev[k] = torch.zeros(example[k].shape, dtype=torch.from_numpy(example[k][:0]).dtype, device=device)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.pinned.values())

    def signature_matches(self, batch):
        return all(k in batch and tuple(batch[k].shape) == s for k, (s, _) in self.shapes.items())

    def load(self, batch=None):
        """batch given: stage it in the free pinned set (waiting only for THAT set's previous copy) and copy it to
        the device; batch None: copy the current set again."""
        if batch is not None:
            self._slot ^= 1
            self.pinned = self._pinned_sets[self._slot]
            if self._copied[self._slot] is not None:
                self._copied[self._slot].synchronize()
            for k, t in self.pinned.items():
                t.copy_(torch.from_numpy(batch[k]))
        for k, t in self.pinned.items():
            self.dev[k].copy_(t, non_blocking=True)
        if batch is not None:
            if self._copied[self._slot] is None:
                self._copied[self._slot] = torch.cuda.Event()
            self._copied[self._slot].record()
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
    if "sample_weight" in dev:      # capacity-padded batch: (inputs, energy, force, (weights, 1 / real molecules))
        return inputs, dev["energy"], RaggedTensor(dev["force"], node_part), (dev["sample_weight"], dev["inv_graphs"])
    return inputs, dev["energy"], RaggedTensor(dev["force"], node_part)


class CapturedForceTrainStep:
    """The whole training step (index shift + plan, forward, dE/dx, loss, weight gradients through the force,
    clipnorm, Adam) captured ONCE in a CUDA graph and replayed per batch of the same signature.

    Sortedness of every relation is detected during an eager warm-up (one sync each), trusted during capture,
    and re-validated after every replay from device-side flags (`check()`), so no host sync sits inside a step."""

    def __init__(self, train_step, example_batch, device, make_inputs=default_model_inputs, warmup=3,
                 device_angles=None, capacity=None, max_nodes=None, keep_weights=False):
        """capacity (BatchCapacity): capture at these static sizes; `run(batch)` then accepts EVERY batch that fits
        (pad_arrays_to_capacity brings it to the captured shapes; a model.fit over shuffled batches of different
        pair / triple counts, kgcnn/training scripts force_*.py, replays one graph).  max_nodes: width of the padded
        per-atom outputs (longest molecule of the data set; default: of the example batch).
        keep_weights: undo the warm-up steps' updates (weights, Adam moments, step counter) after the capture, so
        that the first `run` is the first update (model.fit semantics)."""
        from . import _lib
        from .plan import PlanMode, clear_plan_cache, plan_mode
        self.train_step = train_step
        self.capacity = capacity
        arrays = molecule_batch_to_arrays(example_batch)
        if capacity is not None:
            if device_angles:
                raise ValueError("capacity padding takes the angle triples from the host batch")
            arrays = pad_arrays_to_capacity(arrays, capacity)
        self.static = StaticBatch(arrays, device, device_angles=device_angles)
        self.make_inputs = make_inputs
        self.max_nodes = int(example_batch["node_len"].max()) if max_nodes is None else int(max_nodes)
        self.static.load()
        record = PlanMode("record")
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        snapshot = None
        with torch.cuda.stream(side):
            if keep_weights:
                with torch.no_grad():                            # the layers build their variables lazily
                    train_step.model(self.make_inputs(self.static.dev, self.max_nodes)[0], training=True)
                train_step._ensure_optimizer()
                opt = train_step.optimizer
                snapshot = [t.clone() for t in (opt.flat, opt.m, opt.v, opt.step_count)]
            for i in range(warmup):
                clear_plan_cache()
                mode = PlanMode("record")
                with plan_mode(mode):
                    train_step.step(*self.make_inputs(self.static.dev, self.max_nodes))
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
                self.loss = train_step.step(*self.make_inputs(self.static.dev, self.max_nodes))
        if snapshot is not None:
            opt = train_step.optimizer
            for dst, src in zip((opt.flat, opt.m, opt.v, opt.step_count), snapshot):
                dst.copy_(src)
        self.mode.finalize_mask()
        self.flags = self.mode.flag_buf[:len(self.mode.flags)] if self.mode.flags else None
        self._flags_event, self._flags_pending = torch.cuda.Event(), False
        self.kernels_per_step = _lib.launch_count()      # our kernels recorded in the graph
        self.loss_host = torch.zeros((), dtype=torch.float32).pin_memory()
        # more than one rank: the reduced verdict word (non-zero on EVERY rank when any rank's batch broke its plan)
        self.group_guard = getattr(train_step.optimizer, "reduced_guard", None)
        self.group_guard_host = None if self.group_guard is None else torch.zeros(1, dtype=torch.float32).pin_memory()
        self.flags_host = None if self.flags is None else torch.zeros_like(self.flags, device="cpu").pin_memory()

    def run(self, batch=None, load=True):
        """H2D (if load) -> replay -> async D2H of the loss (+ validation flags).  Returns the pinned loss."""
        arrays = None
        if batch is not None:
            arrays = molecule_batch_to_arrays(batch)
            if self.capacity is not None:
                arrays = pad_arrays_to_capacity(arrays, self.capacity)        # raises when the batch does not fit
            elif not self.static.signature_matches(arrays):
                raise ValueError("batch signature differs from the captured one; capture with capacity=BatchCapacity"
                                 "(...) to serve batches of different sizes")
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
            self.static.load(arrays)
        self.graph.replay()
        self.loss_host.copy_(self.loss.detach(), non_blocking=True)
        if self.flags is not None:
            self.flags_host.copy_(self.flags, non_blocking=True)
            if self.group_guard is not None:
                self.group_guard_host.copy_(self.group_guard, non_blocking=True)
            self._flags_event.record()
            self._flags_pending = True
        return self.loss_host

    def check(self):
        """After a synchronize: were the sortedness / index-range assumptions of the capture still true?"""
        if self.flags is None:
            return True
        if self.group_guard_host is not None and float(self.group_guard_host[0]) != 0.0:
            return False                   # some rank of the group tripped: every rank skipped, every rank reports it
        return not self.mode.violated(list(self.flags_host))


# ------------------------------------------------------------------------------------------------
# model.fit-shaped loop over shuffled batches (force_schnet.py:229-269 / force_painn.py / force_hdnnp2nd.py)
# ------------------------------------------------------------------------------------------------
def collate_molecules(mols):
    """List of per-molecule dicts {z, xyz, index[, angle_index], energy, force} -> one ragged batch dict (the layout of
    gcnn_keras_b200.synthetic.*_batch: flat values + row lengths)."""
    import numpy as np
    out = {"z": np.concatenate([m["z"] for m in mols]), "xyz": np.concatenate([m["xyz"] for m in mols]),
           "index": np.concatenate([m["index"] for m in mols]),
           "node_len": np.array([len(m["z"]) for m in mols], dtype=np.int64),
           "edge_len": np.array([len(m["index"]) for m in mols], dtype=np.int64),
           "energy": np.stack([np.asarray(m["energy"], dtype=np.float32).reshape(-1) for m in mols]),
           "force": np.concatenate([m["force"] for m in mols])}
    if "angle_index" in mols[0]:
        out["angle_index"] = np.concatenate([m["angle_index"] for m in mols])
        out["angle_len"] = np.array([len(m["angle_index"]) for m in mols], dtype=np.int64)
    return out


def linear_learning_rate(epoch, start=1e-3, stop=1e-5, epo_min=0, epo=500, eps=1e-8):
    """LinearLearningRateScheduler.schedule_epoch_lr, kgcnn/training/scheduler.py:275-292."""
    if epoch < epo_min:
        return max(float(start), eps)
    return max(float(start - (start - stop) / (epo - epo_min) * (epoch - epo_min)), eps)


def _worst_case_capacity(mols, batch_size):
    """Capacity no batch of `batch_size` molecules of this data set can exceed: the batch_size largest counts."""
    import numpy as np
    top = lambda v: int(np.sort(np.asarray(v))[::-1][:batch_size].sum())
    angles = top([len(m["angle_index"]) for m in mols]) if "angle_index" in mols[0] else None
    return BatchCapacity(batch_size, top([len(m["z"]) for m in mols]) + GHOST_ATOMS,
                         top([len(m["index"]) for m in mols]), angles)


def predict_energy_force(model, mols, batch_size, device):
    """Eager inference over a list of molecules: (energy [M, k], list of per-molecule force arrays)."""
    import numpy as np
    energies, forces = [], []
    for s in range(0, len(mols), batch_size):
        chunk = mols[s:s + batch_size]
        arrays = molecule_batch_to_arrays(collate_molecules(chunk))
        dev = {k: torch.from_numpy(v).to(device) for k, v in arrays.items()}
        inputs = default_model_inputs(dev, int(max(len(m["z"]) for m in chunk)))[0]
        out = model(inputs, training=False)
        e, f = out[0].detach().cpu().numpy(), out[-1].detach().cpu().numpy()
        energies.append(e.reshape(len(chunk), -1))
        forces += [f[i, :len(m["z"])] for i, m in enumerate(chunk)]
    return np.concatenate(energies), forces


def fit_force_model(train_step, train, batch_size, epochs, device, validation=None, shuffle=True, learning_rate=None,
                    seed=0, verbose=2, capture=True):
    """`model.fit(x, [energy, force], batch_size, epochs, shuffle=True, validation_data=...)` of the reference's
    force_*.py scripts for a ForceTrainStep: every epoch walks the (shuffled) molecules in batches; ONE CUDA graph,
    captured at the worst-case capacity of the data set, serves every batch including the short last one.

    train / validation: lists of per-molecule dicts (collate_molecules).  learning_rate: None or epoch -> rate
    (linear_learning_rate); applied through FlatAdam.set_lr, i.e. without re-capturing.
    Returns a Keras-History-like dict: loss per epoch (mean over batches), val_energy_mae / val_force_mae."""
    import time
    import numpy as np
    rng = np.random.default_rng(seed)
    max_nodes = int(max(len(m["z"]) for m in train))
    first = collate_molecules(train[:batch_size])
    runner = None
    if capture:
        runner = CapturedForceTrainStep(train_step, first, device, capacity=_worst_case_capacity(train, batch_size),
                                        max_nodes=max_nodes, warmup=1, keep_weights=True)
    history = {"loss": [], "val_energy_mae": [], "val_force_mae": [], "lr": [], "time": []}
    for epoch in range(epochs):
        t0 = time.time()
        if learning_rate is not None:
            lr = float(learning_rate(epoch))
            if train_step.optimizer is not None:
                train_step.optimizer.set_lr(lr)
            else:
                train_step.lr = lr
            history["lr"].append(lr)
        order = rng.permutation(len(train)) if shuffle else np.arange(len(train))
        losses = []
        for s in range(0, len(train), batch_size):
            batch = collate_molecules([train[i] for i in order[s:s + batch_size]])
            if runner is not None:
                runner.run(batch)
                losses.append(runner.loss.detach().clone())              # stays on the device until the epoch ends
            else:
                arrays = molecule_batch_to_arrays(batch)
                dev = {k: torch.from_numpy(v).to(device) for k, v in arrays.items()}
                losses.append(train_step.step(*default_model_inputs(dev, max_nodes)).detach())
        history["loss"].append(float(torch.stack(losses).mean()))        # one synchronisation per epoch
        if runner is not None and not runner.check():
            raise RuntimeError("a batch violated the plan assumptions of the captured step")
        if validation:
            e, f = predict_energy_force(train_step.model, validation, batch_size, device)
            e_true = np.stack([np.asarray(m["energy"], dtype=np.float32).reshape(-1) for m in validation])
            history["val_energy_mae"].append(float(np.mean(np.abs(e - e_true))))
            history["val_force_mae"].append(float(np.mean(np.concatenate(
                [np.abs(fp - m["force"]).reshape(-1) for fp, m in zip(f, validation)]))))
        history["time"].append(time.time() - t0)
        if verbose:
            msg = "Epoch %d/%d - %.2fs - loss: %.6f" % (epoch + 1, epochs, history["time"][-1], history["loss"][-1])
            if validation:
                msg += " - val_energy_mae: %.6f - val_force_mae: %.6f" % (history["val_energy_mae"][-1],
                                                                         history["val_force_mae"][-1])
            print(msg, flush=True)
    return history

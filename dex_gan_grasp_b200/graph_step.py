"""The training step as ONE CUDA graph: `update_dexgangrasp` (+ `update_dexevaluator`) of dexgangrasp.py:739-823 captured once
and replayed per batch.

The step is ~600 kernel launches (and, data parallel, ~80 small NCCL collectives) issued from Python; at 4096 / N rows per
GPU the launches, not the kernels, bound the step.  Capturing it removes the host from the loop: static input
buffers, the autograd forward / backward, the gradient and BatchNorm-statistics allreduces and the fused Adam kernels (step
count in device memory, `dexgg_adam_step_dev`) all replay from one `cudaGraphLaunch`.

    step = CapturedTrainStep(model, example_batch)        # 3 eager warm-up steps, then capture
    losses = step(batch)                                   # dict of device tensors (valid until the next call)

Numerics are those of the eager step (same kernels, same order).  The learning rate is a kernel argument frozen into the graph:
every call compares the optimizers' current group learning rates with the captured ones and RE-CAPTURES the graph when a
scheduler (`DexGanGrasp.update_learning_rate`, ReduceLROnPlateau once per epoch in train.py) has changed one.
"""
import torch

from . import train_ops


class CapturedTrainStep:
    def __init__(self, model, example_batch, with_evaluator=True, is_train_gen=True, warmup=3, example_z=None):
        if warmup < 1:
            raise ValueError("at least one eager warm-up step is needed: it creates the Adam moments and the gradient buffers "
                             "outside the capture (a zero-fill captured in the graph would reset the moments at every replay)")
        self.model = model
        self.with_evaluator = bool(with_evaluator)
        self.is_train_gen = bool(is_train_gen)
        dev = model.device
        self.static = {k: (v.to(dev).clone() if torch.is_tensor(v) else v) for k, v in example_batch.items()}
        # optional caller-supplied latent batch (parity tests); otherwise the step draws z from the CUDA generator in the graph
        self.static_z = example_z.to(dev).clone() if example_z is not None else None
        self.opts = [model.optim_dexgangrasp_discriminator] + ([model.optim_dexgangrasp_generator] if is_train_gen else []) + \
                    ([model.optim_dexevaluator] if with_evaluator else [])
        # gradients are zeroed in place instead of dropped, so their addresses (and with them the Adam chunk table built in the
        # warm-up step) stay the same inside the capture: no host -> device table copy gets captured
        model._zero_grad_to_none = False
        for o in self.opts:
            o.device_step = True
            if getattr(o, "_step_dev", None) is None:  # device-side step counter, created outside the capture
                taken = [int(st["step"]) for st in o.state.values() if "step" in st]
                o._step_dev = torch.tensor([taken[0] if taken else 0], dtype=torch.int32, device=dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = None
        self._capture()
        self.eager_steps = warmup
        mods = [model.DexGANGrasp] + ([model.DexEvaluator] if self.with_evaluator else [])
        self._trained_tensors = [t for m in mods for t in list(m.parameters()) + list(m.buffers()) if t.is_floating_point()]

    def _body(self):
        m = self.model
        if self.with_evaluator and hasattr(m, "_use_engine") and m._use_engine() and "label" in self.static:
            # fused engine: GAN + evaluator update as ONE schedule (gradient exchanges overlap the other network's compute)
            m.DexGANGrasp.train()
            m.DexEvaluator.train()
            z = self.static_z
            if z is None:
                z = torch.randn((self.static["bps_object"].shape[0], m.DexGANGrasp.latentD), dtype=torch.float32, device=m.device)
            out = m._engine(int(self.static["bps_object"].shape[0])).combined_step(dict(self.static), z, self.is_train_gen)
            return {k: (v.detach() if torch.is_tensor(v) else v) for k, v in out.items()}
        b = dict(self.static)  # the step mutates the dict (rot_matrix view), never the static tensors
        out = dict(self.model.update_dexgangrasp(b, is_train_gen=self.is_train_gen, Zgen=self.static_z))
        if self.with_evaluator:
            out.update(self.model.update_dexevaluator(dict(self.static)))
        return {k: (v.detach() if torch.is_tensor(v) else v) for k, v in out.items()}

    def _lrs(self):
        return tuple(float(g["lr"]) for o in self.opts for g in o.param_groups)

    def _capture(self):
        from . import _native, dataparallel
        L = _native.lib()
        if getattr(self, "graph", None) is not None:
            torch.cuda.synchronize(self.model.device)
            self.graph.reset()
        n0 = L.dexgg_launch_count()
        c0 = dataparallel.collective_counters()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses = self._body()
        self.kernels_per_replay = int(L.dexgg_launch_count() - n0)  # dexgg kernel nodes in the graph (host counter during capture)
        c1 = dataparallel.collective_counters()
        self.collectives_per_replay = {k: c1[k] - c0[k] for k in c1}  # collectives (calls / bytes per kind) in the graph
        self.captured_lrs = self._lrs()
        for o in self.opts:  # the capture ran the Python side of the step (host counters + 1) without executing a kernel
            for st in o.state.values():
                if "step" in st:
                    st["step"] -= 1

    def __call__(self, batch, z=None):
        if self._lrs() != self.captured_lrs:  # a scheduler changed a learning rate: the frozen kernel argument is stale
            self._capture()
            self.recaptures = getattr(self, "recaptures", 0) + 1
        for k, v in batch.items():
            if torch.is_tensor(v):
                self.static[k].copy_(v, non_blocking=True)
        if z is not None:
            self.static_z.copy_(z, non_blocking=True)
        self.graph.replay()
        for o in self.opts:  # keep torch.optim.Adam's host-side step counters (state_dict layout) in step with the device
            for st in o.state.values():
                if "step" in st:
                    st["step"] += 1
        # the replay rewrote weights and BatchNorm buffers on the device behind torch's back: bump their version counters so
        # the packed eval-mode copies (networks.NativeNet.sync) are rebuilt by the next eval-mode forward
        torch.autograd.graph.increment_version(self._trained_tensors)
        return self.losses

    def close(self):
        """Destroy the graph (it holds references to the NCCL communicator's captured work: `destroy_process_group` must come
        after this).  The model keeps training eagerly afterwards with the device-side step counters."""
        torch.cuda.synchronize(self.model.device)
        self.losses = None
        if self.graph is not None:
            self.graph.reset()
            self.graph = None

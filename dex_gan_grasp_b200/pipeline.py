"""Batched grasp synthesis: point clouds -> BPS -> DexGenerator -> rot6d/offset -> DexEvaluator -> per-object top-k.

The batched form of the reference's per-object loop (dexgangrasp_real_robot.py:131-162, dexgangrasp_offline.py:147-162):
bps.encode -> generate_grasps(n) -> filter_grasps(thresh) -> first k rows.  Everything stays on the device between the
stages; the encoding is used once per object (no np.tile); objects shard over ranks with no collective.
"""
import numpy as np
import torch

from . import _native
from . import bps as bps_mod


class GraspPipeline:
    def __init__(self, gan, evaluator, basis=None, k=10, thresh=0.5, z_offset=0.025, use_graph=False):
        """gan: networks.DexGANGrasp, evaluator: networks.DexEvaluator (both on the same CUDA device, eval mode).
        use_graph: run_device replays ONE captured CUDA graph per (input buffers, shapes) instead of issuing the ~9 launches
        from Python - for small per-call work (one object, or 1024 / N objects per GPU) the host would otherwise bound the
        step.  The returned tensors are then the graph's static outputs: read them before the next call with the same inputs."""
        self.gan, self.eva = gan, evaluator
        self.device = gan.device
        if basis is None:
            basis = np.load(bps_mod.BASIS_PATH)
        self.basis = torch.as_tensor(np.asarray(basis), dtype=torch.float32).to(self.device).contiguous()
        self.k, self.thresh, self.z_offset = int(k), float(thresh), float(z_offset)
        self._ws = bps_mod._Workspace()
        self._stage = {}
        self._copy_stream = None
        self.use_graph = bool(use_graph)
        self._graphs = {}
        gan.eval()
        evaluator.eval()

    # ------------------------------------------------------------------ device-resident path
    def encode(self, clouds):
        """clouds [B,N,3] f32 cuda -> BPS distances [B,4096]"""
        d, _ = bps_mod.bps_encode_device(self.basis, clouds, return_idx=False, ws=self._ws)
        return d

    def run_device(self, clouds, z, samples_per_object, z_ready=None):
        if self.use_graph and z_ready is None:
            return self._run_graph(clouds, z, samples_per_object)
        return self._run_eager(clouds, z, samples_per_object, z_ready)

    def _weights_key(self):
        return tuple((t.data_ptr(), t._version) for net in (self.gan.generator, self.eva) for t in net._native._tensors())

    def _run_graph(self, clouds, z, samples_per_object):
        """One cudaGraphLaunch per call: the graph is captured the first time this (cloud buffer, latent buffer, shape) is seen
        and re-captured when the networks' parameters change."""
        key = (clouds.data_ptr(), tuple(clouds.shape), z.data_ptr(), tuple(z.shape), int(samples_per_object))
        wkey = self._weights_key()
        ent = self._graphs.get(key)
        if ent is None or ent[2] != wkey:
            dev = self.device
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):  # eager warm-up: packs the weights, grows the workspaces
                self._run_eager(clouds, z, samples_per_object, None)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = self._run_eager(clouds, z, samples_per_object, None)
            if len(self._graphs) >= 8:
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._graphs[key] = (graph, out, wkey)
        ent[0].replay()
        return ent[1]

    def _run_eager(self, clouds, z, samples_per_object, z_ready=None):
        """clouds [B,N,3], z [B*S,latentD] on the device.  Returns a dict of device tensors:
        idx [B,k] i32 (global row ids, -1 padded), count [B], score [B,k], rot_matrix [B,k,3,3], transl [B,k,3],
        joint_conf [B,k,nJ] (views of `packed`; rows beyond count are zero), bps [B,4096], p [M,1]."""
        S = int(samples_per_object)
        B = clouds.shape[0]
        M = B * S
        assert z.shape[0] == M
        L = _native.lib()
        dev = self.device
        enc = self.encode(clouds)
        if z_ready is not None:
            torch.cuda.current_stream(dev).wait_event(z_ready)
        rot, transl, joint, _ = self.gan.generator.native_forward(z, enc, self.z_offset, False)
        p = self.eva.native_forward(enc, rot, transl)
        k = min(self.k, S)
        nJ = joint.shape[1]
        idx = torch.empty((B, k), dtype=torch.int32, device=dev)
        cnt = torch.empty((B,), dtype=torch.int32, device=dev)
        packed = torch.empty((B, k, 14 + nJ), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):  # threshold + per-object top-k + gather of the kept rows: one launch (filter_topk.cu)
            _native.check(L.dexgg_filter_topk_gather(p.data_ptr(), M, S, self.thresh, k, rot.data_ptr(), transl.data_ptr(),
                                                     joint.data_ptr(), nJ, idx.data_ptr(), cnt.data_ptr(), packed.data_ptr(),
                                                     _native.stream_ptr(dev)))
        return {
            "idx": idx, "count": cnt, "packed": packed,  # packed [B,k,14+nJ]: score | row id (int32 bits) | rot 9 | transl 3 | joints
            "score": packed[:, :, 0], "rot_matrix": packed[:, :, 2:11].unflatten(2, (3, 3)), "transl": packed[:, :, 11:14],
            "joint_conf": packed[:, :, 14:], "bps": enc, "p": p,
        }

    # ------------------------------------------------------------------ host-buffer path (the e2e number of bench.py)
    def run_host(self, clouds_host, z_host, samples_per_object):
        """clouds_host / z_host: pinned CPU tensors.  Copies them to the device, runs the path and reads the top-k result
        back to the host (one packed [B,k,14+nJ] block: score | row id as int32 bits | rot 9 | transl 3 | joints, + counts;
        `unpack_host` splits it).  Returns (dict of CPU tensors, h2d bytes, d2h bytes)."""
        dev = self.device
        key = (tuple(clouds_host.shape), tuple(z_host.shape))
        if key not in self._stage:
            self._stage = {key: (torch.empty(clouds_host.shape, dtype=torch.float32, device=dev),
                                 torch.empty(z_host.shape, dtype=torch.float32, device=dev))}
        d_clouds, d_z = self._stage[key]
        # clouds first (the BPS stage needs only them); the latent batch is copied on a side stream underneath the BPS kernel
        main = torch.cuda.current_stream(dev)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(dev)
        d_clouds.copy_(clouds_host, non_blocking=True)
        self._copy_stream.wait_stream(main)  # d_z may still be read by the previous step
        with torch.cuda.stream(self._copy_stream):
            d_z.copy_(z_host, non_blocking=True)
            z_ready = torch.cuda.Event()
            z_ready.record(self._copy_stream)
        out = self.run_device(d_clouds, d_z, samples_per_object, z_ready=z_ready)
        packed = out["packed"]
        host = packed.cpu()  # synchronising D2H read of the result
        cnt = out["count"].cpu()
        h2d = clouds_host.numel() * 4 + z_host.numel() * 4
        d2h = host.numel() * 4 + cnt.numel() * 4
        return {"packed": host, "count": cnt}, h2d, d2h

    def run_host_pipelined(self, steps, samples_per_object):
        """Streaming form of run_host for a sequence of steps: `steps` is an iterable of (clouds_host, z_host) pinned CPU
        tensors of one shape.  The inputs of step i+1 cross PCIe (H2D copy engine, own stream) while step i computes, and the
        packed top-k result of step i returns on a third stream (D2H copy engine) while step i+1 computes; two device input
        sets and two pinned result buffers rotate.  Yields ({"packed", "count"} CPU tensors, h2d bytes, d2h bytes) per step,
        in order, each after its result has landed on the host."""
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_h2d", None) is None:
            self._h2d, self._d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        S = int(samples_per_object)
        if getattr(self, "_pl_bufs", None) is None:
            self._pl_bufs = {}    # rotating device input sets / pinned result buffers, kept across calls (keyed by shape below)
        d_in, h_out = [None, None], [None, None]
        done = [None, None]       # compute of the step that last used input set j has finished
        pending = None            # (event, host tensors, byte counts) of the previous step
        for i, (clouds_host, z_host) in enumerate(steps):
            j = i & 1
            if d_in[j] is None:
                key = ("in", j, tuple(clouds_host.shape), tuple(z_host.shape))
                if key not in self._pl_bufs:
                    self._pl_bufs[key] = (torch.empty(clouds_host.shape, dtype=torch.float32, device=dev),
                                          torch.empty(z_host.shape, dtype=torch.float32, device=dev))
                d_in[j] = self._pl_bufs[key]
            d_clouds, d_z = d_in[j]
            if done[j] is not None:
                self._h2d.wait_event(done[j])
            else:
                self._h2d.wait_stream(main)
            with torch.cuda.stream(self._h2d):
                d_clouds.copy_(clouds_host, non_blocking=True)
                d_z.copy_(z_host, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self._h2d)
            main.wait_event(ready)
            out = self.run_device(d_clouds, d_z, S)
            packed = out["packed"]
            cnt = out["count"]
            done[j] = torch.cuda.Event()
            done[j].record(main)
            if h_out[j] is None:
                key = ("out", j, tuple(packed.shape))
                if key not in self._pl_bufs:
                    self._pl_bufs[key] = (torch.empty(packed.shape, dtype=torch.float32).pin_memory(),
                                          torch.empty(cnt.shape, dtype=torch.int32).pin_memory())
                h_out[j] = self._pl_bufs[key]
            self._d2h.wait_event(done[j])
            with torch.cuda.stream(self._d2h):
                h_out[j][0].copy_(packed, non_blocking=True)
                h_out[j][1].copy_(cnt, non_blocking=True)
                landed = torch.cuda.Event()
                landed.record(self._d2h)
            packed.record_stream(self._d2h)
            cnt.record_stream(self._d2h)
            h2d = clouds_host.numel() * 4 + z_host.numel() * 4
            d2h = packed.numel() * 4 + cnt.numel() * 4
            if pending is not None:
                pending[0].synchronize()
                yield pending[1], pending[2], pending[3]
            pending = (landed, {"packed": h_out[j][0], "count": h_out[j][1]}, h2d, d2h)
        if pending is not None:
            pending[0].synchronize()
            yield pending[1], pending[2], pending[3]


def unpack_host(packed, count=None):
    """Splits a packed result block (CPU or device tensor [B,k,14+nJ]) into named views; row ids come back exact (int32)."""
    ids = packed.view(torch.int32)[:, :, 1]
    out = {"score": packed[:, :, 0], "idx": ids, "rot_matrix": packed[:, :, 2:11].unflatten(2, (3, 3)),
           "transl": packed[:, :, 11:14], "joint_conf": packed[:, :, 14:]}
    if count is not None:
        out["count"] = count
    return out

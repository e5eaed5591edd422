"""Host-side mirror of DexGanGrasp/models/networks.py on the sm_100a kernels of libdexgg.

Same class names, constructor arguments, attribute names (so state_dict keys and checkpoints match) and
method surface as the reference (networks.py:26-337).  The nn.Linear / nn.BatchNorm1d children only OWN the
parameters; the arithmetic runs in the hand-written CUDA kernels through the C ABI (include/dexgg.h):
  * eval mode  -> dexgg_{generator,evaluator,discriminator}_forward on packed weights (per-object hoisting when the
                  caller gives one BPS row per object instead of a tiled matrix);
  * train mode -> the autograd Functions of train_ops.py (fused linear, train-mode BatchNorm, ...).
There is no CPU path: modules must live on a CUDA device of compute capability 10.x.
"""
import ctypes
import time

import numpy as np
import torch
from torch import nn
from torch.optim import lr_scheduler

from . import _native
from . import utils

KIND_G, KIND_D, KIND_E = 0, 1, 2
# cfg["precision"]: "fp32" = CUDA-core GEMMs (1e-4 contract), "bf16" = tcgen05 (headline mode, stated tolerance),
# "fp32_tc" = the 1e-4 contract ON the tensor cores (bf16 hi + lo pairs, three products per MAC; inference only - training in this
# mode uses the fp32 path)
PRECISION = {"fp32": 0, "bf16": 1, "fp32_tc": 2}


def get_scheduler(optimizer, cfg):
    """networks.py:12-23"""
    if cfg["lr_policy"] == 'step':
        scheduler = lr_scheduler.StepLR(optimizer, step_size=cfg["lr_decay_iters"], gamma=0.1)
    elif cfg["lr_policy"] == 'plateau':
        scheduler = lr_scheduler.ReduceLROnPlateau(optimizer, mode='min', factor=0.2,
                                                   threshold=cfg["threshold_lr_policy_plateau"],
                                                   patience=cfg["patience_lr_policy_plateau"])
    else:
        raise NotImplementedError('Scheduler not implemented.')
    return scheduler


def _float_params_in_state_dict_order(module):
    """The tensors dexgg_net_create wants: state_dict order without num_batches_tracked."""
    return [v for k, v in module.state_dict(keep_vars=True).items() if v.is_floating_point()]


class _Scratch:
    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        return self.buf


class NativeNet:
    """Owns the dexgg_net handle of one module and keeps its packed weights in step with the module's parameters
    (re-packed when any parameter's version counter or storage changes: optimizer.step, load_state_dict, .to())."""

    def __init__(self, module, kind, latent_dim, n_joints, precision="fp32"):
        self.module = module
        self.kind = kind
        self.latent_dim = latent_dim
        self.n_joints = n_joints
        self.precision = precision
        self.handle = None
        self.key = None
        self.scratch = _Scratch()

    def _tensors(self):
        """The module's float tensors in state_dict order; cached (walking state_dict costs ~0.1 ms per call, as much as a small
        forward) and dropped by invalidate(), which the owning module calls from _apply (.to / .cuda / .float replace buffer
        objects) and load_state_dict."""
        ts = getattr(self, "_ts", None)
        if ts is None:
            ts = _float_params_in_state_dict_order(self.module)
            for t in ts:
                if not t.is_cuda:
                    raise RuntimeError("dexgg: module parameters must be on a CUDA (B200) device; there is no CPU path")
                if t.dtype != torch.float32 or not t.is_contiguous():
                    raise RuntimeError("dexgg: parameters must be contiguous float32")
            self._ts = ts
        return ts

    def invalidate(self):
        self._ts = None
        self.key = None

    def sync(self):
        ts = self._tensors()
        key = (self.precision,) + tuple((t.data_ptr(), t._version) for t in ts)
        if key == self.key:
            return self.handle
        L = _native.lib()
        arr = (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
        dev = ts[0].device
        with torch.cuda.device(dev):
            if self.handle is None or self.key is None or self.key[0] != self.precision:
                self.close()
                h = ctypes.c_void_p()
                _native.check(L.dexgg_net_create(self.kind, arr, len(ts), self.latent_dim, self.n_joints,
                                                 PRECISION[self.precision], _native.stream_ptr(dev), ctypes.byref(h)))
                self.handle = h
            else:
                _native.check(L.dexgg_net_update(self.handle, arr, len(ts), _native.stream_ptr(dev)))
        self.key = key
        return self.handle

    def workspace(self, M, n_obj, device):
        n = _native.lib().dexgg_net_workspace_bytes(self.handle, M, n_obj)
        return self.scratch.get(n, device)

    def close(self):
        if self.handle is not None:
            _native.lib().dexgg_net_destroy(self.handle)
            self.handle = None
            self.key = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _dev_f32(x, device):
    if not torch.is_tensor(x):
        x = torch.as_tensor(np.asarray(x))
    return x.to(device=device, dtype=torch.float32).contiguous()


def _objects_and_rows(bps, M):
    """bps [n_obj,4096] with M rows in total -> (n_obj, rows_per_obj)."""
    n_obj = bps.shape[0]
    if n_obj == 0 or M % n_obj != 0:
        raise ValueError(f"dexgg: {M} rows cannot be split evenly over {n_obj} BPS encodings")
    return n_obj, M // n_obj


class _NativeOwner:
    """Mixin of the three network modules: keeps NativeNet's cached tensor list honest."""

    def _apply(self, fn, *args, **kwargs):
        r = super()._apply(fn, *args, **kwargs)
        if getattr(self, "_native", None) is not None:
            self._native.invalidate()
        return r

    def load_state_dict(self, *args, **kwargs):
        r = super().load_state_dict(*args, **kwargs)
        if getattr(self, "_native", None) is not None:
            self._native.invalidate()
        return r


class ResBlock(nn.Module):
    """networks.py:26-56 — parameter container; forward is the train-mode path (eval mode is fused into the parent)."""

    def __init__(self, Fin, Fout, n_neurons=256):
        super(ResBlock, self).__init__()
        self.Fin = Fin
        self.Fout = Fout
        self.fc1 = nn.Linear(Fin, n_neurons)
        self.bn1 = nn.BatchNorm1d(n_neurons)
        self.fc2 = nn.Linear(n_neurons, Fout)
        self.bn2 = nn.BatchNorm1d(Fout)
        if Fin != Fout:
            self.fc3 = nn.Linear(Fin, Fout)
        self.ll = nn.LeakyReLU(negative_slope=0.2)

    def forward(self, x, final_nl=True):
        from . import train_ops as T
        Xin = x if self.Fin == self.Fout else T.linear(x, self.fc3.weight, self.fc3.bias, act=True)
        Xout = T.linear(x, self.fc1.weight, self.fc1.bias)
        Xout = T.batch_norm(Xout, self.bn1, act=True)
        Xout = T.linear(Xout, self.fc2.weight, self.fc2.bias)
        Xout = T.batch_norm(Xout, self.bn2, act=False, residual=Xin, final_act=final_nl)
        return Xout


class Generator(_NativeOwner, nn.Module):
    """networks.py:58-103"""

    def __init__(self, cfg, n_neurons=512, in_bps=4096, dtype=torch.float32, **kwargs):
        super(Generator, self).__init__()
        self.cfg = cfg
        self.latentD = cfg["latentD"]
        self.gen_bn1 = nn.BatchNorm1d(in_bps)
        self.gen_rb1 = ResBlock(self.latentD + in_bps, n_neurons)
        self.gen_rb2 = ResBlock(n_neurons + self.latentD + in_bps, n_neurons)
        self.gen_joint_conf = nn.Linear(n_neurons, cfg["n_hand_joints"])
        self.gen_rot = nn.Linear(n_neurons, 6)
        self.gen_transl = nn.Linear(n_neurons, 3)
        if self.cfg["is_train"]:
            self.train()
        else:
            self.eval()
        self.dtype = dtype
        self._native = NativeNet(self, KIND_G, self.latentD, cfg["n_hand_joints"], cfg.get("precision", "fp32"))

    def native_forward(self, Zin, bps_objects, z_offset=0.0, want_rot6d=True):
        """Eval-mode forward on the device: Zin [M,latentD], bps_objects [n_obj,4096] (M = n_obj * rows_per_obj).
        Returns rot_matrix [M,3,3], transl [M,3] (moved by z_offset), joint_conf [M,nJ], rot_6D [M,6] or None."""
        dev = self.gen_bn1.weight.device
        z = _dev_f32(Zin, dev)
        bps = _dev_f32(bps_objects, dev).reshape(-1, self.gen_bn1.num_features)
        M = z.shape[0]
        n_obj, rpo = _objects_and_rows(bps, M)
        nJ = self.gen_joint_conf.out_features
        rot = torch.empty((M, 3, 3), dtype=torch.float32, device=dev)
        transl = torch.empty((M, 3), dtype=torch.float32, device=dev)
        joint = torch.empty((M, nJ), dtype=torch.float32, device=dev)
        r6 = torch.empty((M, 6), dtype=torch.float32, device=dev) if want_rot6d else None
        h = self._native.sync()
        ws = self._native.workspace(M, n_obj, dev)
        with torch.cuda.device(dev):
            _native.check(_native.lib().dexgg_generator_forward(
                h, bps.data_ptr(), n_obj, rpo, z.data_ptr(), M, float(z_offset), rot.data_ptr(), transl.data_ptr(),
                joint.data_ptr(), _native.ptr(r6), ws.data_ptr(), ws.numel(), _native.stream_ptr(dev)))
        return rot, transl, joint, r6

    def forward(self, Zin, bps_object):
        if self.training:
            from . import train_ops as T
            return T.generator_forward_train(self, Zin, bps_object)
        _, transl, joint_conf, rot_6D = self.native_forward(Zin, bps_object, 0.0, True)
        return {"rot_6D": rot_6D, "transl": transl, "joint_conf": joint_conf, "z": Zin}


class Discriminator(_NativeOwner, nn.Module):
    """networks.py:105-159"""

    def __init__(self, cfg, in_pose, n_neurons=512, in_bps=4096, dtype=torch.float32, **kwargs):
        super(Discriminator, self).__init__()
        self.cfg = cfg
        self.disc_bn1 = nn.BatchNorm1d(in_bps + in_pose)
        self.disc_rb1 = ResBlock(in_bps + in_pose, n_neurons)
        self.disc_rb2 = ResBlock(n_neurons + in_bps + in_pose, n_neurons)
        self.out_success = nn.Linear(n_neurons, 1)
        self.sigmoid = nn.Sigmoid()
        if self.cfg["is_train"]:
            self.train()
        else:
            self.eval()
        self.dtype = dtype
        self.in_bps = in_bps
        self._native = NativeNet(self, KIND_D, cfg["latentD"], cfg["n_hand_joints"], cfg.get("precision", "fp32"))

    def set_input(self, data):
        """networks.py:132-146"""
        self.rot_matrix = data["rot_matrix"].to(dtype=self.dtype, device=self.device)
        self.transl = data["transl"].to(dtype=self.dtype, device=self.device)
        self.joint_conf = data["joint_conf"].to(dtype=self.dtype, device=self.device)
        self.bps_object = data["bps_object"].to(dtype=self.dtype, device=self.device).contiguous()
        self.rot_matrix = self.rot_matrix.reshape(self.rot_matrix.shape[0], -1)

    def native_forward(self, bps_objects, rot, transl, joint):
        dev = self.out_success.weight.device
        rot = _dev_f32(rot, dev).reshape(-1, 9)
        transl = _dev_f32(transl, dev)
        joint = _dev_f32(joint, dev)
        bps = _dev_f32(bps_objects, dev).reshape(-1, self.in_bps)
        M = rot.shape[0]
        n_obj, rpo = _objects_and_rows(bps, M)
        p = torch.empty((M, 1), dtype=torch.float32, device=dev)
        h = self._native.sync()
        ws = self._native.workspace(M, n_obj, dev)
        with torch.cuda.device(dev):
            _native.check(_native.lib().dexgg_discriminator_forward(
                h, bps.data_ptr(), n_obj, rpo, rot.data_ptr(), transl.data_ptr(), joint.data_ptr(), M, p.data_ptr(),
                ws.data_ptr(), ws.numel(), _native.stream_ptr(dev)))
        return p

    def forward(self, data):
        self.set_input(data)
        if self.training:
            from . import train_ops as T
            if getattr(self, "double_backward", False):  # WGAN-GP interpolates pass (dexgangrasp.py:194)
                return T.discriminator_forward_double_backward(self)
            return T.discriminator_forward_train(self)
        return self.native_forward(self.bps_object, self.rot_matrix, self.transl, self.joint_conf)


class DexGANGrasp(nn.Module):
    """networks.py:161-276 — generator + discriminator container with the sampling helpers."""

    def __init__(self, cfg, n_neurons=512, in_bps=4096, in_pose=9 + 3, dtype=torch.float32, **kwargs):
        super(DexGANGrasp, self).__init__()
        self.cfg = cfg
        in_pose += cfg["n_hand_joints"]
        if not torch.cuda.is_available():
            raise RuntimeError("dexgg: no CUDA device — this implementation runs on a B200 only (no CPU fallback)")
        self.device = torch.device('cuda:{}'.format(cfg["gpu_ids"][0]))
        self.latentD = cfg["latentD"]
        self.discriminator = Discriminator(cfg, in_pose, n_neurons, in_bps, dtype)
        self.discriminator.to(self.device)
        self.generator = Generator(cfg, n_neurons, in_bps, dtype)
        self.generator.to(self.device)
        self.discriminator.device = self.device
        self.generator.device = self.device
        if self.cfg["is_train"]:
            self.train()
        else:
            self.eval()
        self.dtype = dtype

    def forward(self, Zin, bps_object, real_data=None):
        return self.generator(Zin, bps_object)

    def sample_latents(self, n_samples, sample_uniform=False):
        """z exactly where and how the reference draws it (networks.py:236-244) — the "same z" contract."""
        if not sample_uniform:
            return torch.randn((n_samples, self.latentD), dtype=self.dtype, device=self.device)
        return 8 * torch.rand((n_samples, self.latentD), dtype=self.dtype, device=self.device) - 4

    def generate_poses(self, bps_object, return_arr=False, seed=None, sample_uniform=False, z_offset=0.0, n_samples=None):
        """networks.py:223-255.  `bps_object` is either the reference's tiled [n,4096] matrix (n_samples=None) or one
        row per object with `n_samples` rows in total (nothing is tiled; the BPS projection runs once per object)."""
        start = time.time()
        bps_object = _dev_f32(bps_object, self.device)
        if n_samples is None:
            n_samples = bps_object.shape[0]
        self.eval()
        with torch.no_grad():
            Zgen = self.sample_latents(n_samples, sample_uniform)
            rot, transl, joint, _ = self.generator.native_forward(Zgen, bps_object, z_offset, False)
        results = {"transl": transl, "joint_conf": joint, "z": Zgen, "rot_matrix": rot}
        if return_arr:
            for k, v in results.items():
                results[k] = v.cpu().detach().numpy()
        if self.cfg.get("verbose", False):
            print("Sampling took: %.3f" % (time.time() - start))
        return results

    def generate_grasps(self, bps, n_samples, return_arr=True, z_offset=0.025):
        """networks.py:257-276 — one object's encoding [1,4096] or [4096]; no np.tile, no [n,4096] copy."""
        bps = np.asarray(bps.detach().cpu() if torch.is_tensor(bps) else bps, dtype=np.float32)
        if len(bps.shape) > 1:
            bps = bps.squeeze()
        bps_tensor = torch.tensor(bps.reshape(1, -1), dtype=self.dtype, device=self.device)
        return self.generate_poses(bps_tensor, return_arr=return_arr, z_offset=z_offset, n_samples=n_samples)


class DexEvaluator(_NativeOwner, nn.Module):
    """networks.py:278-337"""

    def __init__(self, cfg, n_neurons=512, in_bps=4096, in_pose=9 + 3, dtype=torch.float32, **kwargs):
        super(DexEvaluator, self).__init__()
        self.cfg = cfg
        self.bn1 = nn.BatchNorm1d(in_bps + in_pose)
        self.rb1 = ResBlock(in_bps + in_pose, n_neurons)
        self.rb2 = ResBlock(in_bps + in_pose + n_neurons, n_neurons)
        self.rb3 = ResBlock(in_bps + in_pose + n_neurons, n_neurons)
        self.out_success = nn.Linear(n_neurons, 1)
        self.dout = nn.Dropout(0.3)
        self.sigmoid = nn.Sigmoid()
        if self.cfg["is_train"]:
            self.train()
        else:
            self.eval()
        self.dtype = torch.float32
        if not torch.cuda.is_available():
            raise RuntimeError("dexgg: no CUDA device — this implementation runs on a B200 only (no CPU fallback)")
        self.device = torch.device('cuda:{}'.format(cfg["gpu_ids"][0]))
        self.in_bps = in_bps
        self._native = NativeNet(self, KIND_E, cfg["latentD"], cfg["n_hand_joints"], cfg.get("precision", "fp32"))

    def set_input(self, data):
        """networks.py:307-313"""
        self.rot_matrix = data["rot_matrix"].to(dtype=self.dtype, device=self.device)
        self.transl = data["transl"].to(dtype=self.dtype, device=self.device)
        self.bps_object = data["bps_object"].to(dtype=self.dtype, device=self.device).contiguous()
        if 'label' in data.keys():
            self.gt_label = data["label"].to(dtype=self.dtype, device=self.device).unsqueeze(-1)
        self.rot_matrix = self.rot_matrix.reshape(self.rot_matrix.shape[0], -1)

    def native_forward(self, bps_objects, rot, transl):
        dev = self.out_success.weight.device
        rot = _dev_f32(rot, dev).reshape(-1, 9)
        transl = _dev_f32(transl, dev)
        bps = _dev_f32(bps_objects, dev).reshape(-1, self.in_bps)
        M = rot.shape[0]
        n_obj, rpo = _objects_and_rows(bps, M)
        p = torch.empty((M, 1), dtype=torch.float32, device=dev)
        h = self._native.sync()
        ws = self._native.workspace(M, n_obj, dev)
        with torch.cuda.device(dev):
            _native.check(_native.lib().dexgg_evaluator_forward(
                h, bps.data_ptr(), n_obj, rpo, rot.data_ptr(), transl.data_ptr(), M, p.data_ptr(), ws.data_ptr(),
                ws.numel(), _native.stream_ptr(dev)))
        return p

    def forward(self, data):
        self.set_input(data)
        if self.training:
            from . import train_ops as T
            return T.evaluator_forward_train(self, getattr(self, "dropout_masks", None))
        return self.native_forward(self.bps_object, self.rot_matrix, self.transl)

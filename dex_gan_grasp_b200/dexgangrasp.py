"""Host-side mirror of DexGanGrasp/models/dexgangrasp.py (class DexGanGrasp, :65-822): the same constructor, method names,
arguments, returned dict keys, checkpoint layout and error behaviour; every forward / backward runs in libdexgg's
sm_100a kernels.  refine_grasps / improve_grasps_gradient_based run on the evaluator's input-gradient GEMMs.
"""
from __future__ import division

import os

import numpy as np
import torch

from . import _native
from . import losses
from . import networks
from . import utils
from .networks import DexEvaluator, DexGANGrasp


class EarlyStopping:
    """DexGanGrasp/utils/train_tools.py:4-42 (constructed by the wrapper, never triggered by train.py)."""

    def __init__(self, patience=7, min_delta=0.0):
        self.patience, self.min_delta = patience, min_delta
        self.counter, self.best_loss, self.early_stop = 0, np.inf, False

    def __call__(self, val_loss):
        if self.best_loss - val_loss > self.min_delta:
            self.best_loss, self.counter = val_loss, 0
        else:
            self.counter += 1
            if self.counter >= self.patience:
                self.early_stop = True
        return self.early_stop


def define_losses(loss_type_recon):
    if loss_type_recon == 'transl_rot_6D_l2':
        return None, losses.transl_rot_6D_l2_loss
    raise Exception('Requested loss not available, choose [transl_rot_6D_l2]')


def build_network(cfg, device, is_train):
    """dexgangrasp.py:28-37"""
    gen = DexGANGrasp(cfg)
    eva = DexEvaluator(cfg)
    gen.to(device)
    eva.to(device)
    if is_train:
        init_net(gen)
        init_net(eva)
    return gen, eva


def init_net(net):
    """dexgangrasp.py:39-62"""
    init_type = net.cfg["weight_init_type"]
    init_gain = net.cfg["init_gain"]

    def init_func(m):
        classname = m.__class__.__name__
        if hasattr(m, 'weight') and (classname.find('Conv') != -1 or classname.find('Linear') != -1):
            if init_type == 'normal':
                torch.nn.init.normal_(m.weight.data, 0.0, init_gain)
            elif init_type == 'xavier':
                torch.nn.init.xavier_normal_(m.weight.data, gain=init_gain)
            elif init_type == 'kaiming':
                torch.nn.init.kaiming_normal_(m.weight.data, a=0, mode='fan_in')
            elif init_type == 'orthogonal':
                torch.nn.init.orthogonal_(m.weight.data, gain=init_gain)
            else:
                raise NotImplementedError('initialization method [%s] is not implemented' % init_type)
        elif classname.find('BatchNorm') != -1:
            torch.nn.init.normal_(m.weight.data, 1.0, init_gain)
            torch.nn.init.constant_(m.bias.data, 0.0)

    net.apply(init_func)


class DexGanGrasp(object):
    """Wrapper which houses the network blocks, the losses and training logic (dexgangrasp.py:65)."""

    def __init__(self, cfg):
        self.dtype = torch.float32
        torch.set_default_dtype(self.dtype)
        self.cfg = cfg
        self.is_train = cfg["is_train"]
        self.is_wgan = cfg.get("is_wgan", False)
        if not torch.cuda.is_available():
            raise RuntimeError("dexgg: no CUDA device — this implementation runs on a B200 only (no CPU fallback)")
        self.device = torch.device('cuda:{}'.format(cfg["gpu_ids"][0]))
        _native.lib()  # fail loudly here if the CUDA library is missing

        if cfg["model"] in ("ffhnet", "dexgangrasp"):
            self.DexGANGrasp, self.DexEvaluator = build_network(cfg, self.device, is_train=self.is_train)
        else:
            raise ValueError('Wrong configure model name of', cfg["model"])

        if self.is_train:
            self.bce_weight = cfg["bce_weight"]
            self.kl_coef = cfg.get("kl_coef", 0.0)
            self.transl_coef = 100.
            self.rot_coef = 1.
            self.conf_coef = 10.
            self.train_dexgenerator = cfg["train_ffhgenerator"]
            self.train_dexevaluator = cfg["train_ffhevaluator"]
            from .train_ops import DexggAdam  # torch.optim.Adam state / state_dict, one fused CUDA launch per step
            adam = lambda params, lr: DexggAdam(params, lr=lr, betas=(cfg["beta1"], 0.999),
                                                weight_decay=cfg["weight_decay"])
            self.optim_dexgenerator = adam(self.DexGANGrasp.parameters(), cfg["lr"])
            self.optim_dexgangrasp_generator = adam(self.DexGANGrasp.generator.parameters(), cfg["lr_gen"])
            self.optim_dexgangrasp_discriminator = adam(self.DexGANGrasp.discriminator.parameters(), cfg["lr_dis"])
            self.optim_dexevaluator = adam(self.DexEvaluator.parameters(), cfg["lr"])
            self.scheduler_dexgenerator = networks.get_scheduler(self.optim_dexgenerator, cfg)
            self.scheduler_dexgangrasp_generator = networks.get_scheduler(self.optim_dexgangrasp_generator, cfg)
            self.scheduler_dexgangrasp_discriminator = networks.get_scheduler(self.optim_dexgangrasp_discriminator, cfg)
            self.scheduler_dexevaluator = networks.get_scheduler(self.optim_dexevaluator, cfg)
            self.estop_dexgenerator = EarlyStopping()
            self.estop_dexevaluator = EarlyStopping()

        self.kl_loss, self.rec_pose_loss = define_losses('transl_rot_6D_l2')
        self.L2_loss = torch.nn.MSELoss(reduction='mean')
        self.BCE_loss = torch.nn.BCELoss(reduction='mean')
        self.compute_eva_accuracy = losses.accuracy_evaluator

        # The reference wraps in nn.DataParallel for len(gpu_ids) > 1 (dexgangrasp.py:127-129), which cannot work for this
        # model.  Here data parallelism is one process per GPU: see dataparallel.py.
        self.file_path = os.path.dirname(os.path.abspath(__file__))
        self.logit_thresh = 0.5
        self.last_loss_dict_gen = {}

    # ------------------------------------------------------------------ losses (dexgangrasp.py:142-147,213-232,272-307)
    def compute_loss_dexevaluator(self, pred_success_p):
        bce_loss_val = self.bce_weight * self.BCE_loss(pred_success_p, self.DexEvaluator.gt_label)
        loss_dict = {'total_loss_eva': bce_loss_val, 'bce_loss': bce_loss_val}
        return bce_loss_val, loss_dict

    # -- WGAN-GP variant (`is_wgan`, dexgangrasp.py:149-211,234-270) --
    def calculate_interp(self, real_data, fake_data, alpha=None):
        """dexgangrasp.py:149-158.  alpha ~ randn (NOT rand: reference quirk, kept); `alpha` lets a test supply it."""
        if alpha is None:
            shape = (real_data.size(0), 1, 1) if len(real_data.shape) == 3 else (real_data.size(0), 1)
            alpha = torch.randn(shape, device=self.DexGANGrasp.device)
        return (alpha * real_data + ((1 - alpha) * fake_data)).requires_grad_(True)

    def calculate_grad_interp(self, interpolates, model_interpolates):
        """dexgangrasp.py:160-173: d(sum of scores)/d(interpolates), kept in the graph for the second backward."""
        grad_outputs = torch.ones(model_interpolates.size(), device=self.DexGANGrasp.device, requires_grad=False)
        gradients = torch.autograd.grad(outputs=model_interpolates, inputs=interpolates, grad_outputs=grad_outputs,
                                        create_graph=True, retain_graph=True, only_inputs=True)[0]
        return gradients.view(gradients.size(0), -1)

    def calculate_gradient_penalty(self, real_data, fake_data, alphas=None):
        """dexgangrasp.py:175-202.  The interpolates go through the discriminator's twice-differentiable train-mode pass
        (train_ops.discriminator_forward_double_backward: all GEMMs, of every derivative order, on the dexgg kernels)."""
        dev = self.DexGANGrasp.device
        names = ("rot_matrix", "transl", "joint_conf")
        alphas = alphas if alphas is not None else (None, None, None)
        interp_data = {"bps_object": real_data["bps_object"]}
        for n, a in zip(names, alphas):
            interp_data[n] = self.calculate_interp(real_data[n].to(dev), fake_data[n],
                                                   None if a is None else a.to(dev))
        D = self.DexGANGrasp.discriminator
        D.double_backward = True
        try:
            interpolate_scores = D(interp_data)
        finally:
            D.double_backward = False
        gradient_penalties = []
        for out in names:
            gradients = self.calculate_grad_interp(interp_data[out], interpolate_scores)
            gradient_penalties.append(torch.mean((gradients.norm(2, dim=1) - 1) ** 2))
        return torch.mean(torch.stack(gradient_penalties))

    def compute_loss_dexgangrasp_discriminator_wass(self, real_score, fake_score, real_data, fake_data, penalty_gain=10,
                                                    alphas=None):
        """dexgangrasp.py:204-218 (the key names 'bce_loss_*' are the reference's, the values are mean scores)."""
        bce_loss_real = torch.mean(real_score)
        bce_loss_fake = torch.mean(fake_score)
        gradient_penalty = self.calculate_gradient_penalty(real_data, fake_data, alphas)
        total_loss_disc = -bce_loss_real + bce_loss_fake + (gradient_penalty * penalty_gain)
        loss_dict = {'total_loss_disc': total_loss_disc, 'bce_loss_real': bce_loss_real, 'bce_loss_fake': bce_loss_fake}
        return total_loss_disc, loss_dict

    def compute_loss_dexgangrasp_generator_wass(self, real_data, fake_data, fake_score):
        """dexgangrasp.py:234-270"""
        dev = self.DexGANGrasp.device
        gt = {'transl': real_data['transl'].to(dev).float(), 'rot_matrix': real_data['rot_matrix'].to(dev).float()}
        transl_loss_val, rot_loss_val = self.rec_pose_loss(fake_data, gt, self.L2_loss, self.device)
        conf_loss_val = self.L2_loss(fake_data['joint_conf'].to(dev).float(), real_data['joint_conf'].to(dev).float())
        gen_loss_fake = - torch.mean(fake_score)
        loss_dict = {
            'gen_loss_fake': gen_loss_fake,
            'transl_loss': self.transl_coef * transl_loss_val,
            'rot_loss': self.rot_coef * rot_loss_val,
            'conf_loss': self.conf_coef * conf_loss_val
        }
        total_loss = gen_loss_fake + (loss_dict['transl_loss'] + loss_dict['rot_loss'] + loss_dict['conf_loss'])
        loss_dict["total_loss_gen"] = total_loss
        self.last_loss_dict_gen = loss_dict
        return total_loss, loss_dict

    def compute_loss_dexgangrasp_discriminator(self, real_score, fake_score, real_data=None, fake_data_disc=None):
        bce_loss_real = self.bce_weight * self.BCE_loss(real_score, torch.ones_like(real_score))
        bce_loss_fake = self.bce_weight * self.BCE_loss(fake_score, torch.zeros_like(fake_score))
        total_loss_disc = (bce_loss_real + bce_loss_fake) / 2
        loss_dict = {'total_loss_disc': total_loss_disc, 'bce_loss_real': bce_loss_real, 'bce_loss_fake': bce_loss_fake}
        return total_loss_disc, loss_dict

    def compute_loss_dexgangrasp_generator(self, real_data, fake_data, fake_score):
        dev = self.DexGANGrasp.device
        gt = {'transl': real_data['transl'].to(dev).float(), 'rot_matrix': real_data['rot_matrix'].to(dev).float()}
        transl_loss_val, rot_loss_val = self.rec_pose_loss(fake_data, gt, self.L2_loss, self.device)
        conf_loss_val = self.L2_loss(fake_data['joint_conf'].to(dev).float(), real_data['joint_conf'].to(dev).float())
        gen_loss_fake = self.bce_weight * self.BCE_loss(fake_score, torch.ones_like(fake_score))
        loss_dict = {
            'gen_loss_fake': gen_loss_fake,
            'transl_loss': self.transl_coef * transl_loss_val,
            'rot_loss': self.rot_coef * rot_loss_val,
            'conf_loss': self.conf_coef * conf_loss_val
        }
        total_loss = gen_loss_fake + (loss_dict['transl_loss'] + loss_dict['rot_loss'] + loss_dict['conf_loss'])
        loss_dict["total_loss_gen"] = total_loss
        self.last_loss_dict_gen = loss_dict
        return total_loss, loss_dict

    # ------------------------------------------------------------------ evaluation helpers (:309-365)
    def eval_dexevaluator_accuracy(self, data):
        logits = self.DexEvaluator(data)
        pred_label = utils.class_labels_from_logits(logits, self.logit_thresh)
        pos_acc, neg_acc, acc = self.compute_eva_accuracy(pred_label, self.DexEvaluator.gt_label)
        return pos_acc, neg_acc, pred_label.detach().cpu().numpy(), self.DexEvaluator.gt_label.detach().cpu().numpy()

    def eval_dexevaluator_loss(self, data):
        self.DexEvaluator.eval()
        with torch.no_grad():
            out = self.DexEvaluator(data)
            _, loss_dict_dexevaluator = self.compute_loss_dexevaluator(out)
        return loss_dict_dexevaluator

    def eval_dexgangrasp_generator_loss(self, real_data):
        self.DexGANGrasp.eval()
        with torch.no_grad():
            n_samples = real_data["bps_object"].shape[0]
            Zgen = torch.randn((n_samples, self.DexGANGrasp.latentD), dtype=self.DexGANGrasp.dtype, device=self.DexGANGrasp.device)
            y_fake = self.DexGANGrasp.generator(Zgen, real_data["bps_object"].to(self.DexGANGrasp.device))
            y_fake["rot_matrix"] = utils.rot_matrix_from_ortho6d(y_fake["rot_6D"])
            fake_data = {"bps_object": real_data["bps_object"], "rot_matrix": y_fake["rot_matrix"], "rot_6D": y_fake["rot_6D"],
                         "transl": y_fake["transl"], "joint_conf": y_fake["joint_conf"]}
            fake_score_gen = self.DexGANGrasp.discriminator(fake_data)
            real_data["rot_matrix"] = real_data["rot_matrix"].view(real_data["bps_object"].shape[0], -1)
            gen_loss = self.compute_loss_dexgangrasp_generator_wass if self.is_wgan else self.compute_loss_dexgangrasp_generator
            _, loss_dict_dexgenerator = gen_loss(real_data, fake_data, fake_score_gen)
        return loss_dict_dexgenerator

    # ------------------------------------------------------------------ inference (:367-514)
    def _one_object_bps(self, bps):
        bps = np.asarray(bps.detach().cpu() if torch.is_tensor(bps) else bps, dtype=np.float32)
        if len(bps.shape) > 1:
            bps = bps.squeeze()
        return bps

    def _scores(self, net_forward, bps, grasps, with_joints):
        """Shared head of evaluate_grasps / filter_grasps(_discriminator): same side effect on `grasps` as the reference
        (adds the tiled 'bps_object', dexgangrasp.py:403-408) but the device never sees the tiled matrix."""
        n_samples = grasps['rot_matrix'].shape[0]
        bps = self._one_object_bps(bps)
        grasps['bps_object'] = np.broadcast_to(bps, (n_samples, bps.shape[-1]))  # a view: no n*16 KB host copy
        grasps_t = {k: (v.clone().detach() if torch.is_tensor(v) else torch.tensor(np.asarray(v), dtype=self.dtype, device=self.device))
                    for k, v in grasps.items() if k != 'bps_object'}
        bps_row = torch.tensor(bps.reshape(1, -1), dtype=self.dtype, device=self.device)
        with torch.no_grad():
            if with_joints:
                p = net_forward(bps_row, grasps_t['rot_matrix'], grasps_t['transl'], grasps_t['joint_conf'])
            else:
                p = net_forward(bps_row, grasps_t['rot_matrix'], grasps_t['transl'])
        return p.reshape(-1), grasps_t, bps_row

    def evaluate_grasps(self, bps, grasps, thresh=0.5, return_arr=True):
        self.DexEvaluator.eval()
        p_success, _, _ = self._scores(self.DexEvaluator.native_forward, bps, grasps, False)
        if return_arr:
            p_success = p_success.cpu().detach().numpy()
        return p_success

    def _filter(self, p_success, grasps_t, bps_row, thresh, return_arr):
        n = p_success.shape[0]
        idx = torch.empty((n,), dtype=torch.int32, device=self.device)
        cnt = torch.empty((1,), dtype=torch.int32, device=self.device)
        best = torch.empty((1,), dtype=torch.float32, device=self.device)
        L = _native.lib()
        nws = L.dexgg_filter_topk_workspace_bytes(n, n)  # 0 up to 16384 grasps; longer lists sort in a workspace
        ws = torch.empty((max(nws, 1),), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _native.check(L.dexgg_filter_topk_ws(p_success.data_ptr(), n, n, float(thresh), n, idx.data_ptr(), cnt.data_ptr(),
                                                 best.data_ptr(), None, ws.data_ptr(), ws.numel(), _native.stream_ptr(self.device)))
        best_score, kept = float(best.item()), int(cnt.item())
        if best_score < thresh:
            raise ValueError("In total predicted ", n, " grasps, but best score ", best_score, " is still lower than thresh", thresh)
        indices = idx[:kept].long()
        filt_grasps = {k: v.index_select(0, indices) for k, v in grasps_t.items()}
        filt_grasps['bps_object'] = bps_row.expand(kept, -1)
        if return_arr:
            filt_grasps = {k: v.cpu().detach().numpy() for k, v in filt_grasps.items()}
        return filt_grasps

    def filter_grasps(self, bps, grasps, thresh=0.5, return_arr=True):
        """dexgangrasp.py:393-442: keep p > thresh, best first; ValueError when the best score is below thresh."""
        self.DexEvaluator.eval()
        p, grasps_t, bps_row = self._scores(self.DexEvaluator.native_forward, bps, grasps, False)
        return self._filter(p, grasps_t, bps_row, thresh, return_arr)

    def filter_grasps_discriminator(self, bps, grasps, thresh=0.5, return_arr=True):
        """dexgangrasp.py:444-493"""
        self.DexGANGrasp.eval()
        p, grasps_t, bps_row = self._scores(self.DexGANGrasp.discriminator.native_forward, bps, grasps, True)
        return self._filter(p, grasps_t, bps_row, thresh, return_arr)

    def generate_grasps(self, bps, n_samples, return_arr=True, z_offset=0.025):
        """dexgangrasp.py:495-514"""
        return self.DexGANGrasp.generate_grasps(bps, n_samples, return_arr=return_arr, z_offset=z_offset)

    def improve_grasps_gradient_based(self, data, last_success):
        """dexgangrasp.py:516-541: one gradient-ascent step on the evaluator score w.r.t. transl / rot_matrix, step length
        capped so that transl moves <= 1 cm.  data['rot_matrix'] / data['transl'] are CUDA leaf tensors with
        requires_grad (refine_grasps makes them); as in the reference their .grad is never zeroed, so step i moves along
        the SUM of the gradients of steps 0..i.  DexEvaluator does not read joint_conf (networks.py:307-313), so
        data['joint_conf'].grad stays None; the reference raises TypeError on that line (:539) - here it is the zero
        gradient it mathematically is and joint_conf is left unchanged.  The reference's unused `diff` against
        data['label'] (:528-529) is dropped, so no 'label' key is needed."""
        from . import train_ops
        self.DexEvaluator.eval()
        bps = data['bps_object']
        key = (bps.data_ptr(), tuple(bps.shape), bps._version, id(self.DexEvaluator))
        if getattr(self, '_refine_graph', (None, None))[0] != key:
            self._refine_graph = (key, train_ops.EvaluatorRefineGraph(self.DexEvaluator, bps.to(self.device, torch.float32)))
        with train_ops.fp32_gemms():
            p_success = self._refine_graph[1](data['rot_matrix'], data['transl'])
            p_success.squeeze(-1).backward(torch.ones(p_success.shape[0], device=self.device))
        norm_transl = torch.norm(data['transl'].grad, p=2, dim=-1)
        alpha = torch.min(0.01 / norm_transl, torch.tensor(1.0, device=self.device))
        data['transl'].data += data['transl'].grad * alpha[:, None]
        data['rot_matrix'].data += data['rot_matrix'].grad * alpha[:, None, None]
        if data['joint_conf'].grad is not None:
            data['joint_conf'].data += data['joint_conf'].grad * alpha[:, None]
        return p_success.squeeze(), None

    def refine_grasps(self, data, refine_method, num_refine_steps=10, dtype=torch.float32):
        """dexgangrasp.py:619-660: num_refine_steps evaluator-gradient steps on ranked grasps; returns (refined_data,
        refined_success): the grasp dict as numpy arrays before every step and after the last one, and the scores before
        every step plus the score of the final grasps.  refine_method 'gradient' only: the reference's sampling variant
        (:543-569) reads attributes that do not exist (self.grasp_evaluator, utils.control_points_from_rot_and_trans)."""
        from . import train_ops
        if refine_method != "gradient":
            raise NotImplementedError("refine_grasps: only refine_method='gradient'; the reference's sampling-based variant "
                                      "(dexgangrasp.py:543-569) references attributes that do not exist")
        data = {k: (v.to(dtype) if torch.is_tensor(v) else torch.as_tensor(v, dtype=dtype)) for k, v in data.items()}
        data['bps_object'] = data['bps_object'].to(self.device).contiguous()
        for k in ('rot_matrix', 'transl', 'joint_conf'):
            data[k] = data[k].to(self.device).detach().clone().requires_grad_(True)

        def as_numpy():
            return {k: data[k].cpu().data.numpy() for k in ('rot_matrix', 'transl', 'joint_conf')}
        refined_success, refined_data = [], [as_numpy()]
        last_success = None
        for _ in range(num_refine_steps):
            p_success, last_success = self.improve_grasps_gradient_based(data, last_success)
            refined_success.append(p_success.cpu().data.numpy())
            refined_data.append(as_numpy())
        with torch.no_grad(), train_ops.fp32_gemms():
            refined_success.append(self._refine_graph[1](data['rot_matrix'], data['transl']).squeeze().cpu().numpy())
        self._refine_graph = (None, None)
        return refined_data, refined_success

    # ------------------------------------------------------------------ checkpoints (:571-617,662-712)
    def load_dexevaluator(self, epoch, load_path=None):
        if load_path is None:
            load_path = self.cfg["load_path"]
        load_path = os.path.join(load_path, str(epoch) + '_eva_net.pt')
        ckpt = torch.load(load_path, map_location=self.device)
        self.DexEvaluator.load_state_dict(ckpt['ffhevaluator_state_dict'])
        if self.cfg["is_train"]:
            self.optim_dexevaluator.load_state_dict(ckpt['optim_ffhevaluator_state_dict'])
            self.scheduler_dexevaluator.load_state_dict(ckpt['scheduler_ffhevaluator_state_dict'])
            self.cfg["load_epoch"] = ckpt['epoch']
            self.DexEvaluator.train()
        else:
            self.DexEvaluator.eval()

    def load_dexgenerator(self, epoch, load_path=None):
        if load_path is None:
            load_path = self.cfg["load_path"]
        load_path = os.path.join(load_path, str(epoch) + '_gen_net.pt')
        ckpt = torch.load(load_path, map_location=self.device)
        self.DexGANGrasp.load_state_dict(ckpt['ffhgenerator_state_dict'])
        if self.cfg["is_train"]:
            self.optim_dexgenerator.load_state_dict(ckpt['optim_ffhgenerator_state_dict'])
            self.scheduler_dexgenerator.load_state_dict(ckpt['scheduler_ffhgenerator_state_dict'])
            self.DexGANGrasp.train()
        else:
            self.DexGANGrasp.eval()

    def save_dexevaluator(self, net_name, epoch):
        save_path = os.path.join(self.cfg["save_dir"], net_name + '_eva_net.pt')
        sd = {k: v.detach().cpu() for k, v in self.DexEvaluator.state_dict().items()}
        torch.save({
            'epoch': epoch,
            'ffhevaluator_state_dict': sd,
            'optim_ffhevaluator_state_dict': self.optim_dexevaluator.state_dict(),
            'scheduler_ffhevaluator_state_dict': self.scheduler_dexevaluator.state_dict(),
        }, save_path)

    def save_dexgenerator(self, net_name, epoch):
        save_path = os.path.join(self.cfg["save_dir"], net_name + '_gen_net.pt')
        sd = {k: v.detach().cpu() for k, v in self.DexGANGrasp.state_dict().items()}
        torch.save({
            'epoch': epoch,
            'ffhgenerator_state_dict': sd,
            'optim_ffhgenerator_state_dict': self.optim_dexgenerator.state_dict(),
            'scheduler_ffhgenerator_state_dict': self.scheduler_dexgenerator.state_dict(),
            'optim_ffhgan_generator_state_dict': self.optim_dexgangrasp_generator.state_dict(),
            'scheduler_ffhgan_generator_state_dict': self.scheduler_dexgangrasp_generator.state_dict(),
            'optim_ffhgan_discriminator_state_dict': self.optim_dexgangrasp_discriminator.state_dict(),
            'scheduler_ffhgan_discriminator_state_dict': self.scheduler_dexgangrasp_discriminator.state_dict(),
        }, save_path)

    def update_estop(self, eval_loss_dict):
        if self.train_dexevaluator and self.estop_dexevaluator(eval_loss_dict['total_loss_eva']):
            self.train_dexevaluator = False
        if self.train_dexgenerator and self.estop_dexgenerator(eval_loss_dict['total_loss_gen']):
            self.train_dexgenerator = False

    def update_learning_rate(self, eval_loss_dict):
        if self.train_dexevaluator:
            self.scheduler_dexevaluator.step(eval_loss_dict['total_loss_eva'])
        if self.train_dexgenerator:
            self.scheduler_dexgenerator.step(eval_loss_dict['total_loss_gen'])

    # ------------------------------------------------------------------ training steps (:739-823)
    def _train_precision(self):
        """cfg['precision'] selects the GEMM arithmetic of the training step too: 'fp32' (default, 1e-4 parity with the
        reference) or 'bf16' (tcgen05 tensor cores, fp32 accumulation and fp32 master weights)."""
        from . import train_ops
        prec = self.cfg.get("precision", "fp32")
        train_ops.set_precision("fp32" if prec == "fp32_tc" else prec)

    def train_engine_name(self):
        return ("hand-scheduled fused step (train_engine.py: gemm_umma.cu + train_fused.cu)" if self._use_engine()
                else "autograd Functions over the dexgg kernels (train_ops.py)")

    def _use_engine(self):
        """The fused bf16 step (train_engine.py) runs the BCE variant (any number of ranks); the WGAN-GP variant, the fp32 parity
        mode and cfg['train_engine'] = False use the autograd-Function path of train_ops.py."""
        return self.cfg.get("precision", "fp32") == "bf16" and not self.is_wgan and self.cfg.get("train_engine", True)

    def _engine(self, rows):
        from .train_engine import TrainEngine
        eng = getattr(self, "_train_engine", None)
        if eng is None or eng.B != rows or eng._closed:
            if eng is not None:
                eng.close()  # another batch size (a last, smaller batch of an epoch): the old engine's device buffers go first
            eng = self._train_engine = TrainEngine(self, rows)
        return eng

    def update_dexevaluator(self, data):
        if self._use_engine():
            self.DexEvaluator.train()
            return self._engine(int(data["bps_object"].shape[0])).evaluator_step(data)
        self._train_precision()
        self.DexEvaluator.train()
        out = self.DexEvaluator(data)
        total_loss_dexevaluator, loss_dict_dexevaluator = self.compute_loss_dexevaluator(out)
        self.optim_dexevaluator.zero_grad(set_to_none=getattr(self, '_zero_grad_to_none', True))
        total_loss_dexevaluator.backward()
        self._sync_grads(self.DexEvaluator)
        self.optim_dexevaluator.step()
        return loss_dict_dexevaluator

    def _sync_grads(self, module):
        """Data-parallel hook: average gradients over ranks when torch.distributed is initialised (dataparallel.py)."""
        from . import dataparallel
        dataparallel.allreduce_gradients(module)

    def update_dexgangrasp(self, real_data, is_train_gen=True, Zgen=None, alphas=None):
        """dexgangrasp.py:757-823, both loss selections (:761-763).  `Zgen` / `alphas` (optional) let a test supply the
        latent batch and the three WGAN-GP interpolation weights."""
        is_wgan = self.is_wgan
        gen_loss = self.compute_loss_dexgangrasp_generator_wass if is_wgan else self.compute_loss_dexgangrasp_generator
        self.DexGANGrasp.train()
        G, D = self.DexGANGrasp.generator, self.DexGANGrasp.discriminator
        n_samples = real_data["bps_object"].shape[0]
        if Zgen is None:
            Zgen = torch.randn((n_samples, self.DexGANGrasp.latentD), dtype=self.DexGANGrasp.dtype, device=self.DexGANGrasp.device)
        if self._use_engine():
            out = self._engine(int(n_samples)).gan_step(real_data, is_train_gen, Zgen.to(self.device))
            if is_train_gen:
                real_data["rot_matrix"] = real_data["rot_matrix"].view(n_samples, -1)  # the reference's side effect (:810)
            return out
        self._train_precision()
        y_fake = G(Zgen, real_data["bps_object"].to(self.DexGANGrasp.device))
        fake_rot_6D, fake_transl, fake_joint_conf = y_fake["rot_6D"], y_fake["transl"], y_fake["joint_conf"]
        y_fake["rot_matrix"] = utils.rot_matrix_from_ortho6d(y_fake["rot_6D"])
        fake_data_disc = {"bps_object": real_data["bps_object"], "rot_matrix": y_fake["rot_matrix"].detach(),
                          "transl": fake_transl.detach(), "joint_conf": fake_joint_conf.detach()}
        real_score = D(real_data)
        fake_score = D(fake_data_disc)
        if is_wgan:
            total_loss_disc, loss_dict_disc = self.compute_loss_dexgangrasp_discriminator_wass(
                real_score, fake_score, real_data, fake_data_disc, alphas=alphas)
        else:
            total_loss_disc, loss_dict_disc = self.compute_loss_dexgangrasp_discriminator(real_score, fake_score, real_data, fake_data_disc)
        self.optim_dexgangrasp_discriminator.zero_grad(set_to_none=getattr(self, '_zero_grad_to_none', True))
        total_loss_disc.backward()
        self._sync_grads(D)
        self.optim_dexgangrasp_discriminator.step()

        if is_train_gen:
            fake_data = {"bps_object": real_data["bps_object"], "rot_matrix": y_fake["rot_matrix"], "rot_6D": fake_rot_6D,
                         "transl": fake_transl, "joint_conf": fake_joint_conf}
            # The reference also back-propagates this pass into D's weights and throws the result away at the next
            # zero_grad (dexgangrasp.py:808-816); freezing D here skips those wgrad GEMMs without changing any output.
            d_params = [p for p in D.parameters() if p.requires_grad]
            for p in d_params:
                p.requires_grad_(False)
            try:
                fake_score_gen = D(fake_data)
            finally:
                for p in d_params:
                    p.requires_grad_(True)
            real_data["rot_matrix"] = real_data["rot_matrix"].view(real_data["bps_object"].shape[0], -1)
            total_loss_gen, loss_dict_gen = gen_loss(real_data, fake_data, fake_score_gen)
            self.optim_dexgangrasp_generator.zero_grad(set_to_none=getattr(self, '_zero_grad_to_none', True))
            total_loss_gen.backward()
            self._sync_grads(G)
            self.optim_dexgangrasp_generator.step()
        else:
            loss_dict_gen = self.last_loss_dict_gen
        loss_dict = loss_dict_disc.copy()
        loss_dict.update(loss_dict_gen)
        return loss_dict

"""Train-mode forward / backward of the DexGenerator, Discriminator and DexEvaluator on the kernels of libdexgg.

torch.autograd only records the graph: every Function below calls the C ABI (include/dexgg.h) for its arithmetic —
dexgg_linear_f32 (forward, dgrad with trans_w, wgrad with trans_a), the train-mode BatchNorm kernels, LeakyReLU backward,
column sums for the bias gradients, rot6d forward/backward (utils.py) and the multi-tensor Adam step.

Reference: ResBlock.forward (DexGanGrasp/models/networks.py:43-56), Generator.forward (:88-103), Discriminator.forward
(:148-159), DexEvaluator.forward (:315-337) in training mode, as driven by DexGanGrasp.update_dexgangrasp /
update_dexevaluator (DexGanGrasp/models/dexgangrasp.py:739-823).

Data parallelism: when torch.distributed is initialised with world_size > 1 the BatchNorm statistics
(forward: sum, centred sum of squares; backward: sum g, sum g*xhat) are summed over ranks, so the n-rank step equals the
1-rank step on the concatenated batch.
"""
import ctypes

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import _native

BN_EPS = 1e-5
LRELU_SLOPE = 0.2  # networks.py:41


def _L():
    return _native.lib()


def _st(dev):
    return _native.stream_ptr(dev)


def _world():
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def _allreduce_sum(t, kind="bn_fwd"):
    if _world() > 1:
        from . import dataparallel
        dataparallel.all_reduce_sum(t, kind)


def _check_cuda(x):
    if not x.is_cuda:
        raise RuntimeError("dexgg: training tensors must be on a CUDA (B200) device; there is no CPU path")


def _rowmajor(x):
    """A 2-D fp32 tensor whose rows are contiguous (any row pitch); copies only if needed."""
    if x.dtype != torch.float32:
        x = x.float()
    if x.dim() != 2 or x.stride(1) != 1 or (x.shape[0] > 1 and x.stride(0) < x.shape[1]):
        x = x.contiguous()
    return x


def _ld(x):
    return x.stride(0) if x.shape[0] > 1 else x.shape[1]


_PRECISION = "fp32"


def set_precision(p):
    """'fp32' (CUDA-core GEMMs, the 1e-4 parity mode) or 'bf16' (tcgen05 GEMMs: bf16 operands, fp32 accumulation in TMEM, fp32
    master weights / gradients / BatchNorm / Adam - BASELINE.json configs[4]).  Set by DexGanGrasp from cfg['precision']."""
    global _PRECISION
    if p not in ("fp32", "bf16"):
        raise ValueError(f"precision must be 'fp32' or 'bf16', got {p!r}")
    _PRECISION = p


def _up8(n):
    return (n + 7) // 8 * 8


def _to_bf16(x, transpose, R, C):
    """bf16 copy of the leading [R, C] part of a row-major fp32 matrix (or of its transpose), row pitch padded to 8
    elements (16 bytes, TMA)."""
    dev = x.device
    with torch.cuda.device(dev):
        if transpose:
            out = torch.empty((C, _up8(R)), dtype=torch.bfloat16, device=dev)
            _native.check(_L().dexgg_f32_to_bf16_t(x.data_ptr(), _ld(x), out.data_ptr(), out.shape[1], R, C, _st(dev)))
        else:
            out = torch.empty((R, _up8(C)), dtype=torch.bfloat16, device=dev)
            _native.check(_L().dexgg_f32_to_bf16(x.data_ptr(), _ld(x), out.data_ptr(), out.shape[1], R, C, _st(dev)))
    return out


def _raw_linear_bf16(A, trans_a, W, trans_w, M, N, K, bias, act):
    """Same contract as raw_linear on the tensor cores: both operands are brought to K-major bf16 (a transposing convert
    where the fp32 operand is stored with the contraction index as its row index), fp32 result."""
    dev = A.device
    a16 = _to_bf16(A, True, K, M) if trans_a else _to_bf16(A, False, M, K)   # [M, K]
    w16 = _to_bf16(W, True, K, N) if trans_w else _to_bf16(W, False, N, K)   # [N, K]
    out = torch.empty((M, N), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _native.check(_L().dexgg_linear_bf16(a16.data_ptr(), a16.shape[1], w16.data_ptr(), w16.shape[1], M, N, K, _native.ptr(bias),
                                             None, None, None, 0, int(act), None, out.data_ptr(), N, _st(dev)))
    return out


def raw_linear(A, trans_a, W, trans_w, M, N, K, bias=None, act=0, accumulate=0, out=None):
    """C[M,N] = act(op(A) . op(W)^T + bias) through dexgg_linear_f32 (or the tcgen05 kernel in bf16 mode)."""
    if _PRECISION == "bf16" and not accumulate and out is None and M >= 1:
        return _raw_linear_bf16(A, trans_a, W, trans_w, M, N, K, bias, act)
    dev = A.device
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _native.check(_L().dexgg_linear_f32(A.data_ptr(), _ld(A), int(trans_a), W.data_ptr(), _ld(W), int(trans_w), M, N, K,
                                            _native.ptr(bias), None, None, None, 0, int(act), int(accumulate),
                                            out.data_ptr(), _ld(out), _st(dev)))
    return out


def col_sum(x):
    M, N = x.shape
    s = torch.zeros((N,), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _native.check(_L().dexgg_col_sum(x.data_ptr(), _ld(x), M, N, s.data_ptr(), _st(x.device)))
    return s


def lrelu_backward(g, y):
    M, N = g.shape
    out = torch.empty((M, N), dtype=torch.float32, device=g.device)
    with torch.cuda.device(g.device):
        _native.check(_L().dexgg_lrelu_backward(g.data_ptr(), _ld(g), y.data_ptr(), _ld(y), M, N, out.data_ptr(), N, _st(g.device)))
    return out


class _LinearFn(torch.autograd.Function):
    """y = act(x W^T + b)   (nn.Linear [+ LeakyReLU(0.2)], networks.py:31-41)"""

    @staticmethod
    def forward(ctx, x, weight, bias, act):
        _check_cuda(x)
        x = _rowmajor(x)
        M, K = x.shape
        N = weight.shape[0]
        y = raw_linear(x, 0, weight, 0, M, N, K, bias=bias, act=1 if act else 0)
        ctx.act = bool(act)
        ctx.save_for_backward(x, weight, y if act else None)
        return y

    @staticmethod
    def backward(ctx, g):
        x, weight, y = ctx.saved_tensors
        g = _rowmajor(g)
        if ctx.act:
            g = lrelu_backward(g, y)
        M, K = x.shape
        N = weight.shape[0]
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:  # dgrad: gx[m,k] = sum_n g[m,n] W[n,k]
            gx = raw_linear(g, 0, weight, 1, M, K, N)
        if ctx.needs_input_grad[1]:  # wgrad: gw[n,k] = sum_m g[m,n] x[m,k]
            gw = raw_linear(g, 1, x, 1, N, K, M)
        if ctx.needs_input_grad[2]:
            gb = col_sum(g)
        return gx, gw, gb, None


def linear(x, weight, bias, act=False):
    return _LinearFn.apply(x, weight, bias, act)


class _BatchNormFn(torch.autograd.Function):
    """y = act(gamma * (x - mean_batch) / sqrt(var_batch + eps) + beta [+ residual]); running stats updated in place
    (momentum 0.1, unbiased variance) like nn.BatchNorm1d in training mode."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, momentum, act):
        _check_cuda(x)
        x = _rowmajor(x)
        M, N = x.shape
        dev = x.device
        L = _L()
        count = float(M * _world())
        mean = torch.empty((N,), dtype=torch.float32, device=dev)
        invstd = torch.empty((N,), dtype=torch.float32, device=dev)
        s = col_sum(x)
        _allreduce_sum(s)
        with torch.cuda.device(dev):
            st = _st(dev)
            _native.check(L.dexgg_bn_finalize(s.data_ptr(), 0, count, N, BN_EPS, momentum, mean.data_ptr(), None, None, None, st))
            ss = torch.zeros((N,), dtype=torch.float32, device=dev)
            _native.check(L.dexgg_col_centered_sumsq(x.data_ptr(), _ld(x), mean.data_ptr(), M, N, ss.data_ptr(), st))
            _allreduce_sum(ss)
            _native.check(L.dexgg_bn_finalize(ss.data_ptr(), 1, count, N, BN_EPS, momentum, mean.data_ptr(), invstd.data_ptr(),
                                              _native.ptr(running_mean), _native.ptr(running_var), st))
            y = torch.empty((M, N), dtype=torch.float32, device=dev)
            res = _rowmajor(residual) if residual is not None else None
            _native.check(L.dexgg_bn_apply(x.data_ptr(), _ld(x), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                                           beta.data_ptr(), _native.ptr(res), _ld(res) if res is not None else 0,
                                           1 if act else 0, M, N, y.data_ptr(), N, st))
        if running_mean is not None:  # written through raw pointers: tell torch (packed eval-mode weights key on the version)
            torch.autograd.graph.increment_version([running_mean, running_var])
        ctx.act = bool(act)
        ctx.count = count
        ctx.has_res = residual is not None
        ctx.save_for_backward(x, gamma, mean, invstd, y if act else None)
        return y

    @staticmethod
    def backward(ctx, g):
        x, gamma, mean, invstd, y = ctx.saved_tensors
        g = _rowmajor(g)
        if ctx.act:
            g = lrelu_backward(g, y)
        M, N = x.shape
        dev = x.device
        L = _L()
        sum_g = torch.zeros((N,), dtype=torch.float32, device=dev)
        sum_gx = torch.zeros((N,), dtype=torch.float32, device=dev)
        gx = None
        with torch.cuda.device(dev):
            st = _st(dev)
            _native.check(L.dexgg_bn_backward_reduce(g.data_ptr(), _ld(g), x.data_ptr(), _ld(x), mean.data_ptr(), invstd.data_ptr(),
                                                     M, N, sum_g.data_ptr(), sum_gx.data_ptr(), st))
            g_gamma, g_beta = sum_gx, sum_g  # local sums; the gradient allreduce averages them over ranks
            if ctx.needs_input_grad[0]:
                tot_g, tot_gx = sum_g, sum_gx
                if _world() > 1:
                    both = torch.stack([sum_g, sum_gx])
                    _allreduce_sum(both, "bn_bwd")
                    tot_g, tot_gx = both[0].contiguous(), both[1].contiguous()
                gx = torch.empty((M, N), dtype=torch.float32, device=dev)
                _native.check(L.dexgg_bn_backward_apply(g.data_ptr(), _ld(g), x.data_ptr(), _ld(x), mean.data_ptr(),
                                                        invstd.data_ptr(), gamma.data_ptr(), tot_g.data_ptr(), tot_gx.data_ptr(),
                                                        ctx.count, M, N, gx.data_ptr(), N, st))
        g_res = g if (ctx.has_res and ctx.needs_input_grad[3]) else None
        return gx, (g_gamma if ctx.needs_input_grad[1] else None), (g_beta if ctx.needs_input_grad[2] else None), g_res, None, None, None, None


def batch_norm(x, bn, act=False, residual=None, final_act=True):
    """Train-mode nn.BatchNorm1d `bn` applied to x, optionally followed by `+ residual` and LeakyReLU(0.2).
    act: LeakyReLU directly after the normalisation (ResBlock bn1); residual/final_act: ResBlock tail (networks.py:53-56)."""
    do_act = bool(act) or (residual is not None and bool(final_act))
    if bn.training:
        y = _BatchNormFn.apply(x, bn.weight, bn.bias, residual, bn.running_mean if bn.track_running_stats else None,
                               bn.running_var if bn.track_running_stats else None,
                               0.1 if bn.momentum is None else float(bn.momentum), do_act)
        if bn.track_running_stats and bn.num_batches_tracked is not None:
            bn.num_batches_tracked += 1
        return y
    raise RuntimeError("dexgg: eval-mode BatchNorm goes through the packed forward (dexgg_*_forward), not train_ops")


# ---------------------------------------------------------------------------------------------------- module forwards
def resblock(rb, x, final_nl=True):
    """ResBlock.forward (networks.py:43-56)"""
    Xin = x if rb.Fin == rb.Fout else linear(x, rb.fc3.weight, rb.fc3.bias, act=True)
    h = linear(x, rb.fc1.weight, rb.fc1.bias)
    h = batch_norm(h, rb.bn1, act=True)
    h = linear(h, rb.fc2.weight, rb.fc2.bias)
    return batch_norm(h, rb.bn2, act=False, residual=Xin, final_act=final_nl)


def generator_forward_train(gen, Zin, bps_object):
    """Generator.forward (networks.py:88-103), training mode: z FIRST in X0, X0 FIRST in the rb2 input."""
    dev = gen.gen_bn1.weight.device
    bps = bps_object.to(device=dev, dtype=torch.float32)
    z = Zin.to(device=dev, dtype=torch.float32)
    o_bps = batch_norm(bps, gen.gen_bn1)
    X0 = torch.cat([z, o_bps], dim=1)
    X = gen.gen_rb1(X0, True)
    X = gen.gen_rb2(torch.cat([X0, X], dim=1), True)
    return {
        "rot_6D": linear(X, gen.gen_rot.weight, gen.gen_rot.bias),
        "transl": linear(X, gen.gen_transl.weight, gen.gen_transl.bias),
        "joint_conf": linear(X, gen.gen_joint_conf.weight, gen.gen_joint_conf.bias),
        "z": Zin,
    }


def discriminator_forward_train(disc):
    """Discriminator.forward (networks.py:148-159) on the tensors set_input stored; X0 FIRST in the rb2 input."""
    X = torch.cat([disc.bps_object, disc.rot_matrix, disc.transl, disc.joint_conf], dim=1)
    X0 = batch_norm(X, disc.disc_bn1)
    X = disc.disc_rb1(X0)
    X = disc.disc_rb2(torch.cat([X0, X], dim=1))
    return torch.sigmoid(linear(X, disc.out_success.weight, disc.out_success.bias))


def evaluator_forward_train(eva, dropout_masks=None):
    """DexEvaluator.forward (networks.py:315-337), training mode: X FIRST in the rb2/rb3 inputs, Dropout(0.3) after every
    block.  dropout_masks (optional, three pre-scaled keep masks) replaces torch's Philox stream in parity tests."""
    def dout(x, i):
        if dropout_masks is not None:
            return x * dropout_masks[i]
        return eva.dout(x)
    X = torch.cat([eva.bps_object, eva.rot_matrix, eva.transl], dim=1)
    X0 = batch_norm(X, eva.bn1)
    X = dout(eva.rb1(X0), 0)
    X = dout(eva.rb2(torch.cat([X, X0], dim=1)), 1)
    X = dout(eva.rb3(torch.cat([X, X0], dim=1)), 2)
    return torch.sigmoid(linear(X, eva.out_success.weight, eva.out_success.bias))


# ------------------------------------------------------------------------------ WGAN-GP: twice-differentiable D pass
class _MatmulFn(torch.autograd.Function):
    """C[M,N] = op(A) . op(W)^T on the dexgg GEMM (ta: A stored [K,M]; tw: W stored [K,N]).  Its backward is two more
    _MatmulFn applications, so the op is differentiable to any order: the WGAN-GP gradient penalty
    (dexgangrasp.py:160-202) differentiates the discriminator's input gradient w.r.t. its weights."""

    @staticmethod
    def forward(ctx, A, W, ta, tw):
        _check_cuda(A)
        A = _rowmajor(A)
        W = _rowmajor(W)
        K, M = (A.shape if ta else A.shape[::-1])
        N = W.shape[1] if tw else W.shape[0]
        ctx.ta, ctx.tw = int(ta), int(tw)
        ctx.save_for_backward(A, W)
        return raw_linear(A, int(ta), W, int(tw), M, N, K)

    @staticmethod
    def backward(ctx, G):
        A, W = ctx.saved_tensors
        ta, tw = ctx.ta, ctx.tw
        gA = gW = None
        if ctx.needs_input_grad[0]:  # dA'[m,k] = sum_n G[m,n] W'[n,k]   (stored transposed when ta)
            gA = _MatmulFn.apply(W, G, 1 - tw, 0) if ta else _MatmulFn.apply(G, W, 0, 1 - tw)
        if ctx.needs_input_grad[1]:  # dW'[n,k] = sum_m G[m,n] A'[m,k]   (stored transposed when tw)
            gW = _MatmulFn.apply(A, G, 1 - ta, 1) if tw else _MatmulFn.apply(G, A, 1, 1 - ta)
        return gA, gW, None, None


def _linear_dd(x, lin, act=False):
    y = _MatmulFn.apply(x, lin.weight, 0, 0) + lin.bias
    return F.leaky_relu(y, LRELU_SLOPE) if act else y


class _AllReduceSumFn(torch.autograd.Function):
    """y = sum of x over the ranks.  Every rank differentiates its own loss through y, and the gradient of the ranks' summed loss
    w.r.t. this rank's x is the sum of those: the backward is the same all-reduce - hence differentiable to any order (the
    WGAN-GP penalty differentiates the discriminator's input gradient once more).  Every rank runs the same sequence of them."""

    @staticmethod
    def forward(ctx, x):
        from . import dataparallel
        y = x.contiguous().clone()
        dataparallel.all_reduce_sum(y, "bn_fwd")
        return y

    @staticmethod
    def backward(ctx, g):
        return _AllReduceSumFn.apply(g)


def _batch_norm_dd(x, bn):
    """Train-mode BatchNorm1d with a double backward (elementwise / per-column glue between the dexgg GEMMs).  One process:
    torch's own batch_norm.  Data parallel: the same arithmetic from GLOBAL batch statistics (two all-reduces of per-column sums,
    _AllReduceSumFn, so the N-rank step equals the one-process step on the concatenated batch), running statistics updated like
    nn.BatchNorm1d (momentum, unbiased variance)."""
    momentum = 0.1 if bn.momentum is None else float(bn.momentum)
    if _world() == 1:
        y = F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, True, momentum, BN_EPS)
    else:
        n_total = float(x.shape[0] * _world())
        mean = _AllReduceSumFn.apply(x.sum(0)) / n_total
        xc = x - mean
        var = _AllReduceSumFn.apply((xc * xc).sum(0)) / n_total
        y = xc * torch.rsqrt(var + BN_EPS) * bn.weight + bn.bias
        with torch.no_grad():
            bn.running_mean.mul_(1.0 - momentum).add_(mean.detach(), alpha=momentum)
            bn.running_var.mul_(1.0 - momentum).add_(var.detach() * (n_total / max(n_total - 1.0, 1.0)), alpha=momentum)
    if bn.num_batches_tracked is not None:
        bn.num_batches_tracked += 1
    return y


def _resblock_dd(rb, x):
    Xin = x if rb.Fin == rb.Fout else _linear_dd(x, rb.fc3, act=True)
    h = F.leaky_relu(_batch_norm_dd(_linear_dd(x, rb.fc1), rb.bn1), LRELU_SLOPE)
    h = _batch_norm_dd(_linear_dd(h, rb.fc2), rb.bn2)
    return F.leaky_relu(Xin + h, LRELU_SLOPE)


def discriminator_forward_double_backward(disc):
    """Discriminator.forward (networks.py:148-159), train mode, as a graph torch can differentiate twice: every GEMM of
    the forward, of the input-gradient pass and of the pass through that gradient runs on dexgg_linear_f32 / the
    tcgen05 kernel via _MatmulFn; BatchNorm / LeakyReLU / sigmoid are torch element-wise ops."""
    X = torch.cat([disc.bps_object, disc.rot_matrix, disc.transl, disc.joint_conf], dim=1)
    X0 = _batch_norm_dd(X, disc.disc_bn1)
    X = _resblock_dd(disc.disc_rb1, X0)
    X = _resblock_dd(disc.disc_rb2, torch.cat([X0, X], dim=1))
    return torch.sigmoid(_linear_dd(X, disc.out_success))


# ---------------------------------------------------------------------------------------------------- optimizer
# --------------------------------------------- evaluator-gradient grasp refinement (dexgangrasp.py:516-541, :619-660)
class EvaluatorRefineGraph:
    """Eval-mode DexEvaluator.forward (networks.py:315-337; BatchNorm on running statistics, Dropout inactive) as a graph
    that autograd differentiates w.r.t. rot_matrix / transl - the quantity improve_grasps_gradient_based steps along.
    The BPS columns of the three blocks' fc1 / fc3 products do not change while a grasp is refined, so they are computed
    ONCE here (6 GEMMs of K = 4096); every refinement step then costs GEMMs of K = 12 and K = 512 only.  All GEMMs,
    forward and input-gradient, are _MatmulFn -> dexgg_linear_f32 in fp32 (a <= 1 cm step along the gradient is only
    meaningful at full precision); BatchNorm affines / LeakyReLU / sigmoid between them are torch element-wise ops."""

    def __init__(self, eva, bps_object):
        _check_cuda(bps_object)
        bps = _rowmajor(bps_object.detach())
        n, nb = bps.shape
        nn_ = eva.out_success.weight.shape[1]
        with torch.no_grad(), fp32_gemms():
            s, t = _bn_affine(eva.bn1)
            xb = bps * s[:nb] + t[:nb]
            self.s_pose, self.t_pose = s[nb:].clone(), t[nb:].clone()
            self.blocks = []
            for rb, off in ((eva.rb1, 0), (eva.rb2, nn_), (eva.rb3, nn_)):
                blk = {}
                for name in ("fc1", "fc3"):
                    lin = getattr(rb, name)
                    W = lin.weight.detach()
                    Wb = W[:, off:off + nb].contiguous()
                    blk[name + "_base"] = raw_linear(xb, 0, Wb, 0, n, W.shape[0], nb, bias=lin.bias.detach())
                    blk[name + "_Wp"] = W[:, off + nb:].contiguous()
                    blk[name + "_Wx"] = W[:, :off].contiguous() if off else None
                blk["s1"], blk["t1"] = _bn_affine(rb.bn1)
                blk["s2"], blk["t2"] = _bn_affine(rb.bn2)
                blk["t2"] = blk["t2"] + rb.fc2.bias.detach() * blk["s2"]
                blk["W2"] = rb.fc2.weight.detach().contiguous()
                self.blocks.append(blk)
            self.Wout = eva.out_success.weight.detach().contiguous()
            self.bout = eva.out_success.bias.detach().clone()

    def __call__(self, rot_matrix, transl):
        n = transl.shape[0]
        with fp32_gemms():
            pose = torch.cat([rot_matrix.reshape(n, -1), transl], dim=1) * self.s_pose + self.t_pose
            X = None
            for blk in self.blocks:
                def lin(name):
                    y = blk[name + "_base"] + _MatmulFn.apply(pose, blk[name + "_Wp"], 0, 0)
                    return y if X is None else y + _MatmulFn.apply(X, blk[name + "_Wx"], 0, 0)
                Xin = F.leaky_relu(lin("fc3"), LRELU_SLOPE)
                h = F.leaky_relu(lin("fc1") * blk["s1"] + blk["t1"], LRELU_SLOPE)
                h = _MatmulFn.apply(h, blk["W2"], 0, 0) * blk["s2"] + blk["t2"]
                X = F.leaky_relu(Xin + h, LRELU_SLOPE)
            return torch.sigmoid(_MatmulFn.apply(X, self.Wout, 0, 0) + self.bout)


def _bn_affine(bn):
    """Eval-mode BatchNorm1d as y = x * s + t."""
    s = bn.weight.detach() / torch.sqrt(bn.running_var + bn.eps)
    return s, bn.bias.detach() - bn.running_mean * s


class fp32_gemms:
    """Scope in which raw_linear uses the fp32 CUDA-core GEMM whatever cfg['precision'] says."""

    def __enter__(self):
        global _PRECISION
        self.saved, _PRECISION = _PRECISION, "fp32"

    def __exit__(self, *a):
        global _PRECISION
        _PRECISION = self.saved


class _AdamChunk(ctypes.Structure):
    _fields_ = [("w", ctypes.c_void_p), ("g", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p), ("n", ctypes.c_int)]


class DexggAdam(torch.optim.Adam):
    """torch.optim.Adam (dexgangrasp.py:97-112: betas (beta1, 0.999), coupled weight decay) whose step() is ONE
    dexgg_adam_step launch over all parameters of a group.  State layout ('step', 'exp_avg', 'exp_avg_sq') and state_dict
    are torch's, so optimizer checkpoints written by the reference load unchanged."""

    CHUNK = 1 << 13  # elements per CTA of the fused kernel (many small CTAs keep the HBM pipes full)
    device_step = False  # True: the step count is kept in device memory too (dexgg_adam_step_dev), as CUDA-graph capture needs

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            if group.get("amsgrad") or group.get("maximize"):
                raise NotImplementedError("DexggAdam: amsgrad / maximize are not used by the reference")
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            dev = params[0].device
            _check_cuda(params[0])
            steps = set()
            chunks = []
            for p in params:
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                steps.add(int(st["step"].item()) if torch.is_tensor(st["step"]) else int(st["step"]))
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                if g is not p.grad:
                    p.grad = g
                n = p.numel()
                for o in range(0, n, self.CHUNK):
                    c = min(self.CHUNK, n - o)
                    chunks.append((p.data_ptr() + 4 * o, g.data_ptr() + 4 * o, st["exp_avg"].data_ptr() + 4 * o,
                                   st["exp_avg_sq"].data_ptr() + 4 * o, c))
            if len(steps) != 1:
                raise RuntimeError("DexggAdam: parameters of one group must share the step count")
            key = tuple(chunks)
            if getattr(self, "_table_key", None) != key:  # chunk table lives on the device; rebuilt only when pointers move
                arr = (_AdamChunk * len(chunks))(*[_AdamChunk(*c) for c in chunks])
                # pinned staging + async copy: legal inside a CUDA-graph capture (the pinned buffer is kept alive with the table)
                self._table_host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).pin_memory()
                self._table = self._table_host.to(dev, non_blocking=True)
                self._table_key = key
            b1, b2 = group["betas"]
            lr = group["lr"]
            step_now = steps.pop()
            with torch.cuda.device(dev):
                if self.device_step:
                    if getattr(self, "_step_dev", None) is None:  # created outside any capture (first, eager, step)
                        self._step_dev = torch.tensor([step_now - 1], dtype=torch.int32, device=dev)
                    _native.check(_L().dexgg_adam_step_dev(self._table.data_ptr(), len(chunks), float(lr), float(b1), float(b2),
                                                           float(group["eps"]), float(group["weight_decay"]),
                                                           self._step_dev.data_ptr(), _st(dev)))
                else:
                    _native.check(_L().dexgg_adam_step(self._table.data_ptr(), len(chunks), float(lr), float(b1), float(b2),
                                                       float(group["eps"]), float(group["weight_decay"]), step_now, _st(dev)))
            # parameters changed behind torch's back: bump versions so packed inference weights refresh
            torch.autograd.graph.increment_version(params)
        return loss

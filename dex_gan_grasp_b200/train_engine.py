"""The bf16 training step of DexGanGrasp.update_dexgangrasp / update_dexevaluator (DexGanGrasp/models/dexgangrasp.py:739-823)
as a hand-scheduled sequence of fused sm_100a kernels (csrc/gemm_umma.cu, csrc/train_fused.cu) - no autograd graph, no
torch.cat, no dtype-conversion passes, ~130 launches per GAN + evaluator step (the autograd-Function path of train_ops.py
needs ~600), capturable as one CUDA graph (graph_step.CapturedTrainStep).

What makes the schedule short:
  * every network's input BatchNorm sees the same BPS batch: its statistics and xhat = (bps - mean) / std are computed ONCE
    per step; each network's gamma / beta are folded into its first-layer weights when they are converted to bf16
    (dexgg_prep_weights), and dexgg_unfold_grads maps the gradients of the folded weights back to fc1 / fc3 .weight and to the
    input BatchNorm's own gamma / beta - so no GEMM computes a gradient wrt the 4096 BPS columns;
  * torch.cat([X0, X]) (networks.py:97,153,325) is a second K segment of the GEMM;
  * fc1 and fc3 of a ResBlock are one N = 768 GEMM whose epilogue applies LeakyReLU to the fc3 columns;
  * a train-mode BatchNorm (+ skip, LeakyReLU, dropout) is one column-slab kernel forward and one backward;
  * dgrad / wgrad read weights and activations in their natural layouts (MN-major tcgen05 operands): no transposes;
  * the discriminator's real and fake batches of the D step run as ONE double batch with per-half statistics.
Parameters stay fp32 nn.Parameters (master weights); their .grad tensors are views of one flat buffer per network, filled by
the kernels; the optimizers are the reference's (train_ops.DexggAdam = torch.optim.Adam state + one fused launch).
"""
import ctypes

import torch
import torch.distributed as dist

from . import _native
import os

from ._native import BnBwd, BnFwd, FirstLayer, Gemm, PlainJob, RunEntry, ScoreHead, SlabComm, TailSrc, WeightJobs

BPS, HID, INNER, NCAT = 4096, 512, 256, 768
TAILW = 64          # the tail columns are stored 64 wide (zero padded): 128-byte rows for TMA, N tiles of 64 for the dgrad
LDXT = HID + TAILW  # [X (512) | tail (64)]
EPS = 1e-5


def _L():
    return _native.lib()


def _st(dev):
    return _native.stream_ptr(dev)


def _world():
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


class _Block:
    """Static description + device buffers of one ResBlock inside a network."""

    def __init__(self, rb, cmap, dmap, has_x, dev):
        self.rb, self.has_x = rb, has_x
        self.Fin = rb.fc1.weight.shape[1]
        self.bps_src0 = dmap.index(0)  # the 4096 BPS columns are contiguous in every block's input
        assert dmap[self.bps_src0:self.bps_src0 + BPS] == list(range(BPS)) and cmap[self.bps_src0:self.bps_src0 + BPS] == list(range(BPS))
        self.cmap = torch.tensor(cmap, dtype=torch.int32, device=dev)
        self.dmap = torch.tensor(dmap, dtype=torch.int32, device=dev)
        self.ldxt = LDXT if has_x else TAILW
        self.Wb = torch.zeros((NCAT, BPS), dtype=torch.bfloat16, device=dev)
        self.Wxt = torch.zeros((NCAT, self.ldxt), dtype=torch.bfloat16, device=dev)
        self.W2 = torch.zeros((HID, INNER), dtype=torch.bfloat16, device=dev)
        self.bias13 = torch.zeros((NCAT,), dtype=torch.float32, device=dev)
        self.dWb = torch.zeros((NCAT, BPS), dtype=torch.float32, device=dev)
        self.dWxt = torch.zeros((NCAT, self.ldxt), dtype=torch.float32, device=dev)


class _Net:
    """One network (generator / discriminator / evaluator): blocks, input BatchNorm, bf16 operand buffers, activations."""

    def __init__(self, kind, module, in_bn, blocks, n_tail, tail_bn, head, rows, groups, dev, optimizer):
        self.kind, self.module, self.in_bn, self.n_tail, self.tail_bn, self.head = kind, module, in_bn, n_tail, tail_bn, head
        self.rows, self.groups, self.dev, self.optimizer = rows, groups, dev, optimizer
        self.blocks = blocks
        nb = len(blocks)
        R = rows
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.tail1 = torch.zeros((R, TAILW), **bf)
        self.XT = [None] + [torch.zeros((R, LDXT), **bf) for _ in range(nb - 1)]  # input of block b >= 1: [X_{b-1} | tail]
        self.Xlast = torch.zeros((R, HID), **bf)
        self.T = [torch.zeros((R, NCAT), **bf) for _ in range(nb)]
        self.H = [torch.zeros((R, INNER), **bf) for _ in range(nb)]
        self.Y = [torch.zeros((R, HID), **bf) for _ in range(nb)]
        self.stat1 = [torch.zeros((2, groups, INNER), **f32) for _ in range(nb)]  # [mean | invstd]
        self.stat2 = [torch.zeros((2, groups, HID), **f32) for _ in range(nb)]
        # gradients wrt a block's second / first layer output: one buffer PER BLOCK, so that the weight-gradient GEMMs reading
        # them can run on the network's side stream while the chain moves on to the next block (backward_blocks)
        self.GY = [torch.zeros((R, HID), **bf) for _ in range(nb)]
        self.GH = torch.zeros((R, INNER), **bf)
        self.G13 = [torch.zeros((R, NCAT), **bf) for _ in range(nb)]
        self.wstream = None
        self.gXT = [torch.zeros((R, TAILW), **bf)] + [torch.zeros((R, LDXT), **bf) for _ in range(nb - 1)]
        self.tail_stat = torch.zeros((groups, 2, TAILW), **f32)  # per group: mean | invstd of the tail columns
        # column sums of the first-layer output gradient (fc3 part) and the gradients of the [X | tail] weight columns: one flat
        # scratch zeroed once per backward (bias sums and split contractions accumulate with atomics)
        n_small = nb * NCAT + sum(NCAT * blk.ldxt for blk in blocks)
        self.small_scratch = torch.zeros((n_small,), **f32)
        self.db13 = self.small_scratch[:nb * NCAT].view(nb, NCAT)
        off = nb * NCAT
        for blk in blocks:
            blk.dWxt = self.small_scratch[off:off + NCAT * blk.ldxt].view(NCAT, blk.ldxt)
            off += NCAT * blk.ldxt
        self.dl = torch.zeros((R,), **f32)
        self.p = torch.zeros((R,), **f32)
        # flat gradient buffer; every parameter's .grad is a view into it (one memset to zero, one allreduce to exchange)
        self.params = [p for p in module.parameters()]
        # every slice starts at a multiple of 4 elements (16 bytes): the fused Adam kernel then moves float4s
        self.flat_grad = torch.zeros((sum((p.numel() + 3) // 4 * 4 for p in self.params),), **f32)
        module._dexgg_flat_grad = self.flat_grad  # dataparallel.allreduce_gradients exchanges this buffer in place
        self.attach()

    def attach(self):
        """(Re-)point every parameter's .grad at its slice of the flat buffer (someone may have set it to None)."""
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + 4 * off:
                p.grad = self.flat_grad[off:off + n].view_as(p)
            off += (n + 3) // 4 * 4


def _maps(kind, b, n_tail, Fin):
    """cmap / dmap of block b's input columns in the reference's torch.cat order (networks.py:92-97, 150-153, 318-325)."""
    cmap, dmap = [-1] * Fin, [0] * Fin
    xo = BPS + (HID if b > 0 else 0)  # first tail column inside [bps 4096 | X 512 | tail]
    for c in range(Fin):
        if kind == "G":      # [z | bps] , [z | bps | X]
            if c < n_tail:
                dmap[c] = xo + c
            elif c < n_tail + BPS:
                cmap[c], dmap[c] = c - n_tail, c - n_tail
            else:
                dmap[c] = BPS + (c - n_tail - BPS)
        elif kind == "D":    # [bps | pose] , [bps | pose | X]
            if c < BPS:
                cmap[c], dmap[c] = c, c
            elif c < BPS + n_tail:
                cmap[c], dmap[c] = c, xo + (c - BPS)
            else:
                dmap[c] = BPS + (c - BPS - n_tail)
        else:                # E: [bps | pose] , [X | bps | pose]
            x = HID if b > 0 else 0
            if c < x:
                dmap[c] = BPS + c
            elif c < x + BPS:
                cmap[c], dmap[c] = c - x, c - x
            else:
                cmap[c], dmap[c] = BPS + (c - x - BPS), xo + (c - x - BPS)
    return cmap, dmap


class _DeviceFloats:
    """A float32 device buffer this library allocated (cudaMalloc through dexgg_peer_alloc), exposed to torch by the CUDA array
    interface; `owner` keeps the allocation alive as long as a tensor made from it is."""

    def __init__(self, ptr, numel, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None}


def _wrap_device_floats(ptr, numel, dev, owner=None):
    with torch.cuda.device(dev):
        t = torch.as_tensor(_DeviceFloats(ptr, numel, owner), device=dev)
    assert t.data_ptr() == ptr and t.numel() == numel and t.dtype == torch.float32
    return t


class TrainEngine:
    def __init__(self, model, batch_rows):
        """model: dexgangrasp.DexGanGrasp wrapper in training mode; batch_rows: rows per rank."""
        self.model = model
        dev = model.device
        self.dev = dev
        B = int(batch_rows)
        self.B = B
        G, D, E = model.DexGANGrasp.generator, model.DexGANGrasp.discriminator, model.DexEvaluator
        self.nJ = G.gen_joint_conf.out_features
        self.zD = G.latentD
        nt_pose = 12 + self.nJ

        def blocks(kind, rbs, n_tail):
            return [_Block(rb, *_maps(kind, b, n_tail, rb.fc1.weight.shape[1]), b > 0, dev) for b, rb in enumerate(rbs)]
        self.G = _Net("G", G, G.gen_bn1, blocks("G", [G.gen_rb1, G.gen_rb2], self.zD), self.zD, False, None, B, 1, dev,
                      model.optim_dexgangrasp_generator)
        self.D = _Net("D", D, D.disc_bn1, blocks("D", [D.disc_rb1, D.disc_rb2], nt_pose), nt_pose, True, D.out_success, 2 * B, 2, dev,
                      model.optim_dexgangrasp_discriminator)
        self.E = _Net("E", E, E.bn1, blocks("E", [E.rb1, E.rb2, E.rb3], 12), 12, True, E.out_success, B, 1, dev, model.optim_dexevaluator)
        f32 = dict(dtype=torch.float32, device=dev)
        self.bps_stat = torch.zeros((3, BPS), **f32)  # mean | var | invstd of the step's BPS batch
        self.xhat = torch.zeros((2 * B, BPS), dtype=torch.bfloat16, device=dev)
        # generator heads: rows (joint nJ | rot6d 6 | transl 3) padded to 32
        self.Wh = torch.zeros((32, HID), dtype=torch.bfloat16, device=dev)
        self.bh = torch.zeros((32,), **f32)
        self.OUT = torch.zeros((B, 32), **f32)
        self.gHeads = torch.zeros((B, 32), dtype=torch.bfloat16, device=dev)
        self.dWh = torch.zeros((32, HID), **f32)
        self.dbh = torch.zeros((32,), **f32)
        self.gX = torch.zeros((B, HID), dtype=torch.bfloat16, device=dev)
        self.fake_rot = torch.zeros((B, 9), **f32)
        self.fake_transl = torch.zeros((B, 3), **f32)
        self.fake_joint = torch.zeros((B, self.nJ), **f32)
        self.gPose = torch.zeros((B, nt_pose), **f32)
        self.loss = torch.zeros((16,), **f32)  # 0 bce_real 1 bce_fake 2 gen_fake 3 transl 4 rot 5 conf 6 eva
        self.seed = torch.zeros((1,), dtype=torch.int64, device=dev)  # dropout seed, bumped on the device every evaluator step
        self.seed.fill_(0x1234567)
        mix = torch.zeros((10, 16), **f32)
        bw = float(model.bce_weight)
        mix[0, 0] = mix[0, 1] = 0.5   # total_loss_disc = (real + fake) / 2
        mix[1, 0] = 1.0               # bce_loss_real
        mix[2, 1] = 1.0               # bce_loss_fake
        mix[3, 2] = 1.0               # gen_loss_fake
        mix[4, 3] = 1.0               # transl_loss (coefficient applied in the kernel)
        mix[5, 4] = 1.0               # rot_loss
        mix[6, 5] = 1.0               # conf_loss
        mix[7, 2] = mix[7, 3] = mix[7, 4] = mix[7, 5] = 1.0  # total_loss_gen
        mix[8, 6] = 1.0               # total_loss_eva
        self.mix, self.bw = mix, bw
        # per-launch accumulator slots of the column-slab kernels ([2 groups, 2, 512] floats each; the whole scratch is zeroed once
        # at the start of a step and every launch of the step takes the next slot) + the words of their grid barrier
        self.SLOT = 4096
        self.slab_ctas = 0  # 0: the library's default slab grid; combined_step sets 1 while its two chains run concurrently
        self.world = _world()
        self.rank = dist.get_rank() if self.world > 1 else 0
        self._closed = False
        self._setup_exchange()
        self.grad_ctas = int(os.environ.get("DEXGG_GRAD_CTAS", "0"))
        self.wgrad_streams = os.environ.get("DEXGG_WGRAD_STREAMS", "1") != "0"
        self.two_chains = os.environ.get("DEXGG_TWO_CHAINS", "1") != "0"
        self._setup_grad_exchange()
        self.scratch = torch.zeros((64 * self.SLOT,), **f32)
        self.bar = torch.zeros((4,), dtype=torch.int32, device=dev)
        self._slot = 0

    def _setup_exchange(self):
        """Peer-mapped scratch of the slab kernels' in-kernel exchange (csrc/train_fused.cu: slab_exchange): every rank allocates
        one, the cudaIpc handles travel through torch.distributed, every rank maps the others'.  DEXGG_DP_EXCHANGE=nccl keeps the
        NCCL fallback (one allreduce per BatchNorm between a sums-only and an apply launch)."""
        L = _L()
        self.p2p = (os.environ.get("DEXGG_DP_EXCHANGE", "p2p") != "nccl" and self.world <= 8) or self.world == 1
        self._pslot = 0
        if not self.p2p:
            return
        with torch.cuda.device(self.dev):
            nbytes = L.dexgg_slab_scratch_bytes()
            mine, handle = ctypes.c_void_p(), (ctypes.c_ubyte * 64)()
            _native.check(L.dexgg_peer_alloc(nbytes, ctypes.byref(mine), handle))
            self._peer_mine, self._peer_opened = mine, []
            ptrs = [mine.value]
            if self.world > 1:
                handles = [None] * self.world
                dist.all_gather_object(handles, bytes(handle))
                ptrs = []
                for r, h in enumerate(handles):
                    if r == self.rank:
                        ptrs.append(mine.value)
                        continue
                    q = ctypes.c_void_p()
                    _native.check(L.dexgg_peer_open((ctypes.c_ubyte * 64).from_buffer_copy(h), ctypes.byref(q)))
                    self._peer_opened.append(q)
                    ptrs.append(q.value)
            self.peer_table = torch.tensor(ptrs, dtype=torch.int64, device=self.dev)
            self.parity = torch.zeros((1,), dtype=torch.int32, device=self.dev)
            if self.world > 1:
                dist.barrier()  # every rank has mapped every scratch before anyone launches into it

    def _setup_grad_exchange(self):
        """Data parallel over the peer mapping: the three flat gradient buffers move into ONE peer-mapped allocation per rank, so
        dexgg_grad_allreduce (two-shot over NVLink, one launch, in the step's own stream) replaces the NCCL gradient allreduce."""
        self.grad_p2p = bool(self.p2p and self.world > 1 and os.environ.get("DEXGG_DP_GRADS", "p2p") != "nccl")
        if not self.grad_p2p:
            return
        L = _L()
        nets = (self.G, self.D, self.E)
        sizes = [(n.flat_grad.numel() + 63) // 64 * 64 for n in nets]  # 256-byte aligned sub-buffers
        with torch.cuda.device(self.dev):
            mine, handle = ctypes.c_void_p(), (ctypes.c_ubyte * 64)()
            _native.check(L.dexgg_peer_alloc(4 * sum(sizes), ctypes.byref(mine), handle))
            self._grad_mine = mine
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle))
            bases = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    bases.append(mine.value)
                    continue
                q = ctypes.c_void_p()
                _native.check(L.dexgg_peer_open((ctypes.c_ubyte * 64).from_buffer_copy(h), ctypes.byref(q)))
                self._peer_opened.append(q)
                bases.append(q.value)
            off = 0
            for n, sz in zip(nets, sizes):
                numel = n.flat_grad.numel()
                n.flat_grad = _wrap_device_floats(mine.value + 4 * off, numel, self.dev, owner=self)
                n.module._dexgg_flat_grad = n.flat_grad
                n.grad_peers = torch.tensor([b + 4 * off for b in bases], dtype=torch.int64, device=self.dev)
                n.attach()
                off += sz
            dist.barrier()

    def sync_grads(self, net):
        """Average `net`'s gradients over the ranks (dexgangrasp._sync_grads for the engine's flat buffers)."""
        if self.world == 1:
            return
        if not self.grad_p2p:
            self.model._sync_grads(net.module)
            return
        from . import dataparallel
        c = SlabComm()
        c.peers, c.world, c.rank, c.slot, c.parity = self.peer_table.data_ptr(), self.world, self.rank, self._pslot, self.parity.data_ptr()
        self._pslot += 1
        assert self._pslot <= 64
        dataparallel._COUNTERS["p2p_calls"] = dataparallel._COUNTERS.get("p2p_calls", 0) + 1
        dataparallel._COUNTERS["p2p_grad_bytes"] = dataparallel._COUNTERS.get("p2p_grad_bytes", 0) + 4 * net.flat_grad.numel()
        _native.check(_L().dexgg_grad_allreduce(ctypes.byref(c), net.grad_peers.data_ptr(), net.flat_grad.numel(), 1.0 / self.world,
                                                self.grad_ctas, _st(self.dev)))

    def close(self):
        """Release what this engine allocated outside torch's allocator: the peer-mapped exchange scratch and gradient buffers
        (cudaMalloc + cudaIpc mappings of the other ranks').  Gradients that live in them are detached from the parameters first.
        Every rank must close at the same point of its program (the mappings of a rank's buffer are closed by its peers)."""
        if getattr(self, "_closed", True):
            return
        self._closed = True
        L = _L()
        with torch.cuda.device(self.dev):
            torch.cuda.synchronize(self.dev)
            if getattr(self, "grad_p2p", False):
                for n in (self.G, self.D, self.E):
                    for p_ in n.params:
                        p_.grad = None
                    n.module._dexgg_flat_grad = None
                    n.flat_grad = None
            if self.world > 1 and dist.is_initialized():
                dist.barrier()  # nobody is still inside a kernel that writes into a peer's buffer
            for q in getattr(self, "_peer_opened", []):
                L.dexgg_peer_close(q)
            self._peer_opened = []
            if self.world > 1 and dist.is_initialized():
                dist.barrier()  # every mapping is closed before its owner frees the allocation
            for name in ("_grad_mine", "_peer_mine"):
                ptr = getattr(self, name, None)
                if ptr is not None and ptr.value:
                    L.dexgg_peer_free(ptr)
                    setattr(self, name, None)

    def __del__(self):
        try:
            if self.world == 1:  # multi-rank teardown needs the collective close() call; a finaliser cannot provide it
                self.close()
        except Exception:
            pass

    def _comm(self, a):
        c = a.comm
        from . import dataparallel
        c.peers, c.world, c.rank, c.slot, c.parity = self.peer_table.data_ptr(), self.world, self.rank, self._pslot, self.parity.data_ptr()
        self._pslot += 1
        assert self._pslot <= 64
        if self.world > 1:
            dataparallel._COUNTERS["p2p_calls"] = dataparallel._COUNTERS.get("p2p_calls", 0) + 1

    def _next_slot(self, n=1):
        """index of the next `n` free accumulator slots of this step"""
        i = self._slot
        self._slot += n
        assert self._slot <= 64
        return i

    def _slot_ptr(self, i):
        return self.scratch.data_ptr() + 4 * self.SLOT * i

    def _slot_view(self, i, numel):
        return self.scratch[self.SLOT * i:self.SLOT * i + numel]

    def _allreduce(self, t, kind):
        """Cross-rank sum of a small fp32 buffer: through the peer scratch (one launch) or NCCL."""
        if self.p2p and self.world > 1 and t.numel() <= 4 * 2048 and t.is_contiguous() and t.dtype == torch.float32:
            self._p2p_small(t, sum_out=t)
            return
        from . import dataparallel
        dataparallel.all_reduce_sum(t, kind)

    def _p2p_small(self, src, sum_out=None, gather_out=None):
        from . import dataparallel
        c = SlabComm()
        n = src.numel()
        nb = (n + 2047) // 2048
        c.peers, c.world, c.rank, c.slot, c.parity = self.peer_table.data_ptr(), self.world, self.rank, self._pslot, self.parity.data_ptr()
        self._pslot += nb
        assert self._pslot <= 64
        dataparallel._COUNTERS["p2p_calls"] = dataparallel._COUNTERS.get("p2p_calls", 0) + 1
        _native.check(_L().dexgg_p2p_small(ctypes.byref(c), src.data_ptr(), n, _native.ptr(sum_out), _native.ptr(gather_out), _st(self.dev)))

    def _gather_merge(self, local, n_sets, C, n_local, mean, invstd, var=None, kind="bn_fwd"):
        """local [n_sets, 2, C] = per-rank (mean, M2) -> all_gather -> Chan merge (equal row counts) -> mean / invstd [/ var] [n_sets, C]"""
        from . import dataparallel
        W = self.world
        gathered = torch.empty((W,) + tuple(local.shape), dtype=torch.float32, device=self.dev)
        if self.p2p and local.numel() <= 4 * 2048:
            self._p2p_small(local.contiguous(), gather_out=gathered)
        else:
            dataparallel.all_gather(gathered, local.contiguous(), kind)
        _native.check(_L().dexgg_bn_merge(gathered.data_ptr(), W, n_sets, C, float(n_local), EPS, mean, invstd, var, _st(self.dev)))

    def _begin_step(self):
        self.scratch.zero_()
        self._slot = 0
        if self.p2p:
            self._pslot = 0
            _native.check(_L().dexgg_slab_step_begin(self._peer_mine, self.parity.data_ptr(), _st(self.dev)))

    # ------------------------------------------------------------------------------------------------ kernel wrappers
    def _gemm(self, M, N, segs, bias=None, act_from=None, C16=None, C32=None, ldc=None, split_k=0, accumulate=0, tile_n=0):
        g = Gemm()
        g.M, g.N, g.n_seg = int(M), int(N), len(segs)
        for i, (A, lda, a_mn, Bm, ldb, b_mn, K) in enumerate(segs):
            s = g.seg[i]
            s.A, s.lda, s.a_mn_major, s.B, s.ldb, s.b_mn_major, s.K = A, int(lda), int(a_mn), Bm, int(ldb), int(b_mn), int(K)
        g.bias = bias.data_ptr() if bias is not None else None
        g.act_from = int(N if act_from is None else act_from)
        g.C16 = C16.data_ptr() if C16 is not None else None
        g.C32 = C32.data_ptr() if C32 is not None else None
        g.ldc, g.accumulate, g.split_k, g.tile_n = int(ldc), int(accumulate), int(split_k), int(tile_n)
        _native.check(_L().dexgg_gemm_bf16(ctypes.byref(g), _st(self.dev)))

    def _bn_fwd(self, x, ldx, C, rpg, groups, bn, stat, y, ldy, res=None, ldr=0, drop=None):
        a = BnFwd()
        a.x, a.ldx, a.C, a.rows_per_group, a.n_groups = x, ldx, C, rpg, groups
        a.gamma, a.beta, a.shift = bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr()
        a.mean, a.invstd = stat[0].data_ptr(), stat[1].data_ptr()
        slot = self._next_slot()
        a.acc = self._slot_ptr(slot)
        a.eps, a.count_total = EPS, float(rpg * self.world)
        a.residual, a.ldr, a.act = res, ldr, 1
        if drop is not None:
            mask, p, stream = drop
            if mask is not None:
                a.drop_mask, a.ldm = mask.data_ptr(), mask.stride(0)
            a.drop_p, a.drop_seed, a.drop_stream = float(p), self.seed.data_ptr(), int(stream)
        a.y, a.ldy, a.mode = y, ldy, 0
        a.ctas_per_sm = self.slab_ctas
        if self.p2p:
            self._comm(a)
        elif self.world > 1:  # NCCL fallback: local shifted sums -> ONE allreduce -> normalise with the global statistics
            a.mode = 1
            _native.check(_L().dexgg_bn_fwd_slab(ctypes.byref(a), _st(self.dev)))
            self._allreduce(self._slot_view(slot, groups * 2 * C), "bn_fwd")
            a.mode = 2
        _native.check(_L().dexgg_bn_fwd_slab(ctypes.byref(a), _st(self.dev)))

    def _bn_bwd(self, C, rpg, groups, bn, stat, x, ldx, out, ldo, gx, ldgx, g=None, ldg=0, rank1=None, skip=None, lds=0, gskip=None, ldgs=0,
                dbias_skip=None, want_wgrad=True, drop=None):
        a = BnBwd()
        if g is not None:
            a.g, a.ldg = g, ldg
        else:
            dl, w, dw_out = rank1
            a.dl, a.w, a.dw_out = dl.data_ptr(), w.data_ptr(), (dw_out.data_ptr() if (dw_out is not None and want_wgrad) else None)
        a.out, a.ldo, a.act = out, ldo, 1
        if drop is not None:
            mask, p, stream = drop
            if mask is not None:
                a.drop_mask, a.ldm = mask.data_ptr(), mask.stride(0)
            a.drop_p, a.drop_seed, a.drop_stream = float(p), self.seed.data_ptr(), int(stream)
        a.x, a.ldx = x, ldx
        a.mean, a.invstd, a.gamma = stat[0].data_ptr(), stat[1].data_ptr(), bn.weight.data_ptr()
        a.C, a.rows_per_group, a.n_groups, a.count_total = C, rpg, groups, float(rpg * self.world)
        if want_wgrad:
            a.dgamma, a.dbeta = bn.weight.grad.data_ptr(), bn.bias.grad.data_ptr()
        a.gx, a.ldgx = gx, ldgx
        if skip is not None:
            a.skip, a.lds, a.gskip, a.ldgs = skip, lds, gskip, ldgs
            a.dbias_skip = dbias_skip if want_wgrad else None
        a.mode = 0
        slot = self._next_slot()
        a.acc = self._slot_ptr(slot)
        a.ctas_per_sm = self.slab_ctas
        if self.p2p:
            self._comm(a)
        elif self.world > 1:
            a.mode = 1
            _native.check(_L().dexgg_bn_bwd_slab(ctypes.byref(a), _st(self.dev)))
            self._allreduce(self._slot_view(slot, groups * 2 * C), "bn_bwd")
            a.mode = 2
        _native.check(_L().dexgg_bn_bwd_slab(ctypes.byref(a), _st(self.dev)))

    @staticmethod
    def _tail_src(tensors):
        s = TailSrc()
        f64 = tensors[0].dtype == torch.float64
        for i in range(3):
            if i < len(tensors):
                t = tensors[i]
                assert t.is_contiguous() and (t.dtype == torch.float64) == f64 and t.dtype in (torch.float32, torch.float64)
                s.p[i], s.n[i] = t.data_ptr(), t.numel() // t.shape[0]
            else:
                s.p[i], s.n[i] = None, 0
        s.f64 = int(f64)
        return s

    def _tail_pack(self, net, tensors, row0, rows, group):
        """Normalise (train-mode BatchNorm statistics of this group) and write the tail columns of rows [row0, row0 + rows)."""
        src = self._tail_src(tensors)
        dsts = [net.tail1] + list(net.XT[1:])  # tail-only input of block 0, then the [X | tail] inputs of the later blocks
        P, LD, C0 = (ctypes.c_void_p * 3)(), (ctypes.c_int64 * 3)(), (ctypes.c_int * 3)()
        for i in range(3):
            if i < len(dsts):
                P[i], LD[i], C0[i] = dsts[i][row0:].data_ptr(), dsts[i].stride(0), (0 if i == 0 else HID)
            else:
                P[i], LD[i], C0[i] = None, 0, 0
        st = net.tail_stat[group]
        rm = rv = None
        if net.tail_bn:
            rm = net.in_bn.running_mean.data_ptr() + 4 * BPS
            rv = net.in_bn.running_var.data_ptr() + 4 * BPS
        args = (ctypes.byref(src), rows, int(net.tail_bn), EPS, float(rows * self.world), float(net.in_bn.momentum or 0.1),
                st[0].data_ptr(), st[1].data_ptr(), rm, rv, P, LD, C0)
        if self.world > 1 and net.tail_bn:
            T = net.n_tail
            slot = self._next_slot()
            _native.check(_L().dexgg_tail_pack(*args, 1, self._slot_ptr(slot), _st(self.dev)))
            self._gather_merge(self._slot_view(slot, 2 * T).view(1, 2, T), 1, T, rows, st[0].data_ptr(), st[1].data_ptr())
            _native.check(_L().dexgg_tail_pack(*args, 2, None, _st(self.dev)))
        else:
            _native.check(_L().dexgg_tail_pack(*args, 0, None, _st(self.dev)))

    # ------------------------------------------------------------------------------------------------ weights
    def _weight_jobs(self, net, grads):
        J = WeightJobs()
        J.n_first = len(net.blocks)
        for i, blk in enumerate(net.blocks):
            rb, F = blk.rb, J.first[i]
            F.W1, F.W3, F.b1, F.b3 = rb.fc1.weight.data_ptr(), rb.fc3.weight.data_ptr(), rb.fc1.bias.data_ptr(), rb.fc3.bias.data_ptr()
            F.Fin, F.cmap, F.dmap, F.bps_src0 = blk.Fin, blk.cmap.data_ptr(), blk.dmap.data_ptr(), blk.bps_src0
            F.in_gamma, F.in_beta = net.in_bn.weight.data_ptr(), net.in_bn.bias.data_ptr()
            F.Wb, F.Wxt, F.ldxt, F.bias13 = blk.Wb.data_ptr(), blk.Wxt.data_ptr(), blk.ldxt, blk.bias13.data_ptr()
            if grads:
                F.dWb, F.dWxt, F.db13 = blk.dWb.data_ptr(), blk.dWxt.data_ptr(), net.db13[i].data_ptr()
                F.gW1, F.gW3 = rb.fc1.weight.grad.data_ptr(), rb.fc3.weight.grad.data_ptr()
                F.gb1, F.gb3 = rb.fc1.bias.grad.data_ptr(), rb.fc3.bias.grad.data_ptr()
                F.g_in_gamma, F.g_in_beta = net.in_bn.weight.grad.data_ptr(), net.in_bn.bias.grad.data_ptr()
        n = 0
        if not grads:
            for blk in net.blocks:  # fc2 weights: plain bf16 copies
                P = J.plain[n]
                P.src, P.lds, P.rows, P.cols, P.dst, P.ldd = blk.rb.fc2.weight.data_ptr(), INNER, HID, INNER, blk.W2.data_ptr(), INNER
                n += 1
        if net.kind == "G":
            G = net.module
            r = 0
            for lin, rows in ((G.gen_joint_conf, self.nJ), (G.gen_rot, 6), (G.gen_transl, 3)):
                P = J.plain[n]
                if grads:
                    P.gsrc, P.ldgs, P.rows, P.cols, P.gdst, P.ldgd = self.dWh[r:].data_ptr(), HID, rows, HID, lin.weight.grad.data_ptr(), HID
                else:
                    P.src, P.lds, P.rows, P.cols, P.dst, P.ldd = lin.weight.data_ptr(), HID, rows, HID, self.Wh[r:].data_ptr(), HID
                n += 1
                r += rows
        J.n_plain = n
        return J

    def prep(self, net):
        _native.check(_L().dexgg_prep_weights(ctypes.byref(self._weight_jobs(net, False)), _st(self.dev)))
        if net.kind == "G":
            G = net.module
            torch.cat([G.gen_joint_conf.bias.detach(), G.gen_rot.bias.detach(), G.gen_transl.bias.detach()], out=self.bh[:self.nJ + 9])

    def unfold(self, net):
        _native.check(_L().dexgg_unfold_grads(ctypes.byref(self._weight_jobs(net, True)), _st(self.dev)))
        if net.kind == "G":
            G = net.module
            nJ = self.nJ
            G.gen_joint_conf.bias.grad.copy_(self.dbh[:nJ])
            G.gen_rot.bias.grad.copy_(self.dbh[nJ:nJ + 6])
            G.gen_transl.bias.grad.copy_(self.dbh[nJ + 6:nJ + 9])

    # ------------------------------------------------------------------------------------------------ forward / backward of a block stack
    def _block_in(self, net, b):
        return (net.tail1, TAILW, net.n_tail) if b == 0 else (net.XT[b], LDXT, HID + net.n_tail)

    def forward_blocks(self, net, R, groups, xhat, drops=None):
        rpg = R // groups
        nb = len(net.blocks)
        for b, blk in enumerate(net.blocks):
            rb = blk.rb
            xin, ldx, K2 = self._block_in(net, b)
            self._gemm(R, NCAT, [(xhat.data_ptr(), BPS, 0, blk.Wb.data_ptr(), BPS, 0, BPS), (xin.data_ptr(), ldx, 0, blk.Wxt.data_ptr(), blk.ldxt, 0, K2)],
                       bias=blk.bias13, act_from=INNER, C16=net.T[b], ldc=NCAT)
            self._bn_fwd(net.T[b].data_ptr(), NCAT, INNER, rpg, groups, rb.bn1, net.stat1[b], net.H[b].data_ptr(), INNER)
            self._gemm(R, HID, [(net.H[b].data_ptr(), INNER, 0, blk.W2.data_ptr(), INNER, 0, INNER)], bias=rb.fc2.bias, C16=net.Y[b], ldc=HID)
            out, ldo = (net.XT[b + 1], LDXT) if b + 1 < nb else (net.Xlast, HID)
            self._bn_fwd(net.Y[b].data_ptr(), HID, HID, rpg, groups, rb.bn2, net.stat2[b], out.data_ptr(), ldo,
                         res=net.T[b][:, INNER:].data_ptr(), ldr=NCAT, drop=(drops[b] if drops else None))

    def running_update(self, net, groups, rpg):
        """BatchNorm buffer updates of one pass: hidden layers from the saved statistics, the input BatchNorm's BPS columns from
        the step's shared statistics (its tail columns were updated by dexgg_tail_pack); num_batches_tracked += groups."""
        ents = (RunEntry * 24)()
        n = 0

        def put(bn, mean, second, C, is_var, same_row, nbt=True):
            nonlocal n
            e = ents[n]
            e.running_mean, e.running_var, e.mean, e.second = bn.running_mean.data_ptr(), bn.running_var.data_ptr(), mean, second
            e.num_batches_tracked = bn.num_batches_tracked.data_ptr() if nbt else None
            e.C, e.n_groups, e.second_is_var, e.same_row = C, groups, int(is_var), int(same_row)
            e.count, e.eps, e.momentum = float(rpg * self.world), EPS, float(bn.momentum or 0.1)
            n += 1
        put(net.in_bn, self.bps_stat[0].data_ptr(), self.bps_stat[1].data_ptr(), BPS, True, True)
        for b, blk in enumerate(net.blocks):
            put(blk.rb.bn1, net.stat1[b][0].data_ptr(), net.stat1[b][1].data_ptr(), INNER, False, False)
            put(blk.rb.bn2, net.stat2[b][0].data_ptr(), net.stat2[b][1].data_ptr(), HID, False, False)
        _native.check(_L().dexgg_bn_running_multi(ents, n, _st(self.dev)))

    def backward_blocks(self, net, R, groups, xhat, last, want_wgrad, want_tail_grad, drops=None):
        """last: ('rank1', dl, w, dw_out) or ('mat', tensor, ld) - the gradient wrt the last block's output."""
        rpg = R // groups
        nb = len(net.blocks)
        if want_wgrad:
            net.small_scratch.zero_()  # db13 of every block + the dWxt scratch of every block (split contractions accumulate)
        g_src = last
        # The weight-gradient GEMMs (a third of the step's tensor-core time) are off the critical path: nothing in the backward
        # chain reads them before `unfold`.  They go to the network's side stream behind an event, the chain continues with the
        # data-gradient GEMM / the next slab kernel, and joins the side stream at the end.
        cur = torch.cuda.current_stream(self.dev)
        ws = None
        if want_wgrad and self.wgrad_streams:
            if net.wstream is None:
                net.wstream = torch.cuda.Stream(self.dev)
            ws = net.wstream

        def off_chain(fn):
            if ws is None:
                fn()
                return
            ws.wait_stream(cur)
            with torch.cuda.stream(ws):
                fn()

        for b in range(nb - 1, -1, -1):
            blk, rb = net.blocks[b], net.blocks[b].rb
            GY, G13 = net.GY[b], net.G13[b]
            out, ldo = (net.XT[b + 1], LDXT) if b + 1 < nb else (net.Xlast, HID)
            kw = dict(g=g_src[1].data_ptr(), ldg=g_src[2]) if g_src[0] == "mat" else dict(rank1=(g_src[1], g_src[2], g_src[3]))
            self._bn_bwd(HID, rpg, groups, rb.bn2, net.stat2[b], net.Y[b].data_ptr(), HID, out.data_ptr(), ldo, GY.data_ptr(), HID,
                         skip=net.T[b][:, INNER:].data_ptr(), lds=NCAT, gskip=G13[:, INNER:].data_ptr(), ldgs=NCAT,
                         dbias_skip=net.db13[b][INNER:].data_ptr(), want_wgrad=want_wgrad, drop=(drops[b] if drops else None), **kw)
            if want_wgrad:  # dW2 = GY^T . H  (both operands MN-major, contraction over the rows split over CTAs)
                # [512 x 256] output, contraction over the rows: 64-column tiles x 8-way split = 128 CTAs, 8 atomic adds per element
                off_chain(lambda: self._gemm(HID, INNER, [(GY.data_ptr(), HID, 1, net.H[b].data_ptr(), INNER, 1, R)], C32=rb.fc2.weight.grad,
                                             ldc=INNER, split_k=max(1, min(8, R // 512)), tile_n=64))
            self._gemm(R, INNER, [(GY.data_ptr(), HID, 0, blk.W2.data_ptr(), INNER, 1, HID)], C16=net.GH, ldc=INNER)
            self._bn_bwd(INNER, rpg, groups, rb.bn1, net.stat1[b], net.T[b].data_ptr(), NCAT, net.H[b].data_ptr(), INNER, G13.data_ptr(), NCAT,
                         g=net.GH.data_ptr(), ldg=INNER, want_wgrad=want_wgrad)
            xin, ldx, K2 = self._block_in(net, b)
            if want_wgrad:  # gradients wrt the folded first-layer weights: [768, 4096] and [768, X | tail]
                def first_layer_wgrads(G13=G13, blk=blk, xin=xin, ldx=ldx):
                    self._gemm(NCAT, BPS, [(G13.data_ptr(), NCAT, 1, xhat.data_ptr(), BPS, 1, R)], C32=blk.dWb, ldc=BPS)
                    self._gemm(NCAT, blk.ldxt, [(G13.data_ptr(), NCAT, 1, xin.data_ptr(), ldx, 1, R)], C32=blk.dWxt, ldc=blk.ldxt,
                               split_k=max(1, min(4 if blk.has_x else 16, R // 512)))
                off_chain(first_layer_wgrads)
            if b > 0 or want_tail_grad:
                self._gemm(R, blk.ldxt, [(G13.data_ptr(), NCAT, 0, blk.Wxt.data_ptr(), blk.ldxt, 1, NCAT)], C16=net.gXT[b], ldc=blk.ldxt)
                g_src = ("mat", net.gXT[b], blk.ldxt)
        if ws is not None:
            cur.wait_stream(ws)

    def _score_head(self, net, R, groups, targets=(0.0, 0.0), dscale=(0.0, 0.0), lscale=(0.0, 0.0), labels=None, loss_slot=0, want_db=True):
        a = ScoreHead()
        a.X, a.ldx = net.Xlast.data_ptr(), HID
        a.w, a.b = net.head.weight.data_ptr(), net.head.bias.data_ptr()
        a.rows_per_group, a.n_groups = R // groups, groups
        a.labels = labels.data_ptr() if labels is not None else None
        for i in range(2):
            a.target[i], a.dscale[i], a.lscale[i] = targets[i], dscale[i], lscale[i]
        a.p, a.dl, a.loss = net.p.data_ptr(), net.dl.data_ptr(), self.loss[loss_slot:].data_ptr()
        a.db = net.head.bias.grad.data_ptr() if want_db else None
        a.wass = 0
        _native.check(_L().dexgg_score_head(ctypes.byref(a), _st(self.dev)))

    # ------------------------------------------------------------------------------------------------ the steps
    def _input_norm(self, bps):
        B = self.B
        assert bps.shape == (B, BPS) and bps.dtype == torch.float32 and bps.is_contiguous() and bps.is_cuda
        s = self.bps_stat
        slot = self._next_slot(2)  # [2, 4096] floats = two slots
        _native.check(_L().dexgg_bn_input_stats(bps.data_ptr(), BPS, B, BPS, EPS, s[0].data_ptr(), s[1].data_ptr(), s[2].data_ptr(),
                                                self._slot_ptr(slot), _st(self.dev)))
        if self.world > 1:  # per-rank (mean, M2 = B * var) -> all_gather -> Chan merge
            local = torch.stack([s[0], s[1] * float(B)]).view(1, 2, BPS)
            self._gather_merge(local, 1, BPS, B, s[0].data_ptr(), s[2].data_ptr(), s[1].data_ptr())
        _native.check(_L().dexgg_xhat(bps.data_ptr(), BPS, s[0].data_ptr(), s[2].data_ptr(), B, BPS, self.xhat.data_ptr(), 2, _st(self.dev)))

    def _dev_batch(self, data, keys):
        out = {}
        for k in keys:
            v = data[k]
            v = v if torch.is_tensor(v) else torch.as_tensor(v)
            v = v.to(self.dev, non_blocking=True)
            if k == "bps_object" or v.dtype not in (torch.float32, torch.float64):
                v = v.float()
            out[k] = v.contiguous()
        return out

    def _gan_d_step(self, d, Zgen):
        """Generator forward (dexgangrasp.py:772-776) and the discriminator step up to its gradients (:788-793)."""
        B, dev = self.B, self.dev
        G, D = self.G, self.D
        real_tail = [d["rot_matrix"].reshape(B, 9), d["transl"], d["joint_conf"]]
        self.prep(G)
        self._tail_pack(G, [Zgen.float().contiguous()], 0, B, 0)
        self.forward_blocks(G, B, 1, self.xhat)
        self._gemm(B, self.nJ + 9, [(G.Xlast.data_ptr(), HID, 0, self.Wh.data_ptr(), HID, 0, HID)], bias=self.bh, C32=self.OUT, ldc=32)
        _native.check(_L().dexgg_gen_pose(self.OUT.data_ptr(), 32, self.nJ, B, self.fake_rot.data_ptr(), self.fake_transl.data_ptr(),
                                          self.fake_joint.data_ptr(), _st(dev)))
        self.running_update(G, 1, B)
        fake_tail = [self.fake_rot, self.fake_transl, self.fake_joint]
        # real | fake as one double batch with per-half statistics
        self.prep(D)
        self._tail_pack(D, real_tail, 0, B, 0)
        self._tail_pack(D, fake_tail, B, B, 1)
        self.forward_blocks(D, 2 * B, 2, self.xhat)
        D.flat_grad.zero_()
        bw = self.bw
        self._score_head(D, 2 * B, 2, targets=(1.0, 0.0), dscale=(0.5 * bw / B, 0.5 * bw / B), lscale=(bw / B, bw / B), loss_slot=0)
        self.running_update(D, 2, B)
        self.backward_blocks(D, 2 * B, 2, self.xhat, ("rank1", D.dl, D.head.weight, D.head.weight.grad), True, False)
        self.unfold(D)

    def _gan_g_step(self, d):
        """D(fake) with the UPDATED discriminator, gradients back into the generator, up to the generator's gradients (:796-815)."""
        m, B, dev = self.model, self.B, self.dev
        G, D = self.G, self.D
        real_tail = [d["rot_matrix"].reshape(B, 9), d["transl"], d["joint_conf"]]
        fake_tail = [self.fake_rot, self.fake_transl, self.fake_joint]
        bw = self.bw
        self.prep(D)
        self._tail_pack(D, fake_tail, 0, B, 0)
        self.forward_blocks(D, B, 1, self.xhat)
        self._score_head(D, B, 1, targets=(1.0, 0.0), dscale=(bw / B, 0.0), lscale=(bw / B, 0.0), loss_slot=2, want_db=False)
        self.running_update(D, 1, B)
        self.backward_blocks(D, B, 1, self.xhat, ("rank1", D.dl, D.head.weight, None), False, True)
        src = self._tail_src(fake_tail)
        st = D.tail_stat[0]
        targs = (ctypes.byref(src), B, float(B * self.world), st[0].data_ptr(), st[1].data_ptr(), D.gXT[0].data_ptr(), TAILW, 0,
                 D.gXT[1].data_ptr(), LDXT, HID, self.gPose.data_ptr())
        if self.world > 1:
            slot = self._next_slot()
            _native.check(_L().dexgg_tail_bn_bwd(*targs, 1, self._slot_ptr(slot), _st(dev)))
            self._allreduce(self._slot_view(slot, 2 * D.n_tail), "bn_bwd")
            _native.check(_L().dexgg_tail_bn_bwd(*targs, 2, self._slot_ptr(slot), _st(dev)))
        else:
            _native.check(_L().dexgg_tail_bn_bwd(*targs, 0, None, _st(dev)))
        G.flat_grad.zero_()
        self.dWh.zero_()
        self.dbh.zero_()
        rsrc = self._tail_src(real_tail)
        _native.check(_L().dexgg_gen_loss(self.OUT.data_ptr(), 32, self.nJ, B, float(B), ctypes.byref(rsrc), self.gPose.data_ptr(),
                                          float(m.transl_coef), float(m.rot_coef), float(m.conf_coef), self.gHeads.data_ptr(), 32,
                                          self.loss[3:].data_ptr(), self.dbh.data_ptr(), _st(dev)))
        self._gemm(32, HID, [(self.gHeads.data_ptr(), 32, 1, G.Xlast.data_ptr(), HID, 1, B)], C32=self.dWh, ldc=HID, split_k=max(1, min(32, B // 256)))
        self._gemm(B, HID, [(self.gHeads.data_ptr(), 32, 0, self.Wh.data_ptr(), HID, 1, 32)], C16=self.gX, ldc=HID)
        self.backward_blocks(G, B, 1, self.xhat, ("mat", self.gX, HID), True, False)
        self.unfold(G)

    def _gan_losses(self, is_train_gen):
        m = self.model
        if self.world > 1:  # report the global-batch means
            self._allreduce(self.loss, "grad")
            self.loss.div_(self.world)
        vals = self.mix @ self.loss
        out = {"total_loss_disc": vals[0], "bce_loss_real": vals[1], "bce_loss_fake": vals[2]}
        if is_train_gen:
            m.last_loss_dict_gen = {"gen_loss_fake": vals[3], "transl_loss": vals[4], "rot_loss": vals[5], "conf_loss": vals[6],
                                    "total_loss_gen": vals[7]}
        out.update(m.last_loss_dict_gen)
        return out, vals

    def gan_step(self, real_data, is_train_gen, Zgen):
        m, dev = self.model, self.dev
        G, D = self.G, self.D
        d = self._dev_batch(real_data, ("bps_object", "rot_matrix", "transl", "joint_conf"))
        self.loss.zero_()
        self._begin_step()
        G.attach()
        D.attach()
        with torch.cuda.device(dev), torch.no_grad():
            self._input_norm(d["bps_object"])
            self._gan_d_step(d, Zgen)
            self.sync_grads(D)
            D.optimizer.step()
            if is_train_gen:
                self._gan_g_step(d)
                self.sync_grads(G)
                G.optimizer.step()
            out, _ = self._gan_losses(is_train_gen)
        self._bump_versions(self.G, self.D)
        return out

    def _eva_grads(self, d):
        """Evaluator forward / backward up to its gradients (dexgangrasp.py:739-752)."""
        m, B = self.model, self.B
        E = self.E
        masks = getattr(m.DexEvaluator, "dropout_masks", None)
        p = float(m.DexEvaluator.dout.p)
        drops = [((masks[i] if masks is not None else None), (0.0 if masks is not None else p), i) for i in range(3)]
        if masks is None and p <= 0.0:
            drops = None
        _native.check(_L().dexgg_bump_counter(self.seed.data_ptr(), _st(self.dev)))
        self.prep(E)
        self._tail_pack(E, [d["rot_matrix"].reshape(B, 9), d["transl"]], 0, B, 0)
        self.forward_blocks(E, B, 1, self.xhat, drops)
        E.flat_grad.zero_()
        bw = self.bw
        lab = m.DexEvaluator.gt_label.reshape(-1).contiguous()
        self._score_head(E, B, 1, labels=lab, dscale=(bw / B, 0.0), lscale=(bw / B, 0.0), loss_slot=6)
        self.running_update(E, 1, B)
        self.backward_blocks(E, B, 1, self.xhat, ("rank1", E.dl, E.head.weight, E.head.weight.grad), True, False, drops)
        self.unfold(E)

    def evaluator_step(self, data):
        m, dev = self.model, self.dev
        E = self.E
        d = self._dev_batch(data, ("bps_object", "rot_matrix", "transl", "label"))
        m.DexEvaluator.gt_label = d["label"].float().unsqueeze(-1)
        self.loss[6:7].zero_()
        self._begin_step()
        E.attach()
        with torch.cuda.device(dev), torch.no_grad():
            self._input_norm(d["bps_object"])
            self._eva_grads(d)
            self.sync_grads(E)
            E.optimizer.step()
            val = self.loss[6].clone()
            if self.world > 1:
                self._allreduce(val, "grad")
                val.div_(self.world)
        self._bump_versions(self.E)
        return {"total_loss_eva": val, "bce_loss": val}

    def combined_step(self, data, Zgen, is_train_gen=True):
        """update_dexgangrasp + update_dexevaluator on the SAME batch (what train.py does per iteration when both networks train,
        :249-276, and what C5 times) as ONE schedule of TWO concurrent chains: the updates touch disjoint parameters, so after the
        shared input BatchNorm the evaluator's whole update (forward, backward, gradient exchange, Adam) runs on a second stream
        beside the GAN update.  Neither chain fills the GPU alone - a latency-bound slab / small-GEMM kernel of one chain leaves the
        tensor cores to a big GEMM of the other, HBM-bound Adam / weight preparation of one runs under the GEMMs of the other, and on
        N ranks each chain's gradient exchange crosses NVLink under the other's compute.  The slab kernels spin on an in-kernel
        exchange, so while two chains are in flight their grids are capped at one CTA per SM: any two of them always fit the GPU
        together (2 x 148 CTAs x 32 K registers), and everything else that may hold an SM (GEMMs, NCCL) ends without waiting on
        them.  Results are those of the two calls made one after the other."""
        m, dev = self.model, self.dev
        G, D, E = self.G, self.D, self.E
        d = self._dev_batch(data, ("bps_object", "rot_matrix", "transl", "joint_conf", "label"))
        m.DexEvaluator.gt_label = d["label"].float().unsqueeze(-1)
        self.loss.zero_()
        self._begin_step()
        for n in (G, D, E):
            n.attach()
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_side_stream", None) is None:
            self._side_stream = torch.cuda.Stream(dev)
        side = self._side_stream
        # Two spinning kernels (one per chain) must always fit the GPU together: true for the capped slab grids and the narrow
        # gradient-exchange kernel, NOT for an NCCL kernel next to a slab grid - with NCCL gradients the chains run one after the other.
        two_chains = self.two_chains and (self.world == 1 or self.grad_p2p)
        with torch.cuda.device(dev), torch.no_grad():
            self._input_norm(d["bps_object"])
            if not two_chains:
                self._eva_grads(d)
                self.sync_grads(E)
                E.optimizer.step()
            self.slab_ctas = 1 if two_chains else 0
            try:
                if two_chains:
                    side.wait_stream(main)
                    with torch.cuda.stream(side):   # chain 2: the evaluator update
                        self._eva_grads(d)
                        self.sync_grads(E)
                        E.optimizer.step()
                self._gan_d_step(d, Zgen)      # chain 1: the GAN update
                self.sync_grads(D)
                D.optimizer.step()
                if is_train_gen:
                    self._gan_g_step(d)
                    self.sync_grads(G)
                    G.optimizer.step()
                if two_chains:
                    main.wait_stream(side)
            finally:
                self.slab_ctas = 0
            out, vals = self._gan_losses(is_train_gen)
            out["total_loss_eva"] = out["bce_loss"] = vals[8]
        self._bump_versions(G, D, E)
        return out

    def _bump_versions(self, *nets):
        """Buffers were written through raw pointers: tell torch (the packed eval-mode weights key on the version counters)."""
        ts = []
        for n in nets:
            ts += [b for b in n.module.buffers() if b.is_floating_point()]
        torch.autograd.graph.increment_version(ts)

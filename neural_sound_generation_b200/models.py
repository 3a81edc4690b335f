"""Drop-in for the VQ-VAE part of /root/reference/src/models.py.

``VQEmbedding``, ``ResBlock``, ``VQVAE``, ``weights_init`` and ``to_scalar`` keep
the reference's names, constructor signatures, attribute names and
``state_dict`` keys / shapes (checkpoints interchange both ways), but
``forward`` never calls the torch layers: the ``nn.Conv2d`` / ``nn.BatchNorm2d``
objects are parameter containers only and the arithmetic is the explicit
kernel schedule of ``engine.Engine``.

Returned activations have the reference's logical NCHW shape over NHWC
(channels-last) storage.
"""
import weakref

import os

import torch
from torch import nn
from torch.autograd import Function

from . import ops
from .engine import Engine, RES_BLOCKS
from .engine_tma import EngineTMA
from .vector_quantization import vq, vq_st


def to_scalar(arr):
    """models.py:18-22"""
    if type(arr) == list:
        return [x.item() for x in arr]
    else:
        return arr.item()


def weights_init(m):
    """models.py:25-32: Xavier-uniform weights and zero bias for every module whose class name contains 'Conv'."""
    classname = m.__class__.__name__
    if classname.find('Conv') != -1:
        try:
            nn.init.xavier_uniform_(m.weight.data)
            m.bias.data.fill_(0)
        except AttributeError:
            print("Skipping initialization of ", classname)


def _nhwc(t):
    """(B,C,H,W) logical -> contiguous (B,H,W,C) storage (no copy when already channels-last)."""
    return t.permute(0, 2, 3, 1).contiguous()


def _nchw_view(t):
    return t.permute(0, 3, 1, 2)


class VQEmbedding(nn.Module):
    """models.py:121-142"""

    def __init__(self, z_dim, dim):
        super(VQEmbedding, self).__init__()
        self.embedding = nn.Embedding(z_dim, dim)
        self.embedding.weight.data.uniform_(-1. / z_dim, 1. / z_dim)

    def forward(self, z_e_x):
        z_e_x_ = _nhwc(z_e_x)
        latents = vq(z_e_x_, self.embedding.weight)
        return latents

    def straight_through(self, z_e_x):
        z_e_x_ = _nhwc(z_e_x)
        z_q_x_, indices = vq_st(z_e_x_, self.embedding.weight.detach())
        z_q_x = _nchw_view(z_q_x_)
        z_q_x_bar_ = _GatherRows.apply(self.embedding.weight, indices).view_as(z_e_x_)
        z_q_x_bar = _nchw_view(z_q_x_bar_)
        return z_q_x, z_q_x_bar


def ema_codebook_update(embedding, z_e_x, indices, ema_count, ema_sum, decay=0.99, eps=1e-5, dp=None):
    """EMA codebook training (extension; the reference trains the codebook by Adam through loss_vq).

    ``embedding``: the ``VQEmbedding``; ``z_e_x`` (B,D,H,W) encoder output; ``indices`` flat int64 from ``vq_st``;
    ``ema_count`` (K,) / ``ema_sum`` (K,D) fp32 state, updated in place together with ``embedding.embedding.weight``.
    One scatter-add launch gathers counts[K] + sums[K,D]; under data parallelism (``dp``) they are all-reduce-summed
    before the update so every rank applies the global statistics."""
    w = embedding.embedding.weight
    k, d = w.shape
    z = _nhwc(z_e_x.detach())
    stats = torch.zeros(k * d + k, dtype=torch.float32, device=w.device)
    sums, counts = stats[:k * d].view(k, d), stats[k * d:]
    ops.scatter_add(z, indices.reshape(-1), sums, alpha=1.0, counts=counts)
    if dp is not None:
        dp.reduce_stats(stats)
    with torch.no_grad():
        ops.ema_update(counts, sums, ema_count, ema_sum, w.data, decay, eps)
    return counts, sums


class _GatherRows(Function):
    """torch.index_select(weight, 0, indices) with the index_add backward (models.py:137-138)."""

    @staticmethod
    def forward(ctx, weight, indices):
        ctx.save_for_backward(indices)
        ctx.shape = weight.shape
        out = torch.empty((indices.numel(), weight.shape[1]), dtype=torch.float32, device=weight.device)
        ops.gather_rows(weight.detach().float().contiguous(), indices, out)
        return out

    @staticmethod
    def backward(ctx, grad):
        (indices,) = ctx.saved_tensors
        table = torch.zeros(ctx.shape, dtype=torch.float32, device=grad.device)
        ops.scatter_add(grad.contiguous().float(), indices, table, alpha=1.0)
        return table, None


class ResBlock(nn.Module):
    """models.py:145-158.  ``forward`` = relu(x) + block(relu(x)) (the leading ReLU is in-place)."""

    def __init__(self, dim):
        super(ResBlock, self).__init__()
        self.block = nn.Sequential(
            nn.ReLU(True),
            nn.Conv2d(dim, dim, 3, 1, 1),
            nn.BatchNorm2d(dim),
            nn.ReLU(True),
            nn.Conv2d(dim, dim, 1),
            nn.BatchNorm2d(dim)
        )

    def forward(self, x):
        names = ['block.1.weight', 'block.1.bias', 'block.2.weight', 'block.2.bias',
                 'block.4.weight', 'block.4.bias', 'block.5.weight', 'block.5.bias']
        sd = dict(self.named_parameters())
        return _ResBlockFn.apply(self, x, *[sd[n] for n in names])


class _ResBlockFn(Function):
    NAME = 'decoder.0'   # any ResBlock slot of the engine's BN tables

    @staticmethod
    def forward(ctx, module, x, *params):
        names = ['block.1.weight', 'block.1.bias', 'block.2.weight', 'block.2.bias',
                 'block.4.weight', 'block.4.bias', 'block.5.weight', 'block.5.bias']
        dim = params[0].shape[0]
        eng = Engine(dim, dim, 1, act_dtype=torch.float32)
        pre = _ResBlockFn.NAME + '.'
        P = {pre + n: p.detach() for n, p in zip(names, params)}
        Bufs = {pre + n: b for n, b in module.named_buffers()}
        xin = _nhwc(x.detach().float())
        ectx = eng.begin(xin, module.training)
        y = eng.resblock_fwd(ectx, P, Bufs, _ResBlockFn.NAME, xin)
        ctx.eng, ctx.ectx, ctx.P, ctx.names = eng, ectx, P, names
        # value semantics of the reference: y = relu(x) + block(relu(x)).  The reference also
        # overwrites the caller's x with relu(x) as a side effect of ReLU(True); that mutation is not reproduced.
        return _nchw_view(y)

    @staticmethod
    def backward(ctx, g):
        pre = _ResBlockFn.NAME + '.'
        G = {k: torch.zeros_like(v) for k, v in ctx.P.items()}
        g_x = ctx.eng.resblock_bwd(ctx.ectx, ctx.P, G, _ResBlockFn.NAME, _nhwc(g.float()))
        return (None, _nchw_view(g_x)) + tuple(G[pre + n] for n in ctx.names)


PARAM_ORDER = None


class VQVAE(nn.Module):
    """models.py:161-216"""

    def __init__(self, input_dim, dim, z_dim=512):
        super(VQVAE, self).__init__()
        self.encoder = nn.Sequential(
            nn.Conv2d(input_dim, dim, 4, 2, 1),
            nn.BatchNorm2d(dim),
            nn.ReLU(True),
            nn.Conv2d(dim, dim, 4, 2, 1),
            ResBlock(dim),
            ResBlock(dim),
        )

        self.codebook = VQEmbedding(z_dim, dim)

        self.decoder = nn.Sequential(
            ResBlock(dim),
            ResBlock(dim),
            nn.ReLU(True),
            nn.ConvTranspose2d(dim, dim, 4, 2, 1),
            nn.BatchNorm2d(dim),
            nn.ReLU(True),
            nn.ConvTranspose2d(dim, input_dim, 4, 2, 1),
            nn.Tanh()
        )

        self.apply(weights_init)
        self.input_dim, self.dim, self.z_dim = input_dim, dim, z_dim
        # execution options (not part of the reference API).  None = automatic: the production engine -- bf16
        # operand-ready activations, TMA-fed tcgen05 convolutions (conv_impl 2), tcgen05 split-bf16 search (vq_mode 1)
        # -- wherever the shape is served (dim == 64; search: D in {64, 128, 256}, K <= 8192), the general fp32 kernels otherwise.
        # ``set_precision('fp32')`` pins the fp32 parity engine; explicit values always win.
        self._act_dtype = None              # torch.float32 / torch.bfloat16
        self._conv_impl = None              # 0 = CUDA-core fp32, 1 = register-staged tcgen05, 2 = TMA-fed tcgen05 (bf16)
        self._vq_mode = None                # 0 = exact fp32 CUDA-core search, 1 = tcgen05 split-bf16 search
        self._flat = None
        for p in self.parameters():         # optim.Adam(model.parameters()) finds the model that owns the flat buffers
            p._nsg_owner = weakref.ref(self)

    # ------------------------------------------------------------------ execution options
    def _auto(self):
        """(act_dtype, conv_impl, vq_mode) of the automatic choice for this model's shape."""
        if self.dim == 64:
            vq = 1 if ops.vq_mode_serves(self.z_dim, self.dim, 1) else 0
            return torch.bfloat16, 2, vq
        return torch.float32, 0, (1 if ops.vq_mode_serves(self.z_dim, self.dim, 1) else 0)

    @property
    def conv_impl(self):
        if self._conv_impl is not None:
            return self._conv_impl
        if self._act_dtype is not None:     # an explicit activation type picks the best engine that computes in it
            return 2 if (self._act_dtype == torch.bfloat16 and self.dim == 64) else 0
        return self._auto()[1]

    @conv_impl.setter
    def conv_impl(self, v):
        self._conv_impl = v

    @property
    def act_dtype(self):
        if self._act_dtype is not None:
            return self._act_dtype
        if self._conv_impl is not None:
            return torch.bfloat16 if self._conv_impl == 2 else torch.float32
        return self._auto()[0]

    @act_dtype.setter
    def act_dtype(self, v):
        self._act_dtype = v

    @property
    def vq_mode(self):
        if self._vq_mode is not None:
            return self._vq_mode
        if self._conv_impl == 0 and self._act_dtype in (None, torch.float32):
            return 0                        # the pinned fp32 parity engine keeps the exact CUDA-core search
        return self._auto()[2]

    @vq_mode.setter
    def vq_mode(self, v):
        self._vq_mode = v

    def set_precision(self, mode):
        """'auto' (default: the production engine where served), 'bf16' (TMA engine; raises later if dim != 64) or
        'fp32' (CUDA-core fp32 parity engine, exact search)."""
        if mode == 'auto':
            self._act_dtype = self._conv_impl = self._vq_mode = None
        elif mode == 'bf16':
            self._act_dtype, self._conv_impl, self._vq_mode = torch.bfloat16, 2, None
        elif mode == 'fp32':
            self._act_dtype, self._conv_impl, self._vq_mode = torch.float32, 0, 0
        else:
            raise ValueError("set_precision: 'auto', 'bf16' or 'fp32' (got %r)" % (mode,))
        return self

    # ------------------------------------------------------------------ plumbing
    def engine(self):
        if self.conv_impl == 2:     # bf16 TMA-fed tcgen05 schedule (caches its packed weight images on the model)
            if self.act_dtype != torch.bfloat16:
                raise ops.A.NsgError('conv_impl = 2 (TMA engine) needs act_dtype = torch.bfloat16')
            eng = self.__dict__.get('_eng_tma')
            if eng is None or eng.vq_mode != self.vq_mode:
                eng = EngineTMA(self.input_dim, self.dim, self.z_dim, self.vq_mode)
                self.__dict__['_eng_tma'] = eng
            eng.dp = self.__dict__.get('dp')        # dist.DataParallel (None = single process)
            return eng
        if self.__dict__.get('dp') is not None:
            raise ops.A.NsgError('data parallelism is implemented for the bf16 TMA engine (conv_impl = 2)')
        return Engine(self.input_dim, self.dim, self.z_dim, self.act_dtype, self.conv_impl, self.vq_mode)

    def _named(self):
        names = [n for n, _ in self.named_parameters()]
        params = [p for _, p in self.named_parameters()]
        return names, params

    def _bufs(self):
        return dict(self.named_buffers())

    def flatten_parameters(self):
        """Re-home all parameters (and gradients) as 256-byte aligned views of two flat fp32 buffers so one
        fused Adam launch updates the whole model.  state_dict keys and shapes are unchanged."""
        names, params = self._named()
        dev = params[0].device
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += (p.numel() + 63) // 64 * 64
        if (self._flat is not None and self._flat['params'].device == dev and self._flat['total'] == total
                and all(p.data_ptr() == self._flat['params'].data_ptr() + 4 * o for p, o in zip(params, offs))):
            return self._flat
        flat = torch.zeros(total, dtype=torch.float32, device=dev)
        gflat = torch.zeros(total, dtype=torch.float32, device=dev)
        for p, o in zip(params, offs):
            flat[o:o + p.numel()].copy_(p.data.reshape(-1))
            p.data = flat[o:o + p.numel()].view(p.shape)
        self._flat = dict(params=flat, grads=gflat, offs=offs, total=total, names=names, last_numel=params[-1].numel())
        self.bind_grads()
        return self._flat

    def bind_grads(self):
        """Point every .grad at its slice of the flat gradient buffer (after optimizer.zero_grad(set_to_none))."""
        f = self._flat
        _, params = self._named()
        dropped = False
        for p, o in zip(params, f['offs']):
            if p.grad is None or p.grad.data_ptr() != f['grads'].data_ptr() + 4 * o:
                dropped = dropped or p.grad is None
                p.grad = f['grads'][o:o + p.numel()].view(p.shape)
        if dropped:     # zero_grad(set_to_none=True) semantics: a dropped gradient restarts from zero
            f['grads'].zero_()

    def grad_dict(self):
        f = self._flat
        names, params = self._named()
        return {n: f['grads'][o:o + p.numel()].view(p.shape) for n, p, o in zip(names, params, f['offs'])}

    # ------------------------------------------------------------------ reference API
    def encode(self, x):
        """models.py:188-191"""
        eng = self.engine()
        P = {n: p.detach() for n, p in self.named_parameters()}
        xin = _nhwc(x.detach().float())
        ectx = eng.begin(xin, self.training)
        z_e = eng.encoder_fwd(ectx, P, self._bufs(), xin.to(self.act_dtype))
        idx, _ = eng.quantize(z_e, P['codebook.embedding.weight'], want_codes=False)
        return idx.view(z_e.shape[:-1])

    def decode(self, latents):
        """models.py:193-196"""
        eng = self.engine()
        P = {n: p.detach() for n, p in self.named_parameters()}
        cb = P['codebook.embedding.weight']
        z_q = torch.empty(tuple(latents.shape) + (self.dim,), dtype=self.act_dtype, device=cb.device)
        ops.gather_rows(cb, latents.reshape(-1).contiguous(), z_q)
        ectx = eng.begin(z_q, self.training)
        xt = eng.decoder_fwd(ectx, P, self._bufs(), z_q)
        return _nchw_view(xt.float() if xt.dtype != torch.float32 else xt)

    def forward(self, x):
        """models.py:198-216 -> (x_tilde, z_e_x, z_q_x)"""
        names, params = self._named()
        if torch.is_grad_enabled():
            self.__dict__['_grads_dirty'] = True        # a backward() may follow (GraphedTrainStep clears before replay)
        return _VQVAEFn.apply(self, x, *params)

    # ------------------------------------------------------------------ fused train step (train.py:116-135)
    def loss_and_grads(self, c, beta=1.0, grads=None):
        """One fused forward + loss + backward on mel batch ``c`` (B, n_mels, T) (or (B,1,n_mels,T)).

        Accumulates into the (flat) gradient buffers and returns a float64 device tensor
        [loss_recons, loss_vq, loss_commit] plus the tuple (x_tilde, z_e_x, z_q_x) the reference's forward
        returns.  No host synchronisation.
        """
        if self.input_dim != 1:
            raise ops.A.NsgError('loss_and_grads: the fused ljspeech step needs input_dim == 1')
        if c.dim() == 3:
            c = c.unsqueeze(1)                                              # train.py:115
        eng = self.engine()
        self.flatten_parameters()
        self.bind_grads()
        self.__dict__['_grads_dirty'] = True
        P = {n: p.detach() for n, p in self.named_parameters()}
        G = grads if grads is not None else self.grad_dict()
        Bufs = self._bufs()
        b, _, h, w = c.shape
        c3 = c.detach().float().contiguous().view(b, h, w)
        x = c3.view(b, h, w, 1)
        if self.conv_impl == 2:
            ectx = eng.begin(x, self.training, fused=True)
            sums = ectx.sums                                                # cleared with the step's other sums
        else:
            ectx = eng.begin(x, self.training)
            sums = torch.zeros(2, dtype=torch.float64, device=x.device)    # [recon sq sum, vq sq sum]
        z_e = eng.encoder_fwd(ectx, P, Bufs, x if self.act_dtype == torch.float32 else x.to(self.act_dtype))
        cb = P['codebook.embedding.weight']
        idx, z_q = eng.quantize(z_e, cb, want_codes=True, sqerr=sums[1:2])
        # d loss_vq / d e (the codebook's index_add_) needs only z_q / idx / z_e: it runs on the engine's side stream
        # beside the decoder FORWARD (joined with the weight gradients in finish_backward) instead of in front of the
        # encoder backward: 0.592 -> 0.577 ms per step (NSG_SCATTER_SIDE=1: beside the encoder backward, 0.588; 0: inline)
        scatter_mode = os.environ.get('NSG_SCATTER_SIDE', '2') if hasattr(eng, '_side') else '0'
        cb_grad = lambda: ops.scatter_add(z_q, idx, G['codebook.embedding.weight'], alpha=2.0 / z_e.numel(), sub=z_e)
        if scatter_mode == '2':
            eng._side(ectx, cb_grad, z_q, idx, z_e)
        xt = eng.decoder_fwd(ectx, P, Bufs, z_q)
        wt = xt.shape[2]
        if xt.dtype == torch.bfloat16:
            # loss + d loss / d pre-tanh (bf16, what decoder_bwd reads) + the fp32 x_tilde handed back, one kernel
            xt32 = torch.empty(xt.shape, dtype=torch.float32, device=x.device)
            dpre = torch.empty((b, h, wt), dtype=torch.bfloat16, device=x.device)
            ops.recon_loss_bf16(xt.view(b, h, wt), c3, sums[0:1], dpre, xt32.view(b, h, wt),
                                dpre_sum=G['decoder.6.bias'] if getattr(eng, 'c1', None) is not None else None)
            ectx.dbias6_done = getattr(eng, 'c1', None) is not None
        else:
            xt32 = xt if xt.dtype == torch.float32 else xt.float()
            dpre = torch.empty((b, h, wt), dtype=torch.float32, device=x.device)
            ops.recon_loss(xt32.view(b, h, wt), c3, sums[0:1], dpre)
        g_st = eng.decoder_bwd(ectx, P, G, dpre.view(b, h, wt, 1))
        dp = getattr(eng, 'dp', None)
        if dp is not None:
            # decoder bucket: unpack + all-reduce beside the encoder backward; the loss scalars ride in its padding
            l_now = ops.loss_finalize(sums, 1.0 / (b * h * w), 1.0 / z_e.numel())
            dp.decoder_done(eng, ectx, G, self._flat, losses=l_now)
        n_lat = z_e.numel()
        dz = eng.bottleneck_bwd(ectx, g_st, z_e, z_q, 2.0 * beta / n_lat)  # ST clone + beta * d commit / d z_e
        if scatter_mode == '1':
            eng._side(ectx, cb_grad, z_q, idx, z_e)
        elif scatter_mode != '2':
            cb_grad()
        eng.encoder_bwd(ectx, P, G, dz)
        if dp is not None:
            dp.finish(eng, ectx, G, self._flat)
        else:
            eng.finish_backward(ectx, G)
        losses = ops.loss_finalize(sums, 1.0 / (b * h * w), 1.0 / z_e.numel())
        self.__dict__['_last_ctx'] = ectx       # saved activations of the last step (tests read the ReLU patterns)
        return losses, (_nchw_view(xt32), _nchw_view(z_e), _nchw_view(z_q), idx)

    def relu_patterns(self):
        """{ReLU name: bool (B,C,H,W)} -- the activation pattern of the last ``loss_and_grads`` on the TMA engine, in
        the reference module naming ('encoder.1', '<block>.in', '<block>.block.3', 'decoder.2',
        'decoder.5').  Debug / test aid: lets a reference evaluate the same piecewise-linear branch."""
        ctx = self.__dict__.get('_last_ctx')
        if ctx is None or not hasattr(ctx, 'v0'):
            raise ops.A.NsgError('relu_patterns: no saved TMA-engine step')
        pat = {'encoder.1': ctx.v0, 'decoder.2': ctx.rd, 'decoder.5': ctx.v4}
        for blk in RES_BLOCKS:
            r, _, v1, _ = ctx.saved[blk]
            pat[blk + '.in'], pat[blk + '.block.3'] = r, v1
        return {k: _nchw_view(v > 0) for k, v in pat.items()}


class _VQVAEFn(Function):
    """Autograd boundary of the whole model: forward = Engine schedule, backward = its explicit reverse."""

    @staticmethod
    def forward(ctx, module, x, *params):
        names = [n for n, _ in module.named_parameters()]
        eng = module.engine()
        P = {n: p.detach() for n, p in zip(names, params)}
        Bufs = module._bufs()
        xin = _nhwc(x.detach().float())
        ectx = eng.begin(xin, module.training)
        z_e = eng.encoder_fwd(ectx, P, Bufs, xin if module.act_dtype == torch.float32 else xin.to(module.act_dtype))
        idx, z_q = eng.quantize(z_e, P['codebook.embedding.weight'], want_codes=True)
        xt = eng.decoder_fwd(ectx, P, Bufs, z_q)
        xt32 = xt if xt.dtype == torch.float32 else xt.float()
        ctx.eng, ctx.ectx, ctx.P, ctx.names = eng, ectx, P, names
        ctx.idx, ctx.xt, ctx.need_dx = idx, xt32, x.requires_grad
        ctx.z_shape = z_e.shape
        ze_out = z_e if z_e.dtype == torch.float32 else z_e.float()
        zq_out = z_q if z_q.dtype == torch.float32 else z_q.float()
        return _nchw_view(xt32), _nchw_view(ze_out), _nchw_view(zq_out)

    @staticmethod
    def backward(ctx, g_xt, g_ze, g_zq):
        eng, ectx, P = ctx.eng, ctx.ectx, ctx.P
        G = {k: torch.zeros_like(v) for k, v in P.items()}
        act = eng.act_dtype
        dev = ctx.xt.device
        if g_xt is None:
            g_xt = torch.zeros_like(_nchw_view(ctx.xt))
        dpre = torch.empty_like(ctx.xt)
        ops.tanh_bwd(_nhwc(g_xt.float()), ctx.xt, dpre)
        g_st = eng.decoder_bwd(ectx, P, G, dpre)
        dz = eng.bottleneck_bwd_sum(ectx, g_st, _nhwc(g_ze).to(act) if g_ze is not None else None)
        if g_zq is not None:
            ops.scatter_add(_nhwc(g_zq.float()), ctx.idx, G['codebook.embedding.weight'], alpha=1.0)
        g_x = eng.encoder_bwd(ectx, P, G, dz, need_dx=ctx.need_dx)
        eng.finish_backward(ectx, G)
        gx = _nchw_view(g_x.float()) if g_x is not None else None
        return (None, gx) + tuple(G[n] for n in ctx.names)

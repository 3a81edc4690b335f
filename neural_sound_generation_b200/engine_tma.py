"""bf16 train / inference schedule of the VQ-VAE over the TMA-fed tcgen05 kernels (``csrc/conv_tma.cu``).

Same role and method names as ``engine.Engine`` (the host-side mirror of autograd over
/root/reference/src/models.py:145-216), different data contract: every activation tensor that feeds a
convolution is stored OPERAND-READY in bf16 NHWC -- BatchNorm + ReLU already applied (``bn_act64``), the
ResBlock's in-place leading ReLU already applied by whoever produced the tensor -- so the tensor core reads
what the TMA unit wrote, with no register staging.  Per layer the schedule is

    conv (tcgen05, epilogue: bias + BN moment sums)  ->  bn_act64 (finalise statistics, normalise, ReLU, residual)

and in reverse

    dgrad (tcgen05, epilogue: + skip gradient, ReLU mask, BN-backward sums)  ->  bn_bwd64  ->  wgrad (tcgen05).

Weight images (bf16, UMMA K-major) of all ten 64->64 layers are re-packed by ONE launch per step; their
gradients accumulate in packed fp32 images that ONE launch folds into the reference-layout ``.grad`` tensors.
The single-channel ends (encoder.0, decoder.6) run on the tensor cores too (``c1t_*`` kernels: one
tcgen05.mma per 128 pixels does the 64-channel expansion / reduction); other input_dim use the CUDA-core kernels.
"""
import os

import torch

from . import ops
from .ops import MASK_POS, STATS_MOMENTS, STATS_BNBWD
from .engine import BN_EPS, BN_MOMENTUM, BN_LAYERS, BN_INDEX, RES_BLOCKS, Ctx, _out_hw

BF = torch.bfloat16

# (state_dict prefix, ksize, stride, pad, transposed) of the 64 -> 64 layers
TC_LAYERS = ([('encoder.3', 4, 2, 1, False)]
             + [(b + s, k, 1, p, False) for b in RES_BLOCKS for s, k, p in (('.block.1', 3, 1), ('.block.4', 1, 0))]
             + [('decoder.3', 4, 2, 1, True)])


class _Layer(object):
    __slots__ = ('name', 'k', 's', 'p', 'tr', 'wf', 'wd', 'dwp', 'dwa', 'nslices', 'slice_cache', 'wci_f', 'wco_f', 'wci_d', 'wco_d', 'wci_g', 'wco_g')


class EngineTMA(object):
    """Schedule object; caches the packed weight / gradient images per model (keyed by the flat parameter buffer)."""

    def __init__(self, input_dim, dim, z_dim, vq_mode=1):
        if dim != 64:
            raise ops.A.NsgError('the TMA engine serves dim == 64 (got %d)' % dim)
        self.cin0, self.dim, self.k = input_dim, dim, z_dim
        self.act_dtype = BF
        self.vq_mode = vq_mode
        self.layers = None
        self._key = None
        self.side = None            # second stream: weight gradients run beside the data-gradient chain
        self.overlap_wgrad = os.environ.get('NSG_OVERLAP_WGRAD', '1') != '0'
        # Weight-gradient partials: False = CTAs meet in one packed image per layer through vector atomics; True = one
        # slice per CTA, plain stores, summed in a fixed order by the unpack (bit-reproducible gradients).  Measured on
        # B200 (configs[1], same box): the sliced kernels are 1-2 us faster each in isolation, but their 58 MB of
        # partials push activations out of L2 and the step is slower, 0.824 vs 0.807 ms -- atomics stay the default.
        self.wgrad_sliced = os.environ.get('NSG_WGRAD_SLICED', '0') == '1'
        self.zbuf = None            # begin(fused=True): the step's accumulators, one allocation, one fill per step
        self.dwa_in_zbuf = False
        self.dwa_dirty = False      # the images hold a consumed fused step's sums (cleared by the next fused begin)
        self.dp = None              # dist.DataParallel: cross-rank BatchNorm statistics / bucketed gradient all-reduce
        # BatchNorm fused into the producing convolution across a grid barrier (csrc/conv_tma.cu, BnFuse): one launch
        # instead of conv + bn_act / dgrad + bn_bwd.  MEASURED on B200 (profiles/r2_fuse_bn_ab.txt, configs[1]): the
        # fused step is SLOWER, 0.816 vs 0.667 ms -- after the barrier every CTA re-reads its 2-3 accumulators and
        # stores them with nothing left to overlap (26.7 us against 16.5 + 7.6 us for the 3x3 forward), while the
        # separate streaming pass spreads the same stores over the whole machine under programmatic dependent launch.
        # So training keeps the two-kernel form by default; NSG_FUSE_BN=1 selects the fused launches (they remain the
        # place where synchronised BatchNorm exchanges its statistics between GPUs without a collective, see dist.py).
        # Eval mode needs no barrier -- the running statistics fold into the convolution's output affine -- and is
        # always fused (config 5: 0.111 ms against 0.123 ms per utterance).
        self.fuse_bn = os.environ.get('NSG_FUSE_BN', '0') == '1'
        self.fuse_bn_eval = os.environ.get('NSG_FUSE_BN_EVAL', '1') != '0'
        # Synchronised BatchNorm over peer memory: the exchange is a one-CTA kernel in front of bn_act / bn_bwd by default;
        # NSG_PEER_IN_KERNEL=1 runs it INSIDE them (CTA 0 under the other CTAs' tensor loads, a gate for the rest: no
        # extra launch).  MEASURED (profiles/r2_scale_*): the separate kernel wins -- 0.870 against 0.924 ms at 8 GPUs,
        # 0.82 against 0.86 ms at 2 -- because the gate holds ALL CTAs of the consumer for the NVLink round trip, while
        # the separate kernel's round trip overlaps the consumer's launch under programmatic dependent launch.
        self.peer_in_kernel = os.environ.get('NSG_PEER_IN_KERNEL', '0') == '1'
        # The bias of a convolution that feeds a BatchNorm has an analytically ZERO gradient (the normalisation removes any
        # per-channel constant); the reference computes rounding noise there (~1e-9 of the weight gradients), which Adam
        # then normalises into +-lr steps of a parameter that cannot influence the output.  bn_bwd can form that noise
        # (column sums of dh: one shared-memory reduction per CTA) or leave the exact zero: skipping is the default
        # (MEASURED 0.652 against 0.668 ms per step, profiles/r2_toggles_ab.txt); NSG_PREBN_DBIAS=1 computes it.
        self.prebn_dbias = os.environ.get('NSG_PREBN_DBIAS', '0') == '1'

    # ------------------------------------------------------------------ packed images
    def _prepare(self, P):
        key = tuple(P[n + '.weight'].data_ptr() for n, *_ in TC_LAYERS)
        if self._key == key:
            return
        dev = P['encoder.3.weight'].device
        D = self.dim
        ntaps_total = sum(k * k for _, k, *_ in TC_LAYERS)
        self.wimg = torch.empty(2 * ntaps_total * 8192 + 3 * 8192, dtype=torch.uint8, device=dev)
        self.layers = {}
        self.pack_items, self.unpack_items = [], []
        # wgrad_sliced = False: one packed image per layer, CTAs meet in it through vector atomics, cleared per step
        if self.zbuf is not None and self.zbuf.device == dev:      # fused step: cleared with the step's other sums
            self.dwa = self.zbuf[self.zbuf.numel() - ntaps_total * D * D * 4:].view(torch.float32)
            self.dwa.zero_()
            self.dwa_in_zbuf = True
        else:
            self.dwa = torch.zeros(ntaps_total * D * D, dtype=torch.float32, device=dev)
            self.dwa_in_zbuf = False
        self.dwa_split = 0          # first element of the decoder layers' images (they are laid out last)
        ow, og = 0, 0
        for name, k, s, p, tr in TC_LAYERS:
            if name.startswith('decoder.') and self.dwa_split == 0:
                self.dwa_split = og
            L = _Layer()
            L.name, L.k, L.s, L.p, L.tr = name, k, s, p, tr
            w = P[name + '.weight']
            kk = k * k
            if tr:      # ConvTranspose2d weight (Cin, Cout, k, k)
                L.wci_f, L.wco_f = D * kk, kk           # forward contracts Cin
                L.wci_d, L.wco_d = kk, D * kk           # data gradient contracts Cout
                L.wci_g, L.wco_g = kk, D * kk           # wgrad: a = dy channel (Cout), b = x channel (Cin)
            else:       # Conv2d weight (Cout, Cin, k, k)
                L.wci_f, L.wco_f = kk, D * kk
                L.wci_d, L.wco_d = D * kk, kk
                L.wci_g, L.wco_g = kk, D * kk           # a = x channel (Cin), b = dy channel (Cout)
            L.wf = self.wimg[ow:ow + kk * 8192]
            L.wd = self.wimg[ow + kk * 8192:ow + 2 * kk * 8192]
            ow += 2 * kk * 8192
            # weight-gradient partials [nslices, k*k, D, D]: one slice per CTA of the wgrad kernel (plain stores, no
            # atomics, nothing to clear), sized at the first backward pass of a geometry and summed by unpack_part
            L.dwp, L.nslices, L.slice_cache = None, 0, {}
            L.dwa = self.dwa[og:og + kk * D * D]
            og += kk * D * D
            self.pack_items += [(w, L.wf, kk, L.wci_f, L.wco_f), (w, L.wd, kk, L.wci_d, L.wco_d)]
            self.layers[name] = L
        # single-channel ends on the tensor cores (input_dim == 1): 8 KB operand images
        self.c1 = None
        if self.cin0 == 1:
            self.c1 = dict(enc0=self.wimg[ow:ow + 8192], dec6_fwd=self.wimg[ow + 8192:ow + 16384],
                           dec6_dgrad=self.wimg[ow + 16384:ow + 24576])
            self.pack_items += [(P['encoder.0.weight'], self.c1['enc0'], -1, 1, 16),
                                (P['decoder.6.weight'], self.c1['dec6_fwd'], -2, 16, 1),
                                (P['decoder.6.weight'], self.c1['dec6_dgrad'], -1, 1, 16)]
        self._key = key

    def _unpack_items(self, G, part=None):
        """part: None = every layer, 'decoder' / 'encoder' = that half (the data-parallel gradient buckets)."""
        return [(L.dwp, G[L.name + '.weight'], L.k * L.k, L.wci_g, L.wco_g, L.nslices) for L in self.layers.values()
                if L.nslices > 0 and (part is None or L.name.startswith(part + '.'))]

    def _wgrad_slices(self, L, g, key, dev):
        if not self.wgrad_sliced:
            L.dwp, L.nslices = L.dwa, 1
            return L.dwa
        ns = L.slice_cache.get(key)
        if ns is None:
            ns = L.slice_cache[key] = ops.conv_wgrad_slices(g)
        need = ns * L.k * L.k * self.dim * self.dim
        if L.dwp is None or L.dwp.numel() < need:
            L.dwp = torch.empty(need, dtype=torch.float32, device=dev)
        L.nslices = ns
        return L.dwp

    # ------------------------------------------------------------------ helpers
    def _bn(self, ctx, name):
        i = BN_INDEX[name]
        co = ctx.coef[i]
        return dict(coef=(co[0], co[1], co[2], co[3]), mean=co[0], invstd=co[1], fstats=ctx.stats[i, 0],
                    bstats=ctx.stats[i, 1])

    def begin(self, x_nhwc, training, fused=False):
        ctx = Ctx()
        ctx.training = training
        ctx.saved = {}
        dev = x_nhwc.device
        ctx.coef = torch.empty((len(BN_LAYERS), 4, self.dim), dtype=torch.float32, device=dev)
        nst = len(BN_LAYERS) * 2 * 2 * self.dim
        NBAR = 64                   # grid-barrier counters of the BatchNorm-fused launches (two each, zero at launch)
        ctx.nbar = 0
        if self.dp is not None:
            self.dp.begin_step()
        if fused:
            # everything the step accumulates into lives in ONE buffer cleared by ONE fill at the start of the step:
            # BatchNorm statistic sums (fp64), the two squared-error sums (fp64), the packed weight-gradient images
            if self.zbuf is None or self.zbuf.device != dev:
                nd = sum(k * k for _, k, *_ in TC_LAYERS) * self.dim * self.dim
                self.zbuf = torch.zeros((nst + 2) * 8 + NBAR * 4 + nd * 4, dtype=torch.uint8, device=dev)
                self._key = None                    # the layers' dwa views are re-made by _prepare
            else:
                self.zbuf.zero_()
            self.dwa_dirty = False
            f64 = self.zbuf[:(nst + 2) * 8].view(torch.float64)
            ctx.stats = f64[:nst].view(len(BN_LAYERS), 2, 2 * self.dim)
            ctx.sums = f64[nst:nst + 2]
            ctx.bars = self.zbuf[(nst + 2) * 8:(nst + 2) * 8 + NBAR * 4].view(torch.int32)
        else:
            ctx.stats = torch.zeros((len(BN_LAYERS), 2, 2 * self.dim), dtype=torch.float64, device=dev)
            ctx.sums = None
            ctx.bars = torch.zeros(NBAR, dtype=torch.int32, device=dev)
        ctx.fused = fused
        ctx.packed = False
        ctx.keep = []               # tensors in use on the side stream (kept alive until the join)
        ctx.forked = False
        return ctx

    def _side(self, ctx, fn, *tensors):
        """Run fn() on the side stream after everything issued so far on the current stream."""
        if not self.overlap_wgrad:
            return fn()
        if self.side is None:
            self.side = torch.cuda.Stream()
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            fn()
        ctx.keep.extend(tensors)
        ctx.forked = True

    def _pack(self, ctx, P):
        if not ctx.packed:
            self._prepare(P)
            ops.conv_pack_many(self.pack_items)
            ctx.packed = True

    def _fwd_geom(self, L, x):
        b, h, w, c = x.shape
        oh, ow = _out_hw(h, w, L.k, L.s, L.p, L.tr)
        return ops.conv_geom(b, h, w, c, oh, ow, c, L.k, L.s, L.p, 1 if L.tr else 0, L.wci_f, L.wco_f, ops.BF16), oh, ow

    def _conv_fwd(self, ctx, P, name, x, stats=None, relu_out=False):
        L = self.layers[name]
        g, oh, ow = self._fwd_geom(L, x)
        out = torch.empty((x.shape[0], oh, ow, self.dim), dtype=BF, device=x.device)
        ops.conv_gemm_tma(g, x, L.wf, out, bias=P[name + '.bias'],
                          stats_mode=STATS_MOMENTS if stats is not None else 0, stats=stats, relu_out=relu_out)
        return out

    def _sync(self):
        """True when BatchNorm statistics span the ranks (dist.DataParallel(sync_bn=True))."""
        return self.dp is not None and self.dp.sync_bn and self.dp.world > 1

    def _fusable(self, g):
        # synchronised BatchNorm fuses only when the statistics can be exchanged inside the kernel (peer memory)
        return (self.fuse_bn and (not self._sync() or self.dp.peer_areas is not None) and ops.conv_bn_tma_fits(g))

    def _bar(self, ctx):
        b = ctx.bars[ctx.nbar:ctx.nbar + 2]
        ctx.nbar += 2
        return b

    def _conv_bn_fwd(self, ctx, P, Bufs, name, bn_name, x, res=None, relu_out=True):
        """conv `name` -> BatchNorm `bn_name` [-> + res] [-> ReLU].  Returns (h, v): the pre-normalisation tensor the
        backward pass reads (None in eval mode when fused) and the operand-ready result.  One launch when the fused
        kernel serves the shape, else convolution + bn_act."""
        L = self.layers[name]
        g, oh, ow = self._fwd_geom(L, x)
        bn = self._bn(ctx, bn_name)
        shape = (x.shape[0], oh, ow, self.dim)
        if not ctx.training and self.fuse_bn_eval:
            v = torch.empty(shape, dtype=BF, device=x.device)
            ops.conv_bn_tma(g, x, L.wf, v, ops.BNF_EVAL, P[bn_name + '.weight'], beta=P[bn_name + '.bias'], eps=BN_EPS,
                            running_mean=Bufs[bn_name + '.running_mean'], running_var=Bufs[bn_name + '.running_var'],
                            bias=P[name + '.bias'], add_src=res, relu_out=relu_out)
            return None, v
        if ctx.training and self._fusable(g):
            h = torch.empty(shape, dtype=BF, device=x.device)
            v = torch.empty(shape, dtype=BF, device=x.device)
            sync = self._sync()
            ops.conv_bn_tma(g, x, L.wf, h, ops.BNF_TRAIN, P[bn_name + '.weight'], beta=P[bn_name + '.bias'], eps=BN_EPS,
                            momentum=BN_MOMENTUM, count=h.numel() // self.dim * (self.dp.world if sync else 1),
                            peer=self.dp.next_peer() if sync else None,
                            running_mean=Bufs[bn_name + '.running_mean'], running_var=Bufs[bn_name + '.running_var'],
                            nbt=Bufs[bn_name + '.num_batches_tracked'], coef=ctx.coef[BN_INDEX[bn_name]], res=res,
                            barrier=self._bar(ctx), out2=v, bn_relu=relu_out, bias=P[name + '.bias'], stats=bn['fstats'])
            return h, v
        h = self._conv_fwd(ctx, P, name, x, stats=bn['fstats'] if ctx.training else None)
        return h, self._bn_act(ctx, P, Bufs, bn_name, h, res=res, relu_out=relu_out)

    def _conv_dgrad_bn(self, ctx, P, G, name, dy, in_hw, bn_name, h, want_g, **epi):
        """Data gradient of conv `name` (epilogue ``epi``: + skip gradient, ReLU mask) flowing into the backward of
        BatchNorm `bn_name`, whose pre-normalisation tensor is ``h``.  Returns (g, dh): g = the gradient wrt the
        BatchNorm output (None unless ``want_g``), dh = the gradient wrt ``h``."""
        L = self.layers[name]
        b, oh, ow, c = dy.shape
        ih, iw = in_hw
        g = ops.conv_geom(b, oh, ow, c, ih, iw, c, L.k, L.s, L.p, 0 if L.tr else 1, L.wci_d, L.wco_d, ops.BF16)
        bn = self._bn(ctx, bn_name)
        if self._fusable(g):
            dh = torch.empty((b, ih, iw, c), dtype=BF, device=dy.device)
            gy = torch.empty_like(dh) if want_g else None
            sync = self._sync()
            lead = not sync or self.dp.rank == 0    # the exchanged sums ARE the global dgamma / dbeta: one rank adds them
            ops.conv_bn_tma(g, dy, L.wd, gy, ops.BNF_BWD, P[bn_name + '.weight'],
                            count=dh.numel() // self.dim * (self.dp.world if sync else 1),
                            peer=self.dp.next_peer() if sync else None, barrier=self._bar(ctx),
                            dgamma=G[bn_name + '.weight'] if lead else None,
                            dbeta=G[bn_name + '.bias'] if lead else None, out2=dh,
                            stats=bn['bstats'], xhat_src=h, mean=bn['mean'], invstd=bn['invstd'], **epi)
            return gy, dh
        gy = self._conv_dgrad(name, dy, in_hw, stats_mode=STATS_BNBWD, stats=bn['bstats'], xhat_src=h, mean=bn['mean'],
                              invstd=bn['invstd'], **epi)
        conv_before = {'.block.2': '.block.1', '.block.5': '.block.4'}
        pre = bn_name[:-8] + conv_before[bn_name[-8:]] + '.bias' if bn_name[-8:] in conv_before else None
        if bn_name == 'encoder.1':
            pre = 'encoder.0.bias'
        return gy, self._bn_bwd(ctx, P, G, bn_name, gy, h, G[pre] if pre else None)

    def _conv_dgrad(self, name, dy, in_hw, **epi):
        L = self.layers[name]
        b, oh, ow, c = dy.shape
        ih, iw = in_hw
        g = ops.conv_geom(b, oh, ow, c, ih, iw, c, L.k, L.s, L.p, 0 if L.tr else 1, L.wci_d, L.wco_d, ops.BF16)
        out = torch.empty((b, ih, iw, c), dtype=BF, device=dy.device)
        ops.conv_gemm_tma(g, dy, L.wd, out, **epi)
        return out

    def _conv_wgrad(self, ctx, name, x, dy, q_colsum=None):
        """x = the layer's (operand-ready) forward input, dy = gradient of its output.  Issued on the side stream."""
        L = self.layers[name]
        b, h, w, c = x.shape
        _, oh, ow, _ = dy.shape
        if not ctx.fused and self.dwa_dirty and not self.wgrad_sliced:
            self.dwa.zero_()        # generic autograd backward after a fused step: start from clean images
            self.dwa_dirty = False
        if L.tr:        # P = dy gathered over x's pixels, Q = x
            g = ops.conv_geom(b, oh, ow, c, h, w, c, L.k, L.s, L.p, 0, L.wci_g, L.wco_g, ops.BF16)
            dwp = self._wgrad_slices(L, g, (b, h, w, oh, ow), x.device)
            self._side(ctx, lambda: ops.conv_wgrad_tma(g, dy, x, dwp, q_colsum=q_colsum, sliced=self.wgrad_sliced), x, dy)
        else:           # P = x gathered over dy's pixels, Q = dy
            g = ops.conv_geom(b, h, w, c, oh, ow, c, L.k, L.s, L.p, 0, L.wci_g, L.wco_g, ops.BF16)
            dwp = self._wgrad_slices(L, g, (b, h, w, oh, ow), x.device)
            self._side(ctx, lambda: ops.conv_wgrad_tma(g, x, dy, dwp, q_colsum=q_colsum, sliced=self.wgrad_sliced), x, dy)

    def _stat_exchange(self, ctx, stats):
        """Synchronised BatchNorm: (peer, gate) for a kernel that exchanges ``stats`` itself over peer memory, or
        (None, None) after an ncclAllReduce of them (no peer access) / when nothing is to be exchanged."""
        if not self._sync():
            return None, None
        if self.dp.peer_areas is not None and self.peer_in_kernel:
            return self.dp.next_peer(), self._bar(ctx)[0:1]
        self.dp.reduce_stats(stats)         # one-CTA peer kernel, or ncclAllReduce
        return None, None

    def _bn_act(self, ctx, P, Bufs, name, h, res=None, relu_out=True):
        bn = self._bn(ctx, name)
        out = torch.empty_like(h)
        n = h.numel() // self.dim
        if ctx.training:
            # statistics of the GLOBAL batch, as the single-process reference sees them
            peer, gate = self._stat_exchange(ctx, bn['fstats'])
            if self._sync():
                n *= self.dp.world
            ops.bn_act64(h, out, P[name + '.weight'], P[name + '.bias'], stats=bn['fstats'], count=n, res=res,
                         relu_out=relu_out, eps=BN_EPS, momentum=BN_MOMENTUM,
                         running_mean=Bufs[name + '.running_mean'], running_var=Bufs[name + '.running_var'],
                         nbt=Bufs[name + '.num_batches_tracked'], coef=bn['coef'], peer=peer, gate=gate)
        else:
            ops.bn_act64(h, out, P[name + '.weight'], P[name + '.bias'], stats=None, res=res, relu_out=relu_out,
                         eps=BN_EPS, running_mean=Bufs[name + '.running_mean'],
                         running_var=Bufs[name + '.running_var'], coef=bn['coef'])
        return out

    def _bn_bwd(self, ctx, P, G, name, g, h, dbias_prev):
        bn = self._bn(ctx, name)
        dh = torch.empty_like(h)
        n = h.numel() // self.dim
        dgamma, dbeta = G[name + '.weight'], G[name + '.bias']
        peer, gate = self._stat_exchange(ctx, bn['bstats'])
        if self._sync():
            # the reduced sums ARE the global dgamma / dbeta: only rank 0 contributes them to the gradient all-reduce
            n *= self.dp.world
            if self.dp.rank != 0:
                dgamma = dbeta = None
        ops.bn_bwd64(g, h, dh, bn['bstats'], n, bn['mean'], bn['invstd'], P[name + '.weight'],
                     dgamma=dgamma, dbeta=dbeta, dbias_prev=dbias_prev if self.prebn_dbias else None, peer=peer, gate=gate)
        return dh

    # ------------------------------------------------------------------ ResBlock (models.py:145-158)
    def resblock_fwd(self, ctx, P, Bufs, name, r, relu_out):
        """r = relu(block input) (operand-ready).  Returns relu(y) or y, with y = r + BN5(Conv1x1(relu(BN2(Conv3x3(r)))))."""
        h1, v1 = self._conv_bn_fwd(ctx, P, Bufs, name + '.block.1', name + '.block.2', r, relu_out=True)
        h2, y = self._conv_bn_fwd(ctx, P, Bufs, name + '.block.4', name + '.block.5', v1, res=r, relu_out=relu_out)
        ctx.saved[name] = (r, h1, v1, h2)
        return y

    def resblock_bwd(self, ctx, P, G, name, g_y, g_h2=None, prev_block=None):
        """g_y: gradient wrt the block output y.  g_h2: gradient wrt the block's second convolution output (BN5
        backward already applied by the fused producer of g_y), or None: g_y then carries this block's BN5 backward
        sums (accumulated by its producer) and the separate pass runs here.  Returns (g, g_h2_prev): the gradient wrt
        the block's pre-ReLU input and, when ``prev_block`` names the ResBlock that produced that input and the fused
        kernel serves the shape, the gradient wrt ITS second convolution output (else None, with that block's BN5
        backward sums accumulated)."""
        r, h1, v1, h2 = ctx.saved[name]
        hw = r.shape[1:3]
        if g_h2 is None:
            g_h2 = self._bn_bwd(ctx, P, G, name + '.block.5', g_y, h2, G[name + '.block.4.bias'])
        self._conv_wgrad(ctx, name + '.block.4', v1, g_h2)
        _, g_h1 = self._conv_dgrad_bn(ctx, P, G, name + '.block.4', g_h2, hw, name + '.block.2', h1, False,
                                      mask_mode=MASK_POS, mask_src=v1)
        self._conv_wgrad(ctx, name + '.block.1', r, g_h1)
        if prev_block is not None:
            return self._conv_dgrad_bn(ctx, P, G, name + '.block.1', g_h1, hw, prev_block + '.block.5',
                                       ctx.saved[prev_block][3], True, add_src=g_y, mask_mode=MASK_POS, mask_src=r)
        return self._conv_dgrad(name + '.block.1', g_h1, hw, add_src=g_y, mask_mode=MASK_POS, mask_src=r), None

    # ------------------------------------------------------------------ encoder / decoder (models.py:164-184)
    def encoder_fwd(self, ctx, P, Bufs, x):
        """x (B,H,W,Cin0) bf16 -> z_e (B,H/4,W/4,64) bf16 (raw: the quantiser's input)."""
        self._pack(ctx, P)
        tr = ctx.training
        bn1 = self._bn(ctx, 'encoder.1')
        b, h, w, cin = x.shape
        w0 = P['encoder.0.weight']
        oh, ow = _out_hw(h, w, 4, 2, 1, False)
        a0 = torch.empty((b, oh, ow, self.dim), dtype=BF, device=x.device)
        g = ops.conv_geom(b, h, w, cin, oh, ow, self.dim, 4, 2, 1, 0, 16, cin * 16, ops.BF16)
        if self.c1 is not None and not tr and self.fuse_bn_eval:
            # eval: encoder.1's running statistics fold into the convolution's output affine, one launch
            v0 = ops.c1_conv_bn_eval_tc(g, x, self.c1['enc0'], a0, P['encoder.1.weight'], P['encoder.1.bias'],
                                        Bufs['encoder.1.running_mean'], Bufs['encoder.1.running_var'], eps=BN_EPS,
                                        bias=P['encoder.0.bias'], relu_out=True)
            a0 = None
        else:
            if self.c1 is not None:
                ops.c1_conv_tc(g, x, self.c1['enc0'], a0, bias=P['encoder.0.bias'],
                               stats_mode=STATS_MOMENTS if tr else 0, stats=bn1['fstats'] if tr else None)
            else:
                ops.conv_gemm(g, x, w0, a0, bias=P['encoder.0.bias'], stats_mode=STATS_MOMENTS if tr else 0,
                              stats=bn1['fstats'] if tr else None)
            v0 = self._bn_act(ctx, P, Bufs, 'encoder.1', a0, relu_out=True)
        r4 = self._conv_fwd(ctx, P, 'encoder.3', v0, relu_out=True)            # relu(a1): ResBlock's in-place ReLU
        r5 = self.resblock_fwd(ctx, P, Bufs, 'encoder.4', r4, relu_out=True)
        z_e = self.resblock_fwd(ctx, P, Bufs, 'encoder.5', r5, relu_out=False)
        ctx.x, ctx.a0, ctx.v0 = x, a0, v0
        return z_e

    def decoder_fwd(self, ctx, P, Bufs, z_q):
        """z_q (B,h,w,64) bf16 -> x_tilde (B,4h,4w,Cin0) (tanh applied)."""
        self._pack(ctx, P)
        tr = ctx.training
        rq = torch.empty_like(z_q)
        ops.relu64(z_q, rq)
        r1 = self.resblock_fwd(ctx, P, Bufs, 'decoder.0', rq, relu_out=True)
        rd = self.resblock_fwd(ctx, P, Bufs, 'decoder.1', r1, relu_out=True)   # + decoder.2 ReLU
        u0, v4 = self._conv_bn_fwd(ctx, P, Bufs, 'decoder.3', 'decoder.4', rd, relu_out=True)
        b, h, w, _ = v4.shape
        w6 = P['decoder.6.weight']
        cout = w6.shape[1]
        oh, ow = _out_hw(h, w, 4, 2, 1, True)
        xt = torch.empty((b, oh, ow, cout), dtype=BF, device=v4.device)
        g = ops.conv_geom(b, h, w, self.dim, oh, ow, cout, 4, 2, 1, 1, cout * 16, 16, ops.BF16)
        if self.c1 is not None:
            ops.c1_convT_tc(g, v4, self.c1['dec6_fwd'], xt, bias=P['decoder.6.bias'], tanh_out=True)
        else:
            ops.conv_gemm(g, v4, w6, xt, bias=P['decoder.6.bias'], tanh_out=True)
        ctx.rd, ctx.u0, ctx.v4 = rd, u0, v4
        return xt

    def decoder_bwd(self, ctx, P, G, dpre):
        """dpre (B,H',W',Cin0): gradient wrt the pre-tanh output.  Returns the gradient wrt z_q."""
        rd, u0, v4 = ctx.rd, ctx.u0, ctx.v4
        bn4 = self._bn(ctx, 'decoder.4')
        if dpre.dtype != BF:
            dpre = dpre.to(BF)
        b, oh, ow, cout = dpre.shape
        _, h, w, _ = v4.shape
        w6 = P['decoder.6.weight']
        # decoder.6 weight gradient: P = dpre gathered over v4's pixels, Q = v4
        g = ops.conv_geom(b, oh, ow, cout, h, w, self.dim, 4, 2, 1, 0, 16, cout * 16, ops.BF16)
        gw = g
        if self.c1 is not None:
            if getattr(ctx, 'dbias6_done', False):      # the fused loss kernel already summed dpre into the bias gradient
                self._side(ctx, lambda: ops.c1_wgrad_tc(gw, dpre, v4, G['decoder.6.weight']), dpre, v4)
            else:
                self._side(ctx, lambda: (ops.c1_wgrad_tc(gw, dpre, v4, G['decoder.6.weight']),
                                         ops.colsum(dpre, G['decoder.6.bias'])), dpre, v4)
        else:
            self._side(ctx, lambda: (ops.conv_wgrad(gw, dpre, v4, G['decoder.6.weight']),
                                     ops.colsum(dpre, G['decoder.6.bias'])), dpre, v4)
        # decoder.6 data gradient, masked by the ReLU, with decoder.4's BN-backward sums
        g = ops.conv_geom(b, oh, ow, cout, h, w, self.dim, 4, 2, 1, 0, 16, cout * 16, ops.BF16)
        g_v4 = torch.empty_like(v4)
        if self.c1 is not None:
            ops.c1_conv_tc(g, dpre, self.c1['dec6_dgrad'], g_v4, mask_mode=MASK_POS, mask_src=v4,
                           stats_mode=STATS_BNBWD, stats=bn4['bstats'], xhat_src=u0, mean=bn4['mean'],
                           invstd=bn4['invstd'])
        else:
            ops.conv_gemm(g, dpre, w6, g_v4, mask_mode=MASK_POS, mask_src=v4, stats_mode=STATS_BNBWD,
                          stats=bn4['bstats'], xhat_src=u0, mean=bn4['mean'], invstd=bn4['invstd'])
        g_u0 = self._bn_bwd(ctx, P, G, 'decoder.4', g_v4, u0, G['decoder.3.bias'])
        self._conv_wgrad(ctx, 'decoder.3', rd, g_u0)
        g_d1, g_d1h = self._conv_dgrad_bn(ctx, P, G, 'decoder.3', g_u0, rd.shape[1:3], 'decoder.1.block.5',
                                          ctx.saved['decoder.1'][3], True, mask_mode=MASK_POS, mask_src=rd)
        g_d0, g_d0h = self.resblock_bwd(ctx, P, G, 'decoder.1', g_d1, g_d1h, prev_block='decoder.0')
        return self.resblock_bwd(ctx, P, G, 'decoder.0', g_d0, g_d0h)[0]

    def bottleneck_bwd(self, ctx, g_st, z_e, z_q, coef):
        """dz = g_st + coef*(z_e - z_q) with encoder.5's BN5 backward sums."""
        bn = self._bn(ctx, 'encoder.5.block.5')
        dz = torch.empty_like(z_e)
        ops.vq_bwd_stats64(g_st, z_e, z_q, coef, dz, h=ctx.saved['encoder.5'][3], mean=bn['mean'],
                           invstd=bn['invstd'], stats=bn['bstats'])
        return dz

    def bottleneck_bwd_sum(self, ctx, g_st, g_ze):
        """dz = g_st + g_ze (either may be None) with encoder.5's BN5 backward sums (generic autograd path)."""
        bn = self._bn(ctx, 'encoder.5.block.5')
        a, b = (g_st, g_ze) if g_ze is not None else (None, g_st)
        dz = torch.empty_like(b)
        ops.vq_bwd_stats64(a, b, None, 1.0, dz, h=ctx.saved['encoder.5'][3], mean=bn['mean'], invstd=bn['invstd'],
                           stats=bn['bstats'])
        return dz

    def encoder_bwd(self, ctx, P, G, g_ze, need_dx=False):
        x, a0, v0 = ctx.x, ctx.a0, ctx.v0
        bn1 = self._bn(ctx, 'encoder.1')
        g_r5, g_r5h = self.resblock_bwd(ctx, P, G, 'encoder.5', g_ze, None, prev_block='encoder.4')
        g_a1 = self.resblock_bwd(ctx, P, G, 'encoder.4', g_r5, g_r5h)[0]
        self._conv_wgrad(ctx, 'encoder.3', v0, g_a1, q_colsum=G['encoder.3.bias'])
        # encoder.3's data gradient flows through encoder.2's ReLU into encoder.1's BatchNorm backward: one launch
        # (the 40 x 128 grid is 9 tiles per CTA at batch 32: the surplus tile is re-read from the stored g)
        _, g_a0 = self._conv_dgrad_bn(ctx, P, G, 'encoder.3', g_a1, v0.shape[1:3], 'encoder.1', a0, True,
                                      mask_mode=MASK_POS, mask_src=v0)
        b, h, w, cin = x.shape
        _, oh, ow, _ = a0.shape
        g = ops.conv_geom(b, h, w, cin, oh, ow, self.dim, 4, 2, 1, 0, 16, cin * 16, ops.BF16)
        gw = g
        if self.c1 is not None:
            self._side(ctx, lambda: ops.c1_wgrad_tc(gw, x, g_a0, G['encoder.0.weight']), x, g_a0)
        else:
            self._side(ctx, lambda: ops.conv_wgrad(gw, x, g_a0, G['encoder.0.weight']), x, g_a0)
        dx = None
        if need_dx:
            dx = torch.empty_like(x)
            g = ops.conv_geom(b, oh, ow, self.dim, h, w, cin, 4, 2, 1, 1, cin * 16, 16, ops.BF16)
            ops.conv_gemm(g, g_a0, P['encoder.0.weight'], dx)
        return dx

    def finish_backward(self, ctx, G, part=None):
        """Sum the per-CTA weight-gradient partials into the reference-layout gradients (one launch).
        ``part`` restricts this to one half of the network (see ``unpack_part``)."""
        if ctx.forked:
            torch.cuda.current_stream().wait_stream(self.side)
            ctx.forked = False
        ctx.keep = []
        self.unpack_part(G, part, fused=ctx.fused)

    def unpack_part(self, G, part=None, fused=False):
        """Unpack the gradient partials of 'decoder', 'encoder' or (None) all layers on the current stream.
        ``fused``: the caller is a ``begin(fused=True)`` step, whose next ``begin`` clears the images with the step's
        other accumulators; any other caller (the generic autograd backward) must leave them cleared itself, or the
        next backward would add onto this one's images."""
        ops.wgrad_unpack_many(self._unpack_items(G, part))
        for L in self.layers.values():      # a layer's partials are consumed once
            if part is None or L.name.startswith(part + '.'):
                L.nslices = 0
        if self.dwa_in_zbuf and fused:
            self.dwa_dirty = True
        if not self.wgrad_sliced and not (self.dwa_in_zbuf and fused):
            if part is None:
                self.dwa.zero_()
            elif part == 'decoder':
                self.dwa[self.dwa_split:].zero_()
            else:
                self.dwa[:self.dwa_split].zero_()

    # ------------------------------------------------------------------ bottleneck (models.py:127-142)
    def quantize(self, z_e, codebook, want_codes=True, sqerr=None):
        n = z_e.numel() // z_e.shape[-1]
        idx = torch.empty((n,), dtype=torch.int64, device=z_e.device)
        codes = torch.empty_like(z_e) if want_codes else None
        mode = self.vq_mode if ops.vq_mode_serves(codebook.shape[0], codebook.shape[1], self.vq_mode) else 0
        ops.vq_forward(z_e, codebook, idx, codes, sqerr, mode=mode)
        return idx, codes

"""The reference's other model families on the same C-ABI kernels: the convolutional ``VAE``
(/root/reference/src/models.py:64-118) and the ``GatedPixelCNN`` prior (:218-341).  Same class names, constructor
signatures, attribute names and ``state_dict`` keys as the reference; fp32 NHWC activations.

Kernels: the square <= 4 x 4 layers run on ``nsg_conv_gemm`` / ``nsg_conv_wgrad`` (any channel counts, stride 1-2), the
5 x 5 layers of the VAE and the rectangular masked stacks of the PixelCNN on ``nsg_conv_rect`` / ``nsg_conv_rect_wgrad``;
BatchNorm through ``nsg_bn_finalize / bn_apply / bn_bwd_reduce / bn_bwd_apply``; ``nsg_vae_reparam_kl`` (+ ``_bwd``),
``nsg_sqerr_sum``, ``nsg_gated_act``, ``nsg_gather_rows`` (embeddings), ``nsg_softmax_sample``.  The torch.nn modules below
are parameter containers (and ``print(model)`` / ``state_dict`` fidelity): their ``forward`` is never called, and
nothing here computes in PyTorch except drawing random numbers (``torch.randn`` / ``torch.rand`` on the device -- the
reference's ``rsample`` / ``multinomial`` draws).
"""
import torch
import torch.nn as nn
from torch.autograd import Function

from . import ops
from .ops import PRO_NONE, STATS_MOMENTS  # noqa: F401
from .engine import Engine, BN_EPS, BN_MOMENTUM, _out_hw
from .models import weights_init


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


def _nchw_view(t):
    return t.permute(0, 3, 1, 2)


# ---------------------------------------------------------------------------------------------- convolution helpers
_ENG = Engine(1, 1, 1)        # fp32 CUDA-core helpers of the general square kernels (stateless)


def _conv_fwd(x, w, bias, k, s, p, transposed, stats=None, tanh_out=False):
    if k <= 4:
        return _ENG._conv_fwd(x, w, bias, k, s, p, transposed, ops.prologue(PRO_NONE), stats=stats, tanh_out=tanh_out)
    assert s == 1
    b, h, wd, cin = x.shape
    oh, ow = _out_hw(h, wd, k, 1, p, transposed)
    if transposed:      # (Cin, Cout, k, k): contract Cin, gather like a transposed convolution
        cout, wci, wco = w.shape[1], w.shape[1] * k * k, k * k
    else:
        cout, wci, wco = w.shape[0], k * k, cin * k * k
    out = torch.empty((b, oh, ow, cout), dtype=torch.float32, device=x.device)
    return ops.conv_rect(x, w, out, k, k, p, p, 1 if transposed else 0, wci, wco, bias=bias, act=1 if tanh_out else 0,
                         stats=stats)


def _conv_dgrad(dy, w, k, s, p, transposed, in_hw):
    if k <= 4:
        return _ENG._conv_dgrad(dy, w, k, s, p, transposed, in_hw)
    b = dy.shape[0]
    if transposed:      # forward contracted Cin of (Cin, Cout, k, k); the data gradient contracts Cout, conv-like gather
        cin_f, wci, wco, gather = w.shape[0], k * k, w.shape[1] * k * k, 0
    else:
        cin_f, wci, wco, gather = w.shape[1], w.shape[1] * k * k, k * k, 1
    out = torch.empty((b, in_hw[0], in_hw[1], cin_f), dtype=torch.float32, device=dy.device)
    return ops.conv_rect(dy, w, out, k, k, p, p, gather, wci, wco)


def _conv_wgrad(x, dy, dw, k, s, p, transposed, dbias):
    if k <= 4:
        return _ENG._conv_wgrad(x, dy, dw, k, s, p, transposed, None, dbias=dbias)
    cin, cout = x.shape[-1], dy.shape[-1]
    if transposed:      # dW[ci][co][t]: P = dy gathered over x's pixels, Q = x
        ops.conv_rect_wgrad(dy, x, dw, k, k, p, p, k * k, cout * k * k)
        ops.colsum(dy, dbias)
    else:               # dW[co][ci][t]: P = x gathered over dy's pixels, Q = dy
        ops.conv_rect_wgrad(x, dy, dw, k, k, p, p, k * k, cin * k * k, q_colsum=dbias)


# ---------------------------------------------------------------------------------------------- VAE (models.py:64-118)
# (conv, norm, transposed, k, s, p, relu)
_VAE_LAYERS = [('encoder.0', 'encoder.1', False, 4, 2, 1, True), ('encoder.3', 'encoder.4', False, 4, 2, 1, True),
               ('encoder.6', 'encoder.7', False, 5, 1, 0, True), ('encoder.9', 'encoder.10', False, 3, 1, 0, False),
               ('decoder.0', 'decoder.1', True, 3, 1, 0, True), ('decoder.3', 'decoder.4', True, 5, 1, 0, True),
               ('decoder.6', 'decoder.7', True, 4, 2, 1, True)]


class _VAEFn(Function):
    """Autograd boundary of the whole VAE: forward = explicit kernel schedule, backward = its reverse.  Returns
    (x_tilde NCHW view, kl_div scalar)."""

    @staticmethod
    def forward(ctx, module, x, eps, *params):
        names = [n for n, _ in module.named_parameters()]
        P = {n: p.detach() for n, p in zip(names, params)}
        B = dict(module.named_buffers())
        training = module.training
        dev = x.device
        h = _nhwc(x.detach().float())
        saved = []
        for i, (conv, norm, tr, k, s, p, relu) in enumerate(_VAE_LAYERS):
            if i == 4:      # bottleneck: reparameterisation + KL on the encoder's normalised output (mu | logvar)
                st = h
                zc = st.shape[-1] // 2
                if eps is None:
                    eps = torch.randn(st.shape[:-1] + (zc,), dtype=torch.float32, device=dev)     # rsample, :113
                else:
                    eps = _nhwc(eps.detach().float())
                z = torch.empty_like(eps)
                kl_sum = torch.zeros(1, dtype=torch.float64, device=dev)
                ops.vae_reparam_kl(st, eps, z, kl_sum)
                pixels = z.numel() // zc
                h = z
            c = P[conv + '.weight'].shape[1 if tr else 0]
            stats = torch.zeros(2 * c, dtype=torch.float64, device=dev) if training else None
            pre = _conv_fwd(h, P[conv + '.weight'], P[conv + '.bias'], k, s, p, tr, stats=stats)
            coef = torch.empty((4, c), dtype=torch.float32, device=dev)      # mean | invstd | scale | shift
            if training:
                ops.bn_finalize(stats, pre.numel() // c, P[norm + '.weight'], P[norm + '.bias'], BN_EPS, BN_MOMENTUM,
                                B[norm + '.running_mean'], B[norm + '.running_var'], B[norm + '.num_batches_tracked'],
                                coef[0], coef[1], coef[2], coef[3])
            else:
                ops.bn_eval_coeffs(P[norm + '.weight'], P[norm + '.bias'], B[norm + '.running_mean'],
                                   B[norm + '.running_var'], BN_EPS, coef[2], coef[3])
            y = torch.empty_like(pre)
            ops.bn_apply(pre, coef[2], coef[3], y, relu_out=relu)
            saved.append((h, pre, coef))
            h = y
        xt = _conv_fwd(h, P['decoder.9.weight'], P['decoder.9.bias'], 4, 2, 1, True, tanh_out=True)     # :90-91
        kl = (kl_sum / pixels).to(torch.float32).reshape(())                                         # .sum(1).mean(), :109
        ctx.P, ctx.names, ctx.saved, ctx.h_last, ctx.xt = P, names, saved, h, xt
        ctx.st, ctx.eps, ctx.pixels, ctx.training = st, eps, pixels, training
        return _nchw_view(xt), kl

    @staticmethod
    def backward(ctx, g_xt, g_kl):
        if not ctx.training:
            raise RuntimeError('VAE.backward in eval mode is not implemented (the reference trains in train mode)')
        P, saved = ctx.P, ctx.saved
        G = {k: torch.zeros_like(v) for k, v in P.items()}
        dev = ctx.xt.device
        g = torch.empty_like(ctx.xt)
        if g_xt is None:
            g.zero_()
        else:
            ops.tanh_bwd(_nhwc(g_xt.float()), ctx.xt, g)
        _conv_wgrad(ctx.h_last, g, G['decoder.9.weight'], 4, 2, 1, True, G['decoder.9.bias'])
        g = _conv_dgrad(g, P['decoder.9.weight'], 4, 2, 1, True, ctx.h_last.shape[1:3])
        for i in range(len(_VAE_LAYERS) - 1, -1, -1):
            conv, norm, tr, k, s, p, relu = _VAE_LAYERS[i]
            x_in, pre, coef = saved[i]
            c = pre.shape[-1]
            bst = torch.zeros(2 * c, dtype=torch.float64, device=dev)
            ops.bn_bwd_reduce(g, pre, coef[0], coef[1], bst, scale=coef[2], shift=coef[3], relu_mask=relu)
            dpre = torch.empty_like(pre)
            ops.bn_bwd_apply(g, pre, coef[0], coef[1], P[norm + '.weight'], bst, pre.numel() // c, dpre,
                             dgamma=G[norm + '.weight'], dbeta=G[norm + '.bias'], scale=coef[2], shift=coef[3],
                             relu_mask=relu)
            _conv_wgrad(x_in, dpre, G[conv + '.weight'], k, s, p, tr, G[conv + '.bias'])
            if i == 0:
                break
            g = _conv_dgrad(dpre, P[conv + '.weight'], k, s, p, tr, x_in.shape[1:3])
            if i == 4:      # through the reparameterisation: d(mu | logvar) from dz and from the KL term
                dst = torch.empty_like(ctx.st)
                kl_scale = (float(g_kl) if g_kl is not None else 0.0) / ctx.pixels
                ops.vae_reparam_kl_bwd(ctx.st, ctx.eps, g, kl_scale, dst)
                g = dst
        return (None, None, None) + tuple(G[n] for n in ctx.names)


class VAE(nn.Module):
    """models.py:64-118.  ``forward(x) -> (x_tilde, kl_div)``; ``forward(x, eps=...)`` supplies the noise of ``rsample``
    (NCHW (B, z_dim, h, w)) instead of drawing it."""

    def __init__(self, input_dim, dim, z_dim):
        super(VAE, self).__init__()
        self.encoder = nn.Sequential(
            nn.Conv2d(input_dim, dim, 4, 2, 1), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.Conv2d(dim, dim, 4, 2, 1), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.Conv2d(dim, dim, 5, 1, 0), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.Conv2d(dim, z_dim * 2, 3, 1, 0), nn.BatchNorm2d(z_dim * 2))
        self.decoder = nn.Sequential(
            nn.ConvTranspose2d(z_dim, dim, 3, 1, 0), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.ConvTranspose2d(dim, dim, 5, 1, 0), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.ConvTranspose2d(dim, dim, 4, 2, 1), nn.BatchNorm2d(dim), nn.ReLU(True),
            nn.ConvTranspose2d(dim, input_dim, 4, 2, 1), nn.Tanh())
        self.apply(weights_init)

    def forward(self, x, eps=None):
        if not x.is_cuda:
            raise ops.A.NsgError('VAE runs on the B200 kernels only: move the model and its input to cuda')
        return _VAEFn.apply(self, x, eps, *self.parameters())


class _SqErrFn(Function):
    """sum (a - b)^2 through nsg_sqerr_sum (forward and gradient with respect to a)."""

    @staticmethod
    def forward(ctx, a, b):
        a32, b32 = a.detach().float().contiguous(), b.detach().float().contiguous()
        total = torch.zeros(1, dtype=torch.float64, device=a.device)
        ops.sqerr_sum(a32, b32, total=total)
        ctx.save_for_backward(a32, b32)
        return total.to(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, g):
        a32, b32 = ctx.saved_tensors
        grad = torch.empty_like(a32)
        ops.sqerr_sum(a32, b32, grad=grad, grad_scale=float(g))
        return grad, None


def mse_loss(x_tilde, x, kl_d):
    """loss.py:22-24: ``F.mse_loss(x_tilde, x, size_average=False) / x.size(0) + kl_d``."""
    return _SqErrFn.apply(x_tilde, x) / x.size(0) + kl_d


# ---------------------------------------------------------------------------------------------- PixelCNN (models.py:218-341)
class GatedActivation(nn.Module):
    def __init__(self):
        super(GatedActivation, self).__init__()

    def forward(self, x):
        """models.py:222-225 on an NCHW tensor (stand-alone use; the PixelCNN below calls the kernel on NHWC directly)."""
        xh = _nhwc(x.float())
        out = torch.empty(xh.shape[:-1] + (xh.shape[-1] // 2,), dtype=torch.float32, device=x.device)
        return _nchw_view(ops.gated_act(xh, out))


class GatedMaskedConv2d(nn.Module):
    """models.py:228-277; parameters only -- ``GatedPixelCNN.forward`` schedules the kernels."""

    def __init__(self, mask_type, dim, kernel, residual=True, n_classes=10):
        super(GatedMaskedConv2d, self).__init__()
        assert kernel % 2 == 1, "Kernel size must be odd"
        self.mask_type, self.residual, self.kernel, self.dim = mask_type, residual, kernel, dim
        self.class_cond_embedding = nn.Embedding(n_classes, 2 * dim)
        self.vert_stack = nn.Conv2d(dim, dim * 2, (kernel // 2 + 1, kernel), 1, (kernel // 2, kernel // 2))
        self.vert_to_horiz = nn.Conv2d(2 * dim, 2 * dim, 1)
        self.horiz_stack = nn.Conv2d(dim, dim * 2, (1, kernel // 2 + 1), 1, (0, kernel // 2))
        self.horiz_resid = nn.Conv2d(dim, dim, 1)
        self.gate = GatedActivation()

    def make_causal(self):
        self.vert_stack.weight.data[:, :, -1].zero_()          # mask final row (in place, as the reference: :255-257)
        self.horiz_stack.weight.data[:, :, :, -1].zero_()      # mask final column

    def forward(self, x_v, x_h, h):
        """x_v, x_h: NHWC fp32 (B, H, W, dim); h: (B,) int64 class labels.  Returns (out_v, out_h) NHWC."""
        if self.mask_type == 'A':
            self.make_causal()
        k, d = self.kernel, self.dim
        b, hh, ww, _ = x_v.shape
        dev = x_v.device
        cond = torch.empty((b, 2 * d), dtype=torch.float32, device=dev)
        ops.gather_rows(self.class_cond_embedding.weight.detach(), h, cond)                     # :262
        wv, wh = self.vert_stack.weight.detach(), self.horiz_stack.weight.detach()
        kv = k // 2 + 1
        # vertical stack, only the rows the reference keeps after its crop (:263-264)
        hv = torch.empty((b, hh, ww, 2 * d), dtype=torch.float32, device=dev)
        ops.conv_rect(x_v, wv, hv, kv, k, k // 2, k // 2, 0, kv * k, d * kv * k, bias=self.vert_stack.bias.detach())
        out_v = torch.empty((b, hh, ww, d), dtype=torch.float32, device=dev)
        ops.gated_act(hv, out_v, cond=cond)
        # horizontal stack, cropped to the first W columns (:267-268)
        hz = torch.empty((b, hh, ww, 2 * d), dtype=torch.float32, device=dev)
        ops.conv_rect(x_h, wh, hz, 1, kv, 0, k // 2, 0, kv, d * kv, bias=self.horiz_stack.bias.detach())
        v2h = torch.empty_like(hz)
        ops.conv_rect(hv, self.vert_to_horiz.weight.detach(), v2h, 1, 1, 0, 0, 0, 1, 2 * d,
                      bias=self.vert_to_horiz.bias.detach(), add_src=hz)                        # v2h + h_horiz
        out = torch.empty((b, hh, ww, d), dtype=torch.float32, device=dev)
        ops.gated_act(v2h, out, cond=cond)
        out_h = torch.empty_like(out)
        ops.conv_rect(out, self.horiz_resid.weight.detach(), out_h, 1, 1, 0, 0, 0, 1, d,
                      bias=self.horiz_resid.bias.detach(), add_src=x_h if self.residual else None)
        return out_v, out_h


class GatedPixelCNN(nn.Module):
    """models.py:280-341 (inference: the reference ships no trainer for it).  ``forward(x, label)``: x (B, H, W) int64
    codes -> logits (B, input_dim, H, W).  Unlike the reference (which crops rows with ``size(-1)`` and columns with
    ``size(-2)``, :264,268, and so only runs on square inputs) any H x W works; on square inputs the two agree."""

    def __init__(self, input_dim=256, dim=64, n_layers=15, n_classes=10):
        super(GatedPixelCNN, self).__init__()
        self.dim = dim
        self.embedding = nn.Embedding(input_dim, dim)
        self.layers = nn.ModuleList()
        for i in range(n_layers):
            self.layers.append(GatedMaskedConv2d('A' if i == 0 else 'B', dim, 7 if i == 0 else 3, i != 0, n_classes))
        self.output_conv = nn.Sequential(nn.Conv2d(dim, 512, 1), nn.ReLU(True), nn.Conv2d(512, input_dim, 1))
        self.apply(weights_init)

    @torch.no_grad()
    def _logits_nhwc(self, x, label):
        if not x.is_cuda:
            raise ops.A.NsgError('GatedPixelCNN runs on the B200 kernels only: move the model and its input to cuda')
        b, hh, ww = x.shape
        emb = torch.empty((b, hh, ww, self.dim), dtype=torch.float32, device=x.device)
        ops.gather_rows(self.embedding.weight.detach(), x.reshape(-1).contiguous(), emb.view(-1, self.dim))     # :311-313
        x_v = x_h = emb
        for layer in self.layers:
            x_v, x_h = layer(x_v, x_h, label)
        c0, c2 = self.output_conv[0], self.output_conv[2]
        hid = torch.empty((b, hh, ww, c0.out_channels), dtype=torch.float32, device=x.device)
        ops.conv_rect(x_h, c0.weight.detach(), hid, 1, 1, 0, 0, 0, 1, self.dim, bias=c0.bias.detach(), act=2)
        logits = torch.empty((b, hh, ww, c2.out_channels), dtype=torch.float32, device=x.device)
        ops.conv_rect(hid, c2.weight.detach(), logits, 1, 1, 0, 0, 0, 1, c0.out_channels, bias=c2.bias.detach())
        return logits

    def forward(self, x, label):
        return _nchw_view(self._logits_nhwc(x, label))

    @torch.no_grad()
    def generate(self, label, shape=(8, 8), batch_size=64, uniforms=None):
        """models.py:327-341: raster-scan ancestral sampling.  One categorical draw per (item, position) by inverse CDF
        (``nsg_softmax_sample``) from uniform numbers -- ``uniforms`` (H, W, B) or drawn with ``torch.rand`` on the
        device, standing in for the reference's ``probs.multinomial(1)``.  (The reference's own allocation,
        ``torch.zeros((batch_size, shape))`` with a tuple, raises; the intended (batch_size, H, W) is used.)"""
        dev = next(self.parameters()).device
        hh, ww = shape
        x = torch.zeros((batch_size, hh, ww), dtype=torch.int64, device=dev)
        classes = self.output_conv[2].out_channels
        for i in range(hh):
            for j in range(ww):
                logits = self._logits_nhwc(x, label)
                u = uniforms[i, j].to(dev, torch.float32).contiguous() if uniforms is not None else \
                    torch.rand(batch_size, dtype=torch.float32, device=dev)
                ops.softmax_sample(logits, hh * ww * classes, (i * ww + j) * classes, batch_size, classes, u,
                                   x.view(-1)[i * ww + j:], hh * ww)
        return x

"""The train_vqvae loop body (train.py:109-136) captured once as a CUDA graph and replayed per batch.

The step is ~100 kernel launches of a few microseconds each; driven from Python it is launch-bound,
so the whole iteration -- zero_grad, forward, the three loss terms, backward, [gradient all-reduce], Adam --
is recorded on a capture stream and replayed with one ``cudaGraphLaunch``.  Every entry point of the C ABI is
capturable (no host synchronisation, no allocation: buffers come from the graph's private torch pool).
"""
import torch

from . import optim as _optim


class GraphedTrainStep(object):
    """``step = GraphedTrainStep(model, optimizer, (B, n_mels, T), beta); losses = step(c)``.

    ``c`` may live on the host (pinned for an asynchronous copy) or on the device; it is copied into the
    graph's static input buffer.  Returns the static float64 device tensor [loss_recons, loss_vq, loss_commit]
    (valid until the next call).  ``grad_hook(flat_grads)`` (optional) is captured between backward and the
    optimizer step -- bench.py passes the NCCL all-reduce there.
    """

    def __init__(self, model, optimizer, batch_shape, beta=1.0, grad_hook=None):
        if not isinstance(optimizer, _optim.Adam):
            raise TypeError('GraphedTrainStep needs neural_sound_generation_b200.optim.Adam (device-side step count)')
        self.model, self.optimizer, self.beta = model, optimizer, float(beta)
        self._grad_hook, self._batch_shape = grad_hook, tuple(batch_shape)
        self._copy, self._stage, self._stage_free, self._pending, self._k = None, None, [None, None], None, 0
        self._capture()

    def _hyper(self):
        g = self.optimizer.param_groups[0]
        m = self.model
        return (float(g['lr']), float(g['betas'][0]), float(g['betas'][1]), float(g['eps']),
                float(self.optimizer.grad_scale), m.conv_impl, m.vq_mode, m.act_dtype)

    def _capture(self):
        """(Re)capture.  lr / betas / eps / grad_scale are by-value kernel arguments of the recorded Adam launch and the
        engine choice fixes the recorded kernels, so ``__call__`` re-captures when any of them has changed
        (scheduler, ``load_state_dict``) instead of silently replaying the old values."""
        model, optimizer, grad_hook, batch_shape = self.model, self.optimizer, self._grad_hook, self._batch_shape
        self._captured_hyper = self._hyper()
        self.engine = model.engine() if model.conv_impl == 2 else None      # the graph writes into its buffers
        flat = model.flatten_parameters()
        dev = flat['params'].device
        self.c = torch.zeros(tuple(batch_shape), dtype=torch.float32, device=dev)
        # one eager iteration configures every kernel variant (cudaFuncSetAttribute is not capturable); the
        # model / optimizer state it touches is restored afterwards
        saved = [t.clone() for t in self._state()]
        l0, _ = self._body(grad_hook)
        torch.cuda.synchronize(dev)
        for t, s in zip(self._state(), saved):
            t.copy_(s)
        self._losses_host = torch.empty(l0.shape, dtype=l0.dtype).pin_memory()      # read_losses()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses, self.outputs = self._body(grad_hook)
        torch.cuda.synchronize(dev)
        for t, s in zip(self._state(), saved):
            t.copy_(s)

    def read_losses(self):
        """[loss_recons, loss_vq, loss_commit] of the last replay as Python floats (waits for the step).  The copy goes
        into PINNED host memory and is queued behind the replay while it is still running, so what is left after the
        step is one stream synchronisation -- ``losses.tolist()`` copies into pageable memory through the driver's
        staging buffer, ~10 us more per step on the boxes measured."""
        self._losses_host.copy_(self.losses, non_blocking=True)
        torch.cuda.current_stream(self.c.device).synchronize()
        return self._losses_host.tolist()

    def _state(self):
        f = self.model._flat
        o = self.optimizer
        return [f['params'], o.exp_avg, o.exp_avg_sq, o.step_count] + [b for _, b in self.model.named_buffers()]

    def _body(self, grad_hook):
        # no zero_grad fill in the captured step: Adam clears every gradient right after using it; gradients written
        # by anything else between two calls (eager loss_and_grads / backward) are cleared in __call__
        if self.model.__dict__.get('_grads_dirty', True):
            self.optimizer.zero_grad()
        losses, outs = self.model.loss_and_grads(self.c, beta=self.beta)
        if grad_hook is not None:
            grad_hook(self.model._flat['grads'])
        self.optimizer.step(zero_grads=True)
        return losses, outs

    def close(self):
        """Release the CUDA graph (it holds the NCCL kernels of the captured collectives) -- call before
        ``torch.distributed.destroy_process_group()``, which otherwise can wait on the communicator forever."""
        dev = self.c.device
        torch.cuda.synchronize(dev)
        self.graph.reset()
        self.graph = None
        torch.cuda.synchronize(dev)

    def prefetch(self, c):
        """Start copying the NEXT batch (pinned host tensor) into a staging buffer on a private copy stream, so the
        transfer runs under the step in flight; the following ``step(c)`` with the same tensor object then only waits
        for that copy and moves the batch device-to-device.  Two staging buffers alternate."""
        dev = self.c.device
        if self._copy is None:
            self._copy = torch.cuda.Stream(device=dev)
            self._stage = [torch.empty_like(self.c) for _ in range(2)]
        k = self._k ^ 1
        if self._stage_free[k] is not None:                 # the step that last read this staging buffer
            self._copy.wait_event(self._stage_free[k])
        with torch.cuda.stream(self._copy):
            self._stage[k].copy_(c.reshape(self.c.shape), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy)
        self._pending, self._k = (c, k, ev), k

    def __call__(self, c):
        if self._hyper() != self._captured_hyper:
            torch.cuda.synchronize(self.c.device)
            self.graph.reset()
            self._capture()
        if self.model.__dict__.get('_grads_dirty', True):
            self.optimizer.zero_grad()
        if self._pending is not None and self._pending[0] is c:
            _, k, ev = self._pending
            self._pending = None
            cur = torch.cuda.current_stream(self.c.device)
            cur.wait_event(ev)
            self.c.copy_(self._stage[k])
            free = torch.cuda.Event()
            free.record(cur)
            self._stage_free[k] = free
        else:
            self.c.copy_(c.reshape(self.c.shape), non_blocking=True)
        self.graph.replay()
        return self.losses


class GraphedLossAndGrads(object):
    """Forward + the three loss terms + backward of ``train_vqvae`` as one CUDA graph, for ANY ``torch.optim``
    optimizer: ``fb = GraphedLossAndGrads(model, (B, 1, n_mels, T), beta); losses = fb(c); optimizer.step()``.

    The replay clears and refills the model's flat gradient buffer, whose slices are the parameters' ``.grad``
    (re-bound on every call, so ``optimizer.zero_grad(set_to_none=True)`` between calls is harmless and unnecessary).
    The reference's own ``optim.Adam(model.parameters())`` (main.py:124) therefore trains with one graph launch plus
    its own (foreach) update kernels per iteration."""

    def __init__(self, model, batch_shape, beta=1.0):
        self.model, self.beta = model, float(beta)
        flat = model.flatten_parameters()
        dev = flat['params'].device
        self.c = torch.zeros(tuple(batch_shape), dtype=torch.float32, device=dev)
        bufs = [b for _, b in model.named_buffers()]
        saved = [t.clone() for t in bufs]
        self._body()                        # configures every kernel variant outside the capture
        torch.cuda.synchronize(dev)
        for t, s_ in zip(bufs, saved):
            t.copy_(s_)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses, self.outputs = self._body()
        torch.cuda.synchronize(dev)
        for t, s_ in zip(bufs, saved):
            t.copy_(s_)
        flat['grads'].zero_()

    def _body(self):
        self.model._flat['grads'].zero_()
        return self.model.loss_and_grads(self.c, beta=self.beta)

    def __call__(self, c):
        self.model.bind_grads()
        self.c.copy_(c.reshape(self.c.shape), non_blocking=True)
        self.graph.replay()
        self.model.__dict__['_grads_dirty'] = True
        return self.losses


class GraphedInference(object):
    """``infer = GraphedInference(model, (B, 1, n_mels, T)); latents, x_tilde = infer(x)``.

    The eval-mode ``encode`` -> ``decode`` round trip of /root/reference/src/models.py:188-196 (BatchNorm running
    statistics, no gradient) captured once as a CUDA graph: batch-1 utterances are pure launch latency
    (SURVEY.md section 8d, config 5), so the ~25 launches replay with one ``cudaGraphLaunch``.  Returned tensors are
    static (valid until the next call)."""

    def __init__(self, model, x_shape):
        self.model = model.eval()
        dev = next(model.parameters()).device
        self.x = torch.zeros(tuple(x_shape), dtype=torch.float32, device=dev)
        with torch.no_grad():
            self._body()                    # configures every kernel variant outside the capture
            torch.cuda.synchronize(dev)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.latents, self.x_tilde = self._body()
        torch.cuda.synchronize(dev)

    def _body(self):
        latents = self.model.encode(self.x)
        return latents, self.model.decode(latents)

    def __call__(self, x):
        self.x.copy_(x.reshape(self.x.shape), non_blocking=True)
        self.graph.replay()
        return self.latents, self.x_tilde

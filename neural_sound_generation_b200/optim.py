"""Fused Adam over the model's flat parameter buffer (torch.optim.Adam semantics, main.py:124)."""
import torch

from . import ops


class Adam(torch.optim.Optimizer):
    """Drop-in for ``torch.optim.Adam(model.parameters(), lr=...)`` on a ``VQVAE``.

    Construct with ``Adam(model, lr=1e-3)``.  ``step()`` is ONE kernel over every parameter (plus a
    one-thread step-counter tick), with defaults betas=(0.9, 0.999), eps=1e-8, no weight decay, no amsgrad --
    the configuration the reference uses.  ``grad_scale`` is multiplied into the gradients (1/world for DP).
    """

    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        self.model = model
        flat = model.flatten_parameters()
        super(Adam, self).__init__(list(model.parameters()), dict(lr=lr, betas=betas, eps=eps))
        dev = flat['params'].device
        self.exp_avg = torch.zeros_like(flat['params'])
        self.exp_avg_sq = torch.zeros_like(flat['params'])
        self.step_count = torch.zeros((), dtype=torch.int64, device=dev)
        self.grad_scale = 1.0

    def zero_grad(self, set_to_none=False):
        self.model.flatten_parameters()
        self.model._flat['grads'].zero_()
        self.model.bind_grads()
        self.model.__dict__['_grads_dirty'] = False

    @torch.no_grad()
    def step(self, closure=None, zero_grads=False):
        """zero_grads=True also clears the gradients (what the next iteration's zero_grad() would do) in the same
        pass; GraphedTrainStep uses it so the captured step holds no separate fill."""
        f = self.model.flatten_parameters()
        g = self.param_groups[0]
        ops.adam_step(f['params'], f['grads'], self.exp_avg, self.exp_avg_sq, self.step_count, g['lr'],
                      g['betas'][0], g['betas'][1], g['eps'], self.grad_scale, zero_grads=zero_grads)
        if zero_grads:
            self.model.__dict__['_grads_dirty'] = False

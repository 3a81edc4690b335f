"""Fused Adam over the model's flat parameter buffer (torch.optim.Adam semantics, main.py:124)."""
import torch

from . import ops


class Adam(torch.optim.Optimizer):
    """Drop-in for ``torch.optim.Adam(model.parameters(), lr=...)`` on a ``VQVAE`` (main.py:124).

    Construct exactly as the reference does, ``Adam(model.parameters(), lr=1e-3)`` (the parameters must be ALL the
    parameters of one ``VQVAE``: the model that owns them is found through the back-reference its constructor leaves
    on each parameter), or with the model itself, ``Adam(model, lr=1e-3)``.  ``step()`` is ONE kernel over every
    parameter (plus a one-thread step-counter tick), with defaults betas=(0.9, 0.999), eps=1e-8, no weight decay, no
    amsgrad -- the configuration the reference uses.  ``grad_scale`` is multiplied into the gradients (1/world for DP).
    """

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise ValueError('optim.Adam: weight_decay / amsgrad are not implemented by the fused step '
                             '(the reference uses neither, main.py:124)')
        model = params if isinstance(params, torch.nn.Module) else self._owner(params)
        self.model = model
        flat = model.flatten_parameters()
        # the remaining keys are torch.optim.Adam's own defaults, carried so that state_dict() is interchangeable with
        # the reference's optimizer checkpoint (main.py:61-66, 'optimizer': optimizer.state_dict())
        super(Adam, self).__init__(list(model.parameters()),
                                   dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False,
                                        foreach=None, capturable=False, differentiable=False, fused=None))
        dev = flat['params'].device
        self.exp_avg = torch.zeros_like(flat['params'])
        self.exp_avg_sq = torch.zeros_like(flat['params'])
        self.step_count = torch.zeros((), dtype=torch.int64, device=dev)
        self.ticket = torch.zeros((), dtype=torch.int32, device=dev)    # last-block election inside the Adam kernel
        self.grad_scale = 1.0

    @staticmethod
    def _owner(params):
        params = list(params)
        if len(params) == 0:
            raise ValueError('optim.Adam: empty parameter list')
        if isinstance(params[0], dict):
            raise ValueError('optim.Adam: one parameter group (model.parameters()) expected, as main.py:124 passes')
        ref = getattr(params[0], '_nsg_owner', None)
        model = ref() if ref is not None else None
        if model is None:
            raise TypeError('optim.Adam: these are not the parameters of a neural_sound_generation_b200.VQVAE; '
                            'use torch.optim.Adam for other modules')
        mine = list(model.parameters())
        if len(mine) != len(params) or any(a is not b for a, b in zip(mine, params)):
            raise ValueError('optim.Adam: expected ALL parameters of the model, in model.parameters() order '
                             '(got %d of %d)' % (len(params), len(mine)))
        return model

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
                      g['betas'][0], g['betas'][1], g['eps'], self.grad_scale, zero_grads=zero_grads, ticket=self.ticket)
        if zero_grads:
            self.model.__dict__['_grads_dirty'] = False


    # ------------------------------------------------------------------ checkpoint interchange (main.py:61-66, 218-227)
    def state_dict(self):
        """torch.optim.Adam's layout: ``state[i] = {'step', 'exp_avg', 'exp_avg_sq'}`` per parameter (in
        ``model.parameters()`` order) + ``param_groups``; loads into ``torch.optim.Adam`` and back."""
        sd = super(Adam, self).state_dict()
        step = int(self.step_count.item())
        state = {}
        if step > 0:
            f = self.model.flatten_parameters()
            _, params = self.model._named()
            for i, (p, o) in enumerate(zip(params, f['offs'])):
                n = p.numel()
                state[i] = {'step': torch.tensor(float(step)),
                            'exp_avg': self.exp_avg[o:o + n].view(p.shape).clone(),
                            'exp_avg_sq': self.exp_avg_sq[o:o + n].view(p.shape).clone()}
        sd['state'] = state
        return sd

    def load_state_dict(self, state_dict):
        """Accepts a ``torch.optim.Adam`` state (the reference's checkpoints: 'step' a python int in the torch versions
        the reference targets, a tensor in current ones) or our own.  weight decay / amsgrad / maximize are not
        implemented by the fused kernel and are refused rather than ignored."""
        groups = state_dict['param_groups']
        if len(groups) != 1:
            raise ValueError('Adam.load_state_dict: one parameter group expected (main.py:124), got %d' % len(groups))
        g = groups[0]
        if g.get('weight_decay', 0) != 0 or g.get('amsgrad', False) or g.get('maximize', False):
            raise ValueError('Adam.load_state_dict: weight_decay / amsgrad / maximize are not supported by the fused step')
        f = self.model.flatten_parameters()
        _, params = self.model._named()
        if len(g['params']) != len(params):
            raise ValueError('Adam.load_state_dict: %d parameters in the checkpoint, %d in the model'
                             % (len(g['params']), len(params)))
        mine = self.param_groups[0]
        mine['lr'], mine['eps'] = float(g['lr']), float(g['eps'])
        mine['betas'] = (float(g['betas'][0]), float(g['betas'][1]))
        state = state_dict.get('state', {})
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        steps = set()
        for i, (p, o) in enumerate(zip(params, f['offs'])):
            st = state.get(g['params'][i], state.get(i))
            if st is None:
                continue
            n = p.numel()
            if tuple(st['exp_avg'].shape) != tuple(p.shape):
                raise ValueError('Adam.load_state_dict: moment %d has shape %s, parameter %s'
                                 % (i, tuple(st['exp_avg'].shape), tuple(p.shape)))
            self.exp_avg[o:o + n].copy_(st['exp_avg'].reshape(-1).to(self.exp_avg.dtype))
            self.exp_avg_sq[o:o + n].copy_(st['exp_avg_sq'].reshape(-1).to(self.exp_avg_sq.dtype))
            steps.add(int(st['step'].item()) if torch.is_tensor(st['step']) else int(st['step']))
        if len(steps) > 1:
            raise ValueError('Adam.load_state_dict: parameters at different step counts (%s): one shared counter here'
                             % sorted(steps))
        self.step_count.fill_(steps.pop() if steps else 0)

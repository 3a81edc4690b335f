"""Drop-in for ``train_vqvae`` of /root/reference/src/train.py:104-179.

Same signature, same reads of ``args`` (dataset, beta, log_interval), same two printed line formats, for both
branches of the reference loop: ``ljspeech`` (5-tuples whose ``c`` (B, 80, T) is the mel batch, train.py:108-147)
and the torchvision datasets (``(data, label)`` image batches, train.py:150-179).

What one iteration runs is chosen once per call from the model and the optimizer (``_step_fn``):

* our ``VQVAE`` on one-channel input + ``optim.Adam``  -> the whole iteration (forward, the three MSE terms of
  train.py:129-133, backward, Adam) replays as ONE CUDA graph per batch shape (``GraphedTrainStep``);
* our ``VQVAE`` on one-channel input + any ``torch.optim`` optimizer (what main.py:124 constructs) -> forward + losses
  + backward replay as one graph (``GraphedLossAndGrads``), then the optimizer's own ``step()``;
* anything else (three-channel images, a foreign model, ``args.cuda_graph = False``) -> the eager autograd path.

The zero-pad of x_tilde to the input width (train.py:118-120, a CPU round trip in the reference) happens inside the
loss kernel on the fused paths.  Graphs are cached per batch shape in a bounded LRU (``args.graph_cache``, default 8
shapes: variable-length mel batches would otherwise grow the cache without limit).
"""
from collections import OrderedDict

import torch
import torch.nn.functional as F
from tqdm import tqdm

from .models import to_scalar, VQVAE  # noqa: F401
from . import optim as _optim
from .graph import GraphedTrainStep, GraphedLossAndGrads


def _cached(model, key, limit, make):
    cache = model.__dict__.setdefault('_graphed_steps', OrderedDict())
    step = cache.get(key)
    if step is None:
        while len(cache) >= max(1, limit):
            _, old = cache.popitem(last=False)          # least recently used shape
            if getattr(old, 'graph', None) is not None:
                torch.cuda.synchronize()
                old.graph.reset()
        step = cache[key] = make()
    else:
        cache.move_to_end(key)
    return step


def _step_fn(args, model, optimizer):
    """Returns ``step(batch4d) -> (loss_recons, loss_vq)`` as python floats, having applied one optimizer update."""
    beta = float(args.beta)
    limit = int(getattr(args, 'graph_cache', 8))
    fused = isinstance(model, VQVAE) and model.input_dim == 1
    graph_ok = fused and getattr(args, 'cuda_graph', True)

    if graph_ok and isinstance(optimizer, _optim.Adam):
        def step(c):
            key = ('step', tuple(c.shape), beta, id(optimizer))
            g = _cached(model, key, limit, lambda: GraphedTrainStep(model, optimizer, c.shape, beta))
            g(c)
            lv = g.read_losses()                        # one host sync (reference: two .item())
            return lv[0], lv[1]
    elif graph_ok:
        def step(c):
            key = ('grads', tuple(c.shape), beta)
            lv = _cached(model, key, limit, lambda: GraphedLossAndGrads(model, c.shape, beta))(c).tolist()
            optimizer.step()
            return lv[0], lv[1]
    elif fused:
        def step(c):
            optimizer.zero_grad()
            losses, _ = model.loss_and_grads(c, beta=beta)
            optimizer.step()
            lv = losses.tolist()                        # one host sync (reference: two .item())
            return lv[0], lv[1]
    else:
        def step(c):
            optimizer.zero_grad()
            x_tilde, z_e_x, z_q_x = model(c)
            if x_tilde.size(3) != c.size(3):            # train.py:118-120
                x_tilde = F.pad(x_tilde, (0, c.size(3) - x_tilde.size(3)))
            terms = (F.mse_loss(x_tilde, c), F.mse_loss(z_q_x, z_e_x.detach()), F.mse_loss(z_e_x, z_q_x.detach()))
            (terms[0] + terms[1] + beta * terms[2]).backward()
            optimizer.step()
            return terms[0].item(), terms[1].item()
    return step


def train_vqvae(args, model, optimizer, train_loader, device, epoch):
    model.train()
    train_loss = 0
    step = _step_fn(args, model, optimizer)
    mel = args.dataset == 'ljspeech'
    # the mel loader yields (x, y, c, g, input_lengths) and only c is used (train.py:108-116); image loaders (data, _)
    batches = tqdm(enumerate(train_loader)) if mel else enumerate(train_loader)
    for batch_idx, item in batches:
        c = item[2].unsqueeze(1) if mel else item[0]
        c = c.to(device, non_blocking=True)
        loss_recons, loss_vq = step(c)
        train_loss = loss_recons + loss_vq
        if batch_idx % args.log_interval == 0:
            print('Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(
                epoch, batch_idx * len(c), len(train_loader.dataset),
                args.log_interval * batch_idx / len(train_loader), train_loss))
    print('====> Epoch: {} Average loss: {:.4f}'.format(epoch, train_loss / len(train_loader.dataset)))


def train_vae(args, model, optimizer, train_loader, device, epoch):
    """train.py:41-101: the convolutional VAE loop (row N4).  ``model(c) -> (x_tilde, kl_d)``,
    ``loss = mse_loss(pad(x_tilde), c, kl_d)`` (loss.py:22-24; the zero-pad of train.py:58-60 without its CPU round
    trip), same printed formats.  Eager: one explicit kernel schedule forward, its reverse backward (models_n4.py)."""
    from .models_n4 import mse_loss
    model.train()
    train_loss = 0
    mel = args.dataset == 'ljspeech'
    batches = tqdm(enumerate(train_loader)) if mel else enumerate(train_loader)
    for batch_idx, item in batches:
        optimizer.zero_grad()
        c = (item[2].unsqueeze(1) if mel else item[0]).to(device, non_blocking=True)
        x_tilde, kl_d = model(c)
        if x_tilde.size(3) != c.size(3):
            x_tilde = F.pad(x_tilde, (0, c.size(3) - x_tilde.size(3)))
        loss = mse_loss(x_tilde, c, kl_d)
        loss.backward()
        train_loss += loss.item()
        optimizer.step()
        if batch_idx % args.log_interval == 0:
            print('Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(
                epoch, batch_idx * len(c), len(train_loader.dataset),
                100. * batch_idx / len(train_loader), loss.item() / len(c)))
    print('====> Epoch: {} Average loss: {:.4f}'.format(epoch, train_loss / len(train_loader.dataset)))

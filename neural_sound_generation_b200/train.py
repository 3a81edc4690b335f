"""Drop-in for ``train_vqvae`` of /root/reference/src/train.py:104-179.

Same signature, same reads of ``args`` (dataset, beta, log_interval), same two printed line formats.
For our ``VQVAE`` the loop body runs the fused step (forward, the three MSE terms of train.py:129-133,
backward) through the C-ABI kernels; the zero-pad of x_tilde to the input width (train.py:118-120) is
done inside the loss kernel instead of through a CPU tensor.  Any ``torch.optim`` optimizer works;
``optim.Adam`` (ours) makes the update a single launch.
"""
import torch
import torch.nn.functional as F
from tqdm import tqdm

from .models import to_scalar, VQVAE  # noqa: F401
from . import optim as _optim
from .graph import GraphedTrainStep


def train_vqvae(args, model, optimizer, train_loader, device, epoch):
    model.train()
    train_loss = 0
    fused = isinstance(model, VQVAE)
    # the whole iteration replays as one CUDA graph per batch shape (opt out: args.cuda_graph = False)
    graphed = fused and isinstance(optimizer, _optim.Adam) and getattr(args, 'cuda_graph', True)
    if args.dataset == 'ljspeech':
        for batch_idx, (x, y, c, g, input_lengths) in tqdm(enumerate(train_loader)):
            if not graphed:
                optimizer.zero_grad()       # (the graphed step clears gradients inside Adam: no separate fill)
            # Prepare data (x, y, input_lengths are moved by the reference but never used: train.py:111-114)
            c = c.to(device, non_blocking=True) if c is not None else None
            c = c.unsqueeze(1)
            if graphed:
                key = (tuple(c.shape), float(args.beta), id(optimizer))
                cache = model.__dict__.setdefault('_graphed_steps', {})
                if key not in cache:
                    cache[key] = GraphedTrainStep(model, optimizer, c.shape, args.beta)
                lv = cache[key](c).tolist()
                train_loss = lv[0] + lv[1]
            elif fused:
                losses, _ = model.loss_and_grads(c, beta=args.beta)
                optimizer.step()
                lv = losses.tolist()                                   # one host sync (reference: two .item())
                train_loss = lv[0] + lv[1]
            else:
                x_tilde, z_e_x, z_q_x = model(c)
                target = F.pad(x_tilde, (0, c.size(3) - x_tilde.size(3)))
                loss_recons = F.mse_loss(target, c)
                loss_vq = F.mse_loss(z_q_x, z_e_x.detach())
                loss_commit = F.mse_loss(z_e_x, z_q_x.detach())
                loss = loss_recons + loss_vq + args.beta * loss_commit
                loss.backward()
                optimizer.step()
                train_loss = loss_recons.item() + loss_vq.item()

            if batch_idx % args.log_interval == 0:
                print('Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(
                    epoch, batch_idx * len(c), len(train_loader.dataset),
                    args.log_interval * batch_idx / len(train_loader),
                    train_loss
                    ))

        print('====> Epoch: {} Average loss: {:.4f}'.format(
              epoch, train_loss / len(train_loader.dataset)))

    else:
        for batch_idx, (data, _) in enumerate(train_loader):
            data = data.to(device)
            optimizer.zero_grad()
            x_tilde, z_e_x, z_q_x = model(data)
            loss_recons = F.mse_loss(x_tilde, data)
            loss_vq = F.mse_loss(z_q_x, z_e_x.detach())
            loss_commit = F.mse_loss(z_e_x, z_q_x.detach())
            loss = loss_recons + loss_vq + args.beta * loss_commit
            loss.backward()
            optimizer.step()

            train_loss = loss_recons.item() + loss_vq.item()

            if batch_idx % args.log_interval == 0:
                print('Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(
                    epoch, batch_idx * len(data), len(train_loader.dataset),
                    args.log_interval * batch_idx / len(train_loader),
                    train_loss
                    ))

        print('====> Epoch: {} Average loss: {:.4f}'.format(
              epoch, train_loss / len(train_loader.dataset)))

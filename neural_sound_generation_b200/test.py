"""Drop-in for ``test_vqvae`` of /root/reference/src/test.py:73-106 (eval-mode BatchNorm, no_grad)."""
import torch
import torch.nn.functional as F
from tqdm import tqdm


def test_vqvae(args, model, test_loader, device, epoch):
    model.eval()
    test_loss = 0
    with torch.no_grad():
        if args.dataset == 'ljspeech':
            loss_recons, loss_vq = 0., 0.
            for step, (x, y, c, g, input_lengths) in tqdm(enumerate(test_loader)):
                c = c.to(device) if c is not None else None
                c = c.unsqueeze(1)
                x_tilde, z_e_x, z_q_x = model(c)
                target = F.pad(x_tilde, (0, c.size(3) - x_tilde.size(3)))     # test.py:87-89 without the CPU hop
                loss_recons += F.mse_loss(target, c)
                loss_vq += F.mse_loss(z_q_x, z_e_x)
            loss_recons /= len(test_loader)
            loss_vq /= len(test_loader)
        else:
            loss_recons, loss_vq = 0., 0.
            for i, (data, _) in enumerate(test_loader):
                data = data.to(device)
                x_tilde, z_e_x, z_q_x = model(data)
                loss_recons += F.mse_loss(x_tilde, data)
                loss_vq += F.mse_loss(z_q_x, z_e_x)
            loss_recons /= len(test_loader)
            loss_vq /= len(test_loader)

        test_loss = loss_recons + loss_vq
        print('====> Test set loss: {:.4f}'.format(test_loss))
    return test_loss


def test_vae(args, model, test_loader, device, epoch):
    """test.py:32-70: evaluation loop of the convolutional VAE (row N4); returns the printed mean loss."""
    from .models_n4 import mse_loss
    model.eval()
    test_loss = 0
    mel = args.dataset == 'ljspeech'
    with torch.no_grad():
        for item in test_loader:
            c = (item[2].unsqueeze(1) if mel else item[0]).to(device)
            x_tilde, kl_d = model(c)
            if x_tilde.size(3) != c.size(3):
                x_tilde = F.pad(x_tilde, (0, c.size(3) - x_tilde.size(3)))
            test_loss += mse_loss(x_tilde, c, kl_d).item()
        test_loss /= len(test_loader.dataset)
        print('====> Test set loss: {:.4f}'.format(test_loss))
    return test_loss

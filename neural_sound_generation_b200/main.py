"""``main.py``-compatible driver for the VQ-VAE path of /root/reference/src/main.py:25-229.

Same command-line flags and defaults (main.py:25-58), same checkpoint dict and file name (main.py:61-66, 218-227:
``./models/<model>/checkpoint_<dataset>_<dim>_<z_dim>.pth.tar`` holding ``epoch / arch / state_dict / optimizer``, the
optimizer state in ``torch.optim.Adam``'s own layout), same per-epoch reconstruction dump
(main.py:155-162: ``<sampledir>/<dataset>/reconstruction_<model>_data_<dataset>_dim_<dim>_z_dim_<z_dim>_epoch_<e>.npy``).

Data (main.py:77-99):
  * ``--dataset ljspeech``: ``train.txt`` + mel ``.npy`` files under ``--datadir``, read, length-bucketed, cropped and
    padded as the reference's dataloader.py does, with the batch assembled on the device (``data.py``, row N2);
  * any other name is a ``torchvision.datasets`` class as in the reference (``MNIST`` / ``FashionMNIST``: one channel;
    ``CIFAR10``: ``input_dim = 3``, main.py:100-116), read from ``--datadir/<dataset>/`` -- the files must already be
    there (the reference passes ``download=True``; this build never touches the network);
  * ``--synthetic-batches N`` substitutes N uniform-[0,1] batches per epoch of the dataset's shape (mel 80 x
    ``--synthetic-frames``, 1 x 28 x 28, 3 x 32 x 32).

What is NOT here, and why (SURVEY.md section 8f ranks them after this row):
  * for image datasets the whole per-epoch sample block is skipped: the reference's own block unpacks 5-tuples
    (main.py:138-139) and cannot run on ``(data, label)`` loaders.
The Griffin-Lim audio of the per-epoch sample (main.py:164-197, row N3) runs on the device (``audio.py``) and is written
next to the ``.npy`` dump under the reference's file name.

    python -m neural_sound_generation_b200.main --model vqvae --dataset ljspeech --dim 64 --z-dim 512 \
        --synthetic-batches 100 --epochs 1
"""
import argparse
import os

import numpy as np
import torch

from . import audio
from .models import VQVAE
from .optim import Adam
from .models_n4 import VAE
from .train import train_vqvae, train_vae
from .test import test_vqvae, test_vae


def build_parser():
    """The reference's parser (main.py:25-58): flag names, types and defaults verbatim; our additions come last."""
    parser = argparse.ArgumentParser(description='VQ-VAE on mel spectrograms (B200-native hot path)')
    parser.add_argument('--batch-size', type=int, default=64, metavar='N')
    parser.add_argument('--lr-rate', type=float, default=1e-3, metavar='N')
    parser.add_argument('--dataset', type=str, default='MNIST', metavar='N')
    parser.add_argument('--datadir', type=str, default='./data/', metavar='N')
    parser.add_argument('--sampledir', type=str, default='./results/', metavar='N')
    parser.add_argument('--epochs', type=int, default=3000, metavar='N')
    parser.add_argument('--no-cuda', action='store_true', default=False)
    parser.add_argument('--seed', type=int, default=1, metavar='S')
    parser.add_argument('--log-interval', type=int, default=10, metavar='N')
    parser.add_argument('--model', type=str, default='vae', metavar='N')
    parser.add_argument('--beta', type=float, default=1.0)
    parser.add_argument('--dim', type=int, default=1, metavar='S')
    parser.add_argument('--z-dim', type=int, default=512, metavar='S')
    # ---- not in the reference
    parser.add_argument('--synthetic-batches', type=int, default=0, metavar='N',
                        help='use N synthetic uniform-[0,1] mel batches (80 x --synthetic-frames) per epoch')
    parser.add_argument('--synthetic-frames', type=int, default=256, metavar='T')
    parser.add_argument('--resume', action='store_true', default=False,
                        help='load ./models/<model>/checkpoint_<dataset>_<dim>_<z_dim>.pth.tar before training')
    parser.add_argument('--fp32', action='store_true', default=False,
                        help='fp32 parity engine instead of the automatic choice (bf16 TMA engine when dim == 64)')
    return parser


def parse_args(argv=None):
    args = build_parser().parse_args(argv)
    args.cuda = not args.no_cuda and torch.cuda.is_available()
    return args


def checkpoint_path(args):
    """main.py:61-65"""
    return './models/{}/checkpoint_{}_{}_{}.pth.tar'.format(args.model, args.dataset, args.dim, args.z_dim)


def save_checkpoint(args, state):
    filename = checkpoint_path(args)
    os.makedirs(os.path.dirname(filename), exist_ok=True)      # (the reference assumes the directory exists)
    torch.save(state, filename)


def sample_path(args, epoch):
    """main.py:155-161"""
    return os.path.join(args.sampledir, format(args.dataset),
                        'reconstruction_' + str(args.model) + '_data_' + str(args.dataset) + '_dim_' + str(args.dim)
                        + '_z_dim_' + str(args.z_dim) + '_epoch_' + str(epoch) + '.npy')


def audio_path(args, epoch, fft_size, hop_size):
    """main.py:183-191"""
    return os.path.join(args.sampledir, format(args.dataset),
                        'audio_recon_' + str(args.model) + '_data_' + str(args.dataset) + '_dim_' + str(args.dim)
                        + '_z_dim_' + str(args.z_dim) + '_epoch_' + str(epoch) + '_fftsize_' + str(fft_size)
                        + '_hopsize_' + str(hop_size) + '.wav')


class SyntheticMelLoader(object):
    """Stands in for the LJSpeech loaders: yields the reference's 5-tuples ``(x, y, c, g, input_lengths)`` with only
    ``c`` (B, 80, T) populated -- the one field the VQ-VAE loops read (train.py:108-116)."""

    def __init__(self, batches, batch_size, frames, seed, pin=False):
        g = torch.Generator().manual_seed(seed)
        self.batches = [torch.rand(batch_size, 80, frames, generator=g) for _ in range(batches)]
        if pin:
            self.batches = [b.pin_memory() for b in self.batches]
        self.dataset = range(batches * batch_size)
        self.batch_size = batch_size

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        for c in self.batches:
            b = c.shape[0]
            yield (torch.zeros(b, 1, dtype=torch.long), torch.zeros(b, 1, dtype=torch.long), c, None,
                   torch.full((b,), c.shape[2], dtype=torch.long))


IMAGE_SHAPES = {'MNIST': (1, 28, 28), 'FashionMNIST': (1, 28, 28), 'KMNIST': (1, 28, 28), 'CIFAR10': (3, 32, 32),
                'CIFAR100': (3, 32, 32)}


class SyntheticImageLoader(object):
    """Stands in for the torchvision loaders of main.py:84-99: yields ``(data, label)`` with data (B, C, H, W) in [0, 1]."""

    def __init__(self, batches, batch_size, shape, seed):
        g = torch.Generator().manual_seed(seed)
        self.batches = [torch.rand((batch_size,) + tuple(shape), generator=g) for _ in range(batches)]
        self.dataset = range(batches * batch_size)
        self.batch_size = batch_size

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        for d in self.batches:
            yield d, torch.zeros(d.shape[0], dtype=torch.long)


def get_loaders(args):
    if args.synthetic_batches > 0 and args.dataset != 'ljspeech':
        shape = IMAGE_SHAPES.get(args.dataset, (1, 28, 28))
        return (SyntheticImageLoader(args.synthetic_batches, args.batch_size, shape, args.seed),
                SyntheticImageLoader(max(1, args.synthetic_batches // 8), args.batch_size, shape, args.seed + 1))
    if args.synthetic_batches > 0:
        tr = SyntheticMelLoader(args.synthetic_batches, args.batch_size, args.synthetic_frames, args.seed, args.cuda)
        te = SyntheticMelLoader(max(1, args.synthetic_batches // 8), args.batch_size, args.synthetic_frames,
                                args.seed + 1, args.cuda)
        return tr, te
    if args.dataset == 'ljspeech':
        # the reference's get_audio_data_loaders (dataloader.py:437-489): train.txt under --datadir, 95 / 5 split,
        # length-bucketed batches cropped to hparams.max_time_steps -- assembled on the device (data.py)
        from . import data as D
        tr_set, te_set = D.MelNpyDataset(args.datadir, train=True), D.MelNpyDataset(args.datadir, train=False)
        frames = D.max_time_frames()
        tr = D.DeviceMelLoader(tr_set, args.batch_size, max_frames=frames, seed=args.seed, bucket=True)
        te = D.DeviceMelLoader(te_set, args.batch_size, max_frames=frames, seed=args.seed + 1, bucket=False, shuffle=True)
        return tr, te
    # main.py:84-99 / dataloader.py:45-60: a torchvision dataset under <datadir>/<dataset>/, ToTensor() images
    try:
        from torchvision import datasets, transforms
        cls = getattr(datasets, args.dataset)
    except (ImportError, AttributeError) as e:
        raise SystemExit('--dataset %s: not a torchvision dataset (%s)' % (args.dataset, e))
    root = './{}/{}/'.format(args.datadir, args.dataset)
    kwargs = {'num_workers': 1, 'pin_memory': True}
    try:
        tr = torch.utils.data.DataLoader(cls(root, train=True, download=False, transform=transforms.ToTensor()),
                                         batch_size=args.batch_size, shuffle=True, **kwargs)
        te = torch.utils.data.DataLoader(cls(root, train=False, download=False, transform=transforms.ToTensor()),
                                         batch_size=args.batch_size, shuffle=True, **kwargs)
    except RuntimeError as e:
        raise SystemExit('--dataset %s: files not found under %s and this build does not download (%s); '
                         'use --synthetic-batches N for a smoke run' % (args.dataset, root, e))
    return tr, te


def main(argv=None):
    args = parse_args(argv)
    torch.manual_seed(args.seed)
    if args.model not in ('vqvae', 'vae'):
        raise SystemExit('--model %s: vqvae or vae (main.py:117-121)' % args.model)
    if not args.cuda:
        raise SystemExit('the B200 path has no CPU execution mode (remove --no-cuda / run on a B200)')
    device = torch.device('cuda')
    train_loader, test_loader = get_loaders(args)
    input_dim = 3 if args.dataset == 'CIFAR10' else 1                                  # main.py:100-116
    vae = args.model == 'vae'
    if vae:         # main.py:117-118 (row N4): the convolutional VAE on the general fp32 kernels, torch's own Adam
        model = VAE(input_dim, args.dim, args.z_dim).to(device)
    else:
        model = VQVAE(input_dim, args.dim, args.z_dim).to(device)
        if args.fp32:
            model.set_precision('fp32')
        elif model.conv_impl != 2:
            # the reference's own default --dim 1, or any dim other than 64, is served by the general fp32 kernels
            print('--dim %d: the bf16 tensor-core engine serves dim == 64; running the fp32 CUDA-core engine' % args.dim)
    print(model)
    optimizer = (torch.optim.Adam if vae else Adam)(model.parameters(), lr=args.lr_rate)   # main.py:124
    last_epoch = 0
    if args.resume:
        ck = torch.load(checkpoint_path(args), map_location=device)
        model.load_state_dict(ck['state_dict'])
        optimizer.load_state_dict(ck['optimizer'])
        last_epoch = int(ck.get('epoch', 0))
        print('resumed %s at epoch %d' % (checkpoint_path(args), last_epoch))

    def state():
        return {'epoch': last_epoch, 'arch': args.model, 'state_dict': model.state_dict(),
                'optimizer': optimizer.state_dict()}

    try:
        for epoch in range(last_epoch + 1, args.epochs + 1):
            if vae:                                                                    # main.py:130-135
                train_vae(args, model, optimizer, train_loader, device, epoch)
                test_vae(args, model, test_loader, device, epoch)
            else:
                train_vqvae(args, model, optimizer, train_loader, device, epoch)
                test_vqvae(args, model, test_loader, device, epoch)
            with torch.no_grad():
                if args.dataset != 'ljspeech':      # (the reference's sample block unpacks mel 5-tuples, main.py:138)
                    last_epoch = epoch
                    save_checkpoint(args, state())
                    continue
                x, y, c, g, input_lengths = next(iter(test_loader))
                c = c.to(device).unsqueeze(1)
                print("Evaluating samples")
                reconstruction = model(c)[0]                                           # main.py:148-151
                reconstruction = reconstruction.squeeze(1).float().cpu().numpy()
                path = sample_path(args, epoch)
                os.makedirs(os.path.dirname(path), exist_ok=True)
                np.save(path, reconstruction, allow_pickle=True)
                # main.py:164-197: Griffin-Lim audio of the sample, on the device (audio.py / csrc/audio.cu, row N3).
                # As in the reference, the concatenation of the batch is only printed; the waveform is made from the
                # loop variable, i.e. the LAST item of the batch (main.py:168-181).
                print("Trying audio reconstruction on test set..")
                print((reconstruction.shape[1], reconstruction.shape[0] * reconstruction.shape[2]))
                sampling_rate, fft_size, hop_size, n_mels = 22050, 1024, 256, 80
                assert reconstruction.shape[1] == n_mels
                signal = audio.inv_mel_spectrogram(reconstruction[-1], sampling_rate, fft_size, hop_size, n_mels)
                audio.save_wav(signal, audio_path(args, epoch, fft_size, hop_size))
            last_epoch = epoch
            save_checkpoint(args, state())
    except KeyboardInterrupt:
        print("Interrupted!")
    finally:
        save_checkpoint(args, state())
    return 0


if __name__ == '__main__':
    raise SystemExit(main())

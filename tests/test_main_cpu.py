"""Host logic of the ``main.py``-compatible driver: flags and defaults, checkpoint and sample
file names (reference src/main.py:25-66, 155-161), optimizer-state interchange with ``torch.optim.Adam``.  No GPU."""
import os

import pytest
import torch

import neural_sound_generation_b200 as nsg
from neural_sound_generation_b200 import main as M

# src/main.py:25-58, transcribed: flag -> (type, default)
REFERENCE_FLAGS = {
    'batch_size': (int, 64), 'lr_rate': (float, 1e-3), 'dataset': (str, 'MNIST'), 'datadir': (str, './data/'),
    'sampledir': (str, './results/'), 'epochs': (int, 3000), 'no_cuda': (bool, False), 'seed': (int, 1),
    'log_interval': (int, 10), 'model': (str, 'vae'), 'beta': (float, 1.0), 'dim': (int, 1), 'z_dim': (int, 512),
}


def test_parser_has_the_reference_flags_and_defaults():
    args = M.parse_args([])
    for name, (typ, default) in REFERENCE_FLAGS.items():
        assert hasattr(args, name), name
        assert getattr(args, name) == default and isinstance(getattr(args, name), typ), name
    assert args.cuda == (torch.cuda.is_available())                       # main.py:57
    a = M.parse_args(['--batch-size', '32', '--lr-rate', '3e-4', '--dataset', 'ljspeech', '--model', 'vqvae', '--dim', '64',
                      '--z-dim', '256', '--beta', '0.25', '--no-cuda', '--log-interval', '5', '--epochs', '2',
                      '--seed', '7', '--datadir', '/d', '--sampledir', '/s'])
    assert (a.batch_size, a.lr_rate, a.dataset, a.model, a.dim, a.z_dim, a.beta, a.no_cuda, a.cuda) == \
           (32, 3e-4, 'ljspeech', 'vqvae', 64, 256, 0.25, True, False)


def test_checkpoint_and_sample_names_match_the_reference():
    a = M.parse_args(['--model', 'vqvae', '--dataset', 'ljspeech', '--dim', '64', '--z-dim', '512', '--sampledir', './results/'])
    assert M.checkpoint_path(a) == './models/vqvae/checkpoint_ljspeech_64_512.pth.tar'                  # main.py:61-65
    assert M.sample_path(a, 3) == os.path.join('./results/', 'ljspeech',
                                               'reconstruction_vqvae_data_ljspeech_dim_64_z_dim_512_epoch_3.npy')


def test_refusals_are_loud():
    with pytest.raises(SystemExit):
        M.main(['--model', 'pixelrnn', '--dataset', 'ljspeech', '--synthetic-batches', '1'])
    with pytest.raises(SystemExit):
        M.main(['--model', 'vqvae', '--dataset', 'ljspeech', '--no-cuda', '--synthetic-batches', '1'])


def test_synthetic_loader_yields_the_reference_tuple():
    ld = M.SyntheticMelLoader(3, 4, 32, seed=1)
    assert len(ld) == 3 and len(ld.dataset) == 12
    items = list(ld)
    x, y, c, g, lengths = items[0]
    assert c.shape == (4, 80, 32) and c.dtype == torch.float32 and 0.0 <= float(c.min()) and float(c.max()) <= 1.0
    assert g is None and lengths.tolist() == [32] * 4
    assert not torch.equal(items[0][2], items[1][2])


def test_optimizer_state_interchanges_with_torch_adam():
    """'optimizer': optimizer.state_dict() of the reference's checkpoints (main.py:218-227) loads into ours and back."""
    torch.manual_seed(0)
    m = nsg.VQVAE(1, 64, 512)
    ours = nsg.optim.Adam(m, lr=2e-3)
    assert ours.state_dict()['state'] == {}                              # nothing stepped yet
    params = list(m.parameters())
    ref = torch.optim.Adam(params, lr=1e-3, betas=(0.8, 0.99), eps=1e-7)
    for _ in range(2):
        for p in params:
            p.grad = torch.randn_like(p)
        ref.step()
    ours.load_state_dict(ref.state_dict())
    g = ours.param_groups[0]
    assert int(ours.step_count) == 2 and (g['lr'], g['betas'], g['eps']) == (1e-3, (0.8, 0.99), 1e-7)
    f = m._flat
    for p, o in zip(params, f['offs']):
        assert torch.equal(ours.exp_avg[o:o + p.numel()].view(p.shape), ref.state[p]['exp_avg'])
        assert torch.equal(ours.exp_avg_sq[o:o + p.numel()].view(p.shape), ref.state[p]['exp_avg_sq'])
    back = ours.state_dict()
    assert set(back) == {'state', 'param_groups'} and len(back['state']) == len(params)
    ref2 = torch.optim.Adam(params, lr=5e-3)
    ref2.load_state_dict(back)
    for p in params:
        assert torch.equal(ref2.state[p]['exp_avg'], ref.state[p]['exp_avg'])
        assert torch.equal(ref2.state[p]['exp_avg_sq'], ref.state[p]['exp_avg_sq'])
        assert float(ref2.state[p]['step']) == 2.0
    assert ref2.param_groups[0]['lr'] == 1e-3
    # old-style checkpoints carry python-int steps; unsupported options are refused, not ignored
    old = ref.state_dict()
    for st in old['state'].values():
        st['step'] = 2
    ours.load_state_dict(old)
    assert int(ours.step_count) == 2
    bad = ref.state_dict()
    bad['param_groups'][0]['weight_decay'] = 0.1
    with pytest.raises(ValueError):
        ours.load_state_dict(bad)

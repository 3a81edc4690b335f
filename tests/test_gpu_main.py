"""End-to-end run of the ``main.py``-compatible driver on synthetic mels: one epoch, checkpoint in the reference's
layout (loadable by ``torch.optim.Adam``), per-epoch sample dump, resume."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import main as M


def test_driver_trains_checkpoints_and_resumes(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    argv = ['--model', 'vqvae', '--dataset', 'ljspeech', '--dim', '64', '--z-dim', '512', '--batch-size', '4',
            '--synthetic-batches', '3', '--synthetic-frames', '64', '--log-interval', '1',
            '--sampledir', str(tmp_path / 'results')]
    assert M.main(argv + ['--epochs', '1']) == 0
    ck_path = tmp_path / 'models' / 'vqvae' / 'checkpoint_ljspeech_64_512.pth.tar'
    assert ck_path.exists()
    ck = torch.load(str(ck_path), map_location='cpu')
    assert set(ck) == {'epoch', 'arch', 'state_dict', 'optimizer'} and ck['epoch'] == 1 and ck['arch'] == 'vqvae'
    ref_keys = set(nsg.VQVAE(1, 64, 512).state_dict().keys())
    assert set(ck['state_dict'].keys()) == ref_keys and 'codebook.embedding.weight' in ref_keys
    # the optimizer entry is torch.optim.Adam's own format: three optimizer steps were taken
    probe = nsg.VQVAE(1, 64, 512)
    topt = torch.optim.Adam(probe.parameters(), lr=1e-3)
    topt.load_state_dict(ck['optimizer'])
    assert all(float(topt.state[p]['step']) == 3.0 for p in probe.parameters())
    assert any(float(topt.state[p]['exp_avg'].abs().max()) > 0 for p in probe.parameters())
    sample = tmp_path / 'results' / 'ljspeech' / 'reconstruction_vqvae_data_ljspeech_dim_64_z_dim_512_epoch_1.npy'
    rec = np.load(str(sample), allow_pickle=True)
    assert rec.shape == (4, 80, 64) and np.isfinite(rec).all() and np.abs(rec).max() <= 1.0      # tanh output
    # Griffin-Lim audio of the sample (main.py:164-197, row N3): last item of the batch, 63 hops, int16 RIFF
    from scipy.io import wavfile
    wav = tmp_path / 'results' / 'ljspeech' / ('audio_recon_vqvae_data_ljspeech_dim_64_z_dim_512_epoch_1_fftsize_1024'
                                                '_hopsize_256.wav')
    sr, data = wavfile.read(str(wav))
    assert sr == 22050 and data.dtype == np.int16 and data.shape == (256 * 63,)
    # resume: epoch 2 only, optimizer step count continues
    assert M.main(argv + ['--epochs', '2', '--resume']) == 0
    ck2 = torch.load(str(ck_path), map_location='cpu')
    assert ck2['epoch'] == 2
    assert float(next(iter(ck2['optimizer']['state'].values()))['step']) == 6.0
    assert not torch.equal(ck2['state_dict']['encoder.3.weight'], ck['state_dict']['encoder.3.weight'])

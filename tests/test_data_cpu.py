"""CPU tests of the input pipeline: the oracle restatement of the reference's collate_fn
against outputs of the UNMODIFIED reference (tests/golden/collate.npz, made by oracle/gen_golden_data.py), and the
product's host half (crop windows -> staging buffer, batch order, train.txt reader) against that oracle."""
import os

import numpy as np
import pytest
import torch

import neural_sound_generation_b200 as nsg  # noqa: F401
from neural_sound_generation_b200 import data as D
from oracle import data_ref as R
from oracle.gen_golden_data import CASES, make_items

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'collate.npz')


@pytest.mark.parametrize('name', sorted(CASES))
def test_oracle_collate_matches_reference_fixture(name):
    g = np.load(GOLD)
    hop, steps = int(g['hop_size']), int(g['max_time_steps'])
    frames, seed = CASES[name]
    items = make_items(frames, hop)
    mf = R.max_time_frames(steps, hop)
    assert mf == 42
    c, starts, lengths = R.collate_mel([it[1] for it in items], mf, np.random.RandomState(seed))
    assert c.dtype == np.float32 and np.array_equal(c, g[name + '/c'])               # bit-exact: it is data movement
    assert np.array_equal(lengths * hop, g[name + '/input_lengths'])                 # dataloader.py:366 counts audio samples
    # the audio side is cropped with the same window (dataloader.py:360-361): its checksum pins the drawn starts
    xs = sum(float(it[0][s * hop:(s + ln) * hop].astype(np.float64).sum()) for it, s, ln in zip(items, starts, lengths))
    assert abs(xs - float(g[name + '/x_sum'])) < 1e-6 * max(1.0, abs(xs))


def _mels(lengths, seed=0):
    return [np.random.RandomState(seed + i).rand(t, 80).astype(np.float32) for i, t in enumerate(lengths)]


@pytest.mark.parametrize('lengths', [[30, 50, 42, 61], [5], [43, 44, 200, 861, 42, 1]])
def test_host_fill_matches_oracle(lengths):
    mels = _mels(lengths)
    ld = D.DeviceMelLoader(D.ArrayMelDataset(mels), len(mels), device='cpu', max_frames=42, bucket=False)
    buf = np.full(ld._cap, np.nan, dtype=np.float32)
    b, tmax, total, starts = ld._fill(buf, list(range(len(mels))), np.random.RandomState(9))
    c, st, ln = R.collate_mel(mels, 42, np.random.RandomState(9))
    offs = buf[:ld._hdr].view(np.int32)
    assert (b, tmax, total) == (len(mels), c.shape[2], int(ln.sum())) and list(st) == starts
    assert np.array_equal(np.diff(offs[:b + 1]), ln)
    fr = buf[ld._hdr:ld._hdr + total * 80].reshape(total, 80)
    for i in range(b):                                     # what the device kernel pads and transposes
        assert np.array_equal(fr[offs[i]:offs[i + 1]].T, c[i, :, :ln[i]])
        assert not c[i, :, ln[i]:].any()
    # no cropping when max_frames is None (inference: whole utterances)
    ld2 = D.DeviceMelLoader(D.ArrayMelDataset(mels), len(mels), device='cpu', max_frames=None, bucket=False)
    buf2 = np.zeros(ld2._cap, dtype=np.float32)
    assert ld2._fill(buf2, list(range(len(mels))), np.random.RandomState(9))[1:] == (max(lengths), sum(lengths), [0] * len(mels))


def test_bucket_batches_properties():
    rng = np.random.RandomState(3)
    lengths = rng.randint(5, 900, size=1000)
    for bs, group in ((16, None), (32, 64), (7, 7)):
        o = D.bucket_batches(lengths, bs, group, rng=np.random.RandomState(1))
        assert sorted(o.tolist()) == list(range(1000))                              # a permutation (the reference's is not)
        assert np.array_equal(o, R.bucket_batches(lengths, bs, group, rng=np.random.RandomState(1)))
        g = group if group is not None else min(bs * 32, 1000) // bs * bs
        srt = np.argsort(lengths, kind='stable')
        ranks = np.empty(1000, dtype=np.int64)
        ranks[srt] = np.arange(1000)
        full = 1000 // g * g
        for s in range(0, full, bs):                       # every full batch comes from ONE group of the sorted order
            assert len(set((ranks[o[s:s + bs]] // g).tolist())) == 1
        assert set(o[full:].tolist()) == set(srt[full:].tolist())                   # the tail stays the tail
    assert not np.array_equal(D.bucket_batches(lengths, 16, rng=np.random.RandomState(1)),
                              D.bucket_batches(lengths, 16, rng=np.random.RandomState(2)))
    with pytest.raises(ValueError):
        D.bucket_batches(lengths, 16, 24)


def test_mel_npy_dataset_reads_train_txt(tmp_path):
    rows = []
    for i, t in enumerate([30, 50, 42, 61, 17, 43, 200, 5, 99, 70] * 4):
        np.save(tmp_path / ('mel-%d.npy' % i), np.full((t, 80), i, dtype=np.float32))
        rows.append('audio-%d.npy|mel-%d.npy|%d|some text' % (i, i, t * 256))
    (tmp_path / 'train.txt').write_text('\n'.join(rows) + '\n')
    tr, te = D.MelNpyDataset(str(tmp_path), train=True), D.MelNpyDataset(str(tmp_path), train=False)
    assert len(tr) == 38 and len(te) == 2                                           # test_size = 0.05 (dataloader.py:75)
    names = sorted(os.path.basename(p) for p in tr.paths + te.paths)
    assert names == sorted('mel-%d.npy' % i for i in range(40))
    from sklearn.model_selection import train_test_split
    want_tr, _ = train_test_split(np.arange(40), test_size=0.05, random_state=1234)
    assert [os.path.basename(p) for p in tr.paths] == ['mel-%d.npy' % i for i in want_tr]
    m = tr[0]
    assert m.shape == (tr.lengths[0], 80) and m.dtype == np.float32
    assert D.max_time_frames() == 42 and D.max_time_frames(None) is None and D.max_time_frames(8000, 200) == 40


def test_loader_needs_a_device():
    ld = D.DeviceMelLoader(D.ArrayMelDataset(_mels([30, 20])), 2, device='cpu')
    assert len(ld) == 1
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            next(iter(ld))

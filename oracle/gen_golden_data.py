"""Generate tests/golden/collate.npz by running the UNMODIFIED reference dataloader.collate_fn on CPU.

Run in the authoring container only (needs /root/reference):
    python -m oracle.gen_golden_data

/root/reference/src/dataloader.py imports matplotlib, librosa, lws, nnmnkwii and (through hparams_tacotron) tensorflow
at module level; none of them is on the collate path for raw-audio + mel conditioning (hparams.input_type = "raw",
cin_channels = 80), so they are replaced by empty stub modules (tensorflow's HParams by a plain attribute bag) and the
real collate_fn runs as is.  Inputs come from numpy RandomState streams that tests regenerate (``make_items``); the
fixture holds only reference OUTPUTS: c_batch (B, 80, T), input_lengths, and the hparams the crop depends on.
"""
import importlib.machinery
import os
import sys
import types

import numpy as np

REF_SRC = '/root/reference/src'
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')

CASES = {                      # name -> (frame counts, numpy seed for the crop draws)
    'short': ([30, 17, 42, 5, 41], 11),
    'mixed': ([30, 50, 42, 61, 17, 43, 200], 12),
    'long': ([100, 64, 43, 861], 13),
    'single': ([77], 14),
}


def make_items(frames, hop, n_mels=80):
    """(x, c, g) triples as the LJSpeech dataset yields them: raw audio (T * hop,), mel (T, n_mels), speaker id."""
    items = []
    for i, t in enumerate(frames):
        r = np.random.RandomState(1000 + 7 * t + i)
        items.append((r.rand(t * hop).astype(np.float32), r.rand(t, n_mels).astype(np.float32), 0))
    return items


def import_reference_dataloader():
    import torch          # noqa: F401  (the real modules first: torch inspects sys.modules specs on import)
    import torchvision    # noqa: F401

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__spec__ = importlib.machinery.ModuleSpec(name, None)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    class HParams(object):
        def __init__(self, **kw):
            self.__dict__.update(kw)

    stub('matplotlib', use=lambda *a, **k: None)
    stub('matplotlib.pyplot')
    stub('lws')
    lib = stub('librosa')
    lib.util = stub('librosa.util', pad_center=None, tiny=None)
    lib.filters = stub('librosa.filters', mel=None)
    stub('nnmnkwii')
    stub('nnmnkwii.preprocessing')
    stub('nnmnkwii.datasets', FileSourceDataset=type('FileSourceDataset', (), {}), FileDataSource=type('FileDataSource', (), {}))
    tf = stub('tensorflow')
    tf.contrib = types.SimpleNamespace(training=types.SimpleNamespace(HParams=HParams))
    sys.path.insert(0, REF_SRC)
    import dataloader     # noqa: E402
    return dataloader


def main():
    dl = import_reference_dataloader()
    hop = dl.audio.get_hop_size()
    out = {'hop_size': np.int64(hop), 'max_time_steps': np.int64(dl.hparams.max_time_steps)}
    for name, (frames, seed) in CASES.items():
        np.random.seed(seed)
        x, y, c, g, lengths = dl.collate_fn(make_items(frames, hop))
        assert g is None
        out[name + '/c'] = c.numpy()
        out[name + '/input_lengths'] = lengths.numpy()
        out[name + '/x_sum'] = np.float64(x.double().sum().item())      # (the audio side is cropped with the same window)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, 'collate.npz'), **out)
    print('wrote collate.npz:', {k: getattr(v, 'shape', v) for k, v in out.items()})


if __name__ == '__main__':
    main()

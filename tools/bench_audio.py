#!/usr/bin/env python
"""Row N3 timing: mel -> audio (nsg_mel_to_linear + nsg_griffin_lim, 60 iterations in one launch, + nsg_inv_preemphasis)
and audio -> mel (nsg_melspectrogram) for a 10 s utterance (861 frames, the config-5 length), CUDA events, against the
oracle restatement of the reference's numpy / librosa path on the host cores.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neural_sound_generation_b200 import audio as G      # noqa: E402
from oracle import audio_ref as R                        # noqa: E402


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = []
    for _ in range(n):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        ev.append((s, e))
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2]


def main():
    frames = 861
    rng = np.random.RandomState(0)
    mel = torch.tensor(rng.rand(80, frames).astype(np.float32), device='cuda')
    ang = torch.tensor(rng.rand(513, frames).astype(np.float32), device='cuda')
    wav = torch.tensor((0.1 * rng.randn(256 * (frames - 1))).astype(np.float32), device='cuda')
    out = dict(frames=frames, samples=256 * (frames - 1))
    out['inv_mel_spectrogram_ms'] = timed(lambda: G.inv_mel_spectrogram(mel, 22050, 1024, 256, 80, angles=ang))
    S = torch.rand(513, frames, device='cuda')
    for it in (0, 60):
        out['griffin_lim_%d_iters_ms' % it] = timed(lambda: G._griffin_lim(S, 1024, 256, angles=ang, n_iter=it))
    out['us_per_griffin_lim_iteration'] = 1e3 * (out['griffin_lim_60_iters_ms'] - out['griffin_lim_0_iters_ms']) / 60
    out['melspectrogram_ms'] = timed(lambda: G.melspectrogram(wav, 22050, 1024, 256, 80))
    out['stft_ms'] = timed(lambda: G._stft(wav, 1024, 256))
    out['inv_preemphasis_ms'] = timed(lambda: G.inv_preemphasis(wav, 0.97))
    t0 = time.perf_counter()
    R.inv_mel_spectrogram(mel.cpu().numpy().astype(np.float64), ang.cpu().numpy().astype(np.float64))
    out['cpu_oracle_inv_mel_spectrogram_ms'] = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    R.melspectrogram(wav.cpu().numpy().astype(np.float64))
    out['cpu_oracle_melspectrogram_ms'] = 1e3 * (time.perf_counter() - t0)
    out['audio_seconds_per_second'] = (out['samples'] / 22050.0) / (out['inv_mel_spectrogram_ms'] / 1e3)
    print(json.dumps(out))


if __name__ == '__main__':
    main()

"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's mel batch assembly.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package; the product path
(neural_sound_generation_b200/data.py + csrc/collate.cu) never does.

What is restated (numpy, bit-exact: it is copying, cropping, zero padding and a transpose):

* ``max_time_frames``   dataloader.py:340-346,356-358 -- hparams.max_time_steps (11000) rounded DOWN to a multiple of the
                        hop size (256) and divided by it: 42 frames with the reference's hparams_tacotron defaults.
* ``collate_mel``       dataloader.py:351-364 (crop: an utterance longer than max_time_frames keeps a window of exactly
                        that many frames starting at ``np.random.randint(0, len(c) - max_time_frames)`` -- the upper
                        bound is exclusive, so the last possible window is never chosen), :410-415 (zero padding at the
                        END of the time axis up to the longest item of the batch, then (B, T, 80) -> (B, 80, T)).
                        The draws are made in batch order and only for the items that are cropped, from numpy's legacy
                        global stream, which ``np.random.RandomState(seed)`` reproduces.
* ``bucket_batches``    dataloader.py:158-202 (PartialyRandomizedSimilarTimeLengthSampler): its INTENT -- sort by length,
                        shuffle inside groups of batch_group_size consecutive items, permute whole batches, shuffle the
                        tail.  The reference applies ``random.shuffle`` to views of a torch tensor, whose element swaps
                        alias under current PyTorch and duplicate / drop indices (observed here: 12 items -> indices
                        [2, 2, 11, 2, 3, 3, ...]); that defect is not reproduced.  Parity for the sampler is therefore
                        on its documented properties (a permutation; every batch comes from one length-sorted group),
                        not on a golden index list -- "parity unpinned" for this function only.

Pinned against the reference: tests/golden/collate.npz holds the outputs of the UNMODIFIED dataloader.collate_fn
(oracle/gen_golden_data.py imports it with its plotting / audio dependencies stubbed out).
"""
import numpy as np


def max_time_frames(max_time_steps=11000, hop_size=256, max_time_sec=None, sample_rate=22050):
    """dataloader.py:340-346 + :356-358; None = no cropping."""
    if max_time_sec is not None:
        max_time_steps = int(max_time_sec * sample_rate)
    if max_time_steps is None:
        return None
    max_steps = max_time_steps - max_time_steps % hop_size          # ensure_divisible(..., lower=True), :311-317
    return max_steps // hop_size


def collate_mel(mels, max_frames=42, rng=None):
    """mels: list of (T_i, n_mels) float arrays.  Returns (c_batch (B, n_mels, T) float32, starts (B,) int64,
    lengths (B,) int64) where starts[i] = first kept frame (0 when not cropped), lengths[i] = kept frames."""
    rng = rng if rng is not None else np.random
    kept, starts = [], []
    for c in mels:
        s = 0
        if max_frames is not None and len(c) > max_frames:          # :357 (len(x) > max_steps with len(x) = hop * len(c))
            s = int(rng.randint(0, len(c) - max_frames))             # :359
            c = c[s:s + max_frames, :]                               # :362
        kept.append(c)
        starts.append(s)
    max_len = max(len(c) for c in kept)                              # :411
    out = np.zeros((len(kept), max_len, kept[0].shape[1]), dtype=np.float32)
    for i, c in enumerate(kept):
        out[i, :len(c), :] = c                                       # _pad_2d, :67-70 / :412
    return (np.ascontiguousarray(out.transpose(0, 2, 1)),            # :415
            np.asarray(starts, dtype=np.int64), np.asarray([len(c) for c in kept], dtype=np.int64))


def bucket_batches(lengths, batch_size, batch_group_size=None, permutate=True, rng=None):
    """Index order of one epoch (dataloader.py:158-202, intent; see the module docstring)."""
    rng = rng if rng is not None else np.random.RandomState(0)
    lengths = np.asarray(lengths)
    order = np.argsort(lengths, kind='stable')                       # :166
    n = len(order)
    if batch_group_size is None:                                     # :169-172
        batch_group_size = min(batch_size * 32, n)
        if batch_group_size % batch_size != 0:
            batch_group_size -= batch_group_size % batch_size
    assert batch_group_size % batch_size == 0
    e = 0
    for i in range(n // batch_group_size):                           # :181-184
        s, e = i * batch_group_size, (i + 1) * batch_group_size
        order[s:e] = order[s:e][rng.permutation(e - s)]
    if permutate and e > 0:                                          # :187-190
        perm = rng.permutation(e // batch_size)
        order[:e] = order[:e].reshape(-1, batch_size)[perm, :].reshape(-1)
    if e < n:                                                        # :193-195
        order[e:] = order[e:][rng.permutation(n - e)]
    return order

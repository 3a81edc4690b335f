"""Input pipeline of the VQ-VAE train loop: the LJSpeech mel reader, the length-bucketing batch order and the collate
step of /root/reference/src/dataloader.py, rebuilt so that the batch is assembled ON THE DEVICE.

    reference                                                   here
    ---------                                                   ----
    _NPYDataSource.collect_files (dataloader.py:97-140)         MelNpyDataset: train.txt -> mel .npy paths + lengths, the
                                                                same train/test split (sklearn.train_test_split, 5 %)
    PartialyRandomizedSimilarTimeLengthSampler (:158-202)       bucket_batches: sort by length, shuffle inside groups,
                                                                permute whole batches (the sampler's intent; see below)
    collate_fn crop (:356-362), pad + transpose (:410-415)      DeviceMelLoader: the host only copies each utterance's kept
                                                                window into a pinned staging buffer; one H2D copy and one
                                                                kernel (csrc/collate.cu) produce c (B, 80, T) fp32 and
                                                                its bf16 twin on the device, double-buffered on a copy
                                                                stream so that batch i+1 is assembled under step i

The loader yields the reference's 5-tuples ``(x, y, c, g, input_lengths)`` with ``c`` already on the device (``x`` /
``y``, the raw-audio side that train_vqvae never reads -- train.py:108-116 -- are None), so ``train_vqvae`` /
``test_vqvae`` consume it unchanged.

The reference sampler applies ``random.shuffle`` to views of a torch tensor; under current PyTorch those element swaps
alias and the result is not a permutation (indices are duplicated and dropped).  ``bucket_batches`` implements what
the sampler documents -- that defect is not reproduced.
"""
import os
import threading
import queue

import numpy as np
import torch

from . import ops

MAX_TIME_STEPS, HOP_SIZE = 11000, 256        # hparams_tacotron.py:213 / :83 (the values dataloader.collate_fn reads)


def max_time_frames(max_time_steps=MAX_TIME_STEPS, hop_size=HOP_SIZE, max_time_sec=None, sample_rate=22050):
    """Longest mel window a training item keeps (dataloader.py:340-346,356-358): 42 frames with the defaults."""
    if max_time_sec is not None:
        max_time_steps = int(max_time_sec * sample_rate)
    if max_time_steps is None:
        return None
    return (max_time_steps - max_time_steps % hop_size) // hop_size


class MelNpyDataset(object):
    """The mel column of an LJSpeech ``train.txt`` (``audio.npy|mel.npy|timesteps|text[|speaker]``), split into train
    and test as the reference does (dataloader.py:73-140: ``train_test_split(indices, test_size=0.05,
    random_state=1234)``).  ``lengths`` are in FRAMES (the reference keeps audio time steps; the bucketing only needs
    their order)."""

    def __init__(self, data_root, train=True, test_size=0.05, random_state=1234, hop_size=HOP_SIZE):
        meta = os.path.join(data_root, 'train.txt')
        with open(meta, 'rb') as f:
            lines = [ln.decode('utf-8').split('|') for ln in f.readlines() if ln.strip()]
        if not lines or len(lines[0]) not in (4, 5):
            raise ValueError('%s: expected 4 or 5 |-separated columns per line' % meta)
        paths = [os.path.join(data_root, ln[1]) for ln in lines]
        steps = [int(ln[2]) for ln in lines]
        idx = np.arange(len(paths))
        if len(paths) > 1 and test_size:
            from sklearn.model_selection import train_test_split
            tr, te = train_test_split(idx, test_size=test_size, random_state=random_state)
            idx = tr if train else te
        self.paths = [paths[i] for i in idx]
        self.lengths = [max(1, steps[i] // hop_size) for i in idx]

    def __len__(self):
        return len(self.paths)

    def __getitem__(self, i):
        return np.load(self.paths[i])                       # (T, 80) float32, dataloader.py:142-143


class ArrayMelDataset(object):
    """In-memory list of (T_i, n_mels) arrays (tests, synthetic data)."""

    def __init__(self, mels):
        self.mels = [np.ascontiguousarray(m, dtype=np.float32) for m in mels]
        self.lengths = [len(m) for m in self.mels]

    def __len__(self):
        return len(self.mels)

    def __getitem__(self, i):
        return self.mels[i]


def bucket_batches(lengths, batch_size, batch_group_size=None, permutate=True, rng=None):
    """One epoch's index order (dataloader.py:158-202): items sorted by length, shuffled inside groups of
    ``batch_group_size`` (default min(32 batches, everything), a multiple of the batch size), whole batches permuted,
    the tail that does not fill a group shuffled on its own."""
    rng = rng if rng is not None else np.random.RandomState(0)
    order = np.argsort(np.asarray(lengths), kind='stable')
    n = len(order)
    if batch_group_size is None:
        batch_group_size = min(batch_size * 32, n)
        if batch_group_size % batch_size != 0:
            batch_group_size -= batch_group_size % batch_size
    if batch_group_size <= 0 or batch_group_size % batch_size != 0:
        raise ValueError('batch_group_size must be a positive multiple of the batch size')
    e = 0
    for i in range(n // batch_group_size):
        s, e = i * batch_group_size, (i + 1) * batch_group_size
        order[s:e] = order[s:e][rng.permutation(e - s)]
    if permutate and e > 0:
        perm = rng.permutation(e // batch_size)
        order[:e] = order[:e].reshape(-1, batch_size)[perm, :].reshape(-1)
    if e < n:
        order[e:] = order[e:][rng.permutation(n - e)]
    return order


class _Slot(object):
    __slots__ = ('pinned', 'dev', 'out32', 'out16', 'ready', 'free', 'avail')


class DeviceMelLoader(object):
    """Iterates ``(None, None, c, None, input_lengths)`` with ``c`` (B, n_mels, T) fp32 on ``device`` (``.c_bf16`` of the
    loader holds the bf16 twin of the batch last yielded when ``want_bf16``).

    Per batch the worker thread crops (dataloader.py:356-362: a window of ``max_frames`` frames at a random start for
    longer utterances) by copying only the kept rows into a pinned staging buffer ``[offsets | frames]``; the iterator
    issues ONE host-to-device copy and ONE kernel on a private copy stream and makes the consumer's stream wait for
    them, so batch i+1 is assembled while step i runs.  ``depth`` slots (staging + output buffers) rotate; a slot is
    refilled only after the consumer's stream has passed the point where it asked for the following batch -- the
    tensors of a batch are therefore valid until the NEXT batch is requested, like a DataLoader with ``depth - 1``
    prefetched batches.  ``drop_last`` as torch's DataLoader; ``bucket=False`` keeps dataset order; ``pad_to`` pads
    every batch to a fixed T instead of its longest item (one CUDA-graph shape)."""

    def __init__(self, dataset, batch_size, device='cuda', max_frames=42, seed=0, bucket=True, drop_last=False, depth=3,
                 n_mels=80, want_bf16=False, pad_to=None, hop_size=HOP_SIZE, shuffle=False):
        self.dataset, self.batch_size, self.device = dataset, int(batch_size), torch.device(device)
        self.shuffle = shuffle               # bucket=False only: a plain random order per epoch (the reference's test loader)
        self.hop_size = hop_size             # input_lengths are reported in audio samples, as dataloader.py:366 does
        self.max_frames, self.seed, self.bucket, self.drop_last = max_frames, seed, bucket, drop_last
        self.depth, self.n_mels, self.want_bf16 = max(2, int(depth)), n_mels, want_bf16
        self.pad_to = pad_to
        self.epoch = 0
        self.c_bf16 = None
        self.last_starts = None
        self._tcap = max_frames if max_frames is not None else max(dataset.lengths)
        if pad_to is not None:
            self._tcap = max(self._tcap, pad_to)
        self._hdr = (self.batch_size + 1 + 63) // 64 * 64                   # ints reserved for the offsets (256-B multiple)
        self._cap = self._hdr + self.batch_size * self._tcap * n_mels
        self._slots = None
        self._copy = None

    def __len__(self):
        n = len(self.dataset)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _batches(self):
        n = len(self.dataset)
        rng = np.random.RandomState(self.seed + 7919 * self.epoch)
        order = (bucket_batches(self.dataset.lengths, self.batch_size, rng=rng) if self.bucket
                 else (rng.permutation(n) if self.shuffle else np.arange(n)))
        stop = n - n % self.batch_size if self.drop_last else n
        return [order[s:s + self.batch_size] for s in range(0, stop, self.batch_size)], rng

    def _fill(self, buf, idx, rng):
        """Host half of collate_fn: crop windows -> [offsets | frames] in ``buf`` (a float32 numpy view of the pinned
        staging buffer).  Returns (B, longest kept item, total frames, starts)."""
        offs = buf[:self._hdr].view(np.int32)
        frames = buf[self._hdr:]
        pos, tmax, starts = 0, 0, []
        offs[0] = 0
        for j, i in enumerate(idx):
            c = self.dataset[int(i)]
            s = 0
            if self.max_frames is not None and len(c) > self.max_frames:
                s = int(rng.randint(0, len(c) - self.max_frames))           # dataloader.py:359 (upper bound exclusive)
                c = c[s:s + self.max_frames]
            k = len(c)
            frames[pos * self.n_mels:(pos + k) * self.n_mels] = np.asarray(c, dtype=np.float32).reshape(-1)
            pos += k
            offs[j + 1] = pos
            tmax = max(tmax, k)
            starts.append(s)
        return len(idx), tmax, pos, starts

    def _setup(self):
        if self._slots is not None:
            return
        self._slots = []
        nout = self.batch_size * self.n_mels * self._tcap
        for _ in range(self.depth):
            s = _Slot()
            s.pinned = torch.empty(self._cap, dtype=torch.float32).pin_memory()
            s.dev = torch.empty(self._cap, dtype=torch.float32, device=self.device)
            s.out32 = torch.empty(nout, dtype=torch.float32, device=self.device)
            s.out16 = torch.empty(nout, dtype=torch.bfloat16, device=self.device) if self.want_bf16 else None
            s.ready, s.free, s.avail = torch.cuda.Event(), None, threading.Event()
            s.avail.set()
            self._slots.append(s)
        self._copy = torch.cuda.Stream(device=self.device)

    def __iter__(self):
        self._setup()
        batches, rng = self._batches()
        self.epoch += 1
        q = queue.Queue(maxsize=self.depth)
        stop = threading.Event()

        def work():
            try:
                for k, idx in enumerate(batches):
                    slot = self._slots[k % self.depth]
                    slot.avail.wait()                       # the consumer has asked for the batch after this slot's last
                    slot.avail.clear()
                    if stop.is_set():
                        return
                    if slot.free is not None:
                        slot.free.synchronize()             # ... and its stream has got that far
                    q.put((slot, idx) + self._fill(slot.pinned.numpy(), idx, rng))
                q.put(None)
            except BaseException as e:                      # surfaced in the consumer
                q.put(e)

        th = threading.Thread(target=work, daemon=True)
        th.start()
        prev = None
        try:
            while True:
                item = q.get()
                if prev is not None:                        # the previous batch's slot may be refilled once the consumer's
                    prev.free = torch.cuda.Event()          # stream reaches this point
                    prev.free.record(torch.cuda.current_stream(self.device))
                    prev.avail.set()
                    prev = None
                if item is None:
                    break
                if isinstance(item, BaseException):
                    raise item
                slot, idx, b, tmax, total, starts = item
                t_out = self.pad_to if self.pad_to is not None else tmax
                used = self._hdr + total * self.n_mels
                n = b * self.n_mels * t_out
                c32 = slot.out32[:n].view(b, self.n_mels, t_out)
                c16 = slot.out16[:n].view(b, self.n_mels, t_out) if self.want_bf16 else None
                with torch.cuda.stream(self._copy):
                    slot.dev[:used].copy_(slot.pinned[:used], non_blocking=True)
                    ops.mel_collate(slot.dev[self._hdr:used].view(total, self.n_mels),
                                    slot.dev[:b + 1].view(torch.int32), self.n_mels, t_out, c32, c16)
                    slot.ready.record(self._copy)
                torch.cuda.current_stream(self.device).wait_event(slot.ready)
                self.c_bf16, self.last_starts = c16, starts
                prev = slot
                offs = slot.pinned[:b + 1].view(torch.int32)
                yield None, None, c32, None, (offs[1:] - offs[:-1]).to(torch.long) * self.hop_size
        finally:
            stop.set()
            for s in self._slots:
                s.avail.set()
            while th.is_alive():
                try:
                    q.get_nowait()
                except queue.Empty:
                    pass
                th.join(timeout=0.05)
            torch.cuda.current_stream(self.device).synchronize()
            for s in self._slots:
                s.free = None
                s.avail.set()

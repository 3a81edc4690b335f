"""W-tiled inference of ONE long utterance ("W-tiling of a single long utterance with a
2-latent-column halo"): the eval-mode ``encode`` / ``decode`` of /root/reference/src/models.py:188-196 on column tiles
of the mel spectrogram, one tile per call -- or per GPU -- with results identical to the untiled call.

Why it is exact.  Eval-mode BatchNorm is pointwise and the quantiser is per latent pixel, so the only coupling along
the time axis is the convolutions' finite receptive field: a latent column j depends on input columns
[4j - 11, 4j + 14] (two 4/2/1 convolutions under two ResBlocks with one 3 x 3 each), i.e. on at most 3 neighbouring
latent columns on either side, and an output column of the decoder on at most 4 latent columns on either side.  A tile
that carries that many extra columns ("halo") computes its own columns from exactly the numbers the untiled run uses,
in the same order (same taps, same kernels): the results agree bit for bit, which the tests assert.  At the true ends
of the utterance the tile's zero padding IS the reference's zero padding.

``encode_tiled`` / ``decode_tiled`` run the tiles one after the other on the current device;
``encode_decode_sharded`` gives rank r of a ``torch.distributed`` group the tiles r, r + world, ... and all-gathers the
pieces (NCCL on GPUs, gloo in the CPU tests of the partition logic), so N GPUs share one utterance.  Batch-1 inference of
configs[4] is latency-bound, so tiling pays for LONG inputs only; it is a capacity / latency tool,
not part of the benchmarked path.
"""
import torch

ENC_HALO = 3        # latent columns of context the encoder needs on either side of a tile
DEC_HALO = 4        # latent columns of context the decoder needs on either side of a tile


def tile_bounds(n_cols, n_tiles):
    """Split ``n_cols`` latent columns into ``n_tiles`` contiguous, near-equal, non-empty ranges [(lo, hi), ...]."""
    n_tiles = max(1, min(int(n_tiles), int(n_cols)))
    base, extra = divmod(int(n_cols), n_tiles)
    out, lo = [], 0
    for i in range(n_tiles):
        hi = lo + base + (1 if i < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def halo_range(lo, hi, n_cols, halo):
    """The tile's columns with ``halo`` columns of context, clipped to the tensor: (lo_h, hi_h, skip, keep) where the
    tile's own columns are [skip, skip + keep) of the haloed range."""
    lo_h, hi_h = max(0, lo - halo), min(n_cols, hi + halo)
    return lo_h, hi_h, lo - lo_h, hi - lo


@torch.no_grad()
def encode_tiled(model, x, n_tiles, tiles=None):
    """``model.encode(x)`` for x (B, C, H, W) computed on ``n_tiles`` column tiles.  ``tiles``: indices of the tiles to
    compute (default all); returns (latents (B, H/4, W_l) int64, [(lo, hi), ...]) with only the requested tiles'
    columns filled when ``tiles`` is given."""
    w_lat = x.shape[3] // 4
    bounds = tile_bounds(w_lat, n_tiles)
    out = torch.zeros((x.shape[0], x.shape[2] // 4, w_lat), dtype=torch.int64, device=x.device)
    for t in (range(len(bounds)) if tiles is None else tiles):
        lo, hi = bounds[t]
        lo_h, hi_h, skip, keep = halo_range(lo, hi, w_lat, ENC_HALO)
        # the last tile also takes the (W mod 4) trailing input columns the untiled convolution sees
        x_hi = x.shape[3] if hi_h == w_lat else 4 * hi_h
        lat = model.encode(x[:, :, :, 4 * lo_h:x_hi].contiguous())
        out[:, :, lo:hi] = lat[:, :, skip:skip + keep]
    return out, bounds


@torch.no_grad()
def decode_tiled(model, latents, n_tiles, tiles=None):
    """``model.decode(latents)`` for latents (B, h, w) computed on ``n_tiles`` column tiles -> (B, C, 4h, 4w)."""
    w_lat = latents.shape[2]
    bounds = tile_bounds(w_lat, n_tiles)
    out = None
    for t in (range(len(bounds)) if tiles is None else tiles):
        lo, hi = bounds[t]
        lo_h, hi_h, skip, keep = halo_range(lo, hi, w_lat, DEC_HALO)
        xt = model.decode(latents[:, :, lo_h:hi_h].contiguous())
        if out is None:
            out = torch.zeros((xt.shape[0], xt.shape[1], xt.shape[2], 4 * w_lat), dtype=xt.dtype, device=xt.device)
        out[:, :, :, 4 * lo:4 * hi] = xt[:, :, :, 4 * skip:4 * (skip + keep)]
    return out, bounds


def my_tiles(n_tiles, rank, world):
    """Tiles of rank ``rank``: r, r + world, ..."""
    return list(range(rank, n_tiles, world))


@torch.no_grad()
def encode_decode_sharded(model, x, n_tiles=None, group=None):
    """One utterance over the ranks of ``group``: every rank encodes its tiles, the latent pieces are summed across
    ranks (disjoint columns, so the sum is the concatenation), every rank decodes its tiles, the output pieces are summed
    likewise.  Returns (latents, x_tilde) on every rank, equal to ``model.encode(x)`` / ``model.decode(latents)``."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_tiles = n_tiles or world
    lat, _ = encode_tiled(model, x, n_tiles, tiles=my_tiles(min(n_tiles, x.shape[3] // 4), rank, world))
    if world > 1:
        dist.all_reduce(lat, op=dist.ReduceOp.SUM, group=group)
    xt, _ = decode_tiled(model, lat, n_tiles, tiles=my_tiles(min(n_tiles, lat.shape[2]), rank, world))
    if xt is None:          # more ranks than tiles: this rank contributes nothing
        h, w = lat.shape[1], lat.shape[2]
        xt = torch.zeros((lat.shape[0], x.shape[1], 4 * h, 4 * w), dtype=torch.float32, device=x.device)
    if world > 1:
        dist.all_reduce(xt, op=dist.ReduceOp.SUM, group=group)
    return lat, xt

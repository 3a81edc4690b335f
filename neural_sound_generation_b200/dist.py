"""Batch-sharded data parallelism for the VQ-VAE train step (one process per GPU, NCCL over NVLink / NVSwitch).

The reference is single-process (/root/reference/src/main.py:76), so the semantics are ours to define:

    an N-rank run on a global batch of N*B items  ==  the reference's single process on those N*B items.

Rank r trains items [r*B, (r+1)*B) (``shard_range``); model, codebook and optimiser state are replicated.  Per step:

* gradients: every rank's losses are means over its shard, so the global gradient is the MEAN of the rank gradients:
  all-reduce-sum of the flat gradient buffer, then ``optim.Adam.grad_scale = 1/N``.  The buffer is cut into two
  buckets in reverse layer order (``bucket_slices``): the decoder bucket is reduced on a communication stream as soon
  as the decoder's backward is complete and overlaps the bottleneck + encoder backward; codebook + encoder follow.
* ``sync_bn=True``: BatchNorm batch statistics ([sum | sum sq] forward, [sum g | sum g*xhat] backward; 1 KB of fp64 per
  layer and direction) are all-reduce-summed between the convolution epilogue that accumulates them and the
  normalisation kernel that consumes them, and the element count is multiplied by N -- train-mode BatchNorm then sees
  the global batch exactly as the reference does, and running statistics stay identical on every rank.  With
  ``sync_bn=False`` (default; standard DDP semantics) each rank normalises with its own shard's statistics; that
  is NOT equivalent to the reference on the global batch and is documented as such.
* the three loss scalars are averaged across ranks for logging.

All collectives are issued through ``torch.distributed`` (NCCL on GPUs; the host logic is backend-agnostic and is
tested with gloo on CPU) and are captured into the step's CUDA graph.
"""
import torch
import torch.distributed as dist

from .graph import GraphedTrainStep


def shard_range(rank, world, global_batch):
    """Items [lo, hi) of the global batch owned by ``rank``; the batch must divide evenly (equal shards keep
    mean-of-means == global mean)."""
    if global_batch % world != 0:
        raise ValueError('global batch %d does not divide over %d ranks' % (global_batch, world))
    per = global_batch // world
    return rank * per, (rank + 1) * per


def bucket_slices(names, offsets, total):
    """Cut the flat gradient buffer (parameters in ``named_parameters`` order: encoder.*, codebook.*, decoder.*) into
    the two all-reduce buckets: {'decoder': (lo, hi), 'rest': (lo, hi)}."""
    dec = [o for n, o in zip(names, offsets) if n.startswith('decoder.')]
    if not dec:
        return {'decoder': (total, total), 'rest': (0, total)}
    lo = min(dec)
    if any(o > lo and not n.startswith('decoder.') for n, o in zip(names, offsets)):
        raise ValueError('decoder parameters are not contiguous at the end of the flat buffer')
    return {'decoder': (lo, total), 'rest': (0, lo)}


def allreduce_sum_(t, group=None):
    """In-place all-reduce-sum on the current stream (no-op for a single rank)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def shutdown(step=None, timeout_s=20.0):
    """Orderly end of a data-parallel process: release ``step``'s graph, then destroy the process group; if NCCL does
    not return within ``timeout_s`` the process exits hard (all results must have been written by then)."""
    import os
    import sys
    import threading
    if step is not None and getattr(step, 'graph', None) is not None:
        step.close()
    if not (dist.is_available() and dist.is_initialized()):
        return
    sys.stdout.flush()
    sys.stderr.flush()
    t = threading.Timer(timeout_s, lambda: os._exit(0))
    t.daemon = True
    t.start()
    dist.destroy_process_group()
    t.cancel()


def allreduce_buckets_(flat, slices, order=('decoder', 'rest'), group=None):
    """All-reduce-sum every bucket of ``flat`` (1-D tensor) in ``order``; equals one all-reduce of the whole buffer."""
    for key in order:
        lo, hi = slices[key]
        if hi > lo:
            allreduce_sum_(flat[lo:hi], group)
    return flat


class DataParallel(object):
    """Per-model data-parallel state; attached as ``model.dp`` and consulted by the kernel schedule
    (``engine_tma.EngineTMA``) and by ``VQVAE.loss_and_grads``."""

    def __init__(self, sync_bn=False, overlap=True, group=None, peer_ok=True):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError('DataParallel needs an initialised torch.distributed process group')
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.sync_bn = bool(sync_bn)
        self.overlap = bool(overlap)
        self.comm = None            # communication stream (CUDA only)
        self.collectives = 0        # issued per step (bench.py reports it)
        self.loss_slots = None      # view of the flat gradient buffer's tail padding that carried the loss scalars
        # sync_bn over peer memory: the statistics are exchanged INSIDE the kernels that need them (csrc/conv_tma.cu,
        # PeerSync) instead of one 1 KB ncclAllReduce launch per layer and direction
        self.peer_areas = None      # every rank's exchange area as mapped on this GPU, or None (-> NCCL)
        self.peer_exchanges = 0     # exchange slots used so far in the current step
        if self.sync_bn and self.world > 1 and torch.cuda.is_available() and peer_ok and self.world <= 8:
            self._setup_peer_memory()

    def _setup_peer_memory(self):
        import os
        if os.environ.get('NSG_PEER_SYNC', '1') == '0':
            return
        try:
            import torch.distributed._symmetric_memory as symm
            from ._cabi import PEER_AREA_BYTES
            dev = torch.device('cuda', torch.cuda.current_device())
            area = symm.empty(PEER_AREA_BYTES, dtype=torch.uint8, device=dev)
            hdl = symm.rendezvous(area, self.group if self.group is not None else dist.group.WORLD)
            area.zero_()
            torch.cuda.synchronize(dev)
            dist.barrier(group=self.group)
            self._peer_keep = (area, hdl)
            self.peer_areas = [int(p) for p in hdl.buffer_ptrs]
        except Exception as e:              # no peer access between these GPUs: fall back to NCCL, loudly
            print('[nsg dist] peer-memory statistics unavailable (%s): sync-BN uses ncclAllReduce' % str(e)[:200])
            self.peer_areas = None

    def begin_step(self):
        self.peer_exchanges = 0

    def next_peer(self):
        """nsg_peer_sync of the next exchange of this step (all ranks issue the same sequence)."""
        from . import ops
        if self.peer_exchanges >= 32:
            raise RuntimeError('more than 32 statistic exchanges in one step')
        ps = ops.peer_sync(self.world, self.rank, self.peer_exchanges, self.peer_areas)
        self.peer_exchanges += 1
        return ps

    # -- BatchNorm statistics -------------------------------------------------------------------------------
    def reduce_stats(self, stats):
        """stats: fp64 [2*C] sums accumulated by a convolution epilogue on the current stream."""
        if self.peer_areas is not None:
            from . import ops
            ops.peer_allreduce_stats(stats, self.next_peer())
            return
        allreduce_sum_(stats, self.group)
        self.collectives += 1

    # -- gradient buckets -----------------------------------------------------------------------------------
    def _slices(self, flat):
        s = flat.get('buckets')
        if s is None:
            s = flat['buckets'] = bucket_slices(flat['names'], flat['offs'], flat['total'])
        return s

    def decoder_done(self, eng, ctx, G, flat, losses=None):
        """Called when the decoder's data-gradient chain has been issued on the current stream.  The decoder's
        weight-gradient images are folded and its bucket is all-reduced on the communication stream.  ``losses``
        (3 device scalars, already final after the forward pass) are copied into the unused padding at the end of
        the bucket so that they are summed across ranks by the same collective."""
        if not self.overlap:
            return
        if losses is not None:
            last = flat['offs'][-1] + flat['last_numel']
            if flat['total'] - last >= losses.numel():
                self.loss_slots = flat['grads'][flat['total'] - losses.numel():]
                self.loss_slots.copy_(losses)
        main = torch.cuda.current_stream()
        if self.comm is None:
            self.comm = torch.cuda.Stream()
        self.comm.wait_stream(main)
        if ctx.forked:
            self.comm.wait_stream(eng.side)         # decoder weight gradients issued so far on the side stream
        lo, hi = self._slices(flat)['decoder']
        with torch.cuda.stream(self.comm):
            eng.unpack_part(G, 'decoder', fused=ctx.fused)
            allreduce_sum_(flat['grads'][lo:hi], self.group)
        self.collectives += 1

    def finish(self, eng, ctx, G, flat):
        """End of backward: fold the remaining gradient images, all-reduce the remaining bucket(s), join streams."""
        sl = self._slices(flat)
        if self.overlap:
            eng.finish_backward(ctx, G, part='encoder')
            lo, hi = sl['rest']
            allreduce_sum_(flat['grads'][lo:hi], self.group)
            torch.cuda.current_stream().wait_stream(self.comm)
            self.collectives += 1
        else:
            eng.finish_backward(ctx, G)
            allreduce_sum_(flat['grads'], self.group)
            self.collectives += 1


class DataParallelStep(GraphedTrainStep):
    """``GraphedTrainStep`` for one rank of a data-parallel job: ``step = DataParallelStep(model, optimizer,
    (B_local, n_mels, T), beta, sync_bn=False); losses = step(c_local)``.

    ``losses`` are the GLOBAL-batch means [loss_recons, loss_vq, loss_commit] (averaged over ranks).  The model must
    run the bf16 TMA engine (``conv_impl = 2``); parameters must be identical on every rank at construction (they are
    broadcast from rank 0 to make sure)."""

    def __init__(self, model, optimizer, batch_shape, beta=1.0, sync_bn=False, overlap=True, group=None):
        self.dp = DataParallel(sync_bn=sync_bn, overlap=overlap, group=group)
        model.__dict__['dp'] = self.dp
        optimizer.grad_scale = 1.0 / self.dp.world
        flat = model.flatten_parameters()
        if self.dp.world > 1:
            dist.broadcast(flat['params'], src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            for _, b in model.named_buffers():
                dist.broadcast(b, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            # the in-kernel statistic exchange has a bounded wait: start the first (eager) iteration together
            torch.cuda.synchronize()
            dist.barrier(group=group)
        super(DataParallelStep, self).__init__(model, optimizer, batch_shape, beta)

    def _body(self, grad_hook):
        dp = self.dp
        dp.collectives = 0
        dp.loss_slots = None
        if self.model.__dict__.get('_grads_dirty', True):      # as GraphedTrainStep._body: Adam clears after use
            self.optimizer.zero_grad()
        losses, outs = self.model.loss_and_grads(self.c, beta=self.beta)
        if dp.world > 1:
            if dp.loss_slots is not None:
                # the loss scalars rode in the padding at the end of the decoder gradient bucket (see decoder_done);
                # read them before Adam clears the buffer
                losses = dp.loss_slots.double() * (1.0 / dp.world)
            else:
                allreduce_sum_(losses, dp.group)
                losses.mul_(1.0 / dp.world)
                dp.collectives += 1
        self.optimizer.step(zero_grads=True)
        self.collectives_per_step = dp.collectives
        return losses, outs

"""Data-parallel gradient reduction for the GAN step.

The reference scales with ``nn.DataParallel`` (trainer.py:70-72): one process, scatter
the batch, per-replica BatchNorm statistics, gradients summed on GPU 0.  The B200
design is one process per GPU with the same arithmetic: every rank runs the full
model on ``global_batch / N`` samples (per-replica BN statistics, like DataParallel)
and the gradients of the *batch-mean* losses (trainer.py:228-229,240-241,272) are
averaged with a bucketed all-reduce over NCCL / NVLink that overlaps the rest of the
backward pass:

* parameters are bucketed in reverse registration order (the order backward produces
  their gradients), ~``bucket_mb`` MB of fp32 per bucket; a parameter larger than a
  bucket (the 134 M-element filter of dcgan.py:51) is its own bucket;
* every bucket is ONE persistent flat fp32 buffer and every parameter owns a *slot*
  (a view) in it.  The weight-gradient kernels write straight into the slot
  (``b200conv3d.functional`` asks :func:`grad_slot`), autograd adopts that tensor as
  ``.grad`` without a copy, so a bucket is complete - in place, no ``torch.cat`` - the
  moment its last gradient is written; gradients produced any other way are copied into
  their slot by the post-accumulate hook (and ``.grad`` re-pointed at it);
* a complete bucket is all-reduced IN PLACE with ``ReduceOp.AVG`` on a side stream while
  backward keeps running: no scaling pass and no copy back - ``.grad`` already is the
  reduced buffer;
* :meth:`finish` (before ``optimizer.step()``) joins the side stream.

Several ``backward()`` calls between :meth:`begin` and :meth:`finish` are allowed (StyleGAN2Loss calls it two or three
times per optimizer step, stylegan.py:39-99): the reductions are then deferred to :meth:`finish`, because a bucket that
was averaged once must not be accumulated into again.  Every rank must produce gradients for the same set of
parameters in a pass (a bucket reduced on some ranks only would hang the collective).

With the ``gloo`` backend (CPU tests) the same code runs synchronously (sum, then divide).
"""
import torch
import torch.distributed as dist

_SLOTS = {}          # id(param) -> (reducer, index): consulted by b200conv3d.functional through grad_slot()


def grad_slot(p):
    """The fp32 tensor (shaped like ``p``) a gradient kernel should write ``p``'s gradient into during the current
    backward pass, or None.  Handed out ONCE per pass per parameter: a second gradient of the same parameter in one
    pass (two discriminator passes) must be accumulated by autograd, not overwrite the first."""
    ent = _SLOTS.get(id(p))
    if ent is None:
        return None
    red, i = ent
    if not red.armed or red.taken[i] or not p.requires_grad or p.grad is not None:
        return None
    red.taken[i] = True
    # a FRESH view object: autograd adopts a returned gradient without a copy only when nobody else holds that tensor
    # object (AccumulateGrad's use_count test) - the reducer keeps `slots[i]`, so the kernel gets its own alias
    return red.slots[i].view(red.slots[i].shape)


class GradReducer:
    def __init__(self, module, bucket_mb=32, group=None, overlap=True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in module.parameters()]
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.cuda = dev.type == "cuda"
        self.avg = self.cuda                    # NCCL has ReduceOp.AVG; gloo does not
        self.overlap = overlap
        cap = bucket_mb * (1 << 20)
        buckets, cur, size = [], [], 0
        for i in reversed(range(len(self.params))):
            n = self.params[i].numel() * 4
            if cur and size + n > cap:
                buckets.append(cur)
                cur, size = [], 0
            cur.append(i)
            size += n
        if cur:
            buckets.append(cur)
        self.buckets = buckets                  # lists of parameter indices (reverse order)
        self.flat, self.slots, self.bucket_of = [], [None] * len(self.params), {}
        for bi, b in enumerate(buckets):
            total = sum((self.params[i].numel() + 63) // 64 * 64 for i in b)      # 256-byte aligned slots
            flat = torch.zeros(total, dtype=torch.float32, device=dev)
            off = 0
            for i in b:
                n = self.params[i].numel()
                self.slots[i] = flat[off:off + n].view(self.params[i].shape)
                off += (n + 63) // 64 * 64
                self.bucket_of[i] = bi
            self.flat.append(flat)
        self.index = {id(p): i for i, p in enumerate(self.params)}
        self.taken = [False] * len(self.params)
        self.pending = [0] * len(buckets)
        self.inflight = {}                      # bucket index -> work handle
        self.deferred = False                   # a second backward() was seen: reduce everything in finish()
        self.armed = False
        self.copied = 0
        self.stream = torch.cuda.Stream(device=dev) if self.cuda else None
        self.handles = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        for i, p in enumerate(self.params):
            _SLOTS[id(p)] = (self, i)

    # ------------------------------------------------------------------ pass life cycle
    def begin(self, backwards=1):
        """Call after the parameters' requires_grad flags are set for this pass and ``zero_grad()``
        (trainer.py:209-211,255-258), before backward.  `backwards`: how many ``backward()`` calls of this pass reach
        every trainable parameter (StyleGAN2Loss.step_D: 2) - a bucket is reduced when the last one has written it."""
        self.inflight, self.deferred = {}, False
        self.taken = [False] * len(self.params)
        self.copied = 0                               # gradients that had to be copied into their slot this pass
        self.active = [[i for i in b if self.params[i].requires_grad] for b in self.buckets]
        self.pending = [len(a) * int(backwards) for a in self.active]
        self.armed = self.world > 1

    def _hook(self, p):
        if not self.armed:
            return
        i = self.index[id(p)]
        slot = self.slots[i]
        if p.grad is not None and p.grad.data_ptr() != slot.data_ptr():
            # a gradient that was not written into its slot (dtype / layout / produced by a stock torch op)
            slot.copy_(p.grad)
            p.grad = slot
            self.copied += 1
        bi = self.bucket_of[i]
        self.pending[bi] -= 1
        if self.pending[bi] == 0 and not self.deferred and self.overlap:
            self._launch(bi)
        elif self.pending[bi] < 0 and not self.deferred:
            # a further backward() in this pass: buckets already averaged cannot take more gradients -
            # undo nothing, but from now on everything waits for finish()
            self.deferred = True
            if self.inflight:
                raise RuntimeError(
                    "GradReducer: backward() was called again after gradient buckets of this pass were already "
                    "all-reduced; say begin(backwards=k), or construct the reducer with overlap=False, for losses that "
                    "call backward() several times per optimizer step (StyleGAN2Loss)")

    def _launch(self, bi):
        if not self.active[bi]:
            return
        flat = self.flat[bi]
        if self.cuda:
            ev = torch.cuda.current_stream().record_event()
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ev)
                work = dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        else:
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            work.wait()
            flat.div_(self.world)
        self.inflight[bi] = work

    def finish(self):
        """Join all bucket reductions; afterwards every ``p.grad`` holds the rank-averaged gradient."""
        if self.armed:
            for bi in range(len(self.buckets)):       # same order on every rank: the collectives stay matched
                if bi not in self.inflight and any(self.params[i].grad is not None for i in self.active[bi]):
                    self._launch(bi)
            for bi in sorted(self.inflight):
                self.inflight[bi].wait()              # CUDA: the current stream waits for NCCL
            self.inflight = {}
        self.armed = False

    def remove(self):
        for h in self.handles:
            h.remove()
        for p in self.params:
            _SLOTS.pop(id(p), None)

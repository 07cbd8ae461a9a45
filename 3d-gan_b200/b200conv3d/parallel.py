"""Data-parallel gradient reduction for the GAN step.

The reference scales with ``nn.DataParallel`` (trainer.py:70-72): one process, scatter
the batch, per-replica BatchNorm statistics, gradients summed on GPU 0.  The B200
design is one process per GPU with the same arithmetic: every rank runs the full
model on ``global_batch / N`` samples (per-replica BN statistics, like DataParallel)
and the gradients of the *batch-mean* losses (trainer.py:228-229,240-241,272) are
averaged with a bucketed all-reduce over NCCL / NVLink that overlaps the rest of the
backward pass:

* parameters are bucketed in reverse registration order (the order backward produces
  their gradients), ~``bucket_mb`` MB of fp32 per bucket;
* a post-accumulate-grad hook marks a parameter ready; when a bucket is complete it is
  flattened and all-reduced on a side stream while backward keeps running;
* :meth:`finish` (called before ``optimizer.step()``) joins the side stream and
  scatters the averaged gradients back.

Several ``backward()`` calls between :meth:`begin` and :meth:`finish` are allowed (StyleGAN2Loss): a bucket whose
gradients change after its reduce was launched is reduced again in :meth:`finish`.  Every rank must produce
gradients for the same set of parameters in a pass (a bucket reduced on some ranks only would hang the collective).

With the ``gloo`` backend (CPU tests) the same code runs synchronously.
"""
import torch
import torch.distributed as dist


class GradReducer:
    def __init__(self, module, bucket_mb=32, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in module.parameters()]
        self.buckets = []          # list of lists of params (reverse order)
        cur, size, cap = [], 0, bucket_mb * (1 << 20)
        for p in reversed(self.params):
            cur.append(p)
            size += p.numel() * 4
            if size >= cap:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self.bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for p in b:
                self.bucket_of[p] = bi
        self.pending = [0] * len(self.buckets)
        self.inflight = {}          # bucket index -> (grads, flat tensor, work handle)
        self.dirty = set()          # buckets whose gradients changed after their reduce was launched
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.cuda = dev.type == "cuda"
        self.stream = torch.cuda.Stream(device=dev) if self.cuda else None
        self.handles = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        self._reset()

    def _reset(self):
        for bi, b in enumerate(self.buckets):
            self.pending[bi] = sum(1 for p in b if p.requires_grad)

    def begin(self):
        """Call after the parameters' requires_grad flags are set for this pass
        (trainer.py:209-210,255-256) and before backward."""
        self.inflight = {}
        self.dirty = set()
        self._reset()

    def _hook(self, p):
        if self.world == 1:
            return
        bi = self.bucket_of[p]
        self.pending[bi] -= 1
        if self.pending[bi] == 0:
            self._launch(bi)
        elif self.pending[bi] < 0:
            # a further backward() between begin() and finish() (StyleGAN2Loss.step_D / step_G call it two or
            # three times per optimizer step, stylegan.py:39-99) accumulated into p.grad AFTER this bucket's
            # reduce was launched: the in-flight result is stale, finish() reduces the bucket again
            self.dirty.add(bi)

    def _launch(self, bi):
        grads = [p.grad for p in self.buckets[bi] if p.requires_grad and p.grad is not None]
        if not grads:
            return
        if self.cuda:
            ev = torch.cuda.current_stream().record_event()
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ev)
                flat = torch.cat([g.reshape(-1).float() for g in grads])
                for g in grads:
                    g.record_stream(self.stream)
                work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        else:
            flat = torch.cat([g.reshape(-1).float() for g in grads])
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.inflight[bi] = (grads, flat, work)

    def finish(self):
        """Join all bucket reductions, write back the rank-averaged gradients."""
        if self.world > 1:
            # buckets whose parameters never produced a gradient this pass are skipped;
            # buckets with missing hooks (unused params) are flushed here
            for bi in range(len(self.buckets)):
                if 0 < self.pending[bi] < sum(1 for p in self.buckets[bi] if p.requires_grad):
                    self._launch(bi)
            for bi in sorted(self.dirty):        # same order on every rank: the collectives stay matched
                stale = self.inflight.pop(bi, None)
                if stale is not None:
                    stale[2].wait()
                self._launch(bi)
            inv = 1.0 / self.world
            for bi in sorted(self.inflight):
                grads, flat, work = self.inflight[bi]
                work.wait()                       # CUDA: the current stream waits for NCCL
                if self.cuda:
                    flat.record_stream(torch.cuda.current_stream())
                flat.mul_(inv)
                for g, o in zip(grads, flat.split([g.numel() for g in grads])):
                    g.copy_(o.view_as(g))
            self.inflight = {}
            self.dirty = set()
        self._reset()

    def remove(self):
        for h in self.handles:
            h.remove()

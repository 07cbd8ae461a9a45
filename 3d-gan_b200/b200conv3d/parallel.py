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
* a complete bucket is averaged IN PLACE on a side stream while backward keeps running: no
  scaling pass and no copy back - ``.grad`` already is the reduced buffer.  On CUDA the
  buckets live in peer-addressable memory (CUDA IPC over NVLink / NVSwitch, :class:`PeerComm`)
  and the averaging is ONE small kernel of this library per rank (``b2c_peer_allreduce_avg``,
  csrc/peer.cu: two-shot over peer pointers, 128 blocks x 192 threads, no shared memory), sized
  to run on the SMs NEXT TO the resident blocks of the persistent convolution kernels.  NCCL's
  ring kernels need whole SMs: measured at N = 2 and 8, every conv kernel that overlapped one
  ran its displaced blocks after the others (+8 ms per step, the same at both N).  NCCL
  (``ReduceOp.AVG``) stays as the transport when peer memory cannot be set up
  (``B2C_REDUCE=nccl`` forces it);
* :meth:`finish` (before ``optimizer.step()``) joins the side stream.

Several ``backward()`` calls between :meth:`begin` and :meth:`finish` are allowed (StyleGAN2Loss calls it two or three
times per optimizer step, stylegan.py:39-99): the reductions are then deferred to :meth:`finish`, because a bucket that
was averaged once must not be accumulated into again.  Every rank must produce gradients for the same set of
parameters in a pass (a bucket reduced on some ranks only would hang the collective).

With the ``gloo`` backend (CPU tests) the same code runs synchronously (sum, then divide).
"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

from . import _lib

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


class _DevMem:
    """A raw device allocation presented through __cuda_array_interface__ (torch.as_tensor wraps it without a copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


class PeerComm:
    """Peer-memory plumbing of one process group on one node: flag blocks and gradient buckets allocated by
    ``b2c_peer_alloc``, their IPC handles exchanged through the group, every peer's copy mapped with
    ``b2c_peer_open``.  All averaging calls of the process run on ONE side stream in program order (they share the
    flag blocks; include/b2c.h), with the same sequence numbers on every rank."""

    _by_group = {}
    CTAS = int(os.environ.get("B2C_PEER_CTAS", "128"))

    @classmethod
    def get(cls, group, dev):
        """The communicator of `group`, or None when peer memory cannot be used (decided collectively, so that every
        rank picks the same transport)."""
        key = id(group) if group is not None else 0
        if key not in cls._by_group:
            comm = None
            if os.environ.get("B2C_REDUCE", "peer") != "nccl":
                try:
                    comm = cls(group, dev)
                    comm.self_test()
                except Exception as e:          # noqa: BLE001 - any failure means "use NCCL", on every rank
                    print("b200conv3d.parallel: peer-memory transport unavailable on rank %d (%s: %s)"
                          % (dist.get_rank(group), type(e).__name__, e), file=sys.stderr)
                    comm = None
            ok = torch.tensor([1 if comm is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                if comm is not None or os.environ.get("B2C_REDUCE", "peer") != "nccl":
                    if dist.get_rank(group) == 0:
                        print("b200conv3d.parallel: gradient averaging falls back to NCCL all-reduce", file=sys.stderr)
                comm = None
            cls._by_group[key] = comm
        return cls._by_group[key]

    def __init__(self, group, dev):
        self.group, self.dev = group, dev
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 8:
            raise RuntimeError("peer transport covers one node (<= 8 ranks)")
        self.lib = _lib.lib()
        self.seq = 0
        # high priority: when a kernel with many short blocks (LayerNorm passes) fills the SMs, the averaging kernel's
        # blocks take the next free slots instead of queueing behind the whole grid
        self.stream = torch.cuda.Stream(device=dev, priority=-1)
        self.keep = []                               # local allocations (freed with the process)
        self.flags = self.shared(160 * 16 * 4)       # B2C_PEER_FLAG_BYTES

    def shared(self, nbytes):
        """Allocate `nbytes` on every rank; returns the ctypes array of the `world` device addresses (own at [rank])."""
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.b2c_peer_alloc(nbytes, C.byref(ptr), handle), "peer_alloc")
            self.keep.append(ptr.value)
            handles = [None] * self.world
            dist.all_gather_object(handles, handle.raw, group=self.group)
            ptrs = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    ptrs.append(ptr.value)
                else:
                    q = C.c_void_p()
                    _lib.check(self.lib.b2c_peer_open(h, C.byref(q)), "peer_open")
                    ptrs.append(q.value)
        return (C.c_void_p * self.world)(*ptrs)

    def bucket(self, nfloats):
        """A zero-filled fp32 bucket of `nfloats` on every rank: (local tensor, address table)."""
        table = self.shared(max(int(nfloats), 4) * 4)
        return torch.as_tensor(_DevMem(table[self.rank], nfloats), device=self.dev), table

    def average(self, table, count):
        """Enqueue the in-place average of the bucket `table` on the communicator's stream."""
        self.seq += 1
        _lib.check(self.lib.b2c_peer_allreduce_avg(table, self.flags, self.rank, self.world, int(count), self.seq,
                                                   self.CTAS, C.c_void_p(self.stream.cuda_stream)), "peer_allreduce_avg")

    def self_test(self):
        """One small average end to end (flags, NVLink loads and stores, the tail path) before anything relies on it."""
        n = 4 * 192 * 6 * 3 + 3
        t, table = self.bucket(n)
        t.copy_(torch.arange(n, device=self.dev, dtype=torch.float32) * (self.rank + 1))
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        self.average(table, n)
        self.stream.synchronize()
        want = torch.arange(n, device=self.dev, dtype=torch.float32) * ((self.world + 1) / 2.0)
        if self.lib.b2c_peer_error() != 0 or not torch.allclose(t, want, rtol=1e-6, atol=0):
            raise RuntimeError("peer all-reduce self test failed (error flag %d)" % self.lib.b2c_peer_error())


class GradReducer:
    def __init__(self, module, bucket_mb=32, group=None, overlap=True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in module.parameters()]
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.cuda = dev.type == "cuda"
        self.avg = self.cuda                    # NCCL has ReduceOp.AVG; gloo does not
        self.overlap = overlap
        self.comm = PeerComm.get(group, dev) if self.cuda and self.world > 1 else None
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
        self.flat, self.tables, self.slots, self.bucket_of = [], [], [None] * len(self.params), {}
        for bi, b in enumerate(buckets):
            total = sum((self.params[i].numel() + 63) // 64 * 64 for i in b)      # 256-byte aligned slots
            if self.comm is not None:
                flat, table = self.comm.bucket(total)
                self.tables.append(table)
            else:
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
        self.done = {}                          # bucket index -> event on the side stream (peer transport)
        self.order = []                         # buckets in the order their reductions were enqueued this pass
        self.deferred = False                   # a second backward() was seen: reduce everything in finish()
        self.armed = False
        self.copied = 0
        self.stream = (self.comm.stream if self.comm is not None else torch.cuda.Stream(device=dev)) if self.cuda else None
        self.handles = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        for i, p in enumerate(self.params):
            _SLOTS[id(p)] = (self, i)

    # ------------------------------------------------------------------ pass life cycle
    def begin(self, backwards=1):
        """Call after the parameters' requires_grad flags are set for this pass and ``zero_grad()``
        (trainer.py:209-211,255-258), before backward.  `backwards`: how many ``backward()`` calls of this pass reach
        every trainable parameter (StyleGAN2Loss.step_D: 2) - a bucket is reduced when the last one has written it."""
        self.inflight, self.deferred = {}, False
        self.done, self.order = {}, []
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
        if self.comm is not None:
            self.stream.wait_event(torch.cuda.current_stream().record_event())
            self.comm.average(self.tables[bi], flat.numel())
            self.done[bi] = self.stream.record_event()
            work = None
        elif self.cuda:
            ev = torch.cuda.current_stream().record_event()
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ev)
                work = dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        else:
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            work.wait()
            flat.div_(self.world)
        self.inflight[bi] = work
        self.order.append(bi)

    def _launch_rest(self):
        for bi in range(len(self.buckets)):           # same order on every rank: the collectives stay matched
            if bi not in self.inflight and any(self.params[i].grad is not None for i in self.active[bi]):
                self._launch(bi)

    def step(self, optimizer):
        """``finish()`` + ``optimizer.step()`` with the update running bucket by bucket as the averages arrive: the
        optimizer is stepped on the parameters of one bucket at a time, each after the event of that bucket's
        reduction, so only the update of the LAST bucket waits for the last reduction (the tail that ``finish()``
        exposes: 0.5-1 ms per pass for the 134 MB LayerNorm affine gradients of dcgan.py:161-163, produced last).
        Per-parameter optimizers only (Adam / SGD ...: every parameter's update depends on its own state); the
        optimizer must hold this module's parameters in ONE group, otherwise this is finish() then step()."""
        if not self.armed:
            optimizer.step()
            return
        groups = optimizer.param_groups
        mine = {id(p) for p in self.params}
        if (self.comm is None or len(groups) != 1 or any(id(p) not in mine for p in groups[0]["params"])):
            self.finish()
            optimizer.step()
            return
        self._launch_rest()
        cur = torch.cuda.current_stream()
        listed = {id(p) for p in groups[0]["params"]}
        try:
            for bi in self.order:
                cur.wait_event(self.done[bi])
                ps = [self.params[i] for i in self.active[bi] if id(self.params[i]) in listed]
                if ps:
                    optimizer.param_groups = [dict(groups[0], params=ps)]
                    optimizer.step()
            rest = [p for p in groups[0]["params"] if p.grad is not None and
                    self.bucket_of[self.index[id(p)]] not in self.done]
            if rest:                                  # parameters whose bucket was not reduced (no gradient on any rank)
                optimizer.param_groups = [dict(groups[0], params=rest)]
                optimizer.step()
        finally:
            optimizer.param_groups = groups
        cur.wait_stream(self.stream)
        self.inflight, self.armed = {}, False

    def finish(self):
        """Join all bucket reductions; afterwards every ``p.grad`` holds the rank-averaged gradient."""
        if self.armed:
            self._launch_rest()
            if self.comm is not None:
                torch.cuda.current_stream().wait_stream(self.stream)
            else:
                for bi in sorted(self.inflight):
                    self.inflight[bi].wait()          # CUDA: the current stream waits for NCCL
            self.inflight = {}
        self.armed = False

    def remove(self):
        for h in self.handles:
            h.remove()
        for p in self.params:
            _SLOTS.pop(id(p), None)

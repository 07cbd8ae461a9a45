"""World-size-2 gloo test (CPU) of the data-parallel host logic
(b200conv3d.parallel.GradReducer + b200conv3d.train.GanStep):
rank-averaged gradients must equal the gradients of the concatenated global batch,
frozen parameters must not take part, and both ranks must step identically."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _nets():
    torch.manual_seed(0)
    g = nn.Sequential(nn.ConvTranspose3d(4, 3, 4, 1), nn.ReLU(), nn.ConvTranspose3d(3, 1, 4, 2, 1), nn.Tanh())
    d = nn.Sequential(nn.Conv3d(1, 3, 4, 2, 1, bias=False), nn.LeakyReLU(0.2), nn.Conv3d(3, 1, 4, 1, 0, bias=False))
    return g, d


def _worker(rank, world, port, q):
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200conv3d.parallel import GradReducer
    from b200conv3d.train import GanStep
    g, d = _nets()
    rg, rd = GradReducer(g, bucket_mb=0), GradReducer(d, bucket_mb=0)     # one bucket per parameter
    step = GanStep(g, d, z_size=4, autocast_dtype=None, reducer_g=rg, reducer_d=rd)
    torch.manual_seed(100)
    reals = torch.rand(4, 1, 8, 8, 8) * 2 - 1
    zs = torch.randn(4, 4, 1, 1, 1)
    lo, hi = rank * 2, rank * 2 + 2
    step.noise = lambda n: zs[lo:hi]
    step.D_step(reals[lo:hi])
    gd = [p.grad.clone() for p in d.parameters()]
    assert all(p.grad is None for p in g.parameters())          # G frozen during D_step
    step.G_step(2)
    gg = [p.grad.clone() for p in g.parameters()]
    npy = lambda ts: [t.detach().numpy().copy() for t in ts]          # by value: the worker exits before the read
    q.put((rank, npy(gd), npy(gg), npy(list(d.parameters()) + list(g.parameters()))))
    dist.barrier()
    dist.destroy_process_group()


def test_grad_reducer_world2_matches_global_batch():
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    from b200conv3d.train import GanStep
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    res = [(r[0],) + tuple([torch.from_numpy(a) for a in part] for part in r[1:]) for r in res]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # single-process run on the global batch of 4 (losses are batch means -> same gradients)
    g, d = _nets()
    step = GanStep(g, d, z_size=4, autocast_dtype=None)
    torch.manual_seed(100)
    reals = torch.rand(4, 1, 8, 8, 8) * 2 - 1
    zs = torch.randn(4, 4, 1, 1, 1)
    step.noise = lambda n: zs
    step.D_step(reals)
    gd = [p.grad.clone() for p in d.parameters()]
    step.G_step(4)
    gg = [p.grad.clone() for p in g.parameters()]
    for r in res:
        for a, b in zip(r[1], gd):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-6)
        for a, b in zip(r[2], gg):
            torch.testing.assert_close(a, b, rtol=1e-3, atol=1e-6)   # D differs by one Adam step only through fp noise
    for a, b in zip(res[0][3], res[1][3]):
        torch.testing.assert_close(a, b, rtol=0, atol=0)             # replicas stay bit-identical


def _worker_multi_backward(rank, world, port, q):
    """Three backward() calls between begin() and finish(), as StyleGAN2Loss.step_D / step_G do
    (stylegan.py:39-99): the rank-averaged gradient must be the average of the SUM of all three."""
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200conv3d.parallel import GradReducer
    _, d = _nets()
    red = GradReducer(d, bucket_mb=0)
    torch.manual_seed(7)
    x = torch.rand(4, 1, 8, 8, 8)[rank * 2:rank * 2 + 2]
    for prm in d.parameters():
        prm.grad = None
    red.begin(backwards=3)
    d(x).mean().backward()
    (d(x * 0.5) ** 2).mean().backward()
    (3.0 * d(-x)).mean().backward()
    red.finish()
    q.put((rank, [p.grad.detach().numpy().copy() for p in d.parameters()]))
    dist.barrier()
    dist.destroy_process_group()


def test_grad_reducer_world2_several_backward_calls_per_step():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker_multi_backward, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    _, d = _nets()
    torch.manual_seed(7)
    x = torch.rand(4, 1, 8, 8, 8)
    d(x).mean().backward()
    (d(x * 0.5) ** 2).mean().backward()
    (3.0 * d(-x)).mean().backward()
    for _, grads in res:
        for a, p in zip(grads, d.parameters()):
            torch.testing.assert_close(torch.from_numpy(a), p.grad, rtol=1e-4, atol=1e-6)


def _b200_nets():
    import b200conv3d as B
    torch.manual_seed(0)
    m = B.nn
    g = nn.Sequential(m.ConvTranspose3d(4, 8, 4, 1), m.BatchNorm3d(8), m.ReLU(True), m.ConvTranspose3d(8, 1, 4, 2, 1), m.Tanh())
    d = nn.Sequential(m.Conv3d(1, 8, 4, 2, 1, bias=False), m.LayerNorm([8, 4, 4, 4]), m.LeakyReLU(0.2, inplace=True),
                      m.Conv3d(8, 1, 4, 1, 0, bias=False))
    B.fuse(g), B.fuse(d)
    return g, d


def _worker_slots(rank, world, port, q, oracle):
    """The B200 layers (bound to the host oracle library) write their weight gradients straight into the all-reduce
    bucket slots: after the step every .grad IS its slot (no copy), and equals the global-batch gradient."""
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    from b200conv3d import _lib
    _lib.use_library(oracle, host=True)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200conv3d.parallel import GradReducer
    from b200conv3d.train import GanStep
    g, d = _b200_nets()
    rg, rd = GradReducer(g), GradReducer(d, bucket_mb=0)
    step = GanStep(g, d, z_size=4, autocast_dtype=None, reducer_g=rg, reducer_d=rd, batch_real_fake=False)
    torch.manual_seed(100)
    reals = torch.rand(4, 1, 8, 8, 8) * 2 - 1
    zs = torch.randn(4, 4, 1, 1, 1)
    lo, hi = rank * 2, rank * 2 + 2
    step.noise = lambda n: zs[lo:hi]
    step.D_step(reals[lo:hi])
    copied_d = rd.copied
    adopted_d = [p.grad.data_ptr() == rd.slots[i].data_ptr() for i, p in enumerate(d.parameters())]
    gd = [p.grad.clone() for p in d.parameters()]
    step.G_step(2)
    # G: the two filters were WRITTEN into their slots by the weight-gradient kernel and adopted by autograd without a
    # copy; only the two biases and the BatchNorm affine pair (tiny) go through the hook's copy.  D ran D(real) and
    # D(fake) as two passes here (CPU), so autograd summed two contributions per parameter before accumulating.
    assert rg.copied == 4 and copied_d <= 4, (copied_d, rg.copied)
    adopted_g = [p.grad.data_ptr() == rg.slots[i].data_ptr() for i, p in enumerate(g.parameters())]
    gg = [p.grad.clone() for p in g.parameters()]
    npy = lambda ts: [t.detach().numpy().copy() for t in ts]
    q.put((rank, npy(gd), npy(gg), adopted_d, adopted_g))
    dist.barrier()
    dist.destroy_process_group()


def test_grad_reducer_world2_b200_layers_write_into_bucket_slots(oracle_path):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 33500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker_slots, args=(r, 2, port, q, oracle_path)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(res[0][3]) and all(res[0][4]), (res[0][3], res[0][4])     # every gradient lives in its bucket slot
    # single process on the global batch, same layers through the same oracle library (in a subprocess-free way:
    # stock torch layers with the same weights give the reference gradients)
    torch.manual_seed(0)
    g = nn.Sequential(nn.ConvTranspose3d(4, 8, 4, 1), nn.BatchNorm3d(8), nn.ReLU(True), nn.ConvTranspose3d(8, 1, 4, 2, 1), nn.Tanh())
    d = nn.Sequential(nn.Conv3d(1, 8, 4, 2, 1, bias=False), nn.LayerNorm([8, 4, 4, 4]), nn.LeakyReLU(0.2, inplace=True),
                      nn.Conv3d(8, 1, 4, 1, 0, bias=False))
    sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
    from b200conv3d.train import GanStep
    torch.manual_seed(100)
    reals = torch.rand(4, 1, 8, 8, 8) * 2 - 1
    zs = torch.randn(4, 4, 1, 1, 1)
    # per-replica BatchNorm statistics (DataParallel semantics): average the two half-batch runs
    acc_d, acc_g = None, None
    import copy
    for r in range(2):
        g2, d2 = copy.deepcopy(g), copy.deepcopy(d)
        st = GanStep(g2, d2, z_size=4, autocast_dtype=None, batch_real_fake=False)
        st.noise = lambda n, r=r: zs[r * 2:r * 2 + 2]
        st.D_step(reals[r * 2:r * 2 + 2])
        gd = [p.grad.clone() for p in d2.parameters()]
        acc_d = gd if acc_d is None else [a + b for a, b in zip(acc_d, gd)]
    for a, b in zip(res[0][1], acc_d):
        torch.testing.assert_close(torch.from_numpy(a), b / 2, rtol=1e-4, atol=1e-6)
    for a, b in zip(res[0][1], res[1][1]):
        np_equal = (a == b).all()
        assert np_equal                                                    # both ranks hold the same averaged gradient

"""Gradient averaging over peer memory (include/b2c.h b2c_peer_*; csrc/peer.cu; the reference's counterpart is the
gradient reduce of nn.DataParallel, trainer.py:70-72).

CPU: the host stand-in (one call per "rank" over buffers of one process) gives the mean of the buffers for every
world size / count, every buffer ends identical, and elements beyond `count` are untouched.  GPU (one device): the
kernel with its `world` ranks played by `world` streams of one process - flags, slices, tail - against the host
stand-in bit for bit, several calls in a row on the same flag blocks, plus the IPC handle round trip that a second
process would do.  The N-process path itself is covered by `bench.py --gpus N` (tools/scale_round2.sh) and, on
a box with two GPUs, by tests/test_gpu_multi.py."""
import ctypes as C

import pytest
import torch

from abi import Abi

COUNTS = [1, 3, 4, 7, 1000, 4 * 192 * 6 + 5, 300001]


def _tables(bufs, flags):
    w = len(bufs)
    return (C.c_void_p * w)(*[b.data_ptr() for b in bufs]), (C.c_void_p * w)(*[f.data_ptr() for f in flags])


def _host_average(abi, bufs, count, seq=1):
    flags = [torch.zeros(160 * 16, dtype=torch.int32) for _ in bufs]
    tb, tf = _tables(bufs, flags)
    for r in range(len(bufs)):
        abi._ck(abi.lib.b2c_peer_allreduce_avg(tb, tf, r, len(bufs), count, seq, 8, None), "peer_allreduce_avg")


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("count", COUNTS)
def test_host_standin_averages(oracle_path, world, count):
    cpu = Abi(oracle_path, "cpu")
    g = torch.Generator().manual_seed(world * 1000 + count)
    src = [torch.randn(count + 5, generator=g) for _ in range(world)]
    bufs = [s.clone() for s in src]
    _host_average(cpu, bufs, count)
    want = torch.stack([s[:count] for s in src]).double().mean(0)
    for r in range(world):
        torch.testing.assert_close(bufs[r][:count].double(), want, rtol=1e-6, atol=1e-6)
        assert torch.equal(bufs[r][:count], bufs[0][:count])
        assert torch.equal(bufs[r][count:], src[r][count:])          # nothing beyond `count` is touched


def test_invalid_arguments(product_path, oracle_path):
    for path in (product_path, oracle_path):
        lib = Abi(path, "cpu").lib
        one = (C.c_void_p * 1)(16)
        assert lib.b2c_peer_allreduce_avg(None, None, 0, 2, 4, 1, 8, None) == -1
        assert lib.b2c_peer_allreduce_avg(one, one, 0, 9, 4, 1, 8, None) == -1      # world > B2C_PEER_MAX_WORLD
        assert lib.b2c_peer_allreduce_avg(one, one, 1, 1, 4, 1, 8, None) == -1      # rank >= world
        assert lib.b2c_peer_allreduce_avg(one, one, 0, 1, 4, 0, 8, None) == -1      # seq starts at 1
        assert lib.b2c_peer_allreduce_avg(one, one, 0, 1, 4, 1, 161, None) == -1     # ctas > B2C_PEER_MAX_CTAS
        assert lib.b2c_peer_allreduce_avg(one, one, 0, 1, 4, 1, 8, None) == 0       # world 1: nothing to do
        assert lib.b2c_peer_alloc(0, None, None) == -1
        assert lib.b2c_peer_open(None, None) == -1
        assert lib.b2c_peer_close(None) == -1 and lib.b2c_peer_free(None) == -1


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_gpu_kernel_matches_host_standin(product_path, oracle_path, world):
    gpu, cpu = Abi(product_path, "cuda:0"), Abi(oracle_path, "cpu")
    dev = torch.device("cuda:0")
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    flags = [torch.zeros(160 * 16, dtype=torch.int32, device=dev) for _ in range(world)]
    g = torch.Generator().manual_seed(world)
    seq = 0
    for count in COUNTS + [3_000_003]:
        src = [torch.randn(count + 5, generator=g) for _ in range(world)]
        want = [s.clone() for s in src]
        _host_average(cpu, want, count)
        bufs = [s.to(dev) for s in src]
        tb, tf = _tables(bufs, flags)
        torch.cuda.synchronize()
        seq += 1                                     # the flag blocks are never reset: values only grow
        for r in range(world):                       # every "rank" is a stream; the kernels meet through the flags
            gpu._ck(gpu.lib.b2c_peer_allreduce_avg(tb, tf, r, world, count, seq, 8, C.c_void_p(streams[r].cuda_stream)),
                    "peer_allreduce_avg")
        torch.cuda.synchronize()
        assert gpu.lib.b2c_peer_error() == 0
        for r in range(world):
            assert torch.equal(bufs[r].cpu(), want[r]), (world, count, r)


@pytest.mark.gpu
def test_gpu_alloc_exports_a_handle(product_path):
    lib = Abi(product_path, "cuda:0").lib
    torch.cuda.init()
    ptr, handle = C.c_void_p(), C.create_string_buffer(64)
    assert lib.b2c_peer_alloc(4096, C.byref(ptr), handle) == 0 and ptr.value
    assert any(handle.raw)
    q = C.c_void_p()
    assert lib.b2c_peer_open(handle.raw, C.byref(q)) != 0      # a handle cannot be opened by the process that made it
    assert lib.b2c_peer_free(ptr) == 0

"""One process driving every GPU - the counterparts of ``test_nccl.py`` and ``test_torch_cuda_stream.py``.

* ``collectives()``: all_gather / all_reduce / reduce_scatter of N=128 fp32 per GPU, checked against CPU results
  exactly as the reference does (``test_nccl.py:10,18-19,23-27,30-38``) - but on this framework's peer-memory kernels
  (direct peer access instead of IPC, since everything lives in one process).
* ``streams()``: ONE stream PER DEVICE (the reference creates all of its streams on device 0, ``test_torch_cuda_
  stream.py:6``), ten matmuls ``[128,64] x [64,256]`` per GPU on that device's stream with the tcgen05 GEMM
  (``:31-35``), then the streamed all-reduce of the outputs the reference left as a TODO (``:36-37``), checked
  against a CPU matmul.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from distributed_llm_code_samples_b200.ops import native as nv


class _Raw:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class LocalPeers:
    """Symmetric blocks on every visible GPU, owned by this one process (peer access instead of CUDA IPC)."""

    def __init__(self, n: int, nbytes: int):
        lib = nv.load()
        self.n, self.nbytes = n, (nbytes + 255) // 256 * 256
        for d in range(n):
            torch.cuda.set_device(d)
            for q in range(n):
                if q != d:
                    nv._check(lib.b200_enable_peer_access(q), "b200_enable_peer_access")
        self.ptrs: List[int] = []
        self.views: List[torch.Tensor] = []
        for d in range(n):
            torch.cuda.set_device(d)
            p = C.c_void_p()
            nv._check(lib.b200_symm_alloc(C.byref(p), self.nbytes), "b200_symm_alloc")
            self.ptrs.append(p.value)
            self.views.append(torch.as_tensor(_Raw(p.value, self.nbytes), device=torch.device("cuda", d)))
        self.table = (C.c_void_p * n)(*self.ptrs)

    def tensor(self, d: int, offset: int, shape, dtype) -> torch.Tensor:
        nb = torch.empty((), dtype=dtype).element_size()
        for s in shape:
            nb *= s
        return self.views[d][offset: offset + nb].view(dtype).view(*shape)

    def close(self) -> None:
        lib = nv.load()
        for d, p in enumerate(self.ptrs):
            torch.cuda.set_device(d)
            torch.cuda.synchronize()
            lib.b200_symm_free(C.c_void_p(p))
        self.views, self.ptrs = [], []


def collectives(n: int | None = None, N: int = 128) -> dict:
    lib = nv.load()
    n = n or torch.cuda.device_count()
    data = LocalPeers(n, 4 * N * n + 256)
    flags = LocalPeers(n, int(lib.b200_coll_flag_bytes()))
    streams = [torch.cuda.Stream(device=d) for d in range(n)]
    res = {}
    # ---- all_gather (test_nccl.py:8-19): rank d's slice d is its input
    cpu_in = [torch.rand(N) for _ in range(n)]
    for d in range(n):
        data.tensor(d, 0, (n * N,), torch.float32)[d * N:(d + 1) * N].copy_(cpu_in[d])
    for d in range(n):
        torch.cuda.set_device(d)
        torch.cuda.synchronize()
    epoch = 1
    for d in range(n):                       # one launch per GPU; each kernel waits for its peers' kernels
        torch.cuda.set_device(d)
        nv._check(lib.b200_all_gather(data.table, 0, 4 * N * n, flags.table, d, n, epoch, 0, streams[d].cuda_stream), "b200_all_gather")
    expect = torch.cat(cpu_in)
    res["all_gather"] = all(torch.equal(_sync(d, data.tensor(d, 0, (n * N,), torch.float32)).cpu(), expect) for d in range(n))
    # ---- all_reduce in place (:21-27)
    cpu_in = [torch.rand(N) for _ in range(n)]
    for d in range(n):
        data.tensor(d, 0, (N,), torch.float32).copy_(cpu_in[d])
        _sync(d, None)
    epoch += 1
    for d in range(n):
        torch.cuda.set_device(d)
        nv._check(lib.b200_all_reduce_f32(data.table, 0, 4 * N, flags.table, d, n, epoch, 0, streams[d].cuda_stream), "b200_all_reduce_f32")
    expect = torch.stack(cpu_in).sum(0)
    res["all_reduce"] = all(torch.allclose(_sync(d, data.tensor(d, 0, (N,), torch.float32)).cpu(), expect) for d in range(n))
    # ---- reduce_scatter (:29-38): every GPU holds n*N inputs, GPU d ends with the sum of everyone's chunk d
    cpu_in = [torch.rand(N * n) for _ in range(n)]
    for d in range(n):
        data.tensor(d, 0, (n * N,), torch.float32).copy_(cpu_in[d])
        _sync(d, None)
    epoch += 1
    for d in range(n):
        torch.cuda.set_device(d)
        nv._check(lib.b200_reduce_scatter_f32(data.table, 0, 4 * N * n, flags.table, d, n, epoch, 0, streams[d].cuda_stream), "b200_reduce_scatter_f32")
    total = torch.stack(cpu_in).sum(0)
    res["reduce_scatter"] = all(torch.allclose(_sync(d, data.tensor(d, 0, (n * N,), torch.float32)[d * N:(d + 1) * N]).cpu(),
                                               total[d * N:(d + 1) * N]) for d in range(n))
    data.close()
    flags.close()
    torch.cuda.set_device(0)
    return res


def _sync(d: int, t):
    torch.cuda.set_device(d)
    torch.cuda.synchronize()
    return t


def streams(n: int | None = None, iters: int = 10) -> dict:
    lib = nv.load()
    n = n or torch.cuda.device_count()
    M, K, N = 128, 64, 256
    out = LocalPeers(n, M * N * 2 + 256)
    flags = LocalPeers(n, int(lib.b200_coll_flag_bytes()))
    g = torch.Generator().manual_seed(0)
    xs = [torch.rand(M, K, generator=g).to(torch.bfloat16) for _ in range(n)]
    ws = [torch.rand(N, K, generator=g).to(torch.bfloat16) for _ in range(n)]
    dx = [xs[d].to(f"cuda:{d}") for d in range(n)]
    dw = [ws[d].to(f"cuda:{d}") for d in range(n)]
    ss = [torch.cuda.Stream(device=d) for d in range(n)]      # one stream per DEVICE
    for d in range(n):
        _sync(d, None)
    for d in range(n):                                        # enqueue everything; nothing here blocks the host
        torch.cuda.set_device(d)
        with torch.cuda.stream(ss[d]):
            y = out.tensor(d, 0, (M, N), torch.bfloat16)
            for _ in range(iters):
                nv.gemm(dx[d], nv.MAJOR_K, dw[d], nv.MAJOR_K, y, M, N, K)
            nv._check(lib.b200_all_reduce_bf16(out.table, 0, M * N * 2, flags.table, d, n, 1, 0, ss[d].cuda_stream), "b200_all_reduce_bf16")
    expect = sum(xs[d].float() @ ws[d].float().T for d in range(n))
    ok = True
    for d in range(n):
        torch.cuda.set_device(d)
        ss[d].synchronize()
        got = out.tensor(d, 0, (M, N), torch.bfloat16).float().cpu()
        ok &= bool((got - expect).abs().max() <= 0.02 * expect.abs().max())
    out.close()
    flags.close()
    torch.cuda.set_device(0)
    return {"per_device_streams_gemm_then_allreduce": ok}


if __name__ == "__main__":
    print(collectives())
    print(streams())

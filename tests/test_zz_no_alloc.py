"""``b200_sumsq_bf16_static`` - the entry point for callers that can neither pass scratch memory nor let the library
allocate (a bare C host, a plug-in in a process whose GPU memory is fully owned by another allocator, a step loop
that must stay off the driver's allocation lock).  Called here straight through the C ABI with ctypes, on two calls in
a row with different inputs, and compared with the scratch-taking ``b200_sumsq_bf16``.

Collected last (file name + conftest ordering) so a ``-x`` run has executed every other test before this one."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from distributed_llm_code_samples_b200.ops import native as nv
    lib = C.CDLL(str(nv.lib_path()))
    vp, ll = C.c_void_p, C.c_longlong
    lib.b200_sumsq_bf16.argtypes = [vp, vp, ll, vp, ll, vp]
    lib.b200_sumsq_bf16.restype = C.c_int
    lib.b200_sumsq_bf16_static.argtypes = [vp, vp, ll, vp]
    lib.b200_sumsq_bf16_static.restype = C.c_int
    lib.b200_sumsq_scratch_bytes.restype = ll
    lib.b200_sumsq_static_max_elems.restype = ll
    lib.b200_last_error.restype = C.c_char_p
    return lib


def test_static_scratch_sumsq_matches_allocating_entry_point(cuda_device):
    lib = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.zeros(int(lib.b200_sumsq_scratch_bytes()), dtype=torch.uint8, device=cuda_device)
    ref = torch.zeros(1, dtype=torch.float32, device=cuda_device)
    got = torch.zeros(1, dtype=torch.float32, device=cuda_device)
    # reference values first, from the existing entry point
    cases = []
    for seed, n in ((1, 3_000_003), (2, 777_777), (3, 8)):
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(n, generator=g).to(torch.bfloat16).to(cuda_device)
        assert lib.b200_sumsq_bf16(ref.data_ptr(), x.data_ptr(), n, scratch.data_ptr(), scratch.numel(), stream) == 0
        torch.cuda.synchronize()
        cases.append((x, n, ref.item()))
    # then the new entry point, calls back to back with different lengths and contents: the array they share must not
    # carry anything over from one call to the next
    outs = [torch.zeros(1, dtype=torch.float32, device=cuda_device) for _ in cases]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(cuda_device)          # driver-level: sees any cudaMalloc, torch's or not
    for (x, n, _), got in zip(cases, outs):
        assert lib.b200_sumsq_bf16_static(got.data_ptr(), x.data_ptr(), n, stream) == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(cuda_device)[0] >= free_before
    for (x, n, r), got in zip(cases, outs):
        v = got.item()
        assert abs(v - r) / r < 1e-4, (n, v, r)                     # same tolerance as tests/test_gemm_gpu.py
        assert abs(v - (x.float() ** 2).sum().item()) / r < 1e-4


def test_static_scratch_sumsq_rejects_oversized_input(cuda_device):
    lib = _lib()
    out = torch.zeros(1, dtype=torch.float32, device=cuda_device)
    x = torch.zeros(8, dtype=torch.bfloat16, device=cuda_device)
    too_many = int(lib.b200_sumsq_static_max_elems()) + 1
    rc = lib.b200_sumsq_bf16_static(out.data_ptr(), x.data_ptr(), too_many, torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    assert b"limit" in lib.b200_last_error()

"""Host side of the experimental VMM/multicast runtime: passing file descriptors between processes (SCM_RIGHTS) through
the native library, which loads without a GPU."""
import ctypes as C
import multiprocessing as mp
import os

from distributed_llm_code_samples_b200.build_native import build
from distributed_llm_code_samples_b200.ops import native as nv


def _lib():
    build()
    lib = C.CDLL(str(nv.lib_path()))
    lib.b200_fd_listen.argtypes = [C.c_char_p]
    lib.b200_fd_send.argtypes = [C.c_char_p, C.c_int, C.c_int]
    lib.b200_fd_recv.argtypes = [C.c_int, C.POINTER(C.c_int)]
    lib.b200_last_error.restype = C.c_char_p
    return lib


def _sender(path: str, payload: bytes, tag: int):
    lib = _lib()
    fd = os.memfd_create("b200ffn-test")
    os.write(fd, payload)
    assert lib.b200_fd_send(path.encode(), fd, tag) == 0, lib.b200_last_error()
    os.close(fd)


def test_descriptor_travels_between_processes(tmp_path):
    lib = _lib()
    path = str(tmp_path / "rx.sock")
    sock = lib.b200_fd_listen(path.encode())
    assert sock >= 0, lib.b200_last_error()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_sender, args=(path, f"hello from {r}".encode(), r)) for r in (1, 2)]
    for p in procs:
        p.start()
    got = {}
    for _ in procs:
        tag = C.c_int()
        fd = lib.b200_fd_recv(sock, C.byref(tag))
        assert fd >= 0, lib.b200_last_error()
        os.lseek(fd, 0, os.SEEK_SET)
        got[tag.value] = os.read(fd, 64)        # the SAME open file the other process wrote
        os.close(fd)
    for p in procs:
        p.join()
        assert p.exitcode == 0
    assert got == {1: b"hello from 1", 2: b"hello from 2"}
    lib.b200_fd_close(sock)


def test_bad_socket_path_is_rejected():
    lib = _lib()
    assert lib.b200_fd_listen(b"") == -1
    assert lib.b200_fd_send(("/tmp/" + "x" * 200).encode(), 0, 0) == -1
    assert b"socket path" in lib.b200_last_error()

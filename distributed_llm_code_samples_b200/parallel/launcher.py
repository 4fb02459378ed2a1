"""Process launcher: one process per rank, loopback rendezvous, exit codes checked.

Reference: ``init_process`` (``train_ffns.py:121-127``) and the ``mp.Process`` spawn/join loops (``:184-191``,
``:275-282``, ``:326-333``).  Differences by design: the backend and port are configurable (gloo on CPU for the
plumbing tests), every rank binds its own device before creating the group, the group is destroyed on exit, a
crashed rank fails the whole launch instead of hanging it, and results come back through CPU shared-memory
tensors the parent allocated rather than CUDA-IPC views of parent-owned device memory.
"""
from __future__ import annotations

import multiprocessing.connection as mpc
import os
import socket
import time
import traceback
from datetime import timedelta
from typing import Any, Callable, Sequence

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


PEER_GRACE_S = 5.0     # after the first failed rank, how long the others may keep running before they are terminated


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def pick_backend(backend: str, device: str) -> str:
    if backend != "auto":
        return backend
    return "nccl" if device == "cuda" else "gloo"


def pick_device(device: str) -> str:
    if device != "auto":
        return device
    return "cuda" if torch.cuda.is_available() else "cpu"


def visible_world(world_size: int, device: str) -> int:
    """0 = all visible GPUs, the reference's ``nGPUs = torch.cuda.device_count()`` (``train_ffns.py:25``)."""
    if world_size > 0:
        return world_size
    return max(torch.cuda.device_count(), 1) if device == "cuda" else 1


def init_process(rank: int, world: int, backend: str, port: int, device: str, timeout_s: int = 600) -> torch.device:
    """Join the process group (``train_ffns.py:121-125``) and return this rank's device."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if device == "cuda":
        dev = torch.device("cuda", rank % max(torch.cuda.device_count(), 1))
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    kwargs = {}
    if backend == "nccl":
        kwargs["device_id"] = dev
    dist.init_process_group(backend, rank=rank, world_size=world, timeout=timedelta(seconds=timeout_s), **kwargs)
    return dev


def _entry(rank: int, world: int, backend: str, port: int, device: str, fn: Callable, args: Sequence[Any], errq) -> None:
    try:
        dev = init_process(rank, world, backend, port, device)
        fn(rank, world, dev, *args)
        if dist.is_initialized():
            if device == "cuda":
                torch.cuda.synchronize()
            dist.barrier()
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put((rank, traceback.format_exc()))
        raise


def launch(world: int, fn: Callable, args: Sequence[Any], backend: str = "auto", device: str = "auto",
           port: int = 0, join_timeout_s: int = 3600) -> None:
    """``_launch_once`` with one retry when the rendezvous port picked by ``free_port`` was taken between the probe and
    the bind (EADDRINUSE; seen once in ~100 launches on the GPU boxes)."""
    try:
        _launch_once(world, fn, args, backend, device, port, join_timeout_s)
    except RuntimeError as e:
        if port == 0 and "EADDRINUSE" in str(e):
            _launch_once(world, fn, args, backend, device, 0, join_timeout_s)
        else:
            raise


def _launch_once(world: int, fn: Callable, args: Sequence[Any], backend: str = "auto", device: str = "auto",
                 port: int = 0, join_timeout_s: int = 3600) -> None:
    """Run ``fn(rank, world, device, *args)`` in ``world`` spawned processes and wait for them.

    Raises ``RuntimeError`` with the first failing rank's traceback if any rank exits non-zero."""
    device = pick_device(device)
    backend = pick_backend(backend, device)
    port = port or free_port()
    ctx = mp.get_context("spawn")          # train_ffns.py:375
    errq = ctx.SimpleQueue()
    procs = []
    for rank in range(world):
        p = ctx.Process(target=_entry, args=(rank, world, backend, port, device, fn, tuple(args), errq))
        p.start()
        procs.append(p)
    # Wait on ALL ranks at once: a rank that dies while its peers are blocked in a collective must fail the launch now,
    # not after the peers' own (NCCL / flag) timeouts.  The error queue is drained while waiting so that a child with a
    # large traceback never blocks in put().
    pending = {p.sentinel: (rank, p) for rank, p in enumerate(procs)}
    deadline = time.monotonic() + join_timeout_s
    failed, tbs = [], []

    def drain() -> None:
        while not errq.empty():
            r, t = errq.get()
            tbs.append(f"\n--- rank {r} ---\n{t}")

    grace_until = None
    while pending:
        ready = mpc.wait(list(pending), timeout=0.5)
        drain()
        for s_ in ready:
            rank, p = pending.pop(s_)
            p.join()
            if p.exitcode != 0:
                failed.append((rank, f"exit code {p.exitcode}"))
        now = time.monotonic()
        if failed and grace_until is None:
            grace_until = now + PEER_GRACE_S          # let the peers report their own tracebacks, briefly
        if (grace_until is not None and now > grace_until) or now > deadline:
            break
    if pending:
        for rank, p in pending.values():              # do not leave peers blocked in a collective
            if now > deadline and not failed:
                failed.append((rank, "timeout"))
            if p.is_alive():
                p.terminate()
        for rank, p in pending.values():
            p.join(10)
            if p.is_alive():
                p.kill()
    drain()
    if failed:
        raise RuntimeError(f"ranks failed: {sorted(failed)}{''.join(tbs)}")


def under_torchrun() -> bool:
    return "RANK" in os.environ and "WORLD_SIZE" in os.environ


def init_from_env(device: str = "auto", backend: str = "auto", timeout_s: int = 600) -> tuple[int, int, torch.device]:
    """Initialise from torchrun's environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).  ``B200_INIT_TIMEOUT_S``
    overrides the rendezvous / collective timeout."""
    timeout_s = int(os.environ.get("B200_INIT_TIMEOUT_S", timeout_s))
    device = pick_device(device)
    backend = pick_backend(backend, device)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if device == "cuda":
        dev = torch.device("cuda", local)
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kwargs = {"device_id": dev} if backend == "nccl" else {}
        dist.init_process_group(backend, rank=rank, world_size=world, timeout=timedelta(seconds=timeout_s), **kwargs)
    return rank, world, dev

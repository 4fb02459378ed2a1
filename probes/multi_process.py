"""One process per GPU - the counterparts of ``test_torch_distributed.py`` and ``test_mp_barrier_gpus.py``.

* ``async_allreduce_side_stream``: a sub-group (``dist.new_group``, ``test_torch_distributed.py:11``), ten asynchronous
  all-reduces of ``ones(1)`` on the native peer-memory kernels, each followed by ``+100`` on a side stream that first
  waits for the handle (``:13-20``); every rank must see ``world + 100`` ten times (the reference prints 104 for 4 ranks).
* ``ipc_barrier``: every rank mutates memory it owns that is mapped into its peers (CUDA IPC), all meet at a
  DEVICE-side barrier, then rank 0 reads every peer's block through the peer mapping ("Inside",
  ``test_mp_barrier_gpus.py:15-19``) and the parent process sees the values through CPU shared memory ("Outside", ``:35``).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from distributed_llm_code_samples_b200.parallel.launcher import launch
from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld


def _async_worker(rank, world, dev, out):
    group = dist.new_group(list(range(world)))                          # sub-group API, as :11
    sw = SymmWorld(rank, world, dev, group)
    comm = SymmComm(sw, staging_bytes=1 << 20, group=group)
    side = torch.cuda.Stream(device=dev)
    seen = []
    for _ in range(10):
        t = torch.ones(8, dtype=torch.float32, device=dev)
        h = comm.all_reduce(t, async_op=True)                           # :15
        h.wait()                                                        # :16
        side.wait_stream(torch.cuda.current_stream())                   # :17-18
        with torch.cuda.stream(side):
            t.add_(100)                                                 # :19-20
        side.synchronize()
        seen.append(float(t[0]))
    out[rank] = 1.0 if all(v == world + 100 for v in seen) else 0.0
    sw.close()


def async_allreduce_side_stream(world: int | None = None) -> bool:
    world = world or torch.cuda.device_count()
    out = torch.zeros(world).share_memory_()
    launch(world, _async_worker, (out,), backend="nccl", device="cuda")
    return out.tolist() == [1.0] * world


def _ipc_worker(rank, world, dev, outside):
    sw = SymmWorld(rank, world, dev)
    comm = SymmComm(sw, staging_bytes=1 << 20)
    block = sw.alloc(2 * 2 * 2 * 4)                                     # two [2,2] fp32 tensors per GPU, as :21-28
    a = block.tensor(0, (2, 2), torch.float32)
    b = block.tensor(16, (2, 2), torch.float32)
    a.add_(rank)                                                        # :15-16
    b.add_(rank)
    comm.barrier()                                                      # device-side, where the reference uses mp.Barrier (:17)
    inside = torch.empty(world * 8, dtype=torch.float32, device=dev)
    comm.all_gather_into(inside, block.tensor(0, (8,), torch.float32))  # read every peer's block through the mapping
    torch.cuda.synchronize()
    if rank == 0:
        print("Inside", inside.view(world, 2, 2, 2).cpu().tolist())    # :18-19
        outside.copy_(inside.cpu())
    sw.close()


def ipc_barrier(world: int | None = None) -> bool:
    world = world or torch.cuda.device_count()
    outside = torch.zeros(world * 8).share_memory_()
    launch(world, _ipc_worker, (outside,), backend="nccl", device="cuda")
    print("Outside", outside.view(world, 2, 2, 2).tolist())            # :35
    expect = torch.arange(world, dtype=torch.float32).repeat_interleave(8)
    return torch.equal(outside, expect)


if __name__ == "__main__":
    print("async all-reduce + side stream:", async_allreduce_side_stream())
    print("ipc + barrier:", ipc_barrier())

"""Does this box do NVLS (NVSwitch multicast / in-switch reduction)?  Run under torchrun on >= 2 GPUs:

    NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/nvls_probe.py 2>&1 | grep -iE "nvls|multicast|probe:"

Prints (rank 0): the driver's attribute bits for POSIX-fd export and multicast on every visible device, whether a
multicast object can actually be created / bound / mapped through parallel/vmm.py, and leaves NCCL's own NVLS lines in
the log for comparison (NCCL prints "NVLS multicast support is available" and "Connected NVLS tree" when it uses it)."""
import os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.parallel import vmm
from distributed_llm_code_samples_b200.parallel.launcher import init_from_env

rank, world, dev = init_from_env("cuda", "nccl")
fd_ok, mc_ok = vmm.supported()
t = torch.ones(1 << 22, device=dev, dtype=torch.bfloat16)
dist.all_reduce(t)                       # makes NCCL build its channels (and print its NVLS lines under NCCL_DEBUG=INFO)
torch.cuda.synchronize()
bits = torch.tensor([int(fd_ok), int(mc_ok)], device=dev)
allbits = [torch.zeros_like(bits) for _ in range(world)]
dist.all_gather(allbits, bits)
mapped = "skipped"
if fd_ok and mc_ok and world > 1:
    try:
        buf = vmm.VmmBuffer(rank, world, dev, 1 << 21, multicast=True, tag="probe")
        mapped = "ok" if buf.mc_ptr else "no multicast pointer"
    except Exception as e:  # noqa: BLE001
        mapped = f"failed: {e}"
if rank == 0:
    print("probe: (posix_fd, multicast) per rank =", [tuple(b.tolist()) for b in allbits])
    print("probe: multicast object create/bind/map =", mapped)
dist.barrier()

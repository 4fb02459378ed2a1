"""Substrate probes: B200-native counterparts of the reference's four exploratory scripts (test_nccl.py,
test_torch_distributed.py, test_torch_cuda_stream.py, test_mp_barrier_gpus.py).  Each is runnable on its own
(``python -m probes.<name>``) on a multi-GPU node and is also called from tests/test_multigpu.py."""

"""Synthetic data: the reference's per-seed CPU generator plus pinned-staging and on-device variants.

Reference: ``mock_data`` (``train_ffns.py:144-151``) - for every seed, reseed a CPU generator and draw
``x = randn(T, D)`` then ``dloss_dx = 0.1 * randn(T, D)``; the seed vector is
``randint(100_000, (num_steps,))`` (``:360``); DDP/FSDP give rank ``r`` the seeds ``seeds[r::n]`` (``:182``).
"""
from __future__ import annotations

from typing import Iterator, Optional, Sequence, Tuple

import torch

from ..config import DLOSS_DX_COEF, SEED_RANGE


def make_seeds(num_steps: int, gen: Optional[torch.Generator]) -> torch.Tensor:
    return torch.randint(SEED_RANGE, (num_steps,), generator=gen)


def split_seeds(seeds: torch.Tensor, world: int) -> list[torch.Tensor]:
    """Rank r gets ``seeds[r::world]`` (``train_ffns.py:182``). Requires ``len(seeds) % world == 0`` (``:175``)."""
    if len(seeds) % world != 0:
        raise ValueError(f"num_steps ({len(seeds)}) must be divisible by world size ({world})")
    return [t.reshape(-1) for t in seeds.reshape((-1, world)).chunk(world, dim=1)]


def mock_batch(seed: int, tokens: int, model_size: int, gen: Optional[torch.Generator] = None):
    gen = gen or torch.Generator()
    gen.manual_seed(int(seed))
    x = torch.randn((tokens, model_size), generator=gen)
    dloss_dx = DLOSS_DX_COEF * torch.randn((tokens, model_size), generator=gen)
    return x, dloss_dx


def mock_data(seeds: Sequence[int] | torch.Tensor, tokens: int, model_size: int,
              dtype: torch.dtype = torch.float32, device: Optional[torch.device] = None,
              pin: bool = False) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Yield ``(x, dloss_dx)`` per seed; bit-identical to the reference on CPU/fp32."""
    gen = torch.Generator()
    seed_list = seeds.tolist() if isinstance(seeds, torch.Tensor) else list(seeds)
    for seed in seed_list:
        x, dy = mock_batch(seed, tokens, model_size, gen)
        if dtype != torch.float32:
            x, dy = x.to(dtype), dy.to(dtype)
        if pin and torch.cuda.is_available():
            x, dy = x.pin_memory(), dy.pin_memory()
        if device is not None and device.type != "cpu":
            x, dy = x.to(device, non_blocking=pin), dy.to(device, non_blocking=pin)
        yield x, dy


def device_mock_data(seeds: Sequence[int] | torch.Tensor, tokens: int, model_size: int, dtype: torch.dtype,
                     device: torch.device) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Same contract, generated on the device (Philox): same seed => same batch on every rank, but the values
    differ from the CPU generator's.  bf16 on CUDA uses the native ``b200_randn_bf16`` kernel."""
    seed_list = seeds.tolist() if isinstance(seeds, torch.Tensor) else list(seeds)
    use_native = device.type == "cuda" and dtype == torch.bfloat16
    if use_native:
        from ..ops import native as nv
    for seed in seed_list:
        if use_native:
            x = nv.randn_(torch.empty(tokens, model_size, dtype=dtype, device=device), 2 * int(seed), 1.0)
            dy = nv.randn_(torch.empty(tokens, model_size, dtype=dtype, device=device), 2 * int(seed) + 1, DLOSS_DX_COEF)
        else:
            g = torch.Generator(device=device)
            g.manual_seed(int(seed))
            x = torch.randn((tokens, model_size), generator=g, device=device).to(dtype)
            dy = (DLOSS_DX_COEF * torch.randn((tokens, model_size), generator=g, device=device)).to(dtype)
        yield x, dy


class PinnedPrefetcher:
    """Double-buffered pinned-host -> device staging on a side stream (replaces the reference's synchronous
    pageable ``.cuda()`` at ``train_ffns.py:106,158,231,294``).  ``next()`` returns device tensors whose copy has
    been ordered before the current stream."""

    def __init__(self, batches: Iterator[Tuple[torch.Tensor, torch.Tensor]], device: torch.device,
                 rows: Optional[slice] = None):
        """``rows``: copy only these token rows of each tensor (the rest of the device tensor is left uninitialised).
        Sequence-parallel MP reads nothing but its own token shard of the batch, so it moves 1/world of the bytes."""
        self.it = iter(batches)
        self.device = device
        self.rows = rows
        self.stream = torch.cuda.Stream(device=device)
        self.last_bytes = 0          # host->device bytes of the batch most recently returned by next()
        self._staged_bytes = 0
        self._next = None
        self._preload()

    def _to_device(self, t: torch.Tensor) -> torch.Tensor:
        if self.rows is None:
            self._staged_bytes += t.numel() * t.element_size()
            return t.to(self.device, non_blocking=True)
        out = torch.empty(t.shape, dtype=t.dtype, device=self.device)
        src = t[self.rows]
        out[self.rows].copy_(src, non_blocking=True)
        self._staged_bytes += src.numel() * src.element_size()
        return out

    def _preload(self) -> None:
        try:
            x, dy = next(self.it)
        except StopIteration:
            self._next = None
            return
        if not x.is_pinned():
            x, dy = x.pin_memory(), dy.pin_memory()
        self._staged_bytes = 0
        with torch.cuda.stream(self.stream):
            self._next = (self._to_device(x), self._to_device(dy))

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        torch.cuda.current_stream(self.device).wait_stream(self.stream)
        x, dy = self._next
        self.last_bytes = self._staged_bytes
        x.record_stream(torch.cuda.current_stream(self.device))
        dy.record_stream(torch.cuda.current_stream(self.device))
        self._preload()
        return x, dy


class ShardedPrefetcher:
    """Model-parallel input staging: every rank needs the WHOLE batch, but not from its own host copy.

    Each rank stages 1/world of the token rows host->device into a symmetric slot (side stream, double-buffered) and the
    ranks exchange the rest over NVLink with the native all-gather, so the node moves ONE batch per step over PCIe
    instead of ``world`` copies of it (8 x 512 MiB per step at the 70B shape - more than the host side sustained in
    round 1's measurements, where MP's end-to-end rate was 15 % below its device-timed rate in both arms).
    Every rank must iterate the same batches.  The reference copies the full batch on every rank (``train_ffns.py:294``).

    Selected with ``TrainConfig.mp_input = "sharded"`` (``--mp_input sharded``)."""

    def __init__(self, batches: Iterator[Tuple[torch.Tensor, torch.Tensor]], device: torch.device, sw, rank: int, world: int):
        from ..parallel.symm import CollChannel
        self.it, self.device, self.sw, self.rank, self.world = iter(batches), device, sw, rank, world
        self.stream = torch.cuda.Stream(device=device)
        self.coll = CollChannel(sw)
        self.stage = None
        self.free = [None, None]         # event per slot: the gather that read it (on every peer) has completed
        self.k = 0
        self.last_bytes = 0
        self._next = None
        self._preload()

    def _preload(self) -> None:
        try:
            x, dy = next(self.it)
        except StopIteration:
            self._next = None
            return
        if not x.is_pinned():
            x, dy = x.pin_memory(), dy.pin_memory()
        T = x.shape[0]
        if T % self.world:
            raise ValueError(f"sharded input staging needs tokens ({T}) divisible by the world size ({self.world})")
        ts = T // self.world
        rows = slice(self.rank * ts, (self.rank + 1) * ts)
        nb = [ts * t[0].numel() * t.element_size() for t in (x, dy)]
        if self.stage is None:           # collective: every rank allocates the same block at its first batch
            self.slot_bytes = [(b + 255) // 256 * 256 for b in nb]
            self.stage = self.sw.alloc(2 * sum(self.slot_bytes))
        slot = self.k % 2
        offs = [slot * sum(self.slot_bytes), slot * sum(self.slot_bytes) + self.slot_bytes[0]]
        if self.free[slot] is not None:
            self.stream.wait_event(self.free[slot])
        with torch.cuda.stream(self.stream):
            for t, off in zip((x, dy), offs):
                self.stage.tensor(off, (ts,) + tuple(t.shape[1:]), t.dtype).copy_(t[rows], non_blocking=True)
        self._next = (x.shape, x.dtype, dy.shape, dy.dtype, offs, nb, slot)
        self.k += 1

    def __iter__(self):
        return self

    def __next__(self):
        from ..ops import native as nv
        if self._next is None:
            raise StopIteration
        xs, xd, ds, dd, offs, nb, slot = self._next
        cur = torch.cuda.current_stream(self.device)
        cur.wait_stream(self.stream)
        outs = []
        for shape, dtype, off, b in ((xs, xd, offs[0], nb[0]), (ds, dd, offs[1], nb[1])):
            full = torch.empty(shape, dtype=dtype, device=self.device)
            e = self.coll.next_epoch()
            # opening barrier: every peer's staging copy is complete; closing barrier: every peer has read my slot
            nv._check(nv.load().b200_all_gather_into(full.data_ptr(), self.stage.peers_c, off, b, self.coll.flags.peers_c,
                                                     self.rank, self.world, e, 0, nv._stream()), "b200_all_gather_into")
            nv.launch_count += 1
            outs.append(full)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.free[slot] = ev
        self.last_bytes = sum(nb)
        self._preload()
        return outs[0], outs[1]

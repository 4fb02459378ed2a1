"""Trainers whose communication is fused into the GEMMs (``csrc/fused.cu``) or runs on the native peer-memory
collectives (``csrc/comm.cu``) - no NCCL on the data path.

* ``FusedDDPTrainer``  dW GEMM -> (push partial tiles to their owners) -> fp32 sum -> ``W -= lr * sum`` -> new weight
  tile stored into every replica.  Replaces dW GEMM + all_reduce + SGD (``train_ffns.py:66/68, :165, :172``).
* ``FusedFSDPTrainer`` dW GEMM -> reduce-scatter to the row-shard owner -> sharded SGD in the same kernel
  (``:253-259``); parameters are unsharded one layer ahead on a side stream by the copy engines (P2P DMA between two
  one-warp barrier kernels) straight into the contiguous buffer the GEMM reads (``:203-209``, ``:236-249``).
* ``FusedMPTrainer``   row-parallel GEMM -> all-reduce of the partial activations inside the kernel
  (``:302-303``, ``:308-309``), or (``mp_mode="sp"``) -> reduce-scatter over token shards + native all-gather, or
  (``mp_mode="spf"``) the all-gather fused into the column-parallel GEMM as well.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, List, Optional, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import Params
from ..ops import native as nv
from .base import Optimizer, block_opts, set_update_rounding
from .symm import CollChannel, SymmBuffer, SymmWorld, _Handle, get_symm_world

OWN_INTERLEAVED, OWN_ROW_SHARDED = 0, 1
FOP_SUM, FOP_SGD = 0, 1
TILE_M, TILE_N = 128, 256


def available() -> bool:
    try:
        return hasattr(nv.load(), "b200_gemm_allreduce_bf16")
    except OSError:
        return False


def _tiles(M: int, N: int) -> int:
    return ((M + TILE_M - 1) // TILE_M) * ((N + TILE_N - 1) // TILE_N)


class FusedChannel:
    """Inbox + flags + call counter for one stream of fused GEMM+collective calls."""

    def __init__(self, sw: SymmWorld, owned_cap: int, ready_cap: int, nvls: bool = False):
        """``nvls=True`` (activations of the MP paths, where the box has NVLS): partial tiles stay in a local, multicast-
        bound buffer and the owner sums them inside the NVSwitch, instead of every rank pushing into the owner's inbox."""
        lib = nv.load()
        self.sw, self.owned_cap, self.ready_cap = sw, owned_cap, ready_cap
        self.part = None
        if nvls and sw.multicast_supported():
            self.part = sw.alloc(int(lib.b200_fused_partial_bytes(ready_cap)), multicast=True)
            self.inbox = None
        else:
            self.inbox = sw.alloc(int(lib.b200_fused_inbox_bytes(owned_cap, sw.world)))
        self.flags = sw.alloc(int(lib.b200_fused_flag_bytes(owned_cap, ready_cap)))
        self.seq = 0

    @property
    def nvls(self) -> bool:
        return self.part is not None

    def gemm(self, A: torch.Tensor, a_major: int, B: torch.Tensor, b_major: int, M: int, N: int, K: int,
             out_peers, ldo: int, out_row_offset: int, owner_mode: int, op: int, broadcast: bool, alpha: float = 1.0,
             max_ctas: int = 0, out_mc: Optional[int] = None) -> None:
        self.seq += 1
        sw = self.sw
        if self.part is not None:
            assert not broadcast or out_mc, "the NVLS form broadcasts through the output's multicast address"
            rc = nv.load().b200_gemm_allreduce_nvls_bf16(
                A.data_ptr(), A.stride(0), a_major, B.data_ptr(), B.stride(0), b_major, M, N, K, out_peers, ldo,
                out_row_offset, self.part.ptr, self.part.mc_ptr, out_mc, self.flags.peers_c, self.owned_cap, self.ready_cap,
                sw.rank, sw.world, self.seq, owner_mode, op, int(broadcast), float(alpha), int(max_ctas), nv._stream())
            nv._check(rc, "b200_gemm_allreduce_nvls_bf16")
            nv.launch_count += 1
            return
        rc = nv.load().b200_gemm_allreduce_bf16(
            A.data_ptr(), A.stride(0), a_major, B.data_ptr(), B.stride(0), b_major, M, N, K, out_peers, ldo,
            out_row_offset, self.inbox.peers_c, self.flags.peers_c, self.owned_cap, self.ready_cap, sw.rank, sw.world,
            self.seq, owner_mode, op, int(broadcast), float(alpha), int(max_ctas), nv._stream())
        nv._check(rc, "b200_gemm_allreduce_bf16")
        nv.launch_count += 1

    def adam_state(self, M: int, N: int, owner_mode: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """Zeroed fp32 moments for the tiles this rank owns of an ``[M, N]`` weight (compact per-tile layout)."""
        lib = nv.load()
        cap = int(lib.b200_fused_owned_tiles(M, N, self.sw.world, owner_mode))
        n = int(lib.b200_fused_adam_state_bytes(cap)) // 4
        return (torch.zeros(n, dtype=torch.float32, device=self.sw.device), torch.zeros(n, dtype=torch.float32, device=self.sw.device))

    def gemm_adam(self, A: torch.Tensor, a_major: int, B: torch.Tensor, b_major: int, M: int, N: int, K: int,
                  out_peers, ldo: int, out_row_offset: int, owner_mode: int, broadcast: bool, state, lr: float, step: int,
                  beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, grad_scale: float = 1.0, max_ctas: int = 0) -> None:
        """As ``gemm`` with FOP_SGD, but the owner applies Adam (moments sharded by tile owner)."""
        self.seq += 1
        sw = self.sw
        m, v = state
        rc = nv.load().b200_gemm_allreduce_adam_bf16(
            A.data_ptr(), A.stride(0), a_major, B.data_ptr(), B.stride(0), b_major, M, N, K, out_peers, ldo,
            out_row_offset, self.inbox.peers_c, self.flags.peers_c, self.owned_cap, self.ready_cap, sw.rank, sw.world,
            self.seq, owner_mode, int(broadcast), m.data_ptr(), v.data_ptr(), float(lr), float(beta1), float(beta2), float(eps),
            int(step), float(grad_scale), int(max_ctas), nv._stream())
        nv._check(rc, "b200_gemm_allreduce_adam_bf16")
        nv.launch_count += 1


def _owned_cap(shapes, world: int, owner_mode: int) -> int:
    lib = nv.load()
    return max(int(lib.b200_fused_owned_tiles(M, N, world, owner_mode)) for M, N in shapes)


class GatherChannel:
    """Flags + call counter for one stream of all-gather -> GEMM calls (``csrc/gather_gemm.cu``).

    Validated on B200 in round 2 (2 GPUs and the single-GPU multi-rank emulation, ``tests/test_emu_gpu.py``)."""

    def __init__(self, sw: SymmWorld, max_rows: int):
        lib = nv.load()
        self.sw = sw
        self.started = sw.alloc(256)                                       # symmetric, zero-initialised
        self.ready = torch.zeros(int(lib.b200_gather_gemm_ready_bytes(max_rows)) // 4, dtype=torch.int32, device=sw.device)
        self.max_rows = max_rows
        self.seq = 0

    def gemm(self, a_peers, B: torch.Tensor, b_major: int, out: torch.Tensor, M: int, N: int, K: int, epilogue: int,
             aux: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None, max_ctas: int = 0,
             a_major: int = nv.MAJOR_K, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``out = epilogue(A_full @ B^T)``; ``a_peers[q]`` = rank q's buffer holding its row shard of A: ``[M, K]`` sharded
        along M (``a_major`` K) or ``[K, M]`` sharded along K (``a_major`` MN)."""
        rows = M if a_major == nv.MAJOR_K else K
        assert rows <= self.max_rows and B.dtype == torch.bfloat16 and B.stride(-1) == 1 and out.stride(-1) == 1
        self.seq += 1
        sw = self.sw
        rc = nv.load().b200_gather_gemm_bf16(
            a_peers, a_major, B.data_ptr(), B.stride(0), b_major, out.data_ptr(), out.stride(0), M, N, K, epilogue,
            nv._ptr(aux), aux.stride(0) if aux is not None else 0, nv._ptr(out2), out2.stride(0) if out2 is not None else 0,
            nv._ptr(bias), 1.0, self.started.peers_c, self.ready.data_ptr(), sw.rank, sw.world, self.seq, int(max_ctas), nv._stream())
        nv._check(rc, "b200_gather_gemm_bf16")
        nv.launch_count += 1
        return out


def _owned_tiles(M: int, N: int, world: int, rank: int, owner_mode: int) -> List[Tuple[int, int]]:
    """(m_blk, n_blk) of this rank's owned tiles in inbox-slot order - the Python mirror of ``owned_tile_coords`` in
    csrc/fused.cu (checked against the C mirror ``b200_debug_fused_map`` in tests/test_native_maps_cpu.py)."""
    tm, tn = (M + TILE_M - 1) // TILE_M, (N + TILE_N - 1) // TILE_N
    if owner_mode == OWN_ROW_SHARDED:
        mpo = tm // world
        return [(rank * mpo + k // tn, k % tn) for k in range(mpo * tn)]
    total = tm * tn
    return [divmod(lin, tn) for lin in range(rank, total, world)]


def _export_tiles(state: torch.Tensor, dst: torch.Tensor, M: int, N: int, world: int, rank: int, owner_mode: int, row0: int = 0) -> None:
    """Compact per-owned-tile moments ``[slot][128][256]`` -> the owned tiles of a dense CPU tensor (rows offset by row0)."""
    tiles = state.view(-1, TILE_M, TILE_N).cpu()
    for k, (mb, nb) in enumerate(_owned_tiles(M, N, world, rank, owner_mode)):
        r0, c0 = mb * TILE_M, nb * TILE_N
        r1, c1 = min(r0 + TILE_M, M), min(c0 + TILE_N, N)
        dst[r0 - row0:r1 - row0, c0:c1].copy_(tiles[k, :r1 - r0, :c1 - c0])


def _import_tiles(state: torch.Tensor, src: torch.Tensor, M: int, N: int, world: int, rank: int, owner_mode: int, row0: int = 0) -> None:
    tiles = torch.zeros(state.numel() // (TILE_M * TILE_N), TILE_M, TILE_N, dtype=torch.float32)
    for k, (mb, nb) in enumerate(_owned_tiles(M, N, world, rank, owner_mode)):
        r0, c0 = mb * TILE_M, nb * TILE_N
        r1, c1 = min(r0 + TILE_M, M), min(c0 + TILE_N, N)
        tiles[k, :r1 - r0, :c1 - c0].copy_(src[r0 - row0:r1 - row0, c0:c1])
    state.copy_(tiles.reshape(-1).to(state.device))


class _ReplicatedBiases:
    """Biases of a data-parallel stack under the fused trainers: replicated fp32 vectors (a few KiB per layer) in ONE
    contiguous tensor, their gradients (native column sums of ``dh`` / ``dy``) in one symmetric fp32 buffer, so a step
    costs one native all-reduce and one optimizer kernel for all layers together.  The reference has no bias
    (``train_ffns.py:35``); this is the north-star block's."""

    def __init__(self, sw: SymmWorld, cfg: TrainConfig, biases: List[Tuple[torch.Tensor, torch.Tensor]], device: torch.device):
        F, D, L = cfg.ffn_dim, cfg.model_size, len(biases)
        self.sw, self.L = sw, L
        self.n = L * (F + D)
        self.values = torch.empty(self.n, dtype=torch.float32, device=device)
        self.gbuf = sw.alloc(self.n * 4 + 256)
        self.grads, self.goff = self.gbuf.carve((self.n,), torch.float32)
        self.coll = CollChannel(sw)
        self.b1, self.b2, self.g1, self.g2 = [], [], [], []
        for l, (b1, b2) in enumerate(biases):
            o = l * (F + D)
            self.b1.append(self.values[o:o + F]); self.b2.append(self.values[o + F:o + F + D])
            self.g1.append(self.grads[o:o + F]); self.g2.append(self.grads[o + F:o + F + D])
            self.b1[l].copy_(b1.to(device=device, dtype=torch.float32))
            self.b2[l].copy_(b2.to(device=device, dtype=torch.float32))
        self.adam = cfg.optimizer == "adam"
        self.lr = cfg.lr * cfg.grad_scale
        if self.adam:
            self.m, self.v = torch.zeros_like(self.values), torch.zeros_like(self.values)

    def export_state(self, m_out, v_out) -> None:
        if self.adam and self.sw.rank == 0:
            for l in range(self.L):
                for j, (b, off) in enumerate(((self.b1[l], 0), (self.b2[l], self.b1[l].numel()))):
                    o = self.b1[l].data_ptr() - self.values.data_ptr()
                    lo = o // 4 + off
                    m_out[l][2 + j].copy_(self.m[lo:lo + b.numel()].cpu())
                    v_out[l][2 + j].copy_(self.v[lo:lo + b.numel()].cpu())

    def import_state(self, m_in, v_in) -> None:
        if self.adam:
            for l in range(self.L):
                for j, (b, off) in enumerate(((self.b1[l], 0), (self.b2[l], self.b1[l].numel()))):
                    lo = (self.b1[l].data_ptr() - self.values.data_ptr()) // 4 + off
                    self.m[lo:lo + b.numel()].copy_(m_in[l][2 + j].to(self.m.device))
                    self.v[lo:lo + b.numel()].copy_(v_in[l][2 + j].to(self.v.device))

    def accumulate(self, l: int, dh: torch.Tensor, dy: torch.Tensor) -> None:
        """This rank's bias gradients of layer ``l``: db1 = dh.sum(0), db2 = dy.sum(0) (native column sums)."""
        nv.colsum(dh, out=self.g1[l])
        nv.colsum(dy, out=self.g2[l])

    def reduce_and_update(self, step: int) -> None:
        """SUM over ranks (train_ffns.py:165 semantics) on the native peer-memory all-reduce, then one optimizer kernel."""
        sw = self.sw
        if sw.world > 1:
            nbytes = (self.n * 4 + 15) // 16 * 16
            nv._check(nv.load().b200_all_reduce_f32(self.gbuf.peers_c, self.goff, nbytes, self.coll.flags.peers_c, sw.rank, sw.world,
                                                    self.coll.next_epoch(), 0, nv._stream()), "b200_all_reduce_f32 (bias gradients)")
            nv.launch_count += 1
        if self.adam:
            nv.adam_f32_(self.values, self.grads, self.m, self.v, self.lr, 0.9, 0.999, 1e-8, step)
        else:
            nv.sgd_f32_(self.values, self.grads, self.lr)


# =================================================================================================== DDP
class FusedDDPTrainer:
    comm_name = "fused: dW GEMM + all-reduce + SGD in one kernel (peer stores over NVLink)"

    def __init__(self, ops, comm, cfg: TrainConfig, layers_params: List[Params], device: torch.device):
        self.ops, self.comm, self.cfg = ops, comm, cfg
        self.opts = block_opts(cfg)
        self.lr = cfg.lr * cfg.grad_scale
        self.sw = get_symm_world(comm, device)
        D, F, L = cfg.model_size, cfg.ffn_dim, len(layers_params)
        self.wbuf = self.sw.alloc(L * 2 * D * F * 2 + 4096)
        self.params: List[Params] = []
        self.offsets: List[Tuple[int, int]] = []
        for lp in layers_params:
            w1, o1 = self.wbuf.carve((F, D), torch.bfloat16)
            w2, o2 = self.wbuf.carve((D, F), torch.bfloat16)
            w1.copy_(lp[0].to(device=device, dtype=torch.bfloat16))
            w2.copy_(lp[1].to(device=device, dtype=torch.bfloat16))
            self.params.append([w1, w2])
            self.offsets.append((o1, o2))
        self.biases: Optional[_ReplicatedBiases] = None
        if len(layers_params[0]) == 4:
            self.biases = _ReplicatedBiases(self.sw, cfg, [(lp[2], lp[3]) for lp in layers_params], device)
            for l, p in enumerate(self.params):
                p += [self.biases.b1[l], self.biases.b2[l]]
        self.peer_tabs = [(self.wbuf.peers_at(o1), self.wbuf.peers_at(o2)) for o1, o2 in self.offsets]
        shapes = [(F, D), (D, F)]
        self.chan = FusedChannel(self.sw, _owned_cap(shapes, comm.world, OWN_INTERLEAVED), max(_tiles(*s) for s in shapes))
        self.adam = cfg.optimizer == "adam"
        if self.adam:                          # 1/world of the moments per rank: the owner of a tile owns its state
            self.comm_name = "fused: dW GEMM + all-reduce + Adam (state sharded by tile owner) in one kernel"
            self.state = [[self.chan.adam_state(*shapes[k], OWN_INTERLEAVED) for k in range(2)] for _ in layers_params]
        self.t = 0
        self.sw.host_barrier()

    def _update(self, i: int, k: int, dy: torch.Tensor, xin: torch.Tensor, M: int, N: int, T: int) -> None:
        """``W[i][k] -= update(sum over ranks of dy^T xin)`` and the new weight stored to every replica - one kernel."""
        if self.adam:
            self.chan.gemm_adam(dy, nv.MAJOR_MN, xin, nv.MAJOR_MN, M, N, T, self.peer_tabs[i][k], N, 0, OWN_INTERLEAVED, True,
                                self.state[i][k], self.lr, self.t)
        else:
            self.chan.gemm(dy, nv.MAJOR_MN, xin, nv.MAJOR_MN, M, N, T, self.peer_tabs[i][k], N, 0, OWN_INTERLEAVED, FOP_SGD, True, self.lr)

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        from ..models.ffn import tlayers_ffn_fwd
        cfg, ops = self.cfg, self.ops
        D, F = cfg.model_size, cfg.ffn_dim
        self.t += 1
        set_update_rounding(cfg, ops, self.t)
        y, acts = tlayers_ffn_fwd(ops, self.params, x, self.opts)
        d = dloss_dx
        T = x.shape[0]
        for i in reversed(range(len(self.params))):
            W1, W2 = self.params[i][:2]
            s = acts[i]
            a, pre = s.a, s.pre
            if a is None:
                a, pre = ops.linear_act(s.x, W1, self.opts.activation, self.params[i][2] if self.biases else None)
            dh = ops.grad_input(d, W2, act=self.opts.activation, aux=(pre if pre is not None else a) if self.opts.activation == "relu" else pre)
            if self.biases:
                self.biases.accumulate(i, dh, d)
            # dW2 = d^T a  summed over ranks, W2 -= lr * dW2, new W2 stored to every replica - one kernel
            self._update(i, 1, d, a, D, F, T)
            dn = ops.grad_input(dh, W1, add=d if self.opts.residual else None) if i > 0 else None
            self._update(i, 0, dh, s.x, F, D, T)
            acts[i] = None
            d = dn
        if self.biases:
            self.biases.reduce_and_update(self.t)
        return y

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        return self.params

    # ---- optimizer state in the checkpoint's full layout (parallel/opt_state.py) -------------------------------
    def export_opt_state(self, m_out, v_out) -> int:
        """Every rank writes the moments of the tiles it OWNS (disjoint, interleaved) into the shared full tensors."""
        if self.adam:
            cfg = self.cfg
            shapes = [(cfg.ffn_dim, cfg.model_size), (cfg.model_size, cfg.ffn_dim)]
            for i in range(len(self.params)):
                for k in range(2):
                    m, v = self.state[i][k]
                    _export_tiles(m, m_out[i][k], *shapes[k], self.sw.world, self.sw.rank, OWN_INTERLEAVED)
                    _export_tiles(v, v_out[i][k], *shapes[k], self.sw.world, self.sw.rank, OWN_INTERLEAVED)
            if self.biases:
                self.biases.export_state(m_out, v_out)
        return self.t

    def import_opt_state(self, m_in, v_in, t: int) -> None:
        self.t = int(t)
        if self.adam:
            cfg = self.cfg
            shapes = [(cfg.ffn_dim, cfg.model_size), (cfg.model_size, cfg.ffn_dim)]
            for i in range(len(self.params)):
                for k in range(2):
                    m, v = self.state[i][k]
                    _import_tiles(m, m_in[i][k], *shapes[k], self.sw.world, self.sw.rank, OWN_INTERLEAVED)
                    _import_tiles(v, v_in[i][k], *shapes[k], self.sw.world, self.sw.rank, OWN_INTERLEAVED)
            if self.biases:
                self.biases.import_state(m_in, v_in)


# =================================================================================================== FSDP
class FusedFSDPTrainer:
    comm_name = "fused: dW GEMM + reduce-scatter + sharded SGD in one kernel; P2P-DMA all-gather prefetch"

    def __init__(self, ops, comm, cfg: TrainConfig, shards: List[Params], device: torch.device):
        self.ops, self.comm, self.cfg = ops, comm, cfg
        self.opts = block_opts(cfg)
        self.lr = cfg.lr * cfg.grad_scale
        self.sw = get_symm_world(comm, device)
        n = comm.world
        self.world, self.rank = n, comm.rank
        D, F, L = cfg.model_size, cfg.ffn_dim, len(shards)
        self.L = L
        self.sbuf = self.sw.alloc(L * 2 * D * F * 2 // n + 4096)
        self.shards: List[Params] = []
        self.shard_off: List[Tuple[int, int]] = []
        for lp in shards:
            w1, o1 = self.sbuf.carve((F // n, D), torch.bfloat16)
            w2, o2 = self.sbuf.carve((D // n, F), torch.bfloat16)
            w1.copy_(lp[0])
            w2.copy_(lp[1])
            self.shards.append([w1, w2])
            self.shard_off.append((o1, o2))
        # biases: replicated (they are tiny), exposed to the driver as dim-0 shards like every other parameter (:265-266)
        self.biases: Optional[_ReplicatedBiases] = None
        if len(shards[0]) == 4:
            fullb = []
            for lp in shards:
                pair = []
                for b in lp[2:]:
                    fb = torch.empty(b.numel() * n, dtype=torch.float32, device=device)
                    comm.all_gather_into(fb, b.to(device=device, dtype=torch.float32).contiguous())
                    pair.append(fb)
                fullb.append(tuple(pair))
            self.biases = _ReplicatedBiases(self.sw, cfg, fullb, device)
            for l, sh in enumerate(self.shards):
                sh += [self.biases.b1[l].chunk(n)[self.rank], self.biases.b2[l].chunk(n)[self.rank]]
        self.full = [[torch.empty(F, D, dtype=torch.bfloat16, device=device), torch.empty(D, F, dtype=torch.bfloat16, device=device)]
                     for _ in range(2)]
        self.local_out = [(C.c_void_p * n)(*[w.data_ptr() if r == self.rank else 0 for r in range(n)]) for lp in self.shards for w in lp[:2]]
        shapes = [(F, D), (D, F)]
        self.chan = FusedChannel(self.sw, _owned_cap(shapes, n, OWN_ROW_SHARDED), max(_tiles(*s) for s in shapes))
        self.coll = CollChannel(self.sw)
        self.side = torch.cuda.Stream(device=device)
        self._pending = {}
        self.adam = cfg.optimizer == "adam"
        if self.adam:
            self.comm_name = "fused: dW GEMM + reduce-scatter + sharded Adam in one kernel; P2P-DMA all-gather prefetch"
            self.state = [[self.chan.adam_state(*shapes[k], OWN_ROW_SHARDED) for k in range(2)] for _ in self.shards]
        self.t = 0
        self.sw.host_barrier()

    def _update(self, i: int, k: int, dy: torch.Tensor, xin: torch.Tensor, M: int, N: int, T: int) -> None:
        """Rows ``[r*M/n, (r+1)*M/n)`` of ``dy^T xin`` summed over ranks and applied to MY shard - one kernel, no broadcast."""
        out, row0 = self.local_out[2 * i + k], self.rank * (M // self.world)
        if self.adam:
            self.chan.gemm_adam(dy, nv.MAJOR_MN, xin, nv.MAJOR_MN, M, N, T, out, N, row0, OWN_ROW_SHARDED, False, self.state[i][k], self.lr, self.t)
        else:
            self.chan.gemm(dy, nv.MAJOR_MN, xin, nv.MAJOR_MN, M, N, T, out, N, row0, OWN_ROW_SHARDED, FOP_SGD, False, self.lr)

    def _start_gather(self, layer: int) -> None:
        if not (0 <= layer < self.L) or layer in self._pending:
            return
        lib = nv.load()
        slot = layer % 2
        self.side.wait_stream(torch.cuda.current_stream())
        for k in range(2):
            e = self.coll.next_epoch()
            sh = self.shards[layer][k]
            # copy engines, not SMs: the GEMMs running meanwhile own every SM's registers and shared memory
            nv._check(lib.b200_all_gather_into_dma(self.full[slot][k].data_ptr(), self.sbuf.peers_c, self.shard_off[layer][k],
                                                   sh.numel() * 2, self.coll.flags.peers_c, self.rank, self.world, e,
                                                   self.side.cuda_stream), "b200_all_gather_into_dma")
            nv.launch_count += 2
        ev = torch.cuda.Event()
        ev.record(self.side)
        self._pending[layer] = _Handle(ev)

    def _finish_gather(self, layer: int) -> Params:
        self._pending.pop(layer).wait()
        return self.full[layer % 2]

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        from ..models.ffn import tlayer_ffn_fwd
        cfg, ops, L = self.cfg, self.ops, self.L
        D, F = cfg.model_size, cfg.ffn_dim
        T = x.shape[0]
        self.t += 1
        set_update_rounding(cfg, ops, self.t)
        acts = []
        y = x
        self._start_gather(0)
        params = None
        bias = self.biases
        for l in range(L):
            params = self._finish_gather(l)
            self._start_gather(l + 1)
            y, s = tlayer_ffn_fwd(ops, params + [bias.b1[l], bias.b2[l]] if bias else params, y, self.opts)
            acts.append(s)
        self._start_gather(L - 2)
        d = dloss_dx
        for i in reversed(range(L)):
            if i < L - 1:
                params = self._finish_gather(i)
                self._start_gather(i - 1)
            W1, W2 = params
            s = acts[i]
            a, pre = s.a, s.pre
            if a is None:
                a, pre = ops.linear_act(s.x, W1, self.opts.activation, bias.b1[i] if bias else None)
            dh = ops.grad_input(d, W2, act=self.opts.activation, aux=(pre if pre is not None else a) if self.opts.activation == "relu" else pre)
            if bias:
                bias.accumulate(i, dh, d)
            self._update(i, 1, d, a, D, F, T)
            dn = ops.grad_input(dh, W1, add=d if self.opts.residual else None) if i > 0 else None
            self._update(i, 0, dh, s.x, F, D, T)
            acts[i] = None
            d = dn
        if bias:
            bias.reduce_and_update(self.t)
        return y

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        lib = nv.load()
        out = []
        for l in range(self.L):
            fl = []
            for k in range(2):
                sh = self.shards[l][k]
                full = torch.empty((sh.shape[0] * self.world, sh.shape[1]), dtype=sh.dtype, device=sh.device)
                e = self.coll.next_epoch()
                nv._check(lib.b200_all_gather_into(full.data_ptr(), self.sbuf.peers_c, self.shard_off[l][k], sh.numel() * 2,
                                                   self.coll.flags.peers_c, self.rank, self.world, e, 0, nv._stream()), "b200_all_gather_into")
                fl.append(full)
            if self.biases:
                fl += [self.biases.b1[l].clone(), self.biases.b2[l].clone()]
            out.append(fl)
        return out


    # ---- optimizer state in the checkpoint's full layout: row-sharded owners = this rank's dim-0 shard -----------
    def _state_shapes(self):
        return [(self.cfg.ffn_dim, self.cfg.model_size), (self.cfg.model_size, self.cfg.ffn_dim)]

    def export_opt_state(self, m_out, v_out) -> int:
        if self.adam:
            for i in range(self.L):
                for k, (M, N) in enumerate(self._state_shapes()):
                    m, v = self.state[i][k]
                    _export_tiles(m, m_out[i][k], M, N, self.world, self.rank, OWN_ROW_SHARDED)
                    _export_tiles(v, v_out[i][k], M, N, self.world, self.rank, OWN_ROW_SHARDED)
            if self.biases:
                self.biases.export_state(m_out, v_out)
        return self.t

    def import_opt_state(self, m_in, v_in, t: int) -> None:
        self.t = int(t)
        if self.adam:
            for i in range(self.L):
                for k, (M, N) in enumerate(self._state_shapes()):
                    m, v = self.state[i][k]
                    _import_tiles(m, m_in[i][k], M, N, self.world, self.rank, OWN_ROW_SHARDED)
                    _import_tiles(v, v_in[i][k], M, N, self.world, self.rank, OWN_ROW_SHARDED)
            if self.biases:
                self.biases.import_state(m_in, v_in)


# =================================================================================================== MP
class FusedMPTrainer:
    """Megatron MP with the collectives inside the row-parallel GEMMs.

    ``mp_mode="ar"`` (reference form): activations replicated, ``y`` and ``dx`` all-reduced in the GEMM kernel.
    ``mp_mode="sp"`` (sequence-parallel form): activations live sharded over tokens ``[T/n, D]``; the row-parallel GEMM
    REDUCE-SCATTERS its output to the token-shard owners inside the kernel, and the next column-parallel GEMM's input is
    all-gathered from the peers' shards by the native pull kernel.  Same NVLink bytes (AR = RS + AG), 1/n of the saved
    activation memory."""

    comm_name = "fused: row-parallel GEMM + all-reduce in one kernel (peer stores over NVLink)"

    def __init__(self, ops, comm, cfg: TrainConfig, shards: List[Params], device: torch.device):
        self.ops, self.comm, self.cfg, self.shards = ops, comm, cfg, shards
        self.opts = block_opts(cfg)
        # Residual: the output buffer is pre-filled with the residual and the kernel runs its in-place update mode with
        # alpha = -1 (out = out + sum over ranks), i.e. the DDP / FSDP code path of the same kernel.
        self.residual = cfg.residual
        # dW is shard-local under MP (:311-312): SGD rides in the dW GEMM's epilogue; Adam = dW GEMM + the native Adam kernel.
        # Biases (fp32): b1 follows W1's row shard, b2 is replicated and enters the reduced output exactly once - the
        # output buffer is pre-filled with it (and the residual) before the kernel's in-place mode adds the cross-rank sum.
        self.opt = Optimizer(ops, cfg)
        self.adam = cfg.optimizer == "adam"
        self.has_bias = len(shards[0]) == 4
        self.sp = cfg.mp_mode == "sp"
        self.spf = cfg.mp_mode == "spf"
        if self.sp:
            self.comm_name = "fused: row-parallel GEMM + reduce-scatter in one kernel; native all-gather of the token shards"
        if self.spf:
            self.comm_name = "fused: all-gather + column-parallel GEMM in one kernel; row-parallel GEMM + reduce-scatter in one kernel"
        self.lr = cfg.lr * cfg.grad_scale
        self.sw = get_symm_world(comm, device)
        self.world, self.rank = comm.world, comm.rank
        self.L = len(shards)
        self.device = device
        self.T = 0
        self.chan: Optional[FusedChannel] = None

    def opt_key(self, i: int, j: int, p: torch.Tensor):
        """Key of parameter (layer i, index j) in ``self.opt.state`` (parallel/opt_state.py)."""
        return p.data_ptr()

    def update_local(self, W: torch.Tensor, dy: torch.Tensor, x: torch.Tensor) -> None:
        """``W -= update(dy^T @ x)`` on this rank's shard; no communication."""
        if not self.adam:
            self.ops.grad_weight_sgd_(W, dy, x, self.lr)
        else:
            self.opt.update(W.data_ptr(), W, self.ops.grad_weight(dy, x))

    def update_bias(self, layer: int, dh: torch.Tensor, dy: torch.Tensor) -> None:
        """b1 shard -= update(dh.sum(0)) (local, like dW1); b2 -= update(dy.sum(0)) (every rank computes the same sum)."""
        if not self.has_bias:
            return
        b1, b2 = self.shards[layer][2], self.shards[layer][3]
        self.opt.update(b1.data_ptr(), b1, nv.colsum(dh))
        self.opt.update(b2.data_ptr(), b2, nv.colsum(dy))

    def _b(self, layer: int, j: int) -> Optional[torch.Tensor]:
        return self.shards[layer][2 + j] if self.has_bias else None

    def _prefill(self, out: torch.Tensor, layer: Optional[int], residual: Optional[torch.Tensor]) -> bool:
        """out = b2 (+ residual) where either exists; returns whether the kernel must ADD its sum to `out`."""
        b2 = self._b(layer, 1) if layer is not None else None
        if b2 is None and residual is None:
            return False
        nv.bias_rows_(out, b2, residual)
        return True

    def _ensure_buffers(self, T: int) -> None:
        if self.T == T:
            return
        D, n = self.cfg.model_size, self.world
        if self.spf:
            if T % (TILE_M * n) or D % 64:
                raise ValueError(f"mp_mode=spf needs tokens ({T}) divisible by 128 * world ({TILE_M * n}) and model_size ({D}) by 64")
            # Activations live in FULL-size symmetric buffers; a rank's own token rows are written by the reduce-scatter
            # kernel, the other rows by the next all-gather+GEMM kernel, in place.
            self.abuf = self.sw.alloc((self.L + 3) * T * D * 2 + 16384)
            self.X, self.X_off = [], []
            for _ in range(self.L + 1):
                t, o = self.abuf.carve((T, D), torch.bfloat16)
                self.X.append(t)
                self.X_off.append(o)
            self.DX, self.DX_off = [], []
            for _ in range(2):
                t, o = self.abuf.carve((T, D), torch.bfloat16)
                self.DX.append(t)
                self.DX_off.append(o)
            self.local_tab = lambda t: (C.c_void_p * n)(*[t.data_ptr() if r == self.rank else 0 for r in range(n)])
            self.chan = FusedChannel(self.sw, _owned_cap([(T, D)], n, OWN_ROW_SHARDED), _tiles(T, D))
            self.gchan = GatherChannel(self.sw, T)
            self.coll = CollChannel(self.sw)
        elif self.sp:
            if T % (TILE_M * n):
                raise ValueError(f"sequence-parallel fused MP needs tokens ({T}) divisible by 128 * world ({TILE_M * n})")
            Ts = T // n
            self.abuf: SymmBuffer = self.sw.alloc((self.L + 3) * Ts * D * 2 + 16384)
            self.xs0, self.xs0_off = self.abuf.carve((Ts, D), torch.bfloat16)          # the input batch's token shard
            self.ys, self.ys_off = [], []
            for _ in range(self.L):
                t, o = self.abuf.carve((Ts, D), torch.bfloat16)
                self.ys.append(t)
                self.ys_off.append(o)
            self.dxs, self.dxs_off = [], []
            for _ in range(2):
                t, o = self.abuf.carve((Ts, D), torch.bfloat16)
                self.dxs.append(t)
                self.dxs_off.append(o)
            self.local_tab = lambda t: (C.c_void_p * n)(*[t.data_ptr() if r == self.rank else 0 for r in range(n)])
            self.chan = FusedChannel(self.sw, _owned_cap([(T, D)], n, OWN_ROW_SHARDED), _tiles(T, D))
            self.coll = CollChannel(self.sw)
        else:
            nvls = self.cfg.nvls != "off"
            self.abuf = self.sw.alloc((self.L + 2) * T * D * 2 + 8192, multicast=nvls)
            self.y, self.y_tab, self.y_mc = [], [], []
            for _ in range(self.L):
                t, o = self.abuf.carve((T, D), torch.bfloat16)
                self.y.append(t)
                self.y_tab.append(self.abuf.peers_at(o))
                self.y_mc.append(self.abuf.mc_at(o))
            self.dx, self.dx_tab, self.dx_mc = [], [], []
            for _ in range(2):
                t, o = self.abuf.carve((T, D), torch.bfloat16)
                self.dx.append(t)
                self.dx_tab.append(self.abuf.peers_at(o))
                self.dx_mc.append(self.abuf.mc_at(o))
            self.chan = FusedChannel(self.sw, _owned_cap([(T, D)], n, OWN_INTERLEAVED), _tiles(T, D), nvls=nvls and self.abuf.mc_ptr is not None)
            if self.chan.nvls:
                self.comm_name = "fused: row-parallel GEMM + all-reduce in one kernel, summed in the NVSwitch (multimem.ld_reduce / multimem.st)"
        self.T = T
        self.sw.host_barrier()

    # ---- reference (all-reduce) form -----------------------------------------------------------------------
    def _step_ar(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        ops = self.ops
        T, D = x.shape
        Fs = self.shards[0][0].shape[0]
        saved = []
        y = x
        for l, sh in enumerate(self.shards):
            W1, W2 = sh[0], sh[1]
            a, pre = ops.linear_act(y, W1, self.opts.activation, self._b(l, 0))
            saved.append((y, None if self.opts.recompute else a, None if self.opts.recompute else pre))
            # y_l = sum over ranks of a_r @ W2_r^T (+ b2 + x), delivered to every rank - one kernel (:302-303)
            if self._prefill(self.y[l], l, y if self.residual else None):
                self.chan.gemm(a, nv.MAJOR_K, W2, nv.MAJOR_K, T, D, Fs, self.y_tab[l], D, 0, OWN_INTERLEAVED, FOP_SGD, True, -1.0, out_mc=self.y_mc[l])
            else:
                self.chan.gemm(a, nv.MAJOR_K, W2, nv.MAJOR_K, T, D, Fs, self.y_tab[l], D, 0, OWN_INTERLEAVED, FOP_SUM, True, out_mc=self.y_mc[l])
            y = self.y[l]
        d = dloss_dx
        for i in reversed(range(self.L)):
            W1, W2 = self.shards[i][0], self.shards[i][1]
            xin, a, pre = saved[i]
            if a is None:
                a, pre = ops.linear_act(xin, W1, self.opts.activation, self._b(i, 0))
            dh = ops.grad_input(d, W2, act=self.opts.activation, aux=(pre if pre is not None else a) if self.opts.activation == "relu" else pre)
            self.update_local(W2, d, a)                                   # shard-local, no communication (:311-312)
            dn = None
            if i > 0:
                slot = i % 2
                # dx = sum over ranks of dh_r @ W1_r (+ d), delivered to every rank - one kernel (:308-309)
                if self._prefill(self.dx[slot], None, d if self.residual else None):
                    self.chan.gemm(dh, nv.MAJOR_K, W1, nv.MAJOR_MN, T, D, Fs, self.dx_tab[slot], D, 0, OWN_INTERLEAVED, FOP_SGD, True, -1.0, out_mc=self.dx_mc[slot])
                else:
                    self.chan.gemm(dh, nv.MAJOR_K, W1, nv.MAJOR_MN, T, D, Fs, self.dx_tab[slot], D, 0, OWN_INTERLEAVED, FOP_SUM, True, out_mc=self.dx_mc[slot])
                dn = self.dx[slot]
            self.update_local(W1, dh, xin)
            self.update_bias(i, dh, d)
            saved[i] = None
            d = dn
        return y

    # ---- sequence-parallel form ----------------------------------------------------------------------------
    def _gather_tokens(self, offset_bytes: int, T: int, D: int) -> torch.Tensor:
        """[T, D] <- every rank's [T/n, D] token shard at ``offset_bytes`` of the activation block (native pull)."""
        full = torch.empty(T, D, dtype=torch.bfloat16, device=self.device)
        e = self.coll.next_epoch()
        nv._check(nv.load().b200_all_gather_into(full.data_ptr(), self.abuf.peers_c, offset_bytes, (T // self.world) * D * 2,
                                                 self.coll.flags.peers_c, self.rank, self.world, e, 0, nv._stream()), "b200_all_gather_into")
        nv.launch_count += 1
        return full

    def _step_sp(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        ops, n, r = self.ops, self.world, self.rank
        T, D = x.shape
        Ts = T // n
        Fs = self.shards[0][0].shape[0]
        self.xs0.copy_(x[r * Ts:(r + 1) * Ts])
        saved = []
        cur_off = self.xs0_off
        cur_shard = self.xs0
        for l, sh in enumerate(self.shards):
            W1, W2 = sh[0], sh[1]
            xfull = self._gather_tokens(cur_off, T, D)                    # all-gather -> column-parallel GEMM
            a, pre = ops.linear_act(xfull, W1, self.opts.activation, self._b(l, 0))
            saved.append((cur_off, None if self.opts.recompute else a, None if self.opts.recompute else pre))
            # row-parallel GEMM -> reduce-scatter over token shards, one kernel: I keep rows [r*Ts, (r+1)*Ts)
            if self._prefill(self.ys[l], l, cur_shard if self.residual else None):
                self.chan.gemm(a, nv.MAJOR_K, W2, nv.MAJOR_K, T, D, Fs, self.local_tab(self.ys[l]), D, r * Ts, OWN_ROW_SHARDED, FOP_SGD, False, -1.0)
            else:
                self.chan.gemm(a, nv.MAJOR_K, W2, nv.MAJOR_K, T, D, Fs, self.local_tab(self.ys[l]), D, r * Ts, OWN_ROW_SHARDED, FOP_SUM, False)
            cur_off, cur_shard = self.ys_off[l], self.ys[l]
        # upstream gradient: same layout, token-sharded; stage my shard in a rotating symmetric slot
        d_slot = self.L % 2
        self.dxs[d_slot].copy_(dloss_dx[r * Ts:(r + 1) * Ts])
        d_off, d_shard = self.dxs_off[d_slot], self.dxs[d_slot]
        for i in reversed(range(self.L)):
            W1, W2 = self.shards[i][0], self.shards[i][1]
            x_off, a, pre = saved[i]
            dfull = self._gather_tokens(d_off, T, D)
            xfull = self._gather_tokens(x_off, T, D)
            if a is None:
                a, pre = ops.linear_act(xfull, W1, self.opts.activation, self._b(i, 0))
            dh = ops.grad_input(dfull, W2, act=self.opts.activation, aux=(pre if pre is not None else a) if self.opts.activation == "relu" else pre)
            self.update_local(W2, dfull, a)
            if i > 0:
                slot = i % 2
                if self._prefill(self.dxs[slot], None, d_shard if self.residual else None):
                    self.chan.gemm(dh, nv.MAJOR_K, W1, nv.MAJOR_MN, T, D, Fs, self.local_tab(self.dxs[slot]), D, r * Ts, OWN_ROW_SHARDED, FOP_SGD, False, -1.0)
                else:
                    self.chan.gemm(dh, nv.MAJOR_K, W1, nv.MAJOR_MN, T, D, Fs, self.local_tab(self.dxs[slot]), D, r * Ts, OWN_ROW_SHARDED, FOP_SUM, False)
                d_off, d_shard = self.dxs_off[slot], self.dxs[slot]
            self.update_local(W1, dh, xfull)
            self.update_bias(i, dh, dfull)
            saved[i] = None
        return self._gather_tokens(self.ys_off[self.L - 1], T, D)

    # ---- sequence-parallel form, gathers fused into the column-parallel GEMMs ------------------------------------
    # The schedule itself is `spf_step` (below the class); these are its three primitives on the native kernels.
    def ag_gemm(self, buf: torch.Tensor, W: torch.Tensor, b_major: int, kind: str, aux: Optional[torch.Tensor] = None,
                bias: Optional[torch.Tensor] = None):
        """``out = epilogue(all_gather(buf) @ W^T)`` in one kernel; ``buf`` (a full-size symmetric ``[T, D]`` buffer whose
        own token rows are final) is complete on return.  Returns ``(out, saved_for_backward)``."""
        T, D = buf.shape
        N = W.shape[0] if b_major == nv.MAJOR_K else W.shape[1]
        out = torch.empty(T, N, dtype=torch.bfloat16, device=self.device)
        peers = self.abuf.peers_at(self.abuf.offset_of(buf))
        if kind == "relu" and N % 128 == 0:
            mask = nv.relu_mask_empty(T, N, self.device)
            self.gchan.gemm(peers, W, b_major, out, T, N, D, nv.EPI_RELU_MASK, out2=mask, bias=bias)
            return out, mask
        if kind == "relu":
            self.gchan.gemm(peers, W, b_major, out, T, N, D, nv.EPI_RELU, bias=bias)
            return out, out
        if kind == "gelu":
            pre = torch.empty_like(out)
            self.gchan.gemm(peers, W, b_major, out, T, N, D, nv.EPI_GELU, out2=pre, bias=bias)
            return out, pre
        if kind == "relu_bwd":
            self.gchan.gemm(peers, W, b_major, out, T, N, D, nv.EPI_RELU_BWD_MASK if aux.dtype == torch.uint8 else nv.EPI_RELU_BWD, aux=aux)
            return out, None
        if kind == "gelu_bwd":
            self.gchan.gemm(peers, W, b_major, out, T, N, D, nv.EPI_GELU_BWD, aux=aux)
            return out, None
        raise ValueError(kind)

    def rs_gemm(self, A: torch.Tensor, W: torch.Tensor, b_major: int, dst: torch.Tensor, residual: Optional[torch.Tensor] = None,
                bias: Optional[torch.Tensor] = None) -> None:
        """``dst[own token rows] = sum over ranks of A @ W^T (+ bias + residual[own token rows])`` - row-parallel GEMM +
        reduce-scatter in one kernel."""
        T, D = dst.shape
        Ts = T // self.world
        rows = slice(self.rank * Ts, (self.rank + 1) * Ts)
        own = dst[rows]
        if residual is not None or bias is not None:
            nv.bias_rows_(own, bias, residual[rows] if residual is not None else None)
            self.chan.gemm(A, nv.MAJOR_K, W, b_major, T, D, A.shape[1], self.local_tab(own), D, self.rank * Ts, OWN_ROW_SHARDED, FOP_SGD, False, -1.0)
        else:
            self.chan.gemm(A, nv.MAJOR_K, W, b_major, T, D, A.shape[1], self.local_tab(own), D, self.rank * Ts, OWN_ROW_SHARDED, FOP_SUM, False)

    def gather_out(self, buf: torch.Tensor) -> torch.Tensor:
        """In-place all-gather of the token shards by the barrier-ed native kernel.  Its opening barrier is also what lets
        the next step overwrite X[0] / DX[0]: every rank has finished every kernel that read them."""
        e = self.coll.next_epoch()
        nv._check(nv.load().b200_all_gather(self.abuf.peers_c, self.abuf.offset_of(buf), buf.numel() * 2, self.coll.flags.peers_c,
                                            self.rank, self.world, e, 0, nv._stream()), "b200_all_gather")
        nv.launch_count += 1
        return buf.clone()                                                # the buffer is rewritten by the next step

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self._ensure_buffers(x.shape[0])
        self.opt.begin_step()
        if self.spf:
            return spf_step(self, x, dloss_dx)
        return self._step_sp(x, dloss_dx) if self.sp else self._step_ar(x, dloss_dx)

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        from .mp import MPTrainer
        return MPTrainer.full_params(self)      # same reassembly rule (W1 dim 0, W2 dim 1)


def spf_step(t, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
    """One ``mp_mode="spf"`` optimizer step, written against three primitives so that the schedule (which buffer, whose
    rows, which operand layout) is testable on CPU with gloo stand-ins (``tests/test_spf_schedule_cpu.py``):

    * ``t.ag_gemm(buf, W, b_major, kind, aux)``  all-gather ``buf``'s token shards, then ``epilogue(buf @ W^T)``
    * ``t.rs_gemm(A, W, b_major, dst, res)``     ``dst[own rows] = sum_ranks A @ W^T (+ res[own rows])``
    * ``t.gather_out(buf)``                      all-gather the final output

    ``t`` also provides ``shards, L, rank, world, ops, opts, update_local`` and the full-size activation buffers ``X[0..L]``,
    ``DX[0..1]``.  Per layer, forward: [all-gather x -> x W1^T -> act] one kernel, [a W2^T -> reduce-scatter] one
    kernel; backward: [all-gather dy -> dy W2 -> act'] one kernel, dW2 / dW1 GEMM+SGD local, [dh W1 -> reduce-scatter]
    one kernel.  Same math as the reference's TP loop (``train_ffns.py:296-312``) with AR split into RS + AG.
    Buffer-reuse argument: docs/KERNELS.md 2.5."""
    n, r = t.world, t.rank
    T = x.shape[0]
    Ts = T // n
    own = slice(r * Ts, (r + 1) * Ts)
    act = t.opts.activation
    t.X[0][own].copy_(x[own])
    saved = []
    bias = lambda l, j: t.shards[l][2 + j] if len(t.shards[l]) == 4 else None
    for l, sh in enumerate(t.shards):
        W1, W2 = sh[0], sh[1]
        a, aux = t.ag_gemm(t.X[l], W1, nv.MAJOR_K, act, bias=bias(l, 0))   # completes X[l] as a side effect
        t.rs_gemm(a, W2, nv.MAJOR_K, t.X[l + 1], t.X[l] if t.opts.residual else None, bias=bias(l, 1))   # my token rows of the next layer's input
        saved.append(None if t.opts.recompute else (a, aux))
    cur = 0
    t.DX[cur][own].copy_(dloss_dx[own])
    for i in reversed(range(t.L)):
        W1, W2 = t.shards[i][0], t.shards[i][1]
        if saved[i] is None:                                              # X[i] is complete since the forward pass
            a, pre = t.ops.linear_act(t.X[i], W1, act, bias(i, 0))
            aux = (pre if pre is not None else a) if act == "relu" else pre
        else:
            a, aux = saved[i]
        dh, _ = t.ag_gemm(t.DX[cur], W2, nv.MAJOR_MN, act + "_bwd", aux)  # completes DX[cur] = the full upstream gradient
        t.update_local(W2, t.DX[cur], a)
        if i > 0:
            t.rs_gemm(dh, W1, nv.MAJOR_MN, t.DX[1 - cur], t.DX[cur] if t.opts.residual else None)
        t.update_local(W1, dh, t.X[i])
        t.update_bias(i, dh, t.DX[cur])
        saved[i] = None
        cur = 1 - cur
    return t.gather_out(t.X[t.L])

"""CPU tests of the native library's scheduling maps (the library loads without a GPU): the tile rasterisation is a
bijection, and the fused kernels' tile -> (owner, inbox slot) map is a bijection that the owner inverts exactly."""
import ctypes as C

import pytest

from distributed_llm_code_samples_b200.ops import native as nv


@pytest.fixture(scope="module")
def lib():
    from distributed_llm_code_samples_b200.build_native import build
    build()
    l = C.CDLL(str(nv.lib_path()))
    l.b200_fused_owned_tiles.restype = C.c_int
    return l


@pytest.mark.parametrize("tiles_m,tiles_n,group_m", [(1, 1, 16), (3, 5, 16), (16, 4, 16), (33, 7, 16), (128, 64, 16), (5, 9, 4)])
def test_tile_rasterisation_visits_every_tile_once(lib, tiles_m, tiles_n, group_m):
    n = tiles_m * tiles_n
    order = (C.c_int * n)()
    assert lib.b200_debug_tile_order(tiles_m, tiles_n, group_m, order) == 0
    assert sorted(order) == list(range(n))
    # grouped: the first group_m * tiles_n tiles stay inside the first group_m rows of tiles
    head = list(order)[: min(group_m, tiles_m) * tiles_n]
    assert all(t // tiles_n < group_m for t in head)


@pytest.mark.parametrize("M,N", [(16384, 4096), (4096, 16384), (512, 1024), (1024, 256), (16384, 8192)])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("owner_mode", [0, 1])
def test_fused_owner_map_is_a_bijection(lib, M, N, world, owner_mode):
    tm, tn = (M + 127) // 128, (N + 255) // 256
    if owner_mode == 1 and M % (128 * world):
        pytest.skip("row-sharded owners need M % (128*world) == 0")
    cap = lib.b200_fused_owned_tiles(M, N, world, owner_mode)
    seen = set()
    out = (C.c_int * 4)()
    per_owner = [0] * world
    for m in range(tm):
        for n in range(tn):
            lib.b200_debug_fused_map(M, N, world, owner_mode, m, n, out)
            owner, slot, mb, nb = out[0], out[1], out[2], out[3]
            assert 0 <= owner < world and 0 <= slot < cap
            assert (mb, nb) == (m, n), "the owner must map its slot back to the tile that was pushed"
            assert (owner, slot) not in seen
            seen.add((owner, slot))
            per_owner[owner] += 1
            if owner_mode == 1:
                assert owner == (m * 128) // (M // world), "row-sharded: the owner holds that block of rows (FSDP dim-0 shard)"
    assert max(per_owner) - min(per_owner) <= 1, "ownership is balanced"


def test_argument_validation_without_a_gpu(lib):
    """Bad arguments are rejected on the host, before any CUDA call."""
    lib.b200_last_error.restype = C.c_char_p
    rc = lib.b200_gemm_bf16(None, C.c_longlong(8), 0, None, C.c_longlong(8), 0, None, C.c_longlong(8), 16, 16, 16, 0, None, C.c_longlong(0),
                            None, C.c_longlong(0), None, C.c_float(1.0), 0, None)
    assert rc == -1 and b"null" in lib.b200_last_error()
    assert lib.b200_sumsq_bf16_static(None, None, C.c_longlong(0), None) == -1


@pytest.mark.parametrize("M,K", [(1024, 64), (2048, 512), (16384, 8192), (4096, 192)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gather_gemm_pull_order_covers_every_remote_vector_once(lib, M, K, world):
    """csrc/gather_gemm.cu: the side warps' chunk list covers each remote 16-byte vector exactly once, in ring order
    starting at the next rank, and every 128-row block receives exactly `chunks_per_block` chunks (the count the TMA
    producer waits for)."""
    lib.b200_debug_gather_chunk.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_longlong, C.POINTER(C.c_longlong)]
    out = (C.c_longlong * 6)()
    for rank in range(world):
        assert lib.b200_debug_gather_chunk(M, K, rank, world, -1, out) == 0
        chunk_vecs, cpb, total = out[3], out[4], out[5]
        shard_vecs = (M // world) * K * 2 // 16
        assert chunk_vecs * 16 in (16384, 32768) and cpb * chunk_vecs * 16 == 128 * K * 2
        assert total == (shard_vecs // chunk_vecs) * (world - 1)
        per_block = {}
        covered = []
        last_dist = 1
        for u in range(total):
            lib.b200_debug_gather_chunk(M, K, rank, world, u, out)
            peer, first, block = out[0], out[1], out[2]
            assert peer != rank
            dist = (peer - rank) % world
            assert dist >= last_dist, "ring order: nearer successors first"
            last_dist = dist
            assert peer * shard_vecs <= first and first + chunk_vecs <= (peer + 1) * shard_vecs
            assert block == first * 16 // (128 * K * 2), "the chunk lies inside the block it is counted for"
            assert (first + chunk_vecs - 1) * 16 // (128 * K * 2) == block
            per_block[block] = per_block.get(block, 0) + 1
            covered.append(first)
        assert len(set(covered)) == total
        blocks = M // 128
        own = set(range(rank * blocks // world, (rank + 1) * blocks // world))
        assert set(per_block) == set(range(blocks)) - own
        assert all(v == cpb for v in per_block.values())


def test_gather_gemm_argument_validation_without_a_gpu(lib):
    lib.b200_last_error.restype = C.c_char_p
    pp = C.POINTER(C.c_void_p)
    vp, ll, i, f, u = C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_uint
    lib.b200_gather_gemm_bf16.argtypes = [pp, i, vp, ll, i, vp, ll, i, i, i, i, vp, ll, vp, ll, vp, f, pp, vp, i, i, u, i, vp]
    two = (C.c_void_p * 2)(0x1000, 0x2000)
    args = dict(M=1024, N=256, K=64, seq=1, world=2, rank=0)

    def call(**kw):
        a = dict(args, **kw)
        return lib.b200_gather_gemm_bf16(two, a.get("a_major", 0), 0x3000, 64, 0, 0x4000, a["N"], a["M"], a["N"], a["K"], 0, None, 0, None, 0,
                                         None, 1.0, two, 0x5000, a["rank"], a["world"], a["seq"], 0, None)
    assert call(M=1000) == -1 and b"128 * world" in lib.b200_last_error()
    assert call(K=72) == -1 and b"multiple of 64" in lib.b200_last_error()
    # K-sharded form: the sharded dimension is K, the row length is M
    assert call(a_major=1, K=320) == -1 and b"128 * world" in lib.b200_last_error()
    assert call(a_major=1, K=512, M=1000) == -1 and b"multiple of 64" in lib.b200_last_error()
    assert call(a_major=2) == -1
    assert call(seq=0) == -1 and b"seq" in lib.b200_last_error()
    assert call(rank=2) == -1


def test_fused_adam_entry_validation_without_a_gpu(lib):
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_fused_adam_state_bytes.restype = C.c_longlong
    assert lib.b200_fused_adam_state_bytes(3) == 3 * 128 * 256 * 4
    pp = C.POINTER(C.c_void_p)
    vp, ll, i, f, u = C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_uint
    lib.b200_gemm_allreduce_adam_bf16.argtypes = [vp, ll, i, vp, ll, i, i, i, i, pp, ll, i, pp, pp, i, i, i, i, u, i, i,
                                                  vp, vp, f, f, f, f, i, f, i, vp]
    two = (C.c_void_p * 2)(0x1000, 0x2000)
    # world=3 > table of 2 is never dereferenced: the rank/world and seq checks come first
    assert lib.b200_gemm_allreduce_adam_bf16(0x100, 8, 1, 0x200, 8, 1, 256, 256, 64, two, 256, 0, two, two, 4, 4, 5, 2, 1, 0, 1,
                                             0x300, 0x400, 1e-3, 0.9, 0.999, 1e-8, 1, 1.0, 0, None) == -1
    assert b"rank/world" in lib.b200_last_error()
    assert lib.b200_gemm_allreduce_adam_bf16(0x100, 8, 1, 0x200, 8, 1, 256, 256, 64, two, 256, 0, two, two, 4, 4, 0, 2, 0, 0, 1,
                                             0x300, 0x400, 1e-3, 0.9, 0.999, 1e-8, 1, 1.0, 0, None) == -1
    assert b"seq" in lib.b200_last_error()


@pytest.mark.parametrize("K,M", [(1024, 64), (4096, 512), (16384, 4096)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gather_gemm_k_sharded_walk_follows_the_pull_order(lib, K, M, world):
    """K-sharded form (FSDP backward): every tile's k-loop is a rotation of 0..num_kb-1 that starts at the rank's own
    shard, visits storage blocks in ring order, and that order is the order in which the side warps complete them."""
    lib.b200_debug_gather_kwalk.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    lib.b200_debug_gather_chunk.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_longlong, C.POINTER(C.c_longlong)]
    o = (C.c_int * 6)()
    c = (C.c_longlong * 6)()
    for rank in range(world):
        assert lib.b200_debug_gather_kwalk(K, rank, world, -1, o) == 0
        kb_rot, num_kb, bps = o[3], o[4], o[5]
        assert num_kb == K // 64 and kb_rot == rank * num_kb // world
        seen, block_order, last_ring = [], [], -1
        for kb in range(num_kb):
            lib.b200_debug_gather_kwalk(K, rank, world, kb, o)
            kbe, ring, block = o[0], o[1], o[2]
            seen.append(kbe)
            assert ring in (last_ring, last_ring + 1), "ring positions never jump: one integer of producer state is enough"
            last_ring = ring
            assert (ring < bps) == (block // bps == rank), "ring positions below blocks_per_shard are exactly my own shard"
            if not block_order or block_order[-1] != block:
                block_order.append(block)
        assert sorted(seen) == list(range(num_kb))
        # the side warps' pull order (storage rows = K, row length = M): blocks in order of first appearance
        lib.b200_debug_gather_chunk(K, M, rank, world, -1, c)
        pulled = []
        for u in range(c[5]):
            lib.b200_debug_gather_chunk(K, M, rank, world, u, c)
            if not pulled or pulled[-1] != c[2]:
                pulled.append(c[2])
        assert block_order[bps:] == pulled, "the k-loop needs remote blocks in exactly the order they are pulled"


@pytest.mark.parametrize("M,N", [(16384, 3584), (16384, 14336), (4096, 1024), (2048, 512)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gather_gemm_tile_walk_starts_in_the_own_shard(lib, M, N, world):
    """Row-sharded form: with the rotation and rasterisation group the launcher picks, a rank visits every pair tile once,
    its own shard's tiles first, then the shards in ring order - the order the side warps pull them in."""
    lib.b200_debug_gather_tile_walk.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    cap = ((M // 128 + 3) // 4) * ((N + 255) // 256)
    rows, nbs = (C.c_int * cap)(), (C.c_int * cap)()
    shard = M // world
    for rank in range(world):
        tiles = lib.b200_debug_gather_tile_walk(M, N, rank, world, rows, nbs, cap)
        assert tiles == cap
        assert len({(rows[i], nbs[i]) for i in range(tiles)}) == tiles
        if shard % 512:
            continue                           # a 512-row pair tile then spans two shards: any order is correct, none is ideal
        ring = [((rows[i] // shard) - rank) % world for i in range(tiles)]
        assert ring == sorted(ring), "own shard first, then the ring"
        assert ring[: (shard // 512) * ((N + 255) // 256)] == [0] * ((shard // 512) * ((N + 255) // 256))


@pytest.mark.parametrize("M,N,world,owner_mode", [(1024, 768, 2, 0), (1000, 520, 4, 0), (4096, 16384, 8, 0), (1024, 512, 4, 1), (16384, 4096, 8, 1)])
def test_python_owned_tile_list_matches_the_kernels_map(lib, M, N, world, owner_mode):
    """parallel/fused.py::_owned_tiles (used to move the tile-sharded Adam moments into and out of the checkpoint's full
    layout) must list, for every rank, exactly the (m_blk, n_blk) the kernel assigns to that rank's inbox slots, in slot order."""
    from distributed_llm_code_samples_b200.parallel.fused import _owned_tiles
    out = (C.c_int * 4)()
    tm, tn = (M + 127) // 128, (N + 255) // 256
    by_owner = {r: {} for r in range(world)}
    for m in range(tm):
        for n in range(tn):
            lib.b200_debug_fused_map(M, N, world, owner_mode, m, n, out)
            by_owner[out[0]][out[1]] = (m, n)
    for r in range(world):
        assert _owned_tiles(M, N, world, r, owner_mode) == [by_owner[r][k] for k in range(len(by_owner[r]))]


def test_adam_tile_state_round_trips_through_the_full_layout():
    """_export_tiles / _import_tiles: compact per-owned-tile moments -> dense full-layout tensor -> back, ragged edges included."""
    import torch
    from distributed_llm_code_samples_b200.parallel.fused import _export_tiles, _import_tiles, _owned_tiles
    M, N, world = 1000, 520, 4
    dense = torch.zeros(M, N)
    states = []
    for r in range(world):
        n_tiles = len(_owned_tiles(M, N, world, r, 0))
        st = torch.randn(n_tiles * 128 * 256)
        states.append(st)
        _export_tiles(st, dense, M, N, world, r, 0)
    assert (dense != 0).all(), "every element of the full layout is owned by exactly one rank"
    for r in range(world):
        back = torch.zeros_like(states[r])
        _import_tiles(back, dense, M, N, world, r, 0)
        again = torch.zeros(M, N)
        _export_tiles(back, again, M, N, world, r, 0)
        ref = torch.zeros(M, N)
        _export_tiles(states[r], ref, M, N, world, r, 0)
        assert torch.equal(again, ref)

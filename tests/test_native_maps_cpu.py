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

"""Property tests of the kernel's tile schedule, run on the CPU.

``ddlb_debug_schedule`` / ``ddlb_debug_tile_coords`` evaluate the SAME ``resolve()`` and
``tile_coord()`` the kernels use (``csrc/tp_gemm_host.inl``, ``csrc/tp_gemm_sm100.cuh``), so
these tests cover shapes, world sizes and options that were never launched on a GPU:
every tile is visited exactly once, rows map to the right shard/owner, chunks arrive in
the order the communication produces them, own rows come first (AG) or last (RS).
"""
import ctypes

import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from ddlb_b200 import _native as N

MODES = {"plain": N.MODE_PLAIN, "ag_before": N.MODE_AG_BEFORE, "ag_after": N.MODE_AG_AFTER,
         "rs": N.MODE_RS}


def _cfg(mode, world, rank, m, n, k, cta_group=0, block_n=0, group_m=0, chunk_rows=0, dtype=N.DTYPE_BF16,
         rs_transport=N.RS_PUSH, unrotated=False):
    cfg = N.TpGemmConfig()
    # fp32 operands are K-major for the tensor cores: B is [n,k]
    cfg.mode, cfg.dtype, cfg.b_layout = MODES[mode], dtype, (N.B_NK if dtype == N.DTYPE_F32 else N.B_KN)
    cfg.rs_transport, cfg.unrotated = rs_transport, int(unrotated)
    cfg.cta_group, cfg.block_n, cfg.group_m, cfg.num_sms = cta_group, block_n, group_m, 148
    cfg.rank, cfg.world, cfg.chunk_rows = rank, world, chunk_rows
    cfg.m, cfg.n, cfg.k = m, n, k
    return cfg


def _schedule(cfg):
    lib = N.lib()
    out = (ctypes.c_int32 * 12)()
    rc = lib.ddlb_debug_schedule(ctypes.byref(cfg), ctypes.byref(out))
    if rc != 0:
        raise N.NativeError(lib.ddlb_last_error().decode())
    keys = ("num_m_blocks", "num_n_blocks", "mblks_per_part", "chunk_mblks", "group_m", "tile_m",
            "block_n", "cta_group", "chunk_rows", "part0", "grid", "threads")
    return dict(zip(keys, list(out)))


def _tiles(cfg, sched):
    lib = N.lib()
    total = sched["num_m_blocks"] * sched["num_n_blocks"]
    buf = (ctypes.c_int32 * (4 * total))()
    N.check(lib.ddlb_debug_tile_coords(ctypes.byref(cfg), 0, total, buf), "tile_coords")
    return [tuple(buf[4 * i:4 * i + 4]) for i in range(total)]


shape_strategy = st.tuples(
    st.sampled_from(["plain", "ag_before", "ag_after", "rs"]),
    st.sampled_from([1, 2, 3, 4, 5, 8]),                 # world
    st.integers(1, 24),                                  # tile rows per rank (x tile_m)
    st.integers(1, 9),                                   # n in units of 64
    st.sampled_from([1, 2]),                             # cta_group
    st.sampled_from([0, 128, 256]),                      # block_n
    st.sampled_from([0, 1, 3, 8, 16]),                   # group_m
    st.integers(0, 6),                                   # chunk selector
    st.sampled_from(["push", "pull", "multimem"]),       # tp_rowwise transport
    st.booleans(),                                       # un-rotated shard order
    st.data(),
)


@settings(max_examples=300, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(shape_strategy)
def test_every_tile_once_and_rows_belong_to_their_part(args):
    mode, world, mult, n64, g, bn, group_m, chunk_sel, rs_t, unrot, data = args
    if mode == "plain":
        world = 1
    rank = data.draw(st.integers(0, world - 1))
    tile_m = 128 * g
    part_rows = mult * tile_m
    m, n = part_rows * world, 64 * n64
    k = 64 * world
    chunk_rows = 0
    if mode in ("ag_before", "rs") and chunk_sel:
        divisors = [d for d in range(1, mult + 1) if mult % d == 0]
        chunk_rows = divisors[chunk_sel % len(divisors)] * tile_m
    cfg = _cfg(mode, world, rank, m, n, k, cta_group=g, block_n=bn, group_m=group_m,
               chunk_rows=chunk_rows, unrotated=unrot,
               rs_transport={"push": N.RS_PUSH, "pull": N.RS_PULL, "multimem": N.RS_PULL_MULTICAST}[rs_t])
    sched = _schedule(cfg)
    tiles = _tiles(cfg, sched)
    nm, nn = sched["num_m_blocks"], sched["num_n_blocks"]
    assert sched["tile_m"] == tile_m and sched["cta_group"] == g
    # a permutation of the tile grid
    assert sorted((t[0], t[1]) for t in tiles) == [(i, j) for i in range(nm) for j in range(nn)]
    if mode in ("ag_before", "rs"):
        per_part = sched["mblks_per_part"]
        cm = sched["chunk_mblks"]
        assert per_part * world == nm and per_part % cm == 0
        for m_blk, n_blk, part, chunk in tiles:
            assert part == m_blk // per_part                       # rows belong to that shard
            assert chunk == (m_blk % per_part) // cm               # and to that chunk of it
        # m-blocks appear in arrival order: stage by stage, parts rotated from the rank
        first_seen = {}
        for idx, (m_blk, n_blk, part, chunk) in enumerate(tiles):
            first_seen.setdefault((chunk, part), idx)
        # first shard: AG the own one (already here), RS push the next rank (own last, its
        # partials need no NVLink); pull reductions and the un-rotated order start at 0
        if unrot or (mode == "rs" and rs_t != "push"):
            part0 = 0
        else:
            part0 = (rank + (1 if mode == "rs" else 0)) % world
        assert sched["part0"] == part0
        order = sorted(first_seen, key=first_seen.get)
        expected = [(c, (part0 + pi) % world) for c in range(per_part // cm) for pi in range(world)]
        # grouped rasterisation interleaves tiles of neighbouring chunks, but the first
        # appearance of each (chunk, part) follows the communication order
        assert order == expected
        if not unrot and mode == "rs" and rs_t == "push":
            # the rank's own rows are the last part of every stage
            assert [p for (c, p) in expected if c == 0][-1] == rank
        elif not unrot and mode == "ag_before":
            assert expected[0][1] == rank                           # own rows first
    else:
        assert all(t[2] == rank and t[3] == 0 for t in tiles)


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.sampled_from(["plain", "ag_before", "ag_after", "rs"]), st.sampled_from([1, 2, 4, 8]),
       st.integers(1, 12), st.integers(1, 6), st.sampled_from([0, 1, 3, 8]), st.booleans(), st.data())
def test_wide_tiles_cover_every_pair_of_n_blocks_once(mode, world, mult, n_pairs, group_m, unrot, data):
    """wide_n=1: the scheduled grid is m-blocks x PAIRS of n-blocks; shard / chunk bookkeeping of
    the multi-rank modes is the one of the ordinary schedule."""
    if mode == "plain":
        world = 1
    rank = data.draw(st.integers(0, world - 1))
    part_rows = mult * 256
    m, n, k = part_rows * world, 512 * n_pairs - data.draw(st.sampled_from([0, 64, 192])), 64 * world
    cfg = _cfg(mode, world, rank, m, n, k, cta_group=2, block_n=256, group_m=group_m, unrotated=unrot)
    cfg.wide_n = 1
    sched = _schedule(cfg)
    nm, nn = sched["num_m_blocks"], sched["num_n_blocks"]
    assert nn == 2 * n_pairs and sched["block_n"] == 256 and sched["cta_group"] == 2
    lib = N.lib()
    total = nm * n_pairs
    buf = (ctypes.c_int32 * (4 * total))()
    N.check(lib.ddlb_debug_tile_coords(ctypes.byref(cfg), 0, total, buf), "tile_coords")
    tiles = [tuple(buf[4 * i:4 * i + 4]) for i in range(total)]
    assert sorted((t[0], t[1]) for t in tiles) == [(i, j) for i in range(nm) for j in range(n_pairs)]
    one_more = (ctypes.c_int32 * 4)()
    assert lib.ddlb_debug_tile_coords(ctypes.byref(cfg), total, 1, one_more) != 0     # nothing beyond
    if mode in ("ag_before", "rs"):
        per_part, cm = sched["mblks_per_part"], sched["chunk_mblks"]
        for m_blk, _, part, chunk in tiles:
            assert part == m_blk // per_part and chunk == (m_blk % per_part) // cm


def test_headline_configs_resolve_as_documented():
    s = _schedule(_cfg("ag_before", 8, 3, 65536, 1024, 1024))
    assert (s["cta_group"], s["block_n"], s["tile_m"]) == (2, 256, 256)
    assert s["chunk_rows"] == 512                     # ~1 MiB of bf16 A per chunk at k = 1024
    assert s["threads"] == 384                        # aux warps 8..11 carry the all-gather
    assert s["num_m_blocks"] == 256 and s["num_n_blocks"] == 4 and s["grid"] == 148
    s = _schedule(_cfg("rs", 8, 0, 8192, 1024, 8192, dtype=N.DTYPE_F16))
    assert s["mblks_per_part"] == 4 and s["chunk_mblks"] == 4       # one chunk per owner by default
    s = _schedule(_cfg("ag_before", 8, 0, 16384, 28672, 8192))      # Llama-3-70B MLP up, TP=8
    assert s["num_n_blocks"] == 112 and s["chunk_rows"] % 256 == 0


def test_small_plain_problems_get_smaller_tiles():
    """Auto tile shape, single rank: halve the tile while one wave still holds all tiles."""
    s = _schedule(_cfg("plain", 1, 0, 1024, 1024, 1024))           # 4 x 8 = 32 pair tiles <= 74
    assert (s["cta_group"], s["block_n"], s["tile_m"]) == (2, 128, 256)
    s = _schedule(_cfg("plain", 1, 0, 2048, 1024, 1024))           # 8 x 8 = 64 pair tiles <= 74
    assert (s["cta_group"], s["block_n"], s["tile_m"]) == (2, 128, 256)
    s = _schedule(_cfg("plain", 1, 0, 4096, 1024, 1024))           # 64 full tiles fill one wave
    assert (s["cta_group"], s["block_n"], s["tile_m"]) == (2, 256, 256)
    s = _schedule(_cfg("plain", 1, 0, 1024, 1024, 1024, block_n=256))      # explicit shape wins
    assert (s["cta_group"], s["block_n"]) == (2, 256)
    s = _schedule(_cfg("ag_before", 2, 0, 1024, 1024, 1024))       # multi-rank schedules unchanged
    assert (s["cta_group"], s["block_n"]) == (2, 256)


def test_wide_tiles_resolve():
    """A scheduled tile covers two n-blocks; explicit request only."""
    cfg = _cfg("plain", 1, 0, 8192, 1024, 8192)
    assert len(set(_tiles_raw(cfg, 32 * 4))) == 32 * 4             # auto: 256 x 256 tiles
    cfg.wide_n = 1
    sched = _schedule(cfg)
    tiles = _tiles_raw(cfg, sched["num_m_blocks"] * sched["num_n_blocks"] // 2)
    assert sorted(tiles) == sorted((mb, nb) for mb in range(32) for nb in range(2))
    cfg = _cfg("plain", 1, 0, 1024, 512, 512)                      # small: auto would pick BLOCK_N 128
    assert _schedule(cfg)["block_n"] == 128
    cfg.wide_n = 1                                                 # the request keeps BLOCK_N 256
    assert _schedule(cfg)["block_n"] == 256 and len(set(_tiles_raw(cfg, 4))) == 4
    cfg = _cfg("ag_before", 2, 0, 8192, 1024, 8192)
    cfg.wide_n = 1
    assert len(set(_tiles_raw(cfg, 32 * 2))) == 32 * 2
    cfg = _cfg("plain", 1, 0, 8192, 768, 8192)                     # 3 n-blocks: not eligible
    assert len(set(_tiles_raw(cfg, 32 * 3))) == 32 * 3
    cfg.wide_n = 1
    with pytest.raises(N.NativeError, match="wide_n"):
        _schedule(cfg)


def _tiles_raw(cfg, count):
    lib = N.lib()
    out = (ctypes.c_int32 * (4 * count))()
    rc = lib.ddlb_debug_tile_coords(ctypes.byref(cfg), 0, count, out)
    if rc != 0:
        raise N.NativeError(lib.ddlb_last_error().decode())
    return [(out[4 * i], out[4 * i + 1]) for i in range(count)]


@pytest.mark.parametrize("kwargs,needle", [
    (dict(mode="rs", world=3, rank=0, m=300, n=64, k=192), "multiple of"),          # rows/rank % 128
    (dict(mode="ag_before", world=2, rank=0, m=512, n=100, k=64), "multiple of"),    # n % 64 (and % 8)
    (dict(mode="plain", world=1, rank=0, m=128, n=64, k=60), "multiple of 8"),       # 16-byte rows
    (dict(mode="rs", world=2, rank=0, m=512, n=64, k=65), "divisible"),              # k % world
    (dict(mode="plain", world=9, rank=0, m=1152, n=64, k=64), "rank/world"),
    (dict(mode="ag_before", world=2, rank=0, m=1024, n=64, k=64, chunk_rows=384), "chunk_rows"),
])
def test_bad_configs_are_rejected_with_a_reason(kwargs, needle):
    with pytest.raises(N.NativeError, match=needle):
        _schedule(_cfg(**kwargs))


def test_fp32_geometry_resolves():
    s = _schedule(_cfg("plain", 1, 0, 512, 96, 36, dtype=N.DTYPE_F32))   # n % 4, k % 4 suffice for [n,k]
    assert s["num_n_blocks"] == 1 and s["threads"] == 384                # 3xTF32 split warps
    cfg = _cfg("plain", 1, 0, 512, 96, 36, dtype=N.DTYPE_F32)
    cfg.b_layout = N.B_KN
    with pytest.raises(N.NativeError, match=r"\[n,k\]"):
        _schedule(cfg)


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.sampled_from([2, 3, 4, 8]), st.integers(1, 6), st.sampled_from([8, 72, 256, 1000, 1024]),
       st.sampled_from([1, 2]), st.integers(0, 5), st.data())
def test_comm_units_tile_every_peer_shard_exactly_once(world, mult, k, g, chunk_sel, data):
    """Default AG_before transport: the units produced by the kernel's ``pull_unit`` cover every
    byte of every peer's shard exactly once, in chunk order, and never touch the own shard."""
    rank = data.draw(st.integers(0, world - 1))
    tile_m = 128 * g
    part_rows = mult * tile_m
    chunk_rows = 0
    if chunk_sel:
        divisors = [d for d in range(1, mult + 1) if mult % d == 0]
        chunk_rows = divisors[chunk_sel % len(divisors)] * tile_m
    cfg = _cfg("ag_before", world, rank, part_rows * world, 64, k, cta_group=g, chunk_rows=chunk_rows)
    sched = _schedule(cfg)
    lib = N.lib()
    count = ctypes.c_int64(0)
    N.check(lib.ddlb_debug_pull_units(ctypes.byref(cfg), 0, None, ctypes.byref(count)), "units")
    buf = (ctypes.c_int64 * (5 * count.value))()
    N.check(lib.ddlb_debug_pull_units(ctypes.byref(cfg), count.value, buf, ctypes.byref(count)), "units")
    units = [tuple(buf[5 * i:5 * i + 5]) for i in range(count.value)]
    row_bytes = k * 2
    part_bytes = part_rows * row_bytes
    chunk_bytes = sched["chunk_rows"] * row_bytes
    covered = {}
    last_stage = -1
    for src, chunk, begin, end, base in units:
        assert src != rank and 0 <= src < world
        assert 0 <= begin < end <= chunk_bytes and begin % 16 == 0 and end % 16 == 0
        assert base == src * part_bytes + chunk * chunk_bytes
        assert chunk >= last_stage                 # stage-major: chunk j of all peers before j+1
        last_stage = chunk
        covered.setdefault(src, []).append((base + begin, base + end))
    assert sorted(covered) == [p for p in range(world) if p != rank]
    for src, spans in covered.items():
        spans.sort()
        assert spans[0][0] == src * part_bytes and spans[-1][1] == (src + 1) * part_bytes
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 == b0                        # contiguous, no gap, no overlap

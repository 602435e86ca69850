"""Host-side planning (algorithm choice, chunking, grid) checked without a GPU through
nbd_plan_query: every chunk must fit the staging half of its algorithm's layout, chunks must
tile the message exactly, be equal-sized and unit-aligned, and the size thresholds must hold."""
import pytest

from nbdistributed_b200 import capi

UNIT = 16
SIZES = [1, 15, 16, 4096, 32 << 10, (32 << 10) + 1, 1 << 20, (2 << 20) + 16, 25 << 20, (64 << 20) + 48, 256 << 20,
         (1 << 30) + 1000, 8 << 30]


def units(b):
    return (b + UNIT - 1) // UNIT


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("stage", [0, 8 << 20, 2 << 20])
@pytest.mark.parametrize("kind", ["all_reduce", "reduce_scatter", "all_gather", "broadcast"])
def test_chunks_tile_the_message_and_fit_the_staging_half(world, stage, kind):
    for nbytes in SIZES:
        p = capi.plan_query(world, kind, nbytes, stage_bytes=stage)
        half, chunk, n = p["half_bytes"], p["chunk_bytes"], p["nchunks"]
        assert half % UNIT == 0 and p["ll_off"] >= 2 * half
        assert 1 <= p["nblocks"] <= 512
        if (kind == "all_reduce" and p["algo"] == 3) or (kind == "broadcast" and p["algo"] == 2):
            limit = (256 << 10) if world <= 4 else (128 << 10)
            if stage and stage < (64 << 20):
                limit = 32 << 10  # small arenas carry the 1 MiB LL area only
            assert nbytes <= limit and n == 1  # LL: no staging halves involved
            continue
        assert chunk % UNIT == 0 and chunk > 0
        assert n * chunk >= nbytes > (n - 1) * chunk, (nbytes, chunk, n)
        if n > 1:  # equal chunks: the last one is not much shorter than the others
            assert nbytes - (n - 1) * chunk > chunk - UNIT * world * n
        if kind == "all_reduce" and p["algo"] == 1:      # one-shot: N slots of a whole chunk
            assert world * units(chunk) * UNIT <= half
        elif kind == "all_reduce":                         # two-shot: N shard slots + own reduced shard
            assert p["algo"] == 2
            assert (world + 1) * ((units(chunk) + world - 1) // world) * UNIT <= half
        elif kind == "reduce_scatter":
            assert world * units(chunk) * UNIT <= half
        elif kind == "all_gather":
            assert chunk <= half
        elif p["algo"] == 0:                               # broadcast direct
            assert chunk <= half
        else:                                              # broadcast scatter: N-1 shards
            assert ((units(chunk) + world - 2) // (world - 1)) * UNIT <= half


def test_algorithm_thresholds():
    assert capi.plan_query(8, "all_reduce", 4096)["algo"] == 3            # LL
    assert capi.plan_query(8, "all_reduce", 128 << 10)["algo"] == 3        # ... up to 128 KiB beyond four ranks
    assert capi.plan_query(8, "all_reduce", (128 << 10) + 16)["algo"] == 1  # one-shot
    assert capi.plan_query(4, "all_reduce", 256 << 10)["algo"] == 3        # ... up to 256 KiB up to four ranks
    assert capi.plan_query(4, "all_reduce", (256 << 10) + 16)["algo"] == 1
    assert capi.plan_query(8, "all_reduce", (32 << 10) + 16, stage_bytes=16 << 20)["algo"] == 1  # small arena: 32 KiB LL limit
    assert capi.plan_query(8, "all_reduce", 128 << 10)["nblocks"] == 16    # one unit per thread
    assert capi.plan_query(8, "all_reduce", 2 << 20)["algo"] == 1
    assert capi.plan_query(8, "all_reduce", (2 << 20) + 16)["algo"] == 2  # two-shot
    assert capi.plan_query(2, "all_reduce", 1 << 30)["algo"] == 1         # two ranks: always one-shot
    assert capi.plan_query(8, "broadcast", 4096)["algo"] == 2                # LL
    assert capi.plan_query(8, "broadcast", 128 << 10)["algo"] == 2
    assert capi.plan_query(8, "broadcast", (128 << 10) + 16)["algo"] == 0    # direct
    assert capi.plan_query(8, "broadcast", (32 << 10) + 16, stage_bytes=16 << 20)["algo"] == 0
    assert capi.plan_query(8, "broadcast", 256 << 10)["algo"] == 0
    assert capi.plan_query(8, "broadcast", (256 << 10) + 16)["algo"] == 1
    assert capi.plan_query(2, "broadcast", 1 << 30)["algo"] == 0
    assert capi.plan_query(2, "broadcast", 256 << 10)["algo"] == 2
    # a message that fits one chunk gets a chunk of its own (aligned) size: all CTAs have work
    p = capi.plan_query(8, "all_reduce", (16 << 20) + 4)
    assert p["nchunks"] == 1 and (16 << 20) + 4 <= p["chunk_bytes"] < (16 << 20) + 4 + 16 * 8
    # 256 MiB at 8 ranks: four equal 64 MiB steps (measured best), not 113 + 113 + 30
    p = capi.plan_query(8, "all_reduce", 256 << 20)
    assert p["nchunks"] == 4 and p["chunk_bytes"] == 64 << 20


def test_emulated_ranks_share_the_sms():
    one = capi.plan_query(8, "all_reduce", 256 << 20, emulated=False)["nblocks"]
    emu = capi.plan_query(8, "all_reduce", 256 << 20, emulated=True)["nblocks"]
    assert one == 148 and emu == 148 // 8


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("stage", [0, 8 << 20, 2 << 20])
def test_flat_table_chunks(world, stage):
    """k_bcast_flat: chunks are whole numbers of (N-1) units, one shard fits a staging half, the
    chunks cover the packed table, equal-sized."""
    nsh = world - 1
    for nbytes in [16 * k for k in (1, 7, 1000, 16385, 1 << 20, (40759296 + 15), 1 << 26)] + [652148736 + 16 * 149]:
        nbytes = units(nbytes) * UNIT  # packed tables are whole units
        p = capi.plan_query(world, "flat_table", nbytes, stage_bytes=stage)
        chunk, n, half = p["chunk_bytes"], p["nchunks"], p["half_bytes"]
        assert chunk % (UNIT * nsh) == 0 and chunk > 0
        assert n * chunk >= nbytes > (n - 1) * chunk
        assert (units(chunk) // nsh) * UNIT <= half
        assert 1 <= p["nblocks"] <= 148


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("stage", [0, 16 << 20])
@pytest.mark.parametrize("emulated", [False, True])
def test_tile_pipeline_geometry(world, stage, emulated):
    """plan_pipe: grid x slots x tile fits both staging halves, tiles are 128-byte multiples,
    grid x rounds x tile covers the message, the grid respects the flag rows of the B flags."""
    for nbytes in [(32 << 10) + 16, 100_001, 1 << 20, (5 << 20) + 3, 64 << 20, 256 << 20, 1 << 30]:
        p = capi.plan_query(world, "pipe", nbytes, stage_bytes=stage, emulated=emulated)
        tile, rounds, grid, depth = p["chunk_bytes"], p["nchunks"], p["nblocks"], p["algo"]
        assert tile % 128 == 0 and tile >= 128
        assert 2 <= depth <= 8
        assert grid * depth * tile <= 2 * p["half_bytes"]
        assert grid * rounds * tile >= nbytes
        assert grid * (rounds - 1) * tile < units(nbytes) * UNIT
        assert 1 <= grid <= (148 // world if emulated else 148)

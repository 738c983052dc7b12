"""CPU: bench.py's own parity gates (Expect) accept a correct decode of the tiled shard and reject a wrong byte, and the
batch split / shard arithmetic of the strong-scaling mode (BASELINE.json configs[4]) tiles the table exactly."""
import argparse

import numpy as np

import bench
import fls_writer as W
import shard
import synth
from oracle import oracle


def test_expect_accepts_the_tiled_decode_and_rejects_a_flipped_byte():
    small = synth.lineitem_image(3)
    img = W.tile_rowgroups(small, 8)
    ex = bench.Expect(small, threads=4)
    ora = oracle.OracleTable(img)
    assert ex.check_distinct([ora.decode(c, 0, 3, padded=True) for c in range(ora.n_cols)]) is None
    for g0, g1 in ((0, 8), (2, 7), (5, 6)):
        got = [ora.decode(c, g0, g1, padded=True) for c in range(ora.n_cols)]
        assert ex.check_batch(got, g0, g1) is None
    got = [ora.decode(c, 2, 7, padded=True) for c in range(ora.n_cols)]
    bad = [g if isinstance(g, tuple) else g.copy() for g in got]
    bad[3].view(np.uint8)[70_000 * 8 + 3] ^= 1
    assert "column 3" in ex.check_batch(bad, 2, 7)
    offs, pay = got[15]
    pay2 = pay.copy()
    pay2[len(pay2) // 2] ^= 0x20
    bad = list(got)
    bad[15] = (offs, pay2)
    assert "column 15" in ex.check_batch(bad, 2, 7)
    offs2 = offs.copy()
    offs2[100_000] += 1
    bad[15] = (offs2, pay)
    assert "column 15" in ex.check_batch(bad, 2, 7)


def test_batches_and_shards_tile_the_table():
    for n, cap in ((916, 916), (9156, 916), (1145, 916), (1, 916), (917, 916)):
        r = bench.batches_of(n, cap)
        assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
        assert max(b - a for a, b in r) <= cap
    a = argparse.Namespace(scaling="strong", rowgroups=9156)
    tot = 0
    for world in (1, 2, 4, 8):
        spans = [bench.shard_of(a, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(n for _, n in spans) == 9156
        assert all(spans[r][0] == shard.rowgroup_range(r, world, 9156)[0] for r in range(world))
    a = argparse.Namespace(scaling="weak", rowgroups=916)
    assert bench.shard_of(a, 3, 8) == (3 * 916, 916)

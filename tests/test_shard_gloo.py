"""Multi-GPU host logic on CPU: two `gloo` ranks shard one table by row-group range (shard.rowgroup_range, the
rule bench.py / FlsScanner::setRowGroupRange users follow), decode their range — here with the CPU oracle standing
in for the device — and the gathered pieces must reassemble the full scan in rank order with no collective on the
data path (only counters are reduced)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import shard
import synth
from oracle import oracle


def test_ranges_tile_the_table():
    for n_rg in (0, 1, 7, 916, 9156):
        for world in (1, 2, 3, 4, 8):
            rs = shard.all_ranges(world, n_rg)
            assert rs[0][0] == 0 and rs[-1][1] == n_rg
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.rowgroup_range(2, 2, 10)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, img_path, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    img = np.fromfile(img_path, dtype=np.uint8)
    ora = oracle.OracleTable(img)
    g0, g1 = shard.rowgroup_range(rank, world, ora.n_rg)
    rows = ora.rows(g0, g1)
    ints = ora.decode(0, g0, g1)
    offs, pay = ora.decode(4, g0, g1)
    np.save(os.path.join(out_dir, "ints%d.npy" % rank), ints)
    np.save(os.path.join(out_dir, "lens%d.npy" % rank), np.diff(offs.astype(np.int64)))
    np.save(os.path.join(out_dir, "pay%d.npy" % rank), pay)
    t = torch.tensor([rows, g1 - g0], dtype=torch.int64)
    dist.all_reduce(t)  # counters only
    assert int(t[0]) == ora.n_rows and int(t[1]) == ora.n_rg
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reassemble_the_scan(tmp_path):
    cols, names = synth.basic_usage(5 * 65536 + 777, seed=3)
    path = str(tmp_path / "t.fls")
    img = synth.encode(cols, names, path=path)
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), path, str(tmp_path)), nprocs=world, join=True)
    ora = oracle.OracleTable(img)
    full_ints = ora.decode(0)
    offs, pay = ora.decode(4)
    got_ints = np.concatenate([np.load(str(tmp_path / ("ints%d.npy" % r))) for r in range(world)])
    got_lens = np.concatenate([np.load(str(tmp_path / ("lens%d.npy" % r))) for r in range(world)])
    got_pay = np.concatenate([np.load(str(tmp_path / ("pay%d.npy" % r))) for r in range(world)])
    assert np.array_equal(got_ints, full_ints)
    assert np.array_equal(got_lens, np.diff(offs.astype(np.int64)))
    assert np.array_equal(got_pay, pay)

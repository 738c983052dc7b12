"""GPU: flsgpu_batch_copy_to_device — the optional result gather of SURVEY.md section 8e (the reference merges its
threads' DataChunks on the host, /root/reference/src/scanner/scan_fastlanes.cpp:89-120).  A shard decoded on one GPU
is copied into caller buffers on another (NVLink peer copies; the same device when the box has one GPU) and must
hold the oracle's bytes there."""
import numpy as np
import pytest
import torch

import flsgpu
import synth
from oracle import oracle

pytestmark = pytest.mark.gpu


def _gather(img, src, dst):
    ora = oracle.OracleTable(img)
    with flsgpu.Reader.from_image(img, device=src) as rd, rd.decode_batch(0, rd.n_rowgroups) as b:
        for i in range(b.n):
            c = b.columns[i]
            want = ora.decode(i)
            if c.width:
                vals = torch.empty(c.padded_rows * c.width, dtype=torch.uint8, device="cuda:%d" % dst)
                ms = b.copy_to_device(i, dst, vals.data_ptr())
                got = vals.cpu().numpy()[: c.rows * c.width]
                assert np.array_equal(got, want.view(np.uint8)), "column %d" % i
            else:
                vals = torch.empty(max(int(c.payload_bytes), 1), dtype=torch.uint8, device="cuda:%d" % dst)
                offs = torch.empty(c.padded_rows + 1, dtype=torch.int32, device="cuda:%d" % dst)
                ms = b.copy_to_device(i, dst, vals.data_ptr(), offs.data_ptr())
                o = offs.cpu().numpy().view(np.uint32)[: c.rows + 1]
                assert np.array_equal(o, want[0]), "column %d offsets" % i
                assert np.array_equal(vals.cpu().numpy()[: int(o[-1])], want[1]), "column %d payload" % i
            assert ms >= 0.0


def test_gather_to_the_last_device():
    n = torch.cuda.device_count()
    _gather(synth.lineitem_image(3, tail_rows=501), 0, n - 1)


def test_gather_back_to_device_0():
    n = torch.cuda.device_count()
    _gather(synth.lineitem_image(2, tail_rows=17, seed=3), n - 1, 0)


def test_bad_destination_is_refused():
    img = synth.lineitem_image(1)
    with flsgpu.Reader.from_image(img) as rd, rd.decode_batch(0, 1) as b:
        with pytest.raises(flsgpu.FlsError):
            b.copy_to_device(0, 99, 0)
        with pytest.raises(flsgpu.FlsError):
            b.copy_to_device(b.n + 3, 0, 0)

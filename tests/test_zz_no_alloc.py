"""GPU: flsgpu_select_fixed_noalloc (include/flsgpu.h) — the selection kernels of flsgpu_decode_filtered with their
scratch in a file-scope `static __device__` array instead of pooled cudaMalloc buffers.  No counterpart in the
reference; added for callers that cannot call cudaMalloc.  Checked through the C
ABI against flsgpu_decode_filtered (the existing entry point) on the same decoded column, bit for bit, on calls in a
row that share the scratch array (the second smaller than the first, so stale scratch would show).

This file sorts last on purpose."""
import numpy as np
import pytest
import torch

import flsgpu
import fls_writer as W

pytestmark = pytest.mark.gpu

N_ROWS = 1_500_000 + 333


def _table():
    r = np.random.default_rng(11)
    i64 = r.integers(-1 << 40, 1 << 40, N_ROWS, dtype=np.int64)
    f64 = np.round(r.normal(size=N_ROWS) * 100.0, 2)
    f64[r.integers(0, N_ROWS, 300)] = np.nan
    i32 = r.integers(-5000, 5000, N_ROWS).astype(np.int32)
    u8 = r.integers(0, 256, N_ROWS).astype(np.uint8)
    u64 = r.integers(0, 1 << 63, N_ROWS, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    f32 = r.normal(size=N_ROWS).astype(np.float32)
    return W.write_table([i64, f64, i32, u8, u64, f32], names=["i64", "f64", "i32", "u8", "u64", "f32"])


@pytest.fixture(scope="module")
def img():
    return _table()


def _select(rd, batch, col, op, const, n_rows, out_vals, out_ids, out_cnt, stream=None):
    """One call of the new entry point on rows [0, n_rows) of decoded column `col`; returns (ids, values as bytes)."""
    c = batch.columns[col]
    assert c.width and n_rows <= c.rows
    flsgpu.select_fixed_noalloc(0, c.data, c.type, n_rows, op, const, out_vals.data_ptr(), out_ids.data_ptr(),
                                out_cnt.data_ptr(), stream)
    torch.cuda.synchronize(0)
    n = int(out_cnt.cpu().numpy().view(np.uint64)[0])
    ids = out_ids[:n].cpu().numpy().view(np.uint64)
    vals = out_vals[: n * c.width].cpu().numpy()
    return ids, vals


def _expect(rd, col, op, const, n_rows):
    """The existing entry point (flsgpu_decode_filtered) on the same column, cut to rows [0, n_rows)."""
    ids, got = rd.decode_filtered([(col, op, const)], 0, rd.n_rowgroups, [col])
    keep = ids < n_rows
    return ids[keep], got[0][keep].view(np.uint8)


CALLS = [
    # (column, op, constant, rows): the first call fills the whole scratch range, the later ones use less of it
    (0, "<", 12345, N_ROWS),
    (1, ">=", 50.0, 700_001),
    (1, "=", float("nan"), N_ROWS),
    (2, "!=", 7, 1024),
    (3, "<=", 3, 65_537),
    (4, ">", (1 << 63) + 99, 1),
    (5, "<", -0.5, 999_999),
    (0, ">", 1 << 62, N_ROWS),       # nothing survives
    (2, ">=", -5000, 300_000),       # everything survives
]


def test_matches_decode_filtered_on_calls_in_a_row(img):
    with flsgpu.Reader.from_image(img) as rd, rd.decode_batch(0, rd.n_rowgroups) as b:
        out_vals = torch.empty(N_ROWS * 8, dtype=torch.uint8, device="cuda:0")
        out_ids = torch.empty(N_ROWS, dtype=torch.int64, device="cuda:0")
        out_cnt = torch.full((1,), -1, dtype=torch.int64, device="cuda:0")
        want = [_expect(rd, col, op, const, n) for (col, op, const, n) in CALLS]
        torch.cuda.synchronize(0)
        free0 = None
        for k, (col, op, const, n) in enumerate(CALLS):
            if k == 1:
                free0 = torch.cuda.mem_get_info(0)[0]
            ids, vals = _select(rd, b, col, op, const, n, out_vals, out_ids, out_cnt)
            assert np.array_equal(ids, want[k][0]), "call %d: row ids" % k
            assert np.array_equal(vals, want[k][1]), "call %d: values" % k
        assert len(want[0][0]) > 0 and len(want[7][0]) == 0 and len(want[8][0]) == 300_000
        # no device memory was taken across the calls after the first (the test's own .cpu() copies land in host memory)
        assert torch.cuda.mem_get_info(0)[0] >= free0


def test_two_identical_calls_give_the_same_answer_and_streams_work(img):
    with flsgpu.Reader.from_image(img) as rd, rd.decode_batch(0, rd.n_rowgroups, cols=[2]) as b:
        out_vals = torch.empty(N_ROWS * 4, dtype=torch.uint8, device="cuda:0")
        out_ids = torch.empty(N_ROWS, dtype=torch.int64, device="cuda:0")
        out_cnt = torch.zeros(1, dtype=torch.int64, device="cuda:0")
        want = _expect(rd, 2, "<", -4000, N_ROWS)
        s = torch.cuda.Stream(device=0)
        torch.cuda.synchronize(0)
        for _ in range(2):
            out_cnt.fill_(-1)
            torch.cuda.synchronize(0)
            ids, vals = _select(rd, b, 0, "<", -4000, N_ROWS, out_vals, out_ids, out_cnt, stream=s.cuda_stream)
            assert np.array_equal(ids, want[0]) and np.array_equal(vals, want[1])


def test_empty_input_and_optional_outputs(img):
    with flsgpu.Reader.from_image(img) as rd, rd.decode_batch(0, 1, cols=[0]) as b:
        c = b.columns[0]
        out_cnt = torch.full((1,), 77, dtype=torch.int64, device="cuda:0")
        flsgpu.select_fixed_noalloc(0, None, c.type, 0, "=", 0, None, None, out_cnt.data_ptr())
        torch.cuda.synchronize(0)
        assert int(out_cnt.item()) == 0
        ids, got = rd.decode_filtered([(0, "<", 0)], 0, 1, [0])
        flsgpu.select_fixed_noalloc(0, c.data, c.type, c.rows, "<", 0, None, None, out_cnt.data_ptr())
        torch.cuda.synchronize(0)
        assert int(out_cnt.item()) == len(ids)


def test_sizes_beyond_the_scratch_array_are_refused(img):
    with flsgpu.Reader.from_image(img) as rd, rd.decode_batch(0, 1, cols=[0]) as b:
        c = b.columns[0]
        out_cnt = torch.full((1,), 5, dtype=torch.int64, device="cuda:0")
        with pytest.raises(flsgpu.FlsError) as e:
            flsgpu.select_fixed_noalloc(0, c.data, c.type, flsgpu.NOALLOC_MAX_ROWS + 1, "<", 0, None, None, out_cnt.data_ptr())
        assert e.value.code == 5 and "FLSGPU_NOALLOC_MAX_ROWS" in str(e.value)
        with pytest.raises(flsgpu.FlsError):
            flsgpu.select_fixed_noalloc(0, c.data, 12, 10, "<", 0, None, None, out_cnt.data_ptr())  # STR: not fixed width
        with pytest.raises(flsgpu.FlsError):
            flsgpu.select_fixed_noalloc(0, c.data, c.type, 10, 9, 0, None, None, out_cnt.data_ptr())  # bad comparison
        with pytest.raises(flsgpu.FlsError):
            flsgpu.select_fixed_noalloc(0, c.data, c.type, 10, "<", 0, None, None, None)
        torch.cuda.synchronize(0)
        assert int(out_cnt.item()) == 5  # nothing was launched

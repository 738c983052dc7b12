"""GPU: damaged files through libflsgpu itself.  Every byte flip inside the vector-header tables, the segments or the
footer must end in FLSGPU_E_FORMAT (the host gate, csrc/fls_validate.h) or in a decode that completes; never in a
CUDA fault — on a shared pool an Xid takes the GPU down for everyone.  The context is checked after every damaged
image by decoding the pristine one bit-exactly.  Run under `compute-sanitizer --tool memcheck` for the out-of-bounds
proof (profiles/r02_memcheck_fuzz.log); the CPU twin (tests/test_footer_fuzz.py) runs the same kernel source under
AddressSanitizer.  The reference has no twin: it trusts FastLanes' reader (/root/reference/src/fastlanes_facade.cpp:34)."""
import numpy as np
import pytest

import flsgpu
import synth
from oracle import oracle
from test_footer_fuzz import zoo_image

pytestmark = pytest.mark.gpu


def _segments(img):
    """(file offset, size, is_header_table) of every segment, from the oracle's own footer parser."""
    ora = oracle.OracleTable(img)
    return ora.segments()


def _decode_all(x):
    with flsgpu.Reader.from_image(x) as rd:
        rd.decode()


def _fuzz(img, mode, seed, n):
    rng = np.random.default_rng(seed)
    size = len(img)
    foot_off = int(img[size - 24:size - 16].view(np.uint64)[0])
    segs = _segments(img)
    hdrs = [s for s in segs if s[2]]
    ok = refused = 0
    with flsgpu.Reader.from_image(img) as good:
        want = good.decode()
        for it in range(n):
            x = img.copy()
            if mode == "hdr":
                for _ in range(int(rng.integers(1, 5))):
                    off, sz, _h = hdrs[int(rng.integers(0, len(hdrs)))]
                    at = off + int(rng.integers(0, sz))
                    if rng.integers(0, 2):
                        at = (at & ~15) + int(rng.integers(0, 8))  # ep / meta: the fields that form addresses
                    x[at] = int(rng.integers(0, 256))
            elif mode == "segs":
                off, sz, _h = segs[int(rng.integers(0, len(segs)))]
                for _ in range(int(rng.integers(1, 9))):
                    span = min(sz, 256) if rng.integers(0, 2) else sz
                    x[off + int(rng.integers(0, span))] = 0xFF if rng.integers(0, 4) == 0 else int(rng.integers(0, 256))
            else:
                for _ in range(int(rng.integers(1, 4))):
                    x[int(rng.integers(foot_off, size - 24))] = int(rng.integers(0, 256))
            try:
                _decode_all(x)
                ok += 1
            except flsgpu.FlsError as e:
                assert e.code in (flsgpu.E_FORMAT, flsgpu.E_LIMIT, flsgpu.E_NOMEM), "damaged image: %s (code %d)" % (e, e.code)
                refused += 1
            if it % 8 == 7 or it == n - 1:  # the device is still healthy and still exact
                got = good.decode()
                for a, b in zip(got, want):
                    if isinstance(a, tuple):
                        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
                    else:
                        assert np.array_equal(a.view(np.uint8), b.view(np.uint8))
    return ok, refused


def test_damaged_vector_headers_never_fault(tmp_path):
    img = np.fromfile(zoo_image(str(tmp_path / "zoo.fls")), dtype=np.uint8)
    ok, refused = _fuzz(img, "hdr", 21, 150)
    assert refused > 0 and ok + refused == 150


def test_damaged_segments_never_fault(tmp_path):
    img = np.fromfile(zoo_image(str(tmp_path / "zoo.fls")), dtype=np.uint8)
    ok, refused = _fuzz(img, "segs", 22, 200)
    assert ok + refused == 200


def test_damaged_footer_never_faults(tmp_path):
    img = np.fromfile(zoo_image(str(tmp_path / "zoo.fls")), dtype=np.uint8)
    ok, refused = _fuzz(img, "footer", 23, 200)
    assert refused > 0


def test_damaged_lineitem_shape_never_faults():
    img = synth.lineitem_image(1, tail_rows=777)
    _fuzz(img, "hdr", 24, 60)
    _fuzz(img, "segs", 25, 80)

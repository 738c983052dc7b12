"""CPU: libflsgpu.so loads, exports every symbol include/flsgpu.h declares, and the non-compute half of the ABI
(open / schema / chunk info / error text) behaves like the reference's bind.  No compute calls without a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

import flsgpu
import fls_writer as W
import synth


def test_exports_every_declared_symbol():
    syms = flsgpu.header_symbols()
    assert len(syms) >= 20 and "flsgpu_decode" in syms and "flsgpu_open" in syms
    L = C.CDLL(flsgpu.lib_path())
    for s in syms:
        assert hasattr(L, s), s
    assert flsgpu.lib().flsgpu_abi_version() == 1


def test_missing_file_error_text_matches_reference():
    # /root/reference/test/sql/fastlane.test:9-12
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Reader.open("nonexistent.fls")
    assert "Failed to open FastLanes file: nonexistent.fls" in str(e.value)
    assert e.value.code == flsgpu.E_IO


def test_not_an_fls_file(tmp_path):
    p = tmp_path / "junk.fls"
    p.write_bytes(b"x" * 4096)
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Reader.open(str(p))
    assert e.value.code == flsgpu.E_FORMAT and "Failed to open FastLanes file" in str(e.value)


def test_schema_from_footer(tmp_path):
    cols, names = synth.basic_usage(70_000)
    p = str(tmp_path / "basic.fls")
    img = synth.encode(cols, names, synth.basic_usage_types(), path=p)
    for rd in (flsgpu.Reader.open(p), flsgpu.Reader.from_image(img)):
        with rd:
            assert rd.names == names
            assert rd.types == synth.basic_usage_types()
            assert rd.n_rows == 70_000 and rd.n_rowgroups == 2
            assert rd.rowgroup_rows(0) == 65536 and rd.rowgroup_rows(1) == 70_000 - 65536
            s, b = rd.chunk_info(0, 0)
            assert s == W.DELTA and b > 0
            with pytest.raises(flsgpu.FlsError):
                rd.chunk_info(9, 0)


def test_decode_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    img = synth.encode([np.arange(5000, dtype=np.int32)])
    with flsgpu.Reader.from_image(img) as rd:
        with pytest.raises(flsgpu.FlsError) as e:
            rd.decode()
        assert e.value.code == flsgpu.E_CUDA and "CUDA" in str(e.value)


def test_product_does_not_link_the_oracle():
    out = os.popen("ldd %s; nm -D --defined-only %s" % (flsgpu.lib_path(), flsgpu.lib_path())).read()
    assert "fls_oracle" not in out and "flso_" not in out and "emul_" not in out

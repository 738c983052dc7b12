"""CPU: the footer parser and the header-only bound reader never crash on damaged files.  A table function's bind
runs on whatever bytes the user points it at (/root/reference/src/scan_fastlanes.cpp:36-45 opens the file before
anything else), so every truncation and every byte flip must end in a clean error (or in a file that still parses),
never in a fault.  The damaged images are opened in a child process: a crash there is a failed test here, not a dead
test run."""
import os
import subprocess
import sys

import numpy as np

import synth

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

CHILD = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import flsgpu
img = np.fromfile(sys.argv[3], dtype=np.uint8)
mode, seed, n = sys.argv[4], int(sys.argv[5]), int(sys.argv[6])
rng = np.random.default_rng(seed)
size = len(img)
foot_off = int(img[size - 24:size - 16].view(np.uint64)[0])
ok = bad = 0
for it in range(n):
    x = img.copy()
    if mode == "truncate":
        x = x[: int(rng.integers(0, size))]
    elif mode == "footer":
        for _ in range(int(rng.integers(1, 4))):
            x[int(rng.integers(foot_off, size))] = int(rng.integers(0, 256))
    elif mode == "footer_wide":  # 8 aligned bytes replaced by a large value: offsets and sizes that point far away
        at = foot_off + 8 * int(rng.integers(0, (size - 24 - foot_off) // 8))
        x[at:at + 8] = np.frombuffer(np.uint64(int(rng.integers(0, 2**63))).tobytes(), dtype=np.uint8)
    else:  # headers of the row groups: the bound reader parses them on the host
        for _ in range(8):
            x[int(rng.integers(16, foot_off))] = int(rng.integers(0, 256))
    try:
        with flsgpu.Reader.from_image(x) as rd:
            for g in range(min(rd.n_rowgroups, 4)):
                rd.rowgroup_rows(g)
                for c in range(rd.n_cols):
                    rd.chunk_info(g, c)
                    rd.rowgroup_minmax(g, c)
        ok += 1
    except flsgpu.FlsError:
        bad += 1
print("done", ok, bad)
"""


def _image(tmp_path):
    cols, names = synth.basic_usage(3 * 65536 + 1000)  # integers, doubles and strings; DELTA, FFOR, ALP, string schemes
    p = str(tmp_path / "ok.fls")
    synth.encode(cols, names, synth.basic_usage_types(), path=p)
    return p


def _run(path, mode, seed, n):
    r = subprocess.run([sys.executable, "-c", CHILD, os.path.join(ROOT, "duckdb-fastlane_b200"), ROOT, path, mode, str(seed), str(n)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, "%s: the reader crashed (rc %d)\n%s" % (mode, r.returncode, r.stderr[-2000:])
    tag, ok, bad = r.stdout.split()[-3:]
    assert tag == "done"
    return int(ok), int(bad)


def test_truncated_files_are_errors(tmp_path):
    ok, bad = _run(_image(tmp_path), "truncate", 1, 200)
    assert ok == 0 and bad == 200


def test_flipped_footer_bytes_never_crash(tmp_path):
    ok, bad = _run(_image(tmp_path), "footer", 2, 400)
    assert ok + bad == 400 and bad > 0


def test_wild_footer_offsets_never_crash(tmp_path):
    ok, bad = _run(_image(tmp_path), "footer_wide", 3, 400)
    assert ok + bad == 400 and bad > 0


def test_flipped_header_bytes_never_crash_the_bound_reader(tmp_path):
    ok, bad = _run(_image(tmp_path), "headers", 4, 200)
    assert ok + bad == 200


# ---- the decode path: validation gate (csrc/fls_validate.h) + the device decoders' own clamps, under ASAN ----------
# tests/emul/fls_fuzz_asan compiles the exact kernel source for the host (FLS_HOST_EMUL) with AddressSanitizer and
# decodes every column of damaged images into heap blocks of exactly the size libflsgpu would hand the kernels.
# Whatever is damaged — vector headers, segment contents, footer fields — the decode must end in a clean refusal or
# stay inside its buffers; an out-of-bounds access aborts the child and fails the test.  The GPU twin of this test
# (tests/test_gpu_fuzz.py) drives libflsgpu itself and expects FLSGPU_E_FORMAT or a decode, never a CUDA fault.
FUZZ = os.path.join(HERE, "emul", "fls_fuzz_asan")


def zoo_image(path, n=70_000):
    """One column per scheme the lineitem / basic_usage shapes do not reach."""
    import cases
    import fls_writer as W
    pick = {"rle_i32": None, "dict_i64": None, "alprd_with_specials": None, "alp_many_exceptions": None,
            "str_uncompressed_comment": None, "str_fsst_dict_shipmode": None, "str_dict_500_entries": None,
            "str_fsst_random_0_40": None, "uncompressed_f64": None, "constant_i16": None, "delta_i16": None,
            "str_dict_uniform_len2": None, "str_fsst_dict_500_entries": None}
    for c in cases.delta_cases(n) + cases.rle_cases(n) + cases.dict_cases(n) + cases.misc_fixed_cases(n) + cases.alp_cases(n) + \
            cases.string_cases(n):
        if c[0] in pick:
            pick[c[0]] = c
    cols = [pick[k][1][0] for k in pick]
    force = [pick[k][3][0] for k in pick]
    W.write_table(cols, force=force).tofile(path)
    return path


def _fuzz(path, mode, seed, n):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0:allocator_may_return_null=1")
    r = subprocess.run([FUZZ, path, mode, str(seed), str(n)], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, "%s: out-of-bounds access or crash in the decode path (rc %d)\n%s" % (mode, r.returncode, r.stderr[-4000:])
    tag, ok, refused, unparsed = r.stdout.split()[-4:]
    assert tag == "done" and int(ok) + int(refused) + int(unparsed) == n
    return int(ok), int(refused), int(unparsed)


def test_decode_of_damaged_vector_headers_stays_in_bounds(tmp_path):
    ok, refused, _ = _fuzz(_image(tmp_path), "hdr", 11, 120)
    assert refused > 0  # entry points / widths / counts that leave their segment are refused by the gate
    ok, refused, _ = _fuzz(zoo_image(str(tmp_path / "zoo.fls")), "hdr", 12, 120)
    assert refused > 0


def test_decode_of_damaged_segments_stays_in_bounds(tmp_path):
    zoo = zoo_image(str(tmp_path / "zoo.fls"))
    _fuzz(zoo, "segs", 13, 250)
    _fuzz(_image(tmp_path), "segs", 14, 150)
    _fuzz(zoo, "data", 15, 60)


def test_decode_of_damaged_footers_stays_in_bounds(tmp_path):
    zoo = zoo_image(str(tmp_path / "zoo.fls"))
    ok, refused, unparsed = _fuzz(zoo, "footer", 16, 400)
    assert unparsed > 0
    _fuzz(zoo, "wide", 17, 400)


def test_decode_of_damaged_lineitem_shape_stays_in_bounds(tmp_path):
    p = str(tmp_path / "li.fls")
    synth.lineitem_image(1, tail_rows=777).tofile(p)
    _fuzz(p, "hdr", 18, 60)
    _fuzz(p, "segs", 19, 80)

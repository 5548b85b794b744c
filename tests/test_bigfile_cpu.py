"""CPU tests (no GPU) of the bigfile snapshot / IC reader of the C ABI (shenqi_b200/csrc/sq_bigfile.cu, host
code): pinned to files written by the reference's OWN vendored bigfile library (depends/bigfile/src/bigfile.c compiled
unmodified into oracle/_ref, driven by oracle/ref_bigfile.c the way petaio.cpp writes snapshots)."""
import ctypes
import os
import numpy as np
import pytest
from helpers import ref

pytestmark = pytest.mark.skipif(ref() is None, reason="reference build (oracle/_ref) not present")


def _write(path, block, arr, nfile):
    R = ref()
    arr = np.ascontiguousarray(arr)
    dt = {np.dtype("f8"): b"f8", np.dtype("f4"): b"f4", np.dtype("u8"): b"u8", np.dtype("i4"): b"i4", np.dtype("u1"): b"u1"}[arr.dtype]
    nmemb = 1 if arr.ndim == 1 else arr.shape[1]
    rc = R.ref_bigfile_write_block(path.encode(), block.encode(), dt, ctypes.c_int(nmemb), arr.ctypes.data_as(ctypes.c_void_p),
                                   ctypes.c_int64(len(arr)), ctypes.c_int(nfile))
    assert rc == 0, rc


def _attr(path, block, name, values, dt):
    R = ref()
    v = np.ascontiguousarray(values, {"f8": np.float64, "u8": np.uint64, "i4": np.int32}[dt])
    rc = R.ref_bigfile_set_attr(path.encode(), block.encode(), name.encode(), v.ctypes.data_as(ctypes.c_void_p), dt.encode(), ctypes.c_int(len(v)))
    assert rc == 0, rc


@pytest.mark.parametrize("nfile", [1, 3, 7])
def test_blocks_written_by_the_reference_library(tmp_path, nfile):
    import shenqi_b200 as sq
    path = str(tmp_path / "PART_000")
    rng = np.random.default_rng(nfile)
    n = 10007
    pos = rng.random((n, 3)) * 64000.0
    mass = (0.03 + 0.01 * rng.random(n)).astype(np.float32)
    ids = rng.integers(1, 2 ** 62, n).astype(np.uint64)
    tb = rng.integers(0, 46, n).astype(np.uint8)
    _write(path, "1/Position", pos, nfile)
    _write(path, "1/Mass", mass, nfile)
    _write(path, "1/ID", ids, nfile)
    _write(path, "1/TimeBinGravity", tb, 1)
    dt, nmemb, nrows, nf = sq.bigfile_block_info(path, "1/Position")
    assert (dt.lstrip("<=|"), nmemb, nrows, nf) == ("f8", 3, n, nfile)
    assert np.array_equal(sq.bigfile_read_block(path, "1/Position"), pos)                         # bit for bit
    assert np.array_equal(sq.bigfile_read_block(path, "1/Mass", want="f4")[:, 0], mass)
    assert np.array_equal(sq.bigfile_read_block(path, "1/Mass", want="f8")[:, 0], mass.astype(np.float64))   # widening conversion
    assert np.array_equal(sq.bigfile_read_block(path, "1/ID", want="u8")[:, 0], ids)
    assert np.array_equal(sq.bigfile_read_block(path, "1/TimeBinGravity", want="u1")[:, 0], tb)
    # ranges that start and end inside physical files
    for first, count in ((0, 1), (max(0, n // nfile - 1), 3 if nfile > 1 else 1), (1234, 5000), (n - 1, 1), (n, 0)):
        assert np.array_equal(sq.bigfile_read_block(path, "1/Position", first, count), pos[first:first + count])
    with pytest.raises(sq.ShenqiError):
        sq.bigfile_read_block(path, "1/Position", n - 1, 2)
    with pytest.raises(sq.ShenqiError):
        sq.bigfile_block_info(path, "1/Velocity")


def test_header_attributes_and_checksum(tmp_path):
    import shenqi_b200 as sq
    path = str(tmp_path / "IC")
    n = 4096
    pos = np.random.default_rng(0).random((n, 3)) * 8.0
    _write(path, "1/Position", pos, 2)
    masstable = [0.0, 0.0431, 0.0, 0.0, 0.0, 0.0]
    _attr(path, "Header", "BoxSize", [64000.0], "f8")
    _attr(path, "Header", "MassTable", masstable, "f8")
    _attr(path, "Header", "TotNumPart", [0, n, 0, 0, 0, 0], "u8")
    _attr(path, "Header", "Time", [0.1], "f8")
    assert sq.bigfile_read_attr(path, "Header", "BoxSize")[0] == 64000.0
    assert np.array_equal(sq.bigfile_read_attr(path, "Header", "MassTable"), masstable)
    assert np.array_equal(sq.bigfile_read_attr(path, "Header", "TotNumPart"), [0, n, 0, 0, 0, 0])
    with pytest.raises(sq.ShenqiError):
        sq.bigfile_read_attr(path, "Header", "NoSuchAttribute")
    # a flipped byte in a physical file is caught by the checksum of the header (sysvsum, bigfile.c:1453-1460)
    fn = os.path.join(path, "1", "Position", "000001")
    raw = bytearray(open(fn, "rb").read())
    raw[100] = (raw[100] + 1) % 256
    open(fn, "wb").write(bytes(raw))
    with pytest.raises(sq.ShenqiError) as e:
        sq.bigfile_read_block(path, "1/Position")
    assert "checksum" in str(e.value)
    assert np.array_equal(sq.bigfile_read_block(path, "1/Position", 0, n // 2), pos[:n // 2])   # the first file is intact

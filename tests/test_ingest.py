"""Frame-list ingest: the parallel bzip2 block decoder returns exactly what the
reference's bz2.BZ2File + cPickle.load (Utils/utilities.py:84-96) returns."""
import bz2
import os
import pickle

import numpy as np
import pytest

from rti_b200 import ingest


def _records(n, r, seed):
    rng = np.random.default_rng(seed)
    return [(rng.integers(0, 256, (r, r), dtype=np.uint8), rng.normal(size=3)) for _ in range(n)]


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert type(x) is type(y) and x[0].dtype == y[0].dtype and x[1].dtype == y[1].dtype
        assert np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1])


def _write_like_reference(path, data):
    with bz2.BZ2File(path, "wb") as f:          # Utils/utilities.py:61-64
        pickle.dump(data, f)


@pytest.mark.parametrize("n,r", [(1, 4), (3, 64), (40, 300)])
def test_parallel_reader_equals_stdlib(tmp_path, n, r):
    path = str(tmp_path / "frames.pbz2")
    recs = _records(n, r, seed=n)
    _write_like_reference(path, recs)
    with bz2.BZ2File(path, "rb") as f:
        want = pickle.load(f)
    reader = ingest.ParallelBz2Reader(path)
    assert reader.parallel
    if n == 40:
        assert len(reader._chunks) >= 4         # 3.6 MB of random bytes: several 900 kB blocks
    reader.close()
    _same(ingest.read_pbz2(path), want)
    _same(ingest.read_pbz2(path, threads=1), recs)


def test_blocks_sit_at_unaligned_bit_offsets(tmp_path):
    path = str(tmp_path / "f.pbz2")
    _write_like_reference(path, _records(30, 300, seed=9))
    raw = open(path, "rb").read()
    offs = ingest._find_magic_bits(raw, ingest.BLOCK_MAGIC)
    assert offs[0] == 32 and any(o % 8 for o in offs)      # the case a byte-level search would miss
    eos = ingest._find_magic_bits(raw, ingest.EOS_MAGIC)
    assert len(eos) == 1 and eos[0] > offs[-1]


def test_reader_file_interface(tmp_path):
    payload = b"first line\n" + bytes(range(256)) * 9000 + b"\nlast"
    path = str(tmp_path / "p.bz2")
    with bz2.BZ2File(path, "wb") as f:
        f.write(payload)
    r = ingest.ParallelBz2Reader(path, threads=2)
    assert r.readline() == b"first line\n"
    assert r.read(10) == payload[11:21]
    buf = bytearray(1_500_000)                                # crosses a block boundary
    assert r.readinto(buf) == len(buf) and bytes(buf) == payload[21:21 + len(buf)]
    assert r.read() == payload[21 + len(buf):]
    assert r.read(5) == b""
    r.close()


def test_concatenated_streams_fall_back_to_stdlib(tmp_path):
    path = str(tmp_path / "two.pbz2")
    data = pickle.dumps(_records(2, 32, seed=3))
    with open(path, "wb") as f:
        f.write(bz2.compress(data[:1000]) + bz2.compress(data[1000:]))
    reader = ingest.ParallelBz2Reader(path)
    assert not reader.parallel
    reader.close()
    _same(ingest.read_pbz2(path), pickle.loads(data))


def test_damaged_and_missing_files(tmp_path):
    path = str(tmp_path / "bad.pbz2")
    _write_like_reference(path, _records(4, 200, seed=2))
    raw = bytearray(open(path, "rb").read())
    raw[len(raw) // 2] ^= 0x55
    open(path, "wb").write(bytes(raw))
    with pytest.raises((OSError, ValueError, EOFError, pickle.UnpicklingError)):
        ingest.read_pbz2(path)
    open(path, "wb").write(b"not a bzip2 file at all")
    with pytest.raises(OSError):
        ingest.read_pbz2(path)
    import rti_b200
    with pytest.raises(Exception, match="Storage file not found!"):
        rti_b200.read_from_file(str(tmp_path / "nope"))


def test_read_from_file_roundtrip_with_reference_writer(tmp_path):
    import rti_b200
    recs = _records(5, 100, seed=4)
    rti_b200.write_on_file(recs, str(tmp_path / "frames_results_x"), compressed=True)
    assert os.path.isfile(str(tmp_path / "frames_results_x.pbz2"))
    _same(rti_b200.read_from_file(str(tmp_path / "frames_results_x")), recs)


def test_pinned_stack_layout():
    recs = _records(3, 10, seed=1)
    frames, cams = ingest.pinned_stack(recs, 8, 8)
    assert tuple(frames.shape) == (3, 8, 8) and cams.shape == (3, 3)
    for i, (roi, cam) in enumerate(recs):
        assert np.array_equal(frames[i].numpy(), roi[:8, :8]) and np.array_equal(cams[i], cam)
    with pytest.raises(IndexError):
        ingest.pinned_stack(recs, 12, 12)


@pytest.mark.gpu
def test_gpu_load_frame_stack_feeds_compute(tmp_path, lib):
    """'.pbz2' -> parallel decode -> pinned slab -> HBM equals compute_intensities on the list."""
    import torch
    import rti_b200
    from rti_b200 import analysis
    import rti_oracle as orc
    r, n = 12, 20
    rng = np.random.default_rng(11)
    recs = [(rng.integers(0, 256, (r, r), dtype=np.uint8),
             np.array([rng.uniform(-5, 17), rng.uniform(-5, 17), rng.uniform(9, 27)])) for _ in range(n)]
    rti_b200.write_on_file(recs, str(tmp_path / "frames_results_t"), compressed=True)
    old = rti_b200.constants.ROI_DIAMETER
    rti_b200.constants.ROI_DIAMETER = r
    try:
        stack = ingest.load_frame_stack(str(tmp_path / "frames_results_t"))
        torch.cuda.synchronize()
        assert np.array_equal(stack.d_frames.cpu().numpy(), np.stack([x[0] for x in recs]))
        coef, bad = analysis._fit_ptm_from_stack(stack)
        lx, ly, inten = orc.gather_soa(np.stack([x[0] for x in recs]), np.stack([x[1] for x in recs]))
        ref = orc.ptm_fit_batch(lx, ly, inten)
        got = coef.cpu().numpy().reshape(6, r, r).transpose(1, 2, 0)
        assert int(bad.item()) == 0 and np.abs(got - ref).max() <= 1e-9 * np.abs(ref).max()
    finally:
        rti_b200.constants.ROI_DIAMETER = old


def test_native_scan_and_bit_copy_equal_the_python_forms():
    """csrc/host_bz2.cpp (rti_find_magic48, rti_copy_bits) against ingest.py's pure-Python forms: the magic at
    every bit alignment, at both ends of the buffer, overlapping a thread-chunk seam, in random data; bit
    ranges at every source / destination alignment."""
    import ctypes as C
    from rti_b200 import _capi, ingest
    lib = _capi.load()
    rng = np.random.default_rng(3)
    magic = ingest.BLOCK_MAGIC

    def plant(buf, bit):                       # write the 48-bit magic at an arbitrary bit offset
        bits = np.unpackbits(buf)
        bits[bit:bit + 48] = ingest._bits_of(magic, 48)
        return np.packbits(bits)

    for n, where in ((6, [0]), (7, [3]), (64, [0, 101, 64 * 8 - 48]), (4096, [17, 1000, 4096 * 8 - 49]),
                     (3 << 20, [5, (1 << 20) * 8 - 20, (1 << 20) * 8 + 43, (2 << 20) * 8 - 1, (3 << 20) * 8 - 48])):
        buf = rng.integers(0, 256, n, dtype=np.uint8)
        for bit in where:
            buf = plant(buf, bit)
        raw = buf.tobytes()
        want = ingest._find_magic_bits_py(raw, magic)
        assert set(where) <= set(want)
        for threads in (1, 3, 0):
            out = (C.c_int64 * 64)()
            cnt = lib.rti_find_magic48(raw, len(raw), magic, out, 64, threads)
            assert list(out[:cnt]) == want
        assert ingest._find_magic_bits(raw, magic) == want
        assert lib.rti_find_magic48(raw, len(raw), magic, None, 0, 1) == len(want)      # count only
    src = rng.integers(0, 256, 64, dtype=np.uint8)
    sbits = np.unpackbits(src)
    for s0 in range(0, 19):
        for d0 in range(0, 11):
            for nbits in (1, 5, 8, 13, 64, 200):
                dst = np.full(48, 0xFF, dtype=np.uint8)
                end = lib.rti_copy_bits(src.ctypes.data, s0, nbits, dst.ctypes.data, d0)
                assert end == d0 + nbits
                want_bits = np.unpackbits(np.full(48, 0xFF, dtype=np.uint8))
                want_bits[d0:d0 + nbits] = sbits[s0:s0 + nbits]
                last = (d0 + nbits + 7) // 8
                want_bits[d0 + nbits:last * 8] = 0                     # the tail of the last byte is zeroed
                assert np.array_equal(dst[:last], np.packbits(want_bits)[:last]), (s0, d0, nbits)
                assert np.all(dst[last:] == 0xFF)                      # nothing beyond it is touched

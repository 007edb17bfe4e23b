"""Frame-list ingest: '<name>.pbz2' -> frame records -> pinned stack -> HBM.

The reference reads its stage-1 checkpoint with ``bz2.BZ2File`` + ``cPickle.load``
(Utils/utilities.py:84-96), a single-threaded decode.  A ``.pbz2`` written by
``ut.write_on_file`` (utilities.py:61-64) is ONE bzip2 stream of independent blocks
(<= 900 kB of pickle each); the blocks start at arbitrary *bit* offsets with the 48-bit
magic 0x314159265359 and carry their own CRC.  ``ParallelBz2Reader`` finds the block
boundaries, re-wraps every block as a stand-alone one-block stream and decodes the blocks
on a thread pool (``bz2.decompress`` releases the GIL), handing the bytes to the unpickler in
order while later blocks are still decoding.  The unpickled object is identical to the
reference's; anything unexpected (several concatenated streams, a damaged file) falls back
to ``bz2.BZ2File`` so errors are the standard library's.

``load_frame_stack`` then copies the records into one pinned (N, R, R) uint8 slab and
uploads it on a side stream (a device-side ``FrameStack`` for analysis.py).
"""
from __future__ import annotations

import bz2
import io
import os
import pickle
from concurrent.futures import ThreadPoolExecutor

import numpy as np

BLOCK_MAGIC = 0x314159265359          # BCD pi, starts every compressed block
EOS_MAGIC = 0x177245385090            # BCD sqrt(pi), starts the stream footer


def _native():
    """librti_b200.so's host-only ingest helpers (csrc/host_bz2.cpp), or None when the library is not
    built: the pure-Python forms below give the same answers, only slower."""
    global _NATIVE
    if _NATIVE is False:
        try:
            from . import _capi
            _NATIVE = _capi.load()
        except Exception:
            _NATIVE = None
    return _NATIVE


_NATIVE = False


def _find_magic_bits(buf: bytes, magic: int):
    """Bit offsets of every occurrence of the 48-bit ``magic`` at any bit alignment: one threaded pass in
    the library (rti_find_magic48) instead of sixteen bytes.find passes under the GIL (3.4 s of a 337 MB
    file's read before; the Python form stays as the fallback and as the test's reference)."""
    lib = _native()
    if lib is not None and len(buf) >= 6:
        import ctypes as C
        src = (C.c_char * len(buf)).from_buffer_copy(buf) if not isinstance(buf, bytes) else buf
        cap = 1 << 16
        while True:
            out = (C.c_int64 * cap)()
            n = lib.rti_find_magic48(src, len(buf), magic, out, cap, 0)
            if n <= cap:
                return list(out[:n])
            cap = n
    return _find_magic_bits_py(buf, magic)


def _find_magic_bits_py(buf: bytes, magic: int):
    hits = []
    size = len(buf)
    for s in range(8):
        # the pattern sits s bits into a 7-byte window; its fully determined bytes are
        # searched with bytes.find, the two partial edge bytes are verified afterwards
        val = (magic << (8 - s)) if s else (magic << 8)
        win = val.to_bytes(7, "big")
        if s == 0:
            core, lead = win[:6], 0
        else:
            core, lead = win[1:6], 1
        hi_mask = (0xFF >> s) if s else 0            # bits of window byte 0 owned by the pattern
        lo_mask = (0xFF << (8 - s)) & 0xFF if s else 0
        pos = buf.find(core)
        while pos >= 0:
            start = pos - lead
            ok = start >= 0
            if ok and s:
                ok = (buf[start] & hi_mask) == (win[0] & hi_mask)
                ok = ok and start + 6 < size and (buf[start + 6] & lo_mask) == (win[6] & lo_mask)
            if ok:
                hits.append(start * 8 + s)
            pos = buf.find(core, pos + 1)
    hits.sort()
    return hits


def _bits_of(value: int, nbits: int) -> np.ndarray:
    return np.array([(value >> (nbits - 1 - i)) & 1 for i in range(nbits)], dtype=np.uint8)


_EOS_BITS = _bits_of(EOS_MAGIC, 48)


_EOS_BYTES = EOS_MAGIC.to_bytes(6, "big")


def _decode_block(buf: memoryview, header: bytes, b0: int, b1: int) -> bytes:
    """Decode the block occupying bits [b0, b1) of ``buf`` as a one-block bzip2 stream:
    header + block + end-of-stream magic + combined CRC (= the block's own CRC)."""
    lib = _native()
    if lib is not None:
        import ctypes as C
        first, last = b0 >> 3, (b1 + 7) >> 3
        src = (C.c_ubyte * (last - first)).from_buffer_copy(buf[first:last])
        nbits = b1 - b0
        out = (C.c_ubyte * ((32 + nbits + 80 + 7) >> 3))()
        out[0:4] = header
        rel = b0 - first * 8
        end = lib.rti_copy_bits(src, rel, nbits, out, 32)                          # the block
        end = lib.rti_copy_bits(_EOS_BYTES, 0, 48, out, end)                       # end-of-stream magic
        lib.rti_copy_bits(src, rel + 48, 32, out, end)                             # stream CRC = the block's CRC
        return bz2.decompress(bytes(out))
    return _decode_block_py(buf, header, b0, b1)


def _decode_block_py(buf: memoryview, header: bytes, b0: int, b1: int) -> bytes:
    first, last = b0 >> 3, (b1 + 7) >> 3
    bits = np.unpackbits(np.frombuffer(buf[first:last], dtype=np.uint8))
    bits = bits[b0 - first * 8: b1 - first * 8]
    head = np.unpackbits(np.frombuffer(header, dtype=np.uint8))
    stream = np.concatenate([head, bits, _EOS_BITS, bits[48:80]])
    return bz2.decompress(np.packbits(stream).tobytes())


class ParallelBz2Reader(io.RawIOBase):
    """File-like object over the decompressed bytes of a single-stream bzip2 file whose
    blocks are decoded concurrently.  ``read``/``readinto``/``readline`` are what
    ``pickle.Unpickler`` needs."""

    def __init__(self, path, threads=None):
        super().__init__()
        with open(path, "rb") as f:
            self._raw = f.read()
        raw = self._raw
        if len(raw) < 14 or raw[:3] != b"BZh" or not (0x31 <= raw[3] <= 0x39):
            raise OSError("Invalid data stream")                 # same error BZ2File gives
        blocks = _find_magic_bits(raw, BLOCK_MAGIC)
        eos = _find_magic_bits(raw, EOS_MAGIC)
        # single stream: one footer whose magic + 32-bit CRC (padded to a byte) ends the file
        self.parallel = (len(eos) == 1 and ((eos[0] + 80 + 7) >> 3) == len(raw)
                         and (not blocks or (blocks[0] == 32 and blocks[-1] < eos[0])))
        self._chunks = []
        self._cur = b""
        self._pos = 0
        self._next = 0
        if self.parallel and blocks:
            threads = threads or min(32, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity")
                                     else (os.cpu_count() or 1))
            bounds = blocks + [eos[0]]
            view = memoryview(raw)
            self._pool = ThreadPoolExecutor(max_workers=max(1, threads))
            self._chunks = [self._pool.submit(_decode_block, view, raw[:4], bounds[i], bounds[i + 1])
                            for i in range(len(blocks))]
        elif self.parallel:                                      # empty stream: header + footer only
            self._pool = None
        else:
            self._pool = None
            self._serial = bz2.BZ2File(io.BytesIO(raw), "rb")

    # -- io.RawIOBase
    def readable(self):
        return True

    def _advance(self):
        if self._next >= len(self._chunks):
            return False
        fut = self._chunks[self._next]
        self._chunks[self._next] = None                          # let the decoded bytes go once consumed
        self._next += 1
        self._cur = fut.result()
        self._pos = 0
        return True

    def readinto(self, b):
        if not self.parallel:
            return self._serial.readinto(b)
        mv = memoryview(b).cast("B")
        done = 0
        while done < len(mv):
            if self._pos >= len(self._cur) and not self._advance():
                break
            k = min(len(mv) - done, len(self._cur) - self._pos)
            mv[done:done + k] = self._cur[self._pos:self._pos + k]
            self._pos += k
            done += k
        return done

    def read(self, size=-1):
        if not self.parallel:
            return self._serial.read(size)
        if size is None or size < 0:
            parts = [self._cur[self._pos:]]
            self._pos = len(self._cur)
            while self._advance():
                parts.append(self._cur)
                self._pos = len(self._cur)
            return b"".join(parts)
        if self._pos + size <= len(self._cur):                   # common case: inside one block
            out = self._cur[self._pos:self._pos + size]
            self._pos += size
            return out
        out = bytearray(size)
        k = self.readinto(out)
        return bytes(out[:k])

    def readline(self, size=-1):
        if not self.parallel:
            return self._serial.readline(size)
        parts = []
        while True:
            if self._pos >= len(self._cur) and not self._advance():
                break
            nl = self._cur.find(b"\n", self._pos)
            end = len(self._cur) if nl < 0 else nl + 1
            parts.append(self._cur[self._pos:end])
            self._pos = end
            if nl >= 0:
                break
        return b"".join(parts)

    def close(self):
        if getattr(self, "_pool", None) is not None:
            self._pool.shutdown(wait=False, cancel_futures=True)
            self._pool = None
        super().close()


def read_pbz2(path, threads=None):
    """Unpickle a bzip2-compressed pickle (Utils/utilities.py:94-96) with parallel block
    decoding.  ``path`` includes the '.pbz2' extension."""
    reader = ParallelBz2Reader(path, threads=threads)
    try:
        return pickle.load(reader)
    except (OSError, ValueError, EOFError, pickle.UnpicklingError):
        if not reader.parallel:
            raise
        # a chance occurrence of a magic inside block data: decode the plain way
        with bz2.BZ2File(path, "rb") as f:
            return pickle.load(f)
    finally:
        reader.close()


def roi_as_uint8(roi, index):
    """The ROI intensities of one frame record as uint8 without silent wrap-around.  The reference
    stores intensities[y][x] into an int32 array (analysis.py:219, 232) and so keeps any integer; the
    CUDA path carries uint8 frames (what FeatureMatcher.py:183-184 emits), so anything else must be
    integral and inside [0, 255] -- otherwise it is an error, not a modulo-256 value."""
    roi = np.asarray(roi)
    if roi.dtype == np.uint8:
        return roi
    if roi.dtype == np.bool_ or not np.issubdtype(roi.dtype, np.number) or np.iscomplexobj(roi):
        raise ValueError("frame %d: ROI intensities have dtype %s; uint8 expected" % (index, roi.dtype))
    if roi.size:
        lo, hi = roi.min(), roi.max()
        if not (lo >= 0 and hi <= 255):
            raise ValueError("frame %d: ROI intensities span [%s, %s], outside the uint8 range of the "
                             "frame records (FeatureMatcher.py:183-184)" % (index, lo, hi))
        if not np.issubdtype(roi.dtype, np.integer) and not np.all(roi == np.rint(roi)):
            raise ValueError("frame %d: non-integral ROI intensities (dtype %s)" % (index, roi.dtype))
    return roi.astype(np.uint8)


def pinned_stack(records, rows, cols):
    """list[(uint8 (H,W), float64 (3,))] -> pinned uint8 (N, rows, cols) tensor + float64 (N, 3)
    ndarray (the layout of analysis._stack_frames, written once, straight into pinned memory)."""
    import torch
    n = len(records)
    frames = torch.empty((n, rows, cols), dtype=torch.uint8, pin_memory=torch.cuda.is_available())
    view = frames.numpy()
    cams = np.empty((n, 3), dtype=np.float64)
    for i, rec in enumerate(records):
        roi = np.asarray(rec[0])
        if roi.ndim != 2 or roi.shape[0] < rows or roi.shape[1] < cols:
            raise IndexError("frame %d ROI has shape %s, need at least (%d, %d)" % (i, roi.shape, rows, cols))
        view[i] = roi_as_uint8(roi[:rows, :cols], i)
        cams[i] = np.asarray(rec[1], dtype=np.float64).reshape(3)
    return frames, cams


def load_frame_stack(filename, rows=None, cols=None, compressed=True, threads=None):
    """'<filename>.pbz2' (or '.pickle') -> analysis.FrameStack on the current device.  The upload
    is issued asynchronously from pinned memory on the current stream."""
    from . import analysis
    torch = analysis._torch()
    path = filename + (".pbz2" if compressed else ".pickle")
    if not os.path.isfile(path):
        raise Exception("Storage file not found!")               # Utils/utilities.py:89-90
    if compressed:
        records = read_pbz2(path, threads=threads)
    else:
        with open(path, "rb") as f:
            records = pickle.load(f)
    if records is None or len(records) <= 0:
        raise Exception("Error computing intensities: results are empty")
    r = analysis._cst.ROI_DIAMETER
    rows = r if rows is None else rows
    cols = r if cols is None else cols
    frames, cams = pinned_stack(records, rows, cols)
    d_frames = frames.cuda(non_blocking=True)
    d_cams = torch.from_numpy(cams).cuda()
    stack = analysis.FrameStack.from_device(d_frames, d_cams)
    stack._pinned = frames                                       # keep the source alive until the copy ran
    return stack

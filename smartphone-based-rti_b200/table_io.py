"""Streaming writer for the interpolation-result pickle.

The reference stores ``interpolated_images[ly][lx] -> int32 (R, R)`` with
``ut.write_on_file(results_images, path, compressed=False)`` = ``pickle.dump`` of a list of
lists of 10 000 small ndarrays (Utils/utilities.py:66-68, analysis.py:475; 6.4 GB at R=400)
and ``interactive_relighting.compute`` reads it back with ``pickle.load``
(interactive_relighting.py:95 -> utilities.py:98-99).  ``pickle.dump`` makes a ``tobytes()``
copy of every image and frames it; once the GPU produces the table in milliseconds that
host-side copy loop is the slowest step of ``compute()``.

``dump_table`` writes the SAME object graph -- ``pickle.load`` returns a list of lists of
writable int32 ndarrays, nothing on the reader's side changes -- straight from one
contiguous ``(G, G, R, R)`` host array: the pickle opcodes that surround an image's raw
bytes are taken once from ``pickle.dumps`` of a template array (so they track the installed
numpy's own reduce format), and the image bytes go to the file as memoryviews of the slab,
without intermediate copies.
"""
from __future__ import annotations

import io
import pickle
import pickletools
import struct

import numpy as np

_PROTO = 4          # what pickle.dump uses by default on Python 3.8 - 3.13 (the reference passes no protocol)


def _array_template(shape, dtype):
    """(prefix, suffix): opcode bytes before / after the raw data of one pickled ndarray of
    this shape and dtype, free of memo references so that they can be repeated verbatim."""
    probe = np.zeros(shape, dtype=dtype)
    stream = pickle.dumps(probe, protocol=_PROTO)
    nbytes = probe.nbytes
    ops = list(pickletools.genops(stream))
    pieces, data_at = [], None
    memo, last = {}, None                             # memo index -> (opcode name, raw bytes) of what was memoized
    leaf = ("SHORT_BINUNICODE", "BINUNICODE", "BINUNICODE8", "SHORT_BINBYTES", "BINBYTES", "BININT", "BININT1",
            "BININT2", "NONE", "NEWTRUE", "NEWFALSE")
    for k, (op, arg, pos) in enumerate(ops):
        end = ops[k + 1][2] if k + 1 < len(ops) else len(stream)
        if op.name in ("PROTO", "FRAME", "STOP"):
            continue                                  # re-framed / re-terminated by the caller
        if op.name == "MEMOIZE":                      # dropped: memo references are expanded below
            memo[len(memo)] = last
            continue
        if op.name in ("BINGET", "LONG_BINGET"):
            name, raw = memo.get(arg, (None, None))
            if name not in leaf:
                raise RuntimeError("ndarray pickle refers back to a non-constant object; template not applicable")
            pieces.append(raw)                        # a constant: emit it again instead of the reference
            last = (name, raw)
            continue
        if op.name in ("GET", "BINPUT", "LONG_BINPUT", "PUT"):
            raise RuntimeError("unexpected memo opcode %s in a protocol-%d ndarray pickle" % (op.name, _PROTO))
        last = (op.name, stream[pos:end])
        if op.name in ("BINBYTES", "BINBYTES8", "SHORT_BINBYTES") and isinstance(arg, bytes) and len(arg) == nbytes \
                and data_at is None:
            data_at = len(pieces)
            pieces.append(None)
            continue
        pieces.append(stream[pos:end])
    if data_at is None:
        raise RuntimeError("could not locate the raw data in the ndarray pickle")
    prefix = b"".join(pieces[:data_at]) + b"\x8e" + struct.pack("<Q", nbytes)      # BINBYTES8 + length
    suffix = b"".join(pieces[data_at + 1:])
    return prefix, suffix


def is_table_slab(images):
    """A C-contiguous (G_v, G_u, rows, cols) ndarray over the images' own memory if `images` is a nested
    list of equally shaped views laid out back to back in one buffer (what prepare_images_data
    returns: every images[v][u] is a view of one host slab), else None."""
    try:
        first = images[0][0]
        base = first.base
        gv, gu = len(images), len(images[0])
        if base is None or not isinstance(first, np.ndarray) or not first.flags.c_contiguous or first.ndim != 2:
            return None
        if any(len(row) != gu for row in images):
            return None
        nb = first.nbytes
        addr0 = first.__array_interface__["data"][0]
        for v in range(gv):
            for u in range(gu):
                a = images[v][u]
                if not isinstance(a, np.ndarray) or a.base is not base or a.shape != first.shape \
                        or a.dtype != first.dtype or not a.flags.c_contiguous \
                        or a.__array_interface__["data"][0] != addr0 + (v * gu + u) * nb:
                    return None
        return np.lib.stride_tricks.as_strided(first, shape=(gv, gu) + first.shape,
                                               strides=(gu * nb, nb) + first.strides, writeable=False)
    except Exception:
        return None


def template_ok(image_shape, dtype) -> bool:
    """True when the installed numpy pickles an ndarray of this shape the way the streaming writer
    assumes; callers check BEFORE opening (truncating) the output file and fall back to
    ``pickle.dump`` otherwise."""
    try:
        _array_template(tuple(image_shape), np.dtype(dtype))
        return True
    except RuntimeError:
        return False


class TableStreamWriter:
    """Push-style writer of the protocol-4 pickle of ``[[img(v, u) for u in range(gu)] for v in
    range(gv)]``: ``write_images(block)`` takes the next images in row-major order as a C-contiguous
    (count, ...) array (or any sequence of buffers) and writes them at once, so the source may be a
    pinned staging buffer that the next device-to-host copy refills; ``close()`` terminates the pickle."""

    def __init__(self, gv, gu, image_shape, dtype, fileobj):
        self.gv, self.gu = int(gv), int(gu)
        self.prefix, self.suffix = _array_template(tuple(image_shape), np.dtype(dtype))
        self.w = fileobj.write
        self.count = 0
        self.w(pickle.PROTO + bytes([_PROTO]))
        self.w(pickle.EMPTY_LIST)
        if self.gv:
            self.w(pickle.MARK)
        if self.gu == 0:                       # rows without images
            for _ in range(self.gv):
                self.w(pickle.EMPTY_LIST)

    def write_images(self, block):
        w, gu = self.w, self.gu
        for img in block:
            if self.count >= self.gv * gu:
                raise ValueError("more images than gv * gu")
            if self.count % gu == 0:
                w(pickle.EMPTY_LIST)
                w(pickle.MARK)
            w(self.prefix)
            w(memoryview(np.ascontiguousarray(img)).cast("B"))
            w(self.suffix)
            self.count += 1
            if self.count % gu == 0:
                w(pickle.APPENDS)

    def close(self):
        if self.count != self.gv * self.gu:
            raise ValueError("table incomplete: %d of %d images written" % (self.count, self.gv * self.gu))
        if self.gv:
            self.w(pickle.APPENDS)
        self.w(pickle.STOP)


class TableTileWriter:
    """Random-access writer of the same pickle: the file is laid out up front (every opcode between
    the images written, the image payloads left as holes), then ``write_rows(i0, block, row0)`` drops
    rows [row0, row0 + block.shape[1]) of images i0 .. i0 + len(block) at their fixed offsets with
    ``os.pwrite``.  This is what lets ``compute()`` produce the table in ROW TILES of the ROI through a
    bounded device buffer (a 4096^2 ROI has a 671 GB table) instead of image by image."""

    def __init__(self, gv, gu, image_shape, dtype, path):
        import os
        self.gv, self.gu = int(gv), int(gu)
        self.rows, self.cols = int(image_shape[0]), int(image_shape[1])
        self.itemsize = np.dtype(dtype).itemsize
        prefix, suffix = _array_template(tuple(image_shape), np.dtype(dtype))
        nbytes = self.rows * self.cols * self.itemsize
        head = pickle.PROTO + bytes([_PROTO]) + pickle.EMPTY_LIST + (pickle.MARK if self.gv else b"")
        row_open, row_close = pickle.EMPTY_LIST + pickle.MARK, pickle.APPENDS
        closer = (pickle.APPENDS if self.gv else b"") + pickle.STOP
        per_image = len(prefix) + nbytes + len(suffix)
        row_bytes = len(row_open) + self.gu * per_image + len(row_close)
        if self.gu == 0:
            row_bytes = len(pickle.EMPTY_LIST)
        self.size = len(head) + self.gv * row_bytes + len(closer)
        self.fd = os.open(path, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
        os.ftruncate(self.fd, self.size)
        os.pwrite(self.fd, head, 0)
        self.offsets = np.empty(self.gv * self.gu, dtype=np.int64)
        glue = suffix + prefix                               # between two images of one row
        for v in range(self.gv):
            base = len(head) + v * row_bytes
            if self.gu == 0:
                os.pwrite(self.fd, pickle.EMPTY_LIST, base)
                continue
            os.pwrite(self.fd, row_open + prefix, base)
            first = base + len(row_open) + len(prefix)
            self.offsets[v * self.gu:(v + 1) * self.gu] = first + per_image * np.arange(self.gu)
            for u in range(1, self.gu):
                os.pwrite(self.fd, glue, first + (u - 1) * per_image + nbytes)
            os.pwrite(self.fd, suffix + row_close, first + (self.gu - 1) * per_image + nbytes)
        os.pwrite(self.fd, closer, self.size - len(closer))

    def write_rows(self, i0, block, row0):
        """block: (count, tile_rows * cols) or (count, tile_rows, cols) C-contiguous rows of images i0.."""
        import os
        block = np.asarray(block)
        skip = row0 * self.cols * self.itemsize
        for k in range(block.shape[0]):
            os.pwrite(self.fd, memoryview(np.ascontiguousarray(block[k])).cast("B"), int(self.offsets[i0 + k]) + skip)

    def close(self):
        import os
        if self.fd is not None:
            os.close(self.fd)
            self.fd = None


def dump_table(table, fileobj):
    """Write ``[[table[v, u] for u] for v]`` as a protocol-4 pickle to the binary file object.

    ``table``: C-contiguous ndarray (G_v, G_u, rows, cols).  Equivalent to
    ``pickle.dump([[table[v, u] for u in range(G_u)] for v in range(G_v)], fileobj)``."""
    table = np.ascontiguousarray(table)
    gv, gu = table.shape[:2]
    writer = TableStreamWriter(gv, gu, table.shape[2:], table.dtype, fileobj)
    writer.write_images(table.reshape((gv * gu,) + table.shape[2:]))
    writer.close()


def dumps_table(table) -> bytes:
    buf = io.BytesIO()
    dump_table(table, buf)
    return buf.getvalue()


def load_table(path):
    """Memory-map a result pickle written by TableStreamWriter / dump_table and return the same object
    graph ``pickle.load`` gives -- list[G_v] of list[G_u] of WRITABLE int32 (rows, cols) ndarrays --
    without reading the file: every image is a view into one copy-on-write mapping (``mmap.ACCESS_COPY``),
    so pages are read when an image is first touched and the relighter's in-place clamp
    (interactive_relighting.py:35-36) stays private to the process.  Returns None when the file does
    not have exactly this writer's byte layout (the caller then uses ``pickle.load``)."""
    import mmap
    import os
    size = os.path.getsize(path)
    with open(path, "rb") as f:
        head = f.read(4096)
        opener = pickle.PROTO + bytes([_PROTO]) + pickle.EMPTY_LIST + pickle.MARK + pickle.EMPTY_LIST + pickle.MARK
        if not head.startswith(opener):
            return None
        # the first image's prefix ends with BINBYTES8 + length; recover shape and dtype by unpickling a
        # one-image file made of the first image only
        k = head.find(b"\x8e", len(opener))
        while k >= 0:
            nbytes = struct.unpack("<Q", head[k + 1:k + 9])[0] if k + 9 <= len(head) else 0
            if 0 < nbytes <= size:
                break
            k = head.find(b"\x8e", k + 1)
        if k < 0:
            return None
        prefix = head[len(opener):k + 9]
        f.seek(len(opener) + len(prefix) + nbytes)
        tail = f.read(512)
        try:
            probe_dtype, probe_shape, suffix = _probe_image(prefix, tail, nbytes)
        except Exception:
            return None
        if probe_dtype != np.dtype(np.int32) or len(probe_shape) != 2:
            return None
        per_image = len(prefix) + nbytes + len(suffix)
        row_open = pickle.EMPTY_LIST + pickle.MARK
        row_close = pickle.APPENDS
        closer = pickle.APPENDS + pickle.STOP
        # size = 3 + 1 + gv * (2 + gu * per_image + 1) + 2  with gv, gu unknown: find gu from the first row
        mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_COPY)
    body0 = len(pickle.PROTO) + 1 + 2                       # PROTO n, EMPTY_LIST, MARK
    gu = 0
    pos = body0 + len(row_open)
    while mm[pos:pos + len(prefix)] == prefix:
        pos += per_image
        gu += 1
        if pos >= size:
            return None
    if gu == 0 or mm[pos:pos + 1] != row_close:
        return None
    row_bytes = len(row_open) + gu * per_image + len(row_close)
    gv, rem = divmod(size - body0 - len(closer), row_bytes)
    if rem != 0 or gv <= 0 or mm[size - len(closer):size] != closer:
        return None
    whole = np.frombuffer(mm, dtype=np.uint8)               # writable: private copy-on-write pages
    out = []
    for v in range(gv):
        base = body0 + v * row_bytes
        if mm[base:base + len(row_open)] != row_open or mm[base + row_bytes - 1:base + row_bytes] != row_close:
            return None
        row = []
        for u in range(gu):
            at = base + len(row_open) + u * per_image
            if mm[at:at + len(prefix)] != prefix or \
                    mm[at + len(prefix) + nbytes:at + per_image] != suffix:
                return None
            img = whole[at + len(prefix):at + len(prefix) + nbytes].view(np.int32).reshape(probe_shape)
            row.append(img)
        out.append(row)
    return out


def _probe_image(prefix, tail, nbytes):
    """dtype, shape and suffix bytes of the images of a stream: the suffix is this writer's template
    suffix for the (shape, dtype) that the prefix announces; found by trying the template of every
    2-D int32 shape with nbytes bytes whose prefix matches."""
    n = nbytes // 4
    for rows in range(1, n + 1):
        if n % rows:
            continue
        cols = n // rows
        p, s = _array_template((rows, cols), np.int32)
        if p == prefix and tail.startswith(s):
            return np.dtype(np.int32), (rows, cols), s
    raise ValueError("not a TableStreamWriter image")

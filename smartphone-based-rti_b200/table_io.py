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
    """The (G, G, R, R) base array if `images` is the nested list of views that
    prepare_images_data returns (every images[v][u] is base[v, u]), else None."""
    try:
        first = images[0][0]
        base = first.base
        if base is None or not isinstance(base, np.ndarray) or base.ndim != 4 or not base.flags.c_contiguous:
            return None
        if len(images) != base.shape[0] or any(len(row) != base.shape[1] for row in images):
            return None
        step = base.strides[1]
        addr0 = base.__array_interface__["data"][0]
        for v in (0, len(images) - 1):
            for u in (0, len(images[v]) - 1):
                a = images[v][u]
                if a.base is not base or a.shape != base.shape[2:] or a.dtype != base.dtype \
                        or a.__array_interface__["data"][0] != addr0 + (v * base.shape[1] + u) * step:
                    return None
        return base
    except Exception:
        return None


def dump_table(table, fileobj):
    """Write ``[[table[v, u] for u] for v]`` as a protocol-4 pickle to the binary file object.

    ``table``: C-contiguous ndarray (G_v, G_u, rows, cols).  Equivalent to
    ``pickle.dump([[table[v, u] for u in range(G_u)] for v in range(G_v)], fileobj)``."""
    table = np.ascontiguousarray(table)
    gv, gu = table.shape[:2]
    prefix, suffix = _array_template(table.shape[2:], table.dtype)
    flat = table.reshape(gv, gu, -1)
    w = fileobj.write
    w(pickle.PROTO + bytes([_PROTO]))
    w(pickle.EMPTY_LIST)
    if gv:
        w(pickle.MARK)
    for v in range(gv):
        w(pickle.EMPTY_LIST)
        if gu:
            w(pickle.MARK)
            for u in range(gu):
                w(prefix)
                w(memoryview(flat[v, u]).cast("B"))
                w(suffix)
            w(pickle.APPENDS)
    if gv:
        w(pickle.APPENDS)
    w(pickle.STOP)


def dumps_table(table) -> bytes:
    buf = io.BytesIO()
    dump_table(table, buf)
    return buf.getvalue()

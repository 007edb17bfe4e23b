"""Host-side logic that needs no GPU: the call surface's argument checks and messages,
the grid, the relighting index arithmetic, file formats, row partitioning, synthetic data."""
import os
import pickle

import numpy as np
import pytest

import rti_b200
from rti_b200 import analysis, relighting, sharding, synthetic
import rti_oracle as orc


def test_light_axis_matches_oracle_and_follows_constants(monkeypatch):
    assert np.array_equal(analysis.light_axis(), orc.light_axis())
    monkeypatch.setattr(rti_b200.constants, "INTERPOLATION_PARAM", 0.05)
    assert analysis.light_axis().size == 40


def test_reference_error_messages():
    # same bare Exceptions and texts as analysis.py:205-206, 332-333, 384-385
    for bad in (None, []):
        with pytest.raises(Exception, match="Error computing intensities: results are empty"):
            rti_b200.compute_intensities(bad)
        with pytest.raises(Exception, match="Error preparing images: results are empty"):
            rti_b200.prepare_images_data(bad)
    for bad in (None, (1, 2)):
        with pytest.raises(Exception, match="Error computing interpolation: results are empty or invalid"):
            rti_b200.interpolate_intensities(bad)
    with pytest.raises(Exception, match="Storage file not found!"):
        rti_b200.read_from_file("/nonexistent/frames_results_x")


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    frames, cams = synthetic.random_stack(4, 4, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        rti_b200.compute_intensities(synthetic.as_records(frames, cams), first_only=True)


def test_file_round_trip_uses_reference_layout(tmp_path):
    frames, cams = synthetic.random_stack(5, 5, 3)
    recs = synthetic.as_records(frames, cams)
    base = str(tmp_path / "frames_results_synth")
    rti_b200.write_on_file(recs, base)                       # compressed -> .pbz2 (utilities.py:61-64)
    assert os.path.isfile(base + ".pbz2")
    import bz2
    with bz2.BZ2File(base + ".pbz2", "rb") as f:             # what the reference's reader does
        back = pickle.load(f)
    assert len(back) == 3 and back[0][0].dtype == np.uint8 and back[0][1].shape == (3,)
    back2 = rti_b200.read_from_file(base)
    assert np.array_equal(back2[2][0], frames[2])
    table = [[np.zeros((2, 2), np.int32) for _ in range(3)] for _ in range(3)]
    rti_b200.write_on_file(table, base + "_t", compressed=False)
    with open(base + "_t.pickle", "rb") as f:
        t = pickle.load(f)
    t[1][2][t[1][2] <= 0] = 0                                # consumer mutates in place (interactive_relighting.py:35-36)
    assert t[1][2].flags.writeable


def test_cursor_and_index_arithmetic(relight_golden):
    g = relight_golden
    for x, l, idx in zip(g["cursor"], g["light"], g["index"]):
        lx, ly = relighting.cursor_to_light(int(x), int(x), (400, 400))
        assert lx == l == ly
        assert relighting.light_to_index(lx) == idx


def test_row_partition():
    assert sharding.row_partition(4096, 8) == [(i * 512, 512) for i in range(8)]
    parts = sharding.row_partition(10, 4)
    assert parts == [(0, 2), (2, 2), (4, 3), (7, 3)]          # remainder on the last ranks
    assert sum(n for _, n in parts) == 10
    assert sharding.row_partition(2, 4) == [(0, 0), (0, 0), (0, 1), (1, 1)]
    assert sharding.row_partition(0, 2) == [(0, 0), (0, 0)]
    with pytest.raises(ValueError):
        sharding.row_partition(4, 0)


def test_synthetic_is_seeded_and_slab_consistent():
    f1, c1 = synthetic.lambertian_stack(8, 8, 5, seed=2)
    f2, c2 = synthetic.lambertian_stack(8, 8, 5, seed=2)
    assert np.array_equal(f1, f2) and np.array_equal(c1, c2)
    assert f1.dtype == np.uint8 and c1.dtype == np.float64 and c1.shape == (5, 3)
    # a slab sees the same cameras as the full ROI
    _, c3 = synthetic.lambertian_stack(4, 8, 5, seed=2, row0=4, rows_total=8)
    assert np.array_equal(c1, c3)
    recs = synthetic.as_records(f1, c1)
    assert len(recs) == 5 and recs[0][0].shape == (8, 8)


def test_stack_frames_validates_roi_size():
    recs = [(np.zeros((3, 3), np.uint8), np.zeros(3))]
    with pytest.raises(IndexError):
        analysis._stack_frames(recs, 4, 4)


def test_install_rebinds_reference_module():
    import types
    fake = types.ModuleType("analysis")
    fake.cst = types.SimpleNamespace(ROI_DIAMETER=7, INTERPOLATION_PARAM=0.5)
    old = analysis._cst
    try:
        rti_b200.install(fake)
        assert fake.compute_intensities is rti_b200.compute_intensities
        assert fake.prepare_images_data is rti_b200.prepare_images_data
        assert analysis.light_axis().size == 4               # now reads the reference's constants
    finally:
        analysis._cst = old


# ------------------------------------------------------------------ streaming table writer (8(f1))
def test_dump_table_round_trips_like_pickle_dump(tmp_path):
    import pickle
    from rti_b200 import table_io
    rng = np.random.default_rng(0)
    table = rng.integers(-500, 500, (5, 7, 6, 6)).astype(np.int32)
    images = [[table[v, u] for u in range(7)] for v in range(5)]
    slab = table_io.is_table_slab(images)
    assert slab.shape == table.shape and np.shares_memory(slab, table) and np.array_equal(slab, table)
    # the same images over a flat 2-D host buffer (how prepare_images_data assembles them) are recognised too
    flat = table.reshape(35, 36).copy()
    views = flat.reshape(5, 7, 6, 6)
    slab2 = table_io.is_table_slab([[views[v, u] for u in range(7)] for v in range(5)])
    assert slab2 is not None and np.shares_memory(slab2, flat) and np.array_equal(slab2, table)
    out = pickle.loads(table_io.dumps_table(table))
    ref = pickle.loads(pickle.dumps(images))
    assert type(out) is list and len(out) == 5 and all(type(r) is list and len(r) == 7 for r in out)
    for v in range(5):
        for u in range(7):
            a = out[v][u]
            assert a.dtype == np.int32 and a.shape == (6, 6) and a.flags.writeable and a.flags.c_contiguous
            assert np.array_equal(a, ref[v][u])
    # interactive_relighting.py:35-36 clamps the looked-up image in place
    out[2][3][out[2][3] > 255] = 255
    out[2][3][out[2][3] <= 0] = 0
    assert out[2][3].min() >= 0 and out[2][3].max() <= 255


def test_is_table_slab_rejects_anything_else():
    from rti_b200 import table_io
    table = np.zeros((3, 4, 2, 2), np.int32)
    assert table_io.is_table_slab([[table[v, u].copy() for u in range(4)] for v in range(3)]) is None
    assert table_io.is_table_slab([[table[v, u] for u in range(3)] for v in range(3)]) is None       # ragged vs base
    assert table_io.is_table_slab([[table[u % 3, v] for u in range(4)] for v in range(3)]) is None   # permuted
    assert table_io.is_table_slab([]) is None and table_io.is_table_slab(None) is None
    assert table_io.is_table_slab([(np.zeros((2, 2), np.uint8), np.zeros(3))]) is None               # frame records


def test_write_on_file_uses_streaming_writer_and_reference_reader_loads_it(tmp_path):
    import pickle
    import rti_b200
    table = np.arange(4 * 4 * 3 * 3, dtype=np.int32).reshape(4, 4, 3, 3)
    images = [[table[v, u] for u in range(4)] for v in range(4)]
    base = str(tmp_path / "interpolation_results_x")
    rti_b200.write_on_file(images, base, compressed=False)
    with open(base + ".pickle", "rb") as f:           # what utilities.py:98-99 does
        back = pickle.load(f)
    assert all(np.array_equal(back[v][u], table[v, u]) for v in range(4) for u in range(4))
    back2 = rti_b200.read_from_file(base, compressed=False)
    assert np.array_equal(back2[3][1], table[3, 1])
    # anything that is not the table goes through pickle.dump unchanged
    rti_b200.write_on_file({"a": 1}, str(tmp_path / "other"), compressed=False)
    assert rti_b200.read_from_file(str(tmp_path / "other"), compressed=False) == {"a": 1}
    rti_b200.write_on_file(images, str(tmp_path / "z"), compressed=True)
    assert np.array_equal(rti_b200.read_from_file(str(tmp_path / "z"))[1][2], table[1, 2])


def test_load_table_maps_the_result_pickle(tmp_path):
    """SURVEY 8(f1) loader: interactive_relighting.py:95 un-pickles the 6.4 GB table; a table written by
    this package is memory-mapped instead (copy-on-write), same object graph, in-place clamp allowed."""
    import pickle
    import rti_b200
    from rti_b200 import table_io
    rng = np.random.default_rng(5)
    table = rng.integers(-300, 600, (4, 6, 5, 8)).astype(np.int32)
    path = str(tmp_path / "interpolation_results_x")
    with open(path + ".pickle", "wb") as f:
        table_io.dump_table(table, f)
    got = rti_b200.read_from_file(path, compressed=False)
    ref = pickle.load(open(path + ".pickle", "rb"))
    assert isinstance(got, list) and len(got) == 4 and all(isinstance(r, list) and len(r) == 6 for r in got)
    for v in range(4):
        for u in range(6):
            assert got[v][u].dtype == np.int32 and got[v][u].shape == (5, 8) and got[v][u].flags.writeable
            assert np.array_equal(got[v][u], ref[v][u])
    assert got[0][0].base is not None and table_io.load_table(path + ".pickle") is not None
    values = got[1][2]
    values[values > 255] = 255
    values[values <= 0] = 0                                   # private pages: the file is unchanged
    assert np.array_equal(pickle.load(open(path + ".pickle", "rb"))[1][2], table[1, 2])
    # files of any other shape fall back to pickle.load
    other = str(tmp_path / "other")
    with open(other + ".pickle", "wb") as f:
        pickle.dump([[table[v, u] for u in range(6)] for v in range(4)], f)
    assert table_io.load_table(other + ".pickle") is None
    back = rti_b200.read_from_file(other, compressed=False)
    assert np.array_equal(np.array(back), table)
    with open(other + ".pickle", "wb") as f:
        pickle.dump({"not": "a table"}, f)
    assert rti_b200.read_from_file(other, compressed=False) == {"not": "a table"}
    # a truncated table is not mapped
    raw = open(path + ".pickle", "rb").read()
    open(path + ".pickle", "wb").write(raw[:-7])
    assert table_io.load_table(path + ".pickle") is None


def test_frame_records_must_fit_uint8():
    """an int32 ROI value of 300 must not become 44 silently (the reference keeps it in its
    int32 array, analysis.py:219/232); in-range integral data of any dtype is accepted as is."""
    from rti_b200 import ingest
    roi = np.arange(12, dtype=np.int32).reshape(3, 4) * 20
    assert np.array_equal(ingest.roi_as_uint8(roi, 0), roi.astype(np.uint8))
    assert np.array_equal(ingest.roi_as_uint8(roi.astype(np.float64), 0), roi.astype(np.uint8))
    for bad in (roi + 100, roi - 1, roi.astype(np.float32) + 0.5):
        with pytest.raises(ValueError):
            ingest.roi_as_uint8(bad, 3)
    recs = [(roi + 100, np.zeros(3))]
    with pytest.raises(ValueError):
        analysis._stack_frames(recs, 3, 4)


def test_write_on_file_falls_back_when_the_template_does_not_apply(tmp_path, monkeypatch):
    """A numpy whose ndarray pickle opcodes differ must give the reference's plain pickle.dump, decided
    before the output file is opened."""
    import pickle
    from rti_b200 import table_io
    table = np.arange(2 * 3 * 4 * 5, dtype=np.int32).reshape(2, 3, 4, 5)
    nested = [[table[v, u] for u in range(3)] for v in range(2)]
    assert table_io.template_ok((4, 5), np.int32)

    def broken(shape, dtype):
        raise RuntimeError("unexpected memo opcode")
    monkeypatch.setattr(table_io, "_array_template", broken)
    assert not table_io.template_ok((4, 5), np.int32)
    name = str(tmp_path / "fallback")
    rti_b200.write_on_file(nested, name, compressed=False)
    with open(name + ".pickle", "rb") as f:
        back = pickle.load(f)
    assert np.array_equal(np.array(back), table)


def test_table_tile_writer_equals_pickle_dump(tmp_path):
    """TableTileWriter lays the result pickle out up front and drops ROW TILES of every image at fixed
    offsets (what compute() does with device tiles); the file must be byte-identical to the streaming
    writer's and un-pickle to the reference's object graph (analysis.py:400-409, 475)."""
    import pickle
    from rti_b200 import table_io
    rng = np.random.default_rng(3)
    gv, gu, rows, cols = 3, 4, 9, 5
    table = rng.integers(-2 ** 31, 2 ** 31 - 1, (gv, gu, rows, cols)).astype(np.int32)
    path = str(tmp_path / "tiles.pickle")
    w = table_io.TableTileWriter(gv, gu, (rows, cols), np.int32, path)
    flat = table.reshape(gv * gu, rows, cols)
    for r0, tr in ((4, 3), (0, 4), (7, 2)):                      # tiles in any order
        for i0, cnt in ((0, 5), (5, 7)):                         # image chunks as the pinned ring delivers them
            w.write_rows(i0, flat[i0:i0 + cnt, r0:r0 + tr].reshape(cnt, -1), r0)
    w.close()
    with open(path, "rb") as f:
        raw = f.read()
    assert raw == table_io.dumps_table(table)
    back = pickle.loads(raw)
    assert len(back) == gv and len(back[0]) == gu
    assert all(np.array_equal(back[v][u], table[v, u]) and back[v][u].flags.writeable
               for v in range(gv) for u in range(gu))
    mapped = table_io.load_table(path)
    assert mapped is not None and np.array_equal(mapped[2][3], table[2, 3])


def test_rbf_pcg_entry_validates_arguments_without_gpu(lib):
    import ctypes as C
    from rti_b200 import _capi
    assert lib.rti_rbf_pcg_workspace_bytes(2, 1) == 0 and lib.rti_rbf_pcg_workspace_bytes(1026, 1) == 0
    one, two = lib.rti_rbf_pcg_workspace_bytes(500, 1), lib.rti_rbf_pcg_workspace_bytes(500, 2)
    assert 0 < one < two and one % 256 == 0
    buf = (C.c_double * 64)()
    p = C.cast(buf, C.c_void_p)
    assert lib.rti_rbf_fit_pcg(None, p, p, 4, 4, 40, p, p, one, 8, 0.01, None, None, None) == _capi.RTI_ERR_ARG
    assert lib.rti_rbf_fit_pcg(p, p, p, 4, 4, 2, p, p, one, 8, 0.01, None, None, None) == _capi.RTI_ERR_ARG   # n < 3
    assert lib.rti_rbf_fit_pcg(p, p, p, 4, 4, 40, p, p, 1024, 8, 0.01, None, None, None) == _capi.RTI_ERR_ARG  # workspace
    assert lib.rti_rbf_fit_pcg(p, p, p, 4, 4, 40, p, p, one, 0, 0.01, None, None, None) == _capi.RTI_ERR_ARG   # max_iters
    assert lib.rti_rbf_table_fast(p, p, p, 0, 40, p, 100, p, 4, None, None) == _capi.RTI_ERR_ARG

#!/usr/bin/env python
"""Generate the golden fixtures in this directory by running the UNMODIFIED
reference (imported from /root/reference) on small seeded inputs.

Runs only in the build container (the reference is not present on the GPU
box); the outputs are committed.  Usage:

    python tests/golden/make_golden.py

Writes
    ref_small.npz   frames/cams + compute_intensities, interpolate_intensities
                    (PTM and RBF) and prepare_images_data outputs, R=6, N=40
    ref_rbf_n96.npz two pixels of the RBF path at N=96
    relight.npz     cursor->index table of relighting_event / draw_light_roi_position,
                    and a random HSV image with cv2.cvtColor(HSV2BGR) output
    roi.npz         ut.get_ROI(img, corners, hsv=True) crops of a seeded BGR image
                    (`make_golden.py roi` regenerates only this one)
    relight_rows.npz cv2.cvtColor(HSV2BGR) of images whose width is not a multiple of the
                    SIMD width (the reference's default 400-wide ROI among them): cvtColor's
                    per-row scalar tail rounds where its vector body truncates
                    (`make_golden.py relight_rows` regenerates only this one)
and checks the oracle's HSV->BGR and BGR->HSV restatements exhaustively against cv2.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("RTI_REFERENCE_DIR", "/root/reference")


def import_reference():
    """Import analysis.py with matplotlib stubbed (it is imported at module top
    but only used by the first_only debug plots)."""
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.path.insert(0, REF)
    cwd = os.getcwd()
    os.chdir(REF)
    try:
        import analysis
        import constants
        from Utils import utilities
    finally:
        os.chdir(cwd)
    return analysis, constants, utilities


def synth_frames(R, N, seed):
    """Seeded frame records [(uint8 (R,R), float64 (3,))] (SURVEY.md 8d geometry)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:R, 0:R]
    albedo = 0.55 + 0.4 * np.sin(0.9 * xx + 0.3) * np.cos(0.7 * yy)
    nx = 0.35 * np.sin(0.8 * xx + 0.5 * yy)
    ny = 0.35 * np.cos(0.6 * yy - 0.4 * xx)
    nz = np.sqrt(1 - nx ** 2 - ny ** 2)
    frames = []
    for _ in range(N):
        th = rng.uniform(0, 2 * np.pi)
        rho = rng.uniform(0, 1.5 * R)
        z = rng.uniform(0.75 * R, 2.25 * R)
        cam = np.array([R / 2 + rho * np.cos(th), R / 2 + rho * np.sin(th), z])
        d = np.stack([cam[0] - xx, cam[1] - yy, np.full_like(xx, cam[2], dtype=float)], -1)
        d = d / np.linalg.norm(d, axis=-1, keepdims=True)
        shade = np.clip(nx * d[..., 0] + ny * d[..., 1] + nz * d[..., 2], 0, None)
        img = 255 * albedo * shade + rng.normal(0, 4, (R, R))
        frames.append((np.clip(img, 0, 255).astype(np.uint8), cam))
    return frames


def make_roi_golden(cst, ut, orc):
    """roi.npz: ut.get_ROI(img, corners, hsv=True) (Utils/utilities.py:302-342) on a seeded
    BGR image for corner sets that hit every clamping branch; exhaustive BGR2HSV check."""
    import cv2
    rng = np.random.default_rng(21)
    h, w = 90, 120
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[10:30, 20:50] = rng.integers(0, 256, (20, 30, 1), dtype=np.uint8)      # grey patch: diff = 0
    diam = 40
    cst.ROI_DIAMETER = diam
    corner_sets = [
        [(40, 30), (80, 30), (80, 60), (40, 60)],        # interior
        [(2, 3), (20, 3), (20, 25), (2, 25)],            # x_min, y_min < 0 -> 1
        [(100, 70), (119, 70), (119, 89), (100, 89)],    # x_max > w, y_max > h -> size - 1
        [(41, 30), (80, 31), (80, 60), (41, 61)],        # odd extents: round-half-even centre
        [(0, 40), (40, 40), (40, 50), (0, 50)],          # x_min == 0 exactly (not clamped)
    ]
    out = {"img": img, "diameter": np.array(diam), "corners": np.array(corner_sets)}
    for k, cs in enumerate(corner_sets):
        roi = ut.get_ROI(img, cs, hsv=True)
        out["roi%d" % k] = roi
        mine = orc.get_roi_hsv(img, cs, diam)
        print("roi case", k, "shape", roi.shape, "oracle equal", np.array_equal(roi, mine))
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(HERE, "roi.npz"), **out)
    a = np.arange(1 << 24, dtype=np.uint32)
    cube = np.stack([a & 255, (a >> 8) & 255, (a >> 16) & 255], -1).astype(np.uint8).reshape(4096, 4096, 3)
    bad = np.count_nonzero(cv2.cvtColor(cube, cv2.COLOR_BGR2HSV) != orc.bgr2hsv_u8(cube))
    print("BGR2HSV exhaustive mismatches vs cv2 %s: %d of %d" % (cv2.__version__, bad, cube.size))


def make_relight_rows_golden(orc):
    """relight_rows.npz: what interactive_relighting.py:38 returns for ROI widths with a scalar
    tail (400 = the reference's ROI_DIAMETER, 399, 37, 31) and without one (64)."""
    import cv2
    rng = np.random.default_rng(11)
    out = {"cv2_version": np.array(cv2.__version__)}
    for rows, cols in ((12, 400), (5, 399), (9, 37), (7, 31), (3, 64)):
        hsv = np.stack([rng.integers(0, 180, (rows, cols)), rng.integers(0, 256, (rows, cols)),
                        rng.integers(0, 256, (rows, cols))], -1).astype(np.uint8)
        bgr = cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR)
        out["hsv_%dx%d" % (rows, cols)] = hsv
        out["bgr_%dx%d" % (rows, cols)] = bgr
        print("relight_rows %dx%d: oracle mismatches %d, vector-body-only mismatches %d"
              % (rows, cols, np.count_nonzero(orc.hsv2bgr_u8(hsv) != bgr),
                 np.count_nonzero(orc.hsv2bgr_u8(hsv, simd_lanes=0) != bgr)))
    np.savez_compressed(os.path.join(HERE, "relight_rows.npz"), **out)
    # exhaustive: every (H, S, V) once in the vector body (width 256) and once in the tail (width 31)
    hh, ss, vv = np.meshgrid(np.arange(180, dtype=np.uint8), np.arange(256, dtype=np.uint8),
                             np.arange(256, dtype=np.uint8), indexing="ij")
    flat = np.stack([hh, ss, vv], -1).reshape(-1, 3)
    body = flat.reshape(180 * 256, 256, 3)
    pad = (-flat.shape[0]) % 31
    tail = np.concatenate([flat, np.zeros((pad, 3), np.uint8)]).reshape(-1, 31, 3)
    for name, img in (("vector body", body), ("scalar tail", tail)):
        bad = np.count_nonzero(cv2.cvtColor(img, cv2.COLOR_HSV2BGR) != orc.hsv2bgr_u8(img))
        print("HSV2BGR exhaustive (%s) mismatches vs cv2 %s: %d of %d" % (name, cv2.__version__, bad, img.size))


def main():
    analysis, cst, ut = import_reference()
    sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
    import rti_oracle as orc
    if sys.argv[1:] == ["roi"]:
        return make_roi_golden(cst, ut, orc)
    if sys.argv[1:] == ["relight_rows"]:
        return make_relight_rows_golden(orc)

    # ---- small end-to-end case --------------------------------------------------
    R, N = 6, 40
    cst.ROI_DIAMETER = R
    frames = synth_frames(R, N, seed=1234)
    lx, ly, inten = analysis.compute_intensities(frames)
    ptm = analysis.interpolate_intensities((lx, ly, inten), interpolate_PTM=True)
    rbf = analysis.interpolate_intensities((lx, ly, inten), interpolate_PTM=False)
    tab_ptm = np.array(analysis.prepare_images_data(ptm))        # (100,100,R,R) int32
    tab_rbf = np.array(analysis.prepare_images_data(rbf))
    keep = [(0, 0), (0, R - 1), (2, 3), (R - 1, R - 1)]          # (y, x) pixels kept in fp64
    np.savez_compressed(
        os.path.join(HERE, "ref_small.npz"),
        frames=np.stack([f[0] for f in frames]), cams=np.stack([f[1] for f in frames]),
        lx=lx, ly=ly, inten=inten,
        keep=np.array(keep),
        ptm_grids=np.stack([ptm[y][x] for y, x in keep]),
        rbf_grids=np.stack([rbf[y][x] for y, x in keep]),
        table_ptm=tab_ptm, table_rbf=tab_rbf)

    # ---- RBF at a larger N ------------------------------------------------------
    R2, N2 = 3, 96
    cst.ROI_DIAMETER = R2
    frames2 = synth_frames(R2, N2, seed=99)
    d2 = analysis.compute_intensities(frames2)
    ax = np.around(np.mgrid[-1:1:cst.INTERPOLATION_PARAM], 2)
    yi, xi = np.mgrid[-1:1:cst.INTERPOLATION_PARAM, -1:1:cst.INTERPOLATION_PARAM]
    yi, xi = np.around(yi, 2), np.around(xi, 2)
    pix = [(0, 0), (2, 1)]
    g_rbf = np.stack([analysis._interpolate_RBF(d2[0][y][x], d2[1][y][x], xi, yi, d2[2][y][x])
                      for y, x in pix])
    g_ptm = np.stack([analysis._interpolate_PTM(d2[0][y][x], d2[1][y][x], xi[0], d2[2][y][x])
                      for y, x in pix])
    np.savez_compressed(
        os.path.join(HERE, "ref_rbf_n96.npz"),
        frames=np.stack([f[0] for f in frames2]), cams=np.stack([f[1] for f in frames2]),
        lx=d2[0], ly=d2[1], inten=d2[2], pix=np.array(pix), rbf_grids=g_rbf, ptm_grids=g_ptm,
        axis=ax)

    # ---- relighting index arithmetic and HSV->BGR ----------------------------------
    import cv2
    w = h = 400
    xs = np.arange(0, w + 1)
    light = np.array([ut.draw_light_roi_position(int(x), int(x), (h, w), to_light_vector=True)[0]
                      for x in xs])
    index = np.array([round((1 + l) / 2 * 100) for l in light])
    rng = np.random.default_rng(7)
    hsv = np.stack([rng.integers(0, 180, (64, 64)), rng.integers(0, 256, (64, 64)),
                    rng.integers(0, 256, (64, 64))], -1).astype(np.uint8)
    bgr = cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR)
    np.savez_compressed(os.path.join(HERE, "relight.npz"), cursor=xs, light=light, index=index,
                        hsv=hsv, bgr=bgr, cv2_version=np.array(cv2.__version__))

    # exhaustive HSV2BGR check of the oracle restatement against cv2
    hh, ss, vv = np.meshgrid(np.arange(180, dtype=np.uint8), np.arange(256, dtype=np.uint8),
                             np.arange(256, dtype=np.uint8), indexing="ij")
    cube = np.stack([hh, ss, vv], -1).reshape(180 * 256, 256, 3)
    bad = np.count_nonzero(cv2.cvtColor(cube, cv2.COLOR_HSV2BGR) != orc.hsv2bgr_u8(cube))
    print("HSV2BGR exhaustive mismatches vs cv2 %s: %d of %d" % (cv2.__version__, bad, cube.size))

    # quick self-check of the oracle against what was just written
    o_lx, o_ly, o_in = orc.gather_soa(np.stack([f[0] for f in frames]),
                                      np.stack([f[1] for f in frames]))
    print("gather: lx equal", np.array_equal(o_lx, lx), "ly equal", np.array_equal(o_ly, ly),
          "inten equal", np.array_equal(o_in, inten))
    og = orc.interpolate(lx, ly, inten, ptm=True)
    print("ptm grid max rel diff",
          max(np.max(np.abs(og[y, x] - ptm[y][x])) / np.max(np.abs(ptm[y][x])) for y, x in keep))
    print("ptm table equal", np.array_equal(orc.table_from_grids(og), tab_ptm))
    og = orc.interpolate(lx, ly, inten, ptm=False)
    print("rbf grid max rel diff",
          max(np.max(np.abs(og[y, x] - rbf[y][x])) / np.max(np.abs(rbf[y][x])) for y, x in keep))
    print("rbf table mismatches", np.count_nonzero(orc.table_from_grids(og) != tab_rbf))
    make_roi_golden(cst, ut, orc)
    make_relight_rows_golden(orc)


if __name__ == "__main__":
    main()

"""Seeded synthetic frame records in the reference's layout: a list of
(uint8 (R, R) ROI intensities, float64 (3,) camera position) per frame, as
FeatureMatcher.extractFeatures emits them (FeatureMatcher.py:183-193, 217)."""
from __future__ import annotations

import numpy as np


def camera_positions(rows, cols, n, rng):
    """cam = (cols/2 + rho cos t, rows/2 + rho sin t, z): light directions cover the unit disc.
    Draws are rejected until all positions are distinct (duplicate RBF nodes are singular)."""
    size = max(rows, cols)
    th = rng.uniform(0, 2 * np.pi, n)
    rho = rng.uniform(0, 1.5 * size, n)
    z = rng.uniform(0.75 * size, 2.25 * size, n)
    cams = np.stack([cols / 2 + rho * np.cos(th), rows / 2 + rho * np.sin(th), z], axis=1)
    assert len(np.unique(cams, axis=0)) == n
    return np.ascontiguousarray(cams, dtype=np.float64)


def lambertian_stack(rows, cols, n, seed=0, noise=4.0, row0=0, rows_total=None):
    """(frames uint8 (n, rows, cols), cams float64 (n, 3)) of a smooth synthetic surface lit
    from the camera positions.  rows/row0 select a slab of a rows_total-high ROI."""
    rows_total = rows if rows_total is None else rows_total
    rng = np.random.default_rng(seed)
    cams = camera_positions(rows_total, cols, n, rng)
    yy, xx = np.mgrid[row0:row0 + rows, 0:cols].astype(np.float64)
    albedo = 0.55 + 0.4 * np.sin(0.09 * xx + 0.3) * np.cos(0.07 * yy)
    nx = 0.35 * np.sin(0.08 * xx + 0.05 * yy)
    ny = 0.35 * np.cos(0.06 * yy - 0.04 * xx)
    nz = np.sqrt(1 - nx ** 2 - ny ** 2)
    frames = np.empty((n, rows, cols), dtype=np.uint8)
    noise_rng = np.random.default_rng(seed + 1000003 * (row0 + 1))
    for i in range(n):
        dx, dy, dz = cams[i, 0] - xx, cams[i, 1] - yy, cams[i, 2]
        inv = 1.0 / np.sqrt(dx * dx + dy * dy + dz * dz)
        shade = np.clip((nx * dx + ny * dy + nz * dz) * inv, 0, None)
        img = 255 * albedo * shade + noise_rng.normal(0, noise, (rows, cols))
        frames[i] = np.clip(img, 0, 255).astype(np.uint8)
    return frames, cams


def random_stack(rows, cols, n, seed=0, rows_total=None):
    """Uniform random intensities (pure-throughput runs)."""
    rows_total = rows if rows_total is None else rows_total
    rng = np.random.default_rng(seed)
    cams = camera_positions(rows_total, cols, n, rng)
    frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)
    return frames, cams


def as_records(frames, cams):
    """The reference's list-of-tuples form of a stack."""
    return [(frames[i], cams[i]) for i in range(len(frames))]

"""Synthetic scene builders shared by tests and bench (inputs only; no compute).

`cloth_sheet` reproduces the reference smoke-test sheet (multibody/gpu_mpm/cuda_mpm_test.cc:27-57):
res x res vertices on a square of side `side` starting at (x0, y0, z), two triangles per quad with
the reference's winding.  `stacked_sheets` is the SURVEY 8d cfg-2 scene: `n_sheets` such sheets,
z from z0 in steps of dz, with a seeded xy jitter of U(-0.1, 0.1) * dx per sheet.
"""
import numpy as np

SEED = 20261104


def cloth_sheet(res=100, side=0.5, x0=0.25, y0=0.25, z=0.75, dtype=np.float64):
    dx = side / res
    i, j = np.meshgrid(np.arange(res), np.arange(res), indexing="ij")
    pos = np.stack([x0 + i * dx, y0 + j * dx, np.full_like(i, z, dtype=np.float64)], axis=-1).reshape(-1, 3)
    p = lambda a, b: a * res + b
    ii, jj = np.meshgrid(np.arange(res - 1), np.arange(res - 1), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    t0 = np.stack([p(ii, jj), p(ii + 1, jj), p(ii, jj + 1)], axis=-1)
    t1 = np.stack([p(ii + 1, jj + 1), p(ii, jj + 1), p(ii + 1, jj)], axis=-1)
    tri = np.stack([t0, t1], axis=1).reshape(-1, 3).astype(np.int32)
    # the reference builds positions as T(0.25 + i * dx) with dx of type T
    if np.dtype(dtype) == np.float32:
        dxf = np.float32(np.float32(side) / np.float32(res))
        pos = np.stack([(np.float64(x0) + i.astype(np.float64) * np.float64(dxf)),
                        (np.float64(y0) + j.astype(np.float64) * np.float64(dxf)),
                        np.full(i.shape, z, dtype=np.float64)], axis=-1).reshape(-1, 3)
    pos = pos.astype(dtype)
    vel = np.zeros_like(pos)
    return pos, vel, tri


def stacked_sheets(n_sheets, res=100, side=0.5, z0=0.15, dz=0.02, domain_bits=7, jitter=0.1, dtype=np.float64,
                   seed=SEED, warm=False):
    """Returns a list of (pos, vel, tri) -- one AddQRCloth call each."""
    rng = np.random.default_rng(seed)
    dx = 1.0 / (1 << domain_bits)
    out = []
    for s in range(n_sheets):
        off = rng.uniform(-jitter, jitter, size=2) * dx
        pos, vel, tri = cloth_sheet(res, side, 0.25 + off[0], 0.25 + off[1], z0 + s * dz, dtype)
        if warm:
            vel = rng.normal(0.0, 0.1, size=pos.shape).astype(dtype)
        out.append((pos, vel, tri))
    return out


def box_contacts(pos, lo, hi, rigid_id=1, rigid_v=(0.0, 0.0, 0.0), rigid_p_WB=None):
    """Contact pairs of points against an axis-aligned rigid box [lo, hi], the way
    DeformableDriver::CalcMpmContactPairs builds them (deformable_driver.h:120-194) from
    ComputeSignedDistanceToPoint: every point strictly inside the box is one pair with
    penetration_distance = signed distance (< 0), normal = -grad_W (grad_W = outward normal of the
    nearest face), position = the point, rigid_v, rigid_p_WB (body origin).
    Returns (particle_index, non_mpm_id, distance, normal, position, rigid_v, rigid_p_WB) arrays."""
    pos = np.asarray(pos, np.float64).reshape(-1, 3)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    inside = np.all((pos > lo) & (pos < hi), axis=1)
    idx = np.nonzero(inside)[0]
    p = pos[idx]
    d = np.concatenate([p - lo, hi - p], axis=1)          # distance to the 6 faces
    face = np.argmin(d, axis=1)
    dist = -d[np.arange(len(idx)), face]
    grad = np.zeros((len(idx), 3))
    grad[np.arange(len(idx)), face % 3] = np.where(face < 3, -1.0, 1.0)
    center = 0.5 * (lo + hi) if rigid_p_WB is None else np.asarray(rigid_p_WB, np.float64)
    n = len(idx)
    return (idx.astype(np.uint32), np.full(n, rigid_id, np.uint32), dist, -grad, p,
            np.tile(np.asarray(rigid_v, np.float64), (n, 1)), np.tile(center, (n, 1)))


def particle_block(lo, hi, per_axis=2, domain_bits=7, dx=None, jitter=0.25, seed=SEED, dtype=np.float64):
    """Volumetric body for the particle-only API (SURVEY 8d "1M-particle volumetric block, 8 ppc, jittered lattice"):
    `per_axis`^3 particles per grid cell inside the box [lo, hi], jittered by U(-jitter, jitter) of the particle
    spacing.  Returns (pos, vel, volume_per_particle)."""
    dx = dx if dx is not None else 1.0 / (1 << domain_bits)
    h = dx / per_axis
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    n = np.maximum(1, np.round((hi - lo) / h).astype(int))
    g = [lo[d] + (np.arange(n[d]) + 0.5) * h for d in range(3)]
    pos = np.stack(np.meshgrid(*g, indexing="ij"), axis=-1).reshape(-1, 3)
    rng = np.random.default_rng(seed)
    pos = pos + rng.uniform(-jitter, jitter, pos.shape) * h
    return pos.astype(dtype), np.zeros_like(pos, dtype=dtype), h ** 3

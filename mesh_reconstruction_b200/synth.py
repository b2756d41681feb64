"""Synthetic occupancy grids for the BASELINE configs.

All generators return float32 arrays of exactly 0.0 / 1.0 in the reference storage order
[x][y][z], with the shape kept >= `pad` voxels away from every border (the reference pads
its grids by 6 voxels, grid.cpp:91-114).
"""
import numpy as np


def _coords(n):
    if isinstance(n, int):
        n = (n, n, n)
    return np.meshgrid(*[np.arange(s, dtype=np.float32) for s in n], indexing="ij")


def sphere(n, radius, center=None):
    if isinstance(n, int):
        n = (n, n, n)
    if center is None:
        center = [(s - 1) / 2.0 for s in n]
    x, y, z = _coords(n)
    d2 = (x - center[0]) ** 2 + (y - center[1]) ** 2 + (z - center[2]) ** 2
    return (d2 <= radius * radius).astype(np.float32)


def torus(n, R, r, center=None):
    if isinstance(n, int):
        n = (n, n, n)
    if center is None:
        center = [(s - 1) / 2.0 for s in n]
    x, y, z = _coords(n)
    q = np.sqrt((x - center[0]) ** 2 + (y - center[1]) ** 2) - R
    return ((q * q + (z - center[2]) ** 2) <= r * r).astype(np.float32)


def box(n, half, center=None, rot_z_deg=0.0, rot_x_deg=0.0):
    if isinstance(n, int):
        n = (n, n, n)
    if center is None:
        center = [(s - 1) / 2.0 for s in n]
    x, y, z = _coords(n)
    x = x - center[0]; y = y - center[1]; z = z - center[2]
    a = np.deg2rad(rot_z_deg); b = np.deg2rad(rot_x_deg)
    x1 = np.cos(a) * x + np.sin(a) * y
    y1 = -np.sin(a) * x + np.cos(a) * y
    y2 = np.cos(b) * y1 + np.sin(b) * z
    z2 = -np.sin(b) * y1 + np.cos(b) * z
    return ((np.abs(x1) <= half[0]) & (np.abs(y2) <= half[1]) & (np.abs(z2) <= half[2])).astype(np.float32)


def ellipsoid(n, radii, center=None):
    if isinstance(n, int):
        n = (n, n, n)
    if center is None:
        center = [(s - 1) / 2.0 for s in n]
    x, y, z = _coords(n)
    q = ((x - center[0]) / radii[0]) ** 2 + ((y - center[1]) / radii[1]) ** 2 + ((z - center[2]) / radii[2]) ** 2
    return (q <= 1.0).astype(np.float32)


def union(*vols):
    out = vols[0].copy()
    for v in vols[1:]:
        out = np.maximum(out, v)
    return out


def random_spheres(n, count, rmin, rmax, seed, pad=6):
    """Union of `count` random spheres fully inside the padded interior (config 1)."""
    rng = np.random.default_rng(seed)
    if isinstance(n, int):
        n = (n, n, n)
    out = np.zeros(n, np.float32)
    for _ in range(count):
        r = rng.uniform(rmin, rmax)
        c = [rng.uniform(pad + r + 1, s - pad - r - 2) for s in n]
        out = np.maximum(out, sphere(n, r, c))
    return out


def random_object(seed, n=52, pad=6):
    """Config 5: one random 40^3-class object (sphere, ellipsoid, box, torus or 2-sphere union)."""
    rng = np.random.default_rng(seed)
    kind = int(rng.integers(0, 5))
    inner = n - 2 * pad
    c = [(n - 1) / 2.0] * 3
    if kind == 0:
        return sphere(n, rng.uniform(8, inner / 2 - 2))
    if kind == 1:
        return ellipsoid(n, [rng.uniform(6, inner / 2 - 2) for _ in range(3)])
    if kind == 2:
        return box(n, [rng.uniform(5, inner / 2 - 8) for _ in range(3)], c, rng.uniform(0, 45), rng.uniform(0, 45))
    if kind == 3:
        r = rng.uniform(3, 5)
        return torus(n, rng.uniform(8, inner / 2 - r - 2), r)
    return random_spheres(n, 2, 6, 10, seed + 7, pad)


def config_volume(cfg):
    """The headline synthetic volume of BASELINE config `cfg` (1..4)."""
    if cfg == 1:
        return sphere(52, 17.0)
    if cfg == 2:
        return union(box(172, (50, 40, 30), None, 20.0, 35.0), sphere(172, 40.0))
    if cfg == 3:
        return sphere(256, 110.0)
    if cfg == 4:
        v = sphere(512, 220.0)
        rng = np.random.default_rng(4)
        for _ in range(8):
            r = rng.uniform(20, 40)
            u = rng.normal(size=3); u /= np.linalg.norm(u)
            c = [255.5 + 215.0 * u[i] for i in range(3)]
            c = [min(max(ci, 6 + r + 1), 512 - 6 - r - 2) for ci in c]
            v = np.maximum(v, sphere(512, r, c))
        return v
    raise ValueError(cfg)


def write_binvox(path, occ, translate=(0.0, 0.0, 0.0), scale=1.0):
    """Writes a binvox file (RLE) for occupancy occ[x][y][z]; voxel order x*wxh + z*width + y
    (binvoxToPcl.cpp:9-13)."""
    occ = np.asarray(occ) != 0
    d, w, h = occ.shape[0], occ.shape[1], occ.shape[2]          # depth(x), width(y), height(z)
    flat = np.ascontiguousarray(occ.transpose(0, 2, 1)).reshape(-1).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(b"#binvox 1\n")
        f.write(("dim %d %d %d\n" % (d, h, w)).encode())
        f.write(("translate %g %g %g\n" % tuple(translate)).encode())
        f.write(("scale %g\n" % scale).encode())
        f.write(b"data\n")
        i, n = 0, len(flat)
        out = bytearray()
        while i < n:
            v = flat[i]
            j = i
            while j < n and flat[j] == v and j - i < 255:
                j += 1
            out += bytes([int(v), j - i])
            i = j
        f.write(bytes(out))


def write_pcd_ascii(path, pts):
    pts = np.asarray(pts, np.float32)
    with open(path, "w") as f:
        f.write("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\n"
                "COUNT 1 1 1\nWIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA ascii\n" % (len(pts), len(pts)))
        for p in pts:
            f.write("%.7g %.7g %.7g\n" % (p[0], p[1], p[2]))


def cli_inputs(dirpath, n=40, radius=12.0):
    """Config 1 at CLI level: a synthetic 40^3 binvox completion of a sphere (world frame [0,1]^3) and an
    ASCII PCD of its +z-facing hemisphere sampled at 4x the voxel density.  Returns (binvox, pcd) paths."""
    import os
    occ = sphere(n, radius)
    bv = os.path.join(dirpath, "completion.binvox")
    pcd = os.path.join(dirpath, "partial.pcd")
    write_binvox(bv, occ)
    c = n / 2.0
    step = 1.0 / (4.0 * n)
    g = np.arange(-radius / n, radius / n + step, step)
    xx, yy = np.meshgrid(g, g, indexing="ij")
    rr = (radius / n) ** 2 - xx ** 2 - yy ** 2
    m = rr > 0
    pts = np.stack([xx[m] + c / n, yy[m] + c / n, np.sqrt(rr[m]) + c / n], 1)
    write_pcd_ascii(pcd, pts)
    return bv, pcd


# ---- the same generators on a torch device (the 512^3 volume takes numpy tens of seconds on a busy host) ----------
def sphere_t(n, radius, center=None, device="cpu"):
    """synth.sphere evaluated with torch on `device`: the same fp32 operations in the same order, hence the same voxels."""
    import torch
    if isinstance(n, int):
        n = (n, n, n)
    if center is None:
        center = [(s - 1) / 2.0 for s in n]
    ax = [torch.arange(s, dtype=torch.float32, device=device) for s in n]
    dx = (ax[0] - np.float32(center[0]).item()) ** 2
    dy = (ax[1] - np.float32(center[1]).item()) ** 2
    dz = (ax[2] - np.float32(center[2]).item()) ** 2
    d2 = (dx[:, None, None] + dy[None, :, None]) + dz[None, None, :]
    return (d2 <= np.float32(radius * radius).item()).to(torch.float32)


def config_volume_t(cfg, device="cpu"):
    """config_volume(cfg) for the sphere-built configs (1, 3, 4) as a torch tensor on `device`."""
    import torch
    if cfg == 1:
        return sphere_t(52, 17.0, device=device)
    if cfg == 3:
        return sphere_t(256, 110.0, device=device)
    if cfg == 4:
        v = sphere_t(512, 220.0, device=device)
        rng = np.random.default_rng(4)
        for _ in range(8):
            r = rng.uniform(20, 40)
            u = rng.normal(size=3); u /= np.linalg.norm(u)
            c = [255.5 + 215.0 * u[i] for i in range(3)]
            c = [min(max(ci, 6 + r + 1), 512 - 6 - r - 2) for ci in c]
            v = torch.maximum(v, sphere_t(512, r, c, device=device))
        return v
    raise ValueError(cfg)


def completion_scene(n=40, kind="sphere", seed=0, view_fraction=0.5, observed_step=0.35, noise=0.002, z_shift=0.0,
                     xform=(0.3, -0.12, 0.05, 0.6)):
    """A synthetic input pair of the CLI (mesh_reconstruction.cpp:99-106): the shape-completion network's binvox grid and a
    partial, denser, noisy point cloud of the same object as a depth camera would see it.

    Returns (voxels, dwh, xform, observed): voxel bytes in binvox file order (index = x*width*height + z*width + y),
    dwh = (depth, width, height), xform = (scale, tx, ty, tz), observed float32 [m, 3] in the binvox file's frame.
    z_shift moves the observed points along z by that many binvox cells, which opens the gap fillGrid closes."""
    rng = np.random.default_rng(seed)
    d = w = h = int(n)
    x, z, y = np.meshgrid(np.arange(d, dtype=np.float32), np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    c = (n - 1) / 2.0
    if kind == "sphere":
        inside = (x - c) ** 2 + (y - c) ** 2 + (z - c) ** 2 <= (0.3 * n) ** 2
    elif kind == "box":
        inside = (np.abs(x - c) <= 0.28 * n) & (np.abs(y - c) <= 0.2 * n) & (np.abs(z - c) <= 0.24 * n)
    elif kind == "blobs":
        inside = np.zeros_like(x, dtype=bool)
        for _ in range(4):
            cc = rng.uniform(0.35 * n, 0.65 * n, 3)
            r = rng.uniform(0.12 * n, 0.22 * n)
            inside |= (x - cc[0]) ** 2 + (y - cc[1]) ** 2 + (z - cc[2]) ** 2 <= r * r
    else:
        raise ValueError(kind)
    vox = inside.astype(np.uint8)                     # [x][z][y] = file order
    # observed: surface samples on a finer lattice, the part with x below a cut ("seen from -x"), noisy
    s = np.arange(0.0, n - 1.0, observed_step, dtype=np.float32)
    gx, gy, gz = np.meshgrid(s, s, s, indexing="ij")
    def occ(ax, ay, az):
        ix, iy, iz = np.clip(np.rint(ax).astype(int), 0, n - 1), np.clip(np.rint(ay).astype(int), 0, n - 1), np.clip(np.rint(az).astype(int), 0, n - 1)
        return vox[ix, iz, iy] > 0
    o = occ(gx, gy, gz)
    surf = o & ~(occ(gx - 1, gy, gz) & occ(gx + 1, gy, gz) & occ(gx, gy - 1, gz) & occ(gx, gy + 1, gz) & occ(gx, gy, gz - 1) & occ(gx, gy, gz + 1))
    keep = surf & (gx <= np.quantile(gx[surf], view_fraction))
    p = np.stack([gx[keep], gy[keep], gz[keep] + np.float32(z_shift)], 1).astype(np.float32)
    p += rng.normal(0.0, noise * n, p.shape).astype(np.float32)
    sc, tx, ty, tz = [np.float32(t) for t in xform]
    obs = np.empty_like(p)
    obs[:, 0] = (p[:, 0] + np.float32(0.5)) * sc / np.float32(d) + tx
    obs[:, 1] = (p[:, 1] + np.float32(0.5)) * sc / np.float32(w) + ty
    obs[:, 2] = (p[:, 2] + np.float32(0.5)) * sc / np.float32(h) + tz
    return vox.ravel().copy(), (d, w, h), (sc, tx, ty, tz), np.ascontiguousarray(obs, np.float32)

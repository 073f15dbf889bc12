"""Synthetic scene generators for the BASELINE.json configs.

PRNG = SplitMix64 with an explicit float conversion u = (x >> 40) * 2^-24 (never a library
distribution), so the same scene can be regenerated bit-for-bit anywhere.  numpy only.
"""
from __future__ import annotations

import numpy as np

_GOLDEN = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed: int, index: np.ndarray) -> np.ndarray:
    """SplitMix64 output for stream position `index` (0-based) of a generator seeded with `seed`."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + (index.astype(np.uint64) + np.uint64(1)) * _GOLDEN
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform01(seed: int, index: np.ndarray) -> np.ndarray:
    return ((splitmix64(seed, index) >> np.uint64(40)).astype(np.float32)) * np.float32(2.0 ** -24)


def uniform_spheres(n: int, radius: float = 1.3e-3, seed: int = 0x5EED0001, chunk: int = 1 << 22):
    """C4: n spheres, centres uniform in [-0.5+r, 0.5-r]^3 (overlaps allowed)."""
    r = np.float32(radius)
    lo = np.float32(-0.5) + r
    span = (np.float32(0.5) - r) - lo
    centers = np.empty((n, 3), dtype=np.float32)
    for start in range(0, n, chunk):
        stop = min(n, start + chunk)
        idx = np.arange(start, stop, dtype=np.uint64) * np.uint64(3)
        for a in range(3):
            centers[start:stop, a] = lo + uniform01(seed, idx + np.uint64(a)) * span
    radii = np.full(n, r, dtype=np.float32)
    return centers, radii


def hcp_spheres(n_target: int, phi: float = 0.55, jitter: float = 0.1, seed: int = 0x5EED0005):
    """C5 frame 0: jittered HCP lattice in the unit cube at volume fraction ~phi.
    Returns (centers, radii) with len <= n_target (lattice truncated to the cube)."""
    # HCP packs at 0.7405 when spheres touch; spacing a = 2 r * (0.7405/phi)^(1/3)
    r = (phi * 3.0 / (4.0 * np.pi * n_target)) ** (1.0 / 3.0)
    a = 2.0 * r * (0.74048 / phi) ** (1.0 / 3.0)
    nx = int(np.floor((1.0 - 2 * r) / a)) + 1
    ny = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(3) / 2))) + 1
    nz = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(6) / 3))) + 1
    i, j, k = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    i, j, k = i.ravel(), j.ravel(), k.ravel()
    x = (i + 0.5 * ((j + k) % 2)) * a
    y = (j + ((k % 2) / 3.0)) * (a * np.sqrt(3) / 2)
    z = k * (a * np.sqrt(6) / 3)
    c = np.stack([x, y, z], 1)
    c = c[(c[:, 0] <= 1 - 2 * r) & (c[:, 1] <= 1 - 2 * r) & (c[:, 2] <= 1 - 2 * r)]
    c = c[:n_target]
    n = len(c)
    idx = np.arange(n, dtype=np.uint64) * np.uint64(3)
    jit = np.stack([uniform01(seed, idx + np.uint64(a)) for a in range(3)], 1)
    c = (c + r - 0.5 + (jit * 2.0 - 1.0) * (jitter * r)).astype(np.float32)
    return np.ascontiguousarray(c), np.full(n, r, dtype=np.float32)


def box_walls(open_plus_x: bool = True):
    """Walls of the cube [-0.5,0.5]^3 as 2 triangles each (the +x wall facing the camera is
    omitted for C4).  Returns (vertices[3n,3], normals[3n,3]) — normals point into the cube."""
    faces = []
    h = 0.5
    # (axis, sign)
    for axis, sign in [(0, -1), (1, -1), (1, 1), (2, -1), (2, 1)] + ([] if open_plus_x else [(0, 1)]):
        u, v = [(1, 2), (2, 0), (0, 1)][axis]
        corners = []
        for su, sv in [(-1, -1), (1, -1), (1, 1), (-1, 1)]:
            p = [0.0, 0.0, 0.0]
            p[axis] = sign * h
            p[u] = su * h
            p[v] = sv * h
            corners.append(p)
        n = [0.0, 0.0, 0.0]
        n[axis] = -float(sign)
        faces.append((corners, n))
    verts, norms = [], []
    for corners, n in faces:
        for tri in ((0, 1, 2), (0, 2, 3)):
            for t in tri:
                verts.append(corners[t])
                norms.append(n)
    return np.asarray(verts, dtype=np.float32), np.asarray(norms, dtype=np.float32)


def c4_camera():
    """C4 camera: centre (2.2,0,0) -> target 0, up (0,0,1), OpenGL convention (no up flip)."""
    return dict(center=(2.2, 0.0, 0.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), flip_up=False)


# ------------------------------------------------------------------ C5: DEM packing time series
def hcp_spheres_chunked(n_target: int, phi: float = 0.55, jitter: float = 0.1, seed: int = 0x5EED0005,
                        chunk_i: int = 16):
    """Same lattice, order and jitter as hcp_spheres, generated in slabs of `chunk_i` x-layers so that
    100M spheres need ~2 GB of host memory instead of ~15 GB.  Returns (centers f32[n,3], radii f32[n])."""
    r = (phi * 3.0 / (4.0 * np.pi * n_target)) ** (1.0 / 3.0)
    a = 2.0 * r * (0.74048 / phi) ** (1.0 / 3.0)
    nx = int(np.floor((1.0 - 2 * r) / a)) + 1
    ny = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(3) / 2))) + 1
    nz = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(6) / 3))) + 1
    out = np.empty((n_target, 3), dtype=np.float32)
    n = 0
    lim = 1 - 2 * r
    for i0 in range(0, nx, chunk_i):
        if n >= n_target:
            break
        i, j, k = np.meshgrid(np.arange(i0, min(nx, i0 + chunk_i)), np.arange(ny), np.arange(nz), indexing="ij")
        i, j, k = i.ravel(), j.ravel(), k.ravel()
        x = (i + 0.5 * ((j + k) % 2)) * a
        y = (j + ((k % 2) / 3.0)) * (a * np.sqrt(3) / 2)
        z = k * (a * np.sqrt(6) / 3)
        keep = (x <= lim) & (y <= lim) & (z <= lim)
        c = np.stack([x[keep], y[keep], z[keep]], 1)[: n_target - n]
        m = len(c)
        idx = np.arange(n, n + m, dtype=np.uint64) * np.uint64(3)
        jit = np.stack([uniform01(seed, idx + np.uint64(q)) for q in range(3)], 1)
        out[n:n + m] = (c + r - 0.5 + (jit * 2.0 - 1.0) * (jitter * r)).astype(np.float32)
        n += m
    return np.ascontiguousarray(out[:n]), np.full(n, r, dtype=np.float32)


def hcp_spheres_device(torch, n_target: int, device, phi: float = 0.55, jitter: float = 0.1, seed: int = 0x5EED0005,
                       chunk_i: int = 32):
    """The lattice of hcp_spheres generated ON THE DEVICE with torch (C5 is generated on device):
    same lattice, order and truncation; the +-jitter*r displacement and the fall velocities come from torch's
    seeded generator instead of SplitMix64 (nothing measured on this scene depends on the jitter values).
    Returns (centers f32[n,3], radii f32[n], velocities f32[n,3]) as device tensors."""
    r = (phi * 3.0 / (4.0 * np.pi * n_target)) ** (1.0 / 3.0)
    a = 2.0 * r * (0.74048 / phi) ** (1.0 / 3.0)
    nx = int(np.floor((1.0 - 2 * r) / a)) + 1
    ny = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(3) / 2))) + 1
    nz = int(np.floor((1.0 - 2 * r) / (a * np.sqrt(6) / 3))) + 1
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    out = torch.empty((n_target, 3), dtype=torch.float32, device=device)
    lim = 1 - 2 * r
    n = 0
    jj = torch.arange(ny, device=device, dtype=torch.float64)
    kk = torch.arange(nz, device=device, dtype=torch.float64)
    for i0 in range(0, nx, chunk_i):
        if n >= n_target:
            break
        ii = torch.arange(i0, min(nx, i0 + chunk_i), device=device, dtype=torch.float64)
        i, j, k = torch.meshgrid(ii, jj, kk, indexing="ij")
        i, j, k = i.reshape(-1), j.reshape(-1), k.reshape(-1)
        x = (i + 0.5 * torch.remainder(j + k, 2)) * a
        y = (j + torch.remainder(k, 2) / 3.0) * (a * np.sqrt(3) / 2)
        z = k * (a * np.sqrt(6) / 3)
        keep = (x <= lim) & (y <= lim) & (z <= lim)
        c = torch.stack([x[keep], y[keep], z[keep]], 1)[: n_target - n]
        m = len(c)
        jit = torch.rand((m, 3), generator=gen, device=device, dtype=torch.float32).to(torch.float64)
        out[n:n + m] = (c + r - 0.5 + (jit * 2.0 - 1.0) * (jitter * r)).to(torch.float32)
        n += m
        del i, j, k, x, y, z, keep, c, jit
    centers = out[:n].contiguous()
    radii = torch.full((n,), float(np.float32(r)), dtype=torch.float32, device=device)
    vel = torch.zeros((n, 3), dtype=torch.float32, device=device)
    vel[:, 2] = -2.0 * (1.0 + 0.1 * torch.rand((n,), generator=gen, device=device, dtype=torch.float32))
    return centers, radii, vel


def dem_velocities(n: int, seed: int = 0x5EED00D5, chunk: int = 1 << 22):
    """C5 analytic motion z(t) = z0 - 2 t (1 + 0.1 hash(id)) as a per-particle velocity (0, 0, vz):
    the frame update is then the reference's Mesh-mode translation (velocity * duration / frames * frame)."""
    v = np.zeros((n, 3), dtype=np.float32)
    for start in range(0, n, chunk):
        stop = min(n, start + chunk)
        h = uniform01(seed, np.arange(start, stop, dtype=np.uint64))
        v[start:stop, 2] = np.float32(-2.0) * (np.float32(1.0) + np.float32(0.1) * h)
    return v


def c5_camera(frame: int, dt: float = 0.01):
    """Camera of C5: the C4 view, following the mean fall of the packing (-2.1 t along z)."""
    dz = -2.1 * dt * frame
    return dict(center=[2.2, 0.0, dz], target=[0.0, 0.0, dz], up=[0.0, 0.0, 1.0], flip_up=False)

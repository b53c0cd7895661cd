"""Synthetic inputs shared by bench.py and the tests (SURVEY.md 8d "Synthetic inputs").

phantom3d(Nx, N): ellipsoid Shepp-Logan-style volume (fixed table) plus 32 random spheres
(numpy default_rng(1234)); values in [0, 1], background 0, layout [x][y][z] float32.
"""
import numpy as np

# centre (x,y,z) in [-1,1], semi-axes (a,b,c), rotation about x [deg], amplitude
_ELLIPSOIDS = [
    ((0.00, 0.00, 0.00), (0.90, 0.69, 0.92), 0.0, 0.50),
    ((0.00, -0.02, 0.00), (0.85, 0.62, 0.87), 0.0, -0.25),
    ((0.10, 0.22, 0.00), (0.40, 0.11, 0.31), -18.0, 0.20),
    ((-0.10, -0.22, 0.00), (0.45, 0.16, 0.41), 18.0, 0.20),
    ((0.20, 0.00, 0.35), (0.25, 0.21, 0.25), 0.0, 0.30),
    ((-0.30, 0.00, 0.10), (0.12, 0.05, 0.05), 0.0, 0.35),
    ((0.35, 0.00, -0.10), (0.10, 0.05, 0.05), 0.0, 0.35),
    ((-0.40, -0.08, -0.60), (0.08, 0.05, 0.02), 0.0, 0.40),
    ((0.45, 0.00, -0.60), (0.06, 0.02, 0.02), 0.0, 0.40),
    ((0.00, 0.06, -0.60), (0.06, 0.02, 0.04), 0.0, 0.40),
]


def phantom3d(nx, n, seed=1234, nspheres=32, x0=0, count=None):
    """Slices [x0, x0+count) of the nx-slice phantom (default: all).  The scaling is a fixed
    constant (not the global maximum) so a slab evaluated on one rank equals the same slices of the
    full volume."""
    count = nx - x0 if count is None else count
    xs = (np.arange(x0, x0 + count, dtype=np.float32) + 0.5) / nx * 2 - 1
    ys = (np.arange(n, dtype=np.float32) + 0.5) / n * 2 - 1
    vol = np.zeros((count, n, n), np.float32)
    Y, Z = np.meshgrid(ys, ys, indexing="ij")
    rng = np.random.default_rng(seed)
    spheres = [(rng.uniform(-0.7, 0.7, 3), rng.uniform(0.02, 0.06), rng.uniform(0.2, 1.0)) for _ in range(nspheres)]
    for ix, x in enumerate(xs):
        sl = vol[ix]
        for (cx, cy, cz), (a, b, c), rot, amp in _ELLIPSOIDS:
            rx = ((x - cx) / a) ** 2
            if rx > 1:
                continue
            t = np.deg2rad(rot)
            yr = (Y - cy) * np.cos(t) + (Z - cz) * np.sin(t)
            zr = -(Y - cy) * np.sin(t) + (Z - cz) * np.cos(t)
            sl += np.float32(amp) * ((rx + (yr / b) ** 2 + (zr / c) ** 2) <= 1.0)
        for c3, r, amp in spheres:
            dx2 = (x - c3[0]) ** 2
            if dx2 > r * r:
                continue
            sl += np.float32(amp) * ((dx2 + (Y - c3[1]) ** 2 + (Z - c3[2]) ** 2) <= r * r)
    vol *= np.float32(1.0 / 1.75)
    np.clip(vol, 0.0, 1.0, out=vol)
    return vol


def tilt_angles_deg(na):
    """np.linspace(-90, 90, Na, endpoint=False)."""
    return np.linspace(-90.0, 90.0, na, endpoint=False)


def slab_bounds(nslice, nranks, rank):
    """Contiguous x-slab of rank r; remainder spread over the first ranks (same rule as the
    reference's MPI variant, tomofusion/gpu/utils/mpi_astra_ctvlib.cpp:53-58)."""
    base, rem = divmod(int(nslice), int(nranks))
    first = rank * base + min(rank, rem)
    count = base + (1 if rank < rem else 0)
    return first, count


# modified Shepp-Logan head phantom (Toft 1996): amplitude, semi-axes a, b, centre x0, y0, rotation [deg]
_SHEPP_LOGAN_2D = [
    (1.00, 0.6900, 0.9200, 0.00, 0.0000, 0.0),
    (-0.80, 0.6624, 0.8740, 0.00, -0.0184, 0.0),
    (-0.20, 0.1100, 0.3100, 0.22, 0.0000, -18.0),
    (-0.20, 0.1600, 0.4100, -0.22, 0.0000, 18.0),
    (0.10, 0.2100, 0.2500, 0.00, 0.3500, 0.0),
    (0.10, 0.0460, 0.0460, 0.00, 0.1000, 0.0),
    (0.10, 0.0460, 0.0460, 0.00, -0.1000, 0.0),
    (0.10, 0.0460, 0.0230, -0.08, -0.6050, 0.0),
    (0.10, 0.0230, 0.0230, 0.00, -0.6060, 0.0),
    (0.10, 0.0230, 0.0460, 0.06, -0.6050, 0.0),
]


def shepp_logan_2d(n):
    """n x n modified Shepp-Logan phantom, float32, values in [0, 1] (BASELINE.json configs[0])."""
    c = (np.arange(n, dtype=np.float64) + 0.5) / n * 2 - 1
    X, Y = np.meshgrid(c, -c)                 # x to the right, y up
    img = np.zeros((n, n))
    for amp, a, b, x0, y0, rot in _SHEPP_LOGAN_2D:
        t = np.deg2rad(rot)
        xr = (X - x0) * np.cos(t) + (Y - y0) * np.sin(t)
        yr = -(X - x0) * np.sin(t) + (Y - y0) * np.cos(t)
        img += amp * ((xr / a) ** 2 + (yr / b) ** 2 <= 1.0)
    return np.clip(img, 0.0, 1.0).astype(np.float32)

"""Seeded synthetic inputs of NeRF-synthetic shape (no datasets are available offline).

Everything here is plain numpy so that the same bytes feed the CUDA path, the CPU oracle
and the committed golden fixtures.  Shapes follow the reference's driver:
poses are 4x4 row-major camera-to-world matrices (loader/data_loader.cpp:66-84, used as
`look_at` at main.cu:470), the grid is the dense res^3 cell list of main.cu:154-174 with
primitive index (x*G+y)*G+z, and an "occupied" cell is one that would be an AABB
primitive in the reference's acceleration structure (main.cu:394-399).
"""
import math

import numpy as np

LEGO_CAMERA_ANGLE_X = 0.6911112070083618  # transforms_train.json of the lego scene
CAMERA_RADIUS = 4.031128874  # NeRF-synthetic camera distance


def focal_length(camera_angle_x=LEGO_CAMERA_ANGLE_X):
    """Focal length in the raygen's normalised units: d = R.(u, v, -focal)
    (optixPrograms.cu:62-64).  The reference's own `1/tan(0.5*focal_px)` (main.cu:365) is a
    slip; focal_length is treated as an input scalar."""
    return float(np.float32(1.0 / math.tan(0.5 * camera_angle_x)))


def pose_look_at_origin(theta, phi, radius=CAMERA_RADIUS):
    """Camera-to-world 4x4 (row-major, float32) at spherical position (theta, phi), looking
    at the origin, NeRF/OpenGL convention (camera looks down -z, +y up)."""
    c = np.array([radius * math.cos(phi) * math.cos(theta),
                  radius * math.cos(phi) * math.sin(theta),
                  radius * math.sin(phi)], np.float64)
    back = c / np.linalg.norm(c)
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(up, back)
    right /= np.linalg.norm(right)
    true_up = np.cross(back, right)
    m = np.eye(4)
    m[:3, 0] = right
    m[:3, 1] = true_up
    m[:3, 2] = back
    m[:3, 3] = c
    return m.astype(np.float32)


def poses(n, seed=0, radius=CAMERA_RADIUS):
    """n seeded poses on the upper hemisphere (NeRF-synthetic train split shape)."""
    rng = np.random.default_rng(seed)
    th = rng.uniform(0.0, 2.0 * math.pi, n)
    ph = rng.uniform(math.radians(5.0), math.radians(80.0), n)
    return np.stack([pose_look_at_origin(t, p, radius) for t, p in zip(th, ph)])


def occupancy_bernoulli(G, p, seed=0):
    rng = np.random.default_rng(seed)
    return (rng.random(G ** 3) < p).astype(np.uint8)


def occupancy_lego(G, seed=0):
    """'Lego-shaped' blob: a base plate, a body, a raised cab and studs inside
    [-0.6,0.6]^3 — a few axis-aligned boxes, roughly 5-10 % of the cells."""
    rng = np.random.default_rng(seed)
    occ = np.zeros((G, G, G), bool)
    ax = (np.arange(G) + 0.5) * (2.0 / G) - 1.0
    X, Y, Z = np.meshgrid(ax, ax, ax, indexing="ij")

    def box(lo, hi):
        return ((X >= lo[0]) & (X < hi[0]) & (Y >= lo[1]) & (Y < hi[1]) & (Z >= lo[2]) & (Z < hi[2]))

    occ |= box((-0.6, -0.35, -0.45), (0.6, 0.35, -0.30))      # base plate
    occ |= box((-0.45, -0.28, -0.30), (0.30, 0.28, 0.05))     # body
    occ |= box((-0.10, -0.22, 0.05), (0.28, 0.22, 0.35))      # cab
    occ |= box((0.30, -0.12, -0.30), (0.58, 0.12, -0.12))     # blade arm
    for _ in range(24):                                       # studs / greebles
        c = rng.uniform([-0.55, -0.3, -0.3], [0.55, 0.3, 0.4])
        h = rng.uniform(0.02, 0.06, 3)
        occ |= box(c - h, c + h)
    return occ.reshape(-1).astype(np.uint8)


def pack_bits(occ):
    """uint8 [G^3] -> uint32 words, bit (p & 31) of word (p >> 5) = cell p."""
    occ = np.ascontiguousarray(occ, np.uint8).reshape(-1)
    n = occ.size
    pad = (-n) % 32
    if pad:
        occ = np.concatenate([occ, np.zeros(pad, np.uint8)])
    b = np.packbits(occ.reshape(-1, 32), axis=1, bitorder="little")
    return np.ascontiguousarray(b).view(np.uint32).reshape(-1)


def cell_radiance(G, seed=0):
    """Config-1 fixed per-cell (r,g,b,sigma): RGB ~ U(0,1), sigma ~ U(0,20)."""
    rng = np.random.default_rng(seed)
    out = np.empty((G ** 3, 4), np.float32)
    out[:, :3] = rng.random((G ** 3, 3), dtype=np.float32)
    out[:, 3] = rng.random(G ** 3, dtype=np.float32) * 20.0
    return out


# --------------------------------------------------------------------------- network
def hashgrid_levels(n_levels=16, n_features=2, log2_hashmap_size=19, base_resolution=16,
                    per_level_scale=None, max_resolution=2048):
    """Per-level geometry of the multiresolution hash encoding (instant-ngp):
    scale_l = b^l * N_min - 1, res_l = ceil(scale_l) + 1, size_l = min(roundup8(res^3), T).
    Returns dict of arrays; offsets are in table ENTRIES (each entry = n_features halves)."""
    if per_level_scale is None:
        per_level_scale = math.exp(math.log(max_resolution / base_resolution) / (n_levels - 1))
    b = float(np.float32(per_level_scale))
    scale = np.empty(n_levels, np.float32)
    res = np.empty(n_levels, np.uint32)
    size = np.empty(n_levels, np.uint32)
    off = np.empty(n_levels + 1, np.int64)
    acc = 0
    for l in range(n_levels):
        s = np.float32(2.0 ** (l * math.log2(b)) * base_resolution - 1.0)
        r = int(math.ceil(float(s))) + 1
        dense = (r ** 3 + 7) // 8 * 8
        sz = min(dense, 1 << log2_hashmap_size)
        scale[l], res[l], size[l], off[l] = s, r, sz, acc
        acc += sz
    off[n_levels] = acc
    return dict(n_levels=n_levels, n_features=n_features, log2_hashmap_size=log2_hashmap_size,
                base_resolution=base_resolution, per_level_scale=b, scale=scale, res=res, size=size,
                offset=off, n_entries=acc)


def hashgrid_table(levels, seed=0):
    """Hash-table init U(-1e-4, 1e-4) in fp16 (tiny-cuda-nn's grid initialisation)."""
    rng = np.random.default_rng(seed)
    n = levels["n_entries"] * levels["n_features"]
    return ((rng.random(n, dtype=np.float32) * 2.0 - 1.0) * 1e-4).astype(np.float16)


def mlp_weights(in_dim=32, width=64, n_hidden=2, out_dim=16, seed=0):
    """Xavier-uniform fp16 weights, bias-free, row-major [out][in] per layer, layers
    concatenated: [width][in_dim], (n_hidden-1) x [width][width], [out_dim][width]."""
    rng = np.random.default_rng(seed)
    shapes = [(width, in_dim)] + [(width, width)] * (n_hidden - 1) + [(out_dim, width)]
    parts = []
    for (o, i) in shapes:
        lim = math.sqrt(6.0 / (i + o))
        parts.append(((rng.random((o, i), dtype=np.float32) * 2.0 - 1.0) * lim).astype(np.float16).reshape(-1))
    return np.concatenate(parts)

"""Deterministic synthetic inputs for the AMCL hot path.

Everything is generated from integers so that every host produces identical
maps, scans and particle sets: the reference ships no data for this path (its
tests download rosbags), so the named BASELINE.json shapes are synthesised.
"""
import numpy as np

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised splitmix64 finaliser on uint64 arrays."""
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x


def rooms_map(size, seed=1, room=100, door=(40, 60), clutter_mod=5000, unknown_ring=False):
    """rooms(S, seed): S x S occupancy states (-1 free, 0 unknown, +1 occupied).

    Returned as int8 [size_y, size_x] with x fastest, i.e. the flat order
    map_t::cells uses (MAP_INDEX = i + j*size_x, map.h:144).
    Walls every `room` cells and on the border; each interior wall segment has
    a door gap at offsets door[0]..door[1]; sparse clutter from a hash.
    """
    j, i = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    border = (i == 0) | (j == 0) | (i == size - 1) | (j == size - 1)
    wall_x = (i % room) == 0
    wall_y = (j % room) == 0
    in_door_along_y = ((j % room) >= door[0]) & ((j % room) < door[1])
    in_door_along_x = ((i % room) >= door[0]) & ((i % room) < door[1])
    # a vertical wall (i fixed) has its door where j is in the gap, and vice versa
    wall = (wall_x & ~in_door_along_y) | (wall_y & ~in_door_along_x)
    h = splitmix64(
        (np.uint64(seed) << np.uint64(40))
        ^ (i.astype(np.uint64) << np.uint64(20))
        ^ j.astype(np.uint64)
    )
    clutter = (h % np.uint64(clutter_mod)) == 0
    occ = np.full((size, size), -1, dtype=np.int8)
    occ[wall | clutter | border] = 1
    if unknown_ring:
        ring = (i == 1) | (j == 1) | (i == size - 2) | (j == size - 2)
        occ[ring & (occ != 1)] = 0
    return occ


def scan_bearings(n_beams, fov=1.5 * np.pi):
    """Bearings -fov/2 + i*fov/(B-1) (a 270 degree planar lidar)."""
    return -0.5 * fov + np.arange(n_beams, dtype=np.float64) * (fov / (n_beams - 1))


def room_centre_pose(size, scale, room=100, which=(0.5, 0.5), heading=0.3):
    """World pose at the centre of the room nearest the fractional position."""
    ci = (int(size * which[0]) // room) * room + room // 2
    cj = (int(size * which[1]) // room) * room + room // 2
    x = (ci - size // 2) * scale
    y = (cj - size // 2) * scale
    return np.array([x, y, heading], dtype=np.float64)


def lcg48_uniforms(seed, count):
    """The glibc drand48 stream after srand48(seed): X' = aX + c mod 2^48."""
    a, c, mask = 0x5DEECE66D, 0xB, (1 << 48) - 1
    x = ((seed & 0xFFFFFFFF) << 16) | 0x330E
    out = np.empty(count, dtype=np.float64)
    for k in range(count):
        x = (a * x + c) & mask
        out[k] = x / float(1 << 48)
    return out


def uniform_cloud(occ, scale, n, seed):
    """Global-localisation cloud: uniform over free cells, heading in [-pi, pi).

    A hash-driven stand-in for pf_init_model + uniformPoseGenerator
    (amcl_node.cpp:1050-1060) that is cheap at 16 M particles; parity tests
    use the real generator through the API instead.
    Returns float64 [n, 4] rows (x, y, theta, weight = 1/n).
    """
    size_y, size_x = occ.shape
    free_j, free_i = np.nonzero(occ == -1)
    k = np.arange(n, dtype=np.uint64)
    h1 = splitmix64((np.uint64(seed) << np.uint64(48)) ^ k)
    h2 = splitmix64(h1)
    pick = (h1 % np.uint64(free_i.size)).astype(np.int64)
    out = np.empty((n, 4), dtype=np.float64)
    out[:, 0] = (free_i[pick] - size_x // 2) * scale
    out[:, 1] = (free_j[pick] - size_y // 2) * scale
    out[:, 2] = (h2 >> np.uint64(11)).astype(np.float64) * (2.0 * np.pi / float(1 << 53)) - np.pi
    out[:, 3] = 1.0 / n
    return out


def gaussian_cloud(mean, std, n, seed):
    """Tracking cloud: independent normals around `mean` (x, y, theta)."""
    k = np.arange(n, dtype=np.uint64)
    out = np.empty((n, 4), dtype=np.float64)
    for d in range(3):
        h1 = splitmix64((np.uint64(seed) << np.uint64(48)) ^ (np.uint64(d + 1) << np.uint64(40)) ^ k)
        h2 = splitmix64(h1)
        u1 = ((h1 >> np.uint64(11)).astype(np.float64) + 1.0) / float(1 << 53)
        u2 = (h2 >> np.uint64(11)).astype(np.float64) / float(1 << 53)
        out[:, d] = mean[d] + std[d] * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    out[:, 3] = 1.0 / n
    return out

"""Synthetic cloud pairs of SURVEY §8d (configs 4 and 5) — host-side input generation only, no ICP arithmetic.

Reproducible from the recipe alone (no dependence on a numpy RNG version):
  model  m_i ~ U[-1,1)^3 from splitmix64(seed_m):  coordinate c of point i = 2*(u >> 11)*2^-53 - 1 with u = output 3*i + c
  scene  p_j = R m_i + t + sigma * g_i,  R = `angle_deg` about (1,2,3)/sqrt(14), t = (0.05, -0.03, 0.02),
         g_i = Box-Muller on splitmix64(seed_g): outputs 4*i .. 4*i+3 -> u1,u2,u3,u4 in (0,1];
         g = (r1 cos a1, r1 sin a1, r2 cos a2), r = sqrt(-2 ln u_odd), a = 2 pi u_even
  order  scene position j holds model point perm[j]; perm = stable argsort of splitmix64(seed_perm) outputs
Config 4: seeds (1000 + log2 N, 2000 + log2 N, 3000 + log2 N); config 5, pair k: (1000 + k, 2000 + k, 3000 + k).
"""
import numpy as np

_GOLDEN = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed, n):
    """First n outputs of splitmix64 started at `seed` (vectorised: output i has state seed + (i+1)*golden)."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + _GOLDEN * np.arange(1, n + 1, dtype=np.uint64)
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _unit(u):            # [0, 1)
    return (u >> np.uint64(11)).astype(np.float64) * 2.0 ** -53


def rotation(angle_deg, axis=(1.0, 2.0, 3.0)):
    ax = np.asarray(axis, float)
    ax = ax / np.linalg.norm(ax)
    a = np.deg2rad(angle_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    return np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)


def pair(n, seed_m, seed_g, seed_perm, angle_deg=5.0, noise=1e-3, shift=(0.05, -0.03, 0.02), shuffle=True):
    """-> (model (n,3), scene (n,3), perm) as float64 C-contiguous arrays."""
    m = (2.0 * _unit(splitmix64(seed_m, 3 * n)) - 1.0).reshape(n, 3)
    p = m @ rotation(angle_deg).T + np.asarray(shift, float)
    if noise:
        u = (_unit(splitmix64(seed_g, 4 * n)) + 2.0 ** -53).reshape(n, 4)      # (0, 1]
        r1, r2 = np.sqrt(-2.0 * np.log(u[:, 0])), np.sqrt(-2.0 * np.log(u[:, 2]))
        a1, a2 = 2.0 * np.pi * u[:, 1], 2.0 * np.pi * u[:, 3]
        p = p + noise * np.stack([r1 * np.cos(a1), r1 * np.sin(a1), r2 * np.cos(a2)], axis=1)
    perm = np.argsort(splitmix64(seed_perm, n), kind="stable") if shuffle else np.arange(n)
    return np.ascontiguousarray(m), np.ascontiguousarray(p[perm]), perm


def config4(n):
    """np = nm = n (a power of two), rigid 5 degrees + 1e-3 noise."""
    lg = int(round(np.log2(n)))
    m, p, _ = pair(n, 1000 + lg, 2000 + lg, 3000 + lg)
    return m, p


def config5_pair(k, n=1 << 16):
    """SURVEY §8d config 5, pair k: rotation angle 2..8 degrees varying with k."""
    m, p, _ = pair(n, 1000 + k, 2000 + k, 3000 + k, angle_deg=2.0 + 6.0 * ((k * 37) % 101) / 100.0)
    return m, p

"""Synthetic polydisperse liquid configurations.

N = n*n particles at number density rho in a periodic square box, float32 like HOOMD's GSD output:
jittered square lattice (keeps every pair further apart than (1 - 2*jitter) lattice spacings, so
forces stay in a sane dynamic range) and diameters drawn from P(sigma) ~ sigma^-3 on
[0.73, 1.62] -- the range of the reference's shipped fixture.
"""
import numpy as np

SEEDS = {"cfg2": 20240602, "cfg3": 20240603, "cfg4": 20240604, "cfg5": 20240605}


def synth_frame(n, seed, rho=1.0, jitter=0.3, velocities=False):
    rng = np.random.default_rng(seed)
    L = n / np.sqrt(rho)
    a = L / n
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    pos = np.zeros((n * n, 3), np.float32)
    pos[:, 0] = ((ii.ravel() + 0.5) * a - L / 2 + rng.uniform(-jitter, jitter, n * n) * a).astype(np.float32)
    pos[:, 1] = ((jj.ravel() + 0.5) * a - L / 2 + rng.uniform(-jitter, jitter, n * n) * a).astype(np.float32)
    lo, hi = 0.73, 1.62
    u = rng.uniform(0.0, 1.0, n * n)
    diam = ((lo**-2 - u * (lo**-2 - hi**-2)) ** -0.5).astype(np.float32)
    out = dict(position=pos, diameter=diam, box=np.array([L, L, 1.0, 0, 0, 0], np.float32), N=n * n)
    if velocities:
        v = np.zeros((n * n, 3), np.float32)
        v[:, :2] = rng.standard_normal((n * n, 2)).astype(np.float32)
        out["velocity"] = v
    return out


def linspace_axis(L, nmax):
    """The reference's grid axis (irvingkirkwood/__init__.py:45-46): both ends included."""
    return np.linspace(-L / 2.0, +L / 2.0, nmax)

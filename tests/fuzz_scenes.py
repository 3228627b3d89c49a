"""Adversarial tiny scenes and rays for the intersectors (test infrastructure).

Scenes: 1..16 primitives, planar quads (parallelograms and trapezoids) and triangles including slivers, at three sizes and optionally far
from the origin.  Rays: aimed at interior points, at EDGES and at VERTICES of the primitives, started on other primitives with the
integrator's 1e-4 offset, axis-parallel, and grazing (almost inside a primitive's plane).  These are the cases where a conservative cull
(the flat list's rectangle filter, a BVH box test) can disagree with the exact triangle test that alone decides a hit.
"""
import numpy as np


def tiny_scene(rng, prim_dtype):
    n = int(rng.integers(1, 17))
    offset = (rng.choice([0.0, 0.0, 10.0, 1000.0]) * rng.normal(size=3)).astype(np.float32)
    size = float(rng.choice([0.01, 1.0, 1.0, 50.0]))
    prims = np.zeros(n, prim_dtype)
    for k in range(n):
        c = rng.normal(size=3) * size
        a = rng.normal(size=3) * size * rng.choice([1.0, 1.0, 0.02])
        b = rng.normal(size=3) * size
        if rng.random() < 0.5:
            prims[k]["type"] = 1
            s1, s2 = rng.uniform(0.3, 1.5, 2)
            v = [c, c + a, c + a * s1 + b, c + b * s2] if rng.random() < 0.5 else [c, c + a, c + a + b, c + b]
            prims[k]["v"] = np.array(v, np.float32) + offset
        else:
            prims[k]["type"] = 0
            prims[k]["v"][:3] = np.array([c, c + a, c + b], np.float32) + offset
        nn = np.cross(a, b)
        prims[k]["normal"] = nn / max(np.linalg.norm(nn), 1e-30)
        prims[k]["bsdf"] = 0.5
    return prims, offset, size


def adversarial_rays(rng, prims, offset, size, m):
    n = len(prims)
    V = prims["v"].astype(np.float64)
    k = rng.integers(0, n, m)
    w = rng.dirichlet([1, 1, 1], m)
    edge = rng.random(m) < 0.3
    w[edge, 2] = 0
    w[edge] /= w[edge].sum(1, keepdims=True)
    vert = rng.random(m) < 0.1
    w[vert] = np.eye(3)[rng.integers(0, 3, int(vert.sum()))]
    tgt = (V[k, :3] * w[:, :, None]).sum(1)
    o = offset + rng.normal(size=(m, 3)) * size * 3
    on = rng.random(m) < 0.4
    k2 = rng.integers(0, n, m)
    w2 = rng.dirichlet([1, 1, 1], m)
    p2 = (V[k2, :3] * w2[:, :, None]).sum(1)
    o[on] = p2[on] + prims["normal"][k2[on]] * 1e-4 * rng.choice([-1, 1], int(on.sum()))[:, None]
    d = tgt - o
    rnd = rng.random(m) < 0.2
    d[rnd] = rng.normal(size=(int(rnd.sum()), 3))
    ax = rng.random(m) < 0.05
    d[ax] = np.eye(3)[rng.integers(0, 3, int(ax.sum()))] * rng.choice([-1, 1], int(ax.sum()))[:, None]
    graze = rng.random(m) < 0.1
    e = V[k, 1] - V[k, 0]
    d[graze] = e[graze] + prims["normal"][k[graze]] * np.linalg.norm(e[graze], axis=1, keepdims=True) * rng.normal(size=(int(graze.sum()), 1)) * 1e-4
    d[np.linalg.norm(d, axis=1) < 1e-20] = [1, 0, 0]
    return np.concatenate([o, d], 1).astype(np.float32)


def edge_margin(prim, ray):
    """Double precision: how far inside the primitive the ray's plane crossing lies, as the smallest barycentric coordinate of the best
    of its triangles (negative = outside; about 0 = on an edge or a vertex), and |det| relative to |e1||e2| (about 0 = grazing)."""
    o = ray[:3].astype(np.float64)
    d = ray[3:].astype(np.float64)
    d = d / np.linalg.norm(d)
    V = prim["v"].astype(np.float64)
    best, graze = -np.inf, 1.0
    for (a0, a1, a2) in [(0, 1, 2)] + ([(0, 2, 3)] if prim["type"] == 1 else []):
        v0, e1, e2 = V[a0], V[a1] - V[a0], V[a2] - V[a0]
        pv = np.cross(d, e2)
        det = e1 @ pv
        rel = abs(det) / max(np.linalg.norm(e1) * np.linalg.norm(e2), 1e-300)
        graze = min(graze, rel)
        if rel < 1e-12:
            continue
        tv = o - v0
        u = (tv @ pv) / det
        v = (d @ np.cross(tv, e1)) / det
        best = max(best, min(u, v, 1.0 - u - v))
    return best, graze

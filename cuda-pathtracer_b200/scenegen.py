"""Synthetic scenes of the shapes BASELINE.json names, generated on the box with numpy.

  c3_scene(level)   two radially displaced icospheres (20 * 4^level triangles each; level 9 -> 10 485 760 in total)
                    inside the Cornell enclosure of scenes/cbox_quads.obj; one diffuse, one mirror.
  c4_prims(prims)   Cornell box with the tall box mirror (Kr 0.9) and the short box emissive (Ke 2,2,2).

Everything is float32 and deterministic (hash-based value noise, fixed seeds).  Triangle soups are
produced by subdividing triangles directly: a shared edge midpoint is computed from the same two floats
in both triangles ((a + b) is commutative in IEEE arithmetic), so the meshes are watertight.
"""
import os

import numpy as np

_PHI = (1.0 + 5.0 ** 0.5) / 2.0


def icosahedron():
    v = np.array([[-1, _PHI, 0], [1, _PHI, 0], [-1, -_PHI, 0], [1, -_PHI, 0], [0, -1, _PHI], [0, 1, _PHI], [0, -1, -_PHI], [0, 1, -_PHI],
                  [_PHI, 0, -1], [_PHI, 0, 1], [-_PHI, 0, -1], [-_PHI, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    return v[f].astype(np.float32)          # (20, 3, 3)


def _normalize(p):
    n = np.sqrt((p.astype(np.float32) ** 2).sum(-1, keepdims=True, dtype=np.float32))
    return (p / n).astype(np.float32)


def subdivide_sphere(tris, levels):
    """Each unit-sphere triangle -> 4, midpoints pushed back to the sphere."""
    for _ in range(levels):
        a, b, c = tris[:, 0], tris[:, 1], tris[:, 2]
        ab, bc, ca = _normalize(a + b), _normalize(b + c), _normalize(c + a)
        tris = np.concatenate([np.stack([a, ab, ca], 1), np.stack([ab, b, bc], 1), np.stack([ca, bc, c], 1), np.stack([ab, bc, ca], 1)], 0)
    return tris


def _hash3(ix, iy, iz, seed):
    h = (ix.astype(np.uint32) * np.uint32(0x8DA6B343)) ^ (iy.astype(np.uint32) * np.uint32(0xD8163841)) ^ \
        (iz.astype(np.uint32) * np.uint32(0xCB1AB31F)) ^ np.uint32(seed)
    h ^= h >> np.uint32(16); h *= np.uint32(0x7FEB352D); h ^= h >> np.uint32(15); h *= np.uint32(0x846CA68B); h ^= h >> np.uint32(16)
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def value_noise(p, seed):
    pf = np.floor(p)
    f = (p - pf).astype(np.float32)
    i = pf.astype(np.int64)
    w = f * f * (np.float32(3.0) - np.float32(2.0) * f)
    out = np.zeros(len(p), np.float32)
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                wx = w[:, 0] if dx else 1 - w[:, 0]
                wy = w[:, 1] if dy else 1 - w[:, 1]
                wz = w[:, 2] if dz else 1 - w[:, 2]
                out += wx * wy * wz * _hash3(i[:, 0] + dx, i[:, 1] + dy, i[:, 2] + dz, seed)
    return out


def fbm4(p, seed):
    s, amp, tot = np.zeros(len(p), np.float32), np.float32(0.5), np.float32(0.0)
    q = p.astype(np.float32).copy()
    for o in range(4):
        s += amp * value_noise(q, seed + o)
        tot += amp
        q *= np.float32(2.0); amp *= np.float32(0.5)
    return s / tot


def icosphere_indexed(level):
    """Unit icosphere as (vertices (V,3) f32, faces (F,3) i64); V = 10*4^level + 2, F = 20*4^level."""
    tri = icosahedron().astype(np.float32)
    verts, inv = np.unique(tri.reshape(-1, 3), axis=0, return_inverse=True)
    faces = inv.reshape(-1, 3).astype(np.int64)
    verts = verts.astype(np.float32)
    for _ in range(level):
        n = len(verts)
        e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], 0)
        key = np.minimum(e[:, 0], e[:, 1]) * n + np.maximum(e[:, 0], e[:, 1])
        uk, inv = np.unique(key, return_inverse=True)
        mid = _normalize(verts[uk // n] + verts[uk % n])
        verts = np.concatenate([verts, mid], 0)
        f = len(faces)
        ab, bc, ca = n + inv[:f], n + inv[f:2 * f], n + inv[2 * f:]
        a, b, c = faces[:, 0], faces[:, 1], faces[:, 2]
        faces = np.concatenate([np.stack([a, ab, ca], 1), np.stack([ab, b, bc], 1), np.stack([ca, bc, c], 1), np.stack([ab, bc, ca], 1)], 0)
    return verts, faces


def displacement(unit_verts, seed=1234, amplitude=0.05, frequency=8.0, chunk=1 << 20):
    """Radial factor 1 + amplitude * (2 fbm4(frequency p) - 1) per unit-sphere vertex."""
    out = np.empty(len(unit_verts), np.float32)
    for s in range(0, len(unit_verts), chunk):                             # bounded temporaries
        p = unit_verts[s:s + chunk]
        out[s:s + chunk] = np.float32(1.0) + np.float32(amplitude) * (np.float32(2.0) * fbm4(p * np.float32(frequency), seed) - np.float32(1.0))
    return out


def displaced_icosphere(level, radius, centre, seed=1234, amplitude=0.05, frequency=8.0, _cache={}):
    """Triangle soup (20*4^level, 3, 3) f32 of a radially displaced icosphere; shared vertices are bit-identical."""
    k = (level, seed, amplitude, frequency)
    if k not in _cache:
        _cache.clear()
        v, f = icosphere_indexed(level)
        _cache[k] = (v, f, displacement(v, seed, amplitude, frequency))
    v, f, d = _cache[k]
    pos = v * (np.float32(radius) * d)[:, None] + np.asarray(centre, np.float32)
    return pos[f]


CORNELL_MATS = [
    {"bsdf": (0.8, 0.659341, 0.43956), "Le": (0, 0, 0), "type": 0},       # 0 Khaki   (scenes/cbox.mtl)
    {"bsdf": (0.445, 0.0, 0.0), "Le": (0, 0, 0), "type": 0},               # 1 BloodyRed
    {"bsdf": (0.0, 0.32, 0.0), "Le": (0, 0, 0), "type": 0},                # 2 DarkGreen
    {"bsdf": (0.0, 0.0, 0.0), "Le": (25, 25, 25), "type": 0},              # 3 Light
    {"bsdf": (0.8, 0.66, 0.44), "Le": (0, 0, 0), "type": 0},               # 4 diffuse blob
    {"bsdf": (0.9, 0.9, 0.9), "Le": (0, 0, 0), "type": 1},                 # 5 mirror blob (extension)
]


def enclosure_from_prims(prims, mats=CORNELL_MATS, n_enclosure=6):
    """Walls + light of scenes/cbox_quads.obj as a triangle soup: its first six faces (light, floor, ceiling,
    back wall, red wall, green wall; the two boxes follow, scenes/cbox_quads.obj:75-112).
    `prims` = parsed primitives (binding.parse_scene_file)."""
    tris, mid = [], []
    for p in prims[:n_enclosure]:
        pts = np.asarray(p["v"][:4 if p["type"] == 1 else 3], np.float32)
        emissive = float(np.max(p["Le"])) > 0
        m = 3 if emissive else int(np.argmin([np.abs(np.asarray(mm["bsdf"], np.float32) - p["bsdf"]).sum() for mm in mats[:3]]))
        idx = [(0, 1, 2), (0, 2, 3)] if p["type"] == 1 else [(0, 1, 2)]
        for t in idx:
            tris.append(pts[list(t)]); mid.append(m)
    return np.asarray(tris, np.float32), np.asarray(mid, np.uint32)


def c3_scene(prims, level=9):
    """returns (triangles (n,3,3) f32, material ids (n,) u32, material table)."""
    enc, enc_m = enclosure_from_prims(prims)
    s1 = displaced_icosphere(level, 1.0, (-1.2, 1.2, -3.0), seed=1234)
    s2 = displaced_icosphere(level, 0.8, (1.0, 0.9, -2.0), seed=1234)
    tris = np.concatenate([enc, s1, s2], 0)
    mat = np.concatenate([enc_m, np.full(len(s1), 4, np.uint32), np.full(len(s2), 5, np.uint32)])
    return tris, mat, CORNELL_MATS


def c4_prims(prims):
    """the Cornell box of scenes/cbox_quads.obj with the tall box a mirror (Kr 0.9, the library's mirror
    extension) and the short box emissive (Ke 2, 2, 2).  `prims` = binding.parse_scene_file(cbox_quads.obj): faces 6-10 are
    the short box, 11-15 the tall box (scenes/cbox_quads.obj:97-112).  Returns a modified copy."""
    out = prims.copy()
    assert len(out) == 16
    out["Le"][6:11] = (2.0, 2.0, 2.0)
    out["bsdf"][11:16] = (0.9, 0.9, 0.9)
    out["material_type"][11:16] = 1
    return out


def write_c5_pbrt(path, instances=4883, level=5, n_lights=64, seed=7):
    """a pbrt-format scene written on the box -- one icosphere mesh (20 * 4^level triangles) declared
    once with ObjectBegin/ObjectEnd and instanced `instances` times (ObjectInstance under Translate / Scale / Rotate) on a
    jittered grid inside a large room, plus `n_lights` emissive quads under the ceiling.  The defaults give
    4883 * 20480 = 100 003 840 instanced triangles; the loader flattens the instances (as the reference's loadPBRT does).
    Returns the triangle count."""
    rng = np.random.default_rng(seed)
    v, f = icosphere_indexed(level)
    side = int(np.ceil(instances ** (1.0 / 3.0)))
    room = 4.0 * side
    with open(path, "w") as fp:
        fp.write("# generated by cuda-pathtracer_b200/scenegen.py: write_c5_pbrt\n")
        fp.write(f"LookAt {room / 2} {room / 2} {room * 1.9}  {room / 2} {room / 2} 0  0 1 0\n")
        fp.write('Camera "perspective" "float fov" [40]\nWorldBegin\n')
        fp.write('MakeNamedMaterial "grey" "string type" "matte" "rgb Kd" [0.7 0.7 0.7]\n')
        fp.write('MakeNamedMaterial "chrome" "string type" "mirror" "rgb Kr" [0.9 0.9 0.9]\n')
        idx_txt = " ".join(str(int(i)) for i in f.ravel())
        pts_txt = " ".join(repr(float(x)) for x in v.ravel())
        for name, mat in (("blob", "grey"), ("blob_m", "chrome")):        # an instance keeps the material of its definition
            fp.write(f'AttributeBegin\n  NamedMaterial "{mat}"\n  ObjectBegin "{name}"\n    Shape "trianglemesh" "integer indices" [{idx_txt}] "point P" [{pts_txt}]\n  ObjectEnd\nAttributeEnd\n')
        # room: floor, ceiling, back, left, right as two-triangle quads
        def quad(a, b, c, d):
            pts = " ".join(repr(float(x)) for p in (a, b, c, d) for x in p)
            fp.write(f'Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [{pts}]\n')
        R = room
        fp.write('AttributeBegin\n  NamedMaterial "grey"\n')
        quad((0, 0, 0), (R, 0, 0), (R, 0, R), (0, 0, R)); quad((0, R, 0), (0, R, R), (R, R, R), (R, R, 0))
        quad((0, 0, 0), (0, R, 0), (R, R, 0), (R, 0, 0)); quad((0, 0, 0), (0, 0, R), (0, R, R), (0, R, 0)); quad((R, 0, 0), (R, R, 0), (R, R, R), (R, 0, R))
        fp.write("AttributeEnd\n")
        g = int(np.ceil(np.sqrt(n_lights)))
        for k in range(n_lights):
            cx, cz = (k % g + 0.5) * R / g, (k // g + 0.5) * R / g
            s = 0.15 * R / g
            fp.write('AttributeBegin\n  AreaLightSource "diffuse" "rgb L" [12 12 12]\n  Material "matte" "rgb Kd" [0 0 0]\n')
            quad((cx - s, R - 0.01, cz - s), (cx + s, R - 0.01, cz - s), (cx + s, R - 0.01, cz + s), (cx - s, R - 0.01, cz + s))
            fp.write("AttributeEnd\n")
        n = 0
        for k in range(instances):
            i, j, l = k % side, (k // side) % side, k // (side * side)
            c = (np.array([i, j, l]) + 0.5 + rng.uniform(-0.15, 0.15, 3)) * 4.0
            sc = rng.uniform(0.9, 1.5)
            fp.write(f'AttributeBegin\n  Translate {c[0]:.6f} {c[1]:.6f} {c[2]:.6f}\n  Rotate {rng.uniform(0, 360):.4f} 0 1 0\n'
                     f'  Scale {sc:.6f} {sc:.6f} {sc:.6f}\n  ObjectInstance "{"blob_m" if k % 7 == 0 else "blob"}"\nAttributeEnd\n')
            n += len(f)
        fp.write("WorldEnd\n")
    return n + 10 + 2 * n_lights


def write_tree_obj(path, prims, level=5, seed=1234):
    """A mid-size scene the REFERENCE can load (parity tests): the enclosure of scenes/cbox_quads.obj (its first six faces, as
    quads) plus two displaced icospheres of 20 * 4^level triangles each, written as .obj + .mtl with unshared vertices, no `vn`
    (both loaders then store the geometric normal) and 9 significant digits (float32 round-trips).  Lambertian materials only.
    `prims` = binding.parse_scene_file(cbox_quads.obj).  Returns the primitive count."""
    mats = [("Khaki", CORNELL_MATS[0]), ("BloodyRed", CORNELL_MATS[1]), ("DarkGreen", CORNELL_MATS[2]), ("Light", CORNELL_MATS[3]),
            ("Blob", CORNELL_MATS[4]), ("Blob2", {"bsdf": (0.3, 0.4, 0.75), "Le": (0, 0, 0), "type": 0})]
    mtl = os.path.splitext(path)[0] + ".mtl"
    with open(mtl, "w") as fp:
        for name, m in mats:
            fp.write(f"newmtl {name}\nKd {m['bsdf'][0]:.9g} {m['bsdf'][1]:.9g} {m['bsdf'][2]:.9g}\nKe {m['Le'][0]:.9g} {m['Le'][1]:.9g} {m['Le'][2]:.9g}\n\n")
    n_v = n_p = 0
    with open(path, "w") as fp:
        fp.write(f"# generated by cuda-pathtracer_b200/scenegen.py: write_tree_obj(level={level}, seed={seed})\nmtllib {os.path.basename(mtl)}\n")
        for p in prims[:6]:
            k = 4 if p["type"] == 1 else 3
            emissive = float(np.max(p["Le"])) > 0
            m = 3 if emissive else int(np.argmin([np.abs(np.asarray(mm["bsdf"], np.float32) - p["bsdf"]).sum() for _, mm in mats[:3]]))
            fp.write(f"usemtl {mats[m][0]}\n")
            for v in p["v"][:k]:
                fp.write("v %.9g %.9g %.9g\n" % tuple(float(x) for x in v))
            fp.write("f " + " ".join(str(n_v + 1 + i) for i in range(k)) + "\n")
            n_v += k; n_p += 1
        for name, radius, centre in (("Blob", 1.0, (-1.2, 1.2, -3.0)), ("Blob2", 0.8, (1.0, 0.9, -2.0))):
            tris = displaced_icosphere(level, radius, centre, seed=seed)
            fp.write(f"usemtl {name}\n")
            lines = []
            for t in tris:
                for v in t:
                    lines.append("v %.9g %.9g %.9g\n" % (float(v[0]), float(v[1]), float(v[2])))
                lines.append(f"f {n_v + 1} {n_v + 2} {n_v + 3}\n")
                n_v += 3
            fp.write("".join(lines))
            n_p += len(tris)
    return n_p

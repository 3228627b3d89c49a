"""Writes the PLY fixtures of tests/pbrt/plymesh.pbrt (small, committed): ascii with per-vertex normals, binary little-endian with extra
properties (uv, colours, a uchar list count and ushort indices), binary big-endian with double coordinates, and a quad-faced file the
reference rejects ("only triangles are supported", ext/pbrtparser/pbrtParser/impl/semantic/Geometry.cpp:55-58).
    python tests/pbrt/make_ply_fixtures.py"""
import os, struct
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(3)

# 1. ascii, normals, 2 triangles + a comment and an obj_info line
with open(os.path.join(HERE, "tri_ascii.ply"), "w") as f:
    f.write("ply\nformat ascii 1.0\ncomment made by make_ply_fixtures.py\nobj_info nothing\nelement vertex 4\n"
            "property float x\nproperty float y\nproperty float z\nproperty float nx\nproperty float ny\nproperty float nz\n"
            "element face 2\nproperty list uchar int vertex_indices\nend_header\n")
    f.write("0 0 0 0 0 1\n1.5 0 0.25 0 0.6 0.8\n1.5 1.25 0 0 0 1\n0 1 -0.125 0.1 0.2 0.3\n")
    f.write("3 0 1 2\n3 0 2 3\n")

# 2. binary little endian: a 3 x 3 grid of vertices (8 triangles), uv + colour properties interleaved, ushort indices, int list count
n = 3
xs, ys = np.meshgrid(np.linspace(-1, 1, n), np.linspace(-1, 1, n))
P = np.stack([xs.ravel(), ys.ravel(), 0.3 * rng.standard_normal(n * n)], 1).astype(np.float32)
faces = []
for j in range(n - 1):
    for i in range(n - 1):
        a = j * n + i
        faces += [(a, a + 1, a + n + 1), (a, a + n + 1, a + n)]
with open(os.path.join(HERE, "grid_le.ply"), "wb") as f:
    f.write((f"ply\nformat binary_little_endian 1.0\nelement vertex {len(P)}\nproperty float x\nproperty float u\nproperty float y\nproperty uchar red\n"
             f"property float z\nproperty float v\nelement face {len(faces)}\nproperty uchar flags\nproperty list int ushort vertex_index\nend_header\n").encode())
    for k, p in enumerate(P):
        f.write(struct.pack("<f", p[0]) + struct.pack("<f", k / 9.0) + struct.pack("<f", p[1]) + struct.pack("<B", k * 20) + struct.pack("<f", p[2]) + struct.pack("<f", 1 - k / 9.0))
    for fc in faces:
        f.write(struct.pack("<B", 7) + struct.pack("<i", 3) + struct.pack("<HHH", *fc))

# 3. binary big endian: double coordinates, normals as float, uint indices, face element BEFORE... (no: rply needs vertex first only for our use) after
Pd = (rng.standard_normal((5, 3)) * 0.7 + np.array([0.3, 1.1, -2.0])).astype(np.float64)
Nd = rng.standard_normal((5, 3)).astype(np.float32)
fb = [(0, 1, 2), (2, 3, 4), (4, 0, 2)]
with open(os.path.join(HERE, "blob_be.ply"), "wb") as f:
    f.write((f"ply\nformat binary_big_endian 1.0\nelement vertex {len(Pd)}\nproperty double x\nproperty double y\nproperty double z\n"
             f"property float nx\nproperty float ny\nproperty float nz\nelement face {len(fb)}\nproperty list uchar uint vertex_indices\nend_header\n").encode())
    for p, nn in zip(Pd, Nd):
        f.write(struct.pack(">ddd", *p) + struct.pack(">fff", *nn))
    for fc in fb:
        f.write(struct.pack(">B", 3) + struct.pack(">III", *fc))

# 4. a quad face: rejected by the reference
with open(os.path.join(HERE, "quad_ascii.ply"), "w") as f:
    f.write("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\nelement face 1\n"
            "property list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n1 1 0\n0 1 0\n4 0 1 2 3\n")
print("written")

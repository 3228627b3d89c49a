import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge, oracle_binding
pkg = ge.load_package(); orc = oracle_binding.Oracle()
p = os.path.join(ROOT, "scenes", "cbox_quads.obj")
w = h = 48
with pkg.Context(0) as ctx:
    scene = ctx.load_scene(p); cam = pkg.default_camera(w, h); film = ctx.create_film(w, h)
    st = ctx.render(scene, cam, film, spp=256, max_depth=8, rng_mode=1); lin = film.linear().astype(np.float64)
    st2 = ctx.render(scene, cam, film, spp=256, max_depth=8, rng_mode=0); lin2 = film.linear().astype(np.float64)
print(st); print(st2)
osc = orc.load_obj(p); sm, sq = orc.render_moments(osc, w, h, 256, 8)
mean = sm / 256; var = np.maximum(sq / 256 - mean ** 2, 0) / 256
print("means philox/compat/oracle", lin.mean(), lin2.mean(), mean.mean())
z = (lin - mean) / np.sqrt(2 * var + 1e-8)
idx = np.argsort(-np.abs(z).ravel())[:10]
for i in idx:
    y, x, c = np.unravel_index(i, z.shape); print(y, x, c, z[y, x, c], lin[y, x, c], lin2[y,x,c], mean[y, x, c], var[y, x, c])
print("median z", np.median(z), "mean |z|<10:", z[np.abs(z) < 10].mean(), (np.abs(z) < 5).mean())
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "dbg_philox.npz"), philox=lin, compat=lin2, mean=mean, var=var)

"""Generates tests/golden/*.npz HERE (the container that has /root/reference) from the REFERENCE
ITSELF: oracle/_ref/libref_harness.so compiles the reference headers in place.  The fixtures let
the GPU box -- which has no /root/reference -- check the oracle and the CUDA path against the
reference.  Re-run: python tools/make_golden.py
"""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
assert O.have_ref(), "needs /root/reference"

# 1. the reference's committed golden image (image.ppm, UTF-16LE/CRLF) normalised -> rgb8
txt = open("/root/reference/image.ppm", "rb").read().decode("utf-16").replace("\r\n", "\n")
assert hashlib.md5(txt.encode()).hexdigest() == "594de05470fbb83478df57ad1e6ab351"
tok = txt.split()
assert tok[0] == "P3" and tok[1:4] == ["400", "225", "255"]
img = np.array(tok[4:], dtype=np.uint8).reshape(225, 400, 3)
np.savez_compressed(os.path.join(OUT, "reference_image_30spp.npz"), rgb8=img)

# 2. the scene the reference's own main() builds
s, m, draws = O.ref_scene_main(with_draws=True)
np.savez_compressed(os.path.join(OUT, "reference_scene.npz"), spheres=s, materials=m, draws=draws)

# 3. reference harness, keyed Philox stream, C1 (400x225, 10 spp, depth 50, seed 0), linear list
cam = O.camera_rtiow(400, 10, 50)
r = O.render(s, m, cam, rng_mode=O.RNG_PHILOX, seed=0, impl="ref", use_bvh=False)
np.savez_compressed(os.path.join(OUT, "ref_c1_philox_seed0.npz"), rgb8=r["rgb8"], rays=r["rays"],
                    sum_rows=r["sum"][100:104], sum_sha256=hashlib.sha256(r["sum"].tobytes()).hexdigest())

# 4. closest hit: camera rays + rays leaving surfaces + degenerate ones, reference linear list
rng = np.random.default_rng(7)
n = 1536
rays, _ = O.get_rays(cam, rng.integers(0, 400, n), rng.integers(0, 225, n), rng.integers(0, 10, n), seed=3, impl="ref")
h0 = O.closest_hit(s, rays, impl="ref")
hit = h0["prim"] >= 0
b = np.zeros(hit.sum(), O.RAY)
b["orig"] = h0["p"][hit]
d = rng.normal(size=(hit.sum(), 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
nn = O.vec3s(h0["normal"][hit]); d = d + nn
for k, c in enumerate("xyz"):
    b["dir"][c] = d[:, k]
b["tm"] = rays["tm"][hit]
ax = np.zeros(12, O.RAY)                       # axis-aligned / zero-component directions, origin inside a sphere
for k in range(12):
    ax[k]["orig"] = [(0.0, 5.0, 0.0), (4.0, 1.0, 0.0), (-4.0, 1.0, 3.0), (0.0, 1.0, 0.0)][k % 4]
    ax[k]["dir"] = [(0.0, -1.0, 0.0), (1.0, 0.0, 0.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), (-1.0, 0.0, 0.0), (0.0, 0.0, 1.0)][k % 6]
    ax[k]["tm"] = 0.25 * (k % 4)
allr = np.concatenate([rays, b, ax])
hr = O.closest_hit(s, allr, impl="ref")
np.savez_compressed(os.path.join(OUT, "ref_closest_hit.npz"), rays=allr, hits=hr)

# 5. scatter: every material type on reference hit records
sel = np.where(hr["prim"] >= 0)[0][:1500]
mats = m[s["mat"][hr["prim"][sel]]].copy()
flag, att, out = O.scatter(mats, allr[sel], hr[sel], seed=11, impl="ref")
np.savez_compressed(os.path.join(OUT, "ref_scatter.npz"), materials=mats, rays_in=allr[sel], hits=hr[sel], flag=flag, att=att, out=out)

# 6. camera rays
pi, pj, sm = rng.integers(0, 400, 2000), rng.integers(0, 225, 2000), rng.integers(0, 30, 2000)
gr, _ = O.get_rays(cam, pi, pj, sm, seed=5, impl="ref")
np.savez_compressed(os.path.join(OUT, "ref_get_rays.npz"), px_i=pi, px_j=pj, sample=sm, rays=gr)
for f in sorted(os.listdir(OUT)):
    print(f, os.path.getsize(os.path.join(OUT, f)))

# 8f-3: the demo's compute shader itself (compiled as C++, oracle/glsl_harness.cpp): a small image whole,
# the settings-size image (src/common_objects.h:8-9) by its sha256, and hit() on random rays / spheres
import hashlib
small = O.demo_compute(160, 120, impl="ref")
full = O.demo_compute(1024, 768, impl="ref")
rg = np.random.default_rng(11)
n = 4096
org = (rg.standard_normal((n, 3)) * 0.5).astype("f4")
dirs = rg.standard_normal((n, 3)).astype("f4")
cen = (rg.standard_normal((n, 3)) * 3).astype("f4")
rad = rg.uniform(0.2, 2.5, n).astype("f4")
flag = O.demo_hit(org, dirs, cen, rad, impl="ref")
np.savez_compressed(os.path.join(OUT, "ref_demo_shader.npz"), small=small, full_sha256=np.frombuffer(hashlib.sha256(full.tobytes()).digest(), np.uint8),
                    full_hit_pixels=np.int64((full[..., 0] == 255).sum()), origins=org, directions=dirs, centers=cen, radii=rad, hit=flag)
print("demo shader:", small.shape, int((full[..., 0] == 255).sum()), "hit pixels at 1024x768;", int(flag.sum()), "of", n, "random rays hit")

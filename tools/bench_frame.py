"""Framebuffer passes (the only HBM-bound part of the path) against the measured HBM
bandwidth: python tools/bench_frame.py   -> one JSON line per (image, slices, pass)

  resolve        : reads slices x W x H x 16 B (float4 sums), writes 3 B per pixel           (resolve_kernel)
  resolve+encode : the same + the P3 text (<= 12 B per pixel) written, ONE kernel              (resolve_encode_kernel, round 2)
  with RT_NO_FUSED_ENCODE=1 in the environment the text comes from the round-1 three-kernel encoder instead
  (encode_count / scan / write: the 3 B per pixel read twice more), reported as pass "encode_p3"

Times are rt_stats.resolve_ms (CUDA events on the context's stream around the passes), best of `reps`, with a
256 MB write between repetitions so that nothing is served from the 126 MB L2."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracingproject_b200 as rt

peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))).get("hbm_gbs", 6549.0)
ctx = rt.Context(0)
s, m = rt.scene_rtiow()
ctx.upload_scene(s, m)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
reps = 5


def passes(spp, want_text):
    best = 1e30
    n = C.c_size_t(0)
    for _ in range(reps):
        flush.fill_(1); torch.cuda.synchronize()
        ctx._ck(ctx._lib.rt_readback(ctx._h, spp, None, None, None, C.byref(n) if want_text else None))
        best = min(best, ctx.stats()["resolve_ms"])
    return best, n.value


for width, slices in ((1920, 1), (1920, 8), (3840, 1), (3840, 4), (3840, 8)):
    spp = max(2, slices)
    cam = rt.camera_rtiow(width, spp, 4)
    ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, slices=slices)
    ctx.synchronize()
    st = ctx.stats()
    W, H, sl = st["image_width"], st["image_height"], st["slices"]
    t_res, _ = passes(spp, False)
    t_all, text = passes(spp, True)
    t_enc = max(t_all - t_res, 1e-6)
    b_res = W * H * (16 * sl + 3)
    b_enc = W * H * 3 * 2 + text
    fused = not os.environ.get("RT_NO_FUSED_ENCODE")
    rows = (("resolve", t_res, b_res), ("resolve+encode_p3 (one kernel)", t_all, b_res + text)) if fused else \
           (("resolve", t_res, b_res), ("encode_p3", t_enc, b_enc), ("resolve+encode_p3 (four kernels)", t_all, b_res + text))
    for name, ms, nbytes in rows:
        gbs = nbytes / ms / 1e6
        print(json.dumps({"pass": name, "image": f"{W}x{H}", "slices": sl, "ms": round(ms, 4), "algorithmic_bytes": nbytes,
                          "gb_per_s": round(gbs, 1), "hbm_peak_gb_per_s": peak, "frac": round(gbs / peak, 3)}), flush=True)

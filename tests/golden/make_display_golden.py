"""Golden vectors of the display-step leaves, made in the build container from the reference's own headers through oracle/_ref
(dr::linear_to_srgb of ext/drjit/include/drjit/color.h; clip / mod / interpolate2d of src/mitsuba/util.h).  Run:
    python tests/golden/make_display_golden.py        (needs /root/reference; writes tests/golden/display_vectors.json)"""
import ctypes as C, json, os, sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding as ob

P = lambda a: a.ctypes.data_as(C.c_void_p)
u = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32).ravel().tolist()


def main():
    ref = ob.ref(); assert ref is not None, "oracle/_ref could not be built (needs /root/reference)"
    rng = np.random.default_rng(20261105); g = {}
    x = np.concatenate([rng.uniform(0, 1, 300), np.exp(rng.uniform(-18, 4, 300)), [0, 0.0031308, 0.00313081, 1.0, 0.5]]).astype(np.float32)
    y = np.zeros_like(x); ref.ref_linear_to_srgb(P(x), C.c_int(len(x)), P(y)); g["linear_to_srgb"] = dict(x=u(x), y=u(y))
    a = rng.uniform(-9, 9, 300).astype(np.float32); b = rng.integers(1, 9, 300).astype(np.float32); c = np.zeros_like(a); m = np.zeros_like(a)
    ref.ref_clip_mod(P(a), P(b), C.c_int(300), P(c), P(m)); g["clip_mod"] = dict(x=u(a), y=u(b), clip=u(c), mod=u(m))
    w, h, nch = 29, 17, 3
    img = rng.integers(0, 256, (h, w, nch), dtype=np.uint8)
    xy = np.stack([rng.uniform(0, w - 1, 400), rng.uniform(0, h - 1, 400)], 1).astype(np.float32); xy[:4] = [[w - 1, h - 1], [0, 0], [w - 1, 0], [0.5, 0.5]]
    g["interpolate2d"] = dict(w=w, h=h, nch=nch, img=img.ravel().tolist(), xy=u(xy), out={})
    for near in (0, 1):
        for ch in range(3):
            o = np.zeros(400, np.uint8); ref.ref_interpolate2d_u8(P(img), w, h, nch, P(xy), 400, ch, near, P(o)); g["interpolate2d"]["out"][f"{near}_{ch}"] = o.tolist()
    json.dump(g, open(os.path.join(ROOT, "tests", "golden", "display_vectors.json"), "w"))
    print("wrote", len(g), "entries")


if __name__ == "__main__":
    main()

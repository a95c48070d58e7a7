#!/usr/bin/env python
"""Fixture generator for the golden-render Z-tests (run in the container that has /root/reference; the outputs are committed).

1. tests/golden/render_refs.npz: the reference's own golden renders and variance images of the path-tracer scenes this build can load
   (resources/data/tests/scenes/*/refs/*_ref_rgb.exr + *_ref_rgb_var.exr: 48 x 48 / 32 x 32 float32 PIZ OpenEXR, read with OpenCV), the
   fixtures of src/render/tests/test_renders.py:168-260.
2. tests/golden/scenes/: the scene descriptions those renders belong to, with the meshes and the environment map they reference, copied
   byte for byte in the reference's directory layout so that the XML files load unmodified (they are data fixtures; the GPU box has no
   /root/reference).  Also the Cornell box scene of resources/data/scenes/cbox (BASELINE configs 1-3 use its geometry).
3. tests/golden/spectrum_refs.json: wavelength spectra of those scenes converted to linear sRGB by an independent numpy restatement of
   spectrum_list_to_srgb (src/core/spectrum.cpp:85-130) over the CIE tables parsed from the reference sources; also the measured
   (eta, k) spectra of three conductor `material` presets (resources/data/ior/{Au,Cu,Al}.*.spd, copied to tests/golden/scenes/ior/).
"""
import json, os, re, shutil
os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2
import numpy as np

REF = "/root/reference/resources/data"
HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = {
    "roughplastic_linear": ("tests/scenes/bsdf_spheres/test_roughplastic_linear_path.xml", "tests/scenes/bsdf_spheres/refs/test_roughplastic_linear_ref_rgb"),
    "roughplastic_nonlinear": ("tests/scenes/bsdf_spheres/test_roughplastic_nonlinear_path.xml", "tests/scenes/bsdf_spheres/refs/test_roughplastic_nonlinear_ref_rgb"),
    "roughdielectric": ("tests/scenes/bsdf_spheres/test_roughdielectric_path.xml", "tests/scenes/bsdf_spheres/refs/test_roughdielectric_ref_rgb"),
    "infinite_mirrors": ("tests/scenes/infinite_mirrors/scene_path.xml", "tests/scenes/infinite_mirrors/refs/scene_ref_rgb"),
}
COPY = ["tests/scenes/bsdf_spheres/fragments/single_sphere.xml", "common/meshes/sphere.obj", "common/meshes/rectangle.obj", "common/textures/museum.exr",
        "tests/serialized/rectangle_normals.serialized", "tests/serialized/rectangle_normals_uv.serialized", "tests/serialized/rectangle_uv.serialized",
        "scenes/cbox/cbox-rgb.xml", "scenes/cbox/fragments/base.xml", "scenes/cbox/fragments/bsdfs-rgb.xml", "scenes/cbox/fragments/shapes.xml"]


def main():
    out = {}
    for name, (xml, ref) in SCENES.items():
        for suffix, key in ((".exr", "ref"), ("_var.exr", "var")):
            img = cv2.imread(os.path.join(REF, ref + suffix), cv2.IMREAD_UNCHANGED)
            assert img is not None and img.dtype == np.float32, ref + suffix
            out[f"{name}_{key}"] = np.ascontiguousarray(img[..., ::-1])       # OpenCV returns BGR
        COPY.append(xml)
    np.savez_compressed(os.path.join(HERE, "render_refs.npz"), **out)
    for d in sorted(set(COPY)) + [os.path.join("scenes/cbox/meshes", f) for f in sorted(os.listdir(os.path.join(REF, "scenes/cbox/meshes")))]:
        dst = os.path.join(HERE, "scenes", d); os.makedirs(os.path.dirname(dst), exist_ok=True); shutil.copyfile(os.path.join(REF, d), dst)
    # ---- spectra: independent restatement in numpy float64
    src = open("/root/reference/src/core/spectrum.cpp").read()
    blk = src[src.index("cie1931_tbl[MI_CIE_SAMPLES * 3] = {"):]; blk = blk[:blk.index("};")]
    xyz_tbl = np.array(re.findall(r"Float\(([-0-9.eE+]+)\)", blk), np.float32).astype(np.float64).reshape(3, 95)
    hdr = open("/root/reference/include/mitsuba/core/spectrum.h").read()
    blk = hdr[hdr.index("d65_table[MI_CIE_SAMPLES] = {"):]; blk = blk[blk.index("{") + 1:blk.index("};")]
    d65_tbl = np.array(re.findall(r"([-0-9.eE+]+)f", blk), np.float32).astype(np.float64) / 98.99741751876255
    M = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]], np.float32).astype(np.float64)

    def to_srgb(wl, val, within_emitter, unbounded=False, flags=None):
        wl = np.asarray(wl, np.float64); val = np.asarray(val, np.float64) / 106.7502593994140625
        bounded, d65 = flags if flags is not None else (not (within_emitter or unbounded), not (within_emitter and not unbounded))
        w = 360.0 + (np.arange(1000) / 999.0) * 470.0
        w = w[(w >= wl[0]) & (w <= wl[-1])]
        v = np.interp(w, wl, val)
        t = (w - 360.0) * (94 / 470.0); i0 = np.clip(t.astype(np.int64), 0, 93); w1 = t - i0
        tab = lambda a: (1 - w1) * a[i0] + w1 * a[i0 + 1]
        xyz = np.array([np.sum(tab(xyz_tbl[c]) * v * (tab(d65_tbl) if d65 else 1.0)) for c in range(3)]) * (470.0 / 1000)
        rgb = np.maximum(M @ xyz, 0.0)
        return (np.minimum(rgb, 1.0) if bounded else rgb).tolist()
    cases = [dict(wavelengths=[400, 500, 600, 700], values=[0, 8, 15.6, 18.4], within_emitter=True),
             dict(wavelengths=[400, 500, 600, 700], values=[0.78, 0.78, 0.78, 0.78], within_emitter=False),
             dict(wavelengths=[380, 450, 520, 610, 780], values=[0.1, 0.9, 0.4, 0.7, 0.2], within_emitter=False),
             dict(wavelengths=[360, 830], values=[1.0, 1.0], within_emitter=False),
             dict(wavelengths=[360, 830], values=[2.5, 2.5], within_emitter=True)]
    for c in cases:
        c["rgb"] = to_srgb(c["wavelengths"], c["values"], c["within_emitter"])
    uniform_emitter = (M @ np.ones(3)).tolist()
    # conductor `material` presets (ior_from_file, include/mitsuba/render/ior.h:101-137: bounded = false, d65 = false) of three measured metals
    ior = {}
    for m in ("Au", "Cu", "Al"):
        for part in ("eta", "k"):
            fn = f"ior/{m}.{part}.spd"; dst = os.path.join(HERE, "scenes", fn); os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(os.path.join(REF, fn), dst)
            rows = [l.split() for l in open(dst) if l.strip() and not l.lstrip().startswith("#")]
            ior[f"{m}.{part}"] = to_srgb([float(r[0]) for r in rows], [float(r[1]) for r in rows], False, flags=(False, False))
    json.dump(dict(cases=cases, uniform_emitter_white=uniform_emitter, ior=ior), open(os.path.join(HERE, "spectrum_refs.json"), "w"), indent=1)
    print("wrote render_refs.npz,", len(COPY), "scene files, spectrum_refs.json")


if __name__ == "__main__":
    main()

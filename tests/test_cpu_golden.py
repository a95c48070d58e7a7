"""CPU-only tests of round 2 (no GPU): the pieces that pin the render loop to the reference's own fixtures.

* the oracle's emitter leaves (acos / atan2, uniform-sphere and bilinear warps, Hierarchical2D, DiscreteDistribution, grid placement)
  bit for bit against vectors produced by the reference's headers (tests/golden/reference_vectors.json, made by make_golden.py);
* the scene loader's readers: OpenEXR (PIZ) against OpenCV's decode of the same file, wavelength spectra against an independent numpy
  restatement over the reference's CIE tables (tests/golden/spectrum_refs.json);
* THE Z-TEST of src/render/tests/test_renders.py:152-235 for the CPU oracle on the reference's golden renders
  (resources/data/tests/scenes/{bsdf_spheres,infinite_mirrors}/refs/*_ref_rgb[_var].exr -> tests/golden/render_refs.npz): the scene
  files load unmodified through amvpt_scene_from_xml (host code), the oracle receives the loaded description.
"""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")))
SCENE_DIR = os.path.join(ROOT, "tests", "golden", "scenes")
RENDER_SCENES = {
    "roughplastic_linear": "tests/scenes/bsdf_spheres/test_roughplastic_linear_path.xml",
    "roughplastic_nonlinear": "tests/scenes/bsdf_spheres/test_roughplastic_nonlinear_path.xml",
    "roughdielectric": "tests/scenes/bsdf_spheres/test_roughdielectric_path.xml",
    "infinite_mirrors": "tests/scenes/infinite_mirrors/scene_path.xml",
}


def F(key, sub):
    return np.array(GOLDEN[key][sub], np.uint32).view(np.float32)


def same(a, want_bits, what):
    a = np.ascontiguousarray(a, np.float32).view(np.uint32).reshape(-1); w = np.asarray(want_bits, np.uint32).reshape(-1)
    bad = np.flatnonzero(a != w)
    assert bad.size == 0, f"{what}: {bad.size} of {a.size} values differ from the reference headers, first {bad[:5]}"


# ------------------------------------------------------------------ Z-test of the reference (src/render/tests/test_renders.py:152-235)
RGB_TO_XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]], np.float32)   # spectrum.h:395-401


def z_test_passes(img_rgb, spp, ref_rgb, ref_var_rgb, significance_level=0.01):
    """Returns (fraction of channel tests passed / 3 per pixel as the reference counts them, min p-value).  Images are converted to XYZ like
    the reference's test does for the moment integrator's output and for both fixture images."""
    from scipy.special import erf
    to_xyz = lambda a: np.asarray(a, np.float32) @ RGB_TO_XYZ.T
    mean, ref, var = to_xyz(img_rgb), to_xyz(ref_rgb), np.maximum(to_xyz(ref_var_rgb), 1e-4)
    z = np.abs(mean - ref) * np.sqrt(spp / var)
    cdf = (1.0 - erf(-z.astype(np.float64) / np.sqrt(2.0))) * 0.5
    p = 2.0 * (1.0 - cdf)
    pixel_count = mean.shape[0] * mean.shape[1]
    alpha = 1.0 - (1.0 - significance_level) ** (1.0 / pixel_count)
    success = p > alpha
    return np.count_nonzero(success) / 3 / pixel_count, float(p.min())


def load_render_case(pkg, name):
    refs = np.load(os.path.join(ROOT, "tests", "golden", "render_refs.npz"))
    ref, var = refs[name + "_ref"], refs[name + "_var"]
    spp = int(2e6) // (ref.shape[0] * ref.shape[1])                      # test_renders.py:186-188
    scene = pkg.load_file(os.path.join(SCENE_DIR, RENDER_SCENES[name]), spp=spp)
    return scene, ref, var, spp


@pytest.mark.parametrize("name", sorted(RENDER_SCENES))
def test_oracle_passes_reference_z_test(pkg, ob, name):
    scene, ref, var, spp = load_render_case(pkg, name)
    assert scene.info()["film_w"] == ref.shape[1] and scene.info()["film_h"] == ref.shape[0]
    o = ob.OracleScene.from_product(scene)
    lim = scene.integrator()["spp_pass_lim"]; spp_done = spp // lim * lim                    # mvpath.cpp:33-41 rounding
    film, _, _ = o.render(seed=0, spp=spp_done)
    assert np.abs(film[..., 3] - spp_done).max() <= 2          # pixel + jitter can round up into the next pixel (float32), a handful per render
    frac, pmin = z_test_passes(ob.develop(film), spp_done, ref, var)
    assert frac >= 0.9975, f"{name}: only {frac:.4%} of the pixels pass the reference's Z-test (min p-value {pmin:.3g})"


def test_z_test_has_power(pkg, ob):
    """The same test must reject a render that is wrong by a few percent (otherwise passing it proves nothing)."""
    scene, ref, var, spp = load_render_case(pkg, "roughplastic_linear")
    o = ob.OracleScene.from_product(scene); film, _, _ = o.render(seed=0, spp=spp // 16 * 16)
    img = ob.develop(film)
    assert z_test_passes(img, spp // 16 * 16, ref, var)[0] >= 0.9975
    assert z_test_passes(img * 1.06, spp // 16 * 16, ref, var)[0] < 0.9975
    assert z_test_passes(img[:, ::-1], spp // 16 * 16, ref, var)[0] < 0.9975


# ------------------------------------------------------------------ scene-loader readers
def test_exr_piz_reader_matches_opencv(pkg):
    path = os.path.join(SCENE_DIR, "common", "textures", "museum.exr")
    img = pkg.image_read(path)
    assert img.shape == (256, 512, 3) and img.dtype == np.float32
    # SHA-256 of OpenCV's decode of the same file (cv2.imread(IMREAD_UNCHANGED)[..., ::-1].astype(float32)), recorded when the fixture was made
    assert hashlib.sha256(img.tobytes()).hexdigest() == MUSEUM_SHA256
    try:
        os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
        import cv2
        ref = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    except Exception:
        ref = None
    if ref is not None:
        assert np.array_equal(img, ref[..., ::-1].astype(np.float32))


MUSEUM_SHA256 = "b55a24c28fe07a55e325eaccd7b917c055c5cf8af09412421f4486d2d8807ba5"


def test_exr_reader_rejects_what_it_cannot_decode(pkg, tmp_path):
    p = tmp_path / "bad.exr"; p.write_bytes(b"\x76\x2f\x31\x01\x02\x02\x00\x00" + b"\x00" * 16)        # tiled flag set
    with pytest.raises(pkg.AmvptError, match="tiled"):
        pkg.image_read(p)
    q = tmp_path / "x.png"; q.write_bytes(b"\x89PNG\r\n\x1a\n")
    with pytest.raises(pkg.AmvptError, match="unsupported image format"):
        pkg.image_read(q)
    # PFM round trip, both byte orders
    rng = np.random.default_rng(0); img = rng.uniform(0, 4, (5, 7, 3)).astype(np.float32)
    for scale, dt in (("-1.0", "<f4"), ("1.0", ">f4")):
        f = tmp_path / f"a{scale}.pfm"; f.write_bytes(f"PF\n7 5\n{scale}\n".encode() + img[::-1].astype(dt).tobytes())
        assert np.array_equal(pkg.image_read(f), img)


def test_spectrum_to_srgb_matches_independent_restatement(pkg):
    refs = json.load(open(os.path.join(ROOT, "tests", "golden", "spectrum_refs.json")))
    for c in refs["cases"]:
        got = pkg.spectrum_to_srgb(c["wavelengths"], c["values"], within_emitter=c["within_emitter"])
        assert np.allclose(got, c["rgb"], rtol=1e-6, atol=1e-9), (c, got)
    # "integrating the spectrum against the CIE curves and converting to sRGB yields (1, 1, 1) for D65" (src/core/xml.cpp:1200-1204)
    assert np.allclose(pkg.spectrum_to_srgb([360, 830], [1, 1]), [1, 1, 1], atol=2e-3)
    assert np.allclose(pkg.spectrum_to_srgb([0], [10.0], within_emitter=True), 10.0 * np.asarray(refs["uniform_emitter_white"]), rtol=1e-7)
    assert np.array_equal(pkg.spectrum_to_srgb([0], [0.789]), [0.789] * 3)                     # uniform reflectance / eta stays uniform
    assert np.array_equal(pkg.spectrum_to_srgb([0], [0.789], within_emitter=True, unbounded_name=True), [0.789] * 3)


def test_golden_scenes_load_unmodified(pkg):
    s = pkg.load_file(os.path.join(SCENE_DIR, RENDER_SCENES["roughplastic_linear"]), spp=64)
    assert s.counts() == dict(n_meshes=5, n_bsdfs=5, n_emitters=3)
    em = s.emitters()
    assert [e["type"] for e in em] == ["area", "envmap", "point"]                             # scene order (src/render/scene.cpp:22-63)
    assert np.allclose(em[0]["radiance"], [20.710836, 10.854634, 2.771152], rtol=1e-6)         # <spectrum value="400:0, 500:8, 600:15.6, 700:18.4"/>
    assert np.array_equal(em[2]["position"], [-4, 0, -6]) and np.array_equal(em[2]["radiance"], [300, 300, 300])
    assert s.envmap()["rgb"].shape == (256, 512, 3) and s.integrator()["max_depth"] == 2 and not s.integrator()["is_mvpath"]
    b = s.bsdf(0); assert b["type"] == "roughplastic" and b["distribution"] == "ggx" and abs(b["alpha_u"] - 0.01) < 1e-9
    s = pkg.load_file(os.path.join(SCENE_DIR, RENDER_SCENES["infinite_mirrors"]))
    assert s.counts()["n_emitters"] == 1 and s.integrator()["max_depth"] == 8 and s.info()["sample_count"] == 64
    # the Cornell box of resources/data/scenes/cbox (BASELINE configs 1-3 take its geometry)
    s = pkg.load_file(os.path.join(SCENE_DIR, "scenes", "cbox", "cbox-rgb.xml"))
    assert s.counts()["n_meshes"] == 8 and s.info()["n_prims"] == 38 and s.counts()["n_emitters"] == 1


def test_conductor_material_presets(pkg):
    """`material` of conductor / roughconductor (conductor.cpp:228-241): complex_ior_from_file reads data/ior/<name>.{eta,k}.spd and, in an RGB
    variant, converts each to an unbounded sRGB colour.  Fixtures: three of the reference's measured spectra + an independent numpy restatement."""
    refs = json.load(open(os.path.join(ROOT, "tests", "golden", "spectrum_refs.json")))["ior"]
    xml = """<scene version="3.0.0">
      <sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film></sensor>
      <shape type="obj"><string name="filename" value="meshes/cbox_floor.obj"/><bsdf type="conductor"><string name="material" value="Au"/></bsdf></shape>
      <shape type="obj"><string name="filename" value="meshes/cbox_back.obj"/>
        <bsdf type="roughconductor"><string name="material" value="Cu"/><float name="alpha" value="0.2"/></bsdf></shape>
      <shape type="obj"><string name="filename" value="meshes/cbox_ceiling.obj"/><bsdf type="twosided"><bsdf type="conductor"><string name="material" value="Al"/></bsdf></bsdf></shape>
      <shape type="obj"><string name="filename" value="meshes/cbox_luminaire.obj"/><bsdf type="conductor"><string name="material" value="none"/></bsdf>
        <emitter type="area"><rgb name="radiance" value="1"/></emitter></shape>
    </scene>"""
    base = os.path.join(SCENE_DIR, "scenes", "cbox")
    s = pkg.load_string(xml, base_dir=base)
    for i, m in enumerate(("Au", "Cu", "Al")):
        b = s.bsdf(i)
        assert np.allclose(b["eta"], refs[m + ".eta"], rtol=2e-6) and np.allclose(b["k"], refs[m + ".k"], rtol=2e-6), (m, b["eta"], b["k"])
    assert np.allclose(s.bsdf(0)["eta"], [0.143, 0.375, 1.442], atol=1e-3) and np.allclose(s.bsdf(0)["k"], [3.983, 2.386, 1.603], atol=1e-3)   # gold as the documentation tabulates it
    assert tuple(s.bsdf(3)["eta"]) == (0.0, 0.0, 0.0) and tuple(s.bsdf(3)["k"]) == (1.0, 1.0, 1.0)             # material = none: a perfect mirror
    assert s.bsdf(1)["type"] == "roughconductor" and s.bsdf(2)["twosided"]
    # <spectrum filename="x.spd"> (src/core/xml.cpp:805-815): the file's pairs go through the same conversion as value="w:v ..." pairs
    rows = [l.split() for l in open(os.path.join(SCENE_DIR, "ior", "Cu.k.spd")) if l.strip() and not l.lstrip().startswith("#")]
    want = pkg.spectrum_to_srgb([float(r[0]) for r in rows], [float(r[1]) for r in rows], unbounded_name=True)
    s4 = pkg.load_string(xml.replace('<string name="material" value="Au"/>', '<spectrum name="eta" value="0.2"/><spectrum name="k" filename="../../ior/Cu.k.spd"/>'), base_dir=base)
    assert np.allclose(s4.bsdf(0)["k"], want, rtol=2e-6) and np.allclose(s4.bsdf(0)["eta"], 0.2)
    with pytest.raises(pkg.AmvptError, match="Should specify either"):
        pkg.load_string(xml.replace('<string name="material" value="Au"/>', '<string name="material" value="Au"/><rgb name="eta" value="1"/>'), base_dir=base)
    with pytest.raises(pkg.AmvptError, match="file does not exist"):
        pkg.load_string(xml.replace('value="Au"', 'value="Unobtainium"'), base_dir=base)


def _serialized_blob(meshes, version=4):
    """Writes the container of src/shapes/serialized.cpp:85-178 (independent of the loader under test): per sub-mesh a 4-byte header and one
    zlib stream; the offset dictionary + count at the end."""
    import struct, zlib
    out, offs = b"", []
    for name, pos, nrm, faces, dp in meshes:
        offs.append(len(out))
        flags = (0x2000 if dp else 0x1000) | (0x0001 if nrm is not None else 0)
        ft = "<f8" if dp else "<f4"
        body = struct.pack("<I", flags) + (name.encode() + b"\0" if version == 4 else b"") + struct.pack("<QQ", len(pos), len(faces))
        body += np.asarray(pos, ft).tobytes() + (np.asarray(nrm, ft).tobytes() if nrm is not None else b"") + np.asarray(faces, "<u4").tobytes()
        out += struct.pack("<HH", 0x041C, version) + zlib.compress(body)
    out += b"".join(struct.pack("<Q" if version == 4 else "<I", o) for o in offs) + struct.pack("<I", len(offs))
    return out


def test_serialized_meshes(pkg, tmp_path):
    """`serialized` shape plugin: the reference's own fixtures with the known answers of src/render/tests/test_mesh.py:151-196 (test06), and
    sub-mesh selection / double precision / format 3 on files written here."""
    scene = """<scene version="3.0.0">
      <sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film></sensor>
      <shape type="serialized"><string name="filename" value="$fn"/><boolean name="face_normals" value="$fnrm"/><integer name="shape_index" value="$idx"/>
        <emitter type="area"><rgb name="radiance" value="1"/></emitter></shape></scene>"""
    fixtures = os.path.join(SCENE_DIR, "tests", "serialized")
    for features in ("normals", "uv", "normals_uv"):
        for face_normals in (True, False):
            s = pkg.load_string(scene, base_dir=fixtures, fn=f"rectangle_{features}.serialized", fnrm=str(face_normals).lower(), idx=0)
            m = s.mesh(0); p = m["positions"]
            assert np.allclose(p[0], [-2.85, 0.0, -7.6], atol=1e-3) and np.allclose(p[2], [2.85, 0.0, 0.599999], atol=1e-3) and np.allclose(p[3], [2.85, 0.0, -7.6], atol=1e-3)
            assert (m["normals"] is not None) == (not face_normals)
            if m["normals"] is not None:
                assert np.allclose(m["normals"][[0, 2, 3]], [0.0, 1.0, 0.0])
    rng = np.random.default_rng(5)
    def mesh(nv, nf, normals): return (rng.uniform(-1, 1, (nv, 3)), (lambda n: n / np.linalg.norm(n, axis=1, keepdims=True))(rng.normal(size=(nv, 3))) if normals else None,
                                       rng.integers(0, nv, (nf, 3)))
    a, b, c = mesh(7, 5, True), mesh(300, 411, False), mesh(12, 9, True)
    for version in (4, 3):
        fn = tmp_path / f"multi_v{version}.serialized"
        fn.write_bytes(_serialized_blob([("first", *a, False), ("second", *b, True), ("third", *c, True)], version))
        for idx, (pos, nrm, faces), dp in ((0, a, False), (1, b, True), (2, c, True)):
            m = pkg.load_string(scene, base_dir=str(tmp_path), fn=fn.name, fnrm="false", idx=idx).mesh(0)
            assert np.array_equal(m["positions"], np.asarray(pos, np.float64 if dp else np.float32).astype(np.float32)), (version, idx)
            assert np.array_equal(m["faces"], faces)
            assert m["normals"] is not None                      # given, or recomputed (serialized.cpp:396-401)
            if nrm is not None:
                assert np.allclose(m["normals"], nrm, atol=1e-6)
        with pytest.raises(pkg.AmvptError, match="shape index is out of range"):
            pkg.load_string(scene, base_dir=str(tmp_path), fn=fn.name, fnrm="false", idx=3)
    bad = tmp_path / "bad.serialized"; bad.write_bytes(b"\x1d\x04\x04\x00" + b"garbage")
    with pytest.raises(pkg.AmvptError, match="invalid file format"):
        pkg.load_string(scene, base_dir=str(tmp_path), fn=bad.name, fnrm="false", idx=0)
    bad.write_bytes(_serialized_blob([("x", *a, False)])[:40])
    with pytest.raises(pkg.AmvptError, match="serialized file"):
        pkg.load_string(scene, base_dir=str(tmp_path), fn=bad.name, fnrm="false", idx=0)


# ------------------------------------------------------------------ emitter leaves vs the reference's headers
def test_oracle_emitter_leaves_match_reference_headers(ob):
    o = ob.orc(); P = lambda x, t=ob.f32p: x.ctypes.data_as(t)
    for op in ("12", "13", "14", "15", "16"):
        x = F("math1", op) if False else np.array(GOLDEN["math1"][op]["x"], np.uint32).view(np.float32)
        same(ob.math1(int(op), x), GOLDEN["math1"][op]["y"], f"math op {op}")
    u2 = F("disk", "u").reshape(-1, 2).copy(); n = len(u2)
    o3 = np.zeros((n, 3), np.float32); o.orc_square_to_uniform_sphere(P(u2), C.c_int(n), P(o3)); same(o3, GOLDEN["uniform_sphere"]["y"], "uniform sphere")
    v4 = F("bilinear", "v").reshape(-1, 4).copy(); o2 = np.zeros((n, 2), np.float32); p1 = np.zeros(n, np.float32); p2 = np.zeros(n, np.float32)
    o.orc_square_to_bilinear(P(v4), P(u2), C.c_int(n), P(o2), P(p1), P(p2))
    same(o2, GOLDEN["bilinear"]["y"], "square_to_bilinear"); same(p1, GOLDEN["bilinear"]["pdf"], "square_to_bilinear pdf"); same(p2, GOLDEN["bilinear"]["pdf_eval"], "square_to_bilinear_pdf")
    g = GOLDEN["hier2d"]; data = F("hier2d", "data").copy(); u = F("hier2d", "u").reshape(-1, 2).copy(); m = len(u)
    uv = np.zeros((m, 2), np.float32); pdf = np.zeros(m, np.float32); ev = np.zeros(m, np.float32)
    o.orc_hier2d(P(data), C.c_uint32(g["sx"]), C.c_uint32(g["sy"]), P(u), C.c_int(m), P(uv), P(pdf), P(ev))
    same(uv, g["uv"], "Hierarchical2D sample"); same(pdf, g["pdf"], "Hierarchical2D sample pdf"); same(ev, g["eval"], "Hierarchical2D eval")
    assert np.allclose(pdf, ev, rtol=2e-4)                                  # eval(sample(u)) is the sample's density
    g = GOLDEN["discrete"]; pmf = F("discrete", "pmf").copy(); du = F("discrete", "u").copy(); k = len(du)
    idx = np.zeros(k, np.uint32); re_ = np.zeros(k, np.float32); po = np.zeros(k, np.float32); s1 = np.zeros(1, np.float32); s2 = np.zeros(1, np.float32)
    o.orc_discrete_sample_reuse(P(pmf), C.c_uint32(len(pmf)), P(du), C.c_int(k), idx.ctypes.data_as(ob.u32p), P(re_), P(po), P(s1), P(s2))
    assert idx.tolist() == g["index"]
    same(s1, g["sum"], "DiscreteDistribution sum"); same(s2, g["normalization"], "DiscreteDistribution normalization")
    same(po, g["pmf_out"], "DiscreteDistribution pmf"); same(re_, g["reused"], "DiscreteDistribution reused sample")
    g = GOLDEN["grid_to_world"]; it = F("grid_to_world", "inv_t").copy()
    for mrec in g["mats"]:
        off = np.array(mrec["offset"], np.uint32).view(np.float32).copy(); om = np.zeros(16, np.float32)
        o.orc_grid_to_world(P(it), P(off), P(om)); same(om, mrec["m"], "grid sub-camera to_world")


def test_product_grid_placement_matches_reference_headers(pkg, scenes):
    """The product's own GridSensor placement (csrc/scene.cpp set_sensor_grid) against the matrices the reference's headers produce."""
    g = GOLDEN["grid_to_world"]
    s = pkg.Scene(); b = s.add_bsdf(type="diffuse")
    s.add_mesh(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32), np.array([[0, 1, 2]], np.uint32), b)
    tw = np.eye(4); tw[:3, 3] = [278, 273, -800]; tw[0, 0] = tw[2, 2] = -1
    s.set_sensor_grid(8, 1, 8 * 16, 16, tw, 39.3077, fov_axis="smaller", near_clip=10, far_clip=2800, cam_dir=(1, 0, 0), cam_dist=400.0, cam_center=True)
    for i, mrec in enumerate(g["mats"]):
        _, m = s.sensor(i)
        same(m[:3], np.array(mrec["m"], np.uint32).reshape(4, 4)[:3], f"product sub-camera {i} to_world")


# ------------------------------------------------------------------ chi-square tests of BSDF sampling (the reference's own method)
CHI2_CASES = {
    # src/bsdfs/tests/test_diffuse.py:13-36, test_rough_conductor.py:60-96, test_rough_dielectric.py, test_rough_plastic.py:14-39
    "diffuse": (dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)), (0, 0, 1), {}),
    "roughconductor_beckmann_iso": (dict(type="roughconductor", distribution="beckmann", alpha=0.5, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)), (0.3, 0.1, 0.9), {}),
    "roughconductor_ggx_aniso": (dict(type="roughconductor", distribution="ggx", alpha_u=0.25, alpha_v=0.5, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)), (0.6, 0.0, 0.8), {}),
    "roughdielectric_beckmann_outside": (dict(type="roughdielectric", distribution="beckmann", alpha=0.5, int_ior=1.5, ext_ior=1.0), (0.3, 0.2, 0.93), dict(ires=8, transmissive=True)),
    "roughdielectric_ggx_inside": (dict(type="roughdielectric", distribution="ggx", alpha=0.5, int_ior=1.5, ext_ior=1.0), (0.2, 0.1, -0.97), dict(ires=8, transmissive=True)),
    "roughplastic_beckmann": (dict(type="roughplastic", distribution="beckmann", alpha=0.25, reflectance=(0.1, 0.1, 0.1), specular_reflectance=(0.7, 0.7, 0.7)), (0, 0, 1), {}),
    "roughplastic_ggx_slanted": (dict(type="roughplastic", distribution="ggx", alpha=0.3, reflectance=(0.5, 0.3, 0.1)), (0.7, 0.1, 0.7), {}),
}


def chi2_scene(scenes, cfg):
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=8)
    desc["bsdfs"] = [cfg]
    for m in desc["meshes"]:
        m["bsdf"] = 0
    return desc


@pytest.mark.parametrize("name", sorted(CHI2_CASES))
def test_oracle_bsdf_sampling_chi2(ob, scenes, name):
    from chi2 import bsdf_chi2
    cfg, wi, kw = CHI2_CASES[name]
    o = ob.OracleScene(chi2_scene(scenes, cfg))
    p, info = bsdf_chi2(lambda a, b, c, d: o.bsdf_eval_pdf_sample(0, a, b, c, d), wi, **kw)
    alpha = 1.0 - (1.0 - 0.01) ** (1.0 / len(CHI2_CASES))             # Sidak correction over the cases (chi2.py:239-254, test_count)
    assert p > alpha, f"{name}: chi^2 rejects sample() against pdf(): p = {p:.3g}, {info}"
    assert abs(info["hist_sum"] - info["pdf_sum"]) < 5e-3


def test_chi2_has_power(ob, scenes):
    """The harness rejects a sampler whose density is off: roughconductor sampled with alpha 0.5, density tabulated with alpha 0.45."""
    from chi2 import bsdf_chi2
    a = ob.OracleScene(chi2_scene(scenes, dict(type="roughconductor", distribution="ggx", alpha=0.5, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14))))
    b = ob.OracleScene(chi2_scene(scenes, dict(type="roughconductor", distribution="ggx", alpha=0.45, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14))))

    def mixed(wi, wo, s1, s2):
        r = a.bsdf_eval_pdf_sample(0, wi, wo, s1, s2); r["pdf"] = b.bsdf_eval_pdf_sample(0, wi, wo, s1, s2)["pdf"]; return r
    p, _ = bsdf_chi2(mixed, (0.3, 0.1, 0.9))
    assert p < 1e-6

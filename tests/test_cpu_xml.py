"""CPU-only tests of the scene-description boundary: the XML loader (csrc/xml.cpp) and the mesh plugins
(csrc/meshio.cpp) must produce the same host scene as the programmatic C-ABI calls -- sub-camera matrices, BSDF records and
per-face vertex data bit for bit -- and must reject what the reference rejects.  The scene files are written by
scenes.export_xml in the format of the reference's resources/data/scenes/cbox (XML 3.0.0, <include>, <default>, <ref>)."""
import os

import numpy as np
import pytest


RELEVANT = {"diffuse": ["reflectance"], "conductor": ["eta", "k", "specular_reflectance"],
            "dielectric": ["int_ior", "specular_reflectance", "specular_transmittance"],
            "roughconductor": ["eta", "k", "specular_reflectance", "distribution", "alpha_u", "alpha_v"],
            "roughdielectric": ["int_ior", "specular_reflectance", "specular_transmittance", "distribution", "alpha_u", "alpha_v"],
            "roughplastic": ["int_ior", "reflectance", "specular_reflectance", "distribution", "alpha_u", "nonlinear"]}


def same_scene(a, b, normals_tol=0.0):
    ia, ib = a.info(), b.info()
    for k in ("film_w", "film_h", "n_views", "n_prims", "sample_count"):
        assert ia[k] == ib[k], k
    for i in range(ia["n_views"]):
        (s1, t1), (s2, t2) = a.sensor(i), b.sensor(i)
        assert np.array_equal(s1.view(np.uint32), s2.view(np.uint32)), f"sample_to_camera {i}"
        assert np.array_equal(t1.view(np.uint32), t2.view(np.uint32)), f"to_world {i}"
    ca, cb = a.counts(), b.counts(); assert ca == cb
    for i in range(ca["n_bsdfs"]):
        x, y = a.bsdf(i), b.bsdf(i)
        for k in RELEVANT[x["type"]] + ["type", "twosided"]:       # only the parameters the plugin reads
            assert x[k] == y[k], f"bsdf {i} {k}: {x[k]} != {y[k]}"
        ta, tb = a.bsdf_tables(i), b.bsdf_tables(i)
        assert np.array_equal(ta[0], tb[0]) and np.array_equal(ta[1], tb[1])
    for i in range(ca["n_meshes"]):
        x, y = a.mesh(i), b.mesh(i)
        assert np.array_equal(x["faces"].shape, y["faces"].shape)
        assert np.array_equal(x["positions"][x["faces"]].view(np.uint32), y["positions"][y["faces"]].view(np.uint32)), f"mesh {i} vertex data"
        assert x["bsdf"] == y["bsdf"] and x["flip_normals"] == y["flip_normals"] and x["radiance"] == y["radiance"]
        assert (x["normals"] is None) == (y["normals"] is None)
        if x["normals"] is not None:
            assert np.abs(x["normals"][x["faces"]] - y["normals"][y["faces"]]).max() <= normals_tol
    assert a.integrator() == b.integrator()


@pytest.mark.parametrize("fmt", ["obj", "ply"])
def test_xml_matches_programmatic_cbox(pkg, scenes, tmp_path, fmt):
    desc = scenes.cornell_box(grid_x=4, grid_y=2, view_res=32, cam_dist=300.0, sample_count=48)
    path = scenes.export_xml(desc, str(tmp_path), mesh_format=fmt)
    a = pkg.load_file(path); b = scenes.build(desc, commit=False)
    same_scene(a, b)
    assert a.integrator()["is_mvpath"] and a.integrator()["sample_count"] == 48
    c = pkg.load_file(path, spp=16)          # -Dspp=16 overrides <default>
    assert c.info()["sample_count"] == 16
    # mvpath's sample-reuse parameters (mvpath.h:120-129) travel through the XML like through set_reuse
    desc["reuse"] = dict(sa_reuse=True, sa_mis=True, fast_mis=True, reuse_count=4)
    path = scenes.export_xml(desc, str(tmp_path), mesh_format=fmt, filename="reuse.xml")
    a = pkg.load_file(path); b = scenes.build(desc, commit=False)
    same_scene(a, b)
    assert a.integrator()["sa_reuse"] and a.integrator()["reuse_count"] == 4 and b.n_reuse == 4
    assert b.set_reuse(sa_reuse=True, reuse_count=3) == 8       # 3 does not divide 8 sub-sensors: first divisor >= 8 (mvpath.cpp:196-217)
    assert b.set_reuse(sa_reuse=True, reuse_count=1) == 0 and b.set_reuse(sa_reuse=False, reuse_count=4) == 0


def test_xml_materials_thinlens_cone_and_normals(pkg, scenes, tmp_path):
    desc = scenes.height_field(n=24, grid_x=2, grid_y=2, view_res=16)
    desc["bsdfs"][0]["twosided"] = True
    desc["sensor"].update(thinlens=True, aperture_radius=0.02, cone_deg=12.0)
    for fmt, tol in (("ply", 2e-7), ("obj", 2e-7)):
        d = tmp_path / fmt; d.mkdir()
        a = pkg.load_file(scenes.export_xml(desc, str(d), mesh_format=fmt)); b = scenes.build(desc, commit=False)
        same_scene(a, b, normals_tol=tol)
        assert a.bsdf(0)["twosided"] and a.bsdf(1)["type"] == "roughdielectric"


CBOX_STYLE = """<scene version="3.0.0">
    <default name="res" value="16"/>
    <include filename="fragments/base.xml"/>
    <bsdf type="diffuse" id="white"><rgb name="reflectance" value="0.885809, 0.698859, 0.666422"/></bsdf>
    <alias id="white" as="also_white"/>
    <shape type="obj">
        <string name="filename" value="meshes/quad.obj"/>
        <transform name="to_world"><scale value="2"/><rotate y="1" angle="90"/><translate x="1" y="-0.5" z="3"/></transform>
        <ref id="also_white"/>
        <ref id="area-emitter"/>
    </shape>
    <shape type="cube"><bsdf type="dielectric"><string name="int_ior" value="water"/></bsdf></shape>
    <emitter type="area" id="area-emitter"><rgb name="radiance" value="18.387, 10.9873, 2.75357"/></emitter>
</scene>
"""
BASE = """<scene version="3.0.0">
    <integrator type="path"><integer name="max_depth" value="6"/></integrator>
    <sensor type="perspective">
        <string name="fov_axis" value="smaller"/><float name="near_clip" value="10"/><float name="far_clip" value="2800"/>
        <float name="fov" value="39.3077"/>
        <transform name="to_world"><lookat origin="278, 273, -800" target="278, 273, -799" up="0, 1, 0"/></transform>
        <sampler type="independent"><integer name="sample_count" value="32"/></sampler>
        <film type="hdrfilm"><integer name="width" value="$res"/><integer name="height" value="$res"/><rfilter type="box"/>
            <string name="pixel_format" value="rgb"/></film>
    </sensor>
</scene>
"""
QUAD = "v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n"


def test_reference_style_scene_with_includes(pkg, tmp_path):
    (tmp_path / "fragments").mkdir(); (tmp_path / "meshes").mkdir()
    (tmp_path / "scene.xml").write_text(CBOX_STYLE); (tmp_path / "fragments" / "base.xml").write_text(BASE)
    (tmp_path / "meshes" / "quad.obj").write_text(QUAD)
    s = pkg.load_file(str(tmp_path / "scene.xml"))
    info = s.info(); assert (info["film_w"], info["film_h"], info["n_views"], info["sample_count"]) == (16, 16, 1, 32)
    assert not s.integrator()["is_mvpath"] and s.integrator()["max_depth"] == 6
    assert s.counts() == dict(n_meshes=2, n_bsdfs=2, n_emitters=1)
    q = s.mesh(0)
    # polygon -> fan (obj.cpp:323-327); to_world = translate * rotate_y(90) * scale(2): (x, y, 0) -> (1, 2y - .5, 3 - 2x)
    assert q["faces"].tolist() == [[0, 1, 2], [0, 2, 3]]
    assert np.allclose(q["positions"], [[1, -.5, 3], [1, -.5, 1], [1, 1.5, 1], [1, 1.5, 3]], atol=1e-6)
    assert np.allclose(q["normals"], [[1, 0, 0]] * 4, atol=1e-6) and q["radiance"] == pytest.approx((18.387, 10.9873, 2.75357))
    cube = s.mesh(1)
    assert cube["faces"].shape == (12, 3) and cube["positions"].shape == (24, 3) and cube["faces"][1].tolist() == [3, 0, 2]
    assert np.array_equal(cube["positions"][0], [1, -1, -1]) and np.array_equal(cube["normals"][0], [0, -1, 0])
    b = s.bsdf(cube["bsdf"]); assert b["type"] == "dielectric" and b["int_ior"] == pytest.approx(1.3330 / 1.000277)
    # the centre ray of the cbox camera (reference test: src/sensors/tests/test_perspective.py:79-84)
    s2c, tw = s.sensor(0); assert np.allclose(tw[:3, 3], [278, 273, -800]) and np.allclose(tw[:3, 2], [0, 0, 1])
    s = pkg.load_file(str(tmp_path / "scene.xml"), res=64); assert s.info()["film_w"] == 64


def test_xml_errors(pkg, tmp_path):
    def load(body, **kw):
        return pkg.load_string(f'<scene version="3.0.0">{body}</scene>', str(tmp_path), **kw)
    sensor = ('<sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/>'
              '<rfilter type="box"/></film></sensor>')
    shape = '<shape type="cube"/>'
    assert load(sensor + shape).info()["n_prims"] == 12
    assert load(sensor.replace('<rfilter type="box"/>', '') + shape).info()["n_prims"] == 12      # hdrfilm default: gaussian, stddev 0.5
    assert load(sensor.replace('<rfilter type="box"/>', '<rfilter type="tent"><float name="radius" value="1.5"/></rfilter>') + shape).info()["n_prims"] == 12
    for rf in ('<rfilter type="lanczos"><integer name="lobes" value="4"/></rfilter>', '<rfilter type="catmullrom"/>', '<rfilter type="mitchell"/>',
               '<rfilter type="mitchell"><float name="B" value="0.55"/><float name="C" value="0.2"/></rfilter>'):
        assert load(sensor.replace('<rfilter type="box"/>', rf) + shape).info()["n_prims"] == 12
    with pytest.raises(pkg.AmvptError, match="unknown reconstruction filter"):
        load(sensor.replace('type="box"', 'type="sinc"') + shape)
    with pytest.raises(pkg.AmvptError, match="too large"):
        load(sensor.replace('<rfilter type="box"/>', '<rfilter type="lanczos"><integer name="lobes" value="9"/></rfilter>') + shape)
    with pytest.raises(pkg.AmvptError, match="requires MultiSensor"):           # mvpath.cpp:15-17
        load('<integrator type="mvpath"/>' + sensor + shape)
    with pytest.raises(pkg.AmvptError, match="undefined parameter"):
        load(sensor.replace('"8"', '"$res"') + shape)
    with pytest.raises(pkg.AmvptError, match="unreferenced property"):          # plugin.cpp:190-205
        load(sensor + '<shape type="cube"><float name="radius" value="1"/></shape>')
    with pytest.raises(pkg.AmvptError, match="unknown object"):
        load(sensor + '<shape type="cube"><ref id="nope"/></shape>')
    with pytest.raises(pkg.AmvptError, match="not part of the AMVPT path"):
        load(sensor + '<shape type="sphere"/>')
    with pytest.raises(pkg.AmvptError, match="file not found"):
        load(sensor + '<shape type="obj"><string name="filename" value="missing.obj"/></shape>')
    with pytest.raises(pkg.AmvptError, match="duplicate id"):
        load(sensor + '<bsdf type="diffuse" id="a"/><bsdf type="diffuse" id="a"/>' + shape)
    with pytest.raises(pkg.AmvptError, match="version"):
        pkg.load_string("<scene>" + sensor + shape + "</scene>")
    with pytest.raises(pkg.AmvptError, match="mismatched closing tag"):
        pkg.load_string('<scene version="3.0.0"><shape type="cube"></scene>')
    with pytest.raises(pkg.AmvptError, match="transmission"):                   # twosided.cpp:101-102
        load(sensor + '<shape type="cube"><bsdf type="twosided"><bsdf type="dielectric"/></bsdf></shape>')
    # wavelength spectra are integrated to sRGB (src/core/xml.cpp:777-833, 1170-1268); malformed pairs raise
    sp = load(sensor + '<shape type="cube"><bsdf type="diffuse"><spectrum name="reflectance" value="400:0.1, 700:0.2"/></bsdf></shape>')
    refl = sp.bsdf(sp.mesh(0)["bsdf"])["reflectance"]
    assert 0.1 < min(refl) and max(refl) < 0.2 and refl[0] > refl[2]                           # a ramp rising towards red
    with pytest.raises(pkg.AmvptError, match="invalid spectrum"):
        load(sensor + '<shape type="cube"><bsdf type="diffuse"><spectrum name="reflectance" value="400:0.1:3, 700:0.2"/></bsdf></shape>')
    with pytest.raises(pkg.AmvptError, match="increasing order"):
        load(sensor + '<shape type="cube"><bsdf type="diffuse"><spectrum name="reflectance" value="700:0.1, 400:0.2"/></bsdf></shape>')
    with pytest.raises(pkg.AmvptError, match="divisible"):                      # grid.cpp:100-101
        load('<sensor type="grid"><integer name="grid_x" value="3"/><integer name="res_x" value="16"/><integer name="res_y" value="8"/>'
             '<wrap type="any"><string name="wrap_class" value="sensor"/><string name="wrap_type" value="perspective"/></wrap>'
             '<wrap type="any"><string name="wrap_class" value="film"/><string name="wrap_type" value="hdrfilm"/><rfilter type="box"/></wrap>'
             '</sensor>' + shape)


def test_ply_ascii_and_big_endian(pkg, scenes, tmp_path):
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0.5]], np.float32); faces = np.array([[0, 1, 2], [1, 3, 2]], np.uint32)
    scenes.write_ply(str(tmp_path / "a.ply"), pos, faces, binary=False)
    scenes.write_ply(str(tmp_path / "b.ply"), pos, faces, binary=True)
    # big-endian copy of b.ply
    raw = (tmp_path / "b.ply").read_bytes(); hdr, body = raw.split(b"end_header\n")
    v = np.frombuffer(body[:48], "<f4").astype(">f4").tobytes()
    rec = np.frombuffer(body[48:], np.dtype([("n", "u1"), ("i", "<i4", 3)])); be = np.zeros(2, np.dtype([("n", "u1"), ("i", ">i4", 3)])); be["n"] = 3; be["i"] = rec["i"]
    (tmp_path / "c.ply").write_bytes(hdr.replace(b"binary_little_endian", b"binary_big_endian") + b"end_header\n" + v + be.tobytes())
    sensor = ('<sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/>'
              '<rfilter type="box"/></film></sensor>')
    out = []
    for f in "abc":
        s = pkg.load_string(f'<scene version="3.0.0">{sensor}<shape type="ply"><string name="filename" value="{f}.ply"/></shape></scene>', str(tmp_path))
        m = s.mesh(0); out.append(m)
        assert np.array_equal(m["positions"], pos) and np.array_equal(m["faces"], faces)
        assert m["normals"] is not None and np.allclose(np.linalg.norm(m["normals"], axis=1), 1, atol=1e-6)   # recomputed (ply.cpp:444-449)
    assert np.array_equal(out[0]["normals"], out[1]["normals"]) and np.array_equal(out[1]["normals"], out[2]["normals"])
    assert np.allclose(out[0]["normals"][0], [0, 0, 1])


def test_anisotropic_bsdf_on_uv_mapped_mesh_is_refused(pkg, tmp_path):
    """The reference orients the shading frame of a UV-mapped mesh along its texture coordinates (mesh.cpp:1484-1510); this path drops them, which
    is the same distribution for isotropic lobes only: the loader refuses the anisotropic case instead of rendering it differently."""
    (tmp_path / "uv.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvt 0 0\nvt 1 0\nvt 0 1\nf 1/1 2/2 3/3\n")
    (tmp_path / "plain.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    xml = """<scene version="3.0.0"><sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film></sensor>
      <shape type="obj"><string name="filename" value="$mesh"/>
        <bsdf type="roughconductor"><float name="alpha_u" value="$au"/><float name="alpha_v" value="0.3"/></bsdf>
        <emitter type="area"><rgb name="radiance" value="1"/></emitter></shape></scene>"""
    pkg.load_string(xml, base_dir=str(tmp_path), mesh="uv.obj", au="0.3")            # isotropic: fine
    pkg.load_string(xml, base_dir=str(tmp_path), mesh="plain.obj", au="0.05")        # anisotropic without texture coordinates: fine
    with pytest.raises(pkg.AmvptError, match="anisotropic roughness"):
        pkg.load_string(xml, base_dir=str(tmp_path), mesh="uv.obj", au="0.05")

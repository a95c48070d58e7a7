"""Synthetic scene descriptions for BASELINE.json's configurations, as plain dictionaries that
both the product (`build`) and the CPU oracle (tests/oracle_binding.py) consume, so both sides see identical bytes.

S1: the Cornell box from the published Cornell measurements (quads split into two triangles; 32 triangles, diffuse
    BSDFs with the reflectances of the reference's resources/data/scenes/cbox/fragments/bsdfs-rgb.xml, one area light
    of radiance (18.387, 10.9873, 2.75357) lowered by 0.5 like the reference's shapes.xml), seen by a `grid` sensor
    whose sub-camera is the cbox perspective camera (fov 39.3077 smaller axis, near 10, far 2800, origin (278,273,-800)).
S2: procedural height field z = 0.15 sin(7x) cos(5y) + 0.05 sin(31x + 17y) over [-1,1]^2, n x n quads (n = 708 gives
    1 002 528 triangles), four meshes with roughconductor / roughdielectric / roughplastic / diffuse BSDFs, inside a
    12-triangle diffuse box with a 2-triangle area light of radiance 10.
"""
import os

import numpy as np

CBOX_QUADS = {
    "floor": [[(552.8, 0, 0), (0, 0, 0), (0, 0, 559.2), (549.6, 0, 559.2)]],
    "ceiling": [[(556, 548.8, 0), (556, 548.8, 559.2), (0, 548.8, 559.2), (0, 548.8, 0)]],
    "back": [[(549.6, 0, 559.2), (0, 0, 559.2), (0, 548.8, 559.2), (556, 548.8, 559.2)]],
    "greenwall": [[(0, 0, 559.2), (0, 0, 0), (0, 548.8, 0), (0, 548.8, 559.2)]],
    "redwall": [[(552.8, 0, 0), (549.6, 0, 559.2), (556, 548.8, 559.2), (556, 548.8, 0)]],
    "luminaire": [[(343, 548.8, 227), (343, 548.8, 332), (213, 548.8, 332), (213, 548.8, 227)]],
    "smallbox": [[(130, 165, 65), (82, 165, 225), (240, 165, 272), (290, 165, 114)],
                 [(290, 0, 114), (290, 165, 114), (240, 165, 272), (240, 0, 272)],
                 [(130, 0, 65), (130, 165, 65), (290, 165, 114), (290, 0, 114)],
                 [(82, 0, 225), (82, 165, 225), (130, 165, 65), (130, 0, 65)],
                 [(240, 0, 272), (240, 165, 272), (82, 165, 225), (82, 0, 225)]],
    "largebox": [[(423, 330, 247), (265, 330, 296), (314, 330, 456), (472, 330, 406)],
                 [(423, 0, 247), (423, 330, 247), (472, 330, 406), (472, 0, 406)],
                 [(472, 0, 406), (472, 330, 406), (314, 330, 456), (314, 0, 456)],
                 [(314, 0, 456), (314, 330, 456), (265, 330, 296), (265, 0, 296)],
                 [(265, 0, 296), (265, 330, 296), (423, 330, 247), (423, 0, 247)]],
}
CBOX_BSDFS = {"box": (0.45, 0.30, 0.90), "white": (0.885809, 0.698859, 0.666422), "red": (0.570068, 0.0430135, 0.0443706),
              "green": (0.105421, 0.37798, 0.076425), "light": (0.936461, 0.740433, 0.705267)}
CBOX_SHAPES = [("luminaire", "light"), ("floor", "white"), ("ceiling", "white"), ("back", "white"), ("greenwall", "green"),
               ("redwall", "red"), ("smallbox", "box"), ("largebox", "box")]
CBOX_RADIANCE = (18.387, 10.9873, 2.75357)


def quads_to_mesh(quads, offset=(0.0, 0.0, 0.0)):
    pos, faces = [], []
    for q in quads:
        b = len(pos)
        pos += [tuple(np.float32(c) + np.float32(o) for c, o in zip(v, offset)) for v in q]
        faces += [(b, b + 1, b + 2), (b, b + 2, b + 3)]
    return np.asarray(pos, np.float32), np.asarray(faces, np.uint32)


def look_at(origin, target, up):
    """Transform4d::look_at (reference include/mitsuba/core/transform.h:269-300), double precision."""
    o, t, u = (np.asarray(v, np.float64) for v in (origin, target, up))
    d = (t - o) / np.linalg.norm(t - o); left = np.cross(u, d); left /= np.linalg.norm(left); nu = np.cross(d, left)
    m = np.eye(4); m[:3, 0] = left; m[:3, 1] = nu; m[:3, 2] = d; m[:3, 3] = o
    return m


def cornell_box(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, rr_depth=5, spp_pass_lim=16, sample_count=64,
                sampler_seed=0, bsdf_overrides=None, thinlens=False, aperture_radius=0.0, cone_deg=None):
    bsdfs, ids = [], {}
    for name, refl in CBOX_BSDFS.items():
        kw = dict(type="diffuse", reflectance=refl)
        if bsdf_overrides and name in bsdf_overrides:
            kw = dict(bsdf_overrides[name])
        ids[name] = len(bsdfs); bsdfs.append(kw)
    meshes = []
    for shape, bsdf in CBOX_SHAPES:
        pos, faces = quads_to_mesh(CBOX_QUADS[shape], (0.0, -0.5, 0.0) if shape == "luminaire" else (0.0, 0.0, 0.0))
        meshes.append(dict(name=shape, positions=pos, faces=faces, normals=None, bsdf=ids[bsdf],
                           radiance=CBOX_RADIANCE if shape == "luminaire" else None, flip_normals=False))
    sensor = dict(grid_x=grid_x, grid_y=grid_y, res_x=grid_x * view_res, res_y=grid_y * view_res,
                  to_world=look_at((278, 273, -800), (278, 273, -799), (0, 1, 0)), fov=39.3077, fov_axis="smaller", near_clip=10.0,
                  far_clip=2800.0, focus_distance=1000.0, thinlens=thinlens, aperture_radius=aperture_radius, reverse_x=False,
                  reverse_y=True, cone_deg=cone_deg, cam_dir=(1.0, 0.0, 0.0), cam_dist=cam_dist, cam_center=True, cam_off=(0.0, 0.0, 0.0))
    integrator = dict(max_depth=max_depth, rr_depth=rr_depth, hide_emitters=False, spp_pass_lim=spp_pass_lim,
                      sampler_seed=sampler_seed, sample_count=sample_count)
    return dict(name=f"cbox_{grid_x}x{grid_y}_{view_res}", bsdfs=bsdfs, meshes=meshes, sensor=sensor, integrator=integrator)


CBOX_XML = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes", "scenes", "cbox", "cbox-rgb.xml")


def cornell_box_ref(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, rr_depth=5, spp_pass_lim=16, sample_count=64, sampler_seed=0):
    """S1 of SURVEY.md sec.8d: the geometry and materials of the reference's own resources/data/scenes/cbox/cbox-rgb.xml (8 OBJ meshes, 38
    triangles, 5 diffuse BSDFs, one area emitter; a byte-for-byte copy lives under tests/golden/scenes), read by the product's XML loader, with
    the file's perspective camera replaced by a `grid` of such cameras on a line (the reference has no scene that uses its grid sensor)."""
    import importlib
    pkg = importlib.import_module("mitsuba3-amvpt_b200")
    s = pkg.load_file(CBOX_XML, res=view_res, spp=sample_count, max_depth=max_depth)
    c = s.counts()
    bsdfs = [s.bsdf(i) for i in range(c["n_bsdfs"])]
    for b in bsdfs:
        b["alpha_u"], b["alpha_v"] = float(b["alpha_u"]), float(b["alpha_v"])
    meshes = []
    for i in range(c["n_meshes"]):
        m = s.mesh(i); meshes.append(dict(name=f"cbox{i}", positions=m["positions"], faces=m["faces"], normals=m["normals"], bsdf=m["bsdf"],
                                          radiance=m["radiance"], flip_normals=m["flip_normals"]))
    sensor = dict(grid_x=grid_x, grid_y=grid_y, res_x=grid_x * view_res, res_y=grid_y * view_res,
                  to_world=look_at((278, 273, -800), (278, 273, -799), (0, 1, 0)), fov=39.3077, fov_axis="smaller", near_clip=10.0,
                  far_clip=2800.0, focus_distance=1000.0, thinlens=False, aperture_radius=0.0, reverse_x=False,
                  reverse_y=True, cone_deg=None, cam_dir=(1.0, 0.0, 0.0), cam_dist=cam_dist, cam_center=True, cam_off=(0.0, 0.0, 0.0))
    integrator = dict(max_depth=max_depth, rr_depth=rr_depth, hide_emitters=False, spp_pass_lim=spp_pass_lim, sampler_seed=sampler_seed, sample_count=sample_count)
    return dict(name=f"cboxref_{grid_x}x{grid_y}_{view_res}", bsdfs=bsdfs, meshes=meshes, sensor=sensor, integrator=integrator)


def height_field(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, rr_depth=5, spp_pass_lim=16, sample_count=64,
                 sampler_seed=0, smooth_normals=True):
    """S2 of SURVEY.md sec.8d.  Vertex data are float32 functions of integer grid indices (no RNG)."""
    bsdfs = [dict(type="roughconductor", distribution="ggx", alpha=0.15, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)),
             dict(type="roughdielectric", distribution="beckmann", alpha=0.1, int_ior=1.5046, ext_ior=1.000277),
             dict(type="roughplastic", distribution="beckmann", alpha=0.1, reflectance=(0.5, 0.5, 0.5)),
             dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)),
             dict(type="diffuse", reflectance=(0.7, 0.7, 0.7)),      # enclosure
             dict(type="diffuse", reflectance=(0.0, 0.0, 0.0))]      # light
    meshes = []
    h = n // 2
    i = np.arange(n + 1, dtype=np.float64)
    xs = -1.0 + 2.0 * i / n
    X, Y = np.meshgrid(xs, xs, indexing="xy")             # vertex (iy, ix)
    Z = 0.15 * np.sin(7 * X) * np.cos(5 * Y) + 0.05 * np.sin(31 * X + 17 * Y)
    dZdx = 0.15 * 7 * np.cos(7 * X) * np.cos(5 * Y) + 0.05 * 31 * np.cos(31 * X + 17 * Y)
    dZdy = -0.15 * 5 * np.sin(7 * X) * np.sin(5 * Y) + 0.05 * 17 * np.cos(31 * X + 17 * Y)
    N = np.stack([-dZdx, -dZdy, np.ones_like(Z)], -1); N /= np.linalg.norm(N, axis=-1, keepdims=True)
    P = np.stack([X, Y, Z], -1)
    quads = [(0, h, 0, h), (h, n, 0, h), (0, h, h, n), (h, n, h, n)]     # (x0, x1, y0, y1) in quad indices
    for q, (x0, x1, y0, y1) in enumerate(quads):
        px = P[y0:y1 + 1, x0:x1 + 1].reshape(-1, 3).astype(np.float32)
        nx = N[y0:y1 + 1, x0:x1 + 1].reshape(-1, 3).astype(np.float32)
        w = x1 - x0 + 1
        jy, jx = np.meshgrid(np.arange(y1 - y0), np.arange(x1 - x0), indexing="ij")
        v00 = (jy * w + jx).reshape(-1); v10 = v00 + 1; v01 = v00 + w; v11 = v01 + 1
        faces = np.empty((2 * len(v00), 3), np.uint32)
        faces[0::2] = np.stack([v00, v10, v11], -1); faces[1::2] = np.stack([v00, v11, v01], -1)      # normal +z
        meshes.append(dict(name=f"field{q}", positions=px, faces=faces, normals=nx if smooth_normals else None, bsdf=q,
                           radiance=None, flip_normals=False))
    # enclosure [-1.6,1.6]^2 x [-0.6,2.4], inward-facing normals
    a, z0, z1 = 1.6, -0.6, 2.4
    c = [(-a, -a, z0), (a, -a, z0), (a, a, z0), (-a, a, z0), (-a, -a, z1), (a, -a, z1), (a, a, z1), (-a, a, z1)]
    box_quads = [[c[0], c[1], c[2], c[3]], [c[4], c[7], c[6], c[5]], [c[0], c[4], c[5], c[1]], [c[1], c[5], c[6], c[2]],
                 [c[2], c[6], c[7], c[3]], [c[3], c[7], c[4], c[0]]]
    pos, faces = quads_to_mesh(box_quads)
    meshes.append(dict(name="enclosure", positions=pos, faces=faces, normals=None, bsdf=4, radiance=None, flip_normals=False))
    l, zl = 0.5, 2.39
    pos, faces = quads_to_mesh([[(-l, -l, zl), (-l, l, zl), (l, l, zl), (l, -l, zl)]])       # normal -z
    meshes.append(dict(name="light", positions=pos, faces=faces, normals=None, bsdf=5, radiance=(10.0, 10.0, 10.0), flip_normals=False))
    sensor = dict(grid_x=grid_x, grid_y=grid_y, res_x=grid_x * view_res, res_y=grid_y * view_res,
                  to_world=look_at((0.0, -1.45, 1.6), (0.0, 0.0, 0.0), (0, 0, 1)), fov=60.0, fov_axis="x", near_clip=0.01, far_clip=100.0,
                  focus_distance=2.0, thinlens=False, aperture_radius=0.0, reverse_x=False, reverse_y=True, cone_deg=None,
                  cam_dir=(1.0, 0.0, 0.0), cam_dist=1.0, cam_center=True, cam_off=(0.0, 0.0, 0.0))
    integrator = dict(max_depth=max_depth, rr_depth=rr_depth, hide_emitters=False, spp_pass_lim=spp_pass_lim,
                      sampler_seed=sampler_seed, sample_count=sample_count)
    return dict(name=f"hfield{n}_{grid_x}x{grid_y}_{view_res}", bsdfs=bsdfs, meshes=meshes, sensor=sensor, integrator=integrator)


def build(desc, device=0, commit=True):
    """Instantiate a description in the product library."""
    from . import Scene
    s = Scene()
    for b in desc["bsdfs"]:
        s.add_bsdf(**b)
    for m in desc["meshes"]:
        s.add_mesh(m["positions"], m["faces"], m["bsdf"], normals=m["normals"], radiance=m["radiance"], flip_normals=m["flip_normals"])
    s.set_sensor_grid(**desc["sensor"])
    s.set_integrator(**desc["integrator"])
    if desc.get("rfilter"):
        s.set_rfilter(*desc["rfilter"])
    if desc.get("reuse"):
        r = dict(desc["reuse"]); dbg = r.pop("debug", False)
        s.set_reuse(**r)
        if dbg:
            s.set_reuse_debug(True)
    if commit:
        s.commit(device)
    return s


# ---------------------------------------------------------------------------------------------------------------------
# Export of a description as a Mitsuba 3 scene (XML version 3.0.0 + OBJ / binary PLY meshes): the same file loads in the
# reference (`mitsuba -m scalar_rgb scene.xml`) and in this library (`load_file`), which is how the synthetic workloads
# are handed to an unmodified reference build, and how tests check the XML loader against the programmatic path.
def _fmt(v):
    return repr(float(v))


def _rgb(name, v):
    v = np.broadcast_to(np.asarray(v, np.float64), (3,))
    return f'<rgb name="{name}" value="{_fmt(v[0])}, {_fmt(v[1])}, {_fmt(v[2])}"/>'


def write_obj(path, positions, faces, normals=None):
    with open(path, "w") as f:
        f.write("# written by mitsuba3-amvpt_b200.scenes.write_obj\n")
        for p in positions:
            f.write("v %s %s %s\n" % tuple(repr(float(np.float32(c))) for c in p))
        if normals is not None:
            for n in normals:
                f.write("vn %s %s %s\n" % tuple(repr(float(np.float32(c))) for c in n))
            for a, b, c in np.asarray(faces, np.int64) + 1:
                f.write(f"f {a}//{a} {b}//{b} {c}//{c}\n")
        else:
            for a, b, c in np.asarray(faces, np.int64) + 1:
                f.write(f"f {a} {b} {c}\n")


def write_ply(path, positions, faces, normals=None, binary=True):
    positions = np.ascontiguousarray(positions, np.float32); faces = np.ascontiguousarray(faces, np.uint32)
    props = ["x", "y", "z"] + (["nx", "ny", "nz"] if normals is not None else [])
    hdr = ["ply", "format binary_little_endian 1.0" if binary else "format ascii 1.0", "comment written by mitsuba3-amvpt_b200",
           f"element vertex {len(positions)}"] + [f"property float {p}" for p in props] + \
          [f"element face {len(faces)}", "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode())
        vtx = positions if normals is None else np.concatenate([positions, np.ascontiguousarray(normals, np.float32)], 1)
        if binary:
            f.write(np.ascontiguousarray(vtx, "<f4").tobytes())
            rec = np.zeros(len(faces), dtype=np.dtype([("n", "u1"), ("i", "<i4", 3)])); rec["n"] = 3; rec["i"] = faces
            f.write(rec.tobytes())
        else:
            for row in vtx:
                f.write((" ".join(repr(float(c)) for c in row) + "\n").encode())
            for a, b, c in faces:
                f.write(f"3 {a} {b} {c}\n".encode())


def _rfilter_xml(rf):
    if not rf or rf[0] == "box":
        return '<rfilter type="box"/>'
    if rf[0] == "catmullrom":
        return '<rfilter type="catmullrom"/>'
    if rf[0] == "lanczos":
        return f'<rfilter type="lanczos"><integer name="lobes" value="{int(rf[1])}"/></rfilter>'
    if rf[0] == "mitchell":
        return f'<rfilter type="mitchell"><float name="B" value="{_fmt(rf[1])}"/><float name="C" value="{_fmt(rf[2])}"/></rfilter>'
    key = "stddev" if rf[0] == "gaussian" else "radius"
    return f'<rfilter type="{rf[0]}"><float name="{key}" value="{_fmt(rf[1])}"/></rfilter>'


def _bsdf_xml(b, ident):
    t = b["type"]; lines = []
    if t == "diffuse":
        lines.append(_rgb("reflectance", b.get("reflectance", 0.5)))
    elif t in ("conductor", "roughconductor"):
        lines += [_rgb("eta", b.get("eta", 0.0)), _rgb("k", b.get("k", 1.0))]
        if "specular_reflectance" in b:
            lines.append(_rgb("specular_reflectance", b["specular_reflectance"]))
    elif t in ("dielectric", "roughdielectric"):
        lines += [f'<float name="int_ior" value="{_fmt(b.get("int_ior", 1.5046))}"/>', f'<float name="ext_ior" value="{_fmt(b.get("ext_ior", 1.000277))}"/>']
        for k in ("specular_reflectance", "specular_transmittance"):
            if k in b:
                lines.append(_rgb(k, b[k]))
    elif t == "roughplastic":
        lines += [f'<float name="int_ior" value="{_fmt(b.get("int_ior", 1.5046))}"/>', f'<float name="ext_ior" value="{_fmt(b.get("ext_ior", 1.000277))}"/>',
                  _rgb("diffuse_reflectance", b.get("reflectance", 0.5)), f'<boolean name="nonlinear" value="{str(bool(b.get("nonlinear", False))).lower()}"/>']
        if "specular_reflectance" in b:
            lines.append(_rgb("specular_reflectance", b["specular_reflectance"]))
    else:
        raise KeyError(t)
    if t.startswith("rough"):
        lines.append(f'<string name="distribution" value="{b.get("distribution", "beckmann")}"/>')
        if b.get("alpha_u") is not None and t != "roughplastic":
            lines += [f'<float name="alpha_u" value="{_fmt(b["alpha_u"])}"/>', f'<float name="alpha_v" value="{_fmt(b["alpha_v"])}"/>']
        else:
            lines.append(f'<float name="alpha" value="{_fmt(b.get("alpha", 0.1))}"/>')
    inner = f'<bsdf type="{t}"' + ("" if b.get("twosided") else f' id="{ident}"') + ">\n            " + "\n            ".join(lines) + "\n        </bsdf>"
    if b.get("twosided"):
        return f'<bsdf type="twosided" id="{ident}">\n        {inner}\n    </bsdf>'
    return inner


def export_xml(desc, directory, mesh_format="obj", integrator="mvpath", filename="scene.xml"):
    """Writes `directory/scene.xml` + `directory/meshes/*`; returns the XML path."""
    import os
    os.makedirs(os.path.join(directory, "meshes"), exist_ok=True)
    s, it = desc["sensor"], desc["integrator"]
    gx, gy = s["grid_x"], s["grid_y"]
    out = ['<scene version="3.0.0">', f'    <default name="spp" value="{it["sample_count"]}"/>',
           f'    <integrator type="{integrator}">', f'        <integer name="max_depth" value="{it["max_depth"]}"/>',
           f'        <integer name="rr_depth" value="{it["rr_depth"]}"/>']
    if integrator == "mvpath":
        out.append(f'        <integer name="spp_pass_lim" value="{it["spp_pass_lim"]}"/>')
        ru = desc.get("reuse")
        if ru:
            for key in ("sa_reuse", "sa_mis", "fast_mis"):
                out.append(f'        <boolean name="{key}" value="{str(bool(ru.get(key, key == "sa_reuse"))).lower()}"/>')
            out.append(f'        <integer name="reuse_count" value="{int(ru.get("reuse_count", 0))}"/>')
    out.append("    </integrator>")
    out += ['    <wrap type="any" id="m_film">', '        <string name="wrap_class" value="film"/>', '        <string name="wrap_type" value="hdrfilm"/>',
            f'        <integer name="width" value="{s["res_x"] // gx}"/>', f'        <integer name="height" value="{s["res_y"] // gy}"/>',
            '        <string name="pixel_format" value="rgb"/>', '        ' + _rfilter_xml(desc.get("rfilter")), "    </wrap>",
            '    <wrap type="any" id="m_sampler">', '        <string name="wrap_class" value="sampler"/>', '        <string name="wrap_type" value="independent"/>',
            '        <integer name="sample_count" value="$spp"/>', f'        <integer name="seed" value="{it["sampler_seed"]}"/>', "    </wrap>"]
    m = np.asarray(s["to_world"], np.float64)
    out += ['    <sensor type="grid">', f'        <integer name="grid_x" value="{gx}"/>', f'        <integer name="grid_y" value="{gy}"/>',
            f'        <boolean name="reverse_x" value="{str(bool(s["reverse_x"])).lower()}"/>', f'        <boolean name="reverse_y" value="{str(bool(s["reverse_y"])).lower()}"/>']
    if s.get("cone_deg") is not None:
        out.append(f'        <float name="cone_deg" value="{_fmt(s["cone_deg"])}"/>')
    else:
        d = s["cam_dir"]
        out += [f'        <vector name="cam_dir" value="{_fmt(d[0])}, {_fmt(d[1])}, {_fmt(d[2])}"/>', f'        <float name="cam_dist" value="{_fmt(s["cam_dist"])}"/>',
                f'        <boolean name="cam_center" value="{str(bool(s["cam_center"])).lower()}"/>']
    o = s.get("cam_off", (0, 0, 0))
    out += [f'        <vector name="cam_off" value="{_fmt(o[0])}, {_fmt(o[1])}, {_fmt(o[2])}"/>', f'        <float name="focus_distance" value="{_fmt(s["focus_distance"])}"/>',
            '        <transform name="to_world">', '            <matrix value="' + " ".join(repr(float(x)) for x in m.reshape(-1)) + '"/>', "        </transform>",
            '        <wrap type="any">', '            <string name="wrap_class" value="sensor"/>',
            f'            <string name="wrap_type" value="{"thinlens" if s.get("thinlens") else "perspective"}"/>',
            f'            <float name="fov" value="{_fmt(s["fov"])}"/>', f'            <string name="fov_axis" value="{s["fov_axis"]}"/>',
            f'            <float name="near_clip" value="{_fmt(s["near_clip"])}"/>', f'            <float name="far_clip" value="{_fmt(s["far_clip"])}"/>']
    if s.get("thinlens"):
        out.append(f'            <float name="aperture_radius" value="{_fmt(s["aperture_radius"])}"/>')
    out += ['            <ref id="m_film"/>', '            <ref id="m_sampler"/>', "        </wrap>", '        <ref id="m_film"/>', '        <ref id="m_sampler"/>', "    </sensor>"]
    for i, b in enumerate(desc["bsdfs"]):
        out.append("    " + _bsdf_xml(b, f"bsdf{i}"))
    for i, me in enumerate(desc["meshes"]):
        fn = f"meshes/{i:02d}_{me['name']}.{mesh_format}"
        (write_obj if mesh_format == "obj" else write_ply)(os.path.join(directory, fn), me["positions"], me["faces"], me["normals"])
        out += [f'    <shape type="{mesh_format}">', f'        <string name="filename" value="{fn}"/>']
        if me["normals"] is None:
            out.append('        <boolean name="face_normals" value="true"/>')
        if me["flip_normals"]:
            out.append('        <boolean name="flip_normals" value="true"/>')
        out.append(f'        <ref id="bsdf{me["bsdf"]}"/>')
        if me["radiance"] is not None:
            out += ['        <emitter type="area">', "            " + _rgb("radiance", me["radiance"]), "        </emitter>"]
        out.append("    </shape>")
    out.append("</scene>")
    path = os.path.join(directory, filename)
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    return path

"""Debug aid: the CLI's interlace path against the API's on the same scene (one process)."""
import importlib, sys, tempfile, numpy as np
sys.path.insert(0, ".")
pkg = importlib.import_module("mitsuba3-amvpt_b200"); scenes = importlib.import_module("mitsuba3-amvpt_b200.scenes")
P = dict(center=-0.5, focus=-0.05, pitch=354.548, tilt=-0.114, subp=0.00013, view=(1.0, 1.0), grid=(2, 2))
tmp = tempfile.mkdtemp()
desc = scenes.cornell_box(grid_x=2, grid_y=2, view_res=32, sample_count=4)
xml = scenes.export_xml(desc, tmp)
s = scenes.build(desc, device=0); img = s.render(seed=0, spp=4)
A = s.to_srgb8(); I1 = s.interlace(P, 96, 64, 3); I1b = s.interlace(P, 96, 64, 3, src=A)
s2 = pkg.load_file(str(xml)).commit(0); s2.render(seed=0, spp=4, develop=False, scheduler="auto")
print("tiles", [tuple(int(x) for x in r) for r in s2.shard_tiles(0, 1)])
B0 = s2.to_srgb8(); J0 = s2.interlace(P, 96, 64, 3)
s2.write_pfm(tmp + "/a.pfm"); s2.write_pfm(tmp + "/v.pfm", view=1)
s2.to_srgb8(fetch=False); J1 = s2.interlace(P, 96, 64, 3)
B = s2.to_srgb8(); J2 = s2.interlace(P, 96, 64, 3); J3 = s2.interlace(P, 96, 64, 3, src=B)
for n, a, b in [("A/B0", A, B0), ("A/B", A, B), ("I1/I1b", I1, I1b), ("I1/J0", I1, J0), ("I1/J1", I1, J1), ("I1/J2", I1, J2), ("I1/J3", I1, J3)]:
    print(n, "equal" if np.array_equal(a, b) else f"DIFF {np.count_nonzero(a != b)} of {a.size}")
print("I1 row0", I1[0, :4].tolist(), "J1 row0", J1[0, :4].tolist())

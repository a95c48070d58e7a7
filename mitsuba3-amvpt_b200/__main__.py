"""Command line of the render path, mirroring the options of the reference's `mitsuba` executable that reach this path
(src/mitsuba/mitsuba_render.cpp:165-355: -D -s -n -o -m -t -v; render_internal :105-133 = integrator->render(scene, sensor, seed 0,
spp, develop = false) followed by film->write(filename)).

    python -m mitsuba3-amvpt_b200 scene.xml [-n SPP] [-o OUT] [-D key=value ...] [-s 0] [--views] [--interlace W H]

The output follows the extension of `-o`: `.exr` (the default, as the reference; float16 RGB unless `--component-format float32`,
written as an uncompressed scan-line OpenEXR file) or `.pfm`.  `--views` also writes one file per view of the quilt
(`foo_v<k>.<ext>`); `--interlace W H` writes the lenticular interlace of the quilt as a binary PPM."""
import argparse
import os
import sys
import time


def parse_args(argv):
    ap = argparse.ArgumentParser(prog="python -m mitsuba3-amvpt_b200", description=__doc__.split("\n\n")[0])
    ap.add_argument("scenes", nargs="+", help="scene XML files (version 3.0.0 subset of SURVEY.md sec.8b)")
    ap.add_argument("-D", "--define", action="append", default=[], metavar="key=value", help="substitute $key in the scene file")
    ap.add_argument("-s", "--sensor", type=int, default=0, help="sensor index (the grid / batch sensor is sensor 0)")
    ap.add_argument("-n", "--nspp", type=int, default=0, help="samples per pixel (0 = the sampler's sample_count)")
    ap.add_argument("-o", "--output", default=None, help="output file, .exr or .pfm (default: scene name with .exr)")
    ap.add_argument("-m", "--mode", default="cuda_rgb", help="accepted for compatibility; this build is the CUDA RGB path")
    ap.add_argument("-t", "--threads", type=int, default=0, help="accepted for compatibility (host threads are not used by the render loop)")
    ap.add_argument("-v", "--verbose", action="count", default=0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--scheduler", default="auto")
    ap.add_argument("--views", action="store_true", help="also write one file per view of the quilt")
    ap.add_argument("--component-format", default="float16", choices=("float16", "float32"), help="OpenEXR component format (hdrfilm.cpp:146-147)")
    ap.add_argument("--interlace", nargs=2, type=int, metavar=("W", "H"), help="also write the lenticular interlace (default LKG preset, the scene's grid) as PPM")
    a = ap.parse_args(argv)
    defines = {}
    for d in a.define:
        if "=" not in d:
            ap.error(f"-D expects key=value, got {d!r}")          # mitsuba_render.cpp:238-244
        k, v = d.split("=", 1); defines[k] = v
    if a.mode not in ("cuda_rgb", "scalar_rgb", "llvm_rgb"):
        ap.error(f"mode {a.mode!r}: only the RGB variants exist on this path")
    a.define = defines
    return a


def output_path(scene, output, n_scenes):
    base = output if (output and n_scenes == 1) else os.path.splitext(scene)[0]
    root, ext = os.path.splitext(base)
    ext = ext.lower()
    if ext == ".pfm":
        return root + ".pfm", False
    return root + ".exr", ext not in ("", ".exr")                   # film->write replaces any other extension (hdrfilm.cpp:432-440)


def main(argv=None):
    a = parse_args(sys.argv[1:] if argv is None else argv)
    from . import load_file
    for scene in a.scenes:
        s = load_file(scene, **a.define).commit(a.device)
        if a.sensor != 0:
            raise SystemExit("Specified sensor index is out of bounds!")       # mitsuba_render.cpp:112-113
        out, note = output_path(scene, a.output, len(a.scenes))
        t0 = time.perf_counter()
        s.render(seed=a.seed, spp=a.nspp, develop=False, scheduler=a.scheduler)
        st = s.last_stats
        print(f"Rendering finished. (took {st['seconds'] * 1e3:.1f} ms on the device, {st['paths'] / max(st['seconds'], 1e-12) / 1e6:.1f} Mpaths/s, "
              f"{st['spp_done']} spp, {time.perf_counter() - t0:.2f} s wall)")
        root, ext = os.path.splitext(out)
        write = (lambda p, view=-1: s.write_pfm(p, view=view)) if ext == ".pfm" else (lambda p, view=-1: s.write_exr(p, view=view, component_format=a.component_format))
        write(out)
        print(f"Writing \"{out}\"" + (" (extension replaced by .exr)" if note else ""))
        if a.views:
            for *_, v in s.view_tiles():
                write(f"{root}_v{int(v)}{ext}", view=int(v))
        if a.interlace:
            s.to_srgb8(fetch=False); tiles = s.view_tiles()           # the grid of the quilt from its view rectangles
            gx, gy = len({int(r[0]) for r in tiles}), len({int(r[1]) for r in tiles})
            img = s.interlace(dict(center=-0.5, focus=-0.05, pitch=354.548, tilt=-0.114, subp=0.00013, view=(1.0, 1.0), grid=(gx, gy)), a.interlace[0], a.interlace[1], 3)
            with open(os.path.splitext(out)[0] + "_interlaced.ppm", "wb") as f:
                f.write(b"P6\n%d %d\n255\n" % (a.interlace[0], a.interlace[1])); f.write(img.tobytes())
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""Builds libamvpt_b200.so in-tree with nvcc for sm_100a (no JIT cache, the .so travels with the repo snapshot)."""
import os, subprocess, sys, hashlib

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("AMVPT_OUT", os.path.join(HERE, "libamvpt_b200.so"))
OBJ = os.environ.get("AMVPT_OBJ", os.path.join(HERE, "build"))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# canonical arithmetic: no FMA contraction on device (-fmad=false) or host (-ffp-contract=off); FMA only where written
COMMON = os.environ.get("AMVPT_CFLAGS", "").split() + ["-std=c++17", "-O3", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC,-mfma,-ffp-contract=off,-fno-fast-math,-pthread"]
SOURCES = ["kernels.cu", "wavefront.cu", "reuse.cu", "display.cu", "noalloc.cu", "scene.cpp", "emitters.cpp", "imageio.cpp", "bvh.cpp", "render.cpp", "comm.cpp", "xml.cpp", "meshio.cpp", "capi.cpp"]


def _stamp(paths):
    h = hashlib.sha1()
    for p in sorted(paths):
        h.update(p.encode()); h.update(str(os.path.getmtime(p)).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))] + [os.path.join(HERE, "..", "include", "amvpt.h")]
    hstamp = _stamp(headers)
    objs, jobs = [], []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        if not os.path.exists(sp):
            continue
        op = os.path.join(OBJ, src + ".o"); st = op + ".stamp"
        want = hstamp + _stamp([sp]) + " ".join(COMMON)
        if force or not os.path.exists(op) or not os.path.exists(st) or open(st).read() != want:
            cmd = [NVCC] + ARCH + COMMON + ["-x", "cu", "-c", sp, "-o", op]
            if verbose: print(" ".join(cmd), file=sys.stderr)
            jobs.append((subprocess.Popen(cmd), st, want))      # translation units compile side by side
        objs.append(op)
    rebuilt = bool(jobs)
    for proc, st, want in jobs:
        if proc.wait() != 0:
            raise subprocess.CalledProcessError(proc.returncode, proc.args)
        open(st, "w").write(want)
    if rebuilt or not os.path.exists(OUT):
        cmd = [NVCC] + ARCH + ["-shared", "-o", OUT] + objs + ["-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))

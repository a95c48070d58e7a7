#!/usr/bin/env python
"""Per-phase breakdown of a megakernel capture (ncu --set full --import-source on): warp instructions, thread instructions and active
lanes per instruction for the inner-node loop, the leaf loop, shading and the refill / splat code, from the source page.
Usage: ncu_phases.py report.ncu-rep out.txt.  Phases are source ranges of csrc/dscene.cuh (line numbers of the profiled commit; the
range table below names the functions, so it can be re-anchored) plus whole files; leaf arithmetic in dmath.cuh is attributed by the
SIMD efficiency of its caller where ncu's inlining info allows, else listed apart."""
import collections, csv, io, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
per_line = collections.defaultdict(lambda: [0, 0, 0]); cur = None; hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or cur is None or r[0] == "Function Name": continue
    d = dict(zip(hdr, r))
    try: ln = int(d["Line No"]); a = per_line[(cur, ln)]; a[0] += int(d["Instructions Executed"] or 0); a[1] += int(d["Thread Instructions Executed"] or 0); a[2] += int(d["# Samples"] or 0)
    except (ValueError, KeyError): pass
def find(fname, needle):
    for i, l in enumerate(open(fname), 1):
        if needle in l: return i
    raise SystemExit(f"anchor {needle!r} not found in {fname}")
D = "mitsuba3-amvpt_b200/csrc/dscene.cuh"
a_slab0 = find(D, "struct SlabRay"); a_step_end = find(D, "// Closest hit (ANY = false)"); a_int = find(D, "AM_D Hit intersect("); a_leaf = find(D, "uint32_t code = ~(uint32_t) node"); a_int_end = find(D, "// ---------------------------------------------------------------- surface interaction")
a_mt = find(D, "AM_D bool moeller_trumbore"); a_mt_end = find(D, "#define AM_SLAB_EPS")
def phase(f, ln):
    if f == "dscene.cuh":
        if a_mt <= ln < a_mt_end: return "leaf: Moeller-Trumbore"
        if a_slab0 <= ln < a_step_end: return "inner nodes: slab tests + node_step"
        if a_int <= ln < a_leaf: return "inner nodes: loop control / stack"
        if a_leaf <= ln < a_int_end: return "leaf: loop, tie rule, pop"
        return "shading: surface interaction, emitter sampling (dscene.cuh)"
    if f == "dmath.cuh": return "leaf arithmetic (dmath.cuh: fmadd / rcp / dot / cross / sqrt, called from every phase)"
    if f in ("dbsdf.cuh",): return "shading: BSDF"
    if f in ("dpath.cuh", "kernels.cu"): return "path loop, refill, splat (dpath.cuh, kernels.cu)"
    return "intrinsics / atomics headers"
agg = collections.defaultdict(lambda: [0, 0, 0])
for (f, ln), (i, t, s) in per_line.items():
    a = agg[phase(f, ln)]; a[0] += i; a[1] += t; a[2] += s
tot = sum(a[0] for a in agg.values()); tott = sum(a[1] for a in agg.values()); tots = sum(a[2] for a in agg.values())
with open(out, "w") as f:
    f.write(f"# per-phase breakdown from {rep.split('/')[-1]} (ncu --set full --import-source on): warp instructions, active lanes per instruction, stall samples\n")
    f.write(f"# total: {tot:,} warp instructions, {tott / tot:.2f} of 32 lanes per instruction, {tots:,} samples\n")
    for k, (i, t, s) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        f.write(f"{k:90s} inst {100 * i / tot:5.1f} %  lanes {t / max(i, 1):5.1f} / 32  idle lane-slots {100 * (32 * i - t) / (32 * tot - tott):5.1f} % of all idle  samples {100 * s / tots:5.1f} %\n")
    # dmath lines that are only reached from traversal (fmadd inside slab, fmsub / cross inside Moeller-Trumbore) show the caller's efficiency
    f.write("# dmath.cuh lines with their own lane counts (fmadd ~ slab tests, dot / cross / fmsub ~ Moeller-Trumbore, rcp: normalisation + triangle test)\n")
    for (fn, ln), (i, t, s) in sorted(((k, v) for k, v in per_line.items() if k[0] == "dmath.cuh"), key=lambda kv: -kv[1][0])[:8]:
        f.write(f"  dmath.cuh:{ln:4d} inst {100 * i / tot:5.2f} %  lanes {t / max(i, 1):5.1f}\n")
print(open(out).read())

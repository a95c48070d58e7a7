"""amvpt_intersect_static (csrc/noalloc.cu): the ray tap with its scratch in a file-scope device array, for callers that cannot call cudaMalloc.
The entry point has no counterpart in the reference.  Checked through the C ABI against the
existing amvpt_intersect on the same rays -- bit for bit, the tolerance of the other intersect tests -- on two calls in a row: the second call
is SMALLER than the first and runs the other query type, so stale rays / results / counters of the first call would show.
The file name keeps it last in pytest's collection order."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    d[: n // 50] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, n // 50)]          # axis-parallel rays
    maxt = np.full(n, np.finfo(np.float32).max, np.float32); maxt[::7] = rng.uniform(0, hi - lo, len(maxt[::7])).astype(np.float32)
    return o, d, maxt


def same_bits(a, b, what):
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    assert a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"{what}: amvpt_intersect_static differs from amvpt_intersect"


def test_intersect_static_equals_intersect(gpu, pkg, scenes):
    s = scenes.build(scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, sample_count=4), device=gpu)
    rng = np.random.default_rng(11)
    # call 1: a full-size batch (the stated maximum), closest hit
    o1, d1, m1 = rays(rng, s.STATIC_MAX_RAYS, -50, 600)
    p_ref, t_ref, st_ref = s.intersect(o1, d1, m1)
    p, t, st = s.intersect_static(o1, d1, m1)
    assert (p_ref != 0xffffffff).sum() > 10000
    same_bits(p, p_ref, "call 1 primitive"); same_bits(t, t_ref, "call 1 (t, u, v)")
    assert np.array_equal(st, st_ref) and st[0] > 0, f"call 1 counters {st} vs {st_ref}"
    # call 2, straight after: fewer rays, shadow-ray semantics; the counters must have been cleared and nothing of call 1 may leak
    o2, d2, m2 = rays(rng, 5000, -50, 600)
    q_ref, u_ref, su_ref = s.intersect(o2, d2, m2, any_hit=True)
    q, u, su = s.intersect_static(o2, d2, m2, any_hit=True)
    same_bits(q, q_ref, "call 2 primitive"); same_bits(u, u_ref, "call 2 (t, u, v)")
    assert np.array_equal(su, su_ref) and 0 < su[0] < st[0], f"call 2 counters {su} vs {su_ref} (call 1: {st})"
    # and the first query again gives the first answer (the array holds call 2's leftovers now)
    p3, t3, st3 = s.intersect_static(o1, d1, m1)
    same_bits(p3, p_ref, "call 3 primitive"); same_bits(t3, t_ref, "call 3 (t, u, v)"); assert np.array_equal(st3, st_ref)
    # the compressed 4-wide tree through the same entry point
    p4_ref, t4_ref, _ = s.intersect(o2, d2, m2, q4=True); p4, t4, _ = s.intersect_static(o2, d2, m2, q4=True)
    same_bits(p4, p4_ref, "q4 primitive"); same_bits(t4, t4_ref, "q4 (t, u, v)")
    # empty batch
    pe, te, se = s.intersect_static(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), np.zeros(0, np.float32))
    assert len(pe) == 0 and se[0] == 0 and se[1] == 0


def test_intersect_static_rejects_oversized_batches(gpu, pkg, scenes):
    s = scenes.build(scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, sample_count=4), device=gpu)
    n = s.STATIC_MAX_RAYS + 1
    o, d, m = rays(np.random.default_rng(12), n, -50, 600)
    with pytest.raises(pkg.AmvptError, match="exceed the static scratch"):
        s.intersect_static(o, d, m)
    # the failed call leaves the entry point usable
    p_ref, t_ref, _ = s.intersect(o[:100], d[:100], m[:100]); p, t, _ = s.intersect_static(o[:100], d[:100], m[:100])
    same_bits(p, p_ref, "after the rejected call: primitive"); same_bits(t, t_ref, "after the rejected call: (t, u, v)")

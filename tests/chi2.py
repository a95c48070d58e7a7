"""Pearson chi-square test of a sampling routine against its density -- a numpy restatement of the reference's test harness
(src/python/python/chi2.py: ChiSquareTest.tabulate_histogram :100-173, tabulate_pdf :175-237, run :239-330, SphericalDomain :416-437;
include/mitsuba/core/math.h:325-352 chi2 with pooling of low-valued cells), used the way src/bsdfs/tests/test_rough_*.py use it through
BSDFAdapter (:477-540): sample wo with three uniform numbers at a fixed wi, histogram it over [phi, cos(theta)], integrate pdf(wi, wo) over
the same cells with the trapezoid rule, compare.  TEST INFRASTRUCTURE."""
import numpy as np
from scipy.stats import chi2 as chi2_dist


def pooled_chi2(obs, exp, pool_threshold=5.0):
    """math.h:325-352 on cells sorted by increasing expected frequency."""
    order = np.argsort(exp, kind="stable"); obs = obs[order].astype(np.float64); exp = exp[order].astype(np.float64)
    keep = ~((exp == 0) & (obs == 0)); obs, exp = obs[keep], exp[keep]
    low = exp < pool_threshold
    chsq, dof, pooled_in, pooled_out = 0.0, 0, int(low.sum()), 0
    po = pe = 0.0
    for o, e in zip(obs[low], exp[low]):
        po += o; pe += e
        if pe > pool_threshold:
            chsq += (po - pe) ** 2 / pe; po = pe = 0.0; pooled_out += 1; dof += 1
    d = obs[~low] - exp[~low]
    chsq += float(np.sum(d * d / exp[~low])); dof += int((~low).sum())
    return chsq, dof - 1, pooled_in, pooled_out


def bsdf_chi2(eval_pdf_sample, wi, sample_count=1000000, res=101, ires=4, seed=0, transmissive=False):
    """eval_pdf_sample(wi[n,3], wo[n,3], s1[n], s2[n,2]) -> dict with pdf, bs_wo, weight (the tap of the oracle or of the product).
    Returns (p_value, info)."""
    rng = np.random.default_rng(seed)
    wi = np.asarray(wi, np.float32) / np.linalg.norm(wi)
    # ---- histogram of sampled directions (map_backward: phi = atan2(y, x), y = -cos(theta))
    s = rng.uniform(size=(sample_count, 3)).astype(np.float32)
    r = eval_pdf_sample(np.tile(wi, (sample_count, 1)), np.tile(np.array([0, 0, 1], np.float32), (sample_count, 1)), s[:, 0], s[:, 1:3])
    wo = r["bs_wo"].astype(np.float64); w = np.where(np.all(r["weight"] == 0, axis=1), 0.0, 1.0)
    nx, ny = 2 * res, res
    px = np.arctan2(wo[:, 1], wo[:, 0]); py = -wo[:, 2]
    ix = np.clip(np.floor((px + np.pi) / (2 * np.pi) * nx).astype(np.int64), 0, nx - 1)
    iy = np.clip(np.floor((py + 1) / 2 * ny).astype(np.int64), 0, ny - 1)
    hist = np.bincount(iy * nx + ix, weights=w, minlength=nx * ny)
    # ---- expected frequencies: trapezoid rule with ires x ires points per cell
    cx = 2 * np.pi / nx; cy = 2.0 / ny
    gx = -np.pi + (np.arange(nx)[:, None] + np.arange(ires)[None, :] / (ires - 1)) * cx          # [nx, ires]
    gy = -1.0 + (np.arange(ny)[:, None] + np.arange(ires)[None, :] / (ires - 1)) * cy            # [ny, ires]
    X = np.broadcast_to(gx[None, :, None, :], (ny, nx, ires, ires)).reshape(-1); Y = np.broadcast_to(gy[:, None, :, None], (ny, nx, ires, ires)).reshape(-1)
    cos_t = -Y; sin_t = np.sqrt(np.maximum(1 - cos_t * cos_t, 0))
    d = np.stack([np.cos(X) * sin_t, np.sin(X) * sin_t, cos_t], 1).astype(np.float32)
    n = len(d); z = np.zeros(n, np.float32)
    pdf = eval_pdf_sample(np.tile(wi, (n, 1)), d, z + 0.5, np.stack([z + 0.5, z + 0.5], 1))["pdf"].astype(np.float64)
    if not transmissive:        # BSDF::eval_pdf leaves the density unmasked where the value is masked to zero (e.g. roughplastic.cpp:424-500
        pdf[d[:, 2] * wi[2] <= 0] = 0.0      # returns `value & active, pdf`); BSDF::pdf, which the reference's adapter calls, is zero there
    tw = np.ones(ires); tw[0] = tw[-1] = 0.5
    wgt = (tw[:, None] * tw[None, :]).reshape(-1) * (cx / (ires - 1)) * (cy / (ires - 1)) * sample_count
    expct = (pdf.reshape(ny * nx, ires * ires) * wgt[None, :]).sum(1)
    chsq, dof, pin, pout = pooled_chi2(hist, expct)
    info = dict(chi2=chsq, dof=dof, pooled_in=pin, pooled_out=pout, hist_sum=hist.sum() / sample_count, pdf_sum=expct.sum() / sample_count,
                samples_in_zero_cells=float(hist[(expct == 0)].sum()))
    if dof < 1 or info["samples_in_zero_cells"] > 0:
        return 0.0, info
    return float(chi2_dist.sf(chsq, dof)), info

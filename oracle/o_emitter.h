// oracle/o_emitter.h -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the emitters that are not area lights, in the canonical arithmetic
// of o_math.h: PointLight (src/emitters/point.cpp:117-148), ConstantBackgroundEmitter (src/emitters/constant.cpp:55-157),
// EnvironmentMapEmitter (src/emitters/envmap.cpp:111-214 constructor, :308-320 eval, :419-481 sample_direction / pdf_direction, :528-586
// eval_spectrum) with its Hierarchical2D<Float, 0> sampling hierarchy (include/mitsuba/core/distr_2d.h:376-482 constructor, :490-575
// sample, :668-698 eval), warp::square_to_uniform_sphere / square_to_bilinear / interval_to_linear (include/mitsuba/core/warp.h:250-255,
// 446-453, 478-521), dr::acos / dr::atan2 (ext/drjit/include/drjit/math.h:458-597), and the scene-level dispatch of
// Scene::sample_emitter_direction (src/render/scene.cpp:295-346).  Included by amvpt_oracle.cpp inside namespace orc.
// Pinned: acos / atan2 / the bilinear warps / Hierarchical2D sample + eval bit for bit against the reference's own headers
// (oracle/_ref, tests/golden/reference_vectors.json); the emitters as a whole through the reference's golden renders
// (tests/golden/render_refs.npz, Z-test of src/render/tests/test_renders.py).
#pragma once

// ---------------------------------------------------------------- math.h:458-510, 512-597
static inline float acos_(float x) {
    const float xa = fabs_(x), x2 = x * x;
    const bool big = xa > 0.5f;
    const float x1 = 0.5f * (1.f - xa), x3 = big ? x1 : x2, x4 = big ? std::sqrt(x1) : xa;
    const float c[5] = { 1.666675242e-1f, 7.4953002686e-2f, 4.5470025998e-2f, 2.4181311049e-2f, 4.2163199048e-2f };
    float z1 = estrin(x3, c, 5);
    z1 = fmadd(z1, x3 * x4, x4);
    float z2 = z1 + z1;
    z2 = x < 0.f ? Pi - z2 : z2;
    const float z3 = (Pi * .5f) - copysign_(z1, x);
    return big ? z2 : z3;
}
static inline float safe_acos(float x) { return acos_(clip(x, -1.f, 1.f)); }        // array_router.h:1775
static inline float atan2_(float y, float x) {
    const float ax = fabs_(x), ay = fabs_(y), mn = fmin_(ay, ax), mx = fmax_(ax, ay), sm = mn / mx, z = sm * sm;
    const float c[7] = { 0.99999934166683966009f, -0.33326497518773606976f, 0.19881342388439013552f, -0.13486708938456973185f,
                         0.083863120428809689910f, -0.037006525670417265220f, 0.0078613793713198150252f };
    float t = estrin(z, c, 7) * sm;
    t = ay > ax ? Pi * .5f - t : t;
    t = x < 0.f ? Pi - t : t;
    const float r = y < 0.f ? -t : t;
    return mx != 0.f ? r : 0.f;
}
// ---------------------------------------------------------------- warp.h
static inline V3 square_to_uniform_sphere(V2 u) {                                   // :250-255 (circ :23)
    const float z = fnmadd(2.f, u.y, 1.f), r = safe_sqrt(fnmadd(z, z, 1.f));
    float s, c; sincos(2.f * Pi * u.x, &s, &c);
    return V3(r * c, r * s, z);
}
constexpr float InvFourPi = 0.07957747154594766788f, InvTwoPi = 0.15915494309189533577f;
static inline float interval_to_linear(float v0, float v1, float sample) {         // :446-453
    const float r = (v0 - safe_sqrt(lerp(sqr(v0), sqr(v1), sample))) / (v0 - v1);
    return fabs_(v0 - v1) > 1e-4f * (v0 + v1) ? r : sample;
}
static inline V2 square_to_bilinear(float v00, float v10, float v01, float v11, V2 sample, float *pdf) {   // :478-494
    const float r0 = v00 + v10, r1 = v01 + v11;
    sample.y = interval_to_linear(r0, r1, sample.y);
    const float c0 = lerp(v00, v01, sample.y), c1 = lerp(v10, v11, sample.y);
    sample.x = interval_to_linear(c0, c1, sample.x);
    *pdf = lerp(c0, c1, sample.x);
    return sample;
}
static inline float square_to_bilinear_pdf(float v00, float v10, float v01, float v11, V2 s) {           // :515-521
    return lerp(lerp(v00, v10, s.x), lerp(v01, v11, s.x), s.y);
}

// ---------------------------------------------------------------- Hierarchical2D<Float, 0>
struct Hier2D {
    struct Level { uint32_t size = 0, width = 0; std::vector<float> data;
        uint32_t index(uint32_t x, uint32_t y) const { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + (y & ~1u) * width; } };   // :757-761
    std::vector<Level> levels; float patch_size[2], inv_patch_size[2]; uint32_t max_patch[2];
    void build(const float *data, uint32_t sx, uint32_t sy) {           // :376-482, normalize = true, enable_sampling = true
        const uint32_t npx = sx - 1, npy = sy - 1, nmax = std::max(npx, npy);
        uint32_t max_level = 0; while ((1u << max_level) < nmax) ++max_level;                                 // math::log2i_ceil
        patch_size[0] = 1.f / (float) npx; patch_size[1] = 1.f / (float) npy; inv_patch_size[0] = (float) npx; inv_patch_size[1] = (float) npy;   // :234-235
        max_patch[0] = npx - 1; max_patch[1] = npy - 1;
        levels.clear(); levels.reserve(max_level + 2);
        auto add = [&](uint32_t w, uint32_t h) { Level l; l.size = w * h; l.width = w; l.data.assign(l.size, 0.f); levels.push_back(std::move(l)); };
        add(sx, sy);
        uint32_t lx = npx, ly = npy;
        for (int level = (int) max_level; level >= 0; --level) { lx += lx & 1u; ly += ly & 1u; add(lx, ly); lx >>= 1; ly >>= 1; }
        Level &l0 = levels[0], &l1 = levels[1]; const float *in = data; double sum = 0.0;
        for (uint32_t y = 0; y < npy; ++y) { for (uint32_t x = 0; x < npx; ++x) {
            const float avg = .25f * (in[0] + in[1] + in[sx] + in[sx + 1]); sum += (double) avg; l1.data[l1.index(x, y)] = avg; ++in; } ++in; }
        const float scale = (float) ((double) (npx * npy) / sum);
        for (uint32_t i = 0; i < l0.size; ++i) l0.data[i] = data[i] * scale;
        for (uint32_t i = 0; i < l1.size; ++i) l1.data[i] *= scale;
        lx = npx; ly = npy;
        for (uint32_t level = 2; level <= max_level + 1; ++level) {
            const Level &a = levels[level - 1]; Level &b = levels[level]; lx = (lx + 1u) >> 1; ly = (ly + 1u) >> 1;
            for (uint32_t y = 0; y < ly; ++y) for (uint32_t x = 0; x < lx; ++x) {
                const float *d0 = a.data.data() + a.index(x * 2, y * 2); b.data[b.index(x, y)] = d0[0] + d0[1] + d0[2] + d0[3]; }
        }
    }
    V2 sample(V2 s, float *pdf) const {                                                                        // :490-575
        s.x = clip(s.x, 0.f, 1.f); s.y = clip(s.y, 0.f, 1.f);
        uint32_t ox = 0, oy = 0;
        for (int l = (int) levels.size() - 2; l > 0; --l) {
            const Level &lv = levels[(size_t) l]; ox <<= 1; oy <<= 1;
            uint32_t oi = lv.index(ox, oy);
            const float v00 = lv.data[oi], v10 = lv.data[oi + 1], v01 = lv.data[oi + 2], v11 = lv.data[oi + 3];
            s.x = clip(s.x, 0.f, 1.f); s.y = clip(s.y, 0.f, 1.f);
            const float r0 = v00 + v10, r1 = v01 + v11;
            s.y *= r0 + r1; bool mask = s.y > r0;
            if (mask) { oy += 1u; s.y -= r0; }
            s.y /= mask ? r1 : r0;
            const float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
            s.x *= c0 + c1; mask = s.x > c0;
            if (mask) s.x -= c0;
            s.x /= mask ? c1 : c0;
            if (mask) ox += 1u;
        }
        const Level &l0 = levels[0]; const uint32_t oi = ox + oy * l0.width;
        s = square_to_bilinear(l0.data[oi], l0.data[oi + 1], l0.data[oi + l0.width], l0.data[oi + l0.width + 1], s, pdf);
        return { ((float) (int32_t) ox + s.x) * patch_size[0], ((float) (int32_t) oy + s.y) * patch_size[1] };
    }
    float eval(V2 pos) const {                                                                                 // :668-698
        pos.x = clip(pos.x, 0.f, 1.f) * inv_patch_size[0]; pos.y = clip(pos.y, 0.f, 1.f) * inv_patch_size[1];
        const uint32_t ox = std::min((uint32_t) (int32_t) pos.x, max_patch[0]), oy = std::min((uint32_t) (int32_t) pos.y, max_patch[1]);
        pos.x -= (float) (int32_t) ox; pos.y -= (float) (int32_t) oy;
        const Level &l0 = levels[0]; const uint32_t oi = ox + oy * l0.width;
        return square_to_bilinear_pdf(l0.data[oi], l0.data[oi + 1], l0.data[oi + l0.width], l0.data[oi + l0.width + 1], pos);
    }
};

// ---------------------------------------------------------------- EnvironmentMapEmitter
struct EnvMap {
    bool present = false; uint32_t res_x = 0, res_y = 0; std::vector<float> data; float scale = 1.f; M4 to_world, from_world; Hier2D warp;
    void build(const float *rgb, uint32_t w, uint32_t h, float scale_, const M4 &tw) {                        // envmap.cpp:111-214
        present = true; scale = scale_; to_world = tw; from_world = m4_inverse(tw);
        res_x = w + 1; res_y = h; data.assign((size_t) res_x * res_y * 3, 0.f);
        std::vector<float> lum((size_t) res_x * res_y);
        const float theta_scale = 1.f / (float) (h - 1) * Pi;
        for (uint32_t y = 0; y < h; ++y) {
            float sin_theta, c; sincos((float) y * theta_scale, &sin_theta, &c);
            for (uint32_t x = 0; x < w; ++x) {
                const float *p = rgb + ((size_t) y * w + x) * 3;
                float l = p[0] * 0.212671f + p[1] * 0.715160f + p[2] * 0.072169f;                              // spectrum.h:431-434
                l = fmax_(l - 0.f, 0.f);
                lum[(size_t) y * res_x + x] = l * sin_theta;
                for (int k = 0; k < 3; ++k) data[((size_t) y * res_x + x) * 3 + k] = p[k];
            }
            lum[(size_t) y * res_x + w] = lum[(size_t) y * res_x];
            for (int k = 0; k < 3; ++k) data[((size_t) y * res_x + w) * 3 + k] = data[((size_t) y * res_x) * 3 + k];
        }
        warp.build(lum.data(), res_x, res_y);
    }
    V3 px(uint32_t i) const { return V3(data[3 * (size_t) i], data[3 * (size_t) i + 1], data[3 * (size_t) i + 2]); }
    V3 eval_uv(V2 uv) const {                                                                                  // :528-586 (RGB)
        uv.x -= .5f / (float) (res_x - 1u);
        uv.x -= std::floor(uv.x); uv.y -= std::floor(uv.y);
        uv.x *= (float) (res_x - 1u); uv.y *= (float) (res_y - 1u);
        const uint32_t px_ = std::min((uint32_t) uv.x, res_x - 2u), py = std::min((uint32_t) uv.y, res_y - 2u);
        const float w1x = uv.x - (float) px_, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
        const uint32_t index = py * res_x + px_;
        const V3 v0 = fmadd(px(index), w0x, px(index + 1) * w1x), v1 = fmadd(px(index + res_x), w0x, px(index + res_x + 1) * w1x);
        return fmadd(v0, w0y, v1 * w1y) * scale;
    }
};
static inline V3 xf_vec3(const M4 &m, V3 v) {          // Transform4f::transform_affine(Vector3f), transform.h:150-160
    float r[3]; for (int i = 0; i < 3; ++i) r[i] = fmadd(m.m[i][2], v.z, fmadd(m.m[i][1], v.y, m.m[i][0] * v.x)); return V3(r[0], r[1], r[2]);
}

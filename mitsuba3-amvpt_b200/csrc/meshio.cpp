// meshio.cpp -- mesh shape plugins of the AMVPT path: `obj` (src/shapes/obj.cpp:142-411), `ply` (src/shapes/ply.cpp:160-455),
// `serialized` (src/shapes/serialized.cpp:237-404), `cube` (src/shapes/cube.cpp:101-165).  All of them end in the same Mesh record: float32 positions already in world space
// (Transform4f::transform_affine, include/mitsuba/core/transform.h:119-127), optional world-space vertex normals
// (normalize(inverse_transpose * n), :164-173) and triangle indices.  What matters for parity is the *order* of vertices
// and faces (global primitive indices feed the closest-hit tie rule and emitter sampling) and the float32 arithmetic of
// the vertex transform; both follow the reference.  Texture coordinates are parsed and dropped (no textured BSDF on the path).
#include "host.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <array>
#include <algorithm>
#include <functional>
#include <dlfcn.h>

namespace amvpt {

namespace {

struct XfF { M4 mat, inv_t; };
XfF to_float(const M4d &tw) { return { m4_from_d(tw), m4_transpose(m4_from_d(m4d_inverse(tw))) }; }
// transform.h:119-127: result = column 3, then fmadd(column i, p_i, result) for i = 0, 1, 2
V3 xf_point(const M4 &m, V3 p) {
    float r[3];
    for (int k = 0; k < 3; ++k) r[k] = fmadd(m.m[k][2], p.z, fmadd(m.m[k][1], p.y, fmadd(m.m[k][0], p.x, m.m[k][3])));
    return v3(r[0], r[1], r[2]);
}
// transform.h:164-173: inverse_transpose column 0 * n.x, then fmadd for y and z
V3 xf_normal(const M4 &it, V3 n) {
    float r[3];
    for (int k = 0; k < 3; ++k) r[k] = fmadd(it.m[k][2], n.z, fmadd(it.m[k][1], n.y, it.m[k][0] * n.x));
    return v3(r[0], r[1], r[2]);
}
bool finite3(V3 v) { return finite_(v.x) && finite_(v.y) && finite_(v.z); }

std::string read_file(const std::string &path, const char *what) {
    std::ifstream f(path, std::ios::binary);
    if (!f) AM_THROW("io: Error while loading %s file \"%s\": file not found", what, path.c_str());
    std::ostringstream ss; ss << f.rdbuf(); return ss.str();
}

// Mesh::recompute_vertex_normals (src/render/mesh.cpp:331-386): angle-weighted face normals (Thuermer & Wuethrich).
// The reference evaluates the corner angles with Dr.Jit's asin polynomial; here they go through libm in double and
// are rounded to float, so the result can differ from the reference in the last place (stated in DESIGN.md).
void recompute_vertex_normals(HMesh &m) {
    size_t nv = m.pos.size() / 3; std::vector<V3> acc(nv, v3(0.f));
    auto P = [&](uint32_t i) { return v3(m.pos[3 * i], m.pos[3 * i + 1], m.pos[3 * i + 2]); };
    auto unit_angle = [](V3 a, V3 b) {     // include/mitsuba/core/vector.h unit_angle: 2 asin(|b - a| / 2), mirrored for obtuse angles
        float d = dot(a, b);
        V3 diff = d < 0.f ? b + a : b - a;
        double t = 2.0 * std::asin(0.5 * (double) norm(diff));
        return (float) (d < 0.f ? 3.14159265358979323846 - t : t);
    };
    for (uint32_t f = 0; f < m.nfaces(); ++f) {
        uint32_t i0 = m.faces[3 * f], i1 = m.faces[3 * f + 1], i2 = m.faces[3 * f + 2];
        V3 v[3] = { P(i0), P(i1), P(i2) };
        V3 s0 = v[1] - v[0], s1 = v[2] - v[0], n = cross(s0, s1);
        float l2 = squared_norm(n);
        if (!(l2 > 0.f)) continue;
        n = n * rsqrt(l2);
        V3 a[3] = { s0, v[2] - v[1], v[0] - v[2] }, b[3] = { s1, v[0] - v[1], v[1] - v[2] };
        uint32_t idx[3] = { i0, i1, i2 };
        for (int j = 0; j < 3; ++j) { float w = unit_angle(normalize(a[j]), normalize(b[j])); acc[idx[j]] = acc[idx[j]] + n * w; }
    }
    m.nrm.resize(3 * nv);
    for (size_t i = 0; i < nv; ++i) {
        V3 n = acc[i]; float len = norm(n);
        n = len != 0.f ? n / len : v3(1.f, 0.f, 0.f);
        m.nrm[3 * i] = n.x; m.nrm[3 * i + 1] = n.y; m.nrm[3 * i + 2] = n.z;
    }
}

} // namespace

// ------------------------------------------------------------------------------------------------ OBJ
// obj.cpp:190-341: `v` lines are transformed as they are read; `vn` lines likewise (skipped with face_normals); faces
// may be polygons (fan around the first corner, :323-327) with v, v/vt, v//vn or v/vt/vn corners (negative indices are
// not supported by the reference's strtoul parser either); output vertices are the distinct (v, vt, vn) keys numbered
// by first use (:296-321).
void load_obj(const std::string &path, const M4d &to_world, bool face_normals, bool flip_tex, HMesh &out) {
    (void) flip_tex;
    std::string data = read_file(path, "OBJ");
    XfF xf = to_float(to_world);
    std::vector<V3> vertices, normals; size_t n_tex = 0;
    std::vector<std::array<uint32_t, 3>> keys;           // output vertex -> (v, vt, vn), 1-based, 0 = absent
    std::map<std::array<uint32_t, 3>, uint32_t> key_to_id;
    out.faces.clear(); out.pos.clear(); out.nrm.clear();
    const char *p = data.c_str(), *eof = p + data.size();
    auto fail = [&](const char *what, const std::string &line) { AM_THROW("io: Error while loading OBJ file \"%s\": %s \"%s\"", path.c_str(), what, line.c_str()); };
    while (p < eof) {
        const char *nl = (const char *) memchr(p, '\n', (size_t) (eof - p)); if (!nl) nl = eof;
        std::string line(p, nl); p = nl + 1;
        if (line.size() >= 1024) fail("file contains an excessively long line!", line.substr(0, 32));
        const char *cur = line.c_str();
        while (*cur == ' ' || *cur == '\t' || *cur == '\r') ++cur;
        auto is_sp = [](char c) { return c == ' ' || c == '\t'; };
        auto floats = [&](int n, float *dst) { for (int i = 0; i < n; ++i) { char *end; dst[i] = strtof(cur, &end); if (end == cur) fail("could not parse line", line); cur = end; } };
        if (cur[0] == 'v' && is_sp(cur[1])) {
            cur += 2; float v[3]; floats(3, v);
            V3 q = xf_point(xf.mat, v3(v[0], v[1], v[2]));
            if (!finite3(q)) fail("mesh contains invalid vertex position data", line);
            vertices.push_back(q);
        } else if (cur[0] == 'v' && cur[1] == 'n' && is_sp(cur[2])) {
            if (face_normals) continue;
            cur += 3; float v[3]; floats(3, v);
            V3 n = normalize(xf_normal(xf.inv_t, v3(v[0], v[1], v[2])));
            if (!finite3(n)) fail("mesh contains invalid vertex normal data", line);
            normals.push_back(n);
        } else if (cur[0] == 'v' && cur[1] == 't' && is_sp(cur[2])) {
            cur += 3; float v[2]; floats(2, v); ++n_tex;
        } else if (cur[0] == 'f' && is_sp(cur[1])) {
            cur += 2;
            uint32_t corner = 0, tri[3] = { 0, 0, 0 };
            while (true) {
                while (is_sp(*cur)) ++cur;
                std::array<uint32_t, 3> key = { 0, 0, 0 }; int slot = 0; bool any = false;
                while (true) {               // one corner: a[/b[/c]] with empty slots allowed
                    char *end; unsigned long v = strtoul(cur, &end, 10);
                    if (end != cur) { if (slot > 2) fail("could not parse line", line); key[slot] = (uint32_t) v; any = true; }
                    cur = end;
                    if (*cur == '/') { ++slot; ++cur; continue; }
                    break;
                }
                if (!any) break;
                if (key[0] == 0 || key[0] > vertices.size()) AM_THROW("io: Error while loading OBJ file \"%s\": reference to invalid vertex %u!", path.c_str(), key[0]);
                auto it = key_to_id.find(key); uint32_t id;
                if (it == key_to_id.end()) { id = (uint32_t) keys.size(); key_to_id.emplace(key, id); keys.push_back(key); } else id = it->second;
                if (corner < 3) tri[corner] = id; else { tri[1] = tri[2]; tri[2] = id; }
                if (++corner >= 3) { out.faces.push_back(tri[0]); out.faces.push_back(tri[1]); out.faces.push_back(tri[2]); }
                if (*cur != ' ' && *cur != '\t') break;
            }
        }
    }
    out.had_uv = n_tex != 0;
    size_t nv = keys.size(); out.pos.resize(3 * nv);
    bool want_normals = !face_normals;
    if (want_normals) out.nrm.assign(3 * nv, 0.f);
    for (size_t i = 0; i < nv; ++i) {
        V3 q = vertices[keys[i][0] - 1]; out.pos[3 * i] = q.x; out.pos[3 * i + 1] = q.y; out.pos[3 * i + 2] = q.z;
        if (want_normals && keys[i][2]) {
            if (keys[i][2] > normals.size()) AM_THROW("io: Error while loading OBJ file \"%s\": reference to invalid normal %u!", path.c_str(), keys[i][2]);
            V3 n = normals[keys[i][2] - 1]; out.nrm[3 * i] = n.x; out.nrm[3 * i + 1] = n.y; out.nrm[3 * i + 2] = n.z;
        }
    }
    if (want_normals && normals.empty()) recompute_vertex_normals(out);   // obj.cpp:402-407
}

// ------------------------------------------------------------------------------------------------ PLY
// ply.cpp:186-455: ascii / binary_little_endian / binary_big_endian; the `vertex` element supplies x y z [nx ny nz]
// [u v | s t] (other properties are skipped), the `face` element a list property vertex_index / vertex_indices that
// must hold exactly three indices per face.
namespace {
enum PlyType { P_I8, P_U8, P_I16, P_U16, P_I32, P_U32, P_F32, P_F64, P_NONE };
PlyType ply_type(const std::string &t) {
    static const std::pair<const char *, PlyType> names[] = {
        { "char", P_I8 }, { "int8", P_I8 }, { "uchar", P_U8 }, { "uint8", P_U8 }, { "short", P_I16 }, { "int16", P_I16 },
        { "ushort", P_U16 }, { "uint16", P_U16 }, { "int", P_I32 }, { "int32", P_I32 }, { "uint", P_U32 }, { "uint32", P_U32 },
        { "float", P_F32 }, { "float32", P_F32 }, { "double", P_F64 }, { "float64", P_F64 } };
    for (auto &n : names) if (t == n.first) return n.second;
    return P_NONE;
}
size_t ply_size(PlyType t) { static const size_t s[] = { 1, 1, 2, 2, 4, 4, 4, 8, 0 }; return s[t]; }
struct PlyProp { std::string name; PlyType type = P_NONE, count_type = P_NONE; bool list = false; };
struct PlyElem { std::string name; size_t count = 0; std::vector<PlyProp> props; };
struct PlyReader {
    const std::string &d; size_t pos; int fmt;     // 0 ascii, 1 little, 2 big
    const std::string &path;
    void need(size_t n) { if (pos + n > d.size()) AM_THROW("io: Error while loading PLY file \"%s\": unexpected end of file!", path.c_str()); }
    double read(PlyType t) {
        if (fmt == 0) {
            while (pos < d.size() && isspace((unsigned char) d[pos])) ++pos;
            need(1); char *end; double v = strtod(d.c_str() + pos, &end);
            if (end == d.c_str() + pos) AM_THROW("io: Error while loading PLY file \"%s\": could not parse value!", path.c_str());
            pos = (size_t) (end - d.c_str()); return v;
        }
        size_t n = ply_size(t); need(n); unsigned char b[8];
        for (size_t i = 0; i < n; ++i) b[i] = (unsigned char) d[pos + (fmt == 1 ? i : n - 1 - i)];
        pos += n;
        switch (t) {
            case P_I8: { int8_t v; memcpy(&v, b, 1); return v; }     case P_U8: return b[0];
            case P_I16: { int16_t v; memcpy(&v, b, 2); return v; }   case P_U16: { uint16_t v; memcpy(&v, b, 2); return v; }
            case P_I32: { int32_t v; memcpy(&v, b, 4); return v; }   case P_U32: { uint32_t v; memcpy(&v, b, 4); return v; }
            case P_F32: { float v; memcpy(&v, b, 4); return v; }     default: { double v; memcpy(&v, b, 8); return v; }
        }
    }
};
} // namespace

void load_ply(const std::string &path, const M4d &to_world, bool face_normals, HMesh &out) {
    std::string data = read_file(path, "PLY");
    XfF xf = to_float(to_world);
    auto fail = [&](const char *what) { AM_THROW("io: Error while loading PLY file \"%s\": %s!", path.c_str(), what); };
    // ---- header (ply.cpp:461-590)
    size_t pos = 0; int fmt = -1; bool seen_ply = false, done = false; std::vector<PlyElem> elems;
    while (!done) {
        size_t nl = data.find('\n', pos); if (nl == std::string::npos) fail("invalid PLY header: missing end_header");
        std::string line = data.substr(pos, nl - pos); pos = nl + 1;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ls(line); std::string tok; if (!(ls >> tok)) continue;
        if (tok == "ply") { if (seen_ply) fail("invalid PLY header: duplicate \"ply\" tag"); seen_ply = true; }
        else if (tok == "comment" || tok == "obj_info") continue;
        else if (tok == "format") {
            if (!seen_ply) fail("invalid PLY header: \"format\" before \"ply\" tag");
            std::string f, ver; ls >> f >> ver;
            fmt = f == "ascii" ? 0 : f == "binary_little_endian" ? 1 : f == "binary_big_endian" ? 2 : -1;
            if (fmt < 0) fail("invalid PLY header: invalid token after \"format\"");
            if (ver != "1.0") fail("PLY file has unknown version number");
        } else if (tok == "element") {
            PlyElem e; if (!(ls >> e.name >> e.count)) fail("invalid PLY header: missing token after \"element\""); elems.push_back(e);
        } else if (tok == "property") {
            if (fmt < 0) fail("invalid PLY header: encountered \"property\" before \"format\"");
            if (elems.empty()) fail("invalid PLY header: encountered \"property\" before \"element\"");
            PlyProp pr; std::string t; if (!(ls >> t)) fail("invalid PLY header: missing token after \"property\"");
            if (t == "list") {
                std::string ct, vt; if (!(ls >> ct >> vt >> pr.name)) fail("invalid PLY header: missing token after \"property list\"");
                pr.list = true; pr.count_type = ply_type(ct); pr.type = ply_type(vt);
                if (pr.count_type == P_NONE || pr.type == P_NONE) fail("invalid PLY header: unknown format type");
            } else {
                pr.type = ply_type(t); if (pr.type == P_NONE) fail("invalid PLY header: unknown format type");
                if (!(ls >> pr.name)) fail("invalid PLY header: missing token after \"property\"");
            }
            elems.back().props.push_back(pr);
        } else if (tok == "end_header") done = true;
        else fail("invalid PLY header: unknown token");
    }
    if (fmt < 0) fail("invalid PLY header: missing \"format\"");
    PlyReader rd{ data, pos, fmt, path };
    out.pos.clear(); out.nrm.clear(); out.faces.clear();
    bool have_normals = false, have_vertices = false;
    for (const PlyElem &e : elems) {
        if (e.name == "vertex") {
            int ix[8]; for (int &i : ix) i = -1;       // x y z nx ny nz u v
            for (size_t k = 0; k < e.props.size(); ++k) {
                const std::string &n = e.props[k].name;
                if (e.props[k].list) fail("incompatible contents -- is this a triangle mesh?");
                if (n == "x") ix[0] = (int) k; else if (n == "y") ix[1] = (int) k; else if (n == "z") ix[2] = (int) k;
                else if (n == "nx") ix[3] = (int) k; else if (n == "ny") ix[4] = (int) k; else if (n == "nz") ix[5] = (int) k;
                else if (n == "u" || n == "s" || n == "texture_u") ix[6] = (int) k; else if (n == "v" || n == "t" || n == "texture_v") ix[7] = (int) k;
            }
            out.had_uv = ix[6] >= 0 && ix[7] >= 0;        // ply.cpp:224-243 (u, v or s, t); the values are dropped
            if (ix[0] < 0 || ix[1] < 0 || ix[2] < 0) fail("incompatible contents -- is this a triangle mesh?");
            have_normals = !face_normals && ix[3] >= 0 && ix[4] >= 0 && ix[5] >= 0;
            out.pos.resize(3 * e.count); if (have_normals) out.nrm.resize(3 * e.count);
            std::vector<double> row(e.props.size());
            for (size_t i = 0; i < e.count; ++i) {
                for (size_t k = 0; k < e.props.size(); ++k) row[k] = rd.read(e.props[k].type);
                V3 q = xf_point(xf.mat, v3((float) row[ix[0]], (float) row[ix[1]], (float) row[ix[2]]));
                if (!finite3(q)) fail("mesh contains invalid vertex position data");
                out.pos[3 * i] = q.x; out.pos[3 * i + 1] = q.y; out.pos[3 * i + 2] = q.z;
                if (have_normals) {
                    V3 n = normalize(xf_normal(xf.inv_t, v3((float) row[ix[3]], (float) row[ix[4]], (float) row[ix[5]])));
                    out.nrm[3 * i] = n.x; out.nrm[3 * i + 1] = n.y; out.nrm[3 * i + 2] = n.z;
                }
            }
            have_vertices = true;
        } else if (e.name == "face") {
            int li = -1;
            for (size_t k = 0; k < e.props.size(); ++k)
                if (e.props[k].list && (e.props[k].name == "vertex_index" || e.props[k].name == "vertex_indices")) li = (int) k;
            if (li < 0) fail("vertex_index/vertex_indices property not found");
            out.faces.reserve(3 * e.count);
            for (size_t i = 0; i < e.count; ++i)
                for (size_t k = 0; k < e.props.size(); ++k) {
                    const PlyProp &pr = e.props[k];
                    if (!pr.list) { rd.read(pr.type); continue; }
                    size_t cnt = (size_t) rd.read(pr.count_type);
                    if ((int) k == li && cnt != 3) fail("incompatible contents -- is this a triangle mesh?");
                    for (size_t j = 0; j < cnt; ++j) { double v = rd.read(pr.type); if ((int) k == li) out.faces.push_back((uint32_t) v); }
                }
        } else {
            for (size_t i = 0; i < e.count; ++i)
                for (const PlyProp &pr : e.props) {
                    if (!pr.list) { rd.read(pr.type); continue; }
                    size_t cnt = (size_t) rd.read(pr.count_type); for (size_t j = 0; j < cnt; ++j) rd.read(pr.type);
                }
        }
    }
    if (!have_vertices) fail("incompatible contents -- is this a triangle mesh?");
    if (fmt == 0) while (rd.pos < data.size() && isspace((unsigned char) data[rd.pos])) ++rd.pos;
    if (rd.pos != data.size()) fail("invalid file -- trailing content");
    if (!face_normals && !have_normals) recompute_vertex_normals(out);   // ply.cpp:444-449
}

// ------------------------------------------------------------------------------------------------ serialized
// SerializedMesh (src/shapes/serialized.cpp:237-404): header (0x041C, version 3 | 4), for shape_index > 0 the offset dictionary at the end of
// the file, then ONE zlib stream per sub-mesh: flags, [v4: name], vertex / face counts (u64), positions, [normals], [texcoords], [colours],
// faces (u32 x 3), single or double precision.  zlib is bound at run time (libz.so.1, like NCCL in comm.cpp): the library itself keeps no
// link-time dependency beyond the CUDA runtime.
namespace {
struct ZApi {
    struct z_stream_s { const unsigned char *next_in; unsigned avail_in; unsigned long total_in; unsigned char *next_out; unsigned avail_out; unsigned long total_out;
                        const char *msg; void *state; void *zalloc; void *zfree; void *opaque; int data_type; unsigned long adler; unsigned long reserved; };
    int (*inflateInit2_)(z_stream_s *, int, const char *, int) = nullptr; int (*inflate)(z_stream_s *, int) = nullptr; int (*inflateEnd)(z_stream_s *) = nullptr;
    const char *(*zlibVersion)() = nullptr;
    static ZApi &get() {
        static ZApi z; static bool tried = false;
        if (!tried) { tried = true;
            void *h = dlopen("libz.so.1", RTLD_NOW | RTLD_LOCAL); if (!h) h = dlopen("libz.so", RTLD_NOW | RTLD_LOCAL);
            if (h) { z.inflateInit2_ = (decltype(z.inflateInit2_)) dlsym(h, "inflateInit2_"); z.inflate = (decltype(z.inflate)) dlsym(h, "inflate");
                     z.inflateEnd = (decltype(z.inflateEnd)) dlsym(h, "inflateEnd"); z.zlibVersion = (decltype(z.zlibVersion)) dlsym(h, "zlibVersion"); } }
        return z;
    }
};
// Inflates one zlib stream starting at `src` until `want` more bytes are available or the stream ends.
struct Inflater {
    ZApi &z; ZApi::z_stream_s zs{}; std::vector<unsigned char> buf; size_t rd = 0; bool ended = false; const std::function<void(const std::string &)> &fail;
    Inflater(const unsigned char *src, size_t n, const std::function<void(const std::string &)> &f) : z(ZApi::get()), fail(f) {
        if (!z.inflateInit2_ || !z.inflate || !z.inflateEnd || !z.zlibVersion) fail("zlib (libz.so.1) could not be loaded");
        zs.next_in = src; zs.avail_in = (unsigned) std::min<size_t>(n, 0xffffffffu);
        if (z.inflateInit2_(&zs, 15, z.zlibVersion(), (int) sizeof(zs)) != 0) fail("could not initialize ZLIB");
    }
    ~Inflater() { if (z.inflateEnd) z.inflateEnd(&zs); }
    void read(void *dst, size_t n) {
        while (buf.size() - rd < n && !ended) {
            const size_t old = buf.size(), chunk = std::max<size_t>(1u << 20, n); buf.resize(old + chunk);
            zs.next_out = buf.data() + old; zs.avail_out = (unsigned) chunk;
            const int rc = z.inflate(&zs, 0 /* Z_NO_FLUSH */);
            buf.resize(old + chunk - zs.avail_out);
            if (rc == 1 /* Z_STREAM_END */) ended = true;
            else if (rc != 0 || (zs.avail_in == 0 && zs.avail_out != 0)) { if (rc != 0) fail("inflate(): " + std::string(zs.msg ? zs.msg : "stream error")); ended = true; }
        }
        if (buf.size() - rd < n) fail("read less data than expected");
        if (dst) std::memcpy(dst, buf.data() + rd, n);
        rd += n;
    }
    template <class T> T get() { T v; read(&v, sizeof(T)); return v; }
};
} // namespace

void load_serialized(const std::string &path, int shape_index, const M4d &to_world, bool face_normals, HMesh &out) {
    const std::function<void(const std::string &)> fail = [&](const std::string &d) { AM_THROW("io: Error while loading serialized file \"%s\": %s!", path.c_str(), d.c_str()); };
    std::ifstream f(path, std::ios::binary); if (!f) fail("file not found");
    if (shape_index < 0) fail("shape index must be nonnegative!");
    std::ostringstream ss; ss << f.rdbuf(); const std::string data = ss.str(); const size_t size = data.size();
    auto at = [&](size_t off, void *dst, size_t n) { if (off + n > size) fail("unexpected end of file"); std::memcpy(dst, data.data() + off, n); };
    int16_t format = 0, version = 0; at(0, &format, 2); at(2, &version, 2);
    if (format != 0x041C) fail("encountered an invalid file format");
    if (version != 3 && version != 4) fail("encountered an incompatible file version");
    size_t start = 4;
    if (shape_index != 0) {
        if (size < 8) fail("unexpected end of file");
        uint32_t count = 0; at(size - 4, &count, 4);
        if (shape_index >= (int) count) fail("Unable to unserialize mesh, shape index is out of range! (requested " + std::to_string(shape_index) + " out of 0.." + std::to_string((int) count - 1) + ")");
        if (version == 4) { uint64_t off = 0; at(size - 8 * (size_t) (count - shape_index) - 4, &off, 8); start = (size_t) off + 4; }
        else { uint32_t off = 0; at(size - 4 * (size_t) (count - shape_index + 1), &off, 4); start = (size_t) off + 4; }
        if (start > size) fail("unexpected end of file");
    }
    Inflater in((const unsigned char *) data.data() + start, size - start, fail);
    const uint32_t flags = in.get<uint32_t>();
    if (version == 4) for (;;) { char ch = in.get<char>(); if (ch == 0) break; }
    const uint64_t nv = in.get<uint64_t>(), nf = in.get<uint64_t>();
    if (nv > 0xffffffffull || nf > 0xffffffffull) fail("mesh too large");
    const bool dp = (flags & 0x2000u) != 0, has_normals = (flags & 0x0001u) != 0, has_uv = (flags & 0x0002u) != 0, has_colors = (flags & 0x0008u) != 0;
    auto read_floats = [&](std::vector<float> *dst, size_t dim) {
        const size_t n = (size_t) nv * dim;
        if (dp) { std::vector<double> tmp(n); in.read(tmp.data(), n * 8); if (dst) { dst->resize(n); for (size_t i = 0; i < n; ++i) (*dst)[i] = (float) tmp[i]; } }
        else if (dst) { dst->resize(n); in.read(dst->data(), n * 4); } else in.read(nullptr, n * 4);
    };
    out.pos.clear(); out.nrm.clear(); out.faces.clear();
    read_floats(&out.pos, 3);
    if (has_normals) read_floats(face_normals ? nullptr : &out.nrm, 3);
    if (has_uv) read_floats(nullptr, 2);                 // texture coordinates: no textures on this path
    out.had_uv = has_uv;
    if (has_colors) read_floats(nullptr, 3);
    out.faces.resize((size_t) nf * 3); in.read(out.faces.data(), (size_t) nf * 12);
    for (uint32_t idx : out.faces) if (idx >= nv) fail("face index out of range");
    XfF xf = to_float(to_world);
    for (size_t i = 0; i < (size_t) nv; ++i) {           // serialized.cpp:359-374
        V3 p = xf_point(xf.mat, v3(out.pos[3 * i], out.pos[3 * i + 1], out.pos[3 * i + 2]));
        out.pos[3 * i] = p.x; out.pos[3 * i + 1] = p.y; out.pos[3 * i + 2] = p.z;
        if (has_normals && !face_normals) {
            V3 n = normalize(xf_normal(xf.inv_t, v3(out.nrm[3 * i], out.nrm[3 * i + 1], out.nrm[3 * i + 2])));
            out.nrm[3 * i] = n.x; out.nrm[3 * i + 1] = n.y; out.nrm[3 * i + 2] = n.z;
        }
    }
    if (!face_normals && !has_normals) recompute_vertex_normals(out);      // :396-401
}

// ------------------------------------------------------------------------------------------------ cube
// cube.cpp:101-165: 24 vertices (four per side, each side with its own normal) and 12 triangles of [-1, 1]^3.  The sides
// come in the order -y, +y, +x, +z, -x, -z, and each side's quad (a, b, c, d) becomes the triangles (a, b, c), (d, a, c).
void make_cube(const M4d &to_world, HMesh &out) {
    XfF xf = to_float(to_world);
    static const signed char corner[6][4][3] = {
        { { 1, -1, -1 }, { 1, -1, 1 }, { -1, -1, 1 }, { -1, -1, -1 } },   // -y
        { { 1, 1, -1 }, { -1, 1, -1 }, { -1, 1, 1 }, { 1, 1, 1 } },       // +y
        { { 1, -1, -1 }, { 1, 1, -1 }, { 1, 1, 1 }, { 1, -1, 1 } },       // +x
        { { 1, -1, 1 }, { 1, 1, 1 }, { -1, 1, 1 }, { -1, -1, 1 } },       // +z
        { { -1, -1, 1 }, { -1, 1, 1 }, { -1, 1, -1 }, { -1, -1, -1 } },   // -x
        { { 1, 1, -1 }, { 1, -1, -1 }, { -1, -1, -1 }, { -1, 1, -1 } } }; // -z
    static const signed char side_n[6][3] = { { 0, -1, 0 }, { 0, 1, 0 }, { 1, 0, 0 }, { 0, 0, 1 }, { -1, 0, 0 }, { 0, 0, -1 } };
    out.pos.clear(); out.nrm.clear(); out.faces.clear();
    for (int s = 0; s < 6; ++s) {
        for (int c = 0; c < 4; ++c) {
            V3 q = xf_point(xf.mat, v3(corner[s][c][0], corner[s][c][1], corner[s][c][2]));
            V3 n = normalize(xf_normal(xf.inv_t, v3(side_n[s][0], side_n[s][1], side_n[s][2])));
            out.pos.insert(out.pos.end(), { q.x, q.y, q.z }); out.nrm.insert(out.nrm.end(), { n.x, n.y, n.z });
        }
        uint32_t b = 4u * (uint32_t) s;
        out.faces.insert(out.faces.end(), { b, b + 1, b + 2, b + 3, b, b + 2 });
    }
}

} // namespace amvpt

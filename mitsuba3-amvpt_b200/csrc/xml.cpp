// xml.cpp -- scene description loader for the subset of Mitsuba's XML format that feeds the AMVPT path:
// what xml::load_file / load_string (src/core/xml.cpp) + PluginManager::create_object do for
//   scene | integrator{mvpath, path} | sensor{grid, batch, perspective, thinlens} | wrap{any} | film{hdrfilm} | rfilter{box}
//   | sampler{independent} | shape{obj, ply, cube} | bsdf{diffuse, conductor, dielectric, roughconductor, roughdielectric,
//   roughplastic, twosided} | emitter{area}
// with the property tags float integer boolean string vector point rgb spectrum(uniform) transform{translate rotate scale
// lookat matrix}, and ref / include / default / alias / path, "$name" parameter substitution (-D), object ids.
// Anything outside that subset raises an error naming the tag / plugin, as the reference does for unknown plugins.
// The XML reader itself is a small recursive-descent parser written for this file (the reference uses pugixml).
#include "host.h"
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <set>
#include <sstream>

namespace amvpt {
namespace {

// ------------------------------------------------------------------------------------------------ XML reader
struct Node {
    std::string tag; std::vector<std::pair<std::string, std::string>> attrs; std::vector<std::unique_ptr<Node>> kids;
    std::string file; int line = 0;
    const std::string *attr(const char *n) const { for (auto &a : attrs) if (a.first == n) return &a.second; return nullptr; }
};
struct Reader {
    const std::string &s; std::string file; size_t p = 0; int line = 1;
    [[noreturn]] void fail(const char *what) { AM_THROW("io: Error while loading \"%s\" (line %d): %s", file.c_str(), line, what); }
    bool eof() const { return p >= s.size(); }
    char peek() const { return p < s.size() ? s[p] : '\0'; }
    char next() { char c = s[p++]; if (c == '\n') ++line; return c; }
    bool starts(const char *t) const { return s.compare(p, strlen(t), t) == 0; }
    void skip_to(const char *t) { size_t e = s.find(t, p); if (e == std::string::npos) fail("unterminated comment or declaration"); while (p < e + strlen(t)) next(); }
    void ws() { while (!eof() && isspace((unsigned char) peek())) next(); }
    void misc() { for (;;) { ws(); if (starts("<!--")) skip_to("-->"); else if (starts("<?")) skip_to("?>"); else if (starts("<!")) skip_to(">"); else break; } }
    std::string name() { size_t b = p; while (!eof() && (isalnum((unsigned char) peek()) || strchr("_-.:", peek()))) next(); if (b == p) fail("expected a name"); return s.substr(b, p - b); }
    std::string decode(const std::string &v) {
        std::string o; o.reserve(v.size());
        for (size_t i = 0; i < v.size(); ++i) {
            if (v[i] != '&') { o += v[i]; continue; }
            size_t e = v.find(';', i); if (e == std::string::npos) fail("malformed entity");
            std::string ent = v.substr(i + 1, e - i - 1);
            if (ent == "amp") o += '&'; else if (ent == "lt") o += '<'; else if (ent == "gt") o += '>'; else if (ent == "quot") o += '"'; else if (ent == "apos") o += '\'';
            else if (ent.size() > 1 && ent[0] == '#') o += (char) (ent[1] == 'x' ? strtol(ent.c_str() + 2, nullptr, 16) : strtol(ent.c_str() + 1, nullptr, 10));
            else fail("unknown entity");
            i = e;
        }
        return o;
    }
    std::unique_ptr<Node> element() {
        if (next() != '<') fail("expected '<'");
        auto n = std::make_unique<Node>(); n->file = file; n->line = line; n->tag = name();
        for (;;) {
            ws();
            if (peek() == '/') { next(); if (next() != '>') fail("expected '>'"); return n; }
            if (peek() == '>') { next(); break; }
            if (eof()) fail("unexpected end of file inside a tag");
            std::string k = name(); ws(); if (next() != '=') fail("expected '=' after attribute name"); ws();
            char q = next(); if (q != '"' && q != '\'') fail("attribute value must be quoted");
            size_t b = p; while (!eof() && peek() != q) next(); if (eof()) fail("unterminated attribute value");
            std::string v = s.substr(b, p - b); next();
            for (auto &a : n->attrs) if (a.first == k) fail("duplicate attribute");
            n->attrs.emplace_back(k, decode(v));
        }
        for (;;) {      // content: child elements, comments; character data is ignored (the format carries none)
            while (!eof() && peek() != '<') next();
            if (eof()) fail("unexpected end of file: unclosed element");
            if (starts("<!--")) { skip_to("-->"); continue; }
            if (starts("<![CDATA[")) { skip_to("]]>"); continue; }
            if (starts("</")) { next(); next(); std::string e = name(); ws(); if (next() != '>') fail("expected '>'"); if (e != n->tag) fail("mismatched closing tag"); return n; }
            n->kids.push_back(element());
        }
    }
    std::unique_ptr<Node> document() { misc(); if (eof()) fail("no root element"); auto r = element(); misc(); if (!eof()) fail("content after the root element"); return r; }
};

// ------------------------------------------------------------------------------------------------ property values
struct Obj;
struct Value {
    enum Kind { Float, Int, Bool, String, Vec, Color, Xform, Object } kind = Float;
    double f = 0; long long i = 0; bool b = false; std::string s; double v[3] = { 0, 0, 0 }; M4d m = m4d_identity();
    std::shared_ptr<Obj> obj; bool queried = false;
};
struct Obj {
    std::string cls, type, id, where;
    std::vector<std::pair<std::string, Value>> props;       // in document order (Properties iterates in natural name order = order of insertion for unnamed children)
    Value *find(const std::string &n) { for (auto &p : props) if (p.first == n) return &p.second; return nullptr; }
    bool has(const std::string &n) { return find(n) != nullptr; }
    [[noreturn]] void wrong(const std::string &n, const char *want) { AM_THROW("%s: the property \"%s\" has the wrong type (expected <%s>)", where.c_str(), n.c_str(), want); }
    double f(const std::string &n, double def) { Value *v = find(n); if (!v) return def; v->queried = true; if (v->kind == Value::Float) return v->f; if (v->kind == Value::Int) return (double) v->i; wrong(n, "float"); }
    double f_req(const std::string &n) { if (!has(n)) AM_THROW("%s: property \"%s\" has not been specified!", where.c_str(), n.c_str()); return f(n, 0); }
    long long i(const std::string &n, long long def) { Value *v = find(n); if (!v) return def; v->queried = true; if (v->kind != Value::Int) wrong(n, "integer"); return v->i; }
    bool b(const std::string &n, bool def) { Value *v = find(n); if (!v) return def; v->queried = true; if (v->kind != Value::Bool) wrong(n, "boolean"); return v->b; }
    std::string s(const std::string &n, const std::string &def) { Value *v = find(n); if (!v) return def; v->queried = true; if (v->kind != Value::String) wrong(n, "string"); return v->s; }
    void vec(const std::string &n, float *out) { Value *v = find(n); if (!v) return; v->queried = true; if (v->kind != Value::Vec && v->kind != Value::Color) wrong(n, "vector"); for (int k = 0; k < 3; ++k) out[k] = (float) v->v[k]; }
    M4d xform(const std::string &n) { Value *v = find(n); if (!v) return m4d_identity(); v->queried = true; if (v->kind != Value::Xform) wrong(n, "transform"); return v->m; }
    // Properties::texture<Texture>(name, default) for a constant colour: <rgb>, uniform <spectrum>, or <float>
    void color(const std::string &n, float def, float *out) {
        Value *v = find(n); if (!v) { out[0] = out[1] = out[2] = def; return; } v->queried = true;
        if (v->kind == Value::Color) for (int k = 0; k < 3; ++k) out[k] = (float) v->v[k];
        else if (v->kind == Value::Float) out[0] = out[1] = out[2] = (float) v->f;
        else if (v->kind == Value::Int) out[0] = out[1] = out[2] = (float) v->i;
        else if (v->kind == Value::Object) AM_THROW("%s: property \"%s\": textures are not supported on this path (constant <rgb>/<spectrum>/<float> only)", where.c_str(), n.c_str());
        else wrong(n, "rgb");
    }
    float scalar_tex(const std::string &n, float def) { float c[3]; color(n, def, c); return c[0]; }
    std::vector<std::shared_ptr<Obj>> children(const char *cls_filter = nullptr) {
        std::vector<std::shared_ptr<Obj>> out;
        for (auto &p : props) if (p.second.kind == Value::Object && (!cls_filter || p.second.obj->cls == cls_filter)) { out.push_back(p.second.obj); }
        return out;
    }
    void mark(const std::shared_ptr<Obj> &o) { for (auto &p : props) if (p.second.kind == Value::Object && p.second.obj == o) p.second.queried = true; }
    // PluginManager::create_object: "Unreferenced property" check (src/core/plugin.cpp:190-205)
    void check_unqueried() {
        for (auto &p : props) if (!p.second.queried)
            AM_THROW("%s: unreferenced %s \"%s\" in %s plugin of type \"%s\"", where.c_str(), p.second.kind == Value::Object ? "object" : "property", p.first.c_str(), cls.c_str(), type.c_str());
    }
};

void read_spd(const std::string &path, std::vector<double> &wl, std::vector<double> &val);     // defined below (spectrum_from_file)
struct Loader {
    std::map<std::string, std::string> params; std::set<std::string> cli_params;
    std::vector<std::string> search;
    std::map<std::string, std::shared_ptr<Obj>> by_id; std::map<std::string, std::string> aliases;
    int depth = 0;

    [[noreturn]] static void err(const Node &n, const std::string &msg) { AM_THROW("Error while loading \"%s\" (line %d): %s", n.file.c_str(), n.line, msg.c_str()); }
    static std::string where(const Node &n) { return strfmt("\"%s\" (line %d)", n.file.c_str(), n.line); }

    std::string subst(const Node &n, const std::string &v) {
        if (v.find('$') == std::string::npos) return v;
        std::string o;
        for (size_t i = 0; i < v.size();) {
            if (v[i] != '$') { o += v[i++]; continue; }
            size_t e = i + 1; while (e < v.size() && (isalnum((unsigned char) v[e]) || v[e] == '_')) ++e;
            // longest defined prefix of the identifier
            size_t best = 0; for (size_t l = e - i - 1; l > 0; --l) if (params.count(v.substr(i + 1, l))) { best = l; break; }
            if (!best) err(n, "undefined parameter \"" + v.substr(i, e - i) + "\" (use <default> or pass -Dname=value)");
            o += params[v.substr(i + 1, best)]; i += 1 + best;
        }
        return o;
    }
    std::string attr(const Node &n, const char *name) { auto *a = n.attr(name); if (!a) err(n, std::string("missing attribute \"") + name + "\" in <" + n.tag + ">"); return subst(n, *a); }
    std::string attr_opt(const Node &n, const char *name, const std::string &def) { auto *a = n.attr(name); return a ? subst(n, *a) : def; }
    static double to_f(const Node &n, const std::string &t) {
        const char *c = t.c_str(); char *end; double v = strtod(c, &end); while (*end && isspace((unsigned char) *end)) ++end;
        if (end == c || *end) err(n, "could not parse floating point value \"" + t + "\""); return v;
    }
    static long long to_i(const Node &n, const std::string &t) {
        const char *c = t.c_str(); char *end; long long v = strtoll(c, &end, 10); while (*end && isspace((unsigned char) *end)) ++end;
        if (end == c || *end) err(n, "could not parse integer value \"" + t + "\""); return v;
    }
    static std::vector<std::string> tokens(const std::string &v) {
        std::vector<std::string> o; std::string cur;
        for (char c : v) { if (c == ',' || isspace((unsigned char) c)) { if (!cur.empty()) { o.push_back(cur); cur.clear(); } } else cur += c; }
        if (!cur.empty()) o.push_back(cur); return o;
    }
    // detail::expand_value_to_xyz + parse_vector (src/core/xml.cpp): value="a" | value="x,y,z" | x= y= z= (missing -> def)
    void xyz(const Node &n, double def, double *out) {
        if (n.attr("value")) {
            auto t = tokens(attr(n, "value"));
            if (t.size() == 1) out[0] = out[1] = out[2] = to_f(n, t[0]);
            else if (t.size() == 3) for (int k = 0; k < 3; ++k) out[k] = to_f(n, t[k]);
            else err(n, "\"value\" attribute must have exactly 1 or 3 elements");
            return;
        }
        const char *nm[3] = { "x", "y", "z" };
        for (int k = 0; k < 3; ++k) out[k] = n.attr(nm[k]) ? to_f(n, attr(n, nm[k])) : def;
    }
    void named_vec(const Node &n, const char *name, double *out) {
        auto t = tokens(attr(n, name)); if (t.size() != 3) err(n, std::string("\"") + name + "\" must have three components");
        for (int k = 0; k < 3; ++k) out[k] = to_f(n, t[k]);
    }
    M4d transform(const Node &n) {      // src/core/xml.cpp:838-926: every child left-multiplies the accumulated transform
        M4d acc = m4d_identity();
        for (auto &k : n.kids) {
            M4d op;
            if (k->tag == "translate") { double v[3]; xyz(*k, 0.0, v); op = m4d_translate(v[0], v[1], v[2]); }
            else if (k->tag == "scale") { double v[3]; xyz(*k, 1.0, v); op = m4d_scale(v[0], v[1], v[2]); }
            else if (k->tag == "rotate") { double v[3]; xyz(*k, 0.0, v); if (v[0] == 0 && v[1] == 0 && v[2] == 0) err(*k, "rotate: zero axis"); op = m4d_rotate(v[0], v[1], v[2], to_f(*k, attr(*k, "angle"))); }
            else if (k->tag == "lookat") {
                double o[3], t[3], up[3] = { 0, 0, 0 }; named_vec(*k, "origin", o); named_vec(*k, "target", t); if (k->attr("up")) named_vec(*k, "up", up);
                if (up[0] == 0 && up[1] == 0 && up[2] == 0) {        // coordinate_system(normalize(target - origin)).first
                    V3 d = normalize(v3((float) (t[0] - o[0]), (float) (t[1] - o[1]), (float) (t[2] - o[2]))), s, tt; coordinate_system(d, &s, &tt);
                    up[0] = s.x; up[1] = s.y; up[2] = s.z;
                }
                op = m4d_look_at(o, t, up);
                for (auto &row : op.m) for (double e : row) if (e != e) err(*k, "invalid lookat transformation");
            } else if (k->tag == "matrix") {
                auto t = tokens(attr(*k, "value")); op = m4d_identity();
                if (t.size() == 16) { for (int i = 0; i < 16; ++i) op.m[i / 4][i % 4] = to_f(*k, t[i]); }
                else if (t.size() == 9) { for (int i = 0; i < 9; ++i) op.m[i / 3][i % 3] = to_f(*k, t[i]); }
                else err(*k, "matrix: expected 16 or 9 values");
            } else err(*k, "unexpected <" + k->tag + "> inside <transform>");
            acc = m4d_mul(op, acc);
        }
        return acc;
    }

    std::string resolve(const std::string &fn, const std::string &base) {
        if (!fn.empty() && fn[0] == '/') return fn;
        std::vector<std::string> dirs = { base }; dirs.insert(dirs.end(), search.begin(), search.end());
        for (auto &d : dirs) { std::string c = d.empty() ? fn : d + "/" + fn; std::ifstream f(c); if (f) return c; }
        return base.empty() ? fn : base + "/" + fn;
    }
    static std::string dir_of(const std::string &path) { size_t s = path.rfind('/'); return s == std::string::npos ? "." : path.substr(0, s); }

    std::unique_ptr<Node> parse_text(const std::string &text, const std::string &file) { Reader r{ text, file }; return r.document(); }
    std::unique_ptr<Node> parse_file(const std::string &path) {
        std::ifstream f(path, std::ios::binary); if (!f) AM_THROW("io: \"%s\": file does not exist!", path.c_str());
        std::ostringstream ss; ss << f.rdbuf(); return parse_text(ss.str(), path);
    }
    static void check_root(const Node &root) {
        if (!root.attr("version")) err(root, "missing version attribute in root element \"" + root.tag + "\"");
    }

    static bool is_object_tag(const std::string &t) {
        static const char *tags[] = { "scene", "shape", "bsdf", "emitter", "sensor", "film", "sampler", "rfilter", "integrator", "texture", "wrap", "medium", "phase", "volume", "spectrum_obj" };
        for (auto *x : tags) if (t == x) return true; return false;
    }

    // One object element -> Obj (properties in document order, children by value).  `base` = directory of the file the
    // element came from (relative `filename` properties and <include>s resolve against it first).
    std::shared_ptr<Obj> object(const Node &n, const std::string &base) {
        if (++depth > 64) err(n, "objects nested too deeply (include cycle?)");
        auto o = std::make_shared<Obj>(); o->cls = n.tag; o->where = where(n);
        if (n.tag != "scene") o->type = attr(n, "type");
        o->id = attr_opt(n, "id", "");
        int anon = 0;
        auto put = [&](const Node &k, const std::string &name, Value v) {
            std::string nm = name.empty() ? strfmt("_arg_%d", anon++) : name;
            if (o->has(nm)) err(k, "duplicate property \"" + nm + "\"");
            o->props.emplace_back(nm, std::move(v));
        };
        for (auto &kp : n.kids) consume(*kp, *o, base, put);
        --depth;
        if (!o->id.empty()) {
            if (by_id.count(o->id)) err(n, "\"" + n.tag + "\" has duplicate id \"" + o->id + "\"");
            by_id[o->id] = o;
        }
        return o;
    }
    template <class Put> void consume(const Node &k, Obj &o, const std::string &base, Put &put) {
        const std::string &t = k.tag;
        if (t == "default") { std::string nm = attr(k, "name"); if (!params.count(nm)) params[nm] = attr(k, "value"); return; }
        if (t == "path") { search.push_back(resolve(attr(k, "value"), base)); return; }
        if (t == "alias") { std::string src = attr(k, "id"), dst = attr(k, "as"); if (aliases.count(dst) || by_id.count(dst)) err(k, "alias: id \"" + dst + "\" already exists"); aliases[dst] = src; return; }
        if (t == "include") {
            std::string path = resolve(attr(k, "filename"), base); auto root = parse_file(path); check_root(*root);
            if (root->tag != "scene") {     // a single object: treated as if it stood in place of the <include>
                Value v; v.kind = Value::Object; v.obj = object(*root, dir_of(path)); put(k, attr_opt(k, "name", ""), std::move(v)); return;
            }
            if (++depth > 64) err(k, "includes nested too deeply");
            for (auto &kk : root->kids) consume(*kk, o, dir_of(path), put);
            --depth; return;
        }
        std::string name = attr_opt(k, "name", "");
        Value v;
        if (t == "float") { v.kind = Value::Float; v.f = to_f(k, attr(k, "value")); }
        else if (t == "integer") { v.kind = Value::Int; v.i = to_i(k, attr(k, "value")); }
        else if (t == "boolean") { std::string s = attr(k, "value"); for (auto &c : s) c = (char) tolower(c); v.kind = Value::Bool; if (s == "true") v.b = true; else if (s == "false") v.b = false; else err(k, "could not parse boolean value \"" + s + "\" -- must be \"true\" or \"false\""); }
        else if (t == "string") { v.kind = Value::String; v.s = attr(k, "value"); }
        else if (t == "vector" || t == "point") { v.kind = Value::Vec; xyz(k, 0.0, v.v); }
        else if (t == "rgb") {
            auto tk = tokens(attr(k, "value")); v.kind = Value::Color;
            if (tk.size() == 1) v.v[0] = v.v[1] = v.v[2] = to_f(k, tk[0]); else if (tk.size() == 3) for (int c = 0; c < 3; ++c) v.v[c] = to_f(k, tk[c]);
            else err(k, "'rgb' tag requires one or three values (got \"" + attr(k, "value") + "\")");
        }
        else if (t == "spectrum") {
            // src/core/xml.cpp:777-833 + create_texture_from_spectrum (:1170-1268) in an RGB variant: a uniform value stays uniform
            // (inside an emitter it is the colour of a flat spectrum under the sRGB white point), wavelength:value pairs are integrated
            // against the CIE 1931 observer (reflectances also against D65) and become an sRGB colour
            const bool within_emitter = o.cls == "emitter", unbounded = name == "eta" || name == "k" || name == "int_ior" || name == "ext_ior";
            if (k.attr("filename")) {           // src/core/xml.cpp:805-815: spectrum_from_file, then the same conversion as wavelength:value pairs
                std::vector<double> wl, val; read_spd(resolve(attr(k, "filename"), base), wl, val);
                v.kind = Value::Color; spectrum_to_srgb(wl, val, within_emitter, unbounded, v.v);
                if (name.empty()) err(k, "missing attribute \"name\" in <spectrum>");
                put(k, name, std::move(v)); return;
            }
            auto tk = tokens(attr(k, "value")); v.kind = Value::Color;
            if (tk.empty()) err(k, "'spectrum' tag requires one of \"value\" or \"filename\" attributes");
            if (tk.size() == 1 && tk[0].find(':') == std::string::npos) {
                const double c = to_f(k, tk[0]);
                if (within_emitter && !unbounded) uniform_emitter_spectrum_to_srgb(c, v.v); else v.v[0] = v.v[1] = v.v[2] = c;
            } else {
                std::vector<double> wl, val;
                for (auto &tok : tk) {
                    size_t c = tok.find(':'); if (c == std::string::npos || tok.find(':', c + 1) != std::string::npos) err(k, "invalid spectrum (expected wavelength:value pairs)");
                    wl.push_back(to_f(k, tok.substr(0, c))); val.push_back(to_f(k, tok.substr(c + 1)));
                }
                if (wl.size() < 2) err(k, "invalid spectrum (expected at least two wavelength:value pairs)");
                spectrum_to_srgb(wl, val, within_emitter, unbounded, v.v);
            }
        }
        else if (t == "transform") { v.kind = Value::Xform; v.m = transform(k); }
        else if (t == "ref") {
            std::string id = attr(k, "id"); v.kind = Value::Object; v.s = id;     // resolved after the whole document is read
        }
        else if (is_object_tag(t)) { v.kind = Value::Object; v.obj = object(k, base); }
        else err(k, "unexpected tag <" + t + ">");
        if (name.empty() && v.kind != Value::Object) err(k, "missing attribute \"name\" in <" + t + ">");
        put(k, name, std::move(v));
    }
    void resolve_refs(Obj &o, std::set<Obj *> &seen) {
        if (!seen.insert(&o).second) return;
        for (auto &p : o.props) if (p.second.kind == Value::Object) {
            if (!p.second.obj) {
                std::string id = p.second.s; auto al = aliases.find(id); if (al != aliases.end()) id = al->second;
                auto it = by_id.find(id);
                if (it == by_id.end()) AM_THROW("%s: reference to unknown object \"%s\"!", o.where.c_str(), p.second.s.c_str());
                p.second.obj = it->second;
            }
            resolve_refs(*p.second.obj, seen);
        }
    }
};

// ------------------------------------------------------------------------------------------------ plugin constructors
// include/mitsuba/render/ior.h:27-52 (named indices of refraction, 589 nm)
float ior_by_name(const std::string &name_) {
    std::string n = name_; for (auto &c : n) c = (char) tolower(c);
    static const std::pair<const char *, float> table[] = {
        { "vacuum", 1.0f }, { "helium", 1.000036f }, { "hydrogen", 1.000132f }, { "air", 1.000277f }, { "carbon dioxide", 1.00045f },
        { "water", 1.3330f }, { "acetone", 1.36f }, { "ethanol", 1.361f }, { "carbon tetrachloride", 1.461f }, { "glycerol", 1.4729f },
        { "benzene", 1.501f }, { "silicone oil", 1.52045f }, { "bromine", 1.661f }, { "water ice", 1.31f }, { "fused quartz", 1.458f },
        { "pyrex", 1.470f }, { "acrylic glass", 1.49f }, { "polypropylene", 1.49f }, { "bk7", 1.5046f }, { "sodium chloride", 1.544f },
        { "amber", 1.55f }, { "pet", 1.5750f }, { "diamond", 2.419f } };
    for (auto &e : table) if (n == e.first) return e.second;
    AM_THROW("Unable to find an IOR value for \"%s\"!", n.c_str());
}
float lookup_ior(Obj &o, const char *name, const char *def) {      // ior.h:75-81
    Value *v = o.find(name);
    if (v && (v->kind == Value::Float || v->kind == Value::Int)) return (float) o.f(name, 0);
    return ior_by_name(o.s(name, def));
}
int distribution(Obj &o) {
    if (!o.has("distribution")) return 0;
    std::string d = o.s("distribution", "beckmann"); for (auto &c : d) c = (char) tolower(c);
    if (d == "beckmann") return 0; if (d == "ggx") return 1;
    AM_THROW("%s: Specified an invalid distribution \"%s\", must be \"beckmann\" or \"ggx\"!", o.where.c_str(), d.c_str());
}
void alphas(Obj &o, DBsdf &b) {      // roughconductor.cpp:189-199
    if (o.has("alpha_u") || o.has("alpha_v")) {
        if (!o.has("alpha_u") || !o.has("alpha_v")) AM_THROW("%s: Microfacet model: both 'alpha_u' and 'alpha_v' must be specified.", o.where.c_str());
        if (o.has("alpha")) AM_THROW("%s: Microfacet model: please specify either 'alpha' or 'alpha_u'/'alpha_v'.", o.where.c_str());
        b.alpha_u = o.scalar_tex("alpha_u", 0.1f); b.alpha_v = o.scalar_tex("alpha_v", 0.1f);
    } else b.alpha_u = b.alpha_v = o.scalar_tex("alpha", 0.1f);
}
// spectrum_from_file (src/core/spectrum.cpp:10-55): `.spd` text, "wavelength value" per line, '#' starts a comment that runs to the end of the line
void read_spd(const std::string &path, std::vector<double> &wl, std::vector<double> &val) {
    if (path.size() < 4 || path.compare(path.size() - 4, 4, ".spd") != 0) AM_THROW("io: \"%s\": You need to provide a valid extension like \".spd\" to read the information from an ASCII file.", path.c_str());
    std::ifstream f(path, std::ios::binary); if (!f) AM_THROW("io: \"%s\": file does not exist!", path.c_str());
    std::ostringstream ss; ss << f.rdbuf(); const std::string text = ss.str();
    const char *cur = text.c_str(), *end = cur + text.size(); bool comment = false; size_t counter = 0;
    while (cur < end) {
        const char c = *cur;
        if (c == '#') { comment = true; ++cur; }
        else if (c == '\n') { comment = false; counter = 0; ++cur; }
        else if (!comment && c != ' ' && c != '\r') {
            char *tmp = nullptr; const double v = std::strtod(cur, &tmp);
            if (tmp == cur) AM_THROW("io: \"%s\": could not parse a floating point value", path.c_str());
            cur = tmp;
            if (counter == 0) wl.push_back(v); else if (counter == 1) val.push_back(v);
            else AM_THROW("io: \"%s\": While parsing the file, more than two elements were defined in a line", path.c_str());
            ++counter;
        } else ++cur;
    }
    if (wl.size() != val.size() || wl.size() < 2) AM_THROW("io: \"%s\": expected at least two \"wavelength value\" lines", path.c_str());
}
// ior_from_file in an RGB variant (include/mitsuba/render/ior.h:101-137): values * MI_CIE_Y_NORMALIZATION, spectrum_list_to_srgb(bounded = false,
// d65 = false), handed to an unbounded `srgb` texture, i.e. used as the colour it is
void ior_from_file(const std::string &path, float *rgb) {
    std::vector<double> wl, val; read_spd(path, wl, val);
    for (double &v : val) v *= 1.0 / 106.7502593994140625;
    double c[3]; spectrum_list_to_srgb(wl, val, false, false, c);
    for (int k = 0; k < 3; ++k) rgb[k] = (float) c[k];
}

struct Builder {
    Scene &sc; std::map<Obj *, int> bsdf_index; std::set<Obj *> used_emitters; Loader *loader = nullptr; std::string base_dir;

    // conductor.cpp:228-241 / roughconductor.cpp: (eta, k) given explicitly, `material = none` (eta 0, k 1: a perfect mirror), or a preset read
    // from data/ior/<material>.{eta,k}.spd (complex_ior_from_file, ior.h:138-143).  The reference finds the files through its resource
    // directory; here: the scene's directory, the <path> entries, $AMVPT_DATA_DIR, and `data/ior` / `ior` under each ancestor of the scene's directory.
    void conductor_ior(Obj &o, DBsdf &b) {
        const std::string m = o.s("material", "none");
        if (o.has("eta") || m == "none") {
            if (o.has("material") && m != "none") AM_THROW("%s: Should specify either (eta, k) or material, not both.", o.where.c_str());
            o.color("eta", 0.f, b.eta); o.color("k", 1.f, b.k); return;
        }
        std::vector<std::string> roots; if (const char *e = getenv("AMVPT_DATA_DIR")) roots.push_back(e);
        for (std::string d = base_dir.empty() ? "." : base_dir; ; ) {
            roots.push_back(d + "/data"); roots.push_back(d);
            size_t sl = d.rfind('/'); if (sl == std::string::npos || sl == 0) break; d = d.substr(0, sl);
        }
        if (loader) for (auto &d : loader->search) { roots.push_back(d + "/data"); roots.push_back(d); }
        auto find = [&](const std::string &leaf) -> std::string {
            for (auto &r : roots) { std::string c = r + "/ior/" + leaf; std::ifstream f(c); if (f) return c; }
            AM_THROW("io: %s: \"data/ior/%s\": file does not exist! (conductor material \"%s\"; set AMVPT_DATA_DIR to the reference's resources/data directory)", o.where.c_str(), leaf.c_str(), m.c_str());
        };
        ior_from_file(find(m + ".eta.spd"), b.eta); ior_from_file(find(m + ".k.spd"), b.k);
    }

    int bsdf(const std::shared_ptr<Obj> &op) {
        auto it = bsdf_index.find(op.get()); if (it != bsdf_index.end()) return it->second;
        Obj *o = op.get(); DBsdf b{}; bool twosided = false;
        if (o->type == "twosided") {        // twosided.cpp:69-108
            auto kids = o->children("bsdf");
            if (kids.empty()) AM_THROW("%s: A nested one-sided material is required!", o->where.c_str());
            if (kids.size() > 2) AM_THROW("%s: At most two nested BSDFs can be specified!", o->where.c_str());
            if (kids.size() == 2 && kids[0] != kids[1]) AM_THROW("%s: twosided with two different nested BSDFs is not supported on this path", o->where.c_str());
            for (auto &k : kids) o->mark(k);
            o->check_unqueried(); o = kids[0].get(); twosided = true;
            if (o->type == "twosided") AM_THROW("%s: nested twosided", o->where.c_str());
        }
        for (int c = 0; c < 3; ++c) { b.spec_refl[c] = 1.f; b.spec_trans[c] = 1.f; b.reflectance[c] = .5f; b.k[c] = 1.f; }
        b.ior_eta = 1.f; b.alpha_u = b.alpha_v = 0.1f;
        const std::string &t = o->type;
        if (t == "diffuse") { b.type = BSDF_DIFFUSE; o->color("reflectance", .5f, b.reflectance); }
        else if (t == "conductor" || t == "roughconductor") {
            b.type = t == "conductor" ? BSDF_CONDUCTOR : BSDF_ROUGHCONDUCTOR;
            conductor_ior(*o, b); o->color("specular_reflectance", 1.f, b.spec_refl);
            if (b.type == BSDF_ROUGHCONDUCTOR) { b.distr = distribution(*o); if (!o->b("sample_visible", true)) AM_THROW("%s: sample_visible=false is not supported on this path", o->where.c_str()); alphas(*o, b); }
        }
        else if (t == "dielectric" || t == "roughdielectric") {
            b.type = t == "dielectric" ? BSDF_DIELECTRIC : BSDF_ROUGHDIELECTRIC;
            float ii = lookup_ior(*o, "int_ior", "bk7"), ei = lookup_ior(*o, "ext_ior", "air");
            if (ii < 0 || ei < 0) AM_THROW("%s: The interior and exterior indices of refraction must be positive!", o->where.c_str());
            b.ior_eta = ii / ei;
            o->color("specular_reflectance", 1.f, b.spec_refl); o->color("specular_transmittance", 1.f, b.spec_trans);
            if (b.type == BSDF_ROUGHDIELECTRIC) { b.distr = distribution(*o); if (!o->b("sample_visible", true)) AM_THROW("%s: sample_visible=false is not supported on this path", o->where.c_str()); alphas(*o, b); }
        }
        else if (t == "roughplastic") {
            b.type = BSDF_ROUGHPLASTIC;
            float ii = lookup_ior(*o, "int_ior", "polypropylene"), ei = lookup_ior(*o, "ext_ior", "air");
            if (ii < 0 || ei < 0 || ii == ei) AM_THROW("%s: The interior and exterior indices of refraction must be positive and differ!", o->where.c_str());
            b.ior_eta = ii / ei;
            o->color("diffuse_reflectance", .5f, b.reflectance); o->color("specular_reflectance", 1.f, b.spec_refl);
            b.nonlinear = o->b("nonlinear", false); b.distr = distribution(*o);
            if (!o->b("sample_visible", true)) AM_THROW("%s: sample_visible=false is not supported on this path", o->where.c_str());
            b.alpha_u = b.alpha_v = (float) o->f("alpha", 0.1);
        }
        else AM_THROW("%s: BSDF plugin \"%s\" is not part of the AMVPT path (supported: diffuse, conductor, dielectric, roughconductor, roughdielectric, roughplastic, twosided)", o->where.c_str(), t.c_str());
        if (twosided && (b.type == BSDF_DIELECTRIC || b.type == BSDF_ROUGHDIELECTRIC)) AM_THROW("%s: Only materials without a transmission component can be nested!", o->where.c_str());
        b.twosided = twosided ? 1 : 0;
        o->check_unqueried();
        int idx = sc.add_bsdf(b, op->id); bsdf_index[op.get()] = idx; return idx;
    }

    void shape(const std::shared_ptr<Obj> &op, const std::string &base, Loader &ld) {
        Obj &o = *op; HMesh m; m.id = o.id;
        M4d tw = o.xform("to_world");
        bool face_normals = o.b("face_normals", false); m.flip_normals = o.b("flip_normals", false) ? 1 : 0;
        if (o.type == "obj") load_obj(ld.resolve(o.s("filename", ""), base), tw, face_normals, o.b("flip_tex_coords", true), m);
        else if (o.type == "ply") { o.b("flip_tex_coords", false); load_ply(ld.resolve(o.s("filename", ""), base), tw, face_normals, m); }
        else if (o.type == "serialized") load_serialized(ld.resolve(o.s("filename", ""), base), (int) o.i("shape_index", 0), tw, face_normals, m);
        else if (o.type == "cube") { make_cube(tw, m); if (face_normals) m.nrm.clear(); }
        else AM_THROW("%s: shape plugin \"%s\" is not part of the AMVPT path (supported meshes: obj, ply, serialized, cube)", o.where.c_str(), o.type.c_str());
        if (face_normals) m.nrm.clear();
        std::shared_ptr<Obj> bs, em;
        for (auto &k : o.children()) {
            if (k->cls == "bsdf") { if (bs) AM_THROW("%s: Only a single BSDF child object can be specified per shape.", o.where.c_str()); bs = k; }
            else if (k->cls == "emitter") { if (em) AM_THROW("%s: Only a single Emitter child object can be specified per shape.", o.where.c_str()); em = k; }
            else AM_THROW("%s: unsupported child <%s> of a shape", o.where.c_str(), k->cls.c_str());
            o.mark(k);
        }
        float radiance[3]; bool emits = false;
        if (em) {
            if (em->type != "area") AM_THROW("%s: emitter plugin \"%s\" is not part of the AMVPT path (supported: area)", em->where.c_str(), em->type.c_str());
            if (!used_emitters.insert(em.get()).second) AM_THROW("%s: An area light can be only be attached to a single shape.", em->where.c_str());
            em->color("radiance", 1.f, radiance); em->check_unqueried(); emits = true;     // area.cpp:60: radiance defaults to d65 = 1 in RGB
        }
        if (bs) m.bsdf = bsdf(bs);
        else { DBsdf b{}; b.type = BSDF_DIFFUSE; float r = emits ? 0.f : .5f; b.reflectance[0] = b.reflectance[1] = b.reflectance[2] = r; m.bsdf = sc.add_bsdf(b); }   // shape.cpp:64-70
        // Mesh::compute_surface_interaction (mesh.cpp:1484-1510) derives dp_du / dp_dv -- the orientation of the shading frame about the normal -- from
        // the texture coordinates when the mesh has them; this path drops them and always uses coordinate_system(n).  For isotropic BSDFs the two
        // are the same distribution (the golden Z-tests on the UV-mapped sphere pass); an anisotropic lobe would point elsewhere: refuse it.
        if (bs && m.had_uv) { const DBsdf &b = sc.bsdfs[m.bsdf]; if ((b.type == BSDF_ROUGHCONDUCTOR || b.type == BSDF_ROUGHDIELECTRIC) && b.alpha_u != b.alpha_v)
            AM_THROW("%s: anisotropic roughness (alpha_u != alpha_v) on a mesh with texture coordinates follows the UV parameterisation in the reference, which this path does not keep", o.where.c_str()); }
        o.check_unqueried();
        sc.add_mesh(std::move(m), emits ? radiance : nullptr);
    }

    // film / sampler: either plugin objects or `wrap` objects of that class (src/render/sensor.cpp:25-52)
    static bool is_a(const std::shared_ptr<Obj> &o, const char *cls) { return o->cls == cls || (o->cls == "wrap" && wrap_class(*o) == cls); }
    static std::string wrap_class(Obj &o) { Value *v = o.find("wrap_class"); return v && v->kind == Value::String ? v->s : ""; }
    static std::string plugin_type(Obj &o) { if (o.cls != "wrap") return o.type; o.s("wrap_class", ""); return o.s("wrap_type", ""); }
    void film(Obj *f, uint32_t *w, uint32_t *h) {
        *w = 768; *h = 576;             // hdrfilm defaults (src/render/film.cpp:13-15)
        if (!f) return;
        std::string t = plugin_type(*f);
        if (t != "hdrfilm") AM_THROW("%s: film plugin \"%s\" is not part of the AMVPT path (supported: hdrfilm)", f->where.c_str(), t.c_str());
        *w = (uint32_t) f->i("width", 768); *h = (uint32_t) f->i("height", 576);
        std::string pf = f->s("pixel_format", "rgb"); for (auto &c : pf) c = (char) tolower(c);
        if (pf != "rgb") AM_THROW("%s: pixel_format \"%s\": only \"rgb\" is supported on this path", f->where.c_str(), pf.c_str());
        f->s("component_format", "float16"); f->s("file_format", "openexr"); f->b("compensate", false);      // compensate: Kahan-compensated accumulation, not observable beyond rounding
        if (f->b("sample_border", false)) AM_THROW("%s: sample_border = true enlarges the film by the filter border and shifts the lane map (mvpath.cpp:29-30, 187-188); it is not supported on this path", f->where.c_str());
        for (const char *c : { "crop_offset_x", "crop_offset_y", "crop_width", "crop_height" }) if (f->has(c)) AM_THROW("%s: film crop windows are not supported on this path", f->where.c_str());
        // hdrfilm's default reconstruction filter is the gaussian with stddev 0.5 (src/render/film.cpp:34-40)
        int ftype = 1; float fparam = 0.5f, fparam2 = 0.f; bool have_filter = false;
        for (auto &k : f->children()) {
            if (k->cls != "rfilter") AM_THROW("%s: unexpected child <%s> of a film", f->where.c_str(), k->cls.c_str());
            if (have_filter) AM_THROW("%s: A film can only have one reconstruction filter.", f->where.c_str());
            if (k->type == "box") { ftype = 0; fparam = 0.5f; }
            else if (k->type == "gaussian") { ftype = 1; fparam = (float) k->f("stddev", 0.5); }
            else if (k->type == "tent") { ftype = 2; fparam = (float) k->f("radius", 1.0); }
            else if (k->type == "mitchell") { ftype = 3; fparam = (float) k->f("B", 1.0 / 3.0); fparam2 = (float) k->f("C", 1.0 / 3.0); }
            else if (k->type == "catmullrom") { ftype = 4; fparam = 0.f; }
            else if (k->type == "lanczos") { ftype = 5; fparam = (float) k->i("lobes", 3); }
            else AM_THROW("%s: unknown reconstruction filter \"%s\" (box, gaussian, tent, mitchell, catmullrom, lanczos)", k->where.c_str(), k->type.c_str());
            k->check_unqueried(); f->mark(k); have_filter = true;
        }
        sc.set_rfilter(ftype, fparam, fparam2);
        f->check_unqueried();
    }
    void sampler(Obj *s) {
        sc.sample_count = 4; sc.sampler_seed = 0;        // sensor.cpp:61-66, src/render/sampler.cpp:13-20
        if (!s) return;
        std::string t = plugin_type(*s);
        if (t != "independent") AM_THROW("%s: sampler plugin \"%s\" is not part of the AMVPT path (supported: independent)", s->where.c_str(), t.c_str());
        sc.sample_count = (uint32_t) s->i("sample_count", 4); sc.sampler_seed = (uint32_t) s->i("seed", 0);
        s->check_unqueried();
    }
    static int fov_axis(Obj &o, const std::string &where) {
        std::string a = o.s("fov_axis", "x"); for (auto &c : a) c = (char) tolower(c);
        const char *names[] = { "x", "y", "diagonal", "smaller", "larger" };
        for (int i = 0; i < 5; ++i) if (a == names[i]) return i;
        AM_THROW("%s: parse_fov(): invalid fov_axis value \"%s\", must be one of \"x\", \"y\", \"diagonal\", \"smaller\", or \"larger\"!", where.c_str(), a.c_str());
    }
    // parse_fov (src/render/sensor.cpp:177-231): `fov` + `fov_axis`, or `focal_length` of a 35 mm film (36 x 24 mm, diagonal)
    static void fov_props(Obj &o, double *fov, int *axis) {
        if (o.has("fov") && o.has("focal_length")) AM_THROW("%s: Please specify either a focal length ('focal_length') or a field of view ('fov')!", o.where.c_str());
        if (o.has("fov")) { *fov = o.f("fov", 45.0); *axis = fov_axis(o, o.where); return; }
        std::string fl = o.s("focal_length", "50mm");
        if (fl.size() > 2 && fl.substr(fl.size() - 2) == "mm") fl = fl.substr(0, fl.size() - 2);
        char *end; double mm = strtod(fl.c_str(), &end);
        if (end == fl.c_str() || *end) AM_THROW("%s: Could not parse the focal length (must be of the form <x>mm, where <x> is a positive integer)!", o.where.c_str());
        *fov = 2.0 * std::atan(std::sqrt(36.0 * 36.0 + 24.0 * 24.0) / (2.0 * mm)) * (180.0 / 3.14159265358979323846); *axis = 2;
    }

    void sensor(const std::shared_ptr<Obj> &op, bool need_multi) {
        Obj &o = *op;
        Obj *f = nullptr, *s = nullptr; std::shared_ptr<Obj> sub_wrap; std::vector<std::shared_ptr<Obj>> sub_sensors;
        for (auto &k : o.children()) {
            if (is_a(k, "film")) { if (f) AM_THROW("%s: Only one film can be specified per sensor.", o.where.c_str()); f = k.get(); }
            else if (is_a(k, "sampler")) { if (s) AM_THROW("%s: Only one sampler can be specified per sensor.", o.where.c_str()); s = k.get(); }
            else if (k->cls == "wrap" && wrap_class(*k) == "sensor") { if (sub_wrap) AM_THROW("%s: Only one Wrap of type Sensor can be specified !", o.where.c_str()); sub_wrap = k; }
            else if (k->cls == "sensor") sub_sensors.push_back(k);
            else AM_THROW("%s: unexpected child <%s> of a sensor", o.where.c_str(), k->cls.c_str());
            o.mark(k);
        }
        uint32_t fw, fh; film(f, &fw, &fh); sampler(s);
        { const float so = (float) o.f("shutter_open", 0.0), open_time = (float) o.f("shutter_close", 0.0) - so;       // sensor.cpp:18-22
          if (open_time < 0.f) AM_THROW("%s: Shutter opening time must be less than or equal to the shutter closing time!", o.where.c_str());
          sc.shutter_draw = open_time > 0.f; }
        if (o.type == "grid") {         // grid.cpp:84-236
            if (!sub_wrap) AM_THROW("%s: Need to specify wraps for sensor, film and sampler !", o.where.c_str());
            Obj &sp = *sub_wrap; std::string cam = plugin_type(sp);
            if (cam != "perspective" && cam != "thinlens") AM_THROW("%s: grid sub-sensor \"%s\": MVPathIntegrator requires Perspective or Thinlens camera", sp.where.c_str(), cam.c_str());
            for (auto &k : sp.children()) sp.mark(k);     // nested films / samplers / wraps are dropped (grid.cpp:160-168)
            GridSensorParams g; g.thinlens = cam == "thinlens";
            g.reverse_x = o.b("reverse_x", false); g.reverse_y = o.b("reverse_y", true);
            g.grid_x = (uint32_t) o.i("grid_x", 1); g.grid_y = (uint32_t) o.i("grid_y", 1);
            uint32_t rx = (uint32_t) o.i("res_x", 0), ry = (uint32_t) o.i("res_y", 0);
            g.res_x = rx ? rx : fw * g.grid_x; g.res_y = ry ? ry : fh * g.grid_y;
            if (o.has("cone_deg")) { g.use_cone = true; g.cone_deg = (float) o.f("cone_deg", 0); }
            else if (o.has("cam_dir")) {
                o.vec("cam_dir", g.cam_dir); g.cam_dist = norm(v3(g.cam_dir[0], g.cam_dir[1], g.cam_dir[2]));
                g.cam_dist = (float) o.f("cam_dist", g.cam_dist); g.cam_center = o.b("cam_center", true);
            } else if (o.has("cam_end")) {   // direction towards a world-space end point, expressed in camera space (grid.cpp:120-127)
                float e[3] = { 0, 0, 0 }; o.vec("cam_end", e); M4d tw = o.xform("to_world"); M4 inv = m4_from_d(m4d_inverse(tw)); M4 twf = m4_from_d(tw);
                V3 d = v3(e[0] - twf.m[0][3], e[1] - twf.m[1][3], e[2] - twf.m[2][3]);
                V3 r = v3(fmadd(inv.m[0][2], d.z, fmadd(inv.m[0][1], d.y, inv.m[0][0] * d.x)), fmadd(inv.m[1][2], d.z, fmadd(inv.m[1][1], d.y, inv.m[1][0] * d.x)),
                          fmadd(inv.m[2][2], d.z, fmadd(inv.m[2][1], d.y, inv.m[2][0] * d.x)));
                g.cam_dist = norm(r); V3 u = r / g.cam_dist; g.cam_dir[0] = u.x; g.cam_dir[1] = u.y; g.cam_dir[2] = u.z; g.cam_center = false;
            }
            o.vec("cam_off", g.cam_off);
            g.to_world = o.xform("to_world");
            Obj &fov_src = o.has("fov") ? o : sp; fov_props(fov_src, &g.fov, &g.fov_axis);
            if (&fov_src == &o) { sp.f("fov", 0); sp.s("fov_axis", ""); sp.s("focal_length", ""); }
            g.focus_distance = (float) o.f("focus_distance", sp.f("focus_distance", 1.0));
            g.near_clip = (float) sp.f("near_clip", 1e-2); g.far_clip = (float) sp.f("far_clip", 1e4);
            if (g.thinlens) g.aperture_radius = (float) sp.f_req("aperture_radius");
            if (sp.has("to_world")) sp.xform("to_world");     // overwritten per sub-camera (grid.cpp:196, 214)
            sp.check_unqueried();
            sc.set_sensor_grid(g);
        } else if (o.type == "batch") {   // batch.cpp:94-130
            if (sub_sensors.empty()) AM_THROW("%s: BatchSensor: at least one child sensor must be specified!", o.where.c_str());
            uint32_t n = (uint32_t) sub_sensors.size(), sub = fw / n;
            if (sub * n != fw) AM_THROW("%s: BatchSensor: the horizontal resolution (currently %u) must be divisible by the number of child sensors (%u)!", o.where.c_str(), fw, n);
            std::vector<float> s2c(16 * n), twv(16 * n); float nc = 0, fc = 0;
            for (uint32_t i = 0; i < n; ++i) {
                Obj &c = *sub_sensors[i];
                if (c.type != "perspective") AM_THROW("%s: batch child sensor \"%s\": only perspective children are supported on this path", c.where.c_str(), c.type.c_str());
                for (auto &k : c.children()) c.mark(k);
                double fov; int axis; fov_props(c, &fov, &axis);
                float near_ = (float) c.f("near_clip", 1e-2), far_ = (float) c.f("far_clip", 1e4); c.f("focus_distance", far_);
                if (i == 0) { nc = near_; fc = far_; } else if (near_ != nc || far_ != fc) AM_THROW("%s: batch children must share near_clip / far_clip on this path", c.where.c_str());
                float x_fov = (float) parse_fov_x(fov, axis, sub / (double) fh);
                M4 m = sample_to_camera((int) sub, (int) fh, x_fov, near_, far_, 0.f), t = m4_from_d(c.xform("to_world"));
                memcpy(&s2c[16 * i], m.m, 64); memcpy(&twv[16 * i], t.m, 64);
                c.check_unqueried();
            }
            if (o.b("reverse_y", false)) AM_THROW("%s: batch reverse_y is not supported", o.where.c_str());
            sc.set_sensor_explicit(n, sub, fh, s2c.data(), twv.data(), nc, fc, o.b("reverse_x", false));
        } else if (o.type == "perspective" || o.type == "thinlens") {
            // MVPathIntegrator throws for a non-MultiSensor (mvpath.cpp:15-17); the stock `path` integrator renders it: one view
            if (need_multi) AM_THROW("%s: MVPathIntegrator requires MultiSensor", o.where.c_str());
            GridSensorParams g; g.thinlens = o.type == "thinlens"; g.grid_x = g.grid_y = 1; g.res_x = fw; g.res_y = fh; g.reverse_y = false; g.single = true;
            g.cam_dist = 0.f; g.to_world = o.xform("to_world"); fov_props(o, &g.fov, &g.fov_axis);
            g.near_clip = (float) o.f("near_clip", 1e-2); g.far_clip = (float) o.f("far_clip", 1e4); g.focus_distance = (float) o.f("focus_distance", g.far_clip);
            if (g.thinlens) g.aperture_radius = (float) o.f_req("aperture_radius");
            sc.set_sensor_grid(g);
        } else AM_THROW("%s: sensor plugin \"%s\" is not part of the AMVPT path (supported: grid, batch, perspective, thinlens)", o.where.c_str(), o.type.c_str());
        o.check_unqueried();
    }

    // emitters that stand on their own in the scene (src/emitters/{point,constant,envmap}.cpp constructors)
    void emitter(const std::shared_ptr<Obj> &op, const std::string &base, Loader &ld) {
        Obj &o = *op;
        if (o.type == "area") return;           // declared at the scene level with an id and attached to its shape through <ref> (scene.cpp:46-48: surface emitters join the list with their shape)
        if (o.type == "point") {                // point.cpp:65-85
            float pos[3] = { 0, 0, 0 }, intensity[3];
            if (o.has("position")) { if (o.has("to_world")) AM_THROW("%s: Only one of the parameters 'position' and 'to_world' can be specified at the same time!'", o.where.c_str()); o.vec("position", pos); }
            else { M4 tw = m4_from_d(o.xform("to_world")); pos[0] = tw.m[0][3]; pos[1] = tw.m[1][3]; pos[2] = tw.m[2][3]; }
            o.color("intensity", 1.f, intensity); o.check_unqueried();
            sc.add_emitter_point(pos, intensity);
        } else if (o.type == "constant") {      // constant.cpp:55-68
            float radiance[3]; o.color("radiance", 1.f, radiance); o.check_unqueried();
            sc.add_emitter_constant(radiance);
        } else if (o.type == "envmap") {        // envmap.cpp:111-214
            if (o.b("mis_compensation", false)) AM_THROW("%s: envmap mis_compensation is not supported on this path", o.where.c_str());
            std::string fn = ld.resolve(o.s("filename", ""), base); uint32_t w = 0, h = 0; std::vector<float> rgb;
            load_image_rgb(fn, &w, &h, rgb);
            const float scale = (float) o.f("scale", 1.0); M4d tw = o.xform("to_world"); o.check_unqueried();
            sc.add_emitter_envmap(rgb.data(), w, h, scale, tw);
        } else AM_THROW("%s: emitter plugin \"%s\" is not part of the AMVPT path (supported: area, point, constant, envmap)", o.where.c_str(), o.type.c_str());
    }

    bool integrator(Obj &o_) {
        Obj *op = &o_;
        if (op->type == "moment") {             // src/integrators/moment.cpp: a wrapper that adds second-moment AOVs around one nested integrator; the radiance is the nested one's
            auto kids = op->children("integrator");
            if (kids.size() != 1) AM_THROW("%s: the moment integrator needs exactly one nested integrator on this path", op->where.c_str());
            op->mark(kids[0]); op->check_unqueried(); op = kids[0].get();
        }
        Obj &o = *op;
        IntegratorParams &ip = sc.integrator;
        if (o.type != "mvpath" && o.type != "path") AM_THROW("%s: integrator plugin \"%s\" is not part of the AMVPT path (supported: mvpath, path)", o.where.c_str(), o.type.c_str());
        ip.is_mvpath = o.type == "mvpath";
        ip.max_depth = o.i("max_depth", -1); ip.rr_depth = (uint32_t) o.i("rr_depth", 5); ip.hide_emitters = o.b("hide_emitters", false);
        if (ip.max_depth < 0 && ip.max_depth != -1) AM_THROW("%s: \"max_depth\" must be set to -1 (infinite) or a value >= 0", o.where.c_str());
        if ((long long) o.i("rr_depth", 5) <= 0) AM_THROW("%s: \"rr_depth\" must be set to a value greater than zero!", o.where.c_str());
        ip.timeout = (float) o.f("timeout", -1.0); ip.block_size = (uint32_t) o.i("block_size", 0); o.i("samples_per_pass", -1);
        if (ip.is_mvpath) {          // mvpath.h:120-129
            ip.spp_pass_lim = (uint32_t) o.i("spp_pass_lim", 16); ip.sa_reuse = o.b("sa_reuse", false); ip.sa_mis = o.b("sa_mis", false);
            ip.fast_mis = o.b("fast_mis", false); ip.reuse_count = (uint32_t) o.i("reuse_count", 1); ip.adaptive = (uint32_t) o.i("adaptive", 0);
            ip.force_eval = o.b("force_eval", true); ip.debug = o.b("debug", false);
        }
        o.check_unqueried();
        return ip.is_mvpath;
    }
};

Scene *build_scene(Loader &ld, const std::shared_ptr<Obj> &root, const std::string &base) {
    if (root->cls != "scene") AM_THROW("%s: the root element must be <scene>", root->where.c_str());
    std::set<Obj *> seen; ld.resolve_refs(*root, seen);
    std::unique_ptr<Scene> sc(new Scene()); Builder b{ *sc }; b.loader = &ld; b.base_dir = base;
    std::shared_ptr<Obj> integ, sens;
    for (auto &k : root->children()) {
        if (k->cls == "integrator") { if (integ) AM_THROW("%s: Only one integrator can be specified per scene.", k->where.c_str()); integ = k; }
        else if (k->cls == "sensor") { if (!sens) sens = k; else AM_THROW("%s: more than one sensor: only sensor 0 is rendered by this path; remove the others", k->where.c_str()); }
    }
    bool need_multi = integ ? b.integrator(*integ) : (sc->integrator.is_mvpath = false, false);   // scene.cpp:98-104: default integrator is `path`
    if (!sens) AM_THROW("%s: the scene has no sensor", root->where.c_str());
    b.sensor(sens, need_multi);
    for (auto &k : root->children()) {
        if (k->cls == "shape") b.shape(k, base, ld);
        else if (k->cls == "bsdf") b.bsdf(k);            // instantiated even if unreferenced, like the reference
        else if (k->cls == "emitter") b.emitter(k, base, ld);
        else if (k->cls == "integrator" || k->cls == "sensor" || k->cls == "wrap" || k->cls == "film" || k->cls == "sampler") continue;
        else AM_THROW("%s: <%s> is not part of the AMVPT path", k->where.c_str(), k->cls.c_str());
    }
    if (sc->meshes.empty()) AM_THROW("%s: the scene has no shapes", root->where.c_str());
    return sc.release();
}

} // namespace

Scene *load_xml_file(const std::string &path, const std::vector<std::pair<std::string, std::string>> &defines) {
    Loader ld; for (auto &d : defines) ld.params[d.first] = d.second;
    auto doc = ld.parse_file(path); Loader::check_root(*doc);
    std::string base = Loader::dir_of(path); ld.search.push_back(base);
    auto root = ld.object(*doc, base);
    return build_scene(ld, root, base);
}
Scene *load_xml_string(const std::string &xml, const std::string &base_dir, const std::vector<std::pair<std::string, std::string>> &defines) {
    Loader ld; for (auto &d : defines) ld.params[d.first] = d.second;
    auto doc = ld.parse_text(xml, "<string>"); Loader::check_root(*doc); ld.search.push_back(base_dir);
    auto root = ld.object(*doc, base_dir);
    return build_scene(ld, root, base_dir);
}

} // namespace amvpt

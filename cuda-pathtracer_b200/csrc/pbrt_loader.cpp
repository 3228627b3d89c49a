// pbrt_loader.cpp -- .pbrt scene ingest behind ptx_scene_load / ptx_parse_scene_file.
//
// Replaces the reference's loadPBRT (utils/pbrt_loader.h:178-422), which drives the third-party pbrtParser v2.4.0
// vendored under the reference's ext/ (not part of this repository).  This is an independent reader for the subset of
// the PBRT v3 format that loadPBRT consumes -- triangle meshes, the transform / attribute stacks, (named) materials,
// diffuse area lights, object instancing, Include -- with the conventions of that parser + loader pair, which the
// golden vectors in tests/golden/pbrt_*.npz (produced by the reference's own loadPBRT, tests/golden/make_golden_pbrt.py) pin:
//   * output order: the world's own shapes first (file order), then every ObjectInstance in file order
//     (pbrtParser's SingleLevelFlattener, impl/semantic/Scene.cpp:387-434); nested instances depth-first;
//   * a shape's vertices are taken to world/object space by the CTM at the Shape statement, normals by its inverse
//     transpose (impl/semantic/Geometry.cpp:238-262); instances then apply their CTM (pbrt_loader.h:63-80, 282-344);
//   * `Transform [m]` CONCATENATES like ConcatTransform in that parser version (impl/syntactic/Parser.inl:366-379);
//   * shading normal = "N" of the triangle's FIRST index when the mesh has normals, else the geometric normal
//     (pbrt_loader.h:318-323);
//   * materials fold to one albedo: bsdf = diffuse (1 - metallic) + specular metallic (pbrt_loader.h:43-47, 86-164), so
//     matte/plastic/substrate/uber/translucent -> Kd, glass -> Kt, mirror -> Kr, disney -> color (1-m) + color m^2,
//     metal -> (0,0,0) (its reflectance lands in `diffuse` with metallic = 1), no / unknown material -> 0.8 grey;
//   * AreaLightSource "diffuse" "rgb L" -> Le;
//   * the 2 000 000-triangle cap that swaps the scene for a 12-triangle bounding-box proxy (pbrt_loader.h:205-272) is
//     reproduced only when ptx_load_opts.triangle_cap asks for it (0 = no cap).
// `Shape "plymesh"` reads ascii / binary PLY files like the reference's parser does through rply (load_ply below).
// Not supported (reported as errors rather than silently dropped): "blackbody L", spectrum files.
// Non-triangle shapes (sphere, ...) are skipped like the reference does (pbrt_loader.h:288-292).
#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "common.h"
#include "loader.h"

namespace ptx {

namespace {

struct Affine {                 // columns vx, vy, vz of the linear part + translation (pbrtParser math.h:133-143)
    vec3 vx, vy, vz, p;
};
Affine affine_identity() { return Affine{mk3(1, 0, 0), mk3(0, 1, 0), mk3(0, 0, 1), mk3(0, 0, 0)}; }
// a.vx*b.x + a.vy*b.y + a.vz*b.z, summed left to right (math.h:163)
vec3 mat_vec(const Affine& a, vec3 b) { return ((a.vx * b.x) + (a.vy * b.y)) + (a.vz * b.z); }
Affine affine_mul(const Affine& a, const Affine& b) {         // math.h:166
    Affine r;
    r.vx = mat_vec(a, b.vx); r.vy = mat_vec(a, b.vy); r.vz = mat_vec(a, b.vz);
    r.p = mat_vec(a, b.p) + a.p;
    return r;
}
vec3 xfm_point(const Affine& m, vec3 p) { return mat_vec(m, p) + m.p; }      // math.h:165, 179
vec3 pcross(vec3 a, vec3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }   // math.h:169
float pdot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
vec3 xfm_normal(const Affine& m, vec3 n) {                                    // inverse_transpose(m.l) * n, math.h:173-181
    vec3 c0 = pcross(m.vy, m.vz), c1 = pcross(m.vz, m.vx), c2 = pcross(m.vx, m.vy);
    float k = 1 / pdot(m.vx, pcross(m.vy, m.vz));
    Affine it; it.vx = c0 * k; it.vy = c1 * k; it.vz = c2 * k; it.p = mk3(0, 0, 0);
    return mat_vec(it, n);
}
Affine affine_rotate(vec3 axis, float r) {                                    // math.h:183-190
    vec3 u = axis * (1 / sqrtf(pdot(axis, axis)));
    float s = sinf(r), c = cosf(r);
    Affine a;
    a.vx = mk3(u.x * u.x + (1 - u.x * u.x) * c, u.x * u.y * (1 - c) + u.z * s, u.x * u.z * (1 - c) - u.y * s);
    a.vy = mk3(u.x * u.y * (1 - c) - u.z * s, u.y * u.y + (1 - u.y * u.y) * c, u.y * u.z * (1 - c) + u.x * s);
    a.vz = mk3(u.x * u.z * (1 - c) + u.y * s, u.y * u.z * (1 - c) - u.x * s, u.z * u.z + (1 - u.z * u.z) * c);
    a.p = mk3(0, 0, 0);
    return a;
}
// the reference's transformPoint / transformNormal for instances (pbrt_loader.h:63-80): row sums left to right
vec3 ref_transform_point(const Affine& x, vec3 p) {
    return mk3(x.vx.x * p.x + x.vy.x * p.y + x.vz.x * p.z + x.p.x, x.vx.y * p.x + x.vy.y * p.y + x.vz.y * p.z + x.p.y,
               x.vx.z * p.x + x.vy.z * p.y + x.vz.z * p.z + x.p.z);
}
vec3 ref_transform_normal(const Affine& x, vec3 n) {
    return unit(mk3(x.vx.x * n.x + x.vy.x * n.y + x.vz.x * n.z, x.vx.y * n.x + x.vy.y * n.y + x.vz.y * n.z,
                    x.vx.z * n.x + x.vy.z * n.y + x.vz.z * n.z));
}

struct Mat {                    // the reference's PBRTMaterial after convertMaterial, already folded
    float bsdf[3] = {0.8f, 0.8f, 0.8f};
    bool mirror = false;
};

struct Mesh {
    std::vector<vec3> P, N;     // already transformed by the shape's CTM
    std::vector<int> idx;
    std::shared_ptr<Mat> mat;   // null => default grey
    float Le[3] = {0, 0, 0};
};

struct Object {
    std::vector<Mesh> meshes;
    std::vector<std::pair<std::shared_ptr<Object>, Affine>> instances;
};

struct GState {
    Affine ctm = affine_identity();
    std::shared_ptr<Mat> material;
    std::vector<float> area_L;  // every AreaLightSource in scope, 3 floats each; a shape uses the FIRST (impl/semantic/Geometry.cpp:435-441)
    std::map<std::string, std::shared_ptr<Mat>> named;
};

// --------------------------------------------------------------------------------------------- lexer
struct Token {
    enum Kind { End, Ident, String, Number, LBracket, RBracket } kind = End;
    std::string text;
};

struct Lexer {
    std::string data;
    size_t pos = 0;
    std::string file;
    int line = 1;
    bool load(const std::string& path) {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) return false;
        fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
        data.resize(n > 0 ? (size_t)n : 0);
        if (n > 0 && fread(&data[0], 1, (size_t)n, f) != (size_t)n) { fclose(f); return false; }
        fclose(f);
        file = path;
        return true;
    }
    Token next() {
        Token t;
        for (;;) {
            while (pos < data.size() && (data[pos] == ' ' || data[pos] == '\t' || data[pos] == '\r' || data[pos] == '\n' || data[pos] == '\0')) { if (data[pos] == '\n') line++; pos++; }
            if (pos < data.size() && data[pos] == '#') { while (pos < data.size() && data[pos] != '\n') pos++; continue; }
            break;
        }
        if (pos >= data.size()) return t;
        char c = data[pos];
        if (c == '[') { pos++; t.kind = Token::LBracket; return t; }
        if (c == ']') { pos++; t.kind = Token::RBracket; return t; }
        if (c == '"') {
            size_t e = data.find('"', pos + 1);
            if (e == std::string::npos) e = data.size();
            t.kind = Token::String; t.text = data.substr(pos + 1, e - pos - 1);
            pos = e + 1;
            return t;
        }
        size_t s = pos;
        while (pos < data.size() && !strchr(" \t\r\n[]\"#", data[pos])) pos++;
        t.text = data.substr(s, pos - s);
        char f = t.text[0];
        t.kind = ((f >= '0' && f <= '9') || f == '-' || f == '+' || f == '.') ? Token::Number : Token::Ident;
        return t;
    }
};

struct Param {
    std::string type, name;
    std::vector<float> nums;
    std::vector<long long> ints;       // "integer" parameters, parsed as integers: float32 is exact only up to 2^24, and a mesh with
                                       // more than 16.7 M vertices would otherwise get its indices silently rounded
    std::vector<std::string> strs;
};


// ---------------------------------------------------------------------------------------------------------------
// PLY reader for `Shape "plymesh"`: what the reference's parser extracts through rply (ext/pbrtparser/pbrtParser/impl/semantic/
// Geometry.cpp:25-66 callbacks, :78-207 ply::parse): vertex x / y / z (required), nx / ny / nz (all three or none), the
// face list `vertex_indices` | `vertex_index`; every value goes through a double and is cast to float / int; a face whose
// list length is not 3 is an error ("only triangles are supported"); a file without vertices or faces is an error.
// Formats: ascii, binary_little_endian, binary_big_endian; every scalar type; other elements and properties are skipped.
// ---------------------------------------------------------------------------------------------------------------
struct PlyProp {
    std::string name;
    int type = 0, count_type = -1;      // count_type >= 0: list property
};
struct PlyElem {
    std::string name;
    long long count = 0;
    std::vector<PlyProp> props;
};
int ply_type(const std::string& t) {
    static const char* names[][2] = {{"char", "int8"}, {"uchar", "uint8"}, {"short", "int16"}, {"ushort", "uint16"},
                                     {"int", "int32"}, {"uint", "uint32"}, {"float", "float32"}, {"double", "float64"}};
    for (int i = 0; i < 8; i++) if (t == names[i][0] || t == names[i][1]) return i;
    return -1;
}
const int kPlySize[8] = {1, 1, 2, 2, 4, 4, 4, 8};

struct PlyReader {
    const unsigned char* p; const unsigned char* end;
    int format = 0;                      // 0 ascii, 1 little endian, 2 big endian
    bool ok = true;
    double scalar(int type) {
        if (format == 0) {
            while (p < end && isspace(*p)) p++;
            if (p >= end) { ok = false; return 0.0; }
            char* e = nullptr;
            double v = strtod(reinterpret_cast<const char*>(p), &e);      // the buffer is NUL-terminated by load_ply
            if (e == reinterpret_cast<const char*>(p)) { ok = false; return 0.0; }
            p = reinterpret_cast<const unsigned char*>(e);
            return v;
        }
        const int n = kPlySize[type];
        if (end - p < n) { ok = false; return 0.0; }
        unsigned char b[8];
        for (int i = 0; i < n; i++) b[i] = format == 1 ? p[i] : p[n - 1 - i];      // to little endian (the host's order)
        p += n;
        switch (type) {
            case 0: { int8_t v; memcpy(&v, b, 1); return v; }
            case 1: { uint8_t v; memcpy(&v, b, 1); return v; }
            case 2: { int16_t v; memcpy(&v, b, 2); return v; }
            case 3: { uint16_t v; memcpy(&v, b, 2); return v; }
            case 4: { int32_t v; memcpy(&v, b, 4); return v; }
            case 5: { uint32_t v; memcpy(&v, b, 4); return v; }
            case 6: { float v; memcpy(&v, b, 4); return v; }
            default: { double v; memcpy(&v, b, 8); return v; }
        }
    }
};

bool load_ply(const std::string& path, std::vector<vec3>& P, std::vector<vec3>& N, std::vector<int>& idx, std::string& err) {
    std::vector<unsigned char> buf;
    {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) { err = "Couldn't open PLY file " + path; return false; }
        fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
        buf.resize((size_t)std::max(sz, 0l) + 1);
        size_t got = sz > 0 ? fread(buf.data(), 1, (size_t)sz, f) : 0;
        fclose(f);
        buf.resize(got + 1); buf[got] = 0;
    }
    const unsigned char* p = buf.data(); const unsigned char* end = buf.data() + buf.size() - 1;
    auto line = [&](std::string& out) {
        if (p >= end) return false;
        const unsigned char* e = p;
        while (e < end && *e != '\n') e++;
        out.assign(reinterpret_cast<const char*>(p), e - p);
        if (!out.empty() && out.back() == '\r') out.pop_back();
        p = e < end ? e + 1 : end;
        return true;
    };
    std::string ln;
    if (!line(ln) || ln != "ply") { err = "Unable to read the header of PLY file " + path; return false; }
    std::vector<PlyElem> elems;
    int format = -1;
    bool header_done = false;
    while (line(ln)) {
        std::istringstream is(ln);
        std::string kw; is >> kw;
        if (kw == "end_header") { header_done = true; break; }
        if (kw == "format") {
            std::string f; is >> f;
            format = f == "ascii" ? 0 : (f == "binary_little_endian" ? 1 : (f == "binary_big_endian" ? 2 : -1));
        } else if (kw == "element") {
            PlyElem e; is >> e.name >> e.count;
            elems.push_back(e);
        } else if (kw == "property" && !elems.empty()) {
            PlyProp pr; std::string t; is >> t;
            if (t == "list") { std::string ct, it; is >> ct >> it >> pr.name; pr.count_type = ply_type(ct); pr.type = ply_type(it); if (pr.count_type < 0) pr.type = -1; }
            else { pr.type = ply_type(t); is >> pr.name; }
            if (pr.type < 0) { err = "Unable to read the header of PLY file " + path; return false; }
            elems.back().props.push_back(pr);
        }
    }
    if (!header_done || format < 0) { err = "Unable to read the header of PLY file " + path; return false; }
    long long vertex_count = 0, face_count = 0;
    bool has_pos[3] = {false, false, false}, has_nor[3] = {false, false, false}, has_indices = false;
    for (const PlyElem& e : elems) {
        if (e.name == "vertex") {
            vertex_count = e.count;
            for (const PlyProp& pr : e.props) {
                if (pr.name == "x") has_pos[0] = true; else if (pr.name == "y") has_pos[1] = true; else if (pr.name == "z") has_pos[2] = true;
                else if (pr.name == "nx") has_nor[0] = true; else if (pr.name == "ny") has_nor[1] = true; else if (pr.name == "nz") has_nor[2] = true;
            }
            if (!(has_pos[0] && has_pos[1] && has_pos[2])) { err = path + ": Vertex coordinate property not found!"; return false; }
        } else if (e.name == "face") {
            face_count = e.count;
            for (const PlyProp& pr : e.props) if (pr.name == "vertex_index" || pr.name == "vertex_indices") has_indices = true;
        }
    }
    if (vertex_count <= 0 || face_count <= 0) { err = path + ": PLY file is invalid! No face/vertex elements found!"; return false; }
    const bool has_normals = has_nor[0] && has_nor[1] && has_nor[2];
    P.assign((size_t)vertex_count, mk3(0.f, 0.f, 0.f));
    N.assign(has_normals ? (size_t)vertex_count : 0, mk3(0.f, 0.f, 0.f));
    idx.assign(has_indices ? (size_t)face_count * 3 : 0, 0);
    PlyReader rd; rd.p = p; rd.end = end; rd.format = format;
    for (const PlyElem& e : elems) {
        const bool is_vertex = e.name == "vertex", is_face = e.name == "face";
        for (long long i = 0; i < e.count; i++)
            for (const PlyProp& pr : e.props) {
                if (pr.count_type < 0) {
                    const double v = rd.scalar(pr.type);
                    if (is_vertex) {
                        const float fv = (float)v;
                        if (pr.name == "x") P[(size_t)i].x = fv; else if (pr.name == "y") P[(size_t)i].y = fv; else if (pr.name == "z") P[(size_t)i].z = fv;
                        else if (has_normals && pr.name == "nx") N[(size_t)i].x = fv; else if (has_normals && pr.name == "ny") N[(size_t)i].y = fv;
                        else if (has_normals && pr.name == "nz") N[(size_t)i].z = fv;
                    }
                } else {
                    const long long len = (long long)rd.scalar(pr.count_type);
                    const bool mine = is_face && (pr.name == "vertex_index" || pr.name == "vertex_indices");
                    if (mine && len != 3) { err = "Found face with vertex count different from 3, only triangles are supported"; return false; }
                    if (len < 0 || !rd.ok) { err = path + ": unable to read the contents of PLY file"; return false; }
                    for (long long k = 0; k < len; k++) {
                        const double v = rd.scalar(pr.type);
                        if (mine) idx[(size_t)i * 3 + (size_t)k] = (int)v;
                    }
                }
                if (!rd.ok) { err = path + ": unable to read the contents of PLY file"; return false; }
            }
    }
    return true;
}
}  // namespace

namespace {

struct Parser {
    std::vector<std::unique_ptr<Lexer>> lexers;     // Include stack
    std::string root_dir;
    Token peeked; bool has_peek = false;
    std::string err;

    GState gs;
    std::vector<GState> attr_stack;
    std::vector<Affine> xfm_stack;
    std::shared_ptr<Object> world = std::make_shared<Object>();
    std::vector<std::shared_ptr<Object>> object_stack;
    std::map<std::string, std::shared_ptr<Object>> named_objects;
    bool mirror_ext = false;

    Token next() {
        if (has_peek) { has_peek = false; return peeked; }
        while (!lexers.empty()) {
            Token t = lexers.back()->next();
            if (t.kind != Token::End) return t;
            lexers.pop_back();
        }
        return Token();
    }
    Token peek() { if (!has_peek) { peeked = next(); has_peek = true; } return peeked; }
    bool fail(const std::string& m) {
        if (err.empty()) err = m + (lexers.empty() ? "" : " (" + lexers.back()->file + ":" + std::to_string(lexers.back()->line) + ")");
        return false;
    }
    bool open(const std::string& path) {
        if (lexers.size() >= 32) return fail("Include nested deeper than 32 levels (a file that includes itself?): " + path);
        std::unique_ptr<Lexer> lx(new Lexer());
        if (!lx->load(path)) return fail("cannot open " + path);
        lexers.push_back(std::move(lx));
        return true;
    }
    static float to_float(const std::string& s) { return (float)strtod(s.c_str(), nullptr); }

    bool read_floats(int n, float* out) {
        for (int i = 0; i < n; i++) {
            Token t = next();
            if (t.kind != Token::Number) return fail("number expected, got '" + t.text + "'");
            out[i] = to_float(t.text);
        }
        return true;
    }
    // "type name" value | "type name" [ values ]  ... until the next directive
    bool read_params(std::vector<Param>& out) {
        for (;;) {
            Token t = peek();
            if (t.kind != Token::String) return true;
            next();
            Param p;
            size_t sp = t.text.find_first_of(" \t");
            if (sp == std::string::npos) return fail("malformed parameter declaration '" + t.text + "'");
            p.type = t.text.substr(0, sp);
            size_t ns = t.text.find_first_not_of(" \t", sp);
            p.name = ns == std::string::npos ? "" : t.text.substr(ns);
            Token v = next();
            const bool integer = p.type == "integer";
            auto take = [&](const Token& x) {
                if (x.kind != Token::Number) p.strs.push_back(x.text);
                else if (integer) p.ints.push_back(strtoll(x.text.c_str(), nullptr, 10));
                else p.nums.push_back(to_float(x.text));
            };
            if (v.kind == Token::LBracket) {
                for (;;) {
                    Token x = next();
                    if (x.kind == Token::RBracket) break;
                    if (x.kind == Token::End) return fail("unterminated '[' in parameter '" + p.name + "'");
                    take(x);
                }
            } else if (v.kind == Token::Number || v.kind == Token::String || v.kind == Token::Ident) take(v);
            else return fail("value expected for parameter '" + p.name + "'");
            out.push_back(std::move(p));
        }
    }
    static const Param* find(const std::vector<Param>& ps, const char* name) {
        for (auto& p : ps) if (p.name == name) return &p;
        return nullptr;
    }
    bool rgb(const std::vector<Param>& ps, const char* name, float out[3]) {        // leaves `out` untouched when absent / textured
        const Param* p = find(ps, name);
        if (!p) return true;
        if (p->type == "rgb" || p->type == "color") {
            if (p->nums.size() < 3) return fail(std::string("rgb parameter '") + name + "' needs 3 values");
            out[0] = p->nums[0]; out[1] = p->nums[1]; out[2] = p->nums[2];
        } else if (p->type == "float" && !p->nums.empty()) out[0] = out[1] = out[2] = p->nums[0];
        else if (p->type == "texture") out[0] = out[1] = out[2] = 1.f;      // textured colours read as 1 (impl/semantic/Materials.cpp:50-74, 200-203)
        else if (p->type == "blackbody") return fail("blackbody spectra are not supported");
        else if (p->type == "spectrum") {
            if (!p->strs.empty()) return fail("spectrum files are not supported");
            // inline (lambda, value) pairs keep the material's default, like a texture reference does
        }
        return true;
    }
    float scalar(const std::vector<Param>& ps, const char* name, float dflt) {
        const Param* p = find(ps, name);
        return (p && p->type == "float" && !p->nums.empty()) ? p->nums[0] : dflt;
    }

    // convertMaterial + getBSDF of the reference (pbrt_loader.h:43-47, 86-164) on pbrtParser's defaults (Scene.h:425-721)
    std::shared_ptr<Mat> make_material(const std::string& type, const std::vector<Param>& ps) {
        auto m = std::make_shared<Mat>();
        float diffuse[3] = {0.8f, 0.8f, 0.8f}, specular[3] = {0, 0, 0}, metallic = 0.f;
        auto set3 = [](float* d, float a) { d[0] = d[1] = d[2] = a; };
        if (type == "matte") { set3(diffuse, 0.5f); if (!rgb(ps, "Kd", diffuse)) return nullptr; }
        else if (type == "plastic") { set3(diffuse, 0.25f); set3(specular, 0.25f); if (!rgb(ps, "Kd", diffuse) || !rgb(ps, "Ks", specular)) return nullptr; }
        else if (type == "mirror") { float kr[3]; set3(kr, 0.9f); if (!rgb(ps, "Kr", kr)) return nullptr; set3(diffuse, 0.f); memcpy(specular, kr, 12); metallic = 1.f; m->mirror = true; }
        else if (type == "metal") { metallic = 1.f; /* reflectance goes to `diffuse`, which metallic = 1 multiplies by 0 */ }
        else if (type == "glass") { set3(diffuse, 1.f); if (!rgb(ps, "Kt", diffuse)) return nullptr; }
        else if (type == "substrate") { set3(diffuse, 0.5f); set3(specular, 0.5f); if (!rgb(ps, "Kd", diffuse) || !rgb(ps, "Ks", specular)) return nullptr; }
        else if (type == "uber") { set3(diffuse, 0.25f); set3(specular, 0.25f); if (!rgb(ps, "Kd", diffuse) || !rgb(ps, "Ks", specular)) return nullptr; }
        else if (type == "translucent") { set3(diffuse, 0.25f); if (!rgb(ps, "Kd", diffuse)) return nullptr; }
        else if (type == "disney") {
            set3(diffuse, 0.5f); if (!rgb(ps, "color", diffuse)) return nullptr;
            metallic = scalar(ps, "metallic", 0.f);
            for (int k = 0; k < 3; k++) specular[k] = diffuse[k] * metallic;
        }
        // anything else ("", mix, fourier, hair, subsurface, kdsubsurface, none): the loader's default grey
        for (int k = 0; k < 3; k++) m->bsdf[k] = diffuse[k] * (1.0f - metallic) + specular[k] * metallic;
        return m;
    }

    Object& current_object() { return object_stack.empty() ? *world : *object_stack.back(); }

    bool shape() {
        Token ty = next();
        if (ty.kind != Token::String) return fail("Shape: type string expected");
        std::vector<Param> ps;
        if (!read_params(ps)) return false;
        if (ty.text == "plymesh") {                                     // emitPlyMesh, Geometry.cpp:220-236
            const Param* F = find(ps, "filename");
            if (!F || F->strs.empty()) return fail("plymesh without a 'filename'");
            Mesh m;
            m.mat = gs.material;
            if (gs.area_L.size() >= 3) memcpy(m.Le, gs.area_L.data(), 12);
            std::string perr;
            if (!load_ply(root_dir + "/" + F->strs[0], m.P, m.N, m.idx, perr)) return fail(perr);      // makeGlobalFileName: basePath + name
            for (vec3& v : m.P) v = xfm_point(gs.ctm, v);
            for (vec3& v : m.N) v = xfm_normal(gs.ctm, v);
            for (int v : m.idx) if (v < 0 || (size_t)v >= m.P.size()) return fail("plymesh index out of range");
            current_object().meshes.push_back(std::move(m));
            return true;
        }
        if (ty.text != "trianglemesh") return true;                     // skipped, like the reference's non-triangle shapes
        const Param* P = find(ps, "P"); const Param* I = find(ps, "indices"); const Param* N = find(ps, "N");
        if (!P || P->nums.size() % 3) return fail("trianglemesh without a valid 'P'");
        Mesh m;
        m.mat = gs.material;
        if (gs.area_L.size() >= 3) memcpy(m.Le, gs.area_L.data(), 12);
        size_t nv = P->nums.size() / 3;
        m.P.resize(nv);
        for (size_t i = 0; i < nv; i++) m.P[i] = xfm_point(gs.ctm, mk3(P->nums[3 * i], P->nums[3 * i + 1], P->nums[3 * i + 2]));
        if (N && N->nums.size() % 3 == 0) {
            m.N.resize(N->nums.size() / 3);
            for (size_t i = 0; i < m.N.size(); i++) m.N[i] = xfm_normal(gs.ctm, mk3(N->nums[3 * i], N->nums[3 * i + 1], N->nums[3 * i + 2]));
        }
        if (I) {
            // "integer indices" arrive in I->ints; a file that declares them as "float indices" still works up to 2^24
            const bool as_int = I->type == "integer";
            size_t nt = (as_int ? I->ints.size() : I->nums.size()) / 3;
            m.idx.resize(nt * 3);
            for (size_t i = 0; i < nt * 3; i++) {
                long long v = as_int ? I->ints[i] : (long long)I->nums[i];
                if (v < 0 || (unsigned long long)v >= nv) return fail("trianglemesh index out of range");
                m.idx[i] = (int)v;
            }
        }
        current_object().meshes.push_back(std::move(m));
        return true;
    }

    bool directive(const Token& t) {
        const std::string& d = t.text;
        float f[16];
        if (d == "AttributeBegin") { attr_stack.push_back(gs); return true; }
        if (d == "AttributeEnd") { if (attr_stack.empty()) return fail("unbalanced AttributeEnd"); gs = attr_stack.back(); attr_stack.pop_back(); return true; }
        if (d == "TransformBegin") { xfm_stack.push_back(gs.ctm); return true; }
        if (d == "TransformEnd") { if (xfm_stack.empty()) return fail("unbalanced TransformEnd"); gs.ctm = xfm_stack.back(); xfm_stack.pop_back(); return true; }
        if (d == "Identity") { gs.ctm = affine_identity(); return true; }
        if (d == "Translate") { if (!read_floats(3, f)) return false; Affine a = affine_identity(); a.p = mk3(f[0], f[1], f[2]); gs.ctm = affine_mul(gs.ctm, a); return true; }
        if (d == "Scale") { if (!read_floats(3, f)) return false; Affine a = affine_identity(); a.vx = mk3(f[0], 0, 0); a.vy = mk3(0, f[1], 0); a.vz = mk3(0, 0, f[2]); gs.ctm = affine_mul(gs.ctm, a); return true; }
        if (d == "Rotate") { if (!read_floats(4, f)) return false; gs.ctm = affine_mul(gs.ctm, affine_rotate(mk3(f[1], f[2], f[3]), f[0] * (float)M_PI / 180.f)); return true; }
        if (d == "Transform" || d == "ConcatTransform") {
            if (next().kind != Token::LBracket) return fail(d + ": '[' expected");
            if (!read_floats(16, f)) return false;
            if (next().kind != Token::RBracket) return fail(d + ": ']' expected");
            Affine a; a.vx = mk3(f[0], f[1], f[2]); a.vy = mk3(f[4], f[5], f[6]); a.vz = mk3(f[8], f[9], f[10]); a.p = mk3(f[12], f[13], f[14]);
            gs.ctm = affine_mul(gs.ctm, a);
            return true;
        }
        if (d == "LookAt") return read_floats(9, f);                              // camera only: the reference never reads scene->cameras
        if (d == "ReverseOrientation") return true;
        if (d == "CoordinateSystem" || d == "CoordSysTransform" || d == "ActiveTransform" || d == "NamedMedium") { next(); return true; }
        if (d == "MediumInterface") { next(); if (peek().kind == Token::String) next(); return true; }
        if (d == "WorldBegin") { gs = GState(); attr_stack.clear(); xfm_stack.clear(); return true; }
        if (d == "WorldEnd") return true;
        if (d == "Include") {
            Token fn = next();
            if (fn.kind != Token::String) return fail("Include: file name expected");
            std::string path = fn.text.size() && fn.text[0] == '/' ? fn.text : root_dir + "/" + fn.text;
            has_peek = false;
            return open(path);
        }
        if (d == "Shape") return shape();
        if (d == "Material") {
            Token ty = next(); std::vector<Param> ps;
            if (ty.kind != Token::String || !read_params(ps)) return fail("Material: malformed");
            gs.material = make_material(ty.text, ps);
            return gs.material != nullptr;
        }
        if (d == "MakeNamedMaterial") {
            Token nm = next(); std::vector<Param> ps;
            if (nm.kind != Token::String || !read_params(ps)) return fail("MakeNamedMaterial: malformed");
            const Param* ty = find(ps, "type");
            if (!ty || ty->strs.empty()) return fail("named material '" + nm.text + "' without a 'type'");
            auto m = make_material(ty->strs[0], ps);
            if (!m) return false;
            gs.named[nm.text] = m;
            return true;
        }
        if (d == "NamedMaterial") {
            Token nm = next();
            auto it = gs.named.find(nm.text);
            gs.material = it == gs.named.end() ? nullptr : it->second;
            return true;
        }
        if (d == "AreaLightSource") {
            Token ty = next(); std::vector<Param> ps;
            if (ty.kind != Token::String || !read_params(ps)) return fail("AreaLightSource: malformed");
            float L[3] = {1.f, 1.f, 1.f};                                          // pbrt's default L
            if (!rgb(ps, "L", L)) return false;
            gs.area_L.insert(gs.area_L.end(), L, L + 3);
            return true;
        }
        if (d == "ObjectBegin") {
            Token nm = next();
            auto& o = named_objects[nm.text];
            if (!o) o = std::make_shared<Object>();
            object_stack.push_back(o);
            return true;
        }
        if (d == "ObjectEnd") { if (object_stack.empty()) return fail("unbalanced ObjectEnd"); object_stack.pop_back(); return true; }
        if (d == "ObjectInstance") {
            Token nm = next();
            auto it = named_objects.find(nm.text);
            if (it == named_objects.end()) return fail("ObjectInstance of unknown object '" + nm.text + "'");
            current_object().instances.push_back({it->second, gs.ctm});
            return true;
        }
        // everything else carries a type string (or two, or three for Texture) and a parameter list we do not need:
        // Camera, Film, Sampler, Integrator, PixelFilter, Accelerator, LightSource, Texture, MakeNamedMedium, Option, ...
        static const char* skip[] = {"Camera", "Film", "Sampler", "Integrator", "SurfaceIntegrator", "VolumeIntegrator", "PixelFilter", "Accelerator",
                                     "LightSource", "Texture", "MakeNamedMedium", "Option", "Renderer", "Volume", "ColorSpace", "Attribute"};
        for (const char* s : skip)
            if (d == s) {
                std::vector<Param> ps;
                while (peek().kind == Token::String && peek().text.find(' ') == std::string::npos) next();     // type / name strings
                return read_params(ps);
            }
        return fail("unexpected token '" + d + "'");
    }

    bool run(const std::string& path) {
        size_t sl = path.find_last_of('/');
        root_dir = sl == std::string::npos ? "." : path.substr(0, sl);
        if (!open(path)) return false;
        for (;;) {
            Token t = next();
            if (t.kind == Token::End) return err.empty();
            if (t.kind != Token::Ident) return fail("directive expected, got '" + t.text + "'");
            if (!directive(t)) return false;
        }
    }
};

void emit_mesh(const Mesh& m, const Affine* inst, bool mirror_ext, std::vector<ptx_primitive>& out) {
    const bool has_normals = m.N.size() >= m.P.size();
    for (size_t t = 0; t + 2 < m.idx.size(); t += 3) {
        vec3 v0 = m.P[m.idx[t]], v1 = m.P[m.idx[t + 1]], v2 = m.P[m.idx[t + 2]];
        vec3 n;
        if (inst) { v0 = ref_transform_point(*inst, v0); v1 = ref_transform_point(*inst, v1); v2 = ref_transform_point(*inst, v2); }
        if (has_normals) n = inst ? ref_transform_normal(*inst, m.N[m.idx[t]]) : unit(m.N[m.idx[t]]);
        else n = unit(cross(v1 - v0, v2 - v0));
        ptx_primitive p;
        memset(&p, 0, sizeof p);
        p.type = 0;
        st3(p.v[0], v0); st3(p.v[1], v1); st3(p.v[2], v2);
        st3(p.normal, n);
        const float grey[3] = {0.8f, 0.8f, 0.8f};
        memcpy(p.bsdf, m.mat ? m.mat->bsdf : grey, 12);
        memcpy(p.Le, m.Le, 12);
        p.material_type = (mirror_ext && m.mat && m.mat->mirror) ? 1 : 0;
        out.push_back(p);
    }
}

// SingleLevelFlattener order: an object's own meshes, then its instances depth-first
void flatten(const Object& o, const Affine& xfm, bool mirror_ext, std::vector<ptx_primitive>& out, int depth) {
    if (depth > 64) return;
    // the reference applies the (identity) instance transform to the world's own shapes too: x*1 + y*0 + z*0 + 0 is exact
    for (const Mesh& m : o.meshes) emit_mesh(m, &xfm, mirror_ext, out);
    for (const auto& in : o.instances) flatten(*in.first, affine_mul(xfm, in.second), mirror_ext, out, depth + 1);
}

size_t count_tris(const Object& o, int depth) {
    if (depth > 64) return 0;
    size_t n = 0;
    for (const Mesh& m : o.meshes) n += m.idx.size() / 3;
    for (const auto& in : o.instances) n += count_tris(*in.first, depth + 1);
    return n;
}

}  // namespace

bool load_pbrt(const std::string& path, bool mirror_ext, long long triangle_cap, std::vector<ptx_primitive>& out, std::string& err) {
    Parser ps;
    ps.mirror_ext = mirror_ext;
    if (!ps.run(path)) { err = "PBRT: " + ps.err; return false; }
    out.clear();
    if (triangle_cap > 0 && count_tris(*ps.world, 0) > (size_t)triangle_cap) {
        // the reference's bounding-box proxy (pbrt_loader.h:225-272): 12 triangles, albedo (0.8, 0.2, 0.2)
        std::vector<ptx_primitive> all;
        flatten(*ps.world, affine_identity(), false, all, 0);
        aabb b = aabb_empty();
        for (auto& p : all) for (int k = 0; k < 3; k++) aabb_grow(b, ld3(p.v[k]));
        vec3 c[8];
        for (int i = 0; i < 8; i++) c[i] = mk3((i & 4) ? b.hi.x : b.lo.x, (i & 2) ? b.hi.y : b.lo.y, (i & 1) ? b.hi.z : b.lo.z);
        const int quads[6][4] = {{0, 1, 3, 2}, {4, 6, 7, 5}, {0, 4, 5, 1}, {2, 3, 7, 6}, {0, 2, 6, 4}, {1, 5, 7, 3}};
        for (auto& q : quads) {
            vec3 a = c[q[0]], bb = c[q[1]], cc = c[q[2]], d = c[q[3]];
            vec3 n = unit(cross(bb - a, cc - a));
            for (int h = 0; h < 2; h++) {
                ptx_primitive p; memset(&p, 0, sizeof p);
                st3(p.v[0], a); st3(p.v[1], h ? cc : bb); st3(p.v[2], h ? d : cc); st3(p.normal, n);
                p.bsdf[0] = 0.8f; p.bsdf[1] = 0.2f; p.bsdf[2] = 0.2f;
                out.push_back(p);
            }
        }
        return true;
    }
    flatten(*ps.world, affine_identity(), mirror_ext, out, 0);
    if (out.empty()) { err = "PBRT: no triangles found in " + path; return false; }        // pbrt_loader.h:402-405
    return true;
}

bool load_pbrt_instanced(const std::string& path, bool mirror_ext, InstancedScene& out, std::string& err) {
    Parser ps;
    ps.mirror_ext = mirror_ext;
    err.clear();
    if (!ps.run(path)) { err = "PBRT: " + ps.err; return false; }
    out.world.clear(); out.objects.clear(); out.instances.clear();
    const Object& w = *ps.world;
    if (w.instances.empty()) return false;
    for (const auto& in : w.instances) if (!in.first->instances.empty()) return false;        // nested instancing: flatten
    const Affine id = affine_identity();
    for (const Mesh& m : w.meshes) emit_mesh(m, &id, mirror_ext, out.world);
    std::map<const Object*, int> index;
    for (const auto& in : w.instances) {
        const Object* o = in.first.get();
        auto it = index.find(o);
        if (it == index.end()) {
            it = index.emplace(o, (int)out.objects.size()).first;
            out.objects.emplace_back();
            std::vector<ptx_primitive>& prims = out.objects.back();
            for (const Mesh& m : o->meshes) {
                const size_t first = prims.size();
                emit_mesh(m, nullptr, mirror_ext, prims);
                if (m.N.size() < m.P.size())                                                 // no vertex normals: the instance's
                    for (size_t k = first; k < prims.size(); k++) prims[k].normal[0] = prims[k].normal[1] = prims[k].normal[2] = NAN;   // geometric normal is computed in world space
            }
        }
        if (out.objects[(size_t)it->second].empty()) continue;                               // an object without triangles contributes nothing
        InstancedScene::Inst ii;
        ii.object = it->second;
        const Affine& a = in.second;
        st3(ii.m + 0, a.vx); st3(ii.m + 3, a.vy); st3(ii.m + 6, a.vz); st3(ii.m + 9, a.p);
        out.instances.push_back(ii);
    }
    if (out.instances.empty()) return false;
    return true;
}

}  // namespace ptx

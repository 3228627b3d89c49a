// loader.cpp -- Wavefront .obj / .mtl ingest into the reference's primitive numbering.
//
// Behavioural contract (utils/file_manager.h:39-79 loadMTL, :93-273 loadOBJ):
//   * lines that are empty or start with '#', 'o' or 's' are skipped (:120);
//   * `v x y z`, `vn x y z` (normals are normalised on load, :141), `mtllib` (relative to the .obj
//     directory), `usemtl` (unknown name -> default grey 0.8, no emission, :150-156);
//   * `f` tokens are v, v/vt, v//vn or v/vt/vn; tokens whose vertex index does not parse (e.g. a
//     trailing "# Top" comment) are dropped and the rest of the line is still used (:166-170);
//   * exactly 3 surviving tokens -> triangle, 4 -> quad, anything else is ignored (:193-249);
//     indices are 1-based, 0 / out of range rejects the face; negative (relative) indices are
//     not supported by the reference and are rejected here too;
//   * the stored normal is the vn of the FIRST face vertex when present, else the geometric normal
//     unit((v1-v0) x (v2-v0)); quads use unit((v10-v00) x (v01-v00)) (:206-210, :235-237; quad.h:27-29);
//   * MTL: only newmtl / Kd / Ke are read (:56-70).  Extension (opt-in, ptx_load_opts.mirror_from_mtl):
//     `illum 3` or `illum 5` marks a perfect mirror with reflectance Ks.
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <string>
#include <vector>

#include "../../include/ptx.h"
#include "common.h"
#include "loader.h"

namespace ptx {

namespace {

struct Cursor {
    const char* p;
    explicit Cursor(const std::string& s) : p(s.c_str()) {}
    bool word(std::string& out) {
        while (*p && std::isspace((unsigned char)*p)) p++;
        if (!*p) return false;
        const char* b = p;
        while (*p && !std::isspace((unsigned char)*p)) p++;
        out.assign(b, p);
        return true;
    }
    bool real(float& out) {
        char* e = nullptr;
        float f = std::strtof(p, &e);
        if (e == p) return false;
        out = f; p = e;
        return true;
    }
    bool real3(float* out) { return real(out[0]) && real(out[1]) && real(out[2]); }
};

struct MtlEntry {
    float kd[3] = {0.8f, 0.8f, 0.8f};
    float ke[3] = {0.f, 0.f, 0.f};
    float ks[3] = {0.f, 0.f, 0.f};
    int illum = -1;
};

std::map<std::string, MtlEntry> read_mtl(const std::string& path) {
    std::map<std::string, MtlEntry> table;
    std::ifstream in(path);
    if (!in.is_open()) return table;
    std::string line, key, name;
    MtlEntry cur;
    bool open = false;
    while (std::getline(in, line)) {
        Cursor c(line);
        if (!c.word(key)) continue;
        if (key == "newmtl") {
            if (open && !name.empty()) table[name] = cur;
            cur = MtlEntry();
            name.clear();
            c.word(name);
            open = true;
        } else if (key == "Kd") { float t[3]; if (c.real3(t)) memcpy(cur.kd, t, sizeof t); }
        else if (key == "Ke") { float t[3]; if (c.real3(t)) memcpy(cur.ke, t, sizeof t); }
        else if (key == "Ks") { float t[3]; if (c.real3(t)) memcpy(cur.ks, t, sizeof t); }
        else if (key == "illum") { float t; if (c.real(t)) cur.illum = (int)t; }
    }
    if (open && !name.empty()) table[name] = cur;
    return table;
}

// unsigned index with istream semantics: optional sign, digits; "-3" wraps to a huge value (rejected later)
bool read_index(const char*& q, unsigned long long& out) {
    const char* s = q;
    bool neg = false;
    if (*s == '+' || *s == '-') { neg = (*s == '-'); s++; }
    if (!std::isdigit((unsigned char)*s)) return false;
    unsigned long long v = 0;
    while (std::isdigit((unsigned char)*s)) { v = v * 10ull + (unsigned long long)(*s - '0'); s++; }
    out = neg ? (0ull - v) : v;
    q = s;
    return true;
}

inline void geometric_normal(const float* a, const float* b, const float* c, float* n) {
    st3(n, unit(cross(ld3(b) - ld3(a), ld3(c) - ld3(a))));
}

}  // namespace

bool load_obj(const std::string& path, bool mirror_from_mtl, std::vector<ptx_primitive>& out, std::string& err) {
    std::ifstream in(path);
    if (!in.is_open()) { err = "cannot open OBJ file: " + path; return false; }
    std::string dir;
    size_t cut = path.find_last_of("/\\");
    if (cut != std::string::npos) dir = path.substr(0, cut + 1);

    std::vector<float> pos, nrm;
    std::map<std::string, MtlEntry> mtl;
    MtlEntry active;
    std::string line, key, tok;
    out.clear();
    while (std::getline(in, line)) {
        if (line.empty() || line[0] == '#' || line[0] == 'o' || line[0] == 's') continue;
        Cursor c(line);
        if (!c.word(key)) continue;
        if (key == "v") {
            float t[3]; if (c.real3(t)) pos.insert(pos.end(), t, t + 3);
        } else if (key == "vn") {
            float t[3];
            if (c.real3(t)) { float u[3]; st3(u, unit(ld3(t))); nrm.insert(nrm.end(), u, u + 3); }
        } else if (key == "mtllib") {
            std::string f; c.word(f);
            mtl = read_mtl(dir + f);
        } else if (key == "usemtl") {
            std::string m; c.word(m);
            auto it = mtl.find(m);
            active = (it == mtl.end()) ? MtlEntry() : it->second;
        } else if (key == "f") {
            unsigned long long vi[8], ni[8]; int cnt = 0; bool overflow = false;
            while (c.word(tok)) {
                const char* q = tok.c_str();
                unsigned long long v = 0, vt = 0, vn = 0;
                if (!read_index(q, v)) continue;
                if (*q == '/') {
                    q++;
                    if (*q == '/') { q++; if (!read_index(q, vn)) vn = 0; }
                    else if (read_index(q, vt) && *q == '/') { q++; if (!read_index(q, vn)) vn = 0; }
                }
                if (cnt < 8) { vi[cnt] = v; ni[cnt] = vn; cnt++; } else overflow = true;
            }
            if (overflow || (cnt != 3 && cnt != 4)) continue;
            size_t nv = pos.size() / 3, nn = nrm.size() / 3;
            bool ok = true;
            for (int k = 0; k < cnt; k++) ok = ok && vi[k] != 0 && vi[k] <= nv;
            if (!ok) continue;
            ptx_primitive pr; memset(&pr, 0, sizeof pr);
            pr.type = cnt == 3 ? 0 : 1;
            for (int k = 0; k < cnt; k++) memcpy(pr.v[k], &pos[3 * (vi[k] - 1)], 12);
            if (ni[0] != 0 && ni[0] <= nn) memcpy(pr.normal, &nrm[3 * (ni[0] - 1)], 12);
            else if (cnt == 3) geometric_normal(pr.v[0], pr.v[1], pr.v[2], pr.normal);
            else geometric_normal(pr.v[0], pr.v[1], pr.v[3], pr.normal);
            memcpy(pr.bsdf, active.kd, 12);
            memcpy(pr.Le, active.ke, 12);
            if (mirror_from_mtl && (active.illum == 3 || active.illum == 5)) { pr.material_type = 1; memcpy(pr.bsdf, active.ks, 12); }
            out.push_back(pr);
        }
    }
    if (out.empty()) { err = "no valid primitives were loaded from " + path; return false; }
    return true;
}

// convertQuadsToTriangles (application_state.h:323-365): (v00,v10,v11), (v00,v11,v01); the 4-argument
// Triangle constructor recomputes GEOMETRIC normals (triangle.h:21-29), dropping any vn override.
void split_quads(std::vector<ptx_primitive>& prims) {
    std::vector<ptx_primitive> res;
    res.reserve(prims.size() * 2);
    for (const ptx_primitive& p : prims) {
        if (p.type == 0) { res.push_back(p); continue; }
        for (int h = 0; h < 2; h++) {
            ptx_primitive t = p;
            t.type = 0;
            memcpy(t.v[1], p.v[1 + h], 12); memcpy(t.v[2], p.v[2 + h], 12); memset(t.v[3], 0, 12);
            geometric_normal(t.v[0], t.v[1], t.v[2], t.normal);
            res.push_back(t);
        }
    }
    prims.swap(res);
}


// subdivide_primitives (form_factors.h:475-574): each level splits a triangle into 4 (corner,
// corner, corner, centre) and a quad into 4 quads around its centre; new normals are geometric.
void subdivide(std::vector<ptx_primitive>& prims, int levels) {
    auto mid = [](const float* a, const float* b, float* m) { st3(m, (ld3(a) + ld3(b)) * 0.5f); };
    for (int lv = 0; lv < levels; lv++) {
        std::vector<ptx_primitive> next;
        next.reserve(prims.size() * 4);
        for (const ptx_primitive& p : prims) {
            if (p.type == 0) {
                float m0[3], m1[3], m2[3];
                mid(p.v[0], p.v[1], m0); mid(p.v[1], p.v[2], m1); mid(p.v[2], p.v[0], m2);
                const float* corners[4][3] = {{p.v[0], m0, m2}, {m0, p.v[1], m1}, {m1, p.v[2], m2}, {m0, m1, m2}};
                for (auto& c : corners) {
                    ptx_primitive t = p;
                    for (int k = 0; k < 3; k++) memcpy(t.v[k], c[k], 12);
                    geometric_normal(t.v[0], t.v[1], t.v[2], t.normal);
                    next.push_back(t);
                }
            } else {
                float m01[3], m12[3], m23[3], m30[3], ctr[3];
                mid(p.v[0], p.v[1], m01); mid(p.v[1], p.v[2], m12); mid(p.v[2], p.v[3], m23); mid(p.v[3], p.v[0], m30);
                st3(ctr, (((ld3(p.v[0]) + ld3(p.v[1])) + ld3(p.v[2])) + ld3(p.v[3])) * 0.25f);
                const float* corners[4][4] = {{p.v[0], m01, ctr, m30}, {m01, p.v[1], m12, ctr}, {ctr, m12, p.v[2], m23}, {m30, ctr, m23, p.v[3]}};
                for (auto& c : corners) {
                    ptx_primitive q = p;
                    for (int k = 0; k < 4; k++) memcpy(q.v[k], c[k], 12);
                    geometric_normal(q.v[0], q.v[1], q.v[3], q.normal);
                    next.push_back(q);
                }
            }
        }
        prims.swap(next);
    }
}

}  // namespace ptx

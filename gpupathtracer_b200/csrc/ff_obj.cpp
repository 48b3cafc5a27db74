// ff_obj.cpp -- OBJ / MTL ingest: every per-corner attribute LoadMesh keeps, parsed in parallel.
//
// Replaces LoadMesh (reference utilities.h:781-840), which goes through tinyobjloader on one thread and keeps, per
// face corner, position / normal / uv (Triangle::m_v0..2, m_n0..2, m_uv0..2, utilities.h:163-170) and reads, then
// drops, the per-face material id (utilities.h:837).  This reader keeps all four.  It is written for the big-scene
// ingest row: the file is cut at line boundaries into one chunk per host thread, every chunk is parsed
// independently (attributes into chunk-local arrays, face corners as raw indices, relative indices against the
// chunk-local running counts), a prefix sum over the chunk counts turns them into global indices, and the per-triangle
// gather runs in parallel again.  Output order = file order, identical for any thread count.
//
// Differences from LoadMesh, all deliberate: a corner without `vn` / `vt` gets a zero normal / uv instead of the
// out-of-range read at utilities.h:823-824 (sceneResources/triangle.obj has neither); faces with more than three
// corners are fan-triangulated (the shipped meshes are all triangles); negative (relative) indices are supported
// (triangle.obj uses them).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "ff_internal.h"

namespace ff {

void set_error(const std::string& msg);

namespace {

constexpr long long REL_BASE = 1ll << 40;  // corner index encoding: [0, REL_BASE/2) absolute, else REL_BASE + chunk-relative
constexpr long long NO_INDEX = -1;

bool read_whole_file(const char* path, std::string& out) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? fread(&out[0], 1, (size_t)n, f) : 0;
    fclose(f);
    return got == out.size();
}

// The file as one read-only byte range.  Mapped, not copied (a 200 MB file costs 0.2 s to fread into a zero-filled
// string before the first thread can start; mapped, the page faults happen inside the parallel parse).  The parsers
// (strtof / strtoll) stop at the first byte that cannot continue a number, so the range must END in a newline: a file
// that does not is read into a string with one appended.
struct FileText {
    const char* data = nullptr;
    size_t size = 0;
    void* map = nullptr;
    size_t map_size = 0;
    std::string copy;
    ~FileText() { if (map) munmap(map, map_size); }
    bool open(const char* path) {
        const int fd = ::open(path, O_RDONLY);
        if (fd < 0) return false;
        struct stat st;
        if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) { ::close(fd); return false; }
        const size_t n = (size_t)st.st_size;
        if (n == 0) { ::close(fd); data = ""; size = 0; return true; }
        void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
        ::close(fd);
        if (m != MAP_FAILED && ((const char*)m)[n - 1] == '\n') { map = m; map_size = n; data = (const char*)m; size = n; return true; }
        if (m != MAP_FAILED) munmap(m, n);
        if (!read_whole_file(path, copy)) return false;
        copy.push_back('\n');
        data = copy.data(); size = copy.size();
        return true;
    }
};

struct Chunk {
    const char *begin = nullptr, *end = nullptr;
    std::vector<float> pos, nrm, tex;
    std::vector<long long> corner;  // 9 per triangle: (v, vt, vn) x 3
    std::vector<std::pair<size_t, std::string>> usemtl;  // (first local triangle it applies to, name)
    std::string mtllib;
    std::string error;
    size_t first_line = 0, lines = 0;
    // filled by the prefix pass
    size_t base_v = 0, base_t = 0, base_n = 0, base_tri = 0;
};

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r'; }

std::string token_rest(const char* q, const char* eol) {
    while (q < eol && is_space(*q)) ++q;
    const char* e = eol;
    while (e > q && is_space(e[-1])) --e;
    return std::string(q, e);
}

// Number parsers that never look past `eol`: strtof / strtoll skip leading white space INCLUDING new lines and stop only at a
// non-numeric byte, so on the mmap'd file a short line (`vt u`, a two-coordinate `v`) would take its missing value from the
// next line, and a number that ends the file on a page boundary would be read past the mapping.  The token is copied to a
// small NUL-terminated buffer first.  Return false when no number starts before eol (the missing-value case).
bool parse_float(const char*& q, const char* eol, float& out) {
    while (q < eol && is_space(*q)) ++q;
    char buf[64];
    size_t n = 0;
    while (q + n < eol && !is_space(q[n]) && n + 1 < sizeof(buf)) { buf[n] = q[n]; ++n; }
    if (n == 0) return false;
    buf[n] = 0;
    char* e;
    out = strtof(buf, &e);
    if (e == buf) return false;
    q += (size_t)(e - buf);
    return true;
}
bool parse_index(const char*& q, const char* eol, long long& out) {
    char buf[32];
    size_t n = 0;
    while (q + n < eol && (q[n] == '-' || q[n] == '+' || (q[n] >= '0' && q[n] <= '9')) && n + 1 < sizeof(buf)) { buf[n] = q[n]; ++n; }
    if (n == 0) return false;
    buf[n] = 0;
    char* e;
    out = strtoll(buf, &e, 10);
    if (e == buf) return false;
    q += (size_t)(e - buf);
    return true;
}

void parse_chunk(Chunk& c) {
    const char* p = c.begin;
    std::vector<long long> poly;
    while (p < c.end) {
        ++c.lines;
        const char* eol = (const char*)memchr(p, '\n', (size_t)(c.end - p));
        if (!eol) eol = c.end;
        const char* q = p;
        while (q < eol && is_space(*q)) ++q;
        const size_t left = (size_t)(eol - q);
        if (left >= 2 && q[0] == 'v' && is_space(q[1])) {
            q += 2;
            for (int k = 0; k < 3; ++k) {
                float v;
                if (!parse_float(q, eol, v)) { c.error = "malformed vertex at line " + std::to_string(c.lines); return; }
                c.pos.push_back(v);
            }
        } else if (left >= 3 && q[0] == 'v' && q[1] == 'n' && is_space(q[2])) {
            q += 3;
            for (int k = 0; k < 3; ++k) {
                float v;
                if (!parse_float(q, eol, v)) { c.error = "malformed normal at line " + std::to_string(c.lines); return; }
                c.nrm.push_back(v);
            }
        } else if (left >= 3 && q[0] == 'v' && q[1] == 't' && is_space(q[2])) {
            q += 3;
            for (int k = 0; k < 2; ++k) {
                float v;
                if (!parse_float(q, eol, v)) { if (k == 0) { c.error = "malformed texture coordinate at line " + std::to_string(c.lines); return; } v = 0.f; }
                c.tex.push_back(v);
            }
        } else if (left >= 2 && q[0] == 'f' && is_space(q[1])) {
            q += 2;
            poly.clear();
            const long long cnt[3] = {(long long)(c.pos.size() / 3), (long long)(c.tex.size() / 2), (long long)(c.nrm.size() / 3)};
            while (q < eol) {
                while (q < eol && is_space(*q)) ++q;
                if (q >= eol) break;
                long long idx[3] = {NO_INDEX, NO_INDEX, NO_INDEX};
                bool bad = false;
                for (int k = 0; k < 3; ++k) {
                    if (q < eol && *q != '/' && !is_space(*q)) {
                        long long v;
                        if (!parse_index(q, eol, v) || v == 0) { bad = true; break; }
                        idx[k] = v > 0 ? v - 1 : REL_BASE + cnt[k] + v;
                    }
                    if (q < eol && *q == '/') ++q; else break;
                }
                if (bad || idx[0] == NO_INDEX) { c.error = "malformed face at line " + std::to_string(c.lines); return; }
                while (q < eol && !is_space(*q)) ++q;
                poly.insert(poly.end(), idx, idx + 3);
            }
            const size_t nc = poly.size() / 3;
            for (size_t k = 2; k < nc; ++k) {
                const size_t tri[3] = {0, k - 1, k};
                for (int a = 0; a < 3; ++a) c.corner.insert(c.corner.end(), poly.begin() + 3 * tri[a], poly.begin() + 3 * tri[a] + 3);
            }
        } else if (left >= 7 && memcmp(q, "usemtl", 6) == 0 && is_space(q[6])) {
            c.usemtl.emplace_back(c.corner.size() / 9, token_rest(q + 7, eol));
        } else if (left >= 7 && memcmp(q, "mtllib", 6) == 0 && is_space(q[6])) {
            c.mtllib = token_rest(q + 7, eol);
        }
        p = eol + 1;
    }
}

// n work items on at most `workers` threads, handed out by an atomic counter: the file is cut into several chunks per
// thread because OBJ lines are not equally expensive (an `f` line is nine integer parses, and files keep all their `v`
// lines first), so one static chunk per thread left most threads waiting for the slowest.
template <class F>
void run_threads(size_t n, size_t workers, F f) {
    workers = std::min(workers, n);
    if (workers <= 1) { for (size_t k = 0; k < n; ++k) f(k); return; }
    std::atomic<size_t> next{0};
    std::vector<std::thread> th;
    for (size_t t = 0; t < workers; ++t)
        th.emplace_back([&] { for (size_t k = next.fetch_add(1); k < n; k = next.fetch_add(1)) f(k); });
    for (auto& t : th) t.join();
}

}  // namespace

int obj_default_threads() {
    const char* e = getenv("FF_OBJ_THREADS");
    if (e && atoi(e) > 0) return atoi(e);
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::min<unsigned>(hw ? hw : 1u, 32u);
}

bool load_obj_mesh(const char* path, ObjMesh& out, int threads) {
    FileText text;
    if (!text.open(path)) { set_error(std::string("cannot read OBJ file: ") + path); return false; }
    if (threads <= 0) threads = obj_default_threads();
    const bool trace = getenv("FF_OBJ_TRACE") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto mark = [&](const char* what) {
        if (!trace) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[ff obj] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    mark("read file");
    // eight chunks per thread, at least 1 MiB each (FF_OBJ_CHUNK_BYTES overrides: tests cut small files), cut after a newline
    size_t min_chunk = 1u << 20;
    if (const char* e = getenv("FF_OBJ_CHUNK_BYTES")) if (atoll(e) > 0) min_chunk = (size_t)atoll(e);
    const size_t workers = (size_t)threads;
    size_t n_chunks = std::max<size_t>(1, std::min<size_t>(workers > 1 ? 8 * workers : 1, text.size / min_chunk));
    std::vector<Chunk> chunks(n_chunks);
    const char* base = text.data;
    const char* end = base + text.size;
    const char* cur = base;
    for (size_t k = 0; k < n_chunks; ++k) {
        chunks[k].begin = cur;
        const char* stop = k + 1 == n_chunks ? end : base + (text.size / n_chunks) * (k + 1);
        if (stop < cur) stop = cur;
        if (k + 1 < n_chunks) {
            const char* nl = (const char*)memchr(stop, '\n', (size_t)(end - stop));
            stop = nl ? nl + 1 : end;
        }
        chunks[k].end = stop;
        cur = stop;
    }
    run_threads(n_chunks, workers, [&](size_t k) { parse_chunk(chunks[k]); });
    mark("parse chunks");
    size_t nv = 0, nt = 0, nn = 0, ntri = 0, line = 0;
    for (auto& c : chunks) {
        c.first_line = line; line += c.lines;
        if (!c.error.empty()) {
            // the chunk counted lines from 1; report the file's line number
            const size_t at = c.error.rfind(' ');
            const size_t local = (size_t)atoll(c.error.c_str() + at + 1);
            set_error(std::string(path) + ": " + c.error.substr(0, at + 1) + std::to_string(c.first_line + local));
            return false;
        }
        c.base_v = nv; c.base_t = nt; c.base_n = nn; c.base_tri = ntri;
        nv += c.pos.size() / 3; nt += c.tex.size() / 2; nn += c.nrm.size() / 3; ntri += c.corner.size() / 9;
    }
    // concatenate the attribute pools
    std::vector<float> pos(3 * nv), tex(2 * nt), nrm(3 * nn);
    run_threads(n_chunks, workers, [&](size_t k) {
        const Chunk& c = chunks[k];
        if (!c.pos.empty()) memcpy(&pos[3 * c.base_v], c.pos.data(), sizeof(float) * c.pos.size());
        if (!c.tex.empty()) memcpy(&tex[2 * c.base_t], c.tex.data(), sizeof(float) * c.tex.size());
        if (!c.nrm.empty()) memcpy(&nrm[3 * c.base_n], c.nrm.data(), sizeof(float) * c.nrm.size());
    });
    // materials: names in order of first use; the material in force carries across chunk boundaries
    out = ObjMesh();
    out.xyz9.resize(9 * ntri); out.n9.assign(9 * ntri, 0.f); out.uv6.assign(6 * ntri, 0.f); out.face_material.assign(ntri, -1);
    std::vector<std::vector<std::pair<size_t, int>>> mat_events(n_chunks);
    std::vector<int> mat_at_start(n_chunks, -1);
    int current = -1;
    for (size_t k = 0; k < n_chunks; ++k) {
        mat_at_start[k] = current;
        for (auto& u : chunks[k].usemtl) {
            auto it = std::find(out.material_names.begin(), out.material_names.end(), u.second);
            if (it == out.material_names.end()) { out.material_names.push_back(u.second); it = out.material_names.end() - 1; }
            current = (int)(it - out.material_names.begin());
            mat_events[k].emplace_back(u.first, current);
        }
        if (!chunks[k].mtllib.empty() && out.mtllib.empty()) out.mtllib = chunks[k].mtllib;
    }
    std::vector<std::string> errors(n_chunks);
    std::vector<char> saw_n(n_chunks, 0), saw_t(n_chunks, 0);
    mark("pools + output allocation");
    run_threads(n_chunks, workers, [&](size_t k) {
        const Chunk& c = chunks[k];
        const size_t n = c.corner.size() / 9;
        char seen_n = 0, seen_t = 0;   // (not saw_n[k] / saw_t[k] directly: neighbouring bytes written per corner by different threads)
        const long long base[3] = {(long long)c.base_v, (long long)c.base_t, (long long)c.base_n};
        const long long limit[3] = {(long long)nv, (long long)nt, (long long)nn};
        size_t ev = 0;
        int mat = mat_at_start[k];
        for (size_t t = 0; t < n; ++t) {
            while (ev < mat_events[k].size() && mat_events[k][ev].first <= t) mat = mat_events[k][ev++].second;
            const size_t g = c.base_tri + t;
            out.face_material[g] = mat;
            for (int a = 0; a < 3; ++a) {
                long long idx[3];
                for (int j = 0; j < 3; ++j) {
                    long long v = c.corner[9 * t + 3 * a + j];
                    if (v != NO_INDEX) {
                        if (v >= REL_BASE / 2) v = base[j] + (v - REL_BASE);
                        if (v < 0 || v >= limit[j]) { if (errors[k].empty()) errors[k] = "face index out of range (triangle " + std::to_string(g) + ")"; v = NO_INDEX; if (j == 0) return; }
                    }
                    idx[j] = v;
                }
                memcpy(&out.xyz9[9 * g + 3 * a], &pos[3 * (size_t)idx[0]], 12);
                if (idx[1] != NO_INDEX) { memcpy(&out.uv6[6 * g + 2 * a], &tex[2 * (size_t)idx[1]], 8); seen_t = 1; }
                if (idx[2] != NO_INDEX) { memcpy(&out.n9[9 * g + 3 * a], &nrm[3 * (size_t)idx[2]], 12); seen_n = 1; }
            }
            saw_n[k] = seen_n; saw_t[k] = seen_t;
        }
    });
    mark("gather corners");
    for (auto& e : errors) if (!e.empty()) { set_error(std::string(path) + ": " + e); return false; }
    out.has_normals = std::find(saw_n.begin(), saw_n.end(), 1) != saw_n.end();
    out.has_uvs = std::find(saw_t.begin(), saw_t.end(), 1) != saw_t.end();
    out.bytes = text.size;
    return true;
}

bool load_obj_positions(const char* path, std::vector<float>& xyz9) {
    ObjMesh m;
    if (!load_obj_mesh(path, m, 0)) return false;
    xyz9 = std::move(m.xyz9);
    return true;
}

// ---------------------------------------------------------------------------------------------- MTL
// newmtl / Kd / Ke / Ks / Tf / Ni / illum -> ff_material (BXDF, utilities.h:77-88):
//   Ke non-zero            -> EMITTER, emissive = Ke, intensity 1
//   illum 4, 6, 7 or 9     -> GLASS, ior = Ni, transmittance = Tf (default 1), specular = Ks
//   illum 3, 5 or 8        -> MIRROR, specular = Ks
//   otherwise              -> DIFFUSE, albedo = Kd
bool load_mtl(const char* path, std::vector<std::string>& names, std::vector<ff_material>& mats) {
    std::string text;
    if (!read_whole_file(path, text)) { set_error(std::string("cannot read MTL file: ") + path); return false; }
    struct Raw { float kd[3] = {0.8f, 0.8f, 0.8f}, ke[3] = {0, 0, 0}, ks[3] = {1, 1, 1}, tf[3] = {1, 1, 1}; float ni = 1.5f; int illum = 2; };
    std::vector<Raw> raw;
    const char* p = text.data();
    const char* end = p + text.size();
    auto three = [](const char* q, float* dst) {
        char* e;
        for (int k = 0; k < 3; ++k) { const float v = strtof(q, &e); if (e == q) { if (k) dst[k] = dst[0]; else return; } else dst[k] = v; q = e; }
    };
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const std::string line = token_rest(p, eol);
        const size_t sp = line.find_first_of(" \t");
        const std::string key = line.substr(0, sp), rest = sp == std::string::npos ? "" : token_rest(line.c_str() + sp, line.c_str() + line.size());
        if (key == "newmtl") { names.push_back(rest); raw.emplace_back(); }
        else if (!raw.empty()) {
            Raw& r = raw.back();
            if (key == "Kd") three(rest.c_str(), r.kd);
            else if (key == "Ke") three(rest.c_str(), r.ke);
            else if (key == "Ks") three(rest.c_str(), r.ks);
            else if (key == "Tf") three(rest.c_str(), r.tf);
            else if (key == "Ni") r.ni = strtof(rest.c_str(), nullptr);
            else if (key == "illum") r.illum = atoi(rest.c_str());
        }
        p = eol + 1;
    }
    for (const Raw& r : raw) {
        ff_material m;
        memset(&m, 0, sizeof(m));
        for (int k = 0; k < 3; ++k) { m.albedo[k] = r.kd[k]; m.emissive[k] = r.ke[k]; m.specular[k] = r.ks[k]; m.transmittance[k] = r.tf[k]; }
        m.ior = r.ni > 0.f ? r.ni : 1.5f;
        m.intensity = 0.f;
        if (r.ke[0] > 0.f || r.ke[1] > 0.f || r.ke[2] > 0.f) { m.type = FF_MAT_EMITTER; m.intensity = 1.f; }
        else if (r.illum == 4 || r.illum == 6 || r.illum == 7 || r.illum == 9) m.type = FF_MAT_GLASS;
        else if (r.illum == 3 || r.illum == 5 || r.illum == 8) m.type = FF_MAT_MIRROR;
        else m.type = FF_MAT_DIFFUSE;
        mats.push_back(m);
    }
    return true;
}

}  // namespace ff

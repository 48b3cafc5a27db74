// ff_scenefile.cpp -- text scene description: replaces the literals main() hard-codes (reference kernel.cu:227-259 scene,
// kernel.cu:311-322 camera; "// TODO: Load scene from file", kernel.cu:261).  SURVEY.md section 8(f) row f1.
//
// One directive per line, '#' starts a comment, keywords take their arguments in any order:
//   camera    pos X Y Z [up X Y Z] [yaw D] [pitch D] [fov D] [size W H] [near N] [far F]
//   material  NAME emitter|diffuse|mirror|glass [albedo R G B] [emissive R G B] [intensity I] [specular R G B] [ior N]
//             [transmittance R G B]
//   mesh      PATH.obj [pos X Y Z] [rot X Y Z] [scale X Y Z] [material NAME] [flip] [normals] [uvs] [materials]
//             normals / uvs / materials keep the file's `vn` / `vt` / `usemtl` + `mtllib` (ff_scene_add_obj_ex flags)
//   plane     [pos ..] [rot ..] [scale ..] [material NAME]
//   sphere    RADIUS [pos ..] [rot ..] [scale ..] [material NAME]
//   triangles [pos ..] [rot ..] [scale ..] [material NAME] [flip]     then lines "tri x0 y0 z0 x1 y1 z1 x2 y2 z2", then "end"
//   soup      N [size S] [material NAME]            synthetic triangle soup of BASELINE.json configs 4/5
//   render    [spp N] [bounces N] [seed N] [integrator wavefront|megakernel|ascoded] [nee 0|1] [aa 0|1]
//             [background R G B] [eps E] [builder lbvh|sah]
// Geometry indices follow file order (like the std::vector in kernel.cu:253-258).  Relative mesh paths resolve against
// `asset_dir` when given, else against the directory of the scene file.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/firefly.h"

namespace ff {
void set_error(const std::string& msg);
}

namespace {

struct ParsedGeom {
    int kind;  // 0 sphere, 1 plane, 2 triangles inline, 3 mesh file, 4 soup
    std::string path;
    float pos[3] = {0, 0, 0}, rot[3] = {0, 0, 0}, scale[3] = {1, 1, 1};
    float radius = 1.f, soup_size = 0.f;
    size_t soup_n = 0;
    bool flip = false;
    int obj_flags = 0;  // FF_OBJ_NORMALS | FF_OBJ_UVS | FF_OBJ_MATERIALS, mesh files only
    ff_material mat;
    std::vector<float> xyz9;
};
struct Parsed {
    ff_camera cam;
    ff_render_params rp;
    int builder = FF_BUILDER_LBVH;
    std::vector<ParsedGeom> geoms;
};

ff_material default_mat() {
    ff_material m;
    memset(&m, 0, sizeof(m));
    m.type = FF_MAT_DIFFUSE;
    m.albedo[0] = m.albedo[1] = m.albedo[2] = 0.8f;
    m.specular[0] = m.specular[1] = m.specular[2] = 1.f;
    m.ior = 1.5f;
    m.transmittance[0] = m.transmittance[1] = m.transmittance[2] = 1.f;
    return m;
}

struct Cursor {
    std::vector<std::string> tok;
    size_t i = 0;
    std::string where;
    bool more() const { return i < tok.size(); }
    bool word(std::string& out) { if (!more()) return false; out = tok[i++]; return true; }
    bool number(double& v) {
        if (!more()) return false;
        char* e = nullptr;
        v = strtod(tok[i].c_str(), &e);
        if (e == tok[i].c_str() || *e != '\0' || !std::isfinite(v)) return false;
        ++i;
        return true;
    }
    bool floats(float* out, int n) {
        for (int k = 0; k < n; ++k) { double v; if (!number(v)) return false; out[k] = (float)v; }
        return true;
    }
};

bool fail(const Cursor& c, const std::string& msg) {
    ff::set_error(c.where + ": " + msg);
    return false;
}

// the keywords shared by every geometry directive
bool parse_placement(Cursor& c, ParsedGeom& g, const std::map<std::string, ff_material>& mats, bool allow_flip) {
    std::string k;
    while (c.word(k)) {
        if (k == "pos") { if (!c.floats(g.pos, 3)) return fail(c, "pos needs 3 numbers"); }
        else if (k == "rot") { if (!c.floats(g.rot, 3)) return fail(c, "rot needs 3 numbers (degrees)"); }
        else if (k == "scale") { if (!c.floats(g.scale, 3)) return fail(c, "scale needs 3 numbers"); }
        else if (k == "material") {
            std::string name;
            if (!c.word(name)) return fail(c, "material needs a name");
            auto it = mats.find(name);
            if (it == mats.end()) return fail(c, "unknown material '" + name + "'");
            g.mat = it->second;
        } else if (k == "flip" && allow_flip) g.flip = true;
        else if (k == "normals" && g.kind == 3) g.obj_flags |= FF_OBJ_NORMALS;
        else if (k == "uvs" && g.kind == 3) g.obj_flags |= FF_OBJ_UVS;
        else if (k == "materials" && g.kind == 3) g.obj_flags |= FF_OBJ_MATERIALS;
        else if (k == "size" && g.kind == 4) { if (!c.floats(&g.soup_size, 1)) return fail(c, "size needs a number"); }
        else return fail(c, "unexpected '" + k + "'");
    }
    return true;
}

bool parse_text(const std::string& text, const std::string& name, Parsed& out) {
    ff_camera& cam = out.cam;
    memset(&cam, 0, sizeof(cam));
    cam.pos[2] = 15.f; cam.world_up[1] = 1.f; cam.yaw_deg = -90.f; cam.fov_deg = 70.f; cam.width = cam.height = 800;  // kernel.cu:311-322
    cam.near_clip = 0.1f; cam.far_clip = 1000.f;
    ff_render_params& rp = out.rp;
    memset(&rp, 0, sizeof(rp));
    rp.spp = 1; rp.max_bounces = 4; rp.seed = 1234; rp.integrator = FF_INTEGRATOR_WAVEFRONT; rp.ray_eps = 1e-4f;
    std::map<std::string, ff_material> mats;
    std::istringstream in(text);
    std::string line;
    int line_no = 0;
    ParsedGeom* open_tris = nullptr;
    while (std::getline(in, line)) {
        ++line_no;
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.resize(hash);
        Cursor c;
        c.where = name + ":" + std::to_string(line_no);
        std::istringstream ls(line);
        for (std::string t; ls >> t;) c.tok.push_back(t);
        std::string d;
        if (!c.word(d)) continue;
        if (open_tris) {
            if (d == "tri") {
                float v[9];
                if (!c.floats(v, 9) || c.more()) return fail(c, "tri needs exactly 9 numbers");
                open_tris->xyz9.insert(open_tris->xyz9.end(), v, v + 9);
                continue;
            }
            if (d == "end") { if (open_tris->xyz9.empty()) return fail(c, "empty triangles block"); open_tris = nullptr; continue; }
            return fail(c, "expected 'tri' or 'end' inside a triangles block");
        }
        if (d == "camera") {
            std::string k;
            while (c.word(k)) {
                double v, w;
                if (k == "pos") { if (!c.floats(cam.pos, 3)) return fail(c, "pos needs 3 numbers"); }
                else if (k == "up") { if (!c.floats(cam.world_up, 3)) return fail(c, "up needs 3 numbers"); }
                else if (k == "yaw") { if (!c.number(v)) return fail(c, "yaw needs a number"); cam.yaw_deg = (float)v; }
                else if (k == "pitch") { if (!c.number(v)) return fail(c, "pitch needs a number"); cam.pitch_deg = (float)v; }
                else if (k == "fov") { if (!c.number(v)) return fail(c, "fov needs a number"); cam.fov_deg = (float)v; }
                else if (k == "near") { if (!c.number(v)) return fail(c, "near needs a number"); cam.near_clip = (float)v; }
                else if (k == "far") { if (!c.number(v)) return fail(c, "far needs a number"); cam.far_clip = (float)v; }
                else if (k == "size") {
                    if (!c.number(v) || !c.number(w) || v < 1 || w < 1 || v > 65536 || w > 65536) return fail(c, "size needs W H");
                    cam.width = (int)v; cam.height = (int)w;
                } else return fail(c, "unexpected '" + k + "' in camera");
            }
        } else if (d == "material") {
            std::string nm, ty, k;
            if (!c.word(nm) || !c.word(ty)) return fail(c, "material needs NAME TYPE");
            ff_material m = default_mat();
            if (ty == "emitter") m.type = FF_MAT_EMITTER;
            else if (ty == "diffuse") m.type = FF_MAT_DIFFUSE;
            else if (ty == "mirror") m.type = FF_MAT_MIRROR;
            else if (ty == "glass") m.type = FF_MAT_GLASS;
            else return fail(c, "unknown material type '" + ty + "'");
            while (c.word(k)) {
                if (k == "albedo") { if (!c.floats(m.albedo, 3)) return fail(c, "albedo needs 3 numbers"); }
                else if (k == "emissive") { if (!c.floats(m.emissive, 3)) return fail(c, "emissive needs 3 numbers"); }
                else if (k == "intensity") { if (!c.floats(&m.intensity, 1)) return fail(c, "intensity needs a number"); }
                else if (k == "specular") { if (!c.floats(m.specular, 3)) return fail(c, "specular needs 3 numbers"); }
                else if (k == "ior") { if (!c.floats(&m.ior, 1) || !(m.ior > 0.f)) return fail(c, "ior needs a positive number"); }
                else if (k == "transmittance") { if (!c.floats(m.transmittance, 3)) return fail(c, "transmittance needs 3 numbers"); }
                else return fail(c, "unexpected '" + k + "' in material");
            }
            mats[nm] = m;
        } else if (d == "mesh" || d == "plane" || d == "sphere" || d == "triangles" || d == "soup") {
            ParsedGeom g;
            g.mat = default_mat();
            if (d == "mesh") { g.kind = 3; if (!c.word(g.path)) return fail(c, "mesh needs a path"); }
            else if (d == "plane") g.kind = 1;
            else if (d == "sphere") { g.kind = 0; if (!c.floats(&g.radius, 1) || !(g.radius > 0.f)) return fail(c, "sphere needs a positive radius"); }
            else if (d == "triangles") g.kind = 2;
            else {
                g.kind = 4;
                double n;
                if (!c.number(n) || n < 1 || n > 1e9 || n != std::floor(n)) return fail(c, "soup needs a triangle count");
                g.soup_n = (size_t)n;
            }
            if (!parse_placement(c, g, mats, g.kind == 2 || g.kind == 3)) return false;
            out.geoms.push_back(std::move(g));
            if (d == "triangles") open_tris = &out.geoms.back();
        } else if (d == "render") {
            std::string k, w;
            double v;
            while (c.word(k)) {
                if (k == "spp") { if (!c.number(v) || v < 1) return fail(c, "spp needs a positive integer"); rp.spp = (int)v; }
                else if (k == "bounces") { if (!c.number(v) || v < 0) return fail(c, "bounces needs an integer >= 0"); rp.max_bounces = (int)v; }
                else if (k == "seed") { if (!c.number(v) || v < 0) return fail(c, "seed needs an integer >= 0"); rp.seed = (uint64_t)v; }
                else if (k == "nee") { if (!c.number(v) || (v != 0 && v != 1)) return fail(c, "nee needs 0 or 1"); rp.nee = (int)v; }
                else if (k == "aa") { if (!c.number(v) || (v != 0 && v != 1)) return fail(c, "aa needs 0 or 1"); rp.aa = (int)v; }
                else if (k == "eps") { if (!c.number(v) || v < 0) return fail(c, "eps needs a number >= 0"); rp.ray_eps = (float)v; }
                else if (k == "background") { if (!c.floats(rp.background, 3)) return fail(c, "background needs 3 numbers"); }
                else if (k == "integrator") {
                    if (!c.word(w)) return fail(c, "integrator needs a name");
                    if (w == "wavefront") rp.integrator = FF_INTEGRATOR_WAVEFRONT;
                    else if (w == "megakernel") rp.integrator = FF_INTEGRATOR_MEGAKERNEL;
                    else if (w == "ascoded") rp.integrator = FF_INTEGRATOR_AS_CODED;
                    else return fail(c, "unknown integrator '" + w + "'");
                } else if (k == "builder") {
                    if (!c.word(w)) return fail(c, "builder needs a name");
                    if (w == "lbvh") out.builder = FF_BUILDER_LBVH;
                    else if (w == "sah") out.builder = FF_BUILDER_SAH;
                    else return fail(c, "unknown builder '" + w + "'");
                } else return fail(c, "unexpected '" + k + "' in render");
            }
        } else {
            return fail(c, "unknown directive '" + d + "'");
        }
    }
    if (open_tris) { ff::set_error(name + ": triangles block without 'end'"); return false; }
    if (out.geoms.empty()) { ff::set_error(name + ": the scene has no geometry"); return false; }
    return true;
}

bool read_text(const char* path, std::string& out) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) out.append(buf, n);
    fclose(f);
    return true;
}

void fill_desc(const Parsed& p, ff_scene_file* d) {
    if (!d) return;
    d->camera = p.cam;
    d->params = p.rp;
    d->builder = p.builder;
    d->n_geometries = (int)p.geoms.size();
}

void flip(std::vector<float>& xyz9) {
    for (size_t t = 0; t + 9 <= xyz9.size(); t += 9)
        for (int k = 0; k < 3; ++k) std::swap(xyz9[t + 3 + k], xyz9[t + 6 + k]);
}

}  // namespace

extern "C" {

int ff_scene_file_validate(const char* path, ff_scene_file* out_desc) {
    if (!path) { ff::set_error("ff_scene_file_validate: NULL path"); return FF_ERR_INVALID; }
    std::string text;
    if (!read_text(path, text)) { ff::set_error(std::string("cannot read scene file: ") + path); return FF_ERR_IO; }
    Parsed p;
    if (!parse_text(text, path, p)) return FF_ERR_IO;
    fill_desc(p, out_desc);
    return FF_OK;
}

int ff_scene_load_file(ff_context* ctx, const char* path, const char* asset_dir, ff_scene** out_scene, ff_scene_file* out_desc) {
    if (!ctx || !path || !out_scene) { ff::set_error("ff_scene_load_file: NULL argument"); return FF_ERR_INVALID; }
    *out_scene = nullptr;
    std::string text;
    if (!read_text(path, text)) { ff::set_error(std::string("cannot read scene file: ") + path); return FF_ERR_IO; }
    Parsed p;
    if (!parse_text(text, path, p)) return FF_ERR_IO;
    std::string base = asset_dir ? asset_dir : "";
    if (!asset_dir) {
        const std::string s(path);
        const size_t slash = s.find_last_of('/');
        base = slash == std::string::npos ? "." : s.substr(0, slash);
    }
    ff_scene* sc = nullptr;
    int rc = ff_scene_create(ctx, &sc);
    if (rc) return rc;
    for (ParsedGeom& g : p.geoms) {
        switch (g.kind) {
            case 0: rc = ff_scene_add_sphere(sc, g.radius, g.pos, g.rot, g.scale, &g.mat, nullptr); break;
            case 1: rc = ff_scene_add_plane(sc, g.pos, g.rot, g.scale, &g.mat, nullptr); break;
            case 2:
                if (g.flip) flip(g.xyz9);
                rc = ff_scene_add_triangles(sc, g.xyz9.data(), g.xyz9.size() / 9, g.pos, g.rot, g.scale, &g.mat, nullptr);
                break;
            case 3: {
                const std::string full = (!g.path.empty() && g.path[0] == '/') ? g.path : base + "/" + g.path;
                rc = ff_scene_add_obj_ex(sc, full.c_str(), g.pos, g.rot, g.scale, &g.mat, g.obj_flags | (g.flip ? FF_OBJ_FLIP_WINDING : 0), nullptr);
                break;
            }
            default: {
                std::vector<float> soup(9 * g.soup_n);
                const float s = g.soup_size > 0.f ? g.soup_size : (float)std::pow((double)g.soup_n, -1.0 / 3.0);
                rc = ff_make_soup(ctx, g.soup_n, s, soup.data());
                if (rc == FF_OK) rc = ff_scene_add_triangles(sc, soup.data(), g.soup_n, g.pos, g.rot, g.scale, &g.mat, nullptr);
            }
        }
        if (rc) { ff_scene_destroy(sc); return rc; }
    }
    fill_desc(p, out_desc);
    *out_scene = sc;
    return FF_OK;
}

}  // extern "C"

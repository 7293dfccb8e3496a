// OBJ / MTL / .camera readers with the reference's grammar (MTPC/sceneManagement.cpp:17-262),
// rewritten for Linux: CRLF-safe, whitespace-tolerant, checked.  Host-only; no CUDA here.
#include "scene_loader.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace mcpt {
namespace {

// One text line split on blanks; '\r' (the shipped scenes are CRLF) and tabs count as blanks.
struct Tokens {
    std::vector<std::string> t;
    explicit Tokens(const std::string& line) {
        size_t i = 0, n = line.size();
        while (i < n) {
            while (i < n && (line[i] == ' ' || line[i] == '\t' || line[i] == '\r')) i++;
            size_t j = i;
            while (j < n && !(line[j] == ' ' || line[j] == '\t' || line[j] == '\r')) j++;
            if (j > i) t.emplace_back(line, i, j - i);
            i = j;
        }
    }
    bool is(const char* kw) const { return !t.empty() && t[0] == kw; }
    double num(size_t k) const { return k < t.size() ? std::atof(t[k].c_str()) : 0.0; }  // atof, as the reference
};

std::ifstream open_or_throw(const std::string& file) {
    std::ifstream f(file);
    if (!f) throw std::runtime_error("cannot open " + file);
    return f;
}

bool read_sidecar(const std::string& file, std::vector<uint8_t>& bytes, int& h, int& w) {
    FILE* f = std::fopen(file.c_str(), "rb");
    if (!f) return false;
    int32_t hdr[2];
    bool ok = std::fread(hdr, 4, 2, f) == 2 && hdr[0] > 0 && hdr[1] > 0;
    if (ok) {
        h = hdr[0]; w = hdr[1];
        bytes.resize(size_t(h) * w * 3);
        ok = std::fread(bytes.data(), 1, bytes.size(), f) == bytes.size();
    }
    std::fclose(f);
    return ok;
}

// Material::readinMap (MTPC/sceneManagement.h:134-143) opens map_Kd relative to the CWD through
// cv::imread; no JPEG decoder exists in this image, so the decoded BGR bytes come from a sidecar.
int load_texture(const std::string& dir, const std::string& map_kd, HostScene& s) {
    std::vector<uint8_t> bytes; int h = 0, w = 0;
    std::vector<std::string> tries = {map_kd + ".bgr", dir + map_kd + ".bgr"};
    if (const char* e = std::getenv("MCPT_TEXDIR")) tries.push_back(std::string(e) + "/" + map_kd + ".bgr");
    for (const auto& f : tries)
        if (read_sidecar(f, bytes, h, w)) {
            s.texture_bytes.push_back(std::move(bytes));
            mcpt_texture t{}; t.height = h; t.width = w; t.bgr = nullptr;
            s.textures.push_back(t);
            return int(s.textures.size()) - 1;
        }
    throw std::runtime_error("texture sidecar " + map_kd + ".bgr not found (CWD, scene dir, $MCPT_TEXDIR)");
}

void read_mtl(const std::string& dir, const std::string& file, HostScene& s) {
    std::ifstream fin = open_or_throw(file);
    std::string line;
    int cur = -1;
    while (std::getline(fin, line)) {
        Tokens k(line);
        if (k.is("newmtl") && k.t.size() > 1) {
            auto it = s.material_index.find(k.t[1]);
            mcpt_material fresh{};
            fresh.tex_id = -1;
            if (it == s.material_index.end()) {
                cur = int(s.materials.size());
                s.material_index[k.t[1]] = cur;
                s.materials.push_back(fresh);
                s.material_names.push_back(k.t[1]);
            } else {
                cur = it->second;  // a repeated name replaces the map entry (sceneManagement.cpp:28)
                s.materials[cur] = fresh;
            }
        } else if (cur >= 0 && k.is("Kd")) {
            for (int c = 0; c < 3; c++) s.materials[cur].kd[c] = k.num(1 + c);
        } else if (cur >= 0 && k.is("Ks")) {
            for (int c = 0; c < 3; c++) s.materials[cur].ks[c] = k.num(1 + c);
        } else if (cur >= 0 && k.is("Ns")) {
            s.materials[cur].Ns = k.num(1);
        } else if (cur >= 0 && k.is("Ni")) {
            s.materials[cur].Ni = k.num(1);
        } else if (cur >= 0 && k.is("map_Kd") && k.t.size() > 1) {
            s.materials[cur].tex_id = load_texture(dir, k.t[1], s);
        }
    }
}

// "a/b/c" -> three 0-based indices.  The reference reads the 2nd index into vn[] and the 3rd into
// vt[] (MTPC/sceneManagement.cpp:145-170) — kept, so that files it renders correctly (all shipped
// faces are i/i/i) render identically here.
void parse_corner(const std::string& tok, long idx[3]) {
    const char* p = tok.c_str();
    for (int k = 0; k < 3; k++) {
        idx[k] = std::atol(p) - 1;
        const char* slash = std::strchr(p, '/');
        if (!slash) { for (int r = k + 1; r < 3; r++) idx[r] = idx[0]; break; }
        p = slash + 1;
    }
}

void read_obj(const std::string& file, HostScene& s) {
    std::ifstream fin = open_or_throw(file);
    std::vector<double> v, vn, vt;
    std::string line;
    int cur_mat = -1;
    std::string cur_name;
    while (std::getline(fin, line)) {
        if (line.size() < 2) continue;
        Tokens k(line);
        if (k.is("v")) { for (int c = 0; c < 3; c++) v.push_back(k.num(1 + c)); }
        else if (k.is("vn")) { for (int c = 0; c < 3; c++) vn.push_back(k.num(1 + c)); }
        else if (k.is("vt")) { for (int c = 0; c < 2; c++) vt.push_back(k.num(1 + c)); }
        else if (k.is("usemtl") && k.t.size() > 1) {
            auto it = s.material_index.find(k.t[1]);
            cur_name = k.t[1];
            cur_mat = it == s.material_index.end() ? -1 : it->second;
        } else if (k.is("f")) {
            if (k.t.size() != 4) throw std::runtime_error(file + ": only triangles are supported (as in the reference)");
            if (cur_mat < 0) throw std::runtime_error(file + ": face uses unknown material '" + cur_name + "'");
            for (int c = 0; c < 3; c++) {
                long idx[3];
                parse_corner(k.t[1 + c], idx);
                if (idx[0] < 0 || size_t(idx[0]) * 3 + 2 >= v.size() || idx[1] < 0 || size_t(idx[1]) * 3 + 2 >= vn.size() ||
                    idx[2] < 0 || size_t(idx[2]) * 2 + 1 >= vt.size())
                    throw std::runtime_error(file + ": face index out of range");
                for (int a = 0; a < 3; a++) s.pos.push_back(v[size_t(idx[0]) * 3 + a]);
                for (int a = 0; a < 3; a++) s.nrm.push_back(vn[size_t(idx[1]) * 3 + a]);
                for (int a = 0; a < 2; a++) s.uv.push_back(vt[size_t(idx[2]) * 2 + a]);
            }
            s.mat_id.push_back(cur_mat);
        }
    }
    s.n_vertices = v.size() / 3;
}

void read_camera(const std::string& file, HostScene& s) {
    std::ifstream fin = open_or_throw(file);
    std::string line;
    while (std::getline(fin, line)) {
        Tokens k(line);
        if (k.is("eye")) for (int c = 0; c < 3; c++) s.camera.eye[c] = k.num(1 + c);
        else if (k.is("lookat")) for (int c = 0; c < 3; c++) s.camera.lookat[c] = k.num(1 + c);
        else if (k.is("up")) for (int c = 0; c < 3; c++) s.camera.up[c] = k.num(1 + c);
        else if (k.is("fovy")) s.camera.fovy = k.num(1);
        else if (k.is("width")) s.camera.width = int(k.num(1));
        else if (k.is("height")) s.camera.height = int(k.num(1));
        else if (k.is("mtlname") && k.t.size() > 1) {
            auto it = s.material_index.find(k.t[1]);
            if (it == s.material_index.end()) throw std::runtime_error(file + ": light names unknown material '" + k.t[1] + "'");
            mcpt_light l{};
            l.mat_id = it->second;
            for (int c = 0; c < 3; c++) l.radiance[c] = k.num(2 + c);
            s.lights.push_back(l);
        }
    }
}

}  // namespace

void HostScene::finalize() {
    for (size_t i = 0; i < textures.size(); i++) textures[i].bgr = texture_bytes[i].data();
    desc.n_faces = int32_t(mat_id.size());
    desc.n_materials = int32_t(materials.size());
    desc.n_textures = int32_t(textures.size());
    desc.n_lights = int32_t(lights.size());
    desc.pos = pos.data(); desc.nrm = nrm.data(); desc.uv = uv.data(); desc.mat_id = mat_id.data();
    desc.materials = materials.data(); desc.textures = textures.data(); desc.lights = lights.data();
}

// read_scene with the OBJ left to the GPU: only the tables from .mtl / .camera are read here.  A .camera error is held
// back until the OBJ has been read, to keep the order in which read_scene meets problems.
void load_tables(const std::string& path, const std::string& name, HostScene& out, std::string& camera_error) {
    const std::string base = path + name;
    read_mtl(path, base + ".mtl", out);
    try { read_camera(base + ".camera", out); }
    catch (const std::exception& e) { camera_error = e.what(); out.lights.clear(); }
    out.finalize();
}

void load_scene(const std::string& path, const std::string& name, HostScene& out) {
    const std::string base = path + name;
    read_mtl(path, base + ".mtl", out);     // order of MTPC/sceneManagement.cpp:270-272
    read_obj(base + ".obj", out);
    read_camera(base + ".camera", out);
    if (out.mat_id.empty()) throw std::runtime_error(base + ".obj: no faces");
    out.finalize();
}

}  // namespace mcpt

// ---- C ABI ---------------------------------------------------------------------------------------
struct mcpt_host_scene { mcpt::HostScene s; };
void mcpt_set_error(const char* msg);  // mcpt_error.cpp

extern "C" {
MCPT_API int mcpt_scene_load(const char* path, const char* name, mcpt_host_scene** out) {
    if (!path || !name || !out) { mcpt_set_error("mcpt_scene_load: null argument"); return MCPT_ERR_ARG; }
    mcpt_host_scene* h = new mcpt_host_scene;
    try { mcpt::load_scene(path, name, h->s); }
    catch (const std::exception& e) { mcpt_set_error(e.what()); delete h; return MCPT_ERR_IO; }
    *out = h;
    return MCPT_OK;
}
MCPT_API const mcpt_scene_desc* mcpt_scene_get_desc(const mcpt_host_scene* s) { return s ? &s->s.desc : nullptr; }
MCPT_API const mcpt_camera* mcpt_scene_get_camera(const mcpt_host_scene* s) { return s ? &s->s.camera : nullptr; }
MCPT_API void mcpt_scene_free(mcpt_host_scene* s) { delete s; }
MCPT_API const char* mcpt_scene_get_material_name(const mcpt_host_scene* s, int32_t i) {
    return (s && i >= 0 && size_t(i) < s->s.material_names.size()) ? s->s.material_names[i].c_str() : nullptr;
}
MCPT_API int mcpt_scene_load_gpu(mcpt_ctx* ctx, const char* path, const char* name, mcpt_camera* camera_out, float* ingest_ms_out) {
    if (!ctx || !path || !name) { mcpt_set_error("mcpt_scene_load_gpu: null argument"); return MCPT_ERR_ARG; }
    mcpt::HostScene hs;
    std::string camera_error;
    try { mcpt::load_tables(path, name, hs, camera_error); }
    catch (const std::exception& e) { mcpt_set_error(e.what()); return MCPT_ERR_IO; }
    const std::string base = std::string(path) + name;
    if (hs.materials.empty()) { mcpt_set_error((base + ".mtl: no materials").c_str()); return MCPT_ERR_IO; }
    std::vector<const char*> names;
    for (const auto& n : hs.material_names) names.push_back(n.c_str());
    const int rc = mcpt_ingest_obj_file(ctx, (base + ".obj").c_str(), &hs.desc, names.data(), ingest_ms_out);
    if (rc) return rc;
    if (!camera_error.empty()) { mcpt_set_error(camera_error.c_str()); return MCPT_ERR_IO; }
    if (camera_out) *camera_out = hs.camera;
    return MCPT_OK;
}
}

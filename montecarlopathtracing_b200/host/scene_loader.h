// Host-side scene model: the reference's scene_data (MTPC/sceneManagement.h:173-199) flattened
// into the structure-of-arrays form the C ABI (include/mcpt.h) hands to the GPU.
#pragma once
#include <map>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

namespace mcpt {

struct HostScene {
    // per face, OBJ order
    std::vector<double> pos, nrm, uv;
    std::vector<int32_t> mat_id;
    std::vector<mcpt_material> materials;
    std::vector<std::string> material_names;
    std::vector<std::vector<uint8_t>> texture_bytes;
    std::vector<mcpt_texture> textures;
    std::vector<mcpt_light> lights;
    mcpt_camera camera{};
    mcpt_scene_desc desc{};
    std::map<std::string, int> material_index;
    size_t n_vertices = 0;
    void finalize();  // (re)point desc at the vectors
};

// read_scene of MTPC/sceneManagement.cpp:264-274: <base>.mtl, <base>.obj, <base>.camera.
// Throws std::runtime_error with a message on a missing file, unknown material, bad index or
// missing texture sidecar (the reference checks none of these and crashes later).
void load_scene(const std::string& path, const std::string& name, HostScene& out);
void load_tables(const std::string& path, const std::string& name, HostScene& out, std::string& camera_error);

}  // namespace mcpt

// The reference's entry point (MTPC/MTPC.cpp:35) and the README's 2-argument form (README.md:15).
bool render_scene(std::string path, std::string filename, int N_ray_per_pixel);
bool render_scene(std::string filename, int N_ray_per_pixel);

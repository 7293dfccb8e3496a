// context.h — the opaque mcpt_ctx of include/mcpt.h (one GPU), shared by api.cu and multigpu.cu.
#pragma once
#include <cstddef>
#include <cstdint>
#include "lbvh.h"
#include "render.h"
#include "obj_ingest.h"

#define PIN_RING 4

struct mcpt_comm;   // multigpu.cu: this rank's NCCL communicator (one process per GPU)

struct mcpt_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;  // ev3..ev2 bracket the integrator kernel
    mcpt::BuildStreams build_streams{};                  // side stream + fork/join events of mcpt_build_bvh
    uint64_t seed = 0;
    mcpt::SceneBuffers sb;
    mcpt::DMaterial* d_materials = nullptr; mcpt::DTexture* d_textures = nullptr; unsigned char* d_tex_bytes = nullptr;
    mcpt::DLight* d_lights = nullptr; mcpt::LightTri* d_light_tris = nullptr; float* d_light_cdf = nullptr;
    int n_materials = 0, n_textures = 0, n_lights = 0;
    float area0 = 0.f;
    unsigned long long* d_counters = nullptr;
    float* d_planes = nullptr; size_t planes_cap = 0;   // scratch: sample-chunk planes
    float* d_image = nullptr; size_t image_cap = 0;     // scratch: reduced image
    float* d_dirs = nullptr; size_t dirs_cap = 0;       // scratch: primary ray directions (K7)
    unsigned char* d_png = nullptr; size_t png_cap = 0; // scratch: device-encoded PNG (+64 B of sums/CRC in front)
    char* d_wf = nullptr; size_t wf_cap = 0;            // scratch: wavefront slot state + ray queue
    mcpt::WfHost wfh;                                         // streams/events of the wavefront loop
    void* d_io = nullptr; size_t io_cap = 0;            // scratch: rays in / hits out of mcpt_trace
    char* d_text = nullptr; size_t text_cap = 0;        // scratch: OBJ text of mcpt_ingest_obj
    mcpt::DevArena ingest_text, ingest_stmt;                  // scratch: ingest temporaries
    mcpt::DevArena scene_arena, tables_arena;                 // grow-only: a re-upload of a scene that fits allocates nothing
    char* h_pin[PIN_RING] = {}; cudaEvent_t pin_ev[PIN_RING] = {};   // pinned ring of mcpt_ingest_obj_file (one buffer per reader thread)
    int last_ingest_launches = 0;
    float last_build_ms = 0.f;
    int last_build_launches = 0;
    int coop_blocks[2] = {0, 0};   // co-resident CTAs of wide_collapse_kernel / ploc_rounds_kernel on this device
    bool scene_loaded = false;
    mcpt_comm* comm = nullptr;                          // mcpt_comm_init (multigpu.cu); destroyed with the context
    float* d_stacked = nullptr; size_t stacked_cap = 0; // scratch: R partial images side by side (exchange target on rank 0)
};

void mcpt_comm_release(mcpt_ctx* c);   // multigpu.cu



// Launchers of the traversal / integrator / reduction kernels (render.cu).
#pragma once
#include <vector>
#include "device_types.cuh"

namespace mcpt {
// K5: closest hit for n rays (org_dir = n x 6 floats, device). hits = n x mcpt_hit (device).
int launch_trace(const DeviceScene& sc, const float* d_org_dir, int n, mcpt_hit* d_hits, bool count,
                 unsigned long long* d_counters, cudaStream_t stream);
// K7: primary ray directions (height*width*3 floats), bit-faithful to generateImg's repeated-addition pixel walk.
int launch_raygen(const DCamera& cam, float* d_dirs, double* d_pos, cudaStream_t stream);
// K6: megakernel integrator; fills P.planes (P.n_chunks planes).
int launch_megakernel(const RenderParams& P, int flags, cudaStream_t stream);
// K8+K9: wavefront integrator (wavefront.cu); same planes as the megakernel, bit for bit.  The host reads the number of live path
// slots one batch of iterations behind the GPU, so the stream never idles waiting for it.
int wavefront_slots(const RenderParams& P);                    // < 0: too many slots for the 26-bit ray id
size_t wavefront_scratch_bytes(int n_slots, int n_lights, long long n_pixels);
enum { WF_MAX_PIPES = 8 };
struct WfHost {   // per-context host resources of the wavefront loop (created on first use)
    std::vector<cudaStream_t> side_streams;
    std::vector<cudaEvent_t> events;        // fork, join and per-pipe batch events
    std::vector<cudaEvent_t> trace_events;  // one timing pair per wf_trace launch of the last render
    unsigned* h_alive = nullptr;            // pinned: live-slot counts read back per pipe and batch parity
    int iterations = 0;                     // wf_trace launches of the last render
    void destroy();
};
int launch_wavefront(const RenderParams& P, int flags, void* d_scratch, WfHost& H, cudaStream_t stream, int* launches);
// K12: out[i] = scale * fixed pairwise tree sum over n planes (n in {1,2,4,8}).
int launch_sum_planes(const float* d_planes, int n, long long count, float scale, float* d_out, cudaStream_t stream);
// K13-K15 (png_device.cu): quantise + frame the image as an uncompressed PNG on the device.
size_t png_encoded_size(int w, int h);
int launch_png_encode(const float* d_rgb, int w, int h, unsigned char* d_png, void* d_scratch32, cudaStream_t stream);
}  // namespace mcpt

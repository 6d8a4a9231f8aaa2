// Interface of the headless body renderer (render.cu).  render.cu is
// its own translation unit because it is compiled with -fmad=false: GLSL floats are
// reproduced as single IEEE-rounded fp32 operations (see the header of render.cu).
#pragma once
#include "common.cuh"

namespace nb {

// THREE.PerspectiveCamera (idx:704-705) looking at `target` with +y up, as Matrix4.lookAt
// builds it: za = normalize(position - target), xa = normalize(cross(up, za)), ya =
// cross(za, xa).  Built on the host in double, stored as float (WebGL uniforms are fp32).
struct RenderCamera {
    float pos[3], xa[3], ya[3], za[3];
    float tan_half;  // tan(fov / 2), fov = 45 degrees in the page
    float aspect;    // width / height
    float near;      // 0.1 in the page
};

// Per-context scratch of the tile binning (render.cu): bodies per tile, list offsets, the lists.
struct RenderScratch {
    unsigned* counts = nullptr;
    unsigned* offsets = nullptr;
    unsigned* list = nullptr;
    size_t ntiles_cap = 0, list_cap = 0;
    size_t last_list_entries = 0;
};
void render_scratch_free(RenderScratch& s);

// rgba: width*height float4 (gl_FragColor before tone mapping; background 0,0,0,0); ids: width*height or nullptr.
// Synchronises the stream once per frame (the tile-list length is read back to size the lists).
template <typename Real>
cudaError_t render_launch(const Real* pos, const Real* radius, const Real* temp, const int* type, const float* seed, size_t n,
                          const RenderCamera& cam, int w, int h, RenderScratch& scratch, float4* rgba, int* ids,
                          cudaStream_t stream);

}  // namespace nb

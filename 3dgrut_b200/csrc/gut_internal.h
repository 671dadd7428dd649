// gut_internal.h -- structures shared by the translation units of libgut_b200 (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gut_b200.h"
#include "gut_math.cuh"

// Per-frame constants computed on the host once per call
// (reference: GUTRenderer::renderForward, src/gutRenderer.cu:259-267,282-285,385).
struct GutFrame {
    uint32_t width, height;
    uint32_t grid_x, grid_y;    // tiles
    uint32_t num_tiles;
    GutMat4x3 view;             // world -> sensor at the mid-shutter pose (depth ordering, z cull)
    GutMat4x3 c2w;              // sensor -> world at the mid-shutter pose (ray set-up)
    GutMat4x3 start;            // world -> sensor at the shutter-START pose: what every sigma point of K1 is projected with first
    float3 cam_centre;          // sensor position in world space (SH view direction)
    float ut_delta;             // sqrt(alpha^2 (3 + kappa)), narrowed from double (setup_3dgut.py:45)
};

// What the compositing kernels stage per sorted (tile, particle) intersection.  The sorted particle list of a tile
// (u32 indices, contiguous) is read with coalesced loads two batches ahead; the particle parameters it points to are
// gathered with 16-byte cp.async copies one batch ahead of the walk and expanded in shared memory into:
//   GutRecord (80 B = 5 x float4): ms = diag(1/scale) * R^T maps world offsets to the particle's canonical
//     (unit-Gaussian) frame; go = ms * (frame origin - position) is the canonical ray origin of every ray that
//     starts at the frame origin (GutFrame::cam_centre) -- all of them for pinhole / fisheye ray bundles -- so a
//     ray/particle test needs ONE mat-vec (the direction), not two;
//   GutRecordBwd (32 B, backward only): 1 / scale and scale^2.
struct __align__(16) GutRecord {
    float ms0x, ms0y, ms0z, gox;    // row 0 of ms | go.x
    float ms1x, ms1y, ms1z, goy;    // row 1       | go.y
    float ms2x, ms2y, ms2z, goz;    // row 2       | go.z
    float sx, sy, sz, density;      // scale       | density
    float fr, fg, fb;               // max(feature, 0)
    uint32_t idx;                   // particle index
};
static_assert(sizeof(GutRecord) == 80, "record must be 5 float4");
struct __align__(16) GutRecordBwd {
    float isx, isy, isz, pad0;      // 1 / scale
    float s2x, s2y, s2z, pad1;      // scale^2
};
static_assert(sizeof(GutRecordBwd) == 32, "backward record must be 2 float4");

// Global-memory sources of the per-intersection data.
struct GutParticleSource {
    const uint32_t* sorted_idx;   // [M]  particle index per sorted intersection, tiles contiguous (tile_start)
    const float4* density12;      // [N,3] {pos, density | quat wxyz | scale, pad}
    const float4* features;       // [N]  SH radiance + 0.5 (unclamped) from K1
};

// Device buffers of the per-particle projection stage (K1 outputs).
struct GutProjBuffers {
    uint32_t* tiles_count;   // [N]
    float2* proj_pos;        // [N]
    float4* conic_opacity;   // [N]
    float2* extent;          // [N]
    float* depth;            // [N]
    float4* features;        // [N] xyz = SH radiance + 0.5 (unclamped), w unused
    uint4* emit_rec;         // [N] culled tile box of the particle for the emit pass (gut_math.cuh, GUT_EMIT_*)
};

// ---- launchers implemented in the .cu files (all enqueue on `stream`, return cudaGetLastError()) ----
cudaError_t gut_launch_project(const GutConfig& cfg, const GutCamera& cam, const GutFrame& frame, uint32_t N, int sph_degree,
                               const float* density12, const float* sph, GutProjBuffers out, float* visibility,
                               uint32_t* tile_hist, cudaStream_t stream);

cudaError_t gut_launch_project_backward(const GutConfig& cfg, const GutFrame& frame, uint32_t N, int sph_degree,
                                        const float* density12, const uint32_t* tiles_count, const float4* features,
                                        float* raw_grad, float* density_grad, float* sph_grad,
                                        float* colour_out /* non-null: write the view's masked colour gradient instead of sph_grad */,
                                        int colour_compact /* layout of colour_out: 0 = [N,3], 1 = compact slot (gut_project.cu, colour_store) */,
                                        float* a_sum /* [N,4] sum of A over the views of a deferred step; NULL for the one-view backward */,
                                        int accumulate, cudaStream_t stream);

// deferred SH backward of a multi-view step (gut_project.cu, gut_sh_expand_kernel)
cudaError_t gut_launch_sh_expand(uint32_t N, int sph_degree, int ncoef_all, const float* density12, uint32_t n_views, const float* centres /* device [n_views,3] */,
                                 const float* colour_grads, float* sph_grad, int accumulate, cudaStream_t stream);

// exclusive scan of tile_hist[T*SUB] -> tile_start[T+1] and absolute sub-bucket cursors tile_cursor[T*SUB]; stats[0]=M, stats[1]=max list
cudaError_t gut_launch_sh_expand_ptrs(uint32_t N, int sph_degree, int ncoef_all, const float* density12, uint32_t n_views, const float* centres /* device [n_views,3] */,
                                      const float* const* colour_ptrs /* host array of device pointers */, int colour_compact, float* sph_grad, int accumulate,
                                      cudaStream_t stream);
cudaError_t gut_launch_tile_scan(uint32_t T, const uint32_t* tile_hist, uint32_t* tile_start, uint32_t* tile_cursor,
                                 uint32_t* stats, cudaStream_t stream);

cudaError_t gut_launch_emit(const GutConfig& cfg, const GutFrame& frame, uint32_t N, GutProjBuffers proj,
                            uint32_t* tile_cursor, uint2* entries, uint32_t capacity, cudaStream_t stream);

// per-tile sort of the (depth, idx) buckets -> sorted particle list; scratch = alt buffer for oversize tiles;
// tile_flags [T] = scratch words (buckets handed from the counting sort to the radix sort)
cudaError_t gut_launch_tile_sort(const GutFrame& frame, uint32_t N, uint32_t max_list, const uint32_t* tile_start, uint2* entries, uint2* scratch,
                                 uint32_t* tile_flags, uint32_t* flagged_count /* zeroed by the tile scan */, uint32_t* sorted_idx,
                                 int sm_count /* of the handle's device */, cudaStream_t stream);
// per-device kernel attributes (dynamic shared memory opt-in) + SM count of the CURRENT device; called by gut_create
cudaError_t gut_configure_tile_sort(int* sm_count);

// debug/parity: rebuild the reference-shaped sorted key array and tile ranges
cudaError_t gut_launch_export_keys(const GutFrame& frame, const uint32_t* tile_start, const uint32_t* sorted_idx, const float* depth,
                                   uint64_t* sorted_keys, uint32_t* ranges, cudaStream_t stream);

cudaError_t gut_launch_render(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters /* non-null: count tests/hits instead of rendering */,
                              uint32_t* hit_masks /* nullable: per (block, batch) masks of accepted entries for the backward */,
                              bool block_cull /* conservative per-block cone cull while staging (gut_set_block_cull) */, cudaStream_t stream);
size_t gut_hit_mask_words(uint64_t num_intersections, uint32_t num_tiles);

cudaError_t gut_launch_render_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                       GutParticleSource src, const float* ray_ori, const float* ray_dir,
                                       const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                       float* raw_grad /* [N,16] accumulators, see gut_render.cu backward_walk */,
                                       const uint32_t* hit_masks /* written by the forward */, cudaStream_t stream);

// k_buffer_size > 0 (sorted compositing, gut_render_kbuffer.cu): same arguments as the two launchers above
bool gut_kbuffer_size_supported(int k);
cudaError_t gut_launch_render_kbuffer(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                                      const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                                      cudaStream_t stream);
cudaError_t gut_launch_render_kbuffer_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                               GutParticleSource src, const float* ray_ori, const float* ray_dir, const float* rgba,
                                               const float* rgba_grad, const float* dist, const float* dist_grad, float* raw_grad,
                                               cudaStream_t stream);

// [N,3] positions, [N,1] density, [N,4] rotation, [N,3] scale <-> [N,12] (gut_project.cu)
cudaError_t gut_launch_pack_particles(uint32_t N, const float* pos, const float* dens, const float* rot, const float* scale, float* out12,
                                      cudaStream_t stream);
cudaError_t gut_launch_unpack_gradient(uint32_t N, const float* grad12, float* pos, float* dens, float* rot, float* scale, cudaStream_t stream);

// model-side passes (gut_model.cu): fused activations + packing, feature cat / split, SelectiveAdam, densification statistics
cudaError_t gut_launch_activate_pack(uint32_t N, const float* pos, const float* rot, const float* scale, const float* dens, int scale_act,
                                     int density_act, float* out12, cudaStream_t stream);
cudaError_t gut_launch_activate_backward(uint32_t N, const float* rot, const float* packed12, const float* grad12, int scale_act, int density_act,
                                         float* pos_g, float* rot_g, float* scale_g, float* dens_g, cudaStream_t stream);
cudaError_t gut_launch_pack_features(uint32_t N, uint32_t ncoef, const float* albedo, const float* specular, float* out, cudaStream_t stream);
cudaError_t gut_launch_unpack_features(uint32_t N, uint32_t ncoef, const float* grad, float* albedo_g, float* specular_g, cudaStream_t stream);
cudaError_t gut_launch_selective_adam(uint32_t N, const GutAdamGroup* groups, uint32_t n_groups, const float* visibility, float b1, float b2,
                                      float eps, int zero_grad, cudaStream_t stream);
cudaError_t gut_launch_density_adam(uint32_t N, const float* packed12, const float* grad12, const GutAdamTensors& pos, const GutAdamTensors& rot,
                                    const GutAdamTensors& scale, const GutAdamTensors& dens, int scale_act, int density_act,
                                    const float* visibility, float b1, float b2, float eps, const float* sensor /* host [3] or null */,
                                    float* accum, float* denom, cudaStream_t stream);
cudaError_t gut_launch_sh_expand_adam(uint32_t N, int sph_degree, const float* density12, uint32_t n_views, const float* centres,
                                      const float* const* colour_ptrs, int colour_compact, const GutAdamTensors& albedo, const GutAdamTensors& specular,
                                      const float* visibility, float b1, float b2, float eps, cudaStream_t stream);
cudaError_t gut_launch_densify_stats(uint32_t N, const float* pos, const float* pos_g, float3 sensor, float* accum, float* denom,
                                     cudaStream_t stream);

cudaError_t gut_launch_fill_outputs(uint32_t num_pixels, float* rgba, float* dist, float* hits, cudaStream_t stream);
cudaError_t gut_launch_copy_features3(uint32_t N, const float4* features4, float* features3, cudaStream_t stream);

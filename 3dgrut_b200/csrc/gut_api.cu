// gut_api.cu -- the C ABI of libgut_b200 (include/gut_b200.h): handle, workspace, stage sequencing.
//
// Replaces the reference's SplatRaster (threedgut_tracer/src/splatRaster.cpp) and GUTRenderer host code
// (threedgut_tracer/src/gutRenderer.cu:99-497).  No torch, no pybind: raw device pointers and a stream.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "gut_internal.h"

namespace {

enum Stage { ST_PROJECT = 0, ST_SCAN, ST_EMIT, ST_SORT, ST_RENDER, ST_BWD_RENDER, ST_BWD_PROJECT, ST_COUNT };
const char* kStageNames[ST_COUNT] = {"render::project", "render::prepare-expand", "render::expand", "render::sort",
                                     "render::render",  "render-backward::render", "render-backward::project"};

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
    int tag = 0;
};

}  // namespace

struct GutHandle {
    GutConfig cfg;
    int device = 0;
    std::string error;
    uint64_t launches = 0;

    // per-particle workspace (K1 outputs) -- grow only, like GutRenderForwardContext (gutRenderer.cu:166-203)
    DevBuf tiles_count, proj_pos, conic, extent, depth, features, emit_rec, raw_grad;
    DevBuf a_sum;   // [N,4] sum over the views of a deferred step of K5's A accumulator (gut_project.cu, origin fold)
    // per-tile
    DevBuf tile_hist, tile_start, tile_cursor, stats;
    // per-intersection
    DevBuf entries, scratch, sorted_idx, hit_masks;
    // host-buffer mode (gut_*_host)
    DevBuf h_ray_ori, h_ray_dir, h_rgba, h_dist, h_hits, h_rgba_grad, h_dist_grad;
    uint32_t* pinned_stats = nullptr;  // [2] = {M, max tile list}

    // forward context consumed by backward
    bool fwd_valid = false;
    bool training = true;           // forward records the hit masks the backward walks (gut_set_training)
    bool fwd_masks = false;         // the last forward recorded them
    bool pending_density = false;   // gut_backward_colour left view-summed accumulators for gut_density_gradient_finish
    uint32_t pending_N = 0;         // particle count of that pending accumulation
    int colour_compact = 0;         // layout of the colour-gradient slots this handle writes and reads (gut_set_colour_layout)
    bool block_cull = true;         // conservative per-block cull of the K = 0 forward (gut_set_block_cull: test control)
    int sm_count = 0;               // SMs of the handle's device (grid of the persistent radix-sort kernel)
    cudaStream_t fwd_stream = nullptr;
    uint32_t fwd_N = 0, fwd_W = 0, fwd_H = 0;
    uint64_t fwd_M = 0;
    uint32_t fwd_max_list = 0;
    GutFrame fwd_frame = {};
    const float* fwd_density = nullptr;  // particle parameters of the last forward (borrowed; used by gut_measure_work)

    // timers (SplatRaster::CudaTimer, splatRaster.cpp:108-151): keep the last 256
    std::deque<EventPair> timers;        // tag 0 = forward_render, 1 = backward_render
    std::deque<EventPair> stage_timers;  // tag = Stage
    float last_fwd_ms = -1.0f, last_bwd_ms = -1.0f;
    float last_stage_ms[ST_COUNT];
};

namespace {

int fail(GutHandle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->error = buf;
    return code;
}

#define GUT_CUDA(h, call)                                                                                    \
    do {                                                                                                     \
        cudaError_t e_ = (call);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            return fail(h, e_ == cudaErrorMemoryAllocation ? GUT_ERR_OUT_OF_MEMORY : GUT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, \
                        cudaGetErrorString(e_), __FILE__, __LINE__);                                         \
    } while (0)

cudaError_t ensure(DevBuf& b, size_t bytes, cudaStream_t s, float slack = 1.0f) {
    if (bytes <= b.cap && b.p) return cudaSuccess;
    if (b.p) {
        cudaError_t e = cudaFreeAsync(b.p, s);
        if (e != cudaSuccess) return e;
        b.p = nullptr; b.cap = 0;
    }
    size_t want = (size_t)((double)bytes * slack);
    want = (want + 255) & ~(size_t)255;
    if (want == 0) want = 256;
    cudaError_t e = cudaMallocAsync(&b.p, want, s);
    if (e != cudaSuccess) { b.p = nullptr; return e; }
    b.cap = want;
    return cudaSuccess;
}

void release(DevBuf& b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.cap = 0;
}

// ---- host pose algebra (reference: include/3dgut/sensors/sensors.h:44-73, tcnn quaternion helpers vec.h:1066-1183) ----
struct HQuat { float w, x, y, z; };

void quat_cols(HQuat q, float3& c0, float3& c1, float3& c2) { quat_to_columns(q.w, q.x, q.y, q.z, c0, c1, c2); }

HQuat quat_from_cols(float3 c0, float3 c1, float3 c2) {
    // m[i][j] = column i, row j
    const float m00 = c0.x, m01 = c0.y, m02 = c0.z, m10 = c1.x, m11 = c1.y, m12 = c1.z, m20 = c2.x, m21 = c2.y, m22 = c2.z;
    HQuat q;
    const float tr = m00 + m11 + m22;
    if (tr > 0.0f) {
        const float S = sqrtf(tr + 1.0f) * 2.0f;
        q.w = 0.25f * S; q.x = (m12 - m21) / S; q.y = (m20 - m02) / S; q.z = (m01 - m10) / S;
    } else if (m00 > m11 && m00 > m22) {
        const float S = sqrtf(1.0f + m00 - m11 - m22) * 2.0f;
        q.w = (m12 - m21) / S; q.x = 0.25f * S; q.y = (m10 + m01) / S; q.z = (m20 + m02) / S;
    } else if (m11 > m22) {
        const float S = sqrtf(1.0f + m11 - m00 - m22) * 2.0f;
        q.w = (m20 - m02) / S; q.x = (m10 + m01) / S; q.y = 0.25f * S; q.z = (m21 + m12) / S;
    } else {
        const float S = sqrtf(1.0f + m22 - m00 - m11) * 2.0f;
        q.w = (m01 - m10) / S; q.x = (m20 + m02) / S; q.y = (m21 + m12) / S; q.z = 0.25f * S;
    }
    return q;
}

HQuat host_slerp(HQuat a, HQuat b, float t) {
    float c = (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z);
    if (c < 0.0f) { b.w = -b.w; b.x = -b.x; b.y = -b.y; b.z = -b.z; c = -c; }
    HQuat r;
    if (c > 1.0f - 1.1920928955078125e-07f) {
        const float u = 1.0f - t;
        r.w = a.w * u + b.w * t; r.x = a.x * u + b.x * t; r.y = a.y * u + b.y * t; r.z = a.z * u + b.z * t;
    } else {
        const float ang = acosf(c);
        const float s0 = sinf((1.0f - t) * ang), s1 = sinf(t * ang), sd = sinf(ang);
        r.w = (s0 * a.w + s1 * b.w) / sd; r.x = (s0 * a.x + s1 * b.x) / sd;
        r.y = (s0 * a.y + s1 * b.y) / sd; r.z = (s0 * a.z + s1 * b.z) / sd;
    }
    return r;
}

GutFrame make_frame(const GutConfig& cfg, const GutCamera& cam, uint32_t W, uint32_t H) {
    GutFrame f;
    f.width = W; f.height = H;
    f.grid_x = (W + GUT_TILE - 1) / GUT_TILE;
    f.grid_y = (H + GUT_TILE - 1) / GUT_TILE;
    f.num_tiles = f.grid_x * f.grid_y;
    // mid-shutter pose: lerp translation, slerp rotation at 0.5 (gutRenderer.cu:266)
    const float* p0 = cam.start_pose; const float* p1 = cam.end_pose;
    const HQuat qm = host_slerp(HQuat{p0[6], p0[3], p0[4], p0[5]}, HQuat{p1[6], p1[3], p1[4], p1[5]}, 0.5f);
    const float3 tm = f3(p0[0] * (1.0f - 0.5f) + p1[0] * 0.5f, p0[1] * (1.0f - 0.5f) + p1[1] * 0.5f, p0[2] * (1.0f - 0.5f) + p1[2] * 0.5f);
    quat_cols(qm, f.view.c[0], f.view.c[1], f.view.c[2]);
    f.view.c[3] = tm;
    // shutter-start pose as a matrix (the same quat_to_columns the kernel would run, in the same IEEE arithmetic)
    quat_cols(HQuat{p0[6], p0[3], p0[4], p0[5]}, f.start.c[0], f.start.c[1], f.start.c[2]);
    f.start.c[3] = f3(p0[0], p0[1], p0[2]);
    // inverse pose: transpose -> quaternion -> matrix round trip, exactly as sensorPoseInverse + sensorPoseToMat do
    const float3 r0 = f.view.c[0], r1 = f.view.c[1], r2 = f.view.c[2];
    const float3 i0 = f3(r0.x, r1.x, r2.x), i1 = f3(r0.y, r1.y, r2.y), i2 = f3(r0.z, r1.z, r2.z);
    const HQuat qi = quat_from_cols(i0, i1, i2);
    const float3 ti = rotate_cols(-1.0f * i0, -1.0f * i1, -1.0f * i2, tm);
    quat_cols(qi, f.c2w.c[0], f.c2w.c[1], f.c2w.c[2]);
    f.c2w.c[3] = ti;
    f.cam_centre = ti;
    f.ut_delta = (float)sqrt((double)cfg.ut_alpha * (double)cfg.ut_alpha * (3.0 + (double)cfg.ut_kappa));
    return f;
}

bool degree_supported(int d) { return d == 0 || d == 1 || d == 2 || d == 3 || d == 4 || d == 5 || d == 8; }

void start_timer(std::deque<EventPair>& q, int tag, cudaStream_t s) {
    EventPair ep;
    ep.tag = tag;
    if (cudaEventCreate(&ep.a) != cudaSuccess) return;
    if (cudaEventCreate(&ep.b) != cudaSuccess) { cudaEventDestroy(ep.a); return; }
    cudaEventRecord(ep.a, s);
    q.push_back(ep);
    while (q.size() > 256) {
        cudaEventDestroy(q.front().a); cudaEventDestroy(q.front().b);
        q.pop_front();
    }
}
void stop_timer(std::deque<EventPair>& q, cudaStream_t s) {
    if (!q.empty()) cudaEventRecord(q.back().b, s);
}

struct StageTimer {  // RAII: times one stage when kernel timings are enabled
    GutHandle* h; cudaStream_t s; bool on;
    StageTimer(GutHandle* h_, int tag, cudaStream_t s_) : h(h_), s(s_), on(h_->cfg.enable_kernel_timings != 0) {
        if (on) start_timer(h->stage_timers, tag, s);
    }
    ~StageTimer() { if (on) stop_timer(h->stage_timers, s); }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int check_common(GutHandle* h, const GutCamera* camera, uint32_t N, int sph_degree, uint32_t W, uint32_t H) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!camera) return fail(h, GUT_ERR_BAD_INPUT, "camera is NULL");
    if (camera->model != GUT_CAMERA_OPENCV_PINHOLE && camera->model != GUT_CAMERA_OPENCV_FISHEYE)
        return fail(h, GUT_ERR_BAD_INPUT, "unknown camera model %d", camera->model);
    if (camera->shutter < 0 || camera->shutter > GUT_SHUTTER_GLOBAL) return fail(h, GUT_ERR_BAD_INPUT, "unknown shutter type %d", camera->shutter);
    if (W == 0 || H == 0) return fail(h, GUT_ERR_BAD_INPUT, "empty image %ux%u", W, H);
    if (sph_degree < 0 || sph_degree > h->cfg.sph_max_degree) return fail(h, GUT_ERR_BAD_INPUT, "active SH degree %d outside [0,%d]", sph_degree, h->cfg.sph_max_degree);
    if ((uint64_t)N * 12ull > 0xffffffffull * 4ull) return fail(h, GUT_ERR_BAD_INPUT, "too many particles");
    return GUT_OK;
}

}  // namespace

extern "C" {

void gut_default_config(GutConfig* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->abi_version = GUT_ABI_VERSION;
    cfg->kernel_degree = 2;
    cfg->min_kernel_response = 0.0113f;
    cfg->min_alpha = (float)(1.0 / 255.0);
    cfg->max_alpha = 0.99f;
    cfg->min_transmittance = 0.0001f;
    cfg->sph_max_degree = 3;
    cfg->enable_hitcounts = 1;
    cfg->rect_bounding = 1;
    cfg->tight_opacity_bounding = 1;
    cfg->tile_based_culling = 1;
    cfg->n_rolling_shutter_iterations = 5;
    cfg->ut_alpha = 1.0f; cfg->ut_beta = 2.0f; cfg->ut_kappa = 0.0f;
    cfg->ut_in_image_margin_factor = 0.1f;
    cfg->ut_require_all_sigma_points_valid = 0;
    cfg->k_buffer_size = 0;
    cfg->global_z_order = 1;
    cfg->enable_kernel_timings = 0;
}

int gut_abi_version(void) { return GUT_ABI_VERSION; }
size_t gut_sizeof_config(void) { return sizeof(GutConfig); }
size_t gut_sizeof_camera(void) { return sizeof(GutCamera); }
size_t gut_sizeof_stage_outputs(void) { return sizeof(GutStageOutputs); }
size_t gut_sizeof_frame_stats(void) { return sizeof(GutFrameStats); }
size_t gut_sizeof_adam_group(void) { return sizeof(GutAdamGroup); }
size_t gut_sizeof_adam_tensors(void) { return sizeof(GutAdamTensors); }

int gut_create(const GutConfig* cfg, int device, GutHandle** out) {
    if (!cfg || !out) return GUT_ERR_BAD_INPUT;
    *out = nullptr;
    if (cfg->abi_version != GUT_ABI_VERSION) return GUT_ERR_BAD_INPUT;
    if (!degree_supported(cfg->kernel_degree)) return GUT_ERR_UNSUPPORTED;
    if (cfg->sph_max_degree < 0 || cfg->sph_max_degree > 3) return GUT_ERR_UNSUPPORTED;
    if (cfg->ut_require_all_sigma_points_valid) return GUT_ERR_UNSUPPORTED;  // static_assert in threedgut.cuh:73
    if (cfg->k_buffer_size != 0 && !gut_kbuffer_size_supported(cfg->k_buffer_size)) return GUT_ERR_UNSUPPORTED;  // K in {0, 1, 2, 4, 8, 12, 16}
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return GUT_ERR_CUDA;
    GutHandle* h = new GutHandle();
    h->cfg = *cfg;
    h->device = device;
    for (int i = 0; i < ST_COUNT; ++i) h->last_stage_ms[i] = -1.0f;
    DeviceGuard g(device);
    if (cudaMallocHost(&h->pinned_stats, 2 * sizeof(uint32_t)) != cudaSuccess) { delete h; return GUT_ERR_CUDA; }
    // kernel attributes are per device: set them for THIS handle's device (a second handle on another GPU needs them too)
    if (gut_configure_tile_sort(&h->sm_count) != cudaSuccess || h->sm_count <= 0) {
        cudaGetLastError();
        cudaFreeHost(h->pinned_stats);
        delete h;
        return GUT_ERR_CUDA;
    }
    *out = h;
    return GUT_OK;
}

void gut_destroy(GutHandle* h) {
    if (!h) return;
    DeviceGuard g(h->device);
    cudaDeviceSynchronize();
    DevBuf* all[] = {&h->tiles_count, &h->proj_pos, &h->conic, &h->extent, &h->depth, &h->features, &h->emit_rec, &h->raw_grad, &h->a_sum, &h->tile_hist,
                     &h->tile_start, &h->tile_cursor, &h->stats, &h->entries, &h->scratch, &h->sorted_idx, &h->hit_masks,
                     &h->h_ray_ori, &h->h_ray_dir, &h->h_rgba, &h->h_dist, &h->h_hits, &h->h_rgba_grad, &h->h_dist_grad};
    for (DevBuf* b : all) release(*b);
    if (h->pinned_stats) cudaFreeHost(h->pinned_stats);
    for (auto& t : h->timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    for (auto& t : h->stage_timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    delete h;
}

const char* gut_last_error(GutHandle* h) { return h ? h->error.c_str() : "null handle"; }

int gut_forward(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                const float* d_radiance, uint32_t W, uint32_t H, const float* d_ray_ori, const float* d_ray_dir, float* d_rgba,
                float* d_dist, float* d_hits, float* d_visibility, GutStageOutputs* so) {
    int rc = check_common(h, camera, N, sph_degree, W, H);
    if (rc != GUT_OK) return rc;
    if (!d_ray_ori || !d_ray_dir || !d_rgba || !d_dist || !d_hits) return fail(h, GUT_ERR_BAD_INPUT, "null ray/output pointer");
    if (N > 0 && (!d_density || !d_radiance || !d_visibility)) return fail(h, GUT_ERR_BAD_INPUT, "null particle pointer");
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    h->fwd_valid = false;
    if (h->cfg.enable_kernel_timings) start_timer(h->timers, 0, s);

    const GutFrame fr = make_frame(h->cfg, *camera, W, H);
    const uint32_t T = fr.num_tiles;
    const size_t n1 = N ? N : 1;
    GUT_CUDA(h, ensure(h->tiles_count, n1 * 4, s));
    GUT_CUDA(h, ensure(h->proj_pos, n1 * 8, s));
    GUT_CUDA(h, ensure(h->conic, n1 * 16, s));
    GUT_CUDA(h, ensure(h->extent, n1 * 8, s));
    GUT_CUDA(h, ensure(h->depth, n1 * 4, s));
    GUT_CUDA(h, ensure(h->features, n1 * 16, s));
    GUT_CUDA(h, ensure(h->emit_rec, n1 * 16, s));
    GUT_CUDA(h, ensure(h->tile_hist, (size_t)T * GUT_TILE_SUB * 4, s));
    GUT_CUDA(h, ensure(h->tile_start, (2 * (size_t)T + 1) * 4, s));   // bucket offsets [T+1] + heavy-first tile order [T]
    GUT_CUDA(h, ensure(h->tile_cursor, (size_t)T * GUT_TILE_SUB * 4, s));
    GUT_CUDA(h, ensure(h->stats, 16, s));

    GutProjBuffers pb;
    pb.tiles_count = (uint32_t*)h->tiles_count.p; pb.proj_pos = (float2*)h->proj_pos.p; pb.conic_opacity = (float4*)h->conic.p;
    pb.extent = (float2*)h->extent.p; pb.depth = (float*)h->depth.p; pb.features = (float4*)h->features.p; pb.emit_rec = (uint4*)h->emit_rec.p;

    {
        StageTimer t(h, ST_PROJECT, s);
        GUT_CUDA(h, cudaMemsetAsync(h->tile_hist.p, 0, (size_t)T * GUT_TILE_SUB * 4, s));
        GUT_CUDA(h, gut_launch_project(h->cfg, *camera, fr, N, sph_degree, d_density, d_radiance, pb, d_visibility, (uint32_t*)h->tile_hist.p, s));
        h->launches += N ? 1 : 0;
    }
    {
        StageTimer t(h, ST_SCAN, s);
        GUT_CUDA(h, gut_launch_tile_scan(T, (const uint32_t*)h->tile_hist.p, (uint32_t*)h->tile_start.p, (uint32_t*)h->tile_cursor.p,
                                         (uint32_t*)h->stats.p, s));
        h->launches += 1;
        // the one device->host dependency of the frame: M sizes the per-intersection buffers (reference: gutRenderer.cu:313-321)
        GUT_CUDA(h, cudaMemcpyAsync(h->pinned_stats, h->stats.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    }
    GUT_CUDA(h, cudaStreamSynchronize(s));
    const uint64_t M = h->pinned_stats[0];
    const uint32_t max_list = h->pinned_stats[1];

    h->fwd_stream = s; h->fwd_N = N; h->fwd_W = W; h->fwd_H = H; h->fwd_M = M; h->fwd_max_list = max_list; h->fwd_frame = fr; h->fwd_density = d_density;

    if (M == 0) {
        // nothing intersects the image: outputs keep their initial values (reference early return, gutRenderer.cu:323-325)
        GUT_CUDA(h, gut_launch_fill_outputs(W * H, d_rgba, d_dist, d_hits, s));
        h->launches += 1;
    } else {
        GUT_CUDA(h, ensure(h->entries, M * 8, s, 1.25f));
        GUT_CUDA(h, ensure(h->sorted_idx, M * 4, s, 1.25f));
        if (max_list > 12288) GUT_CUDA(h, ensure(h->scratch, M * 8, s, 1.25f));
        {
            StageTimer t(h, ST_EMIT, s);
            GUT_CUDA(h, gut_launch_emit(h->cfg, fr, N, pb, (uint32_t*)h->tile_cursor.p, (uint2*)h->entries.p, (uint32_t)M, s));
            h->launches += 1;
        }
        {
            StageTimer t(h, ST_SORT, s);
            GUT_CUDA(h, gut_launch_tile_sort(fr, N, max_list, (const uint32_t*)h->tile_start.p, (uint2*)h->entries.p, (uint2*)h->scratch.p,
                                             (uint32_t*)h->tile_hist.p /* free after the scan: bucket flags */, (uint32_t*)h->stats.p + 2, (uint32_t*)h->sorted_idx.p,
                                             h->sm_count, s));
            h->launches += max_list > 8192 ? 3 : 2;
        }
        {
            StageTimer t(h, ST_RENDER, s);
            const GutParticleSource src{(const uint32_t*)h->sorted_idx.p, (const float4*)d_density, (const float4*)h->features.p};
            if (h->cfg.k_buffer_size > 0)
                GUT_CUDA(h, gut_launch_render_kbuffer(h->cfg, fr, (const uint32_t*)h->tile_start.p, src, d_ray_ori, d_ray_dir,
                                                      d_rgba, d_dist, d_hits, s));
            else
            {
                uint32_t* masks = nullptr;
                if (h->training) {
                    GUT_CUDA(h, ensure(h->hit_masks, gut_hit_mask_words(M, fr.num_tiles) * 4, s));
                    masks = (uint32_t*)h->hit_masks.p;
                }
                h->fwd_masks = masks != nullptr;
                GUT_CUDA(h, gut_launch_render(h->cfg, fr, (const uint32_t*)h->tile_start.p, src, d_ray_ori, d_ray_dir,
                                              d_rgba, d_dist, d_hits, nullptr, masks, h->block_cull, s));
            }
            h->launches += 1;
        }
    }
    h->fwd_valid = true;
    if (h->cfg.enable_kernel_timings) stop_timer(h->timers, s);

    if (so) {
        so->num_intersections = M;
        if (so->d_tiles_count) GUT_CUDA(h, cudaMemcpyAsync(so->d_tiles_count, pb.tiles_count, (size_t)N * 4, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_position) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_position, pb.proj_pos, (size_t)N * 8, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_conic_opacity) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_conic_opacity, pb.conic_opacity, (size_t)N * 16, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_extent) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_extent, pb.extent, (size_t)N * 8, cudaMemcpyDeviceToDevice, s));
        if (so->d_global_depth) GUT_CUDA(h, cudaMemcpyAsync(so->d_global_depth, pb.depth, (size_t)N * 4, cudaMemcpyDeviceToDevice, s));
        if (so->d_features) { GUT_CUDA(h, gut_launch_copy_features3(N, pb.features, (float*)so->d_features, s)); h->launches += N ? 1 : 0; }
        if (so->d_sorted_keys || so->d_sorted_particle_idx || so->d_tile_ranges) {
            if ((so->d_sorted_keys || so->d_sorted_particle_idx) && M > so->capacity_intersections)
                return fail(h, GUT_ERR_BAD_INPUT, "stage outputs hold %llu intersections, frame produced %llu",
                            (unsigned long long)so->capacity_intersections, (unsigned long long)M);
            if (M > 0) {
                if (so->d_sorted_particle_idx) GUT_CUDA(h, cudaMemcpyAsync(so->d_sorted_particle_idx, h->sorted_idx.p, M * 4, cudaMemcpyDeviceToDevice, s));
                GUT_CUDA(h, gut_launch_export_keys(fr, (const uint32_t*)h->tile_start.p, (const uint32_t*)h->sorted_idx.p, pb.depth,
                                                   (uint64_t*)so->d_sorted_keys, (uint32_t*)so->d_tile_ranges, s));
                h->launches += 1;
            } else if (so->d_tile_ranges) {
                GUT_CUDA(h, cudaMemsetAsync(so->d_tile_ranges, 0, (size_t)T * 8, s));
            }
        }
    }
    return GUT_OK;
}

static int backward_impl(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                         const float* d_radiance, uint32_t W, uint32_t H, const float* d_ray_ori, const float* d_ray_dir, const float* d_rgba,
                         const float* d_rgba_grad, const float* d_dist, const float* d_dist_grad, float* d_density_grad, float* d_radiance_grad,
                         float* d_colour_grad, int32_t accumulate) {
    int rc = check_common(h, camera, N, sph_degree, W, H);
    if (rc != GUT_OK) return rc;
    (void)d_radiance;
    cudaStream_t s = (cudaStream_t)stream_;
    if (!h->fwd_valid || h->fwd_stream != s)
        return fail(h, GUT_ERR_NO_FORWARD, "cannot render backward: no forward context on this stream (reference: gutRenderer.cu:413-417)");
    if (h->fwd_N != N || h->fwd_W != W || h->fwd_H != H) return fail(h, GUT_ERR_NO_FORWARD, "backward shapes differ from the preceding forward");
    if (h->fwd_M > 0 && h->cfg.k_buffer_size == 0 && !h->fwd_masks)
        return fail(h, GUT_ERR_NO_FORWARD, "the preceding forward ran in inference mode (gut_set_training(h, 0)): no hit masks for the backward");
    if (!d_ray_ori || !d_ray_dir || !d_rgba || !d_rgba_grad || !d_dist || !d_dist_grad) return fail(h, GUT_ERR_BAD_INPUT, "null ray/gradient pointer");
    if (N > 0 && (!d_density || (!d_density_grad && !d_colour_grad) || (!d_radiance_grad && !d_colour_grad))) return fail(h, GUT_ERR_BAD_INPUT, "null particle pointer");
    if (!d_colour_grad && h->pending_density)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_backward after gut_backward_colour: call gut_density_gradient_finish (or gut_reset_gradients) first");
    if (d_colour_grad && h->pending_density && h->pending_N != N)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_backward_colour: %u particles, but the pending accumulation of this step holds %u "
                                          "(particle set changed mid-step: call gut_reset_gradients)", N, h->pending_N);
    if (N == 0) return GUT_OK;
    DeviceGuard guard(h->device);
    if (h->cfg.enable_kernel_timings) start_timer(h->timers, 1, s);
    const GutFrame& fr = h->fwd_frame;

    // per-particle accumulators of K5 ([N,16] floats): zeroed when (re)allocated, reset by K6 after every use
    if ((size_t)N * 64 > h->raw_grad.cap || !h->raw_grad.p) {
        GUT_CUDA(h, ensure(h->raw_grad, (size_t)N * 64, s));
        GUT_CUDA(h, cudaMemsetAsync(h->raw_grad.p, 0, h->raw_grad.cap, s));
    }
    if (h->fwd_M > 0) {
        StageTimer t(h, ST_BWD_RENDER, s);
        const GutParticleSource src{(const uint32_t*)h->sorted_idx.p, (const float4*)d_density, (const float4*)h->features.p};
        if (h->cfg.k_buffer_size > 0)
            GUT_CUDA(h, gut_launch_render_kbuffer_backward(h->cfg, fr, (const uint32_t*)h->tile_start.p, src, d_ray_ori, d_ray_dir, d_rgba,
                                                           d_rgba_grad, d_dist, d_dist_grad, (float*)h->raw_grad.p, s));
        else
            GUT_CUDA(h, gut_launch_render_backward(h->cfg, fr, (const uint32_t*)h->tile_start.p, src, d_ray_ori, d_ray_dir, d_rgba, d_rgba_grad, d_dist, d_dist_grad,
                                                   (float*)h->raw_grad.p, (const uint32_t*)h->hit_masks.p, s));
        h->launches += 1;
    }
    const bool deferred_view = d_colour_grad && !d_density_grad;
    if (deferred_view && ((size_t)N * 16 > h->a_sum.cap || !h->a_sum.p)) {
        GUT_CUDA(h, ensure(h->a_sum, (size_t)N * 16, s));
        GUT_CUDA(h, cudaMemsetAsync(h->a_sum.p, 0, h->a_sum.cap, s));
    }
    {
        StageTimer t(h, ST_BWD_PROJECT, s);
        GUT_CUDA(h, gut_launch_project_backward(h->cfg, fr, N, sph_degree, d_density, (const uint32_t*)h->tiles_count.p,
                                                (const float4*)h->features.p, (float*)h->raw_grad.p, d_density_grad, d_radiance_grad,
                                                d_colour_grad, h->colour_compact, deferred_view ? (float*)h->a_sum.p : nullptr, accumulate, s));
        h->launches += (d_colour_grad && h->colour_compact) ? 2 : 1;   // + the slot's cursor reset
    }
    if (h->cfg.enable_kernel_timings) stop_timer(h->timers, s);
    return GUT_OK;
}

int gut_backward(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                 const float* d_radiance, uint32_t W, uint32_t H, const float* d_ray_ori, const float* d_ray_dir, const float* d_rgba,
                 const float* d_rgba_grad, const float* d_dist, const float* d_dist_grad, float* d_density_grad, float* d_radiance_grad,
                 int32_t accumulate) {
    if (h && N > 0 && !d_radiance_grad) return fail(h, GUT_ERR_BAD_INPUT, "null particle pointer");
    return backward_impl(h, stream_, camera, N, sph_degree, d_density, d_radiance, W, H, d_ray_ori, d_ray_dir, d_rgba, d_rgba_grad, d_dist,
                         d_dist_grad, d_density_grad, d_radiance_grad, nullptr, accumulate);
}

int gut_backward_colour(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                        const float* d_radiance, uint32_t W, uint32_t H, const float* d_ray_ori, const float* d_ray_dir, const float* d_rgba,
                        const float* d_rgba_grad, const float* d_dist, const float* d_dist_grad, float* d_colour_grad, float* view_centre_out) {
    if (h && N > 0 && !d_colour_grad) return fail(h, GUT_ERR_BAD_INPUT, "null colour gradient pointer");
    int rc = backward_impl(h, stream_, camera, N, sph_degree, d_density, d_radiance, W, H, d_ray_ori, d_ray_dir, d_rgba, d_rgba_grad, d_dist,
                           d_dist_grad, nullptr, nullptr, d_colour_grad, 0);
    if (rc == GUT_OK) {
        h->pending_density = true;
        h->pending_N = N;
        if (view_centre_out) {
            view_centre_out[0] = h->fwd_frame.cam_centre.x; view_centre_out[1] = h->fwd_frame.cam_centre.y; view_centre_out[2] = h->fwd_frame.cam_centre.z;
        }
    }
    return rc;
}

int gut_backward_colour_density(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                                const float* d_radiance, uint32_t W, uint32_t H, const float* d_ray_ori, const float* d_ray_dir, const float* d_rgba,
                                const float* d_rgba_grad, const float* d_dist, const float* d_dist_grad, float* d_density_grad,
                                float* d_colour_grad, float* view_centre_out, int32_t accumulate) {
    if (h && N > 0 && (!d_colour_grad || !d_density_grad)) return fail(h, GUT_ERR_BAD_INPUT, "null gradient pointer");
    if (h && h->pending_density)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_backward_colour_density after gut_backward_colour: call gut_density_gradient_finish (or gut_reset_gradients) first");
    int rc = backward_impl(h, stream_, camera, N, sph_degree, d_density, d_radiance, W, H, d_ray_ori, d_ray_dir, d_rgba, d_rgba_grad, d_dist,
                           d_dist_grad, d_density_grad, nullptr, d_colour_grad, accumulate);
    if (rc == GUT_OK && view_centre_out) {
        view_centre_out[0] = h->fwd_frame.cam_centre.x; view_centre_out[1] = h->fwd_frame.cam_centre.y; view_centre_out[2] = h->fwd_frame.cam_centre.z;
    }
    return rc;
}

int gut_density_gradient_finish(GutHandle* h, void* stream_, uint32_t N, const float* d_density, float* d_density_grad, int32_t accumulate) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (N == 0) { h->pending_density = false; return GUT_OK; }
    if (!d_density || !d_density_grad) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (h->pending_density && h->pending_N != N)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_density_gradient_finish: %u particles, but the pending accumulation holds %u "
                                          "(call gut_reset_gradients after changing the particle set)", N, h->pending_N);
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    if ((size_t)N * 64 > h->raw_grad.cap || !h->raw_grad.p) {   // no gut_backward_colour ran: the sum is zero
        GUT_CUDA(h, ensure(h->raw_grad, (size_t)N * 64, s));
        GUT_CUDA(h, cudaMemsetAsync(h->raw_grad.p, 0, h->raw_grad.cap, s));
    }
    if ((size_t)N * 16 > h->a_sum.cap || !h->a_sum.p) {
        GUT_CUDA(h, ensure(h->a_sum, (size_t)N * 16, s));
        GUT_CUDA(h, cudaMemsetAsync(h->a_sum.p, 0, h->a_sum.cap, s));
    }
    GUT_CUDA(h, gut_launch_project_backward(h->cfg, h->fwd_frame, N, 0, d_density, nullptr, nullptr, (float*)h->raw_grad.p, d_density_grad,
                                            nullptr, nullptr, 0, (float*)h->a_sum.p, accumulate, s));
    h->launches += 1;
    h->pending_density = false;
    return GUT_OK;
}

int gut_set_training(GutHandle* h, int32_t training) {
    if (!h) return GUT_ERR_BAD_INPUT;
    h->training = training != 0;
    return GUT_OK;
}

int gut_set_block_cull(GutHandle* h, int32_t enabled) {
    if (!h) return GUT_ERR_BAD_INPUT;
    h->block_cull = enabled != 0;
    return GUT_OK;
}

int gut_reset_gradients(GutHandle* h, void* stream_) {
    if (!h) return GUT_ERR_BAD_INPUT;
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    if (h->raw_grad.p) GUT_CUDA(h, cudaMemsetAsync(h->raw_grad.p, 0, h->raw_grad.cap, s));
    if (h->a_sum.p) GUT_CUDA(h, cudaMemsetAsync(h->a_sum.p, 0, h->a_sum.cap, s));
    h->pending_density = false;
    h->pending_N = 0;
    return GUT_OK;
}

int gut_set_colour_layout(GutHandle* h, int32_t layout) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (layout != GUT_COLOUR_DENSE && layout != GUT_COLOUR_COMPACT) return fail(h, GUT_ERR_BAD_INPUT, "unknown colour-gradient layout %d", layout);
    if (h->pending_density && h->colour_compact != (layout == GUT_COLOUR_COMPACT))
        return fail(h, GUT_ERR_BAD_INPUT, "gut_set_colour_layout in the middle of a step (views pending): finish or reset it first");
    h->colour_compact = layout == GUT_COLOUR_COMPACT ? 1 : 0;
    return GUT_OK;
}

uint64_t gut_colour_slot_floats(uint32_t N, int32_t layout) {
    const uint64_t rows = ((uint64_t)3 * N + 3u) & ~(uint64_t)3u;
    return layout == GUT_COLOUR_COMPACT ? 4u + (uint64_t)gut_colour_table_words(N) + rows : rows;
}

int gut_sh_gradient_expand(GutHandle* h, void* stream_, uint32_t N, int32_t sph_degree, const float* d_density, uint32_t n_views,
                           const float* d_view_centres, const float* d_colour_grads, float* d_radiance_grad, int32_t accumulate) {
    const float* view_centres = d_view_centres;
    if (!h) return GUT_ERR_BAD_INPUT;
    if (sph_degree < 0 || sph_degree > h->cfg.sph_max_degree) return fail(h, GUT_ERR_BAD_INPUT, "sph_degree_active out of range");
    if (N == 0) return GUT_OK;
    if (!d_density || !d_radiance_grad || (n_views > 0 && (!view_centres || !d_colour_grads))) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (h->colour_compact && n_views > 0)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_sh_gradient_expand takes one dense [n_views,N,3] buffer; with the compact colour layout "
                                          "(gut_set_colour_layout) pass one slot pointer per view to gut_sh_gradient_expand_gather");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_sh_expand(N, sph_degree, (h->cfg.sph_max_degree + 1) * (h->cfg.sph_max_degree + 1), d_density, n_views, view_centres, d_colour_grads, d_radiance_grad, accumulate, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_sh_gradient_expand_gather(GutHandle* h, void* stream_, uint32_t N, int32_t sph_degree, const float* d_density, uint32_t n_views,
                                  const float* d_view_centres, const float* const* colour_grad_ptrs, float* d_radiance_grad, int32_t accumulate) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (sph_degree < 0 || sph_degree > h->cfg.sph_max_degree) return fail(h, GUT_ERR_BAD_INPUT, "sph_degree_active out of range");
    if (N == 0) return GUT_OK;
    if (!d_density || !d_radiance_grad || (n_views > 0 && (!d_view_centres || !colour_grad_ptrs))) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    for (uint32_t v = 0; v < n_views; ++v)
        if (!colour_grad_ptrs[v]) return fail(h, GUT_ERR_BAD_INPUT, "colour gradient pointer of view %u is NULL", v);
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_sh_expand_ptrs(N, sph_degree, (h->cfg.sph_max_degree + 1) * (h->cfg.sph_max_degree + 1), d_density, n_views, d_view_centres, colour_grad_ptrs, h->colour_compact, d_radiance_grad, accumulate,
                                          (cudaStream_t)stream_));
    h->launches += n_views == 0 ? 1 : (n_views + GUT_EXPAND_MAX_VIEWS - 1) / GUT_EXPAND_MAX_VIEWS;
    return GUT_OK;
}

static_assert(sizeof(cudaIpcMemHandle_t) == GUT_IPC_HANDLE_BYTES, "CUDA IPC handle size");

int gut_peer_buffer_create(GutHandle* h, size_t bytes, void** d_ptr_out, unsigned char* ipc_handle_out) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!d_ptr_out || !ipc_handle_out || bytes == 0) return fail(h, GUT_ERR_BAD_INPUT, "null pointer or empty buffer");
    DeviceGuard guard(h->device);
    void* p = nullptr;
    GUT_CUDA(h, cudaMalloc(&p, bytes));      // a plain allocation: stream-ordered (pool) memory cannot be exported through CUDA IPC
    cudaIpcMemHandle_t hd;
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return fail(h, GUT_ERR_CUDA, "exporting the peer buffer failed: %s", cudaGetErrorString(e));
    }
    memcpy(ipc_handle_out, &hd, sizeof(hd));
    *d_ptr_out = p;
    return GUT_OK;
}

int gut_peer_buffer_destroy(GutHandle* h, void* d_ptr) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!d_ptr) return GUT_OK;
    DeviceGuard guard(h->device);
    GUT_CUDA(h, cudaFree(d_ptr));
    return GUT_OK;
}

int gut_peer_buffer_open(GutHandle* h, const unsigned char* ipc_handle, void** d_ptr_out) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!ipc_handle || !d_ptr_out) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    DeviceGuard guard(h->device);
    cudaIpcMemHandle_t hd;
    memcpy(&hd, ipc_handle, sizeof(hd));
    void* p = nullptr;
    GUT_CUDA(h, cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
    *d_ptr_out = p;
    return GUT_OK;
}

int gut_peer_buffer_close(GutHandle* h, void* d_ptr) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!d_ptr) return GUT_OK;
    DeviceGuard guard(h->device);
    GUT_CUDA(h, cudaIpcCloseMemHandle(d_ptr));
    return GUT_OK;
}

int gut_pack_particles(GutHandle* h, void* stream_, uint32_t N, const float* d_positions, const float* d_density, const float* d_rotation,
                       const float* d_scale, float* d_particle_density) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (N == 0) return GUT_OK;
    if (!d_positions || !d_density || !d_rotation || !d_scale || !d_particle_density) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (((uintptr_t)d_rotation | (uintptr_t)d_particle_density) & 15) return fail(h, GUT_ERR_BAD_INPUT, "rotation / packed buffers must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_pack_particles(N, d_positions, d_density, d_rotation, d_scale, d_particle_density, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_unpack_particle_gradient(GutHandle* h, void* stream_, uint32_t N, const float* d_particle_density_grad, float* d_positions_grad,
                                 float* d_density_grad, float* d_rotation_grad, float* d_scale_grad) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (N == 0) return GUT_OK;
    if (!d_particle_density_grad || !d_positions_grad || !d_density_grad || !d_rotation_grad || !d_scale_grad) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (((uintptr_t)d_rotation_grad | (uintptr_t)d_particle_density_grad) & 15) return fail(h, GUT_ERR_BAD_INPUT, "rotation / packed buffers must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_unpack_gradient(N, d_particle_density_grad, d_positions_grad, d_density_grad, d_rotation_grad, d_scale_grad,
                                           (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

static bool valid_activation(int32_t a) { return a == GUT_ACT_NONE || a == GUT_ACT_EXP || a == GUT_ACT_SIGMOID; }

int gut_activate_pack(GutHandle* h, void* stream_, uint32_t N, const float* d_positions, const float* d_rotation_raw, const float* d_scale_raw,
                      const float* d_density_raw, int32_t scale_activation, int32_t density_activation, float* d_particle_density) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!valid_activation(scale_activation) || !valid_activation(density_activation)) return fail(h, GUT_ERR_BAD_INPUT, "unknown activation");
    if (N == 0) return GUT_OK;
    if (!d_positions || !d_rotation_raw || !d_scale_raw || !d_density_raw || !d_particle_density) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (((uintptr_t)d_rotation_raw | (uintptr_t)d_particle_density) & 15) return fail(h, GUT_ERR_BAD_INPUT, "rotation / packed buffers must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_activate_pack(N, d_positions, d_rotation_raw, d_scale_raw, d_density_raw, scale_activation, density_activation,
                                         d_particle_density, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_activate_pack_backward(GutHandle* h, void* stream_, uint32_t N, const float* d_rotation_raw, const float* d_particle_density,
                               const float* d_particle_density_grad, int32_t scale_activation, int32_t density_activation,
                               float* d_positions_grad, float* d_rotation_grad, float* d_scale_grad, float* d_density_grad) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!valid_activation(scale_activation) || !valid_activation(density_activation)) return fail(h, GUT_ERR_BAD_INPUT, "unknown activation");
    if (N == 0) return GUT_OK;
    if (!d_rotation_raw || !d_particle_density || !d_particle_density_grad || !d_positions_grad || !d_rotation_grad || !d_scale_grad || !d_density_grad)
        return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (((uintptr_t)d_rotation_raw | (uintptr_t)d_particle_density | (uintptr_t)d_particle_density_grad | (uintptr_t)d_rotation_grad) & 15)
        return fail(h, GUT_ERR_BAD_INPUT, "rotation / packed buffers must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_activate_backward(N, d_rotation_raw, d_particle_density, d_particle_density_grad, scale_activation, density_activation,
                                             d_positions_grad, d_rotation_grad, d_scale_grad, d_density_grad, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_pack_features(GutHandle* h, void* stream_, uint32_t N, const float* d_albedo, const float* d_specular, float* d_particle_radiance) {
    if (!h) return GUT_ERR_BAD_INPUT;
    const uint32_t ncoef = (uint32_t)((h->cfg.sph_max_degree + 1) * (h->cfg.sph_max_degree + 1));
    if (N == 0) return GUT_OK;
    if (!d_albedo || (!d_specular && ncoef > 1) || !d_particle_radiance) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (ncoef == 16 && ((uintptr_t)d_particle_radiance & 15)) return fail(h, GUT_ERR_BAD_INPUT, "[N,48] buffer must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_pack_features(N, ncoef, d_albedo, d_specular, d_particle_radiance, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_unpack_feature_gradient(GutHandle* h, void* stream_, uint32_t N, const float* d_particle_radiance_grad, float* d_albedo_grad,
                                float* d_specular_grad) {
    if (!h) return GUT_ERR_BAD_INPUT;
    const uint32_t ncoef = (uint32_t)((h->cfg.sph_max_degree + 1) * (h->cfg.sph_max_degree + 1));
    if (N == 0) return GUT_OK;
    if (!d_particle_radiance_grad || !d_albedo_grad || (!d_specular_grad && ncoef > 1)) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (ncoef == 16 && ((uintptr_t)d_particle_radiance_grad & 15)) return fail(h, GUT_ERR_BAD_INPUT, "[N,48] buffer must be 16-byte aligned");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_unpack_features(N, ncoef, d_particle_radiance_grad, d_albedo_grad, d_specular_grad, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_selective_adam_step(GutHandle* h, void* stream_, uint32_t N, const GutAdamGroup* groups, uint32_t num_groups, const float* d_visibility,
                            float beta1, float beta2, float eps, int32_t zero_grad) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (num_groups > GUT_ADAM_MAX_GROUPS) return fail(h, GUT_ERR_BAD_INPUT, "at most %d parameter groups per call", GUT_ADAM_MAX_GROUPS);
    if (N == 0 || num_groups == 0) return GUT_OK;
    if (!groups) return fail(h, GUT_ERR_BAD_INPUT, "groups is NULL");
    for (uint32_t i = 0; i < num_groups; ++i) {
        if (!groups[i].param || !groups[i].grad || !groups[i].exp_avg || !groups[i].exp_avg_sq) return fail(h, GUT_ERR_BAD_INPUT, "null pointer in group %u", i);
        if (groups[i].width == 0) return fail(h, GUT_ERR_BAD_INPUT, "group %u has width 0", i);
    }
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_selective_adam(N, groups, num_groups, d_visibility, beta1, beta2, eps, zero_grad, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

static bool adam_tensors_ok(const GutAdamTensors* t) { return t && t->param && t->exp_avg && t->exp_avg_sq; }

int gut_density_adam(GutHandle* h, void* stream_, uint32_t N, const float* d_density, const float* d_density_grad, const GutAdamTensors* positions,
                     const GutAdamTensors* rotation, const GutAdamTensors* scale, const GutAdamTensors* density, int32_t scale_act,
                     int32_t density_act, const float* d_visibility, float b1, float b2, float eps, const float* sensor_position,
                     float* d_accum, float* d_denom) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (N == 0) return GUT_OK;
    if (!d_density || !d_density_grad || !adam_tensors_ok(positions) || !adam_tensors_ok(rotation) || !adam_tensors_ok(scale) ||
        !adam_tensors_ok(density))
        return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    if (((uintptr_t)rotation->param | (uintptr_t)rotation->exp_avg | (uintptr_t)rotation->exp_avg_sq) & 15u)
        return fail(h, GUT_ERR_BAD_INPUT, "rotation tensors must be 16-byte aligned");
    if (scale_act < GUT_ACT_NONE || scale_act > GUT_ACT_SIGMOID || density_act < GUT_ACT_NONE || density_act > GUT_ACT_SIGMOID)
        return fail(h, GUT_ERR_BAD_INPUT, "unknown activation");
    if ((sensor_position != nullptr) != (d_accum != nullptr && d_denom != nullptr) || ((d_accum != nullptr) != (d_denom != nullptr)))
        return fail(h, GUT_ERR_BAD_INPUT, "densification statistics need sensor_position, d_grad_norm_accum and d_grad_norm_denom together");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_density_adam(N, d_density, d_density_grad, *positions, *rotation, *scale, *density, scale_act, density_act,
                                        d_visibility, b1, b2, eps, sensor_position, d_accum, d_denom, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_sh_expand_adam(GutHandle* h, void* stream_, uint32_t N, int32_t sph_degree, const float* d_density, uint32_t n_views,
                       const float* d_view_centres, const float* const* colour_grad_ptrs, const GutAdamTensors* albedo,
                       const GutAdamTensors* specular, const float* d_visibility, float b1, float b2, float eps) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (sph_degree < 0 || sph_degree > h->cfg.sph_max_degree) return fail(h, GUT_ERR_BAD_INPUT, "sph_degree_active out of range");
    if (h->cfg.sph_max_degree != 3) return fail(h, GUT_ERR_UNSUPPORTED, "gut_sh_expand_adam is built for 16 SH coefficients");
    if (n_views > GUT_EXPAND_MAX_VIEWS)
        return fail(h, GUT_ERR_UNSUPPORTED, "gut_sh_expand_adam takes at most %d views (the update is applied once, in one launch)", GUT_EXPAND_MAX_VIEWS);
    if (N == 0) return GUT_OK;
    if (!d_density || !adam_tensors_ok(albedo) || !adam_tensors_ok(specular) || (n_views > 0 && (!d_view_centres || !colour_grad_ptrs)))
        return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    for (uint32_t v = 0; v < n_views; ++v)
        if (!colour_grad_ptrs[v]) return fail(h, GUT_ERR_BAD_INPUT, "colour gradient pointer of view %u is NULL", v);
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_sh_expand_adam(N, sph_degree, d_density, n_views, d_view_centres, colour_grad_ptrs, h->colour_compact, *albedo, *specular,
                                          d_visibility, b1, b2, eps, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_densify_stats(GutHandle* h, void* stream_, uint32_t N, const float* d_positions, const float* d_positions_grad, const float* sensor_position,
                      float* d_accum, float* d_denom) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (N == 0) return GUT_OK;
    if (!d_positions || !d_positions_grad || !sensor_position || !d_accum || !d_denom) return fail(h, GUT_ERR_BAD_INPUT, "null pointer");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, gut_launch_densify_stats(N, d_positions, d_positions_grad, make_float3(sensor_position[0], sensor_position[1], sensor_position[2]),
                                         d_accum, d_denom, (cudaStream_t)stream_));
    h->launches += 1;
    return GUT_OK;
}

int gut_forward_host(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                     const float* d_radiance, uint32_t W, uint32_t H, const float* h_ray_ori, const float* h_ray_dir, float* h_rgba,
                     float* h_dist, float* h_hits, float* d_visibility) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!h_ray_ori || !h_ray_dir || !h_rgba || !h_dist || !h_hits) return fail(h, GUT_ERR_BAD_INPUT, "null host pointer");
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    const size_t P = (size_t)W * H;
    GUT_CUDA(h, ensure(h->h_ray_ori, P * 12, s)); GUT_CUDA(h, ensure(h->h_ray_dir, P * 12, s));
    GUT_CUDA(h, ensure(h->h_rgba, P * 16, s)); GUT_CUDA(h, ensure(h->h_dist, P * 4, s)); GUT_CUDA(h, ensure(h->h_hits, P * 4, s));
    GUT_CUDA(h, cudaMemcpyAsync(h->h_ray_ori.p, h_ray_ori, P * 12, cudaMemcpyHostToDevice, s));
    GUT_CUDA(h, cudaMemcpyAsync(h->h_ray_dir.p, h_ray_dir, P * 12, cudaMemcpyHostToDevice, s));
    int rc = gut_forward(h, stream_, camera, N, sph_degree, d_density, d_radiance, W, H, (const float*)h->h_ray_ori.p,
                         (const float*)h->h_ray_dir.p, (float*)h->h_rgba.p, (float*)h->h_dist.p, (float*)h->h_hits.p, d_visibility, nullptr);
    if (rc != GUT_OK) return rc;
    GUT_CUDA(h, cudaMemcpyAsync(h_rgba, h->h_rgba.p, P * 16, cudaMemcpyDeviceToHost, s));
    GUT_CUDA(h, cudaMemcpyAsync(h_dist, h->h_dist.p, P * 4, cudaMemcpyDeviceToHost, s));
    GUT_CUDA(h, cudaMemcpyAsync(h_hits, h->h_hits.p, P * 4, cudaMemcpyDeviceToHost, s));
    return GUT_OK;
}

int gut_backward_host(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree, const float* d_density,
                      const float* d_radiance, uint32_t W, uint32_t H, const float* h_rgba_grad, const float* h_dist_grad,
                      float* d_density_grad, float* d_radiance_grad, int32_t accumulate) {
    if (!h) return GUT_ERR_BAD_INPUT;
    if (!h_rgba_grad || !h_dist_grad) return fail(h, GUT_ERR_BAD_INPUT, "null host pointer");
    if (!h->fwd_valid || !h->h_rgba.p || h->fwd_W != W || h->fwd_H != H)
        return fail(h, GUT_ERR_NO_FORWARD, "gut_backward_host needs a preceding gut_forward_host of the same view");
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    const size_t P = (size_t)W * H;
    GUT_CUDA(h, ensure(h->h_rgba_grad, P * 16, s)); GUT_CUDA(h, ensure(h->h_dist_grad, P * 4, s));
    GUT_CUDA(h, cudaMemcpyAsync(h->h_rgba_grad.p, h_rgba_grad, P * 16, cudaMemcpyHostToDevice, s));
    GUT_CUDA(h, cudaMemcpyAsync(h->h_dist_grad.p, h_dist_grad, P * 4, cudaMemcpyHostToDevice, s));
    return gut_backward(h, stream_, camera, N, sph_degree, d_density, d_radiance, W, H, (const float*)h->h_ray_ori.p,
                        (const float*)h->h_ray_dir.p, (const float*)h->h_rgba.p, (const float*)h->h_rgba_grad.p, (const float*)h->h_dist.p,
                        (const float*)h->h_dist_grad.p, d_density_grad, d_radiance_grad, accumulate);
}

static void collect(std::deque<EventPair>& q, float* sums, int* counts, int ntags) {
    for (auto& t : q) {
        float ms = 0.0f;
        if (cudaEventSynchronize(t.b) == cudaSuccess && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess && t.tag < ntags) {
            sums[t.tag] += ms; counts[t.tag] += 1;
        }
        cudaEventDestroy(t.a); cudaEventDestroy(t.b);
    }
    q.clear();
}

int gut_timings(GutHandle* h, float* fwd_ms, float* bwd_ms) {
    if (!h) return GUT_ERR_BAD_INPUT;
    DeviceGuard guard(h->device);
    float sums[2] = {0, 0}; int counts[2] = {0, 0};
    collect(h->timers, sums, counts, 2);
    if (counts[0]) h->last_fwd_ms = sums[0] / counts[0];
    if (counts[1]) h->last_bwd_ms = sums[1] / counts[1];
    if (fwd_ms) *fwd_ms = h->last_fwd_ms;
    if (bwd_ms) *bwd_ms = h->last_bwd_ms;
    return GUT_OK;
}

int gut_num_stages(void) { return ST_COUNT; }
const char* gut_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }

int gut_stage_timings(GutHandle* h, float* ms_out) {
    if (!h || !ms_out) return GUT_ERR_BAD_INPUT;
    DeviceGuard guard(h->device);
    float sums[ST_COUNT]; int counts[ST_COUNT];
    for (int i = 0; i < ST_COUNT; ++i) { sums[i] = 0; counts[i] = 0; }
    collect(h->stage_timers, sums, counts, ST_COUNT);
    for (int i = 0; i < ST_COUNT; ++i) {
        if (counts[i]) h->last_stage_ms[i] = sums[i] / counts[i];
        ms_out[i] = h->last_stage_ms[i];
    }
    return GUT_OK;
}

int gut_measure_work(GutHandle* h, void* stream_, const float* d_ray_ori, const float* d_ray_dir, uint64_t* tests, uint64_t* hits) {
    if (!h || !tests || !hits) return GUT_ERR_BAD_INPUT;
    cudaStream_t s = (cudaStream_t)stream_;
    if (!h->fwd_valid || h->fwd_stream != s) return fail(h, GUT_ERR_NO_FORWARD, "gut_measure_work needs a preceding forward on this stream");
    *tests = 0; *hits = 0;
    if (h->fwd_M == 0) return GUT_OK;
    if (h->cfg.k_buffer_size > 0) return fail(h, GUT_ERR_UNSUPPORTED, "gut_measure_work counts the k_buffer_size == 0 walk only");
    DeviceGuard guard(h->device);
    GUT_CUDA(h, ensure(h->stats, 16, s));
    GUT_CUDA(h, cudaMemsetAsync(h->stats.p, 0, 16, s));
    const GutParticleSource src{(const uint32_t*)h->sorted_idx.p, (const float4*)h->fwd_density, (const float4*)h->features.p};
    GUT_CUDA(h, gut_launch_render(h->cfg, h->fwd_frame, (const uint32_t*)h->tile_start.p, src, d_ray_ori, d_ray_dir,
                                  nullptr, nullptr, nullptr, (unsigned long long*)h->stats.p, nullptr, h->block_cull, s));
    unsigned long long host[2] = {0, 0};
    GUT_CUDA(h, cudaMemcpyAsync(host, h->stats.p, 16, cudaMemcpyDeviceToHost, s));
    GUT_CUDA(h, cudaStreamSynchronize(s));
    *tests = host[0]; *hits = host[1];
    return GUT_OK;
}

int gut_frame_stats(GutHandle* h, GutFrameStats* out) {
    if (!h || !out) return GUT_ERR_BAD_INPUT;
    out->num_intersections = h->fwd_M;
    out->num_tiles = h->fwd_frame.num_tiles;
    out->max_tile_list_length = h->fwd_max_list;
    out->kernel_launches = h->launches;
    return GUT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// gut_project_noalloc (include/gut_b200.h): the projection stage with its scratch in a file-scope device array.
// No counterpart in the reference.  Nothing above this line knows about it.
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr size_t NA_N = GUT_NOALLOC_MAX_PARTICLES, NA_T = GUT_NOALLOC_MAX_TILES;
// byte offsets of the sections (all multiples of 256)
constexpr size_t NA_OFF_COUNT = 0;                          // u32  [NA_N]      tiles touched per particle
constexpr size_t NA_OFF_POS = NA_OFF_COUNT + NA_N * 4;      // f32  [NA_N,2]
constexpr size_t NA_OFF_CONIC = NA_OFF_POS + NA_N * 8;      // f32  [NA_N,4]
constexpr size_t NA_OFF_EXTENT = NA_OFF_CONIC + NA_N * 16;  // f32  [NA_N,2]
constexpr size_t NA_OFF_DEPTH = NA_OFF_EXTENT + NA_N * 8;   // f32  [NA_N]
constexpr size_t NA_OFF_FEAT = NA_OFF_DEPTH + NA_N * 4;     // f32  [NA_N,4]
constexpr size_t NA_OFF_EMIT = NA_OFF_FEAT + NA_N * 16;     // u32  [NA_N,4]    (K1 writes it; unused here)
constexpr size_t NA_OFF_HIST = NA_OFF_EMIT + NA_N * 16;     // u32  [GUT_TILE_SUB][T] per-tile sub-counters: MUST START AT ZERO
constexpr size_t NA_BYTES = NA_OFF_HIST + NA_T * GUT_TILE_SUB * 4;
static_assert(NA_BYTES == GUT_NOALLOC_SCRATCH_BYTES, "header constant out of date");
static_assert(NA_OFF_HIST % 256 == 0 && NA_OFF_EMIT % 256 == 0, "sections stay 256-byte aligned");

// The scratch.  In HOST code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: host code gets
// the address from cudaGetSymbolAddress (noalloc_scratch_of below); device code names the array directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// Tile ranges from the sub-counters: one CTA scans the T per-tile totals (T <= 8192).  Reads the array by name.
constexpr int NA_SCAN_THREADS = 1024;
__global__ void __launch_bounds__(NA_SCAN_THREADS) gut_noalloc_ranges_kernel(uint32_t T, uint32_t* __restrict__ ranges) {
    const uint32_t* hist = reinterpret_cast<const uint32_t*>(g_noalloc_scratch + NA_OFF_HIST);
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s;
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < T; base += NA_SCAN_THREADS) {
        const uint32_t t = base + tid;
        uint32_t v = 0;
        if (t < T) {
#pragma unroll
            for (int k = 0; k < GUT_TILE_SUB; ++k) v += hist[GUT_SUB_SLOT(t, k, T)];
        }
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += n;
        }
        if (lane == 31) warp_sums[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t ws = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= (uint32_t)o) ws += n;
            }
            warp_sums[lane] = ws;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t first = carry + (wid ? warp_sums[wid - 1] : 0u) + incl - v;
        if (t < T) {   // (0,0) for an empty tile, as gut_forward's stage outputs have it
            ranges[2 * t] = v ? first : 0u;
            ranges[2 * t + 1] = v ? first + v : 0u;
        }
        __syncthreads();
        if (tid == NA_SCAN_THREADS - 1) carry_s = carry + warp_sums[31];
        __syncthreads();
    }
}

// Device address of the array on the CURRENT device, looked up once per device (every device that loads the module has
// its own copy of the array).
constexpr int NA_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[NA_MAX_DEVICES] = {};
std::mutex g_noalloc_mutex;

cudaError_t noalloc_scratch_of(int device, unsigned char** out) {
    if (device < 0 || device >= NA_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // the caller made `device` current
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" int gut_project_noalloc(GutHandle* h, void* stream_, const GutCamera* camera, uint32_t N, int32_t sph_degree,
                                   const float* d_density, const float* d_radiance, uint32_t W, uint32_t H, float* d_visibility,
                                   GutStageOutputs* so) {
    int rc = check_common(h, camera, N, sph_degree, W, H);
    if (rc != GUT_OK) return rc;
    if (N > GUT_NOALLOC_MAX_PARTICLES)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_project_noalloc: %u particles exceed GUT_NOALLOC_MAX_PARTICLES = %u (fixed scratch; use gut_forward)",
                    N, GUT_NOALLOC_MAX_PARTICLES);
    const uint64_t tiles = (uint64_t)((W + GUT_TILE - 1) / GUT_TILE) * (uint64_t)((H + GUT_TILE - 1) / GUT_TILE);
    if (tiles > GUT_NOALLOC_MAX_TILES)
        return fail(h, GUT_ERR_BAD_INPUT, "gut_project_noalloc: %ux%u has %llu tiles, more than GUT_NOALLOC_MAX_TILES = %u (fixed scratch; use gut_forward)",
                    W, H, (unsigned long long)tiles, GUT_NOALLOC_MAX_TILES);
    if (N > 0 && (!d_density || !d_radiance || !d_visibility)) return fail(h, GUT_ERR_BAD_INPUT, "null particle pointer");
    if (so && (so->d_sorted_keys || so->d_sorted_particle_idx))
        return fail(h, GUT_ERR_BAD_INPUT, "gut_project_noalloc does not sort: d_sorted_keys / d_sorted_particle_idx must be NULL");
    cudaStream_t s = (cudaStream_t)stream_;
    DeviceGuard guard(h->device);
    unsigned char* base = nullptr;
    {
        const cudaError_t e = noalloc_scratch_of(h->device, &base);
        if (e != cudaSuccess) return fail(h, GUT_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", h->device, cudaGetErrorString(e));
    }
    const GutFrame fr = make_frame(h->cfg, *camera, W, H);
    const uint32_t T = fr.num_tiles;
    GutProjBuffers pb;
    pb.tiles_count = (uint32_t*)(base + NA_OFF_COUNT); pb.proj_pos = (float2*)(base + NA_OFF_POS);
    pb.conic_opacity = (float4*)(base + NA_OFF_CONIC); pb.extent = (float2*)(base + NA_OFF_EXTENT);
    pb.depth = (float*)(base + NA_OFF_DEPTH); pb.features = (float4*)(base + NA_OFF_FEAT); pb.emit_rec = (uint4*)(base + NA_OFF_EMIT);
    uint32_t* hist = (uint32_t*)(base + NA_OFF_HIST);
    // the array keeps the last call's contents: the counters are the one section that must start at zero
    GUT_CUDA(h, cudaMemsetAsync(hist, 0, (size_t)T * GUT_TILE_SUB * 4, s));
    GUT_CUDA(h, gut_launch_project(h->cfg, *camera, fr, N, sph_degree, d_density, d_radiance, pb, d_visibility, hist, s));
    h->launches += N ? 1 : 0;
    if (so) {
        if (so->d_tiles_count) GUT_CUDA(h, cudaMemcpyAsync(so->d_tiles_count, pb.tiles_count, (size_t)N * 4, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_position) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_position, pb.proj_pos, (size_t)N * 8, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_conic_opacity) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_conic_opacity, pb.conic_opacity, (size_t)N * 16, cudaMemcpyDeviceToDevice, s));
        if (so->d_proj_extent) GUT_CUDA(h, cudaMemcpyAsync(so->d_proj_extent, pb.extent, (size_t)N * 8, cudaMemcpyDeviceToDevice, s));
        if (so->d_global_depth) GUT_CUDA(h, cudaMemcpyAsync(so->d_global_depth, pb.depth, (size_t)N * 4, cudaMemcpyDeviceToDevice, s));
        if (so->d_features) { GUT_CUDA(h, gut_launch_copy_features3(N, pb.features, (float*)so->d_features, s)); h->launches += N ? 1 : 0; }
        if (so->d_tile_ranges) {
            gut_noalloc_ranges_kernel<<<1, NA_SCAN_THREADS, 0, s>>>(T, (uint32_t*)so->d_tile_ranges);
            GUT_CUDA(h, cudaGetLastError());
            h->launches += 1;
        }
    }
    return GUT_OK;
}

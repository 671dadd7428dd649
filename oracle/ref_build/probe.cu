// probe.cu -- TEST INFRASTRUCTURE.  Runs the reference's own GUTRenderer (src/gutRenderer.cu, included textually and
// unchanged below; the only liberty taken is opening the class's access specifiers so that the forward context can
// be read) and hands its intermediate buffers to the parity tests through a small C ABI:
//   tiles_count / tiles_offset / projected position, conic+opacity, extent, depth, features   (K1, L1)
//   unsorted and sorted (tile<<32 | depth) keys and particle indices, tile ranges              (K2, L2, K3)
// plus the frame outputs and the backward, so that one library without torch or pybind11 covers the whole path.
// What this file adds of its own: building RenderParameters / GUTRenderer::Parameters exactly as
// src/splatRaster.cpp:153-169,194-223,286-312 does, and cudaMemcpy of the context buffers.  No arithmetic.
// Built by oracle/ref_build/Makefile into oracle/_ref/libgut_ref_stages_k0.so ("reference + shim").
#define private public
#define protected public
#include <3dgut/renderer/gutRenderer.h>
#undef private
#undef protected

#include GUT_REF_RENDERER_CU

#include <cstring>
#include <memory>

namespace {

struct ProbeState {
    threedgut::Logger logger;
    std::unique_ptr<threedgut::GUTRenderer> renderer;
    threedgut::GUTRenderer::Parameters parameters;
    uint32_t numIntersections = 0;
    uint32_t numParticles     = 0;
    uint32_t numTiles         = 0;
};
ProbeState* g_state = nullptr;

ProbeState& state() {
    if (!g_state) {
        g_state           = new ProbeState();
        g_state->renderer = std::make_unique<threedgut::GUTRenderer>(nlohmann::json::object(), g_state->logger);
        auto& p           = g_state->parameters;
        // src/splatRaster.cpp:160-167
        p.valuesBuffer.resize(sizeof(p.values), 0, g_state->logger);
        p.parametersBuffer.resize(sizeof(p.parameters), 0, g_state->logger);
        p.gradientsBuffer.resize(sizeof(p.gradients), 0, g_state->logger);
        p.parameters.dptrValuesBuffer = p.valuesBuffer.data();
        p.m_dptrParametersBuffer      = (uint64_t*)p.parametersBuffer.data();
        p.m_dptrGradientsBuffer       = (uint64_t*)p.gradientsBuffer.data();
    }
    return *g_state;
}

} // namespace

extern "C" {

// plain mirror of threedgut::CameraModelParameters + TSensorState (sensors/cameraModels.h:22-58, sensors.h:37-42)
struct GutRefCamera {
    int32_t model;   // 0 = OpenCV pinhole, 1 = OpenCV fisheye
    int32_t shutter; // CameraModelParameters::ShutterType order: T2B, L2R, B2T, R2L, global
    float principal_point[2];
    float focal_length[2];
    float radial[6];
    float tangential[2];
    float thin_prism[4];
    float max_angle;
    float start_pose[7]; // [t3, q xyzw]
    float end_pose[7];
    int64_t start_timestamp;
    int64_t end_timestamp;
};

static threedgut::RenderParameters makeRenderParameters(const GutRefCamera* cam, uint32_t width, uint32_t height) {
    threedgut::RenderParameters rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.id         = 0;
    rp.resolution = tcnn::ivec2{(int)width, (int)height};
    threedgut::CameraModelParameters sm;
    sm.shutterType = static_cast<threedgut::CameraModelParameters::ShutterType>(cam->shutter);
    if (cam->model == 0) { // bindings.cpp:34-58
        sm.modelType                         = threedgut::CameraModelParameters::OpenCVPinholeModel;
        sm.ocvPinholeParams.nearFar          = tcnn::vec2{0.01f, 100.0f};
        sm.ocvPinholeParams.principalPoint   = tcnn::vec2{cam->principal_point[0], cam->principal_point[1]};
        sm.ocvPinholeParams.focalLength      = tcnn::vec2{cam->focal_length[0], cam->focal_length[1]};
        for (int i = 0; i < 6; ++i) sm.ocvPinholeParams.radialCoeffs[i] = cam->radial[i];
        sm.ocvPinholeParams.tangentialCoeffs = tcnn::vec2{cam->tangential[0], cam->tangential[1]};
        sm.ocvPinholeParams.thinPrismCoeffs  = tcnn::vec4{cam->thin_prism[0], cam->thin_prism[1], cam->thin_prism[2], cam->thin_prism[3]};
    } else { // bindings.cpp:60-80
        sm.modelType                       = threedgut::CameraModelParameters::OpenCVFisheyeModel;
        sm.ocvFisheyeParams.principalPoint = tcnn::vec2{cam->principal_point[0], cam->principal_point[1]};
        sm.ocvFisheyeParams.focalLength    = tcnn::vec2{cam->focal_length[0], cam->focal_length[1]};
        sm.ocvFisheyeParams.radialCoeffs   = tcnn::vec4{cam->radial[0], cam->radial[1], cam->radial[2], cam->radial[3]};
        sm.ocvFisheyeParams.maxAngle       = cam->max_angle;
    }
    rp.sensorModel = sm;
    threedgut::TSensorPose sp, ep;
    for (int i = 0; i < 7; ++i) {
        sp[i] = cam->start_pose[i];
        ep[i] = cam->end_pose[i];
    }
    rp.sensorState = threedgut::TSensorState{cam->start_timestamp, sp, cam->end_timestamp, ep}; // splatRaster.cpp:98-106
    rp.objectAABB  = threedgut::BoundingBox{tcnn::vec3{-1e06f, -1e06f, -1e06f}, tcnn::vec3{1e06f, 1e06f, 1e06f}}; // :223
    return rp;
}

// Forward on the default stream.  All pointers are device pointers; rgba/dist/hits/visibility must be initialised by
// the caller the way SplatRaster::trace does (zeros, 1e6, zeros, zeros; splatRaster.cpp:196-199).
// Returns 0 and the number of particle/tile intersections in *num_intersections.
int gutref_forward(const GutRefCamera* cam, uint32_t num_particles, int sph_degree, const float* density12, const float* sph48,
                   uint32_t width, uint32_t height, const float* ray_ori, const float* ray_dir, float* rgba, float* dist,
                   float* hits, float* visibility, uint32_t* num_intersections) {
    ProbeState& s = state();
    cudaStream_t stream = 0;
    auto& p = s.parameters;
    p.values.numParticles               = num_particles;
    p.values.radianceSphDegree          = sph_degree;
    p.parameters.dptrDensityParameters  = (void*)density12;
    p.parameters.dptrRadianceParameters = (void*)sph48;
    p.valuesBuffer.setFromHost(&p.values, sizeof(p.values), 0, s.logger);
    p.parametersBuffer.setFromHost(&p.parameters, sizeof(p.parameters), 0, s.logger);
    const threedgut::RenderParameters rp = makeRenderParameters(cam, width, height);
    // the context keeps its sort buffers' previous contents when M shrinks; clear what the probe will read
    threedgut::Status st = s.renderer->renderForward(rp, (const tcnn::vec3*)ray_ori, (const tcnn::vec3*)ray_dir, hits, dist,
                                                     (tcnn::vec4*)rgba, (int*)visibility, p, 0, stream);
    if (cudaStreamSynchronize(stream) != cudaSuccess) return 2;
    s.numParticles = num_particles;
    s.numTiles     = ((width + 15) / 16) * ((height + 15) / 16);
    uint32_t m     = 0;
    if (num_particles) {
        cudaMemcpy(&m, (const uint32_t*)s.renderer->m_forwardContext->particlesTilesOffset.data() + num_particles - 1,
                   sizeof(uint32_t), cudaMemcpyDeviceToHost);
    }
    s.numIntersections = m;
    if (num_intersections) *num_intersections = m;
    return (ErrorCode)st == ErrorCode::None ? 0 : 1;
}

// Copies one buffer of the forward context to `dst` (device or host pointer, cudaMemcpyDefault).
// Names: tiles_count u32[N], tiles_offset u32[N], proj_pos f32[N,2], conic_opacity f32[N,4], extent f32[N,2],
// depth f32[N], features f32[N,3], unsorted_keys u64[M], unsorted_idx u32[M], sorted_keys u64[M], sorted_idx u32[M],
// ranges u32[T,2].  Returns the number of bytes copied, or -1.
long long gutref_read(const char* name, void* dst, unsigned long long capacity_bytes) {
    ProbeState& s = state();
    auto* c       = s.renderer->m_forwardContext.get();
    if (!c) return -1;
    const size_t N = s.numParticles, M = s.numIntersections, T = s.numTiles;
    const void* src = nullptr;
    size_t bytes    = 0;
    const std::string n(name);
    if (n == "tiles_count") { src = c->particlesTilesCount.data(); bytes = N * 4; }
    else if (n == "tiles_offset") { src = c->particlesTilesOffset.data(); bytes = N * 4; }
    else if (n == "proj_pos") { src = c->particlesProjectedPosition.data(); bytes = N * 8; }
    else if (n == "conic_opacity") { src = c->particlesProjectedConicOpacity.data(); bytes = N * 16; }
    else if (n == "extent") { src = c->particlesProjectedExtent.data(); bytes = N * 8; }
    else if (n == "depth") { src = c->particlesGlobalDepth.data(); bytes = N * 4; }
    else if (n == "features") { src = c->particlesPrecomputedFeatures.data(); bytes = N * 12; }
    else if (n == "unsorted_keys") { src = c->unsortedTileDepthKeys.data(); bytes = M * 8; }
    else if (n == "unsorted_idx") { src = c->unsortedTileParticleIdx.data(); bytes = M * 4; }
    else if (n == "sorted_keys") { src = c->sortedTileDepthKeys.data(); bytes = M * 8; }
    else if (n == "sorted_idx") { src = c->sortedTileParticleIdx.data(); bytes = M * 4; }
    else if (n == "ranges") { src = c->sortedTileRangeIndices.data(); bytes = M ? T * 8 : 0; }
    else return -1;
    if (bytes > capacity_bytes) return -1;
    if (bytes && cudaMemcpy(dst, src, bytes, cudaMemcpyDefault) != cudaSuccess) return -1;
    return (long long)bytes;
}

// Backward of the preceding gutref_forward (same inputs).  density_grad12 / sph_grad48 must be zero-initialised by the
// caller (splatRaster.cpp:273-274).
int gutref_backward(const GutRefCamera* cam, uint32_t num_particles, int sph_degree, const float* density12, const float* sph48,
                    uint32_t width, uint32_t height, const float* ray_ori, const float* ray_dir, const float* rgba,
                    const float* rgba_grad, const float* dist, const float* dist_grad, float* density_grad12, float* sph_grad48) {
    ProbeState& s = state();
    cudaStream_t stream = 0;
    auto& p = s.parameters;
    p.values.numParticles               = num_particles;
    p.values.radianceSphDegree          = sph_degree;
    p.parameters.dptrDensityParameters  = (void*)density12;
    p.parameters.dptrRadianceParameters = (void*)sph48;
    p.gradients.dptrDensityGradients    = density_grad12;
    p.gradients.dptrRadianceGradients   = sph_grad48;
    p.valuesBuffer.setFromHost(&p.values, sizeof(p.values), 0, s.logger);
    p.parametersBuffer.setFromHost(&p.parameters, sizeof(p.parameters), 0, s.logger);
    p.gradientsBuffer.setFromHost(&p.gradients, sizeof(p.gradients), 0, s.logger);
    const threedgut::RenderParameters rp = makeRenderParameters(cam, width, height);
    threedgut::Status st = s.renderer->renderBackward(rp, (const tcnn::vec3*)ray_ori, (const tcnn::vec3*)ray_dir, dist, dist_grad,
                                                      (const tcnn::vec4*)rgba, (const tcnn::vec4*)rgba_grad, nullptr, nullptr, p, 0,
                                                      stream);
    if (cudaStreamSynchronize(stream) != cudaSuccess) return 2;
    return (ErrorCode)st == ErrorCode::None ? 0 : 1;
}

int gutref_k_buffer_size(void) { return GAUSSIAN_K_BUFFER_SIZE; }

} // extern "C"

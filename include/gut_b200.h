/*
 * gut_b200.h -- C ABI of the B200-native 3DGUT rasteriser (libgut_b200.so).
 *
 * This is the drop-in boundary for the reference's native layer: every entry
 * point replaces one piece of the pybind11 module `lib3dgut_cc`
 * (reference: threedgut_tracer/bindings.cpp:79-113, class SplatRaster in
 * threedgut_tracer/include/3dgut/splatRaster.h:47-89).  Plain C types only:
 * raw device pointers, sizes and a cudaStream_t passed as void*.  No torch,
 * no pybind, no C++ types cross this line.
 *
 * Conventions
 *  - every function returns 0 on success or a GutStatus error code; the text
 *    of the last error of a handle is available from gut_last_error().
 *    (The reference logs and swallows errors, splatRaster.cpp:225-237; the
 *    host wrapper here raises instead.)
 *  - all pointers named d_* are DEVICE pointers valid on the handle's device,
 *    16-byte aligned, fp32 unless stated, borrowed for the duration of the
 *    call (the work is enqueued on `stream`; the caller keeps them alive
 *    until the stream reaches that point, as PyTorch's caching allocator
 *    does for tensors used on the current stream).
 *  - a handle is NOT thread safe, like the reference's SplatRaster
 *    (mutable m_parameters / m_forwardContext).
 *  - gut_backward consumes the forward context of the immediately preceding
 *    gut_forward on the same handle and stream
 *    (reference: gutRenderer.cu:413-417).
 */
#ifndef GUT_B200_H
#define GUT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GUT_ABI_VERSION 1

#if defined(__GNUC__)
#define GUT_API __attribute__((visibility("default")))
#else
#define GUT_API
#endif

typedef enum GutStatus {
    GUT_OK                = 0,
    GUT_ERR_BAD_INPUT     = 1, /* null pointer, bad size, unsupported config     */
    GUT_ERR_CUDA          = 2, /* a CUDA runtime call failed (see last_error)     */
    GUT_ERR_NO_FORWARD    = 3, /* backward without a matching forward context     */
    GUT_ERR_OUT_OF_MEMORY = 4,
    GUT_ERR_UNSUPPORTED   = 5
} GutStatus;

/* reference: CameraModelParameters::ShutterType, cameraModels.h:39-45 (same order) */
typedef enum GutShutterType {
    GUT_SHUTTER_ROLLING_TOP_TO_BOTTOM = 0,
    GUT_SHUTTER_ROLLING_LEFT_TO_RIGHT = 1,
    GUT_SHUTTER_ROLLING_BOTTOM_TO_TOP = 2,
    GUT_SHUTTER_ROLLING_RIGHT_TO_LEFT = 3,
    GUT_SHUTTER_GLOBAL                = 4
} GutShutterType;

/* reference: CameraModelParameters::ModelType, cameraModels.h:47-52 */
typedef enum GutCameraModel {
    GUT_CAMERA_OPENCV_PINHOLE = 0,
    GUT_CAMERA_OPENCV_FISHEYE = 1
} GutCameraModel;

/*
 * The compile-time -D defines of the reference (threedgut_tracer/setup_3dgut.py:47-70,
 * consumed in include/3dgut/threedgut.cuh:31-83) become one runtime struct;
 * kernels are template-specialised on the fields that shape register use
 * (kernel_degree, k_buffer_size) and read the rest as launch constants.
 * Defaults = configs/render/3dgut.yaml + configs/render/3dgrt.yaml.
 */
typedef struct GutConfig {
    int32_t abi_version;                 /* must be GUT_ABI_VERSION                              */
    int32_t kernel_degree;               /* particle_kernel_degree: 0,1,2,3,4,5,8  (default 2)    */
    float   min_kernel_response;         /* particle_kernel_min_response            (0.0113)      */
    float   min_alpha;                   /* particle_kernel_min_alpha               (1/255)       */
    float   max_alpha;                   /* particle_kernel_max_alpha               (0.99)        */
    float   min_transmittance;           /* min_transmittance                       (1e-4)        */
    int32_t sph_max_degree;              /* particle_radiance_sph_degree -> (d+1)^2 coeffs (3)    */
    int32_t enable_hitcounts;            /* enable_hitcounts                        (1)           */
    int32_t rect_bounding;               /* splat.rect_bounding                     (1)           */
    int32_t tight_opacity_bounding;      /* splat.tight_opacity_bounding            (1)           */
    int32_t tile_based_culling;          /* splat.tile_based_culling                (1)           */
    int32_t n_rolling_shutter_iterations;/* splat.n_rolling_shutter_iterations      (5)           */
    float   ut_alpha, ut_beta, ut_kappa; /* splat.ut_*                              (1, 2, 0)     */
    float   ut_in_image_margin_factor;   /* splat.ut_in_image_margin_factor         (0.1)         */
    int32_t ut_require_all_sigma_points_valid; /* must be 0 (threedgut.cuh:73)                    */
    int32_t k_buffer_size;               /* splat.k_buffer_size: 0 = unsorted       (0)           */
    int32_t global_z_order;              /* splat.global_z_order                    (1)           */
    int32_t enable_kernel_timings;       /* render.enable_kernel_timings            (0)           */
} GutConfig;

/* Fill *cfg with the reference defaults listed above. */
GUT_API void gut_default_config(GutConfig* cfg);

/*
 * Camera intrinsics + the two shutter poses of one view.
 * reference: CameraModelParameters (cameraModels.h:22-58) built by
 * fromOpenCV{Pinhole,Fisheye}CameraModelParameters (bindings.cpp:33-77) and
 * TSensorState (sensors.h:37-42).  Poses are WORLD->SENSOR, stored
 * [tx,ty,tz,qx,qy,qz,qw] (sensors.h:33).
 */
typedef struct GutCamera {
    int32_t model;               /* GutCameraModel                                   */
    int32_t shutter;             /* GutShutterType                                   */
    float   principal_point[2];
    float   focal_length[2];
    float   radial_coeffs[6];    /* pinhole k1..k6 ; fisheye uses [0..3]             */
    float   tangential_coeffs[2];/* pinhole only                                     */
    float   thin_prism_coeffs[4];/* pinhole only                                     */
    float   max_angle;           /* fisheye only                                     */
    float   start_pose[7];
    float   end_pose[7];
    int64_t start_timestamp_us;  /* carried for API fidelity; kernels never read it  */
    int64_t end_timestamp_us;
} GutCamera;

/*
 * Optional stage outputs for the bit-exact parity tests.  Any
 * pointer may be NULL.  Sizes: per-particle arrays N; sorted arrays
 * `capacity_intersections` entries (the call fails with GUT_ERR_BAD_INPUT if
 * the frame produced more).  All device pointers.
 */
typedef struct GutStageOutputs {
    uint32_t* d_tiles_count;        /* [N]   K1: tiles touched per particle          */
    float*    d_proj_position;      /* [N,2] K1                                      */
    float*    d_proj_conic_opacity; /* [N,4] K1                                      */
    float*    d_proj_extent;        /* [N,2] K1                                      */
    float*    d_global_depth;       /* [N]   K1                                      */
    float*    d_features;           /* [N,3] K1: SH->RGB + 0.5, unclamped            */
    uint64_t* d_sorted_keys;        /* [M]   (tile<<32)|depth bits, sorted           */
    uint32_t* d_sorted_particle_idx;/* [M]   particle index per sorted key           */
    uint32_t* d_tile_ranges;        /* [T,2] first / one-past-last sorted index      */
    uint64_t  capacity_intersections;
    uint64_t  num_intersections;    /* OUT: M of this frame                          */
} GutStageOutputs;

typedef struct GutHandle GutHandle;

/* reference: SplatRaster::SplatRaster(json) splatRaster.cpp:153-168 */
GUT_API int gut_create(const GutConfig* cfg, int device, GutHandle** out);
/* reference: SplatRaster::~SplatRaster + ~GutRenderForwardContext gutRenderer.cu:106-127 */
GUT_API void gut_destroy(GutHandle* h);

/*
 * Forward render of one view.  reference: SplatRaster::trace
 * (splatRaster.cpp:173-244) -> GUTRenderer::renderForward (gutRenderer.cu:241-398).
 *
 *   d_particle_density [N,12] = {pos3, density, quat wxyz, scale3, pad}   (gaussianParticles.cuh:20-26)
 *   d_particle_radiance[N,(sph_max_degree+1)^2,3] SH coefficients
 *   d_ray_origin, d_ray_direction [H,W,3] sensor-space rays
 * outputs (the function initialises them: rgba=0, dist=1e6, hits=0, visibility=0,
 * as torch::zeros/ones do at splatRaster.cpp:196-199):
 *   d_ray_radiance_density [H,W,4], d_ray_hit_distance [H,W], d_ray_hit_count [H,W],
 *   d_particle_visibility [N] (1.0f where the projected conic is valid; the
 *   reference stores int 1 in float storage)
 */
GUT_API int gut_forward(GutHandle* h, void* stream, const GutCamera* camera,
                uint32_t num_particles, int32_t sph_degree_active,
                const float* d_particle_density, const float* d_particle_radiance,
                uint32_t width, uint32_t height,
                const float* d_ray_origin, const float* d_ray_direction,
                float* d_ray_radiance_density, float* d_ray_hit_distance,
                float* d_ray_hit_count, float* d_particle_visibility,
                GutStageOutputs* stage_outputs /* may be NULL */);

/*
 * Backward of the immediately preceding gut_forward.  reference:
 * SplatRaster::traceBwd (splatRaster.cpp:246-332) -> GUTRenderer::renderBackward
 * (gutRenderer.cu:400-497).
 *   d_particle_density_grad  [N,12]  (column 11, the padding, stays 0)
 *   d_particle_radiance_grad [N,(sph_max_degree+1)^2,3]
 * accumulate == 0: the buffers are overwritten (reference semantics: fresh zeros);
 * accumulate != 0: gradients are ADDED to the buffers' current content (used to sum
 *   the views of one rank before the data-parallel allreduce).
 */
GUT_API int gut_backward(GutHandle* h, void* stream, const GutCamera* camera,
                 uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, const float* d_particle_radiance,
                 uint32_t width, uint32_t height,
                 const float* d_ray_origin, const float* d_ray_direction,
                 const float* d_ray_radiance_density, const float* d_ray_radiance_density_grad,
                 const float* d_ray_hit_distance, const float* d_ray_hit_distance_grad,
                 float* d_particle_density_grad, float* d_particle_radiance_grad,
                 int32_t accumulate);

/*
 * training != 0 (the default): gut_forward also records, per 8x8 pixel block and batch of 32 list entries, which entries
 * some ray accepted; gut_backward revisits only those (the reference re-tests the whole list).  training == 0 skips the
 * recording for inference (Tracer.render(train=False)); a gut_backward after such a forward fails with
 * GUT_ERR_NO_FORWARD.
 */
GUT_API int gut_set_training(GutHandle* h, int32_t training);

/*
 * Test control: enabled == 0 makes the k_buffer_size == 0 forward walk every list entry instead of dropping, per 8x8 block,
 * the entries whose response cannot reach the threshold for any ray of the block (a conservative cone bound).  Outputs are
 * bit-identical either way (tests/test_gpu_fullsize.py); gut_measure_work uses it to count the reference's walk.
 */
GUT_API int gut_set_block_cull(GutHandle* h, int32_t enabled);

/*
 * Drops the gradient state a multi-view step keeps inside the handle: zeroes the [N,16] accumulators and forgets a pending
 * gut_backward_colour sum.  Call it when a step is abandoned (an exception between gut_backward_colour and
 * gut_density_gradient_finish) or when the particle set changes mid-step (densification); gut_backward_colour and
 * gut_density_gradient_finish reject a particle count that differs from the pending accumulation's.
 */
GUT_API int gut_reset_gradients(GutHandle* h, void* stream);

/*
 * Deferred gradients for multi-view steps (this build's extension).
 * The backward of one view leaves two things per particle: 12 ray-summed canonical-frame accumulators that are linear
 * and VIEW-INDEPENDENT (they become the [N,12] gradient through the particle's own rotation and scale only), and
 * dL/dcolour, which the reference expands to dL/dSH = basis(view direction) (x) dL/dcolour
 * (gaussianParticles.cuh:120-187): 3 numbers per particle and view blown up to 3 x 16.  So a step of V views can
 * defer both conversions to its end:
 *   gut_backward_colour          per view: compositing backward; writes the view's masked colour gradient
 *                                d_colour_grad[N,3] (zero for particles the view did not touch) and the world-space
 *                                view centre it used (host float[3], may be NULL); the density-part accumulators keep
 *                                summing inside the handle;
 *   gut_density_gradient_finish  once: d_particle_density_grad[N,12] (+)= conversion of the summed accumulators
 *                                (which are reset);
 *   gut_sh_gradient_expand       once: d_particle_radiance_grad[N,48] (+)= sum_v basis(normalize(position - centre_v))
 *                                (x) colour_grad_v over all views (d_colour_grads is [n_views, N, 3]; d_view_centres
 *                                is a DEVICE array [n_views, 3], so the centres of other ranks' views can arrive by
 *                                the same all-gather as their colour gradients).
 * One step therefore writes the 48-byte and 192-byte gradient rows once instead of read-modify-writing them V times,
 * and data-parallel ranks exchange 12 B per particle and view instead of 192 B per particle.
 * gut_backward (the reference-shaped call) must not be interleaved between gut_backward_colour and the finish.
 */
GUT_API int gut_backward_colour(GutHandle* h, void* stream, const GutCamera* camera,
                 uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, const float* d_particle_radiance,
                 uint32_t width, uint32_t height,
                 const float* d_ray_origin, const float* d_ray_direction,
                 const float* d_ray_radiance_density, const float* d_ray_radiance_density_grad,
                 const float* d_ray_hit_distance, const float* d_ray_hit_distance_grad,
                 float* d_colour_grad, float* view_centre_out);
/* A step in which this handle renders ONE view: the view's [N,12] gradient is written at once (as gut_backward does, into
 * d_particle_density_grad, added when accumulate != 0) together with its colour gradient -- no accumulation is left pending
 * and no gut_density_gradient_finish pass is needed; only the SH expansion waits for the other ranks' views. */
GUT_API int gut_backward_colour_density(GutHandle* h, void* stream, const GutCamera* camera,
                 uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, const float* d_particle_radiance,
                 uint32_t width, uint32_t height,
                 const float* d_ray_origin, const float* d_ray_direction,
                 const float* d_ray_radiance_density, const float* d_ray_radiance_density_grad,
                 const float* d_ray_hit_distance, const float* d_ray_hit_distance_grad,
                 float* d_particle_density_grad, float* d_colour_grad, float* view_centre_out, int32_t accumulate);
GUT_API int gut_density_gradient_finish(GutHandle* h, void* stream, uint32_t num_particles,
                 const float* d_particle_density, float* d_particle_density_grad, int32_t accumulate);
GUT_API int gut_sh_gradient_expand(GutHandle* h, void* stream, uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, uint32_t n_views, const float* d_view_centres,
                 const float* d_colour_grads, float* d_particle_radiance_grad, int32_t accumulate);

/*
 * The same expansion with one pointer per view (a HOST array of n_views device pointers, each [N,3]), so that a view's colour
 * gradient may live in a buffer of a PEER GPU: gut_peer_buffer_create makes a device allocation that other processes of the
 * node map with gut_peer_buffer_open (CUDA IPC over NVLink / NVSwitch), and the expansion kernel then reads the other ranks'
 * per-view colour gradients directly while it computes -- the all-gather of the data-parallel step becomes loads inside
 * the kernel that consumes it.  The caller orders the step across ranks (every rank's colour gradients are complete
 * before any rank expands; 3dgrut_b200/dist.py does it with the small all-gather of the view centres).
 * No reference counterpart (the reference is single-GPU); n_views == 0 writes zeros.
 */
#define GUT_EXPAND_MAX_VIEWS 16      /* views per launch; more are expanded in several launches */
#define GUT_IPC_HANDLE_BYTES 64
GUT_API int gut_sh_gradient_expand_gather(GutHandle* h, void* stream, uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, uint32_t n_views, const float* d_view_centres,
                 const float* const* colour_grad_ptrs, float* d_particle_radiance_grad, int32_t accumulate);
/*
 * Layout of the per-view colour-gradient slots (d_colour_grad of gut_backward_colour / gut_backward_colour_density, the
 * entries of colour_grad_ptrs of gut_sh_gradient_expand_gather / gut_sh_expand_adam).  No reference counterpart.
 *   GUT_COLOUR_DENSE    [N,3] floats, zero rows for untouched particles (the default);
 *   GUT_COLOUR_COMPACT  only the rows the view's rays pushed on (about a sixth of the particles at the bench workload), 12 B
 *                       each, behind a table of one (mask, first row) pair per 32 particles: float 0 = rows in use, floats
 *                       4 .. 4 + table the table, rows behind it.  A slot holds gut_colour_slot_floats(N, layout) floats and is
 *                       16-byte aligned.  The expansion on another GPU then pulls ~3 B per particle and view over NVLink
 *                       instead of 12, through two dependent loads instead of one (measured equal to the dense layout at 1 and
 *                       2 GPUs; an option for steps bound by NVLink ingress).  Writers and readers of a slot must agree: every handle that takes part in a step is
 *                       given the same layout (3dgrut_b200/dist.py, PeerExchange(compact=True)); gut_sh_gradient_expand (one
 *                       dense buffer for all views) refuses a handle in compact mode.
 * The layout cannot change while a step is pending (between gut_backward_colour and gut_density_gradient_finish).
 */
#define GUT_COLOUR_DENSE 0
#define GUT_COLOUR_COMPACT 1
GUT_API int gut_set_colour_layout(GutHandle* h, int32_t layout);
GUT_API uint64_t gut_colour_slot_floats(uint32_t num_particles, int32_t layout);
GUT_API int gut_peer_buffer_create(GutHandle* h, size_t bytes, void** d_ptr_out, unsigned char* ipc_handle_out /* [64] */);
GUT_API int gut_peer_buffer_destroy(GutHandle* h, void* d_ptr);
GUT_API int gut_peer_buffer_open(GutHandle* h, const unsigned char* ipc_handle /* [64] */, void** d_ptr_out);
GUT_API int gut_peer_buffer_close(GutHandle* h, void* d_ptr);

/*
 * Packing of the model's parameter tensors into the [N,12] particle_density record and the inverse for its gradient.
 * Replaces: torch.cat([positions, density, rotation, scale, zeros], dim=1) in threedgut_tracer/tracer.py:176-178 and the
 * column split autograd derives from it for the backward (tracer.py:265-286 returns the [N,12] gradient, autograd slices it).
 *   d_positions [N,3], d_density [N,1], d_rotation [N,4] (w,x,y,z; 16-byte aligned), d_scale [N,3]  <->  [N,12]
 */
GUT_API int gut_pack_particles(GutHandle* h, void* stream, uint32_t num_particles, const float* d_positions, const float* d_density,
                 const float* d_rotation, const float* d_scale, float* d_particle_density);
GUT_API int gut_unpack_particle_gradient(GutHandle* h, void* stream, uint32_t num_particles, const float* d_particle_density_grad,
                 float* d_positions_grad, float* d_density_grad, float* d_rotation_grad, float* d_scale_grad);

/*
 * Activation + packing fused at the Tracer boundary: raw model parameters in, packed records out.
 * Replaces, in one pass each way: rotation_activation = torch.nn.functional.normalize, scale_activation (exp),
 * density_activation (sigmoid) of threedgrut/model/model.py:77-93 (table threedgrut/utils/misc.py:48-53), the
 * torch.cat of threedgut_tracer/tracer.py:176-178, get_features() = torch.cat((albedo, specular), dim=1) of
 * model.py:74-75, and the autograd mirror image of all of them.
 *   d_positions [N,3], d_rotation_raw [N,4] (16-byte aligned), d_scale_raw [N,3], d_density_raw [N,1]  ->  [N,12]
 *   d_features_albedo [N,3], d_features_specular [N,45]  <->  [N,48]
 * gut_activate_pack_backward takes the packed record the forward produced (activated scale / density are read from
 * it) and the [N,12] gradient gut_backward wrote; it writes the gradients of the RAW parameters.
 */
enum GutActivation { GUT_ACT_NONE = 0, GUT_ACT_EXP = 1, GUT_ACT_SIGMOID = 2 };
GUT_API int gut_activate_pack(GutHandle* h, void* stream, uint32_t num_particles, const float* d_positions, const float* d_rotation_raw,
                 const float* d_scale_raw, const float* d_density_raw, int32_t scale_activation, int32_t density_activation,
                 float* d_particle_density);
GUT_API int gut_activate_pack_backward(GutHandle* h, void* stream, uint32_t num_particles, const float* d_rotation_raw,
                 const float* d_particle_density, const float* d_particle_density_grad, int32_t scale_activation,
                 int32_t density_activation, float* d_positions_grad, float* d_rotation_grad, float* d_scale_grad,
                 float* d_density_grad);
GUT_API int gut_pack_features(GutHandle* h, void* stream, uint32_t num_particles, const float* d_features_albedo,
                 const float* d_features_specular, float* d_particle_radiance);
GUT_API int gut_unpack_feature_gradient(GutHandle* h, void* stream, uint32_t num_particles, const float* d_particle_radiance_grad,
                 float* d_features_albedo_grad, float* d_features_specular_grad);

/*
 * The optimiser step after the gradient exchange.
 * gut_selective_adam_step replaces SelectiveAdam.step (threedgrut/optimizers/__init__.py:88-133) and its kernel
 * selective_adam_update_kernel (threedgrut/optimizers/optimizers.cu:47-84), for every parameter group in ONE launch:
 *     m = b1 m + (1 - b1) g ;  v = b2 v + (1 - b2) g g ;  param += -lr m / (sqrt(v) + eps)        (no bias correction)
 * applied to the rows whose d_visibility entry is non-zero (the tracer's mog_visibility [N,1]; NULL = every row).
 * Invisible rows keep param, exp_avg and exp_avg_sq.  zero_grad != 0 also clears every gradient (optimizer.zero_grad(),
 * threedgrut/trainer.py:751).  Each group is a contiguous fp32 [N, width] tensor with its own learning rate.
 * gut_densify_stats replaces GSStrategy.update_gradient_buffer (threedgrut/strategy/gs.py:104-115): for particles whose
 * position gradient is non-zero, accum += |grad * |position - sensor_position|| / 2 and denom += 1.
 */
#define GUT_ADAM_MAX_GROUPS 8
typedef struct GutAdamGroup {
    float*       param;       /* [N, width] */
    const float* grad;        /* [N, width]; written (zeroed) when zero_grad != 0 */
    float*       exp_avg;     /* [N, width] */
    float*       exp_avg_sq;  /* [N, width] */
    uint32_t     width;
    float        lr;
} GutAdamGroup;
GUT_API int gut_selective_adam_step(GutHandle* h, void* stream, uint32_t num_particles, const GutAdamGroup* groups, uint32_t num_groups,
                 const float* d_visibility, float beta1, float beta2, float eps, int32_t zero_grad);
GUT_API int gut_densify_stats(GutHandle* h, void* stream, uint32_t num_particles, const float* d_positions, const float* d_positions_grad,
                 const float* sensor_position /* host [3] */, float* d_grad_norm_accum, float* d_grad_norm_denom);

/*
 * The optimiser step as the EPILOGUE of a step's gradient kernels (no reference counterpart as a fusion:
 * the reference runs autograd's activation adjoints, the feature split and SelectiveAdam as separate kernels per group).
 *   gut_density_adam    the summed [N,12] gradient (after the data-parallel all-reduce) -> activation adjoints -> SelectiveAdam
 *                       on positions / rotation / scale / density in one pass, plus -- when sensor_position and both
 *                       statistics buffers are given -- GSStrategy.update_gradient_buffer on the position gradient and the
 *                       positions BEFORE the update (gs.py:104-115 runs in post_backward, before optimizer.step);
 *   gut_sh_expand_adam  the per-view colour gradients of ALL views of the step (one device pointer per view, local or
 *                       peer memory, as gut_sh_gradient_expand_gather; n_views <= GUT_EXPAND_MAX_VIEWS) are expanded per
 *                       particle into the 48 SH gradients in registers and applied to features_albedo [N,3] /
 *                       features_specular [N,45] at once: the [N,48] gradient (192 B per particle written, read by the
 *                       feature split, written again, read by Adam) never exists.
 * Both are masked by d_visibility [N] (NULL = all rows) like SelectiveAdam.step, and leave bit for bit what
 * gut_activate_pack_backward / gut_sh_gradient_expand_gather + gut_unpack_feature_gradient -> gut_selective_adam_step ->
 * gut_densify_stats leave on the same inputs.
 */
typedef struct GutAdamTensors {
    float* param;        /* raw parameter, updated in place */
    float* exp_avg;
    float* exp_avg_sq;
    float  lr;
} GutAdamTensors;
GUT_API int gut_density_adam(GutHandle* h, void* stream, uint32_t num_particles, const float* d_particle_density,
                 const float* d_particle_density_grad, const GutAdamTensors* positions, const GutAdamTensors* rotation,
                 const GutAdamTensors* scale, const GutAdamTensors* density, int32_t scale_activation, int32_t density_activation,
                 const float* d_visibility, float beta1, float beta2, float eps,
                 const float* sensor_position /* host [3] or NULL */, float* d_grad_norm_accum, float* d_grad_norm_denom);
GUT_API int gut_sh_expand_adam(GutHandle* h, void* stream, uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, uint32_t n_views, const float* d_view_centres,
                 const float* const* colour_grad_ptrs, const GutAdamTensors* features_albedo,
                 const GutAdamTensors* features_specular, const float* d_visibility, float beta1, float beta2, float eps);

/*
 * Same two calls with HOST buffers: inputs are copied host->device and results
 * device->host inside the call on `stream` (the end-to-end path bench.py times).
 * Particle parameters stay device resident (they are the model, not a per-view
 * input).  Pointers named h_* are host pointers, ideally pinned.
 */
GUT_API int gut_forward_host(GutHandle* h, void* stream, const GutCamera* camera,
                     uint32_t num_particles, int32_t sph_degree_active,
                     const float* d_particle_density, const float* d_particle_radiance,
                     uint32_t width, uint32_t height,
                     const float* h_ray_origin, const float* h_ray_direction,
                     float* h_ray_radiance_density, float* h_ray_hit_distance,
                     float* h_ray_hit_count, float* d_particle_visibility);

GUT_API int gut_backward_host(GutHandle* h, void* stream, const GutCamera* camera,
                      uint32_t num_particles, int32_t sph_degree_active,
                      const float* d_particle_density, const float* d_particle_radiance,
                      uint32_t width, uint32_t height,
                      const float* h_ray_radiance_density_grad, const float* h_ray_hit_distance_grad,
                      float* d_particle_density_grad, float* d_particle_radiance_grad,
                      int32_t accumulate);

/* reference: SplatRaster::collectTimes splatRaster.cpp:334-364 -- mean ms of the
 * forward/backward calls since the last collection; -1 when none was timed. */
GUT_API int gut_timings(GutHandle* h, float* forward_render_ms, float* backward_render_ms);

/*
 * Per-stage mean device times (ms) since the last call, in the order of
 * gut_stage_name(i), i < gut_num_stages().  Stage names follow the reference's
 * DeviceLaunchesLogger tags (gutRenderer.cu:256-378).
 */
GUT_API int         gut_num_stages(void);
GUT_API const char* gut_stage_name(int i);
GUT_API int         gut_stage_timings(GutHandle* h, float* ms_out /* [gut_num_stages()] */);

/* Statistics of the last forward (what defines the work). */
typedef struct GutFrameStats {
    uint64_t num_intersections;     /* M                                    */
    uint32_t num_tiles;             /* T                                    */
    uint32_t max_tile_list_length;
    uint64_t kernel_launches;       /* kernels launched by this handle so far */
} GutFrameStats;
GUT_API int gut_frame_stats(GutHandle* h, GutFrameStats* out);

/*
 * Instrumented re-run of the compositing walk of the last forward (same rays): returns the number of
 * ray x particle tests and of accepted hits.  These are the units the FP32 roofline of K4/K5 is
 * expressed in; not part of the render path, never called inside a timed region.
 */
GUT_API int gut_measure_work(GutHandle* h, void* stream, const float* d_ray_origin, const float* d_ray_direction,
                             uint64_t* num_tests, uint64_t* num_hits);

/*
 * gut_project_noalloc -- NO COUNTERPART IN THE REFERENCE; added for callers that
 * cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * The projection stage alone (K1 of gut_forward: unscented projection, conic / extent, SH colour, per-tile culling and
 * counting) for a small particle set, with ALL of its scratch -- the per-particle projected records and the per-tile
 * sub-counters -- in one file-scope `static __device__` array of the library (GUT_NOALLOC_SCRATCH_BYTES = 5 MiB, held by
 * every GPU that loads the library until the process exits).  The call allocates nothing: no cudaMalloc / cudaMallocAsync /
 * cudaFree, no allocating library routine; it clears the tile counters, launches the kernels and copies the requested
 * results into the caller's buffers on `stream`, without synchronising.
 *   inputs : as gut_forward (no rays).  num_particles <= GUT_NOALLOC_MAX_PARTICLES and
 *            ceil(width/16) * ceil(height/16) <= GUT_NOALLOC_MAX_TILES, else GUT_ERR_BAD_INPUT with a message naming the limit.
 *   outputs: d_particle_visibility [N] (required); the K1 fields of `outputs` (d_tiles_count, d_proj_position,
 *            d_proj_conic_opacity, d_proj_extent, d_global_depth, d_features; any may be NULL) with the same values, bit
 *            for bit, as gut_forward's stage outputs; d_tile_ranges [T,2] = first / one-past-last slot of every tile's list
 *            in the sorted order gut_forward would build ((0,0) for empty tiles).  d_sorted_keys / d_sorted_particle_idx
 *            are not produced (they must be NULL) and num_intersections is not written (no device->host read-back).
 * Every call on a device shares the one array, which keeps the previous call's contents (the call clears what must start
 * at zero): two calls that overlap on the GPU -- different streams, or different handles on one device -- RACE.
 * The forward context of `h` is not touched: a gut_backward still refers to the last gut_forward.
 */
#define GUT_NOALLOC_MAX_PARTICLES 65536u
#define GUT_NOALLOC_MAX_TILES     8192u       /* 1920 x 1080 has 8160 tiles of 16 x 16 */
#define GUT_NOALLOC_SCRATCH_BYTES (GUT_NOALLOC_MAX_PARTICLES * 72u + GUT_NOALLOC_MAX_TILES * 64u)
GUT_API int gut_project_noalloc(GutHandle* h, void* stream, const GutCamera* camera,
                 uint32_t num_particles, int32_t sph_degree_active,
                 const float* d_particle_density, const float* d_particle_radiance,
                 uint32_t width, uint32_t height,
                 float* d_particle_visibility, GutStageOutputs* outputs);

GUT_API const char* gut_last_error(GutHandle* h);
GUT_API int         gut_abi_version(void);
/* struct sizes as compiled into the library, so a foreign-language binding can verify its mirror of the structs */
GUT_API size_t      gut_sizeof_config(void);
GUT_API size_t      gut_sizeof_camera(void);
GUT_API size_t      gut_sizeof_stage_outputs(void);
GUT_API size_t      gut_sizeof_frame_stats(void);
GUT_API size_t      gut_sizeof_adam_group(void);
GUT_API size_t      gut_sizeof_adam_tensors(void);

#ifdef __cplusplus
}
#endif
#endif /* GUT_B200_H */

/* fewtensorf.h — C ABI of the B200-native Few-TensoRF per-ray hot path.
 *
 * One shared library (fewtensorf_b200/libfewtensorf_b200.so), plain pointers and sizes only, no
 * torch / ATen types.  Every pointer is a DEVICE pointer unless its comment says "host".
 * All functions enqueue on the caller's stream and never synchronise, never throw, never exit();
 * they return 0 on success or a negative FTRF_E_* code, with text from ftrf_last_error().
 *
 * What each entry point replaces in the reference (paths relative to the reference root;
 * the reference has no FFI of its own, these are the torch call sites a maintainer re-points):
 *   ftrf_forward            models/tensorBase.py:321-395  TensorBase.forward (whole call, no autograd)
 *   ftrf_backward           train.py:374-376              autograd replay of that forward
 *   ftrf_stage_march        models/tensorBase.py:178-208, :349-354, :41-48   sample_ray(+ndc) + alphaMask cull
 *   ftrf_stage_density      models/tensoRF.py:207-227, :358-386; tensorBase.py:291-295, :361-365
 *   ftrf_stage_integrate    models/tensorBase.py:21-28, :346, :368-370, :377, :386-390
 *   ftrf_stage_appearance   models/tensoRF.py:230-263, :388-415  (+ basis_mat)
 *   ftrf_stage_shade        models/mlp.py:8-19, :41-69, :88-107, :127-155; models/sh.py:87-112
 *   ftrf_stage_composite    models/tensorBase.py:375-384
 *   ftrf_mse_loss           train.py:338
 *   ftrf_pack_alpha_mask    models/tensorBase.py:30-39   (AlphaGridMask volume -> bit field)
 *   ftrf_build_coarse_mask  (none: empty-space acceleration of the alphaMask cull, same decisions)
 *   ftrf_density_at         models/tensorBase.py:298-318 compute_alpha's sigma (dense-lattice eval)
 *   ftrf_dense_alpha_bits, ftrf_dilate_alpha_mask, ftrf_unpack_alpha_mask
 *                           models/tensorBase.py:215-256 getDenseAlpha / updateAlphaMask
 *   ftrf_filter_rays        models/tensorBase.py:259-288 filtering_rays
 *   ftrf_reg_factors        models/tensoRF.py:189-205 + loss.py:120-141   TV / L1 regularisers (value + gradient)
 *   ftrf_ortho_reg          models/tensoRF.py:175-187   vector_comp_diffs (value + gradient)
 *   ftrf_resample_factor, ftrf_crop_factor   models/tensoRF.py:267-327, :418-466   upsample_volume_grid / shrink
 *   ftrf_density_feature, ftrf_app_feature, ftrf_shade_points
 *                           models/tensoRF.py:207-263, :358-415; models/mlp.py:15-19, :41-155   the module methods on points
 *   ftrf_adam_step          train.py:272-273, :374-376   torch.optim.Adam over all parameter groups
 *   ftrf_generate_rays      dataLoader/ray_utils.py:24-85, dataLoader/blender.py:59   camera rays from a pose
 *   ftrf_zero_after (+ ftrf_event_record, ftrf_stream_wait_event)   train.py:374 optimizer.zero_grad()
 *
 * Factor layout: CHANNEL-LAST.  A VM plane the reference holds as (1,C,H,W) is read as [H][W][C]
 * (torch.channels_last memory of the same Parameter), a line (1,C,L,1) as [L][C].  C % 4 == 0.
 */
#ifndef FEWTENSORF_H
#define FEWTENSORF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FTRF_ABI_VERSION 2

enum { FTRF_MODEL_VM = 0, FTRF_MODEL_CP = 1 };
enum { FTRF_SHADE_MLP_FEA = 0, FTRF_SHADE_MLP_PE = 1, FTRF_SHADE_MLP = 2, FTRF_SHADE_SH = 3 };
enum { FTRF_ACT_SOFTPLUS = 0, FTRF_ACT_RELU = 1 };

enum {
  FTRF_OK = 0,
  FTRF_E_ARG = -1,        /* bad argument / unsupported shape            */
  FTRF_E_WORKSPACE = -2,  /* workspace buffer smaller than the layout    */
  FTRF_E_CUDA = -3        /* a CUDA runtime call failed                  */
};

/* One decomposition (density or appearance): three planes (VM only) and three lines. */
typedef struct FtrfFactorSet {
  int32_t n_comp[3];        /* channels per plane/line k (CP: all equal)                    */
  int32_t plane_w[3];       /* VM: gridSize[matMode[k][0]]                                   */
  int32_t plane_h[3];       /* VM: gridSize[matMode[k][1]]                                   */
  int32_t line_len[3];      /* gridSize[vecMode[k]]                                          */
  const float* plane[3];    /* [H][W][C]   (NULL for CP)                                     */
  const float* line[3];     /* [L][C]                                                        */
  float* g_plane[3];        /* gradient accumulators, same layout (backward only, else NULL) */
  float* g_line[3];
  const float* mask[3];     /* FreeNeRF decomposition mask per k; NULL = none. CP reads [0]   */
  int32_t mask_stride[3];   /* 0 = one scalar broadcast over channels, 1 = per-channel vector */
} FtrfFactorSet;

/* The field: hyper-parameters + borrowed parameter pointers.  Re-filled by the host every call
 * (parameters are owned by the Python module and may be replaced between calls). */
typedef struct FtrfField {
  int32_t model;            /* FTRF_MODEL_*                                                  */
  int32_t shading;          /* FTRF_SHADE_*                                                  */
  int32_t act;              /* FTRF_ACT_*   (fea2denseAct)                                   */
  int32_t app_dim;          /* data_dim_color (27)                                           */
  int32_t feature_c;        /* MLP width (128)                                               */
  int32_t pos_pe, view_pe, fea_pe;
  int32_t in_dim;           /* MLP input width (mlp.py:31/:75/:113)                          */
  float aabb[6];            /* min xyz, max xyz — fp32 values of the torch tensor            */
  float inv_aabb[3];        /* 2/(max-min) as torch computed it (tensorBase.py:106)          */
  float step_size;          /* tensorBase.py:109, the fp32 torch scalar                      */
  float near_, far_;
  float distance_scale, density_shift, weight_thres;
  /* alphaMask (tensorBase.py:30-48); dims are X,Y,Z = volume.shape[-1],[-2],[-3] */
  int32_t has_alpha_mask;
  int32_t am_dim[3];
  float am_aabb0[3];
  float am_inv[3];          /* AlphaGridMask.invgridSize                                     */
  const uint32_t* am_bits;  /* ftrf_pack_alpha_mask output                                   */
  const uint32_t* am_coarse;/* ftrf_build_coarse_mask output (8^3-block occupancy) or NULL   */
  FtrfFactorSet den, app;
  const float* basis;       /* basis_mat.weight [app_dim][sum n_comp]                        */
  float* g_basis;
  const float* w1; const float* b1;   /* renderModule.mlp.0  [feature_c][in_dim], [feature_c] */
  const float* w2; const float* b2;   /* renderModule.mlp.2  [feature_c][feature_c]           */
  const float* w3; const float* b3;   /* renderModule.mlp.4  [3][feature_c], [3]              */
  float* g_w1; float* g_b1; float* g_w2; float* g_b2; float* g_w3; float* g_b3;
  const float* enc_mask[3]; /* FreeNeRF encoding masks: pos, view, fea; NULL = none          */
  int32_t enc_mask_len[3];  /* 1 = scalar broadcast, else 2*pe*channels                      */
} FtrfField;

/* One call on a ray chunk. */
typedef struct FtrfCall {
  int64_t n_rays;
  int32_t ray_cols;         /* columns per ray row (>= 6); depth uses the LAST column (tensorBase.py:388) */
  int32_t n_samples;        /* S                                                            */
  int32_t is_train;         /* jitter present                                               */
  int32_t ndc;              /* sample_ray_ndc path                                          */
  int32_t add_bg;           /* white_bg or (is_train and coin), resolved by the host (:380) */
  int32_t save_for_backward;/* keep activations in the workspace                            */
  const float* rays;        /* [n_rays][ray_cols]                                           */
  const float* jitter;      /* non-ndc train: [n_rays] U[0,1) (:201).  ndc: z table [S] (:181-183), always */
  float* rgb_map;           /* out [n_rays][3]                                              */
  float* depth_map;         /* out [n_rays]                                                 */
  const float* g_rgb_map;   /* backward in: dL/d rgb_map [n_rays][3]                        */
  int64_t cap_samples;      /* capacity of the packed valid-sample list                     */
  int64_t cap_app;          /* capacity of the packed shaded-sample list                    */
  /* Library-owned hand-off between the calls that share this struct (zero-initialise, then leave alone):
   * which weight images the forward already prepared for the backward, and which side-stream work a split
   * backward (ftrf_backward_part 1 then 2, possibly from different host threads) still has to join.  This is
   * what keeps the entry points re-entrant: the library holds no process-wide state about a call. */
  uint64_t lib_flags;
  void* lib_ctx;
  /* backward, optional (NULL = off): device float that receives 1.0 if this call's packed lists overflowed, else 0.0.
   * A multi-GPU step points it at the tail slot of the gradient buffer, so the flag travels with the all-reduce and
   * every rank learns that some rank has to repeat the step (fewtensorf_b200/module.py: train_step). */
  float* overflow_out;
} FtrfCall;

/* Workspace sections (offsets in bytes from the workspace base, filled by ftrf_workspace_layout). */
enum {
  FTRF_WS_COUNTERS = 0, /* int32[8]: [0] valid samples required, [1] shaded samples required,
                                     [2] overflow flags (bit0 samples, bit1 shaded)          */
  FTRF_WS_TMIN,         /* float[n_rays]                                                     */
  FTRF_WS_RAY_OFF,      /* int32[n_rays]  first packed sample of the ray                     */
  FTRF_WS_RAY_CNT,      /* int32[n_rays]                                                     */
  FTRF_WS_ACC,          /* float[n_rays]  acc_map                                            */
  FTRF_WS_RGB_PRE,      /* float[n_rays][3] rgb_map before clamp                             */
  FTRF_WS_S_RAY,        /* int32[cap_samples] ray id                                         */
  FTRF_WS_S_IDX,        /* int32[cap_samples] sample index i along the ray                   */
  FTRF_WS_S_SIGMA,      /* float[cap_samples]                                                */
  FTRF_WS_S_W,          /* float[cap_samples] weight                                         */
  FTRF_WS_S_T,          /* float[cap_samples] transmittance before the sample                */
  FTRF_WS_S_APP,        /* int32[cap_samples] shaded-list slot or -1                         */
  FTRF_WS_S_GSIG,       /* float[cap_samples] dL/d sigma_feature (backward)                  */
  FTRF_WS_A_SLOT,       /* int32[cap_app] packed-sample slot of the shaded sample            */
  FTRF_WS_A_FEAT,       /* float[cap_app][app_dim]                                           */
  FTRF_WS_A_RGB,        /* float[cap_app][3]                                                 */
  FTRF_WS_A_H1,         /* float[cap_app][feature_c]  (save_for_backward)                    */
  FTRF_WS_A_H2,         /* float[cap_app][feature_c]                                         */
  FTRF_WS_A_GRGB,       /* float[cap_app][3]          backward                               */
  FTRF_WS_A_GH1,        /* float[cap_app][feature_c]  backward                               */
  FTRF_WS_A_GH2,        /* float[cap_app][feature_c]  backward                               */
  FTRF_WS_A_GFEAT,      /* float[cap_app][app_dim]    backward                               */
  FTRF_WS_A_X,          /* float[cap_app][ldx]  MLP input rows, ldx = in_dim + 1 rounded up to 8  */
  FTRF_WS_A_GX,         /* float[cap_app][ldx]  backward                                         */
  FTRF_WS_WIMG,         /* pre-split tf32 hi/lo weight images streamed by the fused MLP kernels      */
  FTRF_WS_COUNT
};

int ftrf_abi_version(void);
const char* ftrf_last_error(void);          /* host string, thread-local */

/* Returns total bytes; offsets_out (host, may be NULL) receives FTRF_WS_COUNT byte offsets. */
size_t ftrf_workspace_layout(const FtrfField* field, const FtrfCall* call, int64_t* offsets_out);

/* `call` is const for the caller's fields; lib_flags / lib_ctx are updated in place (see FtrfCall). */
int ftrf_forward(const FtrfField* field, const FtrfCall* call, void* ws, size_t ws_bytes, void* stream);
/* ftrf_forward that also hands the call's counters to the host as soon as they are final (after the
 * integration stage; the appearance / shading / compositing kernels follow on the stream): counters_host
 * (pinned int32[8], may be NULL) receives {valid samples needed, shaded samples needed, overflow flags, ...} by
 * an async copy and counters_event (a cudaEvent_t, may be NULL) is recorded behind it.  The Python wrapper
 * waits on that event instead of the stream: the float the reference's renderer returns (renderer.py:24) no
 * longer costs the tail of the forward. */
int ftrf_forward_ex(const FtrfField* field, const FtrfCall* call, void* ws, size_t ws_bytes, void* stream,
                    int32_t* counters_host, void* counters_event);
int ftrf_backward(const FtrfField* field, const FtrfCall* call, void* ws, size_t ws_bytes, void* stream);
/* The same work in two halves, for callers that overlap a gradient all-reduce with it (dist.py):
 * part 1 = everything but the density backward (all gradients except the density factors' are final after
 * it), part 2 = the density backward, part 0 = both (ftrf_backward). */
int ftrf_backward_part(const FtrfField* field, const FtrfCall* call, void* workspace, size_t workspace_bytes,
                       void* stream, int part);

int ftrf_stage_march(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);
int ftrf_stage_density(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);
int ftrf_stage_integrate(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);
int ftrf_stage_appearance(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);
int ftrf_stage_shade(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);
int ftrf_stage_composite(const FtrfField*, const FtrfCall*, void* ws, size_t ws_bytes, void* stream);

/* loss_out[0] = mean((rgb-target)^2); g_rgb_out = d loss / d rgb (may be NULL). n = element count. */
int ftrf_mse_loss(const float* rgb, const float* target, int64_t n, float* loss_out, float* g_rgb_out,
                  void* stream);

/* volume: float (Z,Y,X) of {0,1}; bits: uint32[ftrf_alpha_mask_words(X,Y,Z)]. */
size_t ftrf_alpha_mask_words(int32_t x, int32_t y, int32_t z);
int ftrf_pack_alpha_mask(const float* volume, int32_t x, int32_t y, int32_t z, uint32_t* bits, void* stream);
/* Optional acceleration structure of the march kernel (no reference counterpart): one bit per 8x8x8
 * texel block (dilated by one texel), coarse: uint32[ftrf_alpha_mask_coarse_words(X,Y,Z)].  It never
 * changes a keep / cull decision, it only answers the empty-space ones early. */
size_t ftrf_alpha_mask_coarse_words(int32_t x, int32_t y, int32_t z);
int ftrf_build_coarse_mask(const uint32_t* bits, int32_t x, int32_t y, int32_t z, uint32_t* coarse, void* stream);

/* sigma_out[i] = feature2density(compute_densityfeature(normalize(xyz[i]))) for n world-space points,
 * or 0 where the field's alphaMask rejects the point (tensorBase.py:298-313). */
int ftrf_density_at(const FtrfField* field, const float* xyz, int64_t n, float* sigma_out, void* stream);

/* Occupancy rebuild, all bit-packed like ftrf_pack_alpha_mask's output.
 * ftrf_dense_alpha_bits: getDenseAlpha + compute_alpha + the threshold of updateAlphaMask
 *   (models/tensorBase.py:215-230, :298-318, :245).  lin (device) = linspace(0,1,x), linspace(0,1,y),
 *   linspace(0,1,z) back to back; lattice point = aabb0*(1-s) + aabb1*s; raw bit (x,y,z) =
 *   clamp(1 - exp(-sigma*length), 0, 1) >= thres, sigma as ftrf_density_at.  alpha_out (NULL or float
 *   [z][y][x]) additionally receives the clamped alpha lattice (getDenseAlpha's first result, transposed
 *   as updateAlphaMask :237 does).
 * ftrf_dilate_alpha_mask: max_pool3d(kernel 3, padding 1) + threshold assignments (:243-246) = 3^3
 *   dilation of the raw bits.  stats (device int32[8]) receives min x,y,z / max x,y,z of the set texels
 *   (INT_MAX / -1 when none) and their count as a 64-bit value in [6..7] (:248-257).
 * ftrf_unpack_alpha_mask: bits -> float (Z,Y,X) volume of {0,1} (AlphaGridMask.alpha_volume, :31-39). */
int ftrf_dense_alpha_bits(const FtrfField* field, const float* lin, int32_t x, int32_t y, int32_t z, float length,
                          float thres, uint32_t* raw_bits, float* alpha_out, void* stream);
int ftrf_dilate_alpha_mask(const uint32_t* raw_bits, int32_t x, int32_t y, int32_t z, uint32_t* bits, int32_t* stats,
                           void* stream);
int ftrf_unpack_alpha_mask(const uint32_t* bits, int32_t x, int32_t y, int32_t z, float* volume, void* stream);
/* filtering_rays (models/tensorBase.py:259-288): keep[i] = 1 iff ray i ([n][ray_cols] floats, origin then
 * direction) survives.  bbox_only = 0: any of the n_samples un-jittered march points of sample_ray(:189-208)
 * reads alphaMask > 0 (:285-286; needs field->has_alpha_mask); bbox_only != 0: slab test t_max > t_min (:277-283). */
int ftrf_filter_rays(const FtrfField* field, const float* rays, int32_t ray_cols, int64_t n, int32_t n_samples,
                     int32_t bbox_only, uint8_t* keep, void* stream);

/* The module methods the reference exposes on POINTS instead of rays (models/tensoRF.py:207-263, :358-415
 * compute_densityfeature / compute_appfeature; models/mlp.py:15-19, :41-69, :88-107, :127-155 renderModule), on
 * the same kernels as the per-ray path.  xn / viewdirs / features are device arrays of n rows; xn holds
 * NORMALISED coordinates (normalize_coord's output, tensorBase.py:130-131), as those methods take them.
 *   ftrf_density_feature : sigma_feature[n]   (before feature2density; decomposition mask from field->den.mask)
 *   ftrf_app_feature     : app_features[n][app_dim]   (gather + basis_mat; mask from field->app.mask)
 *   ftrf_shade_points    : rgb[n][3] = renderModule(xn, viewdirs, features, field->enc_mask)
 * The last two need a scratch buffer of ftrf_points_workspace_bytes(field, n) bytes (device memory). */
size_t ftrf_points_workspace_bytes(const FtrfField* field, int64_t n);
int ftrf_density_feature(const FtrfField* field, const float* xn, int64_t n, float* sigma_feature_out, void* stream);
int ftrf_app_feature(const FtrfField* field, const float* xn, int64_t n, float* features_out, void* ws, size_t ws_bytes,
                     void* stream);
int ftrf_shade_points(const FtrfField* field, const float* xn, const float* viewdirs, const float* features, int64_t n,
                      float* rgb_out, void* ws, size_t ws_bytes, void* stream);
/* ftrf_shade_points for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs, outputs and
 * kernels as ftrf_shade_points, but the scratch (counters, the MLP's tf32 weight images, row buffers) is a file-scope
 * `static __device__` array of FTRF_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library
 * holds until exit; the call performs no cudaMalloc / cudaFree and calls no allocating routine.
 *   limits : n <= FTRF_STATIC_MAX_POINTS (8192) and ftrf_points_workspace_bytes(field, n) <= FTRF_STATIC_SCRATCH_BYTES
 *            (16 MiB; the shipped MLP shapes need 0.8 MB of weight images + ~1.8 KB per point: 15.2 MB at the limit);
 *            beyond either the call returns FTRF_E_ARG / FTRF_E_WORKSPACE with text and launches nothing.
 *   sharing: all calls (all streams, all host threads) use the ONE array of the current device, and it keeps the last
 *            call's contents; each call rewrites what it needs (counters, weight images), so calls ordered on one stream
 *            are independent — calls that overlap on the GPU race and are not supported. */
#define FTRF_STATIC_SCRATCH_BYTES (16u << 20)
#define FTRF_STATIC_MAX_POINTS 8192
int ftrf_shade_points_static(const FtrfField* field, const float* xn, const float* viewdirs, const float* features, int64_t n,
                             float* rgb_out, void* stream);

/* Per-stage device timing (bench / profiling aid; new, no reference counterpart).  When enabled,
 * ftrf_forward / ftrf_backward record CUDA events on the caller's stream around every stage.
 * ftrf_profile_read must be called after the stream has been synchronised; ms_out receives
 * FTRF_PROF_STAGES durations in milliseconds (-1 for a stage that did not run since enabling).
 * The switch and the events belong to the CALLING HOST THREAD (thread-local), like ftrf_last_error. */
#define FTRF_PROF_STAGES 13
int ftrf_profile_enable(int on);
int ftrf_profile_read(float* ms_out /* host */);
const char* ftrf_profile_stage_name(int stage);
/* All-reduce(sum) of a symmetric-memory fp32 buffer through its NVSwitch multicast address (NVLS): this rank
 * reduces slice `rank` of `world` with multimem.ld_reduce and broadcasts it with multimem.st.  The caller
 * issues a cross-GPU barrier on the stream before and after (fewtensorf_b200/dist.py).  New; the reference
 * has no multi-GPU code. */
int ftrf_multimem_allreduce(void* multicast_ptr, size_t n_floats, int rank, int world, void* stream);
/* The same exchange with plain peer loads / stores over NVLink: peer_ptrs[r] is rank r's copy of the symmetric
 * buffer (host array of `world` device pointers, world = 2, 4 or 8); same barrier contract. */
int ftrf_p2p_allreduce(void* const* peer_ptrs, size_t n_floats, int rank, int world, void* stream);
/* The same two exchanges over the parts of the buffer that can hold a gradient: a region is n_rows runs of row_len
 * floats, row_stride floats apart, starting at float `offset` of the buffer (all multiples of 4; lines and the dense
 * basis / MLP tail are single rows).  Rows and columns of a plane outside the alphaMask's texel bounding box never
 * receive a gradient (their samples are culled in stage 1), so they are exactly zero on every rank and need no
 * exchange (fewtensorf_b200/module.py: TensorBase.grad_regions).  regions: HOST array, at most FTRF_MAX_REGIONS. */
#define FTRF_MAX_REGIONS 24
typedef struct FtrfRegion {
  int64_t offset;
  int64_t row_len;
  int64_t row_stride;
  int64_t n_rows;
} FtrfRegion;
int ftrf_p2p_allreduce_regions(void* const* peer_ptrs, const FtrfRegion* regions, int n_regions, int rank, int world,
                               void* stream);
int ftrf_multimem_allreduce_regions(void* multicast_ptr, const FtrfRegion* regions, int n_regions, int rank, int world,
                                    void* stream);
/* The region exchange with both cross-GPU barriers inside the kernel (no bracketing barrier launches): flag_ptrs[r] is
 * rank r's copy of a zero-initialised symmetric int32[128] array, epoch = 1, 2, 3, ... per call (the same on every
 * rank).  multicast_ptr != NULL selects the NVSwitch multimem path, NULL the peer-load path. */
int ftrf_allreduce_regions_fused(void* const* peer_ptrs, void* multicast_ptr, void* const* flag_ptrs, const FtrfRegion* regions,
                                 int n_regions, int rank, int world, uint32_t epoch, void* stream);
/* Regularisers of one factor set, value and gradient in one pass.
 *   TV : sum_k 1e-2 * TVLoss(plane_k)      models/tensoRF.py:195-205 with loss.py:120-141 (VM planes only; the
 *        reference's CP variant, tensoRF.py:474-484, is 0/0 = NaN and is refused with FTRF_E_ARG)
 *   L1 : sum_k mean|plane_k| (+ mean|line_k| when with_lines)      models/tensoRF.py:189-193, :468-472
 * loss_out (device double[2], caller-zeroed, may be NULL) receives += TV and += L1, unweighted, as the reference
 * methods return them.  Where fs->g_plane / g_line are non-NULL, d(w_tv*TV + w_l1*L1)/d factor is ADDED to them
 * (train.py:345-371 adds the weighted losses to total_loss); dev_scale (device float, may be NULL) multiplies
 * both weights (autograd's upstream gradient).  `model` is FTRF_MODEL_*. */
int ftrf_reg_factors(const FtrfFactorSet* fs, int model, float w_tv, float w_l1, const float* dev_scale,
                     double* loss_out, int with_lines, void* stream);
/* Factor maintenance between training stages, on the channel-last factor layout.
 *   ftrf_resample_factor : F.interpolate(mode='bilinear', align_corners=True) of one plane ([H][W][C] -> [H2][W2][C])
 *                          or line (W = W2 = 1)      models/tensoRF.py:267-288, :418-435 upsample_volume_grid
 *   ftrf_crop_factor     : rows [y0,y1) x columns [x0,x1) of a plane / line into a dense [y1-y0][x1-x0][C] buffer
 *                          models/tensoRF.py:290-327, :437-466 shrink
 *   ftrf_ortho_reg       : vector_comp_diffs, models/tensoRF.py:175-187 — for each of n_lines lines ([L][C]) the mean
 *                          |off-diagonal| of its C x C Gram matrix is ADDED to loss_out (device double, may be NULL);
 *                          where g_lines is non-NULL, weight * dev_scale * d(that mean)/d line is ADDED to g_lines[k].
 *                          lines / g_lines / n_comp / line_len are HOST arrays of n_lines entries. */
int ftrf_resample_factor(const float* src, int32_t C, int32_t H, int32_t W, float* dst, int32_t H2, int32_t W2, void* stream);
int ftrf_crop_factor(const float* src, int32_t C, int32_t H, int32_t W, int32_t y0, int32_t y1, int32_t x0, int32_t x1,
                     float* dst, void* stream);
int ftrf_ortho_reg(const float* const* lines, float* const* g_lines, const int32_t* n_comp, const int32_t* line_len,
                   int n_lines, float weight, const float* dev_scale, double* loss_out, void* stream);
/* torch.optim.Adam(betas, eps; no weight decay / amsgrad) step `step` (1-based) over a host table of segments
 * in ONE launch (train.py:272-273 optimizer, :374-376 step; lr per segment = the param group's current lr,
 * which train.py:389-392 decays on the host).  All four arrays of a segment share one element order. */
typedef struct FtrfAdamSegment {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t n;
  float lr;
  int32_t pad_;
} FtrfAdamSegment;
int ftrf_adam_step(const FtrfAdamSegment* segs /* host */, int n_segs, float beta1, float beta2, float eps, int step,
                   void* stream);
/* The same step, skipped as a whole (parameters and moments untouched) when the device float `skip_if_nonzero` is not 0
 * at launch time on the device — NULL = never.  Point it at the gradient buffer's exchanged overflow flag
 * (FtrfCall.overflow_out) and a step whose sample lists overflowed on any rank is dropped on every rank without a host
 * round trip; the analogue of a loss scaler's found-inf skip.  No reference counterpart (the reference has no lists). */
int ftrf_adam_step_guarded(const FtrfAdamSegment* segs /* host */, int n_segs, float beta1, float beta2, float eps, int step,
                           const float* skip_if_nonzero /* device, may be NULL */, void* stream);
/* Camera rays on the device: rows row0, row0+row_step, ... (n_rows of them) of an H x W
 * pin-hole image, [n_rows*W][6] = (origin, direction).  dataLoader/ray_utils.py:24-41 get_ray_directions (pixel
 * centre + 0.5), :44-63 blender variant (y, z negated), dataLoader/blender.py:59 normalisation,
 * ray_utils.py:67-85 get_rays (d @ c2w[:3,:3]^T, origin c2w[:3,3]).  row_step = world size interleaves the
 * rows of a frame over the GPUs. */
typedef struct FtrfCamera {
  int32_t H, W;
  float fx, fy, cx, cy;
  float c2w[12];            /* row-major 3x4 camera-to-world                                  */
  int32_t normalize;        /* divide the camera-space direction by its norm (blender.py:59)  */
  int32_t blender;          /* get_ray_directions_blender signs                               */
} FtrfCamera;
int ftrf_generate_rays(const FtrfCamera* cam /* host */, int32_t row0, int32_t row_step, int32_t n_rows,
                       float* rays_out, void* stream);
/* Side-stream zeroing of the flat gradient buffer (new; the reference's optimizer.zero_grad(), train.py:374, is a
 * per-tensor loop).  ftrf_event_record marks the fork point on the caller's stream; ftrf_zero_after makes
 * side_stream wait for it, memsets `bytes` at ptr there and records done_event; ftrf_stream_wait_event joins
 * (the backward's stream waits for done_event).  The memset thus runs beside the forward kernels. */
int ftrf_event_record(void* event, void* stream);
int ftrf_stream_wait_event(void* stream, void* event);
int ftrf_zero_after(void* ptr, size_t bytes, void* side_stream, void* wait_event, void* done_event);
/* Number of CUDA kernels this library has launched since it was loaded (bench.py's gpu_launches). */
unsigned long long ftrf_launch_count(void);
/* Adjust that count by `delta`: a host that captured a launch sequence in a CUDA graph subtracts what the capture
 * counted (nothing ran) and adds the sequence's launches on every replay, so the count stays "kernels that ran". */
void ftrf_launch_count_add(long long delta);

#ifdef __cplusplus
}
#endif
#endif /* FEWTENSORF_H */

// abi.cu — the extern "C" boundary: argument checks, workspace carving, stage sequencing.
// No torch types.  Mutable state: the thread-local error string, the thread-local profiling switch / events, each
// host thread's cached side stream (k_shade_fused.cu) and one atomic launch counter; everything that belongs to a
// CALL travels in the caller's FtrfCall (lib_flags / lib_ctx), so the entry points are re-entrant.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "common.cuh"
#include "fused.cuh"

namespace ftrf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};   // kernels this library has launched (every launch is followed by check_launch)

int check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return FTRF_E_CUDA;
  }
  return FTRF_OK;
}

// ---- optional per-stage timing --------------------------------------------------------------
static const char* kStageNames[FTRF_PROF_STAGES] = {
    "march", "density_fwd", "integrate_fwd", "app_fwd", "shade_fwd", "composite",
    "integrate_bwd", "shade_bwd_out", "shade_bwd_dgrad", "shade_wgrad", "shade_wgrad1_unfused", "app_bwd", "density_bwd"};
static thread_local bool g_prof_on = false;
static thread_local cudaEvent_t g_ev[FTRF_PROF_STAGES][2];
static thread_local bool g_ev_made = false, g_ev_used[FTRF_PROF_STAGES];
bool prof_enabled() { return g_prof_on; }

void prof_begin(int stage, cudaStream_t st) {
  if (g_prof_on) cudaEventRecord(g_ev[stage][0], st);
}
void prof_end(int stage, cudaStream_t st) {
  if (g_prof_on) { cudaEventRecord(g_ev[stage][1], st); g_ev_used[stage] = true; }
}

static inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

size_t carve(const FtrfField& f, const FtrfCall& c, int64_t* off, void* base, Ws* ws) {
  const size_t N = (size_t)c.n_rays, V = (size_t)c.cap_samples, A = (size_t)c.cap_app;
  const size_t D = (size_t)f.app_dim, H = (size_t)f.feature_c;
  const bool train = c.save_for_backward != 0;
  const bool mlp = f.shading != FTRF_SHADE_SH;
  size_t sz[FTRF_WS_COUNT];
  sz[FTRF_WS_COUNTERS] = 8 * 4;
  sz[FTRF_WS_TMIN] = N * 4;
  sz[FTRF_WS_RAY_OFF] = N * 4;
  sz[FTRF_WS_RAY_CNT] = N * 4;
  sz[FTRF_WS_ACC] = N * 4;
  sz[FTRF_WS_RGB_PRE] = N * 12;
  sz[FTRF_WS_S_RAY] = V * 4;
  sz[FTRF_WS_S_IDX] = V * 4;
  sz[FTRF_WS_S_SIGMA] = V * 4;
  sz[FTRF_WS_S_W] = V * 4;
  sz[FTRF_WS_S_T] = train ? V * 4 : 0;
  sz[FTRF_WS_S_APP] = V * 4;
  sz[FTRF_WS_S_GSIG] = train ? V * 4 : 0;
  sz[FTRF_WS_A_SLOT] = A * 4;
  sz[FTRF_WS_A_FEAT] = A * D * 4;
  sz[FTRF_WS_A_RGB] = A * 12;
  const size_t LDX = (size_t)((f.in_dim + 1 + 7) & ~7);
  // the fused MLP kernels may save activations in padded blocks of 16 rows (fused.cuh: act16_off): 17/16 of the row-major size
  const size_t A16 = (A + 15) / 16;
  sz[FTRF_WS_A_H1] = mlp ? A16 * act16_block_floats((int)H) * 4 : 0;
  sz[FTRF_WS_A_H2] = mlp ? A16 * act16_block_floats((int)H) * 4 : 0;
  sz[FTRF_WS_A_X] = mlp ? A16 * act16_block_floats((int)LDX) * 4 : 0;
  sz[FTRF_WS_A_GX] = (train && mlp) ? A * LDX * 4 : 0;
  sz[FTRF_WS_A_GRGB] = train ? A * 12 : 0;
  sz[FTRF_WS_A_GH1] = (train && mlp) ? A16 * act16_block_floats((int)H) * 4 : 0;
  sz[FTRF_WS_A_GH2] = (train && mlp) ? A * H * 4 : 0;
  sz[FTRF_WS_A_GFEAT] = train ? A * D * 4 : 0;
  sz[FTRF_WS_WIMG] = mlp ? fused_wimg_bytes() : 0;
  size_t cur = 0, offs[FTRF_WS_COUNT];
  for (int i = 0; i < FTRF_WS_COUNT; ++i) {
    offs[i] = cur;
    cur += align_up(sz[i]);
    if (off) off[i] = (int64_t)offs[i];
  }
  if (ws && base) {
    char* b = (char*)base;
    ws->counters = (int32_t*)(b + offs[FTRF_WS_COUNTERS]);
    ws->tmin = (float*)(b + offs[FTRF_WS_TMIN]);
    ws->ray_off = (int32_t*)(b + offs[FTRF_WS_RAY_OFF]);
    ws->ray_cnt = (int32_t*)(b + offs[FTRF_WS_RAY_CNT]);
    ws->acc = (float*)(b + offs[FTRF_WS_ACC]);
    ws->rgb_pre = (float*)(b + offs[FTRF_WS_RGB_PRE]);
    ws->s_ray = (int32_t*)(b + offs[FTRF_WS_S_RAY]);
    ws->s_idx = (int32_t*)(b + offs[FTRF_WS_S_IDX]);
    ws->s_sigma = (float*)(b + offs[FTRF_WS_S_SIGMA]);
    ws->s_w = (float*)(b + offs[FTRF_WS_S_W]);
    ws->s_T = (float*)(b + offs[FTRF_WS_S_T]);
    ws->s_app = (int32_t*)(b + offs[FTRF_WS_S_APP]);
    ws->s_gsig = (float*)(b + offs[FTRF_WS_S_GSIG]);
    ws->a_slot = (int32_t*)(b + offs[FTRF_WS_A_SLOT]);
    ws->a_feat = (float*)(b + offs[FTRF_WS_A_FEAT]);
    ws->a_rgb = (float*)(b + offs[FTRF_WS_A_RGB]);
    ws->a_h1 = (float*)(b + offs[FTRF_WS_A_H1]);
    ws->a_h2 = (float*)(b + offs[FTRF_WS_A_H2]);
    ws->a_grgb = (float*)(b + offs[FTRF_WS_A_GRGB]);
    ws->a_gh1 = (float*)(b + offs[FTRF_WS_A_GH1]);
    ws->a_gh2 = (float*)(b + offs[FTRF_WS_A_GH2]);
    ws->a_gfeat = (float*)(b + offs[FTRF_WS_A_GFEAT]);
    ws->a_x = (float*)(b + offs[FTRF_WS_A_X]);
    ws->a_gx = (float*)(b + offs[FTRF_WS_A_GX]);
    ws->wimg = mlp ? (float*)(b + offs[FTRF_WS_WIMG]) : nullptr;
    ws->ldx = (int)LDX;
    ws->direct_xn = nullptr;
    ws->direct_view = nullptr;
  }
  return cur;
}

static int check_factor_set(const FtrfFactorSet& s, int model, const char* name, bool need_grad) {
  for (int k = 0; k < 3; ++k) {
    if (s.n_comp[k] <= 0 || s.n_comp[k] % 4) { set_error("%s: n_comp[%d]=%d must be a positive multiple of 4", name, k, s.n_comp[k]); return FTRF_E_ARG; }
    if (!s.line[k] || s.line_len[k] < 1) { set_error("%s: line %d missing", name, k); return FTRF_E_ARG; }
    if (model == FTRF_MODEL_VM && (!s.plane[k] || s.plane_w[k] < 1 || s.plane_h[k] < 1)) { set_error("%s: plane %d missing", name, k); return FTRF_E_ARG; }
    if (model == FTRF_MODEL_CP && s.n_comp[k] != s.n_comp[0]) { set_error("%s: CP needs equal n_comp", name); return FTRF_E_ARG; }
    if (need_grad && (!s.g_line[k] || (model == FTRF_MODEL_VM && !s.g_plane[k]))) { set_error("%s: gradient pointer %d missing", name, k); return FTRF_E_ARG; }
  }
  if (model == FTRF_MODEL_VM) {
    // one resolution per axis (tensoRF.py:186-198): plane k is [size(mat1 k)][size(mat0 k)], line k is size(vecm k)
    const int size[3] = {s.plane_w[0], s.plane_h[0], s.line_len[0]};
    for (int k = 0; k < 3; ++k)
      if (s.plane_w[k] != size[mat0(k)] || s.plane_h[k] != size[mat1(k)] || s.line_len[k] != size[vecm(k)]) {
        set_error("%s: plane / line %d does not match the per-axis resolutions %d x %d x %d", name, k, size[0], size[1], size[2]);
        return FTRF_E_ARG;
      }
  }
  return FTRF_OK;
}

static int check_args(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, bool backward, Ws* ws) {
  if (!f || !c || !wsb) { set_error("null argument"); return FTRF_E_ARG; }
  if (f->model != FTRF_MODEL_VM && f->model != FTRF_MODEL_CP) { set_error("unknown model kind %d", f->model); return FTRF_E_ARG; }
  if (f->shading < 0 || f->shading > FTRF_SHADE_SH) { set_error("unknown shading kind %d", f->shading); return FTRF_E_ARG; }
  if (c->n_rays < 0 || c->n_rays > 0x7fffffff || c->ray_cols < 6 || c->n_samples < 1) { set_error("bad ray batch (n_rays %lld, cols %d, S %d)", (long long)c->n_rays, c->ray_cols, c->n_samples); return FTRF_E_ARG; }
  if (c->cap_samples < 0 || c->cap_samples > 0x7fffffff || c->cap_app < 0 || c->cap_app > 0x7fffffff) { set_error("capacity out of range"); return FTRF_E_ARG; }
  if (c->n_rays && (!c->rays || !c->rgb_map || !c->depth_map)) { set_error("null ray/output pointer"); return FTRF_E_ARG; }
  if ((c->ndc || c->is_train) && !c->jitter) { set_error("jitter / z table missing"); return FTRF_E_ARG; }
  if (f->has_alpha_mask && !f->am_bits) { set_error("alpha mask bits missing"); return FTRF_E_ARG; }
  if (f->app_dim < 1 || f->app_dim > 64) { set_error("app_dim %d unsupported", f->app_dim); return FTRF_E_ARG; }
  if (int e = check_factor_set(f->den, f->model, "density", backward)) return e;
  if (int e = check_factor_set(f->app, f->model, "appearance", backward)) return e;
  if (!f->basis || (backward && !f->g_basis)) { set_error("basis_mat missing"); return FTRF_E_ARG; }
  if (f->shading != FTRF_SHADE_SH) {
    if (!f->w1 || !f->b1 || !f->w2 || !f->b2 || !f->w3 || !f->b3) { set_error("MLP weights missing"); return FTRF_E_ARG; }
    if (backward && (!f->g_w1 || !f->g_b1 || !f->g_w2 || !f->g_b2 || !f->g_w3 || !f->g_b3)) { set_error("MLP gradient pointers missing"); return FTRF_E_ARG; }
  }
  if (backward && (!c->save_for_backward || !c->g_rgb_map)) { set_error("backward needs save_for_backward and g_rgb_map"); return FTRF_E_ARG; }
  size_t need = carve(*f, *c, nullptr, wsb, ws);
  if (ws_bytes < need) { set_error("workspace too small: %zu < %zu", ws_bytes, need); return FTRF_E_WORKSPACE; }
  return FTRF_OK;
}

__global__ void mse_kernel(const float* __restrict__ rgb, const float* __restrict__ tgt, int64_t n, float inv_n,
                           float* __restrict__ loss, float* __restrict__ g) {
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float d = rgb[i] - tgt[i];
    acc = fmaf(d, d, acc);
    if (g) g[i] = 2.f * d * inv_n;
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, d);
  __shared__ float part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += part[w];
    atomicAdd(loss, s * inv_n);
  }
}

}  // namespace ftrf

using namespace ftrf;

extern "C" int ftrf_profile_enable(int on) {
  if (on && !g_ev_made) {
    for (int i = 0; i < FTRF_PROF_STAGES; ++i)
      for (int j = 0; j < 2; ++j)
        if (cudaEventCreate(&g_ev[i][j]) != cudaSuccess) return check_launch("profile event");
    g_ev_made = true;
  }
  for (int i = 0; i < FTRF_PROF_STAGES; ++i) g_ev_used[i] = false;
  g_prof_on = on != 0;
  return FTRF_OK;
}
extern "C" int ftrf_profile_read(float* ms_out) {
  if (!ms_out) { set_error("profile_read: null"); return FTRF_E_ARG; }
  for (int i = 0; i < FTRF_PROF_STAGES; ++i) {
    ms_out[i] = -1.f;
    if (g_ev_made && g_ev_used[i]) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, g_ev[i][0], g_ev[i][1]) == cudaSuccess) ms_out[i] = ms;
      else cudaGetLastError();
    }
  }
  return FTRF_OK;
}
extern "C" const char* ftrf_profile_stage_name(int stage) {
  return (stage >= 0 && stage < FTRF_PROF_STAGES) ? kStageNames[stage] : "";
}

extern "C" unsigned long long ftrf_launch_count(void) { return ftrf::g_launches.load(std::memory_order_relaxed); }
extern "C" void ftrf_launch_count_add(long long delta) { ftrf::g_launches.fetch_add((unsigned long long)delta, std::memory_order_relaxed); }
extern "C" int ftrf_abi_version(void) { return FTRF_ABI_VERSION; }
extern "C" const char* ftrf_last_error(void) { return g_err; }

extern "C" size_t ftrf_workspace_layout(const FtrfField* f, const FtrfCall* c, int64_t* offsets_out) {
  if (!f || !c) { set_error("null argument"); return 0; }
  return carve(*f, *c, offsets_out, nullptr, nullptr);
}

#define STAGE_FN(name, launcher)                                                                          \
  extern "C" int name(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, void* stream) {   \
    Ws ws;                                                                                                 \
    if (int e = check_args(f, c, wsb, ws_bytes, false, &ws)) return e;                                     \
    return launcher(*f, *c, ws, (cudaStream_t)stream);                                                     \
  }

static int march_with_reset(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (cudaMemsetAsync(ws.counters, 0, 32, st) != cudaSuccess) return check_launch("memset counters");
  return launch_march(f, c, ws, st);
}

STAGE_FN(ftrf_stage_march, march_with_reset)
STAGE_FN(ftrf_stage_density, launch_density)
STAGE_FN(ftrf_stage_integrate, launch_integrate)
STAGE_FN(ftrf_stage_appearance, launch_appearance)
STAGE_FN(ftrf_stage_shade, launch_shade)
STAGE_FN(ftrf_stage_composite, launch_composite)

extern "C" int ftrf_forward_ex(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, void* stream,
                               int32_t* counters_host, void* counters_event) {
  Ws ws;
  if (int e = check_args(f, c, wsb, ws_bytes, false, &ws)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  const bool ahead = shade_path_is_fused();
  if (ahead)
    if (int e = shade_prep_ahead(*f, *c, ws, st, 0)) return e;
  prof_begin(0, st);
  if (int e = march_with_reset(*f, *c, ws, st)) return e;
  if (ahead)
    if (int e = shade_prep_ahead(*f, *c, ws, st, 1)) return e;
  prof_end(0, st); prof_begin(1, st);
  if (int e = launch_density(*f, *c, ws, st)) return e;
  prof_end(1, st); prof_begin(2, st);
  if (int e = launch_integrate(*f, *c, ws, st)) return e;
  prof_end(2, st);
  // the list counters and overflow flags are final here: hand them to the host now, so that its one wait
  // of the call ends while the appearance / shading / compositing kernels are still running
  if (counters_host && cudaMemcpyAsync(counters_host, ws.counters, 32, cudaMemcpyDeviceToHost, st) != cudaSuccess)
    return check_launch("counters copy");
  if (counters_event && cudaEventRecord((cudaEvent_t)counters_event, st) != cudaSuccess) return check_launch("counters event");
  prof_begin(3, st);
  if (int e = launch_appearance(*f, *c, ws, st)) return e;
  prof_end(3, st); prof_begin(4, st);
  if (int e = launch_shade(*f, *c, ws, st)) return e;
  prof_end(4, st); prof_begin(5, st);
  int e = launch_composite(*f, *c, ws, st);
  prof_end(5, st);
  return e;
}

extern "C" int ftrf_forward(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, void* stream) {
  return ftrf_forward_ex(f, c, wsb, ws_bytes, stream, nullptr, nullptr);
}

// part 1: everything but the density backward (after it every gradient except the density factors' is
// final); part 2: the density backward; part 0: both.
extern "C" int ftrf_backward_part(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, void* stream, int part) {
  Ws ws;
  if (part < 0 || part > 2) { set_error("backward: part %d", part); return FTRF_E_ARG; }
  if (int e = check_args(f, c, wsb, ws_bytes, true, &ws)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  if (part != 2) {
    prof_begin(6, st);
    // MLP shading on the fused path: the MLP backward only needs dL/d rgb of the shaded samples (one small kernel);
    // the reverse scan, whose sigma gradients only the density backward reads, runs beside it on the side stream
    cudaStream_t scan = (f->shading != FTRF_SHADE_SH && shade_path_is_fused() && !g_prof_on) ? side_scan_begin(*c, st) : nullptr;
    if (scan) {
      if (int e = launch_integrate_bwd_sigma(*f, *c, ws, scan)) return e;
      side_scan_end(*c);
      if (int e = launch_grgb(*f, *c, ws, st)) return e;
    } else {
      if (int e = launch_integrate_bwd(*f, *c, ws, st)) return e;
    }
    prof_end(6, st);
    if (int e = launch_shade_bwd(*f, *c, ws, st, 0)) return e;     // stages 7-10 marked inside
    prof_begin(11, st);
    if (int e = launch_appearance_bwd(*f, *c, ws, st)) return e;
    prof_end(11, st);
  }
  if (part != 1) {
    if (int e = side_scan_join(*c, st)) return e;
    prof_begin(12, st);
    int e = launch_density_bwd(*f, *c, ws, st);
    prof_end(12, st);
    if (e) return e;
  }
  return side_tail_join(*c, st);     // the un-permute of the first-layer weight gradient ran beside the kernels above
}

extern "C" int ftrf_backward(const FtrfField* f, const FtrfCall* c, void* wsb, size_t ws_bytes, void* stream) {
  return ftrf_backward_part(f, c, wsb, ws_bytes, stream, 0);
}

extern "C" int ftrf_mse_loss(const float* rgb, const float* target, int64_t n, float* loss_out, float* g_rgb_out,
                             void* stream) {
  if (!rgb || !target || !loss_out || n <= 0) { set_error("mse_loss: bad argument"); return FTRF_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemsetAsync(loss_out, 0, sizeof(float), st) != cudaSuccess) return check_launch("memset loss");
  int64_t blocks = (n + 255) / 256;
  int grid = (int)(blocks < FTRF_SMS ? blocks : FTRF_SMS);
  mse_kernel<<<grid, 256, 0, st>>>(rgb, target, n, 1.f / (float)n, loss_out, g_rgb_out);
  return check_launch("mse_loss");
}

// Stream plumbing of the gradient-buffer memset (runs beside the forward on a side stream): thin wrappers so the
// Python wrapper pays ~1 us per call instead of torch's stream context manager.
extern "C" int ftrf_event_record(void* event, void* stream) {
  if (!event) { set_error("event_record: null event"); return FTRF_E_ARG; }
  if (cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream) != cudaSuccess) return check_launch("event_record");
  return FTRF_OK;
}
extern "C" int ftrf_stream_wait_event(void* stream, void* event) {
  if (!event) { set_error("stream_wait_event: null event"); return FTRF_E_ARG; }
  if (cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0) != cudaSuccess) return check_launch("stream_wait_event");
  return FTRF_OK;
}
extern "C" int ftrf_zero_after(void* ptr, size_t bytes, void* side_stream, void* wait_event, void* done_event) {
  if (!ptr || !wait_event || !done_event) { set_error("zero_after: null argument"); return FTRF_E_ARG; }
  cudaStream_t ss = (cudaStream_t)side_stream;
  if (cudaStreamWaitEvent(ss, (cudaEvent_t)wait_event, 0) != cudaSuccess) return check_launch("zero_after wait");
  if (cudaMemsetAsync(ptr, 0, bytes, ss) != cudaSuccess) return check_launch("zero_after memset");
  if (cudaEventRecord((cudaEvent_t)done_event, ss) != cudaSuccess) return check_launch("zero_after record");
  return FTRF_OK;
}

extern "C" int ftrf_density_at(const FtrfField* f, const float* xyz, int64_t n, float* sigma_out, void* stream) {
  if (!f || !xyz || !sigma_out || n < 0) { set_error("density_at: bad argument"); return FTRF_E_ARG; }
  if (int e = check_factor_set(f->den, f->model, "density", false)) return e;
  if (f->has_alpha_mask && !f->am_bits) { set_error("alpha mask bits missing"); return FTRF_E_ARG; }
  return launch_density_at(*f, xyz, n, sigma_out, (cudaStream_t)stream);
}

// ---- the module methods on points (compute_densityfeature / compute_appfeature / renderModule) ---------------
namespace ftrf {
__global__ void set_counters_kernel(int32_t* counters, int32_t n_app) {
  if (threadIdx.x < 8) counters[threadIdx.x] = threadIdx.x == 1 ? n_app : 0;
}
static FtrfCall points_call(int64_t n) {
  FtrfCall c;
  memset(&c, 0, sizeof(c));
  c.ray_cols = 6;
  c.n_samples = 1;
  c.cap_app = n;
  return c;
}
static int points_setup(const FtrfField* f, int64_t n, void* wsb, size_t ws_bytes, FtrfCall* c, Ws* ws, cudaStream_t st) {
  if (!f || !wsb || n < 0 || n > 0x7fffffff) { set_error("points call: bad argument (n %lld)", (long long)n); return FTRF_E_ARG; }
  if (f->model != FTRF_MODEL_VM && f->model != FTRF_MODEL_CP) { set_error("unknown model kind %d", f->model); return FTRF_E_ARG; }
  if (f->app_dim < 1 || f->app_dim > 64) { set_error("app_dim %d unsupported", f->app_dim); return FTRF_E_ARG; }
  *c = points_call(n);
  const size_t need = carve(*f, *c, nullptr, wsb, ws);
  if (ws_bytes < need) { set_error("points workspace too small: %zu < %zu", ws_bytes, need); return FTRF_E_WORKSPACE; }
  set_counters_kernel<<<1, 32, 0, st>>>(ws->counters, (int32_t)n);
  return check_launch("set_counters");
}
}  // namespace ftrf

extern "C" size_t ftrf_points_workspace_bytes(const FtrfField* f, int64_t n) {
  if (!f || n < 0) { set_error("points_workspace_bytes: bad argument"); return 0; }
  FtrfCall c = points_call(n);
  return carve(*f, c, nullptr, nullptr, nullptr);
}

extern "C" int ftrf_density_feature(const FtrfField* f, const float* xn, int64_t n, float* out, void* stream) {
  if (!f || !xn || !out || n < 0) { set_error("density_feature: bad argument"); return FTRF_E_ARG; }
  if (int e = check_factor_set(f->den, f->model, "density", false)) return e;
  return launch_density_feature(*f, xn, n, out, (cudaStream_t)stream);
}

extern "C" int ftrf_app_feature(const FtrfField* f, const float* xn, int64_t n, float* out, void* wsb, size_t ws_bytes,
                                void* stream) {
  if (!xn || !out) { set_error("app_feature: null argument"); return FTRF_E_ARG; }
  if (n == 0) return FTRF_OK;
  FtrfCall c;
  Ws ws;
  cudaStream_t st = (cudaStream_t)stream;
  if (int e = points_setup(f, n, wsb, ws_bytes, &c, &ws, st)) return e;
  if (int e = check_factor_set(f->app, f->model, "appearance", false)) return e;
  if (!f->basis) { set_error("basis_mat missing"); return FTRF_E_ARG; }
  ws.direct_xn = xn;
  ws.a_feat = out;
  return launch_appearance(*f, c, ws, st);
}

extern "C" int ftrf_shade_points(const FtrfField* f, const float* xn, const float* viewdirs, const float* features, int64_t n,
                                 float* rgb_out, void* wsb, size_t ws_bytes, void* stream) {
  if (!xn || !viewdirs || !features || !rgb_out) { set_error("shade_points: null argument"); return FTRF_E_ARG; }
  if (n == 0) return FTRF_OK;
  FtrfCall c;
  Ws ws;
  cudaStream_t st = (cudaStream_t)stream;
  if (int e = points_setup(f, n, wsb, ws_bytes, &c, &ws, st)) return e;
  if (f->shading < 0 || f->shading > FTRF_SHADE_SH) { set_error("unknown shading kind %d", f->shading); return FTRF_E_ARG; }
  if (f->shading != FTRF_SHADE_SH && (!f->w1 || !f->b1 || !f->w2 || !f->b2 || !f->w3 || !f->b3)) { set_error("MLP weights missing"); return FTRF_E_ARG; }
  ws.direct_xn = xn;
  ws.direct_view = viewdirs;
  ws.a_feat = const_cast<float*>(features);     // read-only here: the forward shading kernels never write it
  ws.a_rgb = rgb_out;
  return launch_shade(*f, c, ws, st);
}

// ---- ftrf_shade_points without a caller-provided (or any) allocation ---------------------------------------------
// Added; the reference has no counterpart.  Callers that must not call
// cudaMalloc (CUDA graph capture, tight serving loops) get the scratch of ftrf_shade_points — the counters, the tf32
// hi / lo weight images prep_weights_kernel writes and the fused forward reads, the carved row buffers — out of ONE
// file-scope device array that every GPU loading this library holds until exit.  In host code `g_static_scratch` is only
// the host-side stand-in of the symbol: its device address is looked up once per device (that device current) with
// cudaGetSymbolAddress and that pointer is what the kernels get.  All calls share the array and it keeps the last
// call's contents: what must start at a known value (the counters) is rewritten by set_counters_kernel at the top of
// every call, the weight images are rebuilt by every call; two calls that overlap on the GPU (different streams,
// no ordering) race on it.
namespace ftrf {
static __device__ __align__(256) unsigned char g_static_scratch[FTRF_STATIC_SCRATCH_BYTES];
constexpr int STATIC_MAX_DEVICES = 64;
static std::atomic<void*> g_static_scratch_addr[STATIC_MAX_DEVICES];

static int static_scratch(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= STATIC_MAX_DEVICES) {
    cudaGetLastError();
    set_error("static scratch: no current device (%d)", dev);
    return FTRF_E_CUDA;
  }
  void* p = g_static_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      set_error("static scratch: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return FTRF_E_CUDA;
    }
    g_static_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return FTRF_OK;
}
}  // namespace ftrf

extern "C" int ftrf_shade_points_static(const FtrfField* f, const float* xn, const float* viewdirs, const float* features,
                                        int64_t n, float* rgb_out, void* stream) {
  if (!f) { set_error("shade_points_static: null field"); return FTRF_E_ARG; }
  if (n < 0 || n > FTRF_STATIC_MAX_POINTS) {
    set_error("shade_points_static: %lld points exceed the static scratch's limit of %d (use ftrf_shade_points with your own buffer)",
              (long long)n, (int)FTRF_STATIC_MAX_POINTS);
    return FTRF_E_ARG;
  }
  const size_t need = ftrf_points_workspace_bytes(f, n);
  if (need == 0 || need > (size_t)FTRF_STATIC_SCRATCH_BYTES) {
    set_error("shade_points_static: this field needs %zu scratch bytes for %lld points, the static array holds %zu",
              need, (long long)n, (size_t)FTRF_STATIC_SCRATCH_BYTES);
    return FTRF_E_WORKSPACE;
  }
  void* scratch = nullptr;
  if (int e = static_scratch(&scratch)) return e;
  return ftrf_shade_points(f, xn, viewdirs, features, n, rgb_out, scratch, (size_t)FTRF_STATIC_SCRATCH_BYTES, stream);
}

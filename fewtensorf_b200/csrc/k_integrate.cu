// k_integrate.cu — stage 4: volume integration as a warp scan over each ray's packed samples.
// Replaces raw2alpha (models/tensorBase.py:21-28), the dists construction (:346 / :340-342), the
// shading threshold `weight > rayMarch_weight_thres` (:370), acc/depth (:377, :386-388),
// num_valid_samples (:390) and, in composite, rgb accumulation + background + clamp (:375-384).
//
// Culled samples have sigma = 0 => alpha = 0 => their transmittance factor (1 - 0 + 1e-10) rounds
// to exactly 1.0f, so scanning only the packed survivors is exact w.r.t. the dense cumprod up to
// the association order of the fp32 products.
// Algorithmic bytes: 12 B per packed sample forward (idx, sigma in; w out [+T when training]).
#include "common.cuh"

namespace ftrf {

constexpr int INT_WARPS = 8;

// dist_i * distance_scale exactly as forward builds it (:346, :340-342, :368)
__device__ __forceinline__ float sample_dist(const FtrfField& f, const FtrfCall& c, const Ray& ray, int i, float z,
                                             float dnorm) {
  if (i >= c.n_samples - 1) return 0.f;
  float d = __fsub_rn(sample_z(f, c, ray, i + 1), z);
  if (c.ndc) d = __fmul_rn(d, dnorm);
  return __fmul_rn(d, f.distance_scale);
}

__device__ __forceinline__ float ray_dnorm(const FtrfCall& c, const Ray& ray) {
  return c.ndc ? sqrtf(ray.d[0] * ray.d[0] + ray.d[1] * ray.d[1] + ray.d[2] * ray.d[2]) : 1.f;
}

// One warp per ray.  Pass 1 scans the ray's packed samples 32 at a time (the next chunk's loads are issued
// before this chunk's scan), writes w / T and keeps the "shaded" ballots in shared memory; then ONE atomicAdd
// per ray allocates its run of the shaded list (a per-chunk atomic put a ~1 us round trip on the critical
// path of every chunk of the longest ray), and pass 2 hands out the slots.
__global__ void __launch_bounds__(INT_WARPS * 32)
integrate_fwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  extern __shared__ uint32_t s_shade[];              // [INT_WARPS][chunks]
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int chunks = (c.n_samples + 31) >> 5;
  uint32_t* bal = s_shade + warp * chunks;
  __shared__ int s_cnt[INT_WARPS], s_base[INT_WARPS];
  const int64_t stride = (int64_t)gridDim.x * INT_WARPS;
  // (the trip count is the same for every warp of the CTA: the shaded-list space of its rays is allocated with ONE atomic
  // per CTA and pass — an atomicAdd per ray on the one cursor serialises in L2 at ~3.7 ns each, which for the 4096 rays
  // of a training batch was the whole 15 us of this kernel)
  for (int64_t r0 = (int64_t)blockIdx.x * INT_WARPS; r0 < c.n_rays; r0 += stride) {
    const int64_t r = r0 + warp;
    const bool have = r < c.n_rays;
    const int off = have ? ws.ray_off[r] : 0, cnt = have ? ws.ray_cnt[r] : 0;
    Ray ray = load_ray_ws(f, c, ws, have ? r : r0);
    const float dnorm = ray_dnorm(c, ray);
    float carry = 1.f, acc = 0.f, depth = 0.f;
    int n_shade = 0;
    int i_n = 0;
    float sg_n = 0.f;
    if (lane < cnt) { i_n = ws.s_idx[off + lane]; sg_n = ws.s_sigma[off + lane]; }
    for (int j0 = 0; j0 < cnt; j0 += 32) {
      const int j = j0 + lane;
      const bool live = j < cnt;
      const int i = i_n;
      const float sg = sg_n;
      if (j + 32 < cnt) { i_n = ws.s_idx[off + j + 32]; sg_n = ws.s_sigma[off + j + 32]; }
      float alpha = 0.f, z = 0.f;
      if (live) {
        z = sample_z(f, c, ray, i);
        float sd = sg * sample_dist(f, c, ray, i, z, dnorm);
        alpha = 1.f - expf(-sd);
      }
      float fac = (1.f - alpha) + 1e-10f;
      // inclusive product scan over the 32 lanes
      float incl = fac;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        float up = __shfl_up_sync(FULL_MASK, incl, d);
        if (lane >= d) incl *= up;
      }
      float excl = __shfl_up_sync(FULL_MASK, incl, 1);
      float T = carry * (lane == 0 ? 1.f : excl);
      carry *= __shfl_sync(FULL_MASK, incl, 31);
      float w = alpha * T;
      const bool shade = live && (w > f.weight_thres);
      const uint32_t b = __ballot_sync(FULL_MASK, shade);
      if (lane == 0) bal[j0 >> 5] = b;
      n_shade += __popc(b);
      if (live) {
        ws.s_w[off + j] = w;
        if (c.save_for_backward) ws.s_T[off + j] = T;
        acc += w;
        depth = fmaf(w, z, depth);
      }
    }
    if (lane == 0) s_cnt[warp] = n_shade;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
#pragma unroll
      for (int w = 0; w < INT_WARPS; ++w) { s_base[w] = tot; tot += s_cnt[w]; }
      const int b0 = tot ? atomicAdd(ws.counters + 1, tot) : 0;
#pragma unroll
      for (int w = 0; w < INT_WARPS; ++w) s_base[w] += b0;
    }
    __syncthreads();
    int base = s_base[warp];
    for (int j0 = 0; j0 < cnt; j0 += 32) {
      const int j = j0 + lane;
      const uint32_t b = bal[j0 >> 5];
      if (j < cnt) {
        int slot = -1;
        if ((b >> lane) & 1u) {
          slot = base + __popc(b & ((1u << lane) - 1u));
          if (slot < c.cap_app) ws.a_slot[slot] = off + j;
          else { slot = -1; atomicOr(ws.counters + 2, 2); }
        }
        ws.s_app[off + j] = slot;
      }
      base += __popc(b);
    }
    __syncwarp();
#pragma unroll
    for (int d = 16; d; d >>= 1) {
      acc += __shfl_xor_sync(FULL_MASK, acc, d);
      depth += __shfl_xor_sync(FULL_MASK, depth, d);
    }
    if (lane == 0 && have) {
      ws.acc[r] = acc;
      float last = __ldg(c.rays + r * c.ray_cols + (c.ray_cols - 1));     // rays_chunk[..., -1], :388
      c.depth_map[r] = depth + (1.f - acc) * last;
    }
  }
}

// rgb_map = sum_i w_i rgb_i (+ 1 - acc) -> clamp(0,1); one warp per ray over its packed samples.
__global__ void __launch_bounds__(INT_WARPS * 32)
composite_kernel(const FtrfField f, const FtrfCall c, const Ws ws) {
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int64_t stride = (int64_t)gridDim.x * INT_WARPS;
  for (int64_t r = (int64_t)blockIdx.x * INT_WARPS + warp; r < c.n_rays; r += stride) {
    const int off = ws.ray_off[r], cnt = ws.ray_cnt[r];
    float rgb[3] = {0.f, 0.f, 0.f};
    // four chunks in flight: all slot loads first, then the dependent rgb gathers, then the sums
    for (int j0 = lane; j0 < cnt; j0 += 128) {
      int a[4];
      float w[4], q[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + 32 * u;
        a[u] = -1; w[u] = 0.f;
        if (j < cnt) { a[u] = ws.s_app[off + j]; w[u] = ws.s_w[off + j]; }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        q[u][0] = q[u][1] = q[u][2] = 0.f;
        if (a[u] >= 0) {
          const float* p = ws.a_rgb + (size_t)a[u] * 3;
          q[u][0] = p[0]; q[u][1] = p[1]; q[u][2] = p[2];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (a[u] >= 0) {
          rgb[0] = fmaf(w[u], q[u][0], rgb[0]); rgb[1] = fmaf(w[u], q[u][1], rgb[1]); rgb[2] = fmaf(w[u], q[u][2], rgb[2]);
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1)
#pragma unroll
      for (int k = 0; k < 3; ++k) rgb[k] += __shfl_xor_sync(FULL_MASK, rgb[k], d);
    if (lane < 3) {
      float v = rgb[lane];
      if (c.add_bg) v = v + (1.f - ws.acc[r]);
      ws.rgb_pre[r * 3 + lane] = v;
      c.rgb_map[r * 3 + lane] = fminf(fmaxf(v, 0.f), 1.f);
    }
  }
}

// ------------------------------------------------------------------------------------ backward
// Inputs: g_rgb_map.  Outputs: a_grgb (dL/d rgb of every shaded sample) and s_gsig
// (dL/d sigma_feature of every packed sample).
//   g_w_i     = g_acc + [shaded] <g_map, rgb_i>,          g_acc = -sum_c g_map[c] if background added
//   g_alpha_i = T_i g_w_i - (sum_{j>i} g_w_j w_j) / (1 - alpha_i + 1e-10)
//   g_sigma_i = g_alpha_i * dist_i * (1 - alpha_i);  g_feat_i = g_sigma_i * act'(.)
__global__ void __launch_bounds__(INT_WARPS * 32)
integrate_bwd_kernel(const FtrfField f, const FtrfCall c, const Ws ws, const bool write_grgb) {
  const int warp = threadIdx.x >> 5, lane = lane_id();
  if (write_grgb && c.overflow_out && blockIdx.x == 0 && threadIdx.x == 0) *c.overflow_out = ws.counters[2] ? 1.f : 0.f;
  const int64_t stride = (int64_t)gridDim.x * INT_WARPS;
  for (int64_t r = (int64_t)blockIdx.x * INT_WARPS + warp; r < c.n_rays; r += stride) {
    const int off = ws.ray_off[r], cnt = ws.ray_cnt[r];
    if (cnt == 0) continue;
    Ray ray = load_ray_ws(f, c, ws, r);
    const float dnorm = ray_dnorm(c, ray);
    float gm[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float pre = ws.rgb_pre[r * 3 + k];
      gm[k] = (pre >= 0.f && pre <= 1.f) ? __ldg(c.g_rgb_map + r * 3 + k) : 0.f;   // clamp backward
    }
    const float g_acc = c.add_bg ? -(gm[0] + gm[1] + gm[2]) : 0.f;
    float suffix = 0.f;                     // sum over later samples of g_w * w
    const int last0 = ((cnt - 1) >> 5) << 5;
    // the loads of the NEXT chunk (list entries and the shaded sample's rgb) are issued before this chunk's scan
    struct Pre { int i, a; float sigma, T, w, q0, q1, q2; };
    auto fetch = [&](int j0) {
      Pre p{0, -1, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const int j = j0 + lane;
      if (j0 >= 0 && j < cnt) {
        p.i = ws.s_idx[off + j];
        p.sigma = ws.s_sigma[off + j];
        p.T = ws.s_T[off + j];
        p.w = ws.s_w[off + j];
        p.a = ws.s_app[off + j];
        if (p.a >= 0) {
          const float* q = ws.a_rgb + (size_t)p.a * 3;
          p.q0 = q[0]; p.q1 = q[1]; p.q2 = q[2];
        }
      }
      return p;
    };
    Pre nx = fetch(last0);
    for (int j0 = last0; j0 >= 0; j0 -= 32) {
      const int j = j0 + lane;
      const bool live = j < cnt;
      const Pre cur = nx;
      nx = fetch(j0 - 32);
      float gw = 0.f, w = 0.f, T = 0.f, alpha = 0.f, dist = 0.f, sigma = 0.f;
      if (live) {
        const int i = cur.i;
        float z = sample_z(f, c, ray, i);
        dist = sample_dist(f, c, ray, i, z, dnorm);
        sigma = cur.sigma;
        alpha = 1.f - expf(-sigma * dist);
        T = cur.T;
        w = cur.w;
        gw = g_acc;
        if (cur.a >= 0) {
          gw += gm[0] * cur.q0 + gm[1] * cur.q1 + gm[2] * cur.q2;
          if (write_grgb) {
            float* o = ws.a_grgb + (size_t)cur.a * 3;
            o[0] = w * gm[0]; o[1] = w * gm[1]; o[2] = w * gm[2];
          }
        }
      }
      // exclusive suffix sum of gw*w across lanes (higher lanes = later samples)
      float v = gw * w, incl = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        float dn = __shfl_down_sync(FULL_MASK, incl, d);
        if (lane + d < 32) incl += dn;
      }
      float after = (incl - v) + suffix;
      suffix += __shfl_sync(FULL_MASK, incl, 0);
      if (live) {
        float one_m = 1.f - alpha;
        float g_alpha = T * gw - after / (one_m + 1e-10f);
        float g_sigma = g_alpha * dist * one_m;
        float dact = (f.act == FTRF_ACT_SOFTPLUS) ? (sigma > 20.f ? 1.f : 1.f - expf(-sigma))
                                                  : (sigma > 0.f ? 1.f : 0.f);
        ws.s_gsig[off + j] = g_sigma * dact;
      }
    }
  }
}

static int ray_grid(int64_t n_rays) {
  int64_t blocks = (n_rays + INT_WARPS - 1) / INT_WARPS;
  int64_t cap = (int64_t)FTRF_SMS * 8;
  return (int)(blocks < cap ? blocks : cap);
}

int launch_integrate(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  const size_t smem = (size_t)INT_WARPS * ((c.n_samples + 31) / 32) * sizeof(uint32_t);
  if (smem > 200 * 1024) { set_error("integrate: n_samples %d too large", c.n_samples); return FTRF_E_ARG; }
  if (smem > 48 * 1024) cudaFuncSetAttribute(integrate_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  integrate_fwd_kernel<<<ray_grid(c.n_rays), INT_WARPS * 32, smem, st>>>(f, c, ws);
  return check_launch("integrate_fwd");
}
int launch_composite(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  composite_kernel<<<ray_grid(c.n_rays), INT_WARPS * 32, 0, st>>>(f, c, ws);
  return check_launch("composite");
}
// dL/d rgb of every shaded sample on its own: a_grgb[a] = w * (clamp-masked g_rgb_map of the sample's ray), the
// same values integrate_bwd_kernel writes.  The MLP backward only needs these, so with this kernel in front of it
// the reverse scan (whose sigma gradients nothing reads before the density backward) can run on a side stream.
__global__ void __launch_bounds__(256)
grgb_kernel(const FtrfCall c, const Ws ws) {
  const int64_t n = min((int64_t)ws.counters[1], c.cap_app);
  if (c.overflow_out && blockIdx.x == 0 && threadIdx.x == 0) *c.overflow_out = ws.counters[2] ? 1.f : 0.f;
  for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n; a += (int64_t)gridDim.x * blockDim.x) {
    const int slot = ws.a_slot[a];
    const int r = ws.s_ray[slot];
    const float w = ws.s_w[slot];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float pre = ws.rgb_pre[(int64_t)r * 3 + k];
      const float gm = (pre >= 0.f && pre <= 1.f) ? __ldg(c.g_rgb_map + (int64_t)r * 3 + k) : 0.f;   // clamp backward
      ws.a_grgb[a * 3 + k] = w * gm;
    }
  }
}

int launch_integrate_bwd(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  integrate_bwd_kernel<<<ray_grid(c.n_rays), INT_WARPS * 32, 0, st>>>(f, c, ws, true);
  return check_launch("integrate_bwd");
}

// the two halves of that kernel's job as separate launches (see grgb_kernel)
int launch_grgb(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0 || c.cap_app == 0) return FTRF_OK;
  grgb_kernel<<<FTRF_SMS * 4, 256, 0, st>>>(c, ws);
  return check_launch("grgb");
}
int launch_integrate_bwd_sigma(const FtrfField& f, const FtrfCall& c, const Ws& ws, cudaStream_t st) {
  if (c.n_rays == 0) return FTRF_OK;
  integrate_bwd_kernel<<<ray_grid(c.n_rays), INT_WARPS * 32, 0, st>>>(f, c, ws, false);
  return check_launch("integrate_bwd");
}

}  // namespace ftrf

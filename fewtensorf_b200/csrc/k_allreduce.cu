// k_allreduce.cu — all-reduce(sum) of the flat gradient buffer over NVSwitch multicast (NVLS), the one
// collective of a multi-GPU training step.  The buffer lives in symmetric memory (same
// offset on every GPU, one multicast address that maps all copies).  Rank r owns the r-th slice:
// `multimem.ld_reduce` pulls the element-wise fp32 sum of all ranks' copies out of the switch and
// `multimem.st` writes the result back to every copy, so each GPU moves ~N bytes in and ~N bytes out
// regardless of the number of ranks.  The caller brackets the launch with two cross-GPU barriers (all
// gradients written before; all slices written after) — fewtensorf_b200/dist.py uses the symmetric-memory
// signal pads for them.
#include "common.cuh"

namespace ftrf {

__global__ void __launch_bounds__(512)
multimem_allreduce_kernel(float* __restrict__ mc, size_t lo4, size_t hi4) {   // [lo4, hi4) in float4 units
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  constexpr int U = 4;                                                       // independent round trips in flight
  size_t i = lo4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < hi4; i += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float* p = mc + (i + u * stride) * 4;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(p) : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float* p = mc + (i + u * stride) * 4;
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                   ::"l"(p), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
    }
  }
  for (; i < hi4; i += stride) {
    float4 v;
    float* p = mc + i * 4;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
}

// Same exchange with plain peer loads / stores over NVLink (no switch reduction): rank r sums slice r of every
// GPU's copy in rank order and writes the result into every copy.  W loads in flight per element hide the
// NVLink round trip; per GPU (W-1)/W of the buffer travels in each direction.
struct PeerPtrs { float* p[16]; };

template <int W>
__global__ void __launch_bounds__(512)
p2p_allreduce_kernel(const PeerPtrs peers, size_t lo4, size_t hi4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = lo4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi4; i += stride) {
    float4 v[W];
#pragma unroll
    for (int r = 0; r < W; ++r)
      asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v[r].x), "=f"(v[r].y), "=f"(v[r].z), "=f"(v[r].w) : "l"(peers.p[r] + i * 4) : "memory");
    float4 a = v[0];
#pragma unroll
    for (int r = 1; r < W; ++r) { a.x += v[r].x; a.y += v[r].y; a.z += v[r].z; a.w += v[r].w; }
#pragma unroll
    for (int r = 0; r < W; ++r)
      asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                   ::"l"(peers.p[r] + i * 4), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
  }
}

// ---- the same two exchanges restricted to the parts of the buffer that can hold a gradient -------------------
// A plane's gradient is non-zero only inside the texel bounding box of the alphaMask (+ the interpolation ring):
// samples outside it are culled before the density gather, so nothing is ever reduced into the other rows and
// columns — exactly zero on every rank, nothing to exchange.  In the un-shrunk benchmark field that is 16 % of the
// planes.  A region is a 2-D run of float4 units inside the flat buffer: n_rows rows of row_len4 units, row_stride4
// apart, starting at off4 (lines and the dense basis / MLP tail are single rows).  The union of the regions is
// numbered row-major, region after region, and split into `world` equal shares like the dense kernels' slices.
struct RegionTable {
  int n;
  uint32_t first[FTRF_MAX_REGIONS + 1];        // virtual index of each region's first unit; first[n] = total
  uint32_t off4[FTRF_MAX_REGIONS], row_len4[FTRF_MAX_REGIONS], row_stride4[FTRF_MAX_REGIONS];
};

__device__ __forceinline__ size_t region_unit(const RegionTable& t, uint32_t v) {
  int r = 0;
#pragma unroll 1
  while (r + 1 < t.n && v >= t.first[r + 1]) ++r;
  const uint32_t e = v - t.first[r];
  const uint32_t row = e / t.row_len4[r], col = e - row * t.row_len4[r];
  return (size_t)t.off4[r] + (size_t)row * t.row_stride4[r] + col;
}

template <int W>
__global__ void __launch_bounds__(512)
p2p_allreduce_regions_kernel(const PeerPtrs peers, const RegionTable tab, uint32_t lo, uint32_t hi) {
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t v = lo + blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += stride) {
    const size_t i = region_unit(tab, v);
    float4 x[W];
#pragma unroll
    for (int r = 0; r < W; ++r)
      asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(x[r].x), "=f"(x[r].y), "=f"(x[r].z), "=f"(x[r].w) : "l"(peers.p[r] + i * 4) : "memory");
    float4 a = x[0];
#pragma unroll
    for (int r = 1; r < W; ++r) { a.x += x[r].x; a.y += x[r].y; a.z += x[r].z; a.w += x[r].w; }
#pragma unroll
    for (int r = 0; r < W; ++r)
      asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                   ::"l"(peers.p[r] + i * 4), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
  }
}

__global__ void __launch_bounds__(512)
multimem_allreduce_regions_kernel(float* __restrict__ mc, const RegionTable tab, uint32_t lo, uint32_t hi) {
  const uint32_t stride = gridDim.x * blockDim.x;
  constexpr int U = 4;
  uint32_t v = lo + blockIdx.x * blockDim.x + threadIdx.x;
  for (; v < hi; v += U * stride) {
    float4 x[U];
    float* p[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t vv = v + u * stride;
      p[u] = vv < hi ? mc + region_unit(tab, vv) * 4 : nullptr;
      if (p[u])
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x[u].x), "=f"(x[u].y), "=f"(x[u].z), "=f"(x[u].w) : "l"(p[u]) : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (p[u])
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(p[u]), "f"(x[u].x), "f"(x[u].y), "f"(x[u].z), "f"(x[u].w) : "memory");
  }
}

// ---- the same exchange with the two cross-GPU barriers INSIDE the kernel ------------------------------------
// flags[r] is rank r's copy of a small symmetric int32 array: [0, 16) arrival flags of barrier A (one slot per sender),
// [16, 32) of barrier B, [64] "go" word CTA 0 releases this GPU's other CTAs with, [65] finished-CTA counter.
// `epoch` grows by one per call on every rank.  CTA 0 (always the first scheduled) signals and awaits barrier A — every
// rank's backward has written its gradients — then lets the grid run; the last CTA to finish its share signals and
// awaits barrier B — every share is back in every copy — so the kernel's end is the collective's end.  Two launches
// and their gaps fewer than bracketing the exchange with separate barrier kernels.
struct FlagPtrs { uint32_t* p[16]; };

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int W, bool MULTIMEM>
__global__ void __launch_bounds__(512)
allreduce_regions_fused_kernel(const PeerPtrs peers, float* __restrict__ mc, const FlagPtrs flags, const RegionTable tab,
                               uint32_t lo, uint32_t hi, int rank, uint32_t epoch) {
  uint32_t* mine = flags.p[rank];
  if (blockIdx.x == 0) {
    if (threadIdx.x < W) {
      __threadfence_system();
      st_release_sys(flags.p[threadIdx.x] + rank, epoch);                 // barrier A: tell peer `tid` we are here
      while (ld_acquire_sys(mine + threadIdx.x) < epoch) {}               //            and wait for peer `tid`
    }
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(mine + 64), "r"(epoch) : "memory");
  } else {
    if (threadIdx.x == 0) while (ld_acquire_gpu(mine + 64) < epoch) {}
    __syncthreads();
  }
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t v = lo + blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += stride) {
    const size_t i = region_unit(tab, v);
    if (MULTIMEM) {
      float4 x;
      float* p = mc + i * 4;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "l"(p) : "memory");
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                   ::"l"(p), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
    } else {
      float4 x[W];
#pragma unroll
      for (int r = 0; r < W; ++r)
        asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(x[r].x), "=f"(x[r].y), "=f"(x[r].z), "=f"(x[r].w) : "l"(peers.p[r] + i * 4) : "memory");
      float4 a = x[0];
#pragma unroll
      for (int r = 1; r < W; ++r) { a.x += x[r].x; a.y += x[r].y; a.z += x[r].z; a.w += x[r].w; }
#pragma unroll
      for (int r = 0; r < W; ++r)
        asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(peers.p[r] + i * 4), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
    }
  }
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(mine + 65, 1u);
    last = done == gridDim.x - 1;
    if (last) mine[65] = 0;                                               // (next launch starts after this kernel ends)
  }
  __syncthreads();
  if (last && threadIdx.x < W) {
    __threadfence_system();
    st_release_sys(flags.p[threadIdx.x] + 16 + rank, epoch);              // barrier B
    while (ld_acquire_sys(mine + 16 + threadIdx.x) < epoch) {}
  }
}

static int make_region_table(const FtrfRegion* regions, int n, RegionTable* t, const char* who) {
  if (!regions || n < 1 || n > FTRF_MAX_REGIONS) { set_error("%s: 1..%d regions", who, FTRF_MAX_REGIONS); return FTRF_E_ARG; }
  uint64_t total = 0;
  t->n = n;
  for (int r = 0; r < n; ++r) {
    const FtrfRegion& g = regions[r];
    if (g.offset < 0 || g.row_len < 4 || g.n_rows < 1 || (g.offset & 3) || (g.row_len & 3) || (g.row_stride & 3) ||
        (g.n_rows > 1 && g.row_stride < g.row_len) || (uint64_t)g.offset + (uint64_t)(g.n_rows - 1) * (uint64_t)g.row_stride + (uint64_t)g.row_len > (1ull << 33)) {
      set_error("%s: region %d is not a run of whole float4 units", who, r);
      return FTRF_E_ARG;
    }
    t->first[r] = (uint32_t)total;
    t->off4[r] = (uint32_t)(g.offset / 4);
    t->row_len4[r] = (uint32_t)(g.row_len / 4);
    t->row_stride4[r] = (uint32_t)(g.row_stride / 4);
    total += (uint64_t)(g.row_len / 4) * (uint64_t)g.n_rows;
    if (total >= (1ull << 32)) { set_error("%s: regions too large", who); return FTRF_E_ARG; }
  }
  t->first[n] = (uint32_t)total;
  return FTRF_OK;
}

}  // namespace ftrf

extern "C" int ftrf_p2p_allreduce_regions(void* const* peer_ptrs, const FtrfRegion* regions, int n_regions, int rank, int world,
                                          void* stream) {
  if (!peer_ptrs || (world != 2 && world != 4 && world != 8) || rank < 0 || rank >= world) {
    ftrf::set_error("p2p_allreduce_regions: bad argument (world must be 2, 4 or 8)");
    return FTRF_E_ARG;
  }
  ftrf::RegionTable tab;
  if (int e = ftrf::make_region_table(regions, n_regions, &tab, "p2p_allreduce_regions")) return e;
  ftrf::PeerPtrs pp{};
  for (int r = 0; r < world; ++r) {
    if (!peer_ptrs[r] || (((size_t)peer_ptrs[r]) & 15)) { ftrf::set_error("p2p_allreduce_regions: bad peer pointer %d", r); return FTRF_E_ARG; }
    pp.p[r] = (float*)peer_ptrs[r];
  }
  const uint32_t n4 = tab.first[tab.n], per = (n4 + world - 1) / world;
  const uint32_t lo = (uint64_t)per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
  if (hi > lo) {
    const uint32_t want = (hi - lo + 511) / 512;
    const int grid = (int)(want < (uint32_t)FTRF_SMS * 4 ? want : (uint32_t)FTRF_SMS * 4);
    cudaStream_t st = (cudaStream_t)stream;
    if (world == 2) ftrf::p2p_allreduce_regions_kernel<2><<<grid, 512, 0, st>>>(pp, tab, lo, hi);
    else if (world == 4) ftrf::p2p_allreduce_regions_kernel<4><<<grid, 512, 0, st>>>(pp, tab, lo, hi);
    else ftrf::p2p_allreduce_regions_kernel<8><<<grid, 512, 0, st>>>(pp, tab, lo, hi);
  }
  return ftrf::check_launch("p2p_allreduce_regions");
}

extern "C" int ftrf_allreduce_regions_fused(void* const* peer_ptrs, void* multicast_ptr, void* const* flag_ptrs,
                                            const FtrfRegion* regions, int n_regions, int rank, int world, uint32_t epoch,
                                            void* stream) {
  if (!peer_ptrs || !flag_ptrs || (world != 2 && world != 4 && world != 8) || rank < 0 || rank >= world || epoch == 0) {
    ftrf::set_error("allreduce_regions_fused: bad argument (world must be 2, 4 or 8)");
    return FTRF_E_ARG;
  }
  ftrf::RegionTable tab;
  if (int e = ftrf::make_region_table(regions, n_regions, &tab, "allreduce_regions_fused")) return e;
  ftrf::PeerPtrs pp{};
  ftrf::FlagPtrs fp{};
  for (int r = 0; r < world; ++r) {
    if (!peer_ptrs[r] || (((size_t)peer_ptrs[r]) & 15) || !flag_ptrs[r]) { ftrf::set_error("allreduce_regions_fused: bad peer pointer %d", r); return FTRF_E_ARG; }
    pp.p[r] = (float*)peer_ptrs[r];
    fp.p[r] = (uint32_t*)flag_ptrs[r];
  }
  const uint32_t n4 = tab.first[tab.n], per = (n4 + world - 1) / world;
  const uint32_t lo = (uint64_t)per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
  const uint32_t want = hi > lo ? (hi - lo + 511) / 512 : 1;
  const int grid = (int)(want < (uint32_t)FTRF_SMS * 2 ? want : (uint32_t)FTRF_SMS * 2);
  cudaStream_t st = (cudaStream_t)stream;
  float* mc = (float*)multicast_ptr;
  if (mc) {
    if (world == 2) ftrf::allreduce_regions_fused_kernel<2, true><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
    else if (world == 4) ftrf::allreduce_regions_fused_kernel<4, true><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
    else ftrf::allreduce_regions_fused_kernel<8, true><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
  } else {
    if (world == 2) ftrf::allreduce_regions_fused_kernel<2, false><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
    else if (world == 4) ftrf::allreduce_regions_fused_kernel<4, false><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
    else ftrf::allreduce_regions_fused_kernel<8, false><<<grid, 512, 0, st>>>(pp, mc, fp, tab, lo, hi, rank, epoch);
  }
  return ftrf::check_launch("allreduce_regions_fused");
}

extern "C" int ftrf_multimem_allreduce_regions(void* multicast_ptr, const FtrfRegion* regions, int n_regions, int rank, int world,
                                               void* stream) {
  if (!multicast_ptr || world < 1 || rank < 0 || rank >= world || (((size_t)multicast_ptr) & 15)) {
    ftrf::set_error("multimem_allreduce_regions: bad argument");
    return FTRF_E_ARG;
  }
  ftrf::RegionTable tab;
  if (int e = ftrf::make_region_table(regions, n_regions, &tab, "multimem_allreduce_regions")) return e;
  const uint32_t n4 = tab.first[tab.n], per = (n4 + world - 1) / world;
  const uint32_t lo = (uint64_t)per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
  if (hi > lo) {
    const uint32_t want = (hi - lo + 512 * 4 - 1) / (512 * 4);
    const int grid = (int)(want < (uint32_t)FTRF_SMS * 4 ? want : (uint32_t)FTRF_SMS * 4);
    ftrf::multimem_allreduce_regions_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>((float*)multicast_ptr, tab, lo, hi);
  }
  return ftrf::check_launch("multimem_allreduce_regions");
}

// peer_ptrs: `world` device pointers (host array) to the same symmetric buffer on every GPU, in rank order.
extern "C" int ftrf_p2p_allreduce(void* const* peer_ptrs, size_t n_floats, int rank, int world, void* stream) {
  if (!peer_ptrs || (world != 2 && world != 4 && world != 8) || rank < 0 || rank >= world || (n_floats & 3)) {
    ftrf::set_error("p2p_allreduce: bad argument (world must be 2, 4 or 8)");
    return FTRF_E_ARG;
  }
  ftrf::PeerPtrs pp{};
  for (int r = 0; r < world; ++r) {
    if (!peer_ptrs[r] || (((size_t)peer_ptrs[r]) & 15)) { ftrf::set_error("p2p_allreduce: bad peer pointer %d", r); return FTRF_E_ARG; }
    pp.p[r] = (float*)peer_ptrs[r];
  }
  const size_t n4 = n_floats / 4, per = (n4 + world - 1) / world;
  const size_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
  if (hi > lo) {
    const size_t want = (hi - lo + 511) / 512;
    const int grid = (int)(want < (size_t)FTRF_SMS * 4 ? want : (size_t)FTRF_SMS * 4);
    cudaStream_t st = (cudaStream_t)stream;
    if (world == 2) ftrf::p2p_allreduce_kernel<2><<<grid, 512, 0, st>>>(pp, lo, hi);
    else if (world == 4) ftrf::p2p_allreduce_kernel<4><<<grid, 512, 0, st>>>(pp, lo, hi);
    else ftrf::p2p_allreduce_kernel<8><<<grid, 512, 0, st>>>(pp, lo, hi);
  }
  return ftrf::check_launch("p2p_allreduce");
}

// multicast_ptr: multicast address of the symmetric buffer; n_floats: multiple of 4; this rank reduces and
// broadcasts slice `rank` of `world` equal slices.
extern "C" int ftrf_multimem_allreduce(void* multicast_ptr, size_t n_floats, int rank, int world, void* stream) {
  if (!multicast_ptr || world < 1 || rank < 0 || rank >= world || (n_floats & 3) || (((size_t)multicast_ptr) & 15)) {
    ftrf::set_error("multimem_allreduce: bad argument");
    return FTRF_E_ARG;
  }
  const size_t n4 = n_floats / 4, per = (n4 + world - 1) / world;
  const size_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
  if (hi > lo) {
    const size_t want = (hi - lo + 512 * 4 - 1) / (512 * 4);
    const int grid = (int)(want < (size_t)FTRF_SMS * 4 ? want : (size_t)FTRF_SMS * 4);
    ftrf::multimem_allreduce_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>((float*)multicast_ptr, lo, hi);
  }
  return ftrf::check_launch("multimem_allreduce");
}

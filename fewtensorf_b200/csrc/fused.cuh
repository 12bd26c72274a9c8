// fused.cuh — shared definitions of the fused shading-MLP kernels (k_shade_fused.cu) and the kernels
// that consume what they save for backward.
//
// INTERNAL COLUMN ORDER of the MLP input.  The first layer is linear, so the columns of X (and of W1)
// may be permuted freely.  The fused kernels use an order in which sin(a) and cos(a) of the same
// argument are neighbours, so one sincosf() yields two columns and every 16-byte unit of the operand
// tile is complete when it is written:
//     [ feat(0..D) | view(0..3) | (pad to even) | block0 pairs | block1 pairs | ... | zero pad to 8 ]
// pair q = ch*F + f of a PE block holds (sin, cos) of src[ch] * 2^f  (models/mlp.py:8-13 flattens
// channel-major / frequency-minor, [sin | cos]).  `orig[]` maps an internal column to the reference's
// column (mlp.py:41-64 / :88-101 / :127-149 concat order), -1 for padding.
#pragma once
#include "shade_common.cuh"

namespace ftrf {

constexpr int FZ_MAX_K = 192;        // widest internal MLP input supported by the fused path
constexpr int FZ_MAX_SRC = 40;       // D + 6 staged source values per sample

enum { PK_ZERO = 0, PK_COPY2 = 1, PK_COPY1 = 2, PK_TRIG = 3 };

struct PairRec {          // two neighbouring internal columns
  uint8_t kind;           // PK_*
  uint8_t sa, sb;         // source slots ([feat | view | pos] staged row); TRIG uses sa only
  uint8_t fexp;           // TRIG: argument scale 2^fexp
  uint8_t msrc;           // TRIG: encoding-mask table (SRC_*)
  uint8_t pad_;
  int16_t ma, mb;         // TRIG: index of the sin / cos column inside that mask vector
};

struct FusedSpec {
  int kp;                 // internal width, multiple of 8
  int n1;                 // 32-column chunks of layer 1
  int nsrc;               // D + 6
  int D;
  int need_pos;
  int blk_icol0[3];       // internal column of pair 0 of each PE block (InSpec block order)
  int16_t orig[FZ_MAX_K];
  PairRec pair[FZ_MAX_K / 2];
};

// Returns false when the fused path does not cover this configuration (caller falls back).
inline bool make_fused_spec(const FtrfField& f, const InSpec& sp, FusedSpec& fs) {
  fs = FusedSpec{};
  const int D = sp.D;
  if (D + 6 > FZ_MAX_SRC || f.feature_c != FC) return false;
  int col = 0;
  auto push = [&](PairRec r, int o0, int o1) {
    if (col + 2 > FZ_MAX_K) { col += 2; return; }
    fs.pair[col / 2] = r;
    fs.orig[col] = (int16_t)o0;
    fs.orig[col + 1] = (int16_t)o1;
    col += 2;
  };
  const int ncopy = D + 3;
  for (int i = 0; i + 1 < ncopy; i += 2) push(PairRec{PK_COPY2, (uint8_t)i, (uint8_t)(i + 1), 0, 0, 0, 0, 0}, i, i + 1);
  if (ncopy & 1) push(PairRec{PK_COPY1, (uint8_t)(ncopy - 1), 0, 0, 0, 0, 0, 0}, ncopy - 1, -1);
  for (int b = 0; b < sp.nblk; ++b) {
    const PeBlock& pb = sp.b[b];
    const int half = pb.F * pb.C;
    const int s0 = pb.src == SRC_FEA ? 0 : (pb.src == SRC_VIEW ? D : D + 3);
    fs.blk_icol0[b] = col;
    for (int ch = 0; ch < pb.C; ++ch)
      for (int fq = 0; fq < pb.F; ++fq) {
        int q = ch * pb.F + fq;
        push(PairRec{PK_TRIG, (uint8_t)(s0 + ch), 0, (uint8_t)fq, (uint8_t)pb.src, 0, (int16_t)q, (int16_t)(half + q)},
             pb.col0 + q, pb.col0 + half + q);
      }
  }
  if (col > FZ_MAX_K) return false;
  while (col & 7) push(PairRec{PK_ZERO, 0, 0, 0, 0, 0, 0, 0}, -1, -1);
  fs.kp = col;
  fs.n1 = (col + 31) / 32;
  fs.nsrc = D + 6;
  fs.D = D;
  fs.need_pos = sp.need_pos;
  return true;
}

// Weight images in the workspace (FTRF_WS_WIMG), all pre-split into tf32 hi / lo and laid out as
// UMMA K-major no-swizzle [128 rows x 32 k] chunks (hi 16 KB then lo 16 KB), ready for bulk copies:
//   forward : n1 chunks of W1 (rows = hidden unit, k = internal input column), then 4 chunks of W2.
constexpr size_t FZ_CHUNK_BYTES = 2u * 128u * 32u * 4u;     // 32 KB
inline size_t fused_w1p_offset_floats() { return (size_t)(FZ_MAX_K / 32 + 4) * (FZ_CHUNK_BYTES / 4); }
inline size_t fused_scratch_offset_floats() { return fused_w1p_offset_floats() + (size_t)FC * FZ_MAX_K; }
inline size_t fused_bwd_offset_floats() { return fused_scratch_offset_floats() + (size_t)128 * 48 * 4; }
inline size_t fused_wimg_bytes() { return (fused_bwd_offset_floats() + (size_t)4 * 8192 + (size_t)4 * 2 * 160 * 32) * 4; }

int launch_shade_fused(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t, bool* handled);
int launch_shade_dgrad_fused(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t, bool* handled);
bool shade_path_is_fused();
bool shade_dgrad_is_fused(const FtrfField&, const FtrfCall&, const Ws&);   // launch_shade_dgrad_fused will handle this call
bool prep_ahead_take_scratch(const FtrfCall&);        // true: this call's forward already zeroed the wgrad scratch
size_t wgrad_fused_scratch_bytes();
cudaStream_t side_tail_begin(const FtrfCall&, cudaStream_t st);   // side stream ordered after `st`, or nullptr
void side_tail_end(const FtrfCall&);
cudaStream_t side_scan_begin(const FtrfCall&, cudaStream_t st);   // the same pair for the integration backward's reverse scan
void side_scan_end(const FtrfCall&);
int side_scan_join(const FtrfCall&, cudaStream_t st);
int side_tail_join(const FtrfCall&, cudaStream_t st);             // `st` waits for the tail (no-op when none is pending)
int shade_prep_ahead(const FtrfField&, const FtrfCall&, const Ws&, cudaStream_t, int phase);   // weight images on a side stream

// argument block of the fused weight-gradient kernel (k_wgrad_fused.cu)
// Layout of the activations the fused forward / dgrad kernels save for the weight-gradient kernel (X, h1, h2, g_h1) when
// ws.counters[WS_ACT_LAYOUT] == 1: blocks of 16 rows; inside a block the 8-column groups are outermost and every group
// (16 rows x 8 floats = 512 bytes) is followed by 32 bytes of padding —
//     (row, col) of a W-column array  ->  (row / 16) (W / 8) 136  +  (col / 8) 136  +  (row % 16) 8  +  col % 8
// A thread of those kernels owns a sample row, so with plain row-major arrays a warp's 32-byte store touched 32 different
// lines; here the 16 rows of a block are contiguous per 8-column group: 4x fewer store wavefronts (10 us of the forward, 5
// of the dgrad kernel).  The weight-gradient kernel consumes whole blocks with ONE bulk copy per array and chunk, and the
// padding — which travels with the copy — puts the four groups a warp reads from shared memory into different banks.
// 0 = row-major (the per-layer GEMM path, and every array when the fused dgrad kernel does not cover the field).
constexpr int WS_ACT_LAYOUT = 6;          // index in ws.counters
constexpr int ACT16_G8 = 16 * 8 + 8;      // floats of one 8-column group of a block
__host__ __device__ __forceinline__ size_t act16_block_floats(int W) { return (size_t)(W / 8) * ACT16_G8; }
__host__ __device__ __forceinline__ size_t act16_off(int64_t row, int col, int W) {
  return (size_t)(row >> 4) * act16_block_floats(W) + (size_t)(col >> 3) * ACT16_G8 + (size_t)(((int)(row & 15) << 3) + (col & 7));
}

struct WgradFusedArgs {
  const float* g2; const float* h1; const float* g1;   // [M][128] row-major
  const float* x; int ldx; int kp;                     // [M][ldx], kp <= 152 data columns (multiple of 4)
  const int32_t* count; int64_t cap;
  float* dw2;                                          // [128][128]  +=
  float* dw1i;                                         // [128][ngt*4] zeroed scratch, internal column order; col kp = db1
  int ngt;                                             // X operand groups (N of the dW1 product = 4 ngt)
  int dw2_vec;                                         // dw2 is 16-byte aligned
  int variant;                                         // bring-up switches
  // fold3 != 0: `g2` holds h2 (the layer-2 activations) and `go4` [M] = (dL/d pre-sigmoid output, 0) per sample as the
  // fused dgrad kernel left it; g_h2 = relu'(h2) * (go W3) is rebuilt by the producers, which also accumulate the
  // output layer's gradients dW3 = go^T h2, db3 = sum go and db2 = sum g_h2 (flushed with atomics at the end)
  int fold3;
  const float4* go4; const float* w3;
  float* dw3; float* db3; float* db2;
};
// `call`: the FtrfCall the hand-off flags live in (nullptr for the unit-test entry: everything in line)
int launch_wgrad_fused(const WgradFusedArgs&, const FusedSpec&, float* g_w1, int ldw, float* g_b1, cudaStream_t, const FtrfCall* call);

}  // namespace ftrf

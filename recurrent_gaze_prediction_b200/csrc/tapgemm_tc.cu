// tapgemm_tc.cu — the tcgen05 / TMEM / TMA engine of the tap GEMM (bf16 operands, fp32 accumulate).
//
// Persistent, warp-specialised kernel, one CTA per SM:
//   warp 0      : TMA producer.  Per k-step it issues two 2-D tile loads (SWIZZLE_128B): a 128 x 64 box of
//                 A at row (m0 + tap_off[tap]) — a conv tap is only a row shift in the p-layout — and a
//                 BN x 64 box of the tap's weight matrix.  4..6-stage mbarrier ring.
//   warp 1      : TMEM allocator + MMA issuer: one elected thread issues tcgen05.mma (M=128, N=BN, K=16)
//                 x4 per stage, accumulating all taps x K/64 chunks into one TMEM accumulator; tcgen05.commit
//                 releases the smem stage and finally signals the epilogue.  Two accumulator stages in TMEM
//                 so the epilogue of tile i overlaps the main loop of tile i+1.
//   warps 2..5  : epilogue.  tcgen05.ld 32 lanes x 32 columns -> registers -> fused epilogue (common.cuh).
#include <map>
#include <vector>
#include <tuple>
#include <algorithm>
#include <mutex>
#include <tuple>
#include "tc_common.cuh"

namespace rgp {

namespace {

constexpr int BM = 128, BK = 64;
constexpr int NTHREADS = 192;

using namespace tc;

// single-MUFU gates for the bf16 engine (rel. error ~2^-11, below the bf16 operand rounding of 2^-9;
// the fp32 engine keeps the libm forms)
__device__ __forceinline__ float tanh_fast(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }

// ---- vectorised row-chunk epilogue helpers (one thread = one output row, 32 consecutive columns) ----
// 256-bit accesses (sm_100: LDG/STG.E.ENL2.256): a thread-per-row warp access touches 32 different lines per instruction,
// so an epilogue array costs its instruction count x 32 L1 tag lookups - 256-bit words halve that against float4
// (measured in the persistent kernel: c-epilogue 20 -> 14 us).  The `vec` condition of the callers guarantees 32-byte
// aligned row starts.
__device__ __forceinline__ void ld8(const float* p, float* v) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p)
               : "memory");
}
__device__ __forceinline__ void st8(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
               "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
               : "memory");
}
__device__ __forceinline__ void ldf32(const float* p, float v[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) ld8(p + 8 * i, v + 8 * i);
}
__device__ __forceinline__ void stf32(float* p, const float v[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) st8(p + 8 * i, v + 8 * i);
}
__device__ __forceinline__ void stbf32(__nv_bfloat16* p, const float v[32]) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    uint32_t u[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      __nv_bfloat162 b = __floats2bfloat162_rn(v[16 * i + 2 * j], v[16 * i + 2 * j + 1]);
      u[j] = *reinterpret_cast<uint32_t*>(&b);
    }
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p + 16 * i), "r"(u[0]), "r"(u[1]), "r"(u[2]),
                 "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7])
                 : "memory");
  }
}

// 32 consecutive columns [n0, n0+32) of row r.  Fast paths need N % 32 == 0 and 32-element-aligned leading dims.
__device__ __forceinline__ void epi_row32(const rgp_tapgemm_args& a, const RowInfo& r, int n0, float v[32], bool vec) {
  using TA = __nv_bfloat16;
  if (!r.inrange) return;
  if (vec) {
    switch (a.epi) {
      case RGP_EPI_STORE_F32:
        if (!a.accumulate) { stf32(a.out_f32 + (long long)r.m * a.ldo + n0, v); return; }
        break;
      case RGP_EPI_PROJ: {
        float b[32];
        ldf32(a.bias + n0, b);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += b[i];
        stbf32((TA*)a.out_op + r.orow * a.ldo + n0, v);
        return;
      }
      case RGP_EPI_GRU_ZR: {
        float x[32];
        ldf32(a.xg + (long long)r.m * a.ldx + n0, x);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = sigmoid_fast(v[i] + x[i]);
        if (n0 < a.Ch) {
          stf32(a.z_f32 + (long long)r.m * a.Ch + n0, v);
        } else {
          int j = n0 - a.Ch;
          if (a.save_r) stf32(a.save_r + (long long)r.m * a.Ch + j, v);
          if (r.valid) {
            ldf32(a.hprev_f32 + r.prow * a.Ch + j, x);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] *= x[i];
            stbf32((TA*)a.out_op + r.prow * a.Ch + j, v);
          }
        }
        return;
      }
      case RGP_EPI_GRU_C: {
        float x[32];
        ldf32(a.xg + (long long)r.m * a.ldx + 2 * a.Ch + n0, x);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = tanh_fast(v[i] + x[i]);
        if (a.save_c) stf32(a.save_c + (long long)r.m * a.Ch + n0, v);
        if (r.valid) {
          float z[32];
          ldf32(a.z_f32 + (long long)r.m * a.Ch + n0, z);
          ldf32(a.hprev_f32 + r.prow * a.Ch + n0, x);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = z[i] * x[i] + (1.0f - z[i]) * v[i];
          stf32(a.hnext_f32 + r.prow * a.Ch + n0, v);
          stbf32((TA*)a.out_op + r.prow * a.Ch + n0, v);
          if (a.out_op2) {
            if (a.bias) {
              ldf32(a.bias + n0, x);
              ldf32(a.bias2 + n0, z);
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = x[i] * (v[i] * a.alpha) + z[i];
            }
            stbf32((TA*)a.out_op2 + r.orow * a.Ch + n0, v);
          }
        }
        return;
      }
      case RGP_EPI_BWD_RH: {   // acc = dL/d(r.h): dpre_r and the direct part of dL/dh_{t-1}  (common.cuh epi_elem)
        if (!r.valid) return;
        float rr[32], x[32];
        ldf32(a.save_r + (long long)r.m * a.Ch + n0, rr);
        ldf32(a.hprev_f32 + r.prow * a.Ch + n0, x);
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = v[i] * x[i] * rr[i] * (1.0f - rr[i]);
        stbf32((TA*)a.dg + (long long)r.m * a.ldg + a.Ch + n0, x);
        float* dp = a.dhpart + (long long)r.m * a.Ch + n0;
        ldf32(dp, x);
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = fmaf(v[i], rr[i], x[i]);
        stf32(dp, x);
        return;
      }
      case RGP_EPI_BWD_H: {    // acc = convT(dpre_z, dpre_r) of step t: full dL/dh_{t-1}, then the gates of step t-1
        if (!r.valid) return;
        float x[32], y[32];
        ldf32(a.dhs + (long long)r.m * a.Ch + n0, x);
        if (a.bias) {
          ldf32(a.bias + n0, y);
#pragma unroll
          for (int i = 0; i < 32; ++i) x[i] *= y[i] * a.alpha;
        }
        float* dp = a.dhpart + (long long)r.m * a.Ch + n0;
        ldf32(dp, y);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += x[i] + y[i];                  // dh
        ldf32(a.save_z + (long long)r.m * a.Ch + n0, x);                   // z
#pragma unroll
        for (int i = 0; i < 32; ++i) y[i] = v[i] * x[i];                   // dh * z -> dhpart
        stf32(dp, y);
        float c[32];
        ldf32(a.save_c + (long long)r.m * a.Ch + n0, c);
        ldf32(a.hprev_f32 + r.prow * a.Ch + n0, y);                        // h_{t-2}
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float dz = v[i] * (y[i] - c[i]), dc = v[i] * (1.0f - x[i]);
          y[i] = dz * x[i] * (1.0f - x[i]);
          c[i] = dc * (1.0f - c[i] * c[i]);
        }
        stbf32((TA*)a.dg + (long long)r.m * a.ldg + n0, y);
        stbf32((TA*)a.dg + (long long)r.m * a.ldg + 2 * a.Ch + n0, c);
        return;
      }
      case RGP_EPI_DHS:
        if (r.valid) stf32(a.out_f32 + r.orow * a.ldo + n0, v);
        return;
      case RGP_EPI_DX_DENSE:
        if (r.valid) stbf32((TA*)a.out_op + r.orow * a.ldo + n0, v);
        return;
      case RGP_EPI_STORE_OP:
        stbf32((TA*)a.out_op + (long long)r.m * a.ldo + n0, v);
        return;
      case RGP_EPI_UP7: {      // 32 consecutive channels of one pixel (P % 32 == 0): one 64-byte store
        if (a.P % 32 != 0) break;
        const int pix = n0 / a.P;
        if (pix >= 49) return;
        const int co = n0 - pix * a.P;
        const int ry = pix / 7, rx = pix - ry * 7;
        const int Y = r.hy + ry, X = r.hx + rx;
        if (Y >= 2 && Y < 51 && X >= 2 && X < 51)
          stbf32((TA*)a.out_op + (r.obase + (long long)Y * a.Wo + X) * a.ldo + co, v);
        return;
      }
      case RGP_EPI_STORE_OP_P:
        if (r.valid) stbf32((TA*)a.out_op + r.prow * a.ldo + n0, v);
        return;
      default:
        break;
    }
  }
#pragma unroll 4
  for (int i = 0; i < 32; ++i) epi_elem<TA>(a, r, n0 + i, v[i]);
}


// ---- coalesced quad epilogue: after a transpose through shared memory a lane owns 4 consecutive columns of
// a row, so that one warp-wide 16-byte access touches 4 rows x 128 bytes (4 cache lines instead of 32) ----
struct RowLite { int m; int flags; long long prow; long long orow; };   // flags: 1 = in range, 2 = valid position


__device__ __forceinline__ void st_bf16x4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 b0 = __floats2bfloat162_rn(v.x, v.y), b1 = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&b0); u.y = *reinterpret_cast<uint32_t*>(&b1);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__device__ __forceinline__ bool quad_mode(int epi, int accumulate) {
  // measured on B200 (profiles/r01_notes.md): the transposed quad path wins for the pure-store epilogues; the
  // gate epilogues are faster thread-per-row (8 independent 16-byte loads per array in flight per thread)
  return (epi == RGP_EPI_STORE_F32 && !accumulate) || epi == RGP_EPI_PROJ;
}

// The quad epilogue is split in two so that the global loads of several rows are all in flight before the
// first dependent instruction (the stores of one row may alias the loads of the next as far as the compiler
// knows, which would otherwise serialise one L2 round trip per row).
struct QuadIn { float4 x, z, h; };

__device__ __forceinline__ void quad_load(const rgp_tapgemm_args& a, const RowLite& q, int n, QuadIn& in) {
  const bool valid = (q.flags & 2) != 0;
  switch (a.epi) {
    case RGP_EPI_GRU_ZR:
      in.x = ld4(a.xg + (long long)q.m * a.ldx + n);
      if (n >= a.Ch && valid) in.h = ld4(a.hprev_f32 + q.prow * a.Ch + (n - a.Ch));
      break;
    case RGP_EPI_GRU_C:
      in.x = ld4(a.xg + (long long)q.m * a.ldx + 2 * a.Ch + n);
      if (valid) {
        in.z = ld4(a.z_f32 + (long long)q.m * a.Ch + n);
        in.h = ld4(a.hprev_f32 + q.prow * a.Ch + n);
      }
      break;
    default:
      break;
  }
}

__device__ __forceinline__ void quad_store(const rgp_tapgemm_args& a, const RowLite& q, int n, float4 v,
                                           const QuadIn& in, const float4& g, const float4& b) {
  using TA = __nv_bfloat16;
  const bool valid = (q.flags & 2) != 0;
  switch (a.epi) {
    case RGP_EPI_STORE_F32:
      st4(a.out_f32 + (long long)q.m * a.ldo + n, v);
      break;
    case RGP_EPI_PROJ:   // g = bias[n..n+3]
      v.x += g.x; v.y += g.y; v.z += g.z; v.w += g.w;
      st_bf16x4((TA*)a.out_op + q.orow * a.ldo + n, v);
      break;
    case RGP_EPI_GRU_ZR: {
      v.x = sigmoid_fast(v.x + in.x.x); v.y = sigmoid_fast(v.y + in.x.y);
      v.z = sigmoid_fast(v.z + in.x.z); v.w = sigmoid_fast(v.w + in.x.w);
      if (n < a.Ch) {
        st4(a.z_f32 + (long long)q.m * a.Ch + n, v);
      } else {
        const int j = n - a.Ch;
        if (a.save_r) st4(a.save_r + (long long)q.m * a.Ch + j, v);
        if (valid) {
          v.x *= in.h.x; v.y *= in.h.y; v.z *= in.h.z; v.w *= in.h.w;
          st_bf16x4((TA*)a.out_op + q.prow * a.Ch + j, v);
        }
      }
      break;
    }
    case RGP_EPI_GRU_C: {
      v.x = tanh_fast(v.x + in.x.x); v.y = tanh_fast(v.y + in.x.y);
      v.z = tanh_fast(v.z + in.x.z); v.w = tanh_fast(v.w + in.x.w);
      if (a.save_c) st4(a.save_c + (long long)q.m * a.Ch + n, v);
      if (valid) {
        const float4 z = in.z, h = in.h;
        v.x = z.x * h.x + (1.0f - z.x) * v.x; v.y = z.y * h.y + (1.0f - z.y) * v.y;
        v.z = z.z * h.z + (1.0f - z.z) * v.z; v.w = z.w * h.w + (1.0f - z.w) * v.w;
        st4(a.hnext_f32 + q.prow * a.Ch + n, v);
        st_bf16x4((TA*)a.out_op + q.prow * a.Ch + n, v);
        if (a.out_op2) {
          if (a.bias) {
            v.x = g.x * (v.x * a.alpha) + b.x; v.y = g.y * (v.y * a.alpha) + b.y;
            v.z = g.z * (v.z * a.alpha) + b.z; v.w = g.w * (v.w * a.alpha) + b.w;
          }
          st_bf16x4((TA*)a.out_op2 + q.orow * a.Ch + n, v);
        }
      }
      break;
    }
    default:
      break;
  }
}

constexpr int kScrLd = 36;                          // floats per scratch row (32 + 4 pad: conflict-free 16-byte accesses)
constexpr int kEpiScratchBytes = 4 * 32 * kScrLd * 4 + 4 * 32 * (int)sizeof(RowLite);

template <int BN>
struct Cfg {
  static constexpr int kStageBytes = BM * BK * 2 + BN * BK * 2;
  static constexpr int kStages = (BN >= 256) ? 4 : ((BN >= 128) ? 6 : 8);
  static constexpr int kAccStride = (BN <= 32) ? 32 : ((BN <= 64) ? 64 : ((BN <= 128) ? 128 : 256));
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + kEpiScratchBytes;
};

// Resident-A-slab mode (SLAB = true).  Every tap of a tap GEMM reads the SAME rows shifted by tap_off, so per
// (tile, k-chunk) the rows [m0 + min_off, m0 + max_off + 128) are fetched ONCE as one TMA box ("slab", <= 256 rows x
// 128 bytes, SWIZZLE_128B) and the MMAs of all taps take their A operand from inside it: descriptor start address =
// slab + (tap_off - min_off) * 128.  A start row that is not a multiple of the 8-row swizzle atom is legal because the
// hardware applies the 128-byte swizzle to absolute shared-memory address bits (probed on B200: csrc/debug_umma.cu,
// tools/umma_rowoff_probe.py - exact results for every offset with the descriptor's base-offset field left 0).
// Only the per-tap weight boxes still stream: L2->SMEM traffic per k-chunk drops from ntaps*(16 KB + W) to
// slab + ntaps*W (i2h: 432 -> 307 KB, head: 352 -> 117 KB).
struct SlabCfg {
  int rows;      // slab rows (multiple of 8 * nbox)
  int nbox;      // TMA boxes per slab (a box is <= 256 rows): rows / nbox rows each, stacked in shared memory
  int boxrows;
  int bytes;     // slab slot bytes (rows * 128 rounded up to 1024)
  int min_off;   // smallest tap offset
  int nw;        // weight-ring stages
  int slots;     // slab slots (k-chunks of A in flight)
  int tg;        // taps per weight stage (narrow tiles: one TMA box of tg*BN rows and ONE barrier round trip per tg taps)
  int pm_rot;    // position-major conv mode: rotation of the position order per frame group (load balance)
};

// Position-major tiles (im2col == 3: 3x3 SAME conv forward, 4: its transpose, 5: the composed 4x4-tap head GEMM): an M
// tile is ONE position (y, x) of 128 consecutive frames, so the taps that fall outside the frame are outside it for every
// row of the tile and are skipped - (3G-2)^2 of 9 G^2 tap-tiles for the convs (0.82 at 7x7, 0.91 at 14x14), 27^2 = 729 of
// 1600 row-taps per frame for the head.  Tile order: frame group, position (rotated by pm_rot per group so that the
// cheap corner / edge tiles spread over the CTAs), column tile.  The source row of tap (ky, kx) is (y + oy) * Wp + x + ox
// inside the frame's S rows; it is real data iff lo <= y + oy, x + ox < lo + n:
//   3: p-layout operand,  (oy, ox) = (ky, kx),         lo = 1, n = G     4: q-layout gradient, (1 - ky, 1 - kx), lo = 0, n = G
//   5: head-input layout (data at +3, +3 of a 10 x 10 frame), (3 - ky, 3 - kx) for the 4 x 4 taps, lo = 3, n = G - 2
struct PmTile {
  int fg, p, y, x, n0;
  uint32_t mask;   // valid taps
};
__device__ __forceinline__ int pm_ntaps(const rgp_tapgemm_args& a) { return a.im2col == 5 ? 16 : 9; }
__device__ __forceinline__ bool pm_tap(const rgp_tapgemm_args& a, int y, int x, int tap, int& srow) {
  int oy, ox, lo, n;
  if (a.im2col == 5) { oy = 3 - (tap >> 2); ox = 3 - (tap & 3); lo = 3; n = a.G - 2; }
  else if (a.im2col == 4) { oy = 1 - tap / 3; ox = 1 - tap % 3; lo = 0; n = a.G; }
  else { oy = tap / 3; ox = tap % 3; lo = 1; n = a.G; }
  const int sy = y + oy, sx = x + ox;
  srow = sy * a.Wp + sx;
  return sy >= lo && sy < lo + n && sx >= lo && sx < lo + n;
}
__device__ __forceinline__ int pm_frames(const rgp_tapgemm_args& a) { return a.im2col == 5 ? a.M / a.S : a.M / (a.G * a.G); }
__device__ __forceinline__ PmTile pm_decode(const rgp_tapgemm_args& a, int tile, int ntiles, int BNv, int rot) {
  PmTile t;
  const int P = a.G * a.G, per = P * ntiles;
  t.fg = tile / per;
  const int rem = tile - t.fg * per;
  const int j = rem / ntiles;
  t.n0 = (rem - j * ntiles) * BNv;
  t.p = (int)(((long long)j + (long long)t.fg * rot) % P);
  t.y = t.p / a.G;
  t.x = t.p - t.y * a.G;
  t.mask = 0;
  const int nt = pm_ntaps(a);
  for (int tap = 0; tap < nt; ++tap) {
    int srow;
    if (pm_tap(a, t.y, t.x, tap, srow)) t.mask |= 1u << tap;
  }
  return t;
}

constexpr int kSlabMaxSlots = 4;
constexpr int kSlabBarBytes = 512;     // (2 * nw + 8) mbarriers + the TMEM slot, nw <= kSlabMaxW
constexpr int kSlabMaxW = 24;
constexpr int kSmemBudget = 227 * 1024;

template <int BN, bool SLAB>
__global__ void __launch_bounds__(NTHREADS, 1)
tapgemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const rgp_tapgemm_args a, const SlabCfg sc) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int kWBytes = BN * BK * 2;
  const int wstage = SLAB ? sc.tg * kWBytes : kWBytes;   // bytes of one weight stage (tg taps)
  const int nstage = SLAB ? sc.nw : C::kStages;          // barrier pairs of the (weight) ring
  const int ring_bytes = SLAB ? sc.nw * wstage + sc.slots * sc.bytes : C::kStages * C::kStageBytes;
  uint8_t* slab0 = smem + (SLAB ? sc.nw * wstage : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ring_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + nstage;
  uint64_t* tfull = bars + 2 * nstage;
  uint64_t* tempty = tfull + 2;
  uint64_t* afull = tempty + 2;                          // SLAB only
  uint64_t* aempty = afull + kSlabMaxSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aempty + kSlabMaxSlots);
  float* scr_base = reinterpret_cast<float*>(smem + ring_bytes + (SLAB ? kSlabBarBytes : 256));
  RowLite* rl_base = reinterpret_cast<RowLite*>(scr_base + 4 * 32 * kScrLd);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mtiles = (a.M + BM - 1) / BM, ntiles = a.N / BN;
  const bool pm = !SLAB && a.im2col >= 3;
  const int total_tiles = pm ? ((pm_frames(a) + BM - 1) / BM) * (a.G * a.G) * ntiles : mtiles * ntiles;
  const int kchunks = a.K / BK;
  const int ksteps = a.ntaps * kchunks;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    for (int i = 0; i < nstage; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    if (SLAB)
      for (int i = 0; i < kSlabMaxSlots; ++i) { mbar_init(&afull[i], 1); mbar_init(&aempty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (SLAB && lane == 0) {
      int ws = 0; uint32_t wph = 0;
      int as = 0; uint32_t aph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int mt = tile / ntiles, nt = tile - mt * ntiles;
        const int m0 = mt * BM, n0 = nt * BN;
        for (int kc = 0; kc < kchunks; ++kc) {
          mbar_wait(&aempty[as], aph ^ 1);
          mbar_expect_tx(&afull[as], (uint32_t)sc.rows * BK * 2);
          for (int bx = 0; bx < sc.nbox; ++bx)     // boxrows % 8 == 0: every box starts on a swizzle-atom boundary
            tma_load_2d(slab0 + as * sc.bytes + bx * sc.boxrows * (BK * 2), &tmA, &afull[as], kc * BK,
                        m0 + sc.min_off + bx * sc.boxrows);
          for (int tap = 0; tap < a.ntaps; tap += sc.tg) {     // tg > 1 only when N == BN: the taps' rows are contiguous
            mbar_wait(&empty[ws], wph ^ 1);
            mbar_expect_tx(&full[ws], (uint32_t)wstage);       // rows past the last tap are out of bounds: zero filled
            tma_load_2d(smem + ws * wstage, &tmB, &full[ws], kc * BK, tap * a.N + n0);
            if (++ws == sc.nw) { ws = 0; wph ^= 1; }
          }
          if (++as == sc.slots) { as = 0; aph ^= 1; }
        }
      }
    } else if (lane == 0 && pm) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const PmTile t = pm_decode(a, tile, ntiles, BN, sc.pm_rot);
        const int nt = pm_ntaps(a);
        // RGP_EPI_UP7X: one weight set per class of the block position (first / inner / last row and column)
        const int wcls = a.epi == RGP_EPI_UP7X
                             ? 9 * (3 * (t.y == 0 ? 0 : (t.y == a.G - 1 ? 2 : 1)) + (t.x == 0 ? 0 : (t.x == a.G - 1 ? 2 : 1))) : 0;
        // accumulation order = the dense paths' (results equal bit for bit): convs tap-major, head k-chunk-major (slab mode)
        const bool kmajor = a.im2col == 5;
        const int n_out = kmajor ? kchunks : nt, n_in = kmajor ? nt : kchunks;
        for (int io = 0; io < n_out; ++io) {
          for (int ii = 0; ii < n_in; ++ii) {
            const int tap = kmajor ? ii : io, kc = kmajor ? io : ii;
            int srow;
            if (!pm_tap(a, t.y, t.x, tap, srow)) continue;
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* sa = smem + stage * C::kStageBytes;
            mbar_expect_tx(&full[stage], C::kStageBytes);
            tma_load_3d(sa, &tmA, &full[stage], kc * BK, srow, t.fg * BM);
            tma_load_2d(sa + BM * BK * 2, &tmB, &full[stage], kc * BK, (wcls + tap) * a.N + t.n0);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    } else if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int mt = tile / ntiles, nt = tile - mt * ntiles;
        const int m0 = mt * BM, n0 = nt * BN;
        for (int ks = 0; ks < ksteps; ++ks) {
          const int tap = ks / kchunks, kc = ks - tap * kchunks;
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + BM * BK * 2;
          mbar_expect_tx(&full[stage], C::kStageBytes);
          if (a.im2col) {            // dense-M conv: 128 consecutive real positions from m0, tap (ky, kx) as the im2col offset
            const int P = a.G * a.G;
            const int fn = m0 / P, rem = m0 - fn * P;
            const int hy = rem / a.G, wx = rem - hy * a.G;
            // im2col == 2: the transposed conv of the backward pass - tap (ky, kx) reads pixel (x + 1 - kx, y + 1 - ky)
            const int ox = a.im2col == 2 ? 2 - tap % 3 : tap % 3, oy = a.im2col == 2 ? 2 - tap / 3 : tap / 3;
            tma_load_im2col_4d(sa, &tmA, &full[stage], kc * BK, wx - 1, hy - 1, fn, (uint16_t)ox, (uint16_t)oy);
          } else {
            tma_load_2d(sa, &tmA, &full[stage], kc * BK, m0 + a.tap_off[tap]);
          }
          tma_load_2d(sb, &tmB, &full[stage], kc * BK, tap * a.N + n0);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // instruction descriptor: D=f32, A=B=bf16, both K-major, N=BN, M=128
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    int stage = 0; uint32_t phase = 0;
    int aslot = 0; uint32_t aphase = 0;
    int acc = 0; uint32_t accphase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      mbar_wait(&tempty[acc], accphase ^ 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + (uint32_t)(acc * C::kAccStride);
      if (SLAB) {
        // stage / phase walk the weight ring here; aslot / aphase the slab slots
        for (int kc = 0; kc < kchunks; ++kc) {
          mbar_wait(&afull[aslot], aphase);
          for (int tap0 = 0; tap0 < a.ntaps; tap0 += sc.tg) {
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            if (lane == 0) {
              const int tend = (tap0 + sc.tg < a.ntaps) ? tap0 + sc.tg : a.ntaps;
              for (int tap = tap0; tap < tend; ++tap) {
                const uint32_t sa = smem_u32(slab0 + aslot * sc.bytes) + (uint32_t)(a.tap_off[tap] - sc.min_off) * (BK * 2);
                const uint32_t sb = smem_u32(smem + stage * wstage) + (uint32_t)(tap - tap0) * kWBytes;
                const uint64_t adesc = make_sdesc(sa), bdesc = make_sdesc(sb);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)
                  tc_mma_bf16(tacc, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kc | tap | k) ? 1u : 0u);
              }
              tc_commit(&empty[stage]);
              if (tend == a.ntaps) {
                tc_commit(&aempty[aslot]);
                if (kc == kchunks - 1) tc_commit(&tfull[acc]);
              }
            }
            __syncwarp();
            if (++stage == sc.nw) { stage = 0; phase ^= 1; }
          }
          if (++aslot == sc.slots) { aslot = 0; aphase ^= 1; }
        }
        if (++acc == 2) { acc = 0; accphase ^= 1; }
        continue;
      }
      const int ksteps_t = pm ? __popc(pm_decode(a, tile, ntiles, BN, sc.pm_rot).mask) * kchunks : ksteps;
      for (int ks = 0; ks < ksteps_t; ++ks) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + BM * BK * 2;
          const uint64_t adesc = make_sdesc(sa), bdesc = make_sdesc(sb);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)   // +32 bytes (>>4 = 2) per K=16 slice inside the 128-byte swizzle row
            tc_mma_bf16(tacc, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (ks | k) ? 1u : 0u);
          tc_commit(&empty[stage]);
          if (ks == ksteps_t - 1) tc_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; accphase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..5 -> TMEM lane groups (warp % 4)) =====================
    const int lg = warp & 3;
    int acc = 0; uint32_t accphase = 0;
    const bool vec = (a.N % 32 == 0) && (a.ldo % 16 == 0) && (a.ldx % 8 == 0) && (a.Ch % 32 == 0) && (a.ldg % 16 == 0);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int n0, mrow;
      if (pm) {           // row i of the tile = frame fg*128 + i at position p: the dense row make_row() expects
        const PmTile t = pm_decode(a, tile, ntiles, BN, sc.pm_rot);
        n0 = t.n0;
        const int f = t.fg * BM + lg * 32 + lane;
        mrow = a.im2col == 5 ? f * a.S + t.y * a.Wp + t.x : f * (a.G * a.G) + t.p;
      } else {
        const int mt = tile / ntiles, nt = tile - mt * ntiles;
        n0 = nt * BN;
        mrow = mt * BM + lg * 32 + lane;
      }
      const RowInfo r = make_row(a, mrow);
      // While the MMAs of this tile are still running: pull the epilogue's inputs (pre-activations from the
      // input-half GEMM, z, h_{t-1}) from HBM into L2, one 128-byte line per prefetch, this thread's row.
      if (r.inrange && (a.epi == RGP_EPI_GRU_ZR || a.epi == RGP_EPI_GRU_C)) {
        const float* px = a.xg + (long long)r.m * a.ldx + (a.epi == RGP_EPI_GRU_C ? 2 * a.Ch : 0) + n0;
        for (int c = 0; c < BN; c += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(px + c));
        if (r.valid) {
          const int jb = (a.epi == RGP_EPI_GRU_ZR) ? (n0 >= a.Ch ? n0 - a.Ch : -1) : n0;
          if (jb >= 0) {
            const float* ph = a.hprev_f32 + r.prow * a.Ch + jb;
            for (int c = 0; c < BN && jb + c < a.Ch; c += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(ph + c));
          }
          if (a.epi == RGP_EPI_GRU_C) {
            const float* pz = a.z_f32 + (long long)r.m * a.Ch + n0;
            for (int c = 0; c < BN; c += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(pz + c));
          }
        }
      }
      mbar_wait(&tfull[acc], accphase);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * C::kAccStride);
      if constexpr (BN % 32 == 0) {
        const bool quad = vec && quad_mode(a.epi, a.accumulate);
        float* scr = scr_base + (warp - 2) * 32 * kScrLd;
        RowLite* rl = rl_base + (warp - 2) * 32;
        if (quad) {
          RowLite q;
          q.m = r.m; q.flags = (r.inrange ? 1 : 0) | (r.valid ? 2 : 0); q.prow = r.prow; q.orow = r.orow;
          rl[lane] = q;
        }
        if (a.epi == RGP_EPI_UP7X) {
          // fused up-sampling + top input conv: 32 columns = 2 pixels of this row's 7x7 block x 16 gate columns
#pragma unroll 1
          for (int c = 0; c < BN; c += 32) {
            float v[32];
            tc_ld32(tbase + (uint32_t)c, v);
            if (r.inrange) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int ph = (n0 + c) / 16 + h;
                if (ph < 49) {
                  const int py = ph / 7, px = ph - py * 7;
                  float* o = a.out_f32 + (r.obase + (long long)(r.hy + py) * a.Wo + r.hx + px) * a.ldo;
#pragma unroll
                  for (int q = 0; q < 4; ++q)
                    st4(o + 4 * q, make_float4(v[16 * h + 4 * q], v[16 * h + 4 * q + 1], v[16 * h + 4 * q + 2], v[16 * h + 4 * q + 3]));
                }
              }
            }
          }
        } else if (quad && (a.epi == RGP_EPI_STORE_F32 || a.epi == RGP_EPI_PROJ)) {
          // Pure-store epilogues (i2h pre-activations, projection): the 8 rows a lane serves are the same for every
          // 32-column chunk of the tile, so their destination pointers are formed once per tile and a chunk is
          // TMEM -> scratch transpose -> 8 x (16-byte scratch read, [bias,] one 16-/8-byte store): ~110 instructions
          // instead of the ~460 of the generic quad path (row records re-read from shared memory and a switch on the
          // epilogue kind per store) - the projection was bound by exactly that (ncu: 70 % of all samples there).
          __syncwarp();
          const int rr = lane >> 3, cg = (lane & 7) * 4;
          const bool proj = a.epi == RGP_EPI_PROJ;
          char* rowp[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const RowLite q = rl[4 * j + rr];
            rowp[j] = nullptr;
            if (q.flags & 1)
              rowp[j] = proj ? reinterpret_cast<char*>((__nv_bfloat16*)a.out_op + q.orow * a.ldo + n0 + cg)
                             : reinterpret_cast<char*>(a.out_f32 + (long long)q.m * a.ldo + n0 + cg);
          }
          if (proj) {
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
              float v[32];
              tc_ld32(tbase + (uint32_t)c, v);
#pragma unroll
              for (int i = 0; i < 8; ++i)
                st4(scr + lane * kScrLd + 4 * i, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
              __syncwarp();
              const float4 g = a.bias ? ld4(a.bias + n0 + c + cg) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 t = ld4(scr + (4 * j + rr) * kScrLd + cg);
                t.x += g.x; t.y += g.y; t.z += g.z; t.w += g.w;
                if (rowp[j]) st_bf16x4(reinterpret_cast<__nv_bfloat16*>(rowp[j]) + c, t);
              }
              __syncwarp();
            }
          } else {
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
              float v[32];
              tc_ld32(tbase + (uint32_t)c, v);
#pragma unroll
              for (int i = 0; i < 8; ++i)
                st4(scr + lane * kScrLd + 4 * i, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 t = ld4(scr + (4 * j + rr) * kScrLd + cg);
                if (rowp[j]) st4(reinterpret_cast<float*>(rowp[j]) + c, t);
              }
              __syncwarp();
            }
          }
        } else
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          float v[32];
          tc_ld32(tbase + (uint32_t)c, v);
          if (quad) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              st4(scr + lane * kScrLd + 4 * i, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
            __syncwarp();
            const int rr = lane >> 3, cg = (lane & 7) * 4;
            const int n = n0 + c + cg;
            float4 g = make_float4(0.f, 0.f, 0.f, 0.f), bb = g;
            if ((a.epi == RGP_EPI_PROJ || a.epi == RGP_EPI_GRU_C) && a.bias) g = ld4(a.bias + n);
            if (a.epi == RGP_EPI_GRU_C && a.bias) bb = ld4(a.bias2 + n);
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              QuadIn in[4];
              RowLite q[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                q[i] = rl[4 * (4 * half + i) + rr];
                if (q[i].flags & 1) quad_load(a, q[i], n, in[i]);
              }
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 t = ld4(scr + (4 * (4 * half + i) + rr) * kScrLd + cg);
                if (q[i].flags & 1) quad_store(a, q[i], n, t, in[i], g, bb);
              }
            }
            __syncwarp();
          } else {
            epi_row32(a, r, n0 + c, v, vec);
          }
        }
      } else {
        if (BN == 48 && a.epi == RGP_EPI_HEAD && a.N == 48) {
          // composed head (gaze_grcn.py:326-366): the 36 classes of a row are the 6x6 pixel block at (6a-3, 6b-3).
          // All 48 accumulators are read first, then the block is written with compile-time offsets - no per-element
          // switch / division (the generic element path made this GEMM epilogue-bound).
          float v[48];
          tc_ld16(tbase, v);
          tc_ld16(tbase + 16u, v + 16);
          tc_ld16(tbase + 32u, v + 32);
          if (r.inrange && r.valid) {
            const float bias = a.bias[0];
            float* o = a.out_f32 + r.obase;
#pragma unroll
            for (int ry = 0; ry < 6; ++ry) {
              const int Y = r.hy + ry;
              if (Y >= 0 && Y < 49) {
#pragma unroll
                for (int rx = 0; rx < 6; ++rx) {
                  const int X = r.hx + rx;
                  if (X >= 0 && X < 49) o[Y * 49 + X] = v[ry * 6 + rx] + bias;
                }
              }
            }
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < BN; c += 16) {
            float v[16];
            tc_ld16(tbase + (uint32_t)c, v);
#pragma unroll 4
            for (int i = 0; i < 16; ++i) epi_elem<__nv_bfloat16>(a, r, n0 + c + i, v[i]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; accphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::kTmemCols)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side: tensor maps (cuTensorMapEncodeTiled through the runtime's driver entry point) + launch
// ---------------------------------------------------------------------------------------------
int make_map(const void* base, long long rows, int ld, int K, int boxrows, CUtensorMap* out) {
  return tc::encode_map_bf16(base, K, rows, (long long)ld * 2, BK, boxrows, out);
}

// RGP_SLAB=0 switches the resident-A-slab mode off (A/B timing)
bool slab_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("RGP_SLAB");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

long long pm_frames_host(const rgp_tapgemm_args& a) { return a.im2col == 5 ? a.M / a.S : a.M / (a.G * a.G); }

// taps per position of the position-major modes (host mirror of pm_tap)
void pm_costs(const rgp_tapgemm_args& a, std::vector<int>& cost) {
  const int G = a.G;
  cost.assign(G * G, 0);
  const int nt = a.im2col == 5 ? 16 : 9;
  for (int y = 0; y < G; ++y)
    for (int x = 0; x < G; ++x)
      for (int tap = 0; tap < nt; ++tap) {
        int oy, ox, lo, n;
        if (a.im2col == 5) { oy = 3 - (tap >> 2); ox = 3 - (tap & 3); lo = 3; n = G - 2; }
        else if (a.im2col == 4) { oy = 1 - tap / 3; ox = 1 - tap % 3; lo = 0; n = G; }
        else { oy = tap / 3; ox = tap % 3; lo = 1; n = G; }
        const int sy = y + oy, sx = x + ox;
        if (sy >= lo && sy < lo + n && sx >= lo && sx < lo + n) ++cost[y * G + x];
      }
}

// Rotation of the position order per frame group that minimises the heaviest CTA of the static tile schedule
// (tile i runs on CTA i % grid): exhaustive over the G*G candidates, cached per shape.
int pm_rotation(const rgp_tapgemm_args& a, int fgroups, int ntiles, int grid) {
  static std::map<std::tuple<int, int, int, int, int>, int> cache;
  static std::mutex mu;
  const auto key = std::make_tuple(a.im2col, a.G, fgroups, ntiles, grid);
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
  }
  std::vector<int> cost;
  pm_costs(a, cost);
  const int P = a.G * a.G;
  const long long tiles = (long long)fgroups * P * ntiles;
  int best = 0;
  long long best_max = -1;
  std::vector<long long> load(grid);
  for (int rot = 0; rot < P; ++rot) {
    std::fill(load.begin(), load.end(), 0);
    for (long long i = 0; i < tiles; ++i) {
      const long long fg = i / ((long long)P * ntiles), j = (i % ((long long)P * ntiles)) / ntiles;
      load[i % grid] += cost[(j + fg * rot) % P];
    }
    const long long mx = *std::max_element(load.begin(), load.end());
    if (best_max < 0 || mx < best_max) { best_max = mx; best = rot; }
    if (tiles * P > 40000000LL) break;     // very long inputs: keep rot = 0 rather than search for seconds
  }
  std::lock_guard<std::mutex> g(mu);
  cache[key] = best;
  return best;
}

template <int BN>
int launch(const rgp_tapgemm_args& a, cudaStream_t s) {
  using C = Cfg<BN>;
  // slab geometry: all taps of the tile inside one TMA box of <= 256 rows
  SlabCfg sc{0, 1, 0, 0, 0, 0, 2, 1};
  bool slab = false;
  // Narrow tiles (the N = 48 head GEMM) keep the per-tap boxes: they are bound by the cost of their small MMAs (~150 clk per
  // 128x48x16 MMA, 580 clk per k-step whatever the operand traffic), measured 0.87 ms per-tap vs 0.95 ms slab with an 8-stage
  // weight ring and 0.93 ms with 24 stages and 4 slab slots (profiles/r01_notes.md passes 7 and 18).
  // ... except many-tap narrow GEMMs (the cascade's 5x5 top cell: 25 taps, N = 16, K = 64): there the per-tap A boxes
  // (25 x 16 KB per 128 rows) are the whole cost and the slab (128 + 4*51+4 rows, two TMA boxes) removes them.
  // Narrow tiles (N = 48 head, N = 16 top cell of the cascade) are bound by the MMA warp's per-k-step barrier round trip,
  // not by bytes (measured: per-tap boxes 0.87 ms, slab with one barrier per tap 0.93-0.95 ms): when N == BN the weight rows
  // of consecutive taps are contiguous, so ONE TMA box / barrier serves tg = 256 / BN taps (16 or 64 MMAs per wait).
  const bool group = BN <= 64 && a.N == BN && BN % 8 == 0 && a.ntaps >= 4;
  if (a.ntaps > 1 && (BN >= 128 || group) && slab_enabled() && !a.im2col) {
    int mn = a.tap_off[0], mx = a.tap_off[0];
    for (int i = 1; i < a.ntaps; ++i) { mn = a.tap_off[i] < mn ? a.tap_off[i] : mn; mx = a.tap_off[i] > mx ? a.tap_off[i] : mx; }
    sc.rows = (BM + (mx - mn) + 7) & ~7;
    if (sc.rows > 256) { sc.nbox = 2; sc.rows = (sc.rows + 15) & ~15; }
    sc.boxrows = sc.rows / sc.nbox;
    sc.bytes = (sc.rows * BK * 2 + 1023) & ~1023;
    sc.min_off = mn;
    sc.slots = (BN <= 64) ? kSlabMaxSlots : 2;      // narrow tiles finish a k-chunk faster than a slab load takes
    if (group) { sc.tg = 256 / BN; if (sc.tg > a.ntaps) sc.tg = a.ntaps; }
    if (sc.slots * sc.bytes > 120 * 1024) sc.slots = 2;
    const int fixed = 1024 + kSlabBarBytes + kEpiScratchBytes + sc.slots * sc.bytes;
    sc.nw = (kSmemBudget - fixed) / (sc.tg * BN * BK * 2);
    if (sc.nw > kSlabMaxW) sc.nw = kSlabMaxW;
    slab = sc.boxrows <= 256 && sc.nw >= 3;
  }
  CUtensorMap tmA, tmB;
  int rc;
  if (a.im2col >= 3) {
    const long long frames = pm_frames_host(a);
    rc = tc::encode_map_frames_bf16(a.A, a.K, a.S, frames, a.lda, BM, &tmA);
    sc.pm_rot = pm_rotation(a, (int)((frames + BM - 1) / BM), a.N / BN, num_sms());
  } else if (a.im2col) {
    const long long frames = a.M / (a.G * a.G);
    // im2col == 1: A is a p-layout operand (pixel (0,0) at row Wp+1); == 2: a q-layout gradient (pixel (0,0) at row 0)
    const long long row0 = a.im2col == 2 ? 0 : (a.Wp + 1);
    rc = tc::encode_map_im2col_bf16((const __nv_bfloat16*)a.A + row0 * a.lda, a.K, a.G, frames, a.lda, a.Wp, a.S, 1, BM, &tmA);
  } else {
    rc = make_map(a.A, a.rowsA, a.lda, a.K, slab ? sc.boxrows : BM, &tmA);
  }
  if (rc) return rc;
  rc = make_map(a.Bw, (long long)(a.epi == RGP_EPI_UP7X ? 81 : a.ntaps) * a.N, a.K, a.K, (slab ? sc.tg : 1) * BN, &tmB);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(tapgemm_tc_kernel<BN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(tapgemm_tc_kernel<BN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(smem=%d): %s", C::kSmemBytes, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.M + BM - 1) / BM, ntiles = a.N / BN;
  long long tiles = (long long)mtiles * ntiles;
  if (a.im2col >= 3) tiles = ((pm_frames_host(a) + BM - 1) / BM) * (long long)(a.G * a.G) * ntiles;
  int grid = tiles > num_sms() ? num_sms() : (int)tiles;
  if (slab) {
    const int smem_bytes = sc.nw * sc.tg * BN * BK * 2 + sc.slots * sc.bytes + 1024 + kSlabBarBytes + kEpiScratchBytes;
    tapgemm_tc_kernel<BN, true><<<grid, NTHREADS, smem_bytes, s>>>(tmA, tmB, a, sc);
  } else {
    tapgemm_tc_kernel<BN, false><<<grid, NTHREADS, C::kSmemBytes, s>>>(tmA, tmB, a, sc);
  }
  return check_launch("tapgemm_tc");
}

}  // namespace

namespace tc {

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

struct MapKey {
  const void* p; long long d0, d1, st; int b0, b1;
  bool operator<(const MapKey& o) const {
    return std::tie(p, d0, d1, st, b0, b1) < std::tie(o.p, o.d0, o.d1, o.st, o.b0, o.b1);
  }
};

int encode_map_bf16(const void* base, long long dim0, long long dim1, long long stride1_bytes, int box0, int box1,
                    CUtensorMap* out) {
  static std::map<MapKey, CUtensorMap> cache;
  static std::mutex mu;
  MapKey key{base, dim0, dim1, stride1_bytes, box0, box1};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return 0; }
  }
  auto enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return -38; }
  cuuint64_t dims[2] = {(cuuint64_t)dim0, (cuuint64_t)dim1};
  cuuint64_t strides[1] = {(cuuint64_t)stride1_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box0, (cuuint32_t)box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) base=%p dims=%lld x %lld stride=%lld box=%d x %d", (int)r, base, dim0,
              dim1, stride1_bytes, box0, box1);
    return -5;
  }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 4096) cache.clear();
  cache[key] = *out;
  return 0;
}

int encode_map_frames_bf16(const void* base, int C, int rows_per_frame, long long frames, int ld, int box_frames,
                           CUtensorMap* out) {
  auto enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return -38; }
  cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)rows_per_frame, (cuuint64_t)frames};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)rows_per_frame * ld * 2};
  cuuint32_t box[3] = {64, 1, (cuuint32_t)box_frames};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(3d) failed (%d) base=%p C=%d rows=%d frames=%lld ld=%d", (int)r, base, C, rows_per_frame,
              frames, ld);
    return -5;
  }
  return 0;
}

int encode_map_im2col_bf16(const void* base, int C, int G, long long frames, int ld, int Wp, int S, int pad, int pixels,
                           CUtensorMap* out) {
  static PFN_cuTensorMapEncodeIm2col_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeIm2col_v12000>(p);
  });
  if (!fn) { set_error("cuTensorMapEncodeIm2col not available from the driver"); return -38; }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)G, (cuuint64_t)G, (cuuint64_t)frames};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)Wp * ld * 2, (cuuint64_t)S * ld * 2};
  // bounding box of the base pixel (the filter's top-left position): [-pad, G - 1 + pad - (k-1)] = [-pad, G-1-pad]
  int lower[2] = {-pad, -pad}, upper[2] = {-pad, -pad};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper, 64,
                  (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeIm2col failed (%d) base=%p C=%d G=%d frames=%lld ld=%d", (int)r, base, C, G, frames, ld);
    return -5;
  }
  return 0;
}

int num_sms() {
  static int cache[kMaxDevices] = {0};
  const int dev = current_device();
  int n = cache[dev];
  if (!n) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cache[dev] = n;
  }
  return n;
}

}  // namespace tc

int tapgemm_tc(const rgp_tapgemm_args& a, cudaStream_t s) {
  RGP_CHECK_ARG(a.K % BK == 0, "tapgemm_tc: K=%d must be a multiple of %d", a.K, BK);
  if (a.im2col)
    RGP_CHECK_ARG(a.ntaps == 9 && !a.accumulate && a.G > 0 && a.M % (a.G * a.G) == 0 && a.N % 32 == 0 &&
                  a.S >= (a.G + 1) * a.Wp - 1 && a.Wp > a.G && a.im2col <= 4 &&
                  ((a.im2col % 2 == 1 && a.epi == RGP_EPI_STORE_F32) || (a.im2col % 2 == 0 && a.epi == RGP_EPI_DX_DENSE) ||
                   (a.im2col == 3 && a.epi == RGP_EPI_UP7X)),
                  "tapgemm_tc: im2col mode is a 3x3 SAME conv: 1 + RGP_EPI_STORE_F32 (forward) or 2 + RGP_EPI_DX_DENSE (its "
                  "transpose) (im2col=%d ntaps=%d epi=%d M=%d G=%d)", a.im2col, a.ntaps, a.epi, a.M, a.G);
  RGP_CHECK_ARG(a.epi != RGP_EPI_UP7X || (a.im2col == 3 && a.G == 7 && a.N == 1024 && a.ldo >= 16 && a.ldo % 4 == 0 && a.out_f32),
                "tapgemm_tc: RGP_EPI_UP7X is the position-major (im2col = 3) GEMM over the 7x7 grid with N = 1024 (im2col=%d G=%d N=%d)",
                a.im2col, a.G, a.N);
  RGP_CHECK_ARG(a.lda % 8 == 0, "tapgemm_tc: lda=%d must be a multiple of 8 (16-byte TMA stride)", a.lda);
  RGP_CHECK_ARG(((uintptr_t)a.A & 15) == 0 && ((uintptr_t)a.Bw & 15) == 0, "tapgemm_tc: operands must be 16-byte aligned");
  if (a.im2col == 1 || a.im2col == 2) {
    // position-major tiles (im2col + 2) when skipping the out-of-frame taps outweighs the ragged last frame group:
    // ceil(F / 128) * (3G-2)^2 tap-tiles against ceil(F * G^2 / 128) * 9.  RGP_PMAJOR=0 keeps the dense-M order.
    const char* pe = getenv("RGP_PMAJOR");
    const bool pm_on = !(pe && pe[0] == '0');
    const long long F = a.M / (a.G * a.G), e3 = 3 * a.G - 2;
    const long long cost_pm = ((F + BM - 1) / BM) * e3 * e3, cost_dense = ((long long)(a.M + BM - 1) / BM) * 9;
    if (pm_on && a.N % 256 == 0 && cost_pm * 100 < cost_dense * 95) {
      rgp_tapgemm_args b = a;
      b.im2col = a.im2col + 2;
      return launch<256>(b, s);
    }
  }
  if (!a.im2col && a.epi == RGP_EPI_HEAD && a.ntaps == 16 && a.N == 48 && a.G == 9 && a.Wp == 10 && a.S == 100 && !a.accumulate) {
    // the composed head (taps (3-dy)*Wp + (3-dx)) in position-major tiles: 729 instead of 1600 row-taps per frame
    bool pattern = true;
    for (int t = 0; t < 16; ++t) pattern = pattern && a.tap_off[t] == (3 - (t >> 2)) * a.Wp + (3 - (t & 3));
    const char* pe = getenv("RGP_PMAJOR");
    const long long F = a.M / a.S;
    const long long cost_pm = ((F + BM - 1) / BM) * 729 * 5 / 4, cost_dense = ((long long)(a.M + BM - 1) / BM) * 16;
    if (pattern && !(pe && pe[0] == '0') && cost_pm * 100 < cost_dense * 90) {
      rgp_tapgemm_args b = a;
      b.im2col = 5;
      return launch<48>(b, s);
    }
  }
  const int mtiles = (a.M + BM - 1) / BM;
  // less than half a wave of 128x256 tiles (small-batch recurrence steps): 128x128 tiles fill twice the SMs
  if (a.N % 256 == 0 && (long long)mtiles * (a.N / 256) * 2 > num_sms()) return launch<256>(a, s);
  if (a.N % 128 == 0) return launch<128>(a, s);
  if (a.N % 64 == 0) return launch<64>(a, s);
  if (a.N % 48 == 0) return launch<48>(a, s);
  if (a.N % 32 == 0) return launch<32>(a, s);
  if (a.N % 16 == 0) return launch<16>(a, s);
  set_error("tapgemm_tc: N=%d must be a multiple of 16", a.N);
  return -22;
}

}  // namespace rgp

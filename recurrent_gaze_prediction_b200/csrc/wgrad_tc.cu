// wgrad_tc.cu — tcgen05 / TMEM / TMA engine of the weight-gradient contraction over rows (bf16 operands, fp32 accumulate):
//   D[tap][i, j] += sum_{m<M} Y[m, ycol0+i] * X[m + tap_off[tap], xcol0+j]
// (dW of every conv / projection / the composed head: base.py:278-281 applied to gaze_grcn.py:108-125,239-242,326-358).
//
// The contraction index is the ROW index m, so both operands are "MN-major" for the tensor core: a 128-byte shared
// memory row holds 64 consecutive i (or j) of ONE m.  A 2-D TMA box {64 columns, 64 rows} with SWIZZLE_128B lands in
// exactly the canonical MN-major SW128 layout (8 m-rows = one 1024-byte swizzle atom), so Y and X are fed to
// tcgen05.mma straight from their row-major HBM layout — no transposed copies.  A conv tap is a row shift of the X box.
//
// Work decomposition: the row range is cut into slabs; unit u = slab * ntiles + tile is `kps` k-steps (64 rows each)
// long.  Units are dealt ROUND-ROBIN to the persistent CTAs (round r: CTA c owns unit r*grid + c), so the CTAs that run
// at the same time work on neighbouring units = the same one or two row slabs, which are sized to stay in the 126 MB
// L2 while all tiles and all taps consume them (a contiguous stream-K split of the whole space would spread the CTAs
// over ALL slabs at once: measured 15x the algorithmic DRAM reads).  The units left over after the last whole round
// are split stream-K style (equal contiguous k-step shares), so the load is balanced to one k-step.  Partial tiles
// are added to D with red.global.add.v4.f32.
//
// Warp roles as in tapgemm_tc.cu: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..5 epilogue;
// two TMEM accumulator stages so the flush of one segment overlaps the main loop of the next.
#include "tc_common.cuh"

namespace rgp {
namespace {
using namespace tc;

constexpr int BM = 128;       // i-tile (rows of D) = MMA M
constexpr int KB = 64;        // m-rows per k-step (4 MMAs of K=16)
constexpr int NTHREADS = 192;
constexpr int CH = 64;        // columns per TMA box (128 bytes)

template <int BN>
struct WCfg {
  static constexpr int kABytes = BM * KB * 2;              // 2 boxes of 64 cols x 64 rows
  static constexpr int kBBytes = BN * KB * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : 6;
  static constexpr int kAccStride = (BN <= 64) ? 64 : ((BN <= 128) ? 128 : 256);
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

struct Sched {
  int kps;          // k-steps per unit
  int ksteps_total; // ceil(M / KB)
  int ntiles;       // ntaps * tiles_i * tiles_j
  int tiles_j, tiles_ij;
  int rounds;       // whole-unit rounds: in round r CTA c owns unit r*grid + c (neighbouring units = same row slab -> L2 hits)
  long long tail0;  // first linear k-step index of the stream-K tail (= rounds*grid*kps)
  long long total;  // units * kps
  long long per;    // tail k-steps per CTA
};

struct Seg { int tile, kb, ke; };   // output tile, absolute k-step range [kb, ke)

// Iterates the segments of one CTA: `rounds` whole units dealt round-robin, then an equal contiguous share of the
// remaining units' k-steps (stream-K tail).  All three warp roles walk the identical sequence.
struct SegIter {
  int round;
  long long g, g1;
  __device__ __forceinline__ SegIter(const Sched& sc) {
    round = 0;
    g = sc.tail0 + (long long)blockIdx.x * sc.per;
    g1 = g + sc.per;
    if (g1 > sc.total) g1 = sc.total;
  }
  __device__ __forceinline__ bool next(const Sched& sc, Seg& s) {
    long long lin, len;
    if (round < sc.rounds) {
      lin = ((long long)round * gridDim.x + blockIdx.x) * sc.kps;
      len = sc.kps;
      ++round;
    } else {
      if (g >= g1) return false;
      lin = g;
      len = sc.kps - (g % sc.kps);
      if (len > g1 - g) len = g1 - g;
      g += len;
    }
    const long long u = lin / sc.kps;
    const int k0 = (int)(lin - u * sc.kps);
    const int split = (int)(u / sc.ntiles);
    s.tile = (int)(u - (long long)split * sc.ntiles);
    s.kb = split * sc.kps + k0;
    s.ke = s.kb + (int)len;
    if (s.ke > sc.ksteps_total) s.ke = sc.ksteps_total;   // the last split may be short: skipped steps do no work
    return true;
  }
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmX,
                const rgp_wgrad_args a, const Sched sc) {
  using C = WCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* tfull = bars + 2 * C::kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmY) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
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
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      SegIter it(sc);
      Seg s;
      while (it.next(sc, s)) {
        const int tap = s.tile / sc.tiles_ij;
        const int rem = s.tile - tap * sc.tiles_ij;
        const int ti = rem / sc.tiles_j, tj = rem - ti * sc.tiles_j;
        const int i0 = ti * BM, j0 = tj * BN, off = a.tap_off[tap];
        for (int k = s.kb; k < s.ke; ++k) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          mbar_expect_tx(&full[stage], C::kStageBytes);
          const int m0 = k * KB;
#pragma unroll
          for (int c = 0; c < BM / CH; ++c) tma_load_2d(sa + c * (CH * KB * 2), &tmY, &full[stage], i0 + c * CH, m0);
#pragma unroll
          for (int c = 0; c < BN / CH; ++c) tma_load_2d(sb + c * (CH * KB * 2), &tmX, &full[stage], j0 + c * CH, m0 + off);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // instruction descriptor: D=f32, A=B=bf16, both MN-major (bits 15,16), N=BN, M=128
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                               ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    constexpr uint32_t kLbo = CH * KB * 2;    // next 64 columns: the next TMA box (8192 bytes)
    constexpr uint32_t kSbo = 1024;           // next 8 m-rows: the next swizzle atom
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t accphase = 0;
    SegIter it(sc);
    Seg s;
    while (it.next(sc, s)) {
      if (s.kb >= s.ke) continue;
      mbar_wait(&tempty[acc], accphase ^ 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + (uint32_t)(acc * C::kAccStride);
      for (int k = s.kb; k < s.ke; ++k) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + C::kABytes;
          const uint64_t adesc = make_sdesc_mn(sa, kLbo, kSbo), bdesc = make_sdesc_mn(sb, kLbo, kSbo);
#pragma unroll
          for (int q = 0; q < KB / 16; ++q)   // 16 m-rows = 2 atoms = 2048 bytes (>>4 = 128) per MMA
            tc_mma_bf16(tacc, adesc + (uint64_t)(128 * q), bdesc + (uint64_t)(128 * q), idesc, (k > s.kb || q) ? 1u : 0u);
          tc_commit(&empty[stage]);
          if (k == s.ke - 1) tc_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; accphase ^= 1; }
    }
  } else {
    // ===================== epilogue: TMEM -> red.global.add into D =====================
    const int lg = warp & 3;
    int acc = 0; uint32_t accphase = 0;
    SegIter it(sc);
    Seg s;
    while (it.next(sc, s)) {
      if (s.kb >= s.ke) continue;
      const int tap = s.tile / sc.tiles_ij;
      const int rem = s.tile - tap * sc.tiles_ij;
      const int ti = rem / sc.tiles_j, tj = rem - ti * sc.tiles_j;
      const int i = ti * BM + lg * 32 + lane, j0 = tj * BN;
      float* drow = a.D + ((long long)tap * a.Ny + i) * a.Nx + j0;
      mbar_wait(&tfull[acc], accphase);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * C::kAccStride);
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        float v[32];
        tc_ld32(tbase + (uint32_t)c, v);
        if (i < a.Ny) {
#pragma unroll
          for (int q = 0; q < 8; ++q)
            if (j0 + c + 4 * q < a.Nx) red_add_v4(drow + c + 4 * q, v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
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

template <int BN>
int launch(const rgp_wgrad_args& a, cudaStream_t s) {
  using C = WCfg<BN>;
  CUtensorMap tmY, tmX;
  const __nv_bfloat16* Y = (const __nv_bfloat16*)a.Y + a.ycol0;
  const __nv_bfloat16* X = (const __nv_bfloat16*)a.X + a.xcol0;
  int rc = encode_map_bf16(Y, a.Ny, a.M, (long long)a.ldy * 2, CH, KB, &tmY);
  if (rc) return rc;
  rc = encode_map_bf16(X, a.Nx, a.rowsX, (long long)a.ldx * 2, CH, KB, &tmX);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(wgrad smem=%d): %s", C::kSmemBytes, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  Sched sc;
  sc.ksteps_total = (a.M + KB - 1) / KB;
  sc.tiles_j = a.Nx / BN;
  sc.tiles_ij = ((a.Ny + BM - 1) / BM) * sc.tiles_j;
  sc.ntiles = a.ntaps * sc.tiles_ij;
  // Row slabs: CTAs of one round work on ~grid/ntiles + 1 neighbouring slabs at the same time; size the slab so that
  // this concurrent set of Y and X rows (~64 MB) stays in the 126 MB L2 while every tile (and every tap) consumes it.
  const int nsm = num_sms();
  const double concurrent = (double)nsm / sc.ntiles + 1.0;
  long long slab_rows = (long long)(64.0 * 1024 * 1024 / (2.0 * (a.Ny + a.Nx) * concurrent));
  int kps = (int)(slab_rows / KB);
  if (kps < 32) kps = 32;
  if (kps > sc.ksteps_total) kps = sc.ksteps_total;
  const int splits = (sc.ksteps_total + kps - 1) / kps;
  sc.kps = (sc.ksteps_total + splits - 1) / splits;
  const long long units = (long long)splits * sc.ntiles;
  sc.total = units * sc.kps;
  long long grid = nsm;
  if (grid > units) grid = units;
  sc.rounds = (int)(units / grid);
  sc.tail0 = (long long)sc.rounds * grid * sc.kps;
  const long long tail = sc.total - sc.tail0;
  sc.per = (tail + grid - 1) / grid;
  wgrad_tc_kernel<BN><<<(int)grid, NTHREADS, C::kSmemBytes, s>>>(tmY, tmX, a, sc);
  return check_launch("wgrad_tc");
}

}  // namespace

int wgrad_tc(const rgp_wgrad_args& a, cudaStream_t s) {
  const bool ok = (a.ldy % 8 == 0) && (a.ldx % 8 == 0) && (a.ycol0 % 8 == 0) && (a.xcol0 % 8 == 0) &&
                  (((uintptr_t)a.Y & 15) == 0) && (((uintptr_t)a.X & 15) == 0) && (((uintptr_t)a.D & 15) == 0) &&
                  (a.Nx % 64 == 0) && (a.Ny % 8 == 0);
  if (!ok) return wgrad_simt(a, s);     // shapes the TMA boxes cannot express (tiny test sizes): CUDA-core engine
  for (int t = 0; t < a.ntaps; ++t)
    RGP_CHECK_ARG(a.tap_off[t] >= 0, "wgrad_tc: negative tap offset %d", a.tap_off[t]);
  if (a.Nx % 256 == 0) return launch<256>(a, s);
  if (a.Nx % 128 == 0) return launch<128>(a, s);
  return launch<64>(a, s);
}

}  // namespace rgp

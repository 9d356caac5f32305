// lstm_seq.cu — persistent ConvLSTM recurrence: all T steps of gaze_lstm.py:103-133 (unrolled by :260-286) in ONE launch,
// for any grid (BASELINE config 4: 14x14, T = 128, where a frame is 225 q-rows and straddles the 128-row MMA tiles).
//
// CTA (tile, split) owns 128 q-rows x NC = 16 channels of ALL THREE hidden convolutions (W_hi, W_hf, W_ho - the
// candidate reuses W_hi and W_hc is dead in the reference, gaze_lstm.py:112-133) for the whole sequence:
//   * its slice of W_h - 9 taps x Ch/64 k-chunks x (3 gates x 16 rows) x 128 B = 110 KB at Ch = 128 - is fetched ONCE by TMA
//     and stays in shared memory ("recurrent weights resident in SMEM");
//   * per step: the 128 + 2*Wp+2 state rows the tile's 9 taps touch arrive as one TMA slab per k-chunk (tap = constant row
//     shift inside the slab, same trick as tapgemm_tc.cu / gru_seq.cu), 9 x Ch/16 tcgen05 MMAs (M = 128, N = 48, K = 16)
//     accumulate i | f | o into 48 TMEM columns, and the epilogue (one thread = one row) applies the gate maths with the
//     cell state kept in REGISTERS for all T steps and the peepholes in shared memory;
//   * the halo rows of a step come from the two neighbouring tiles.  Every CTA bumps its tile's counter in `flags`
//     (release, gpu scope) after its rows of h_t are in global memory; the producer of tile i waits until tiles i-1, i,
//     i+1 have all nsplit splits of step t (acquire), then fences the async proxy and issues the slab TMA.  No grid-wide
//     barrier, no per-step launch.  All CTAs are co-resident (cooperative launch), so the waits cannot dead-lock; a
//     bounded spin reports a time-out through flags[mtiles] instead of hanging.
// Warp roles: warp 0 = dependency wait + TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer, warps 2..5 =
// epilogue (TMEM lane group = warp % 4).
#include "tc_common.cuh"

namespace rgp {
namespace lstmseq {

using namespace tc;

constexpr int BM = 128, BK = 64;
constexpr int NC = 16;                    // channels per CTA
constexpr int NG = 3 * NC;                // MMA N: i | f | o
constexpr int NTHREADS = 192;
constexpr int NTHREADS_MULTI = 192;      // (two alternating epilogue groups of four warps were measured SLOWER: 5.21 vs 4.85 ms)
constexpr int kMaxSlabRows = 160;         // 128 + 2*15 + 2 at G = 14
constexpr int kMaxKch = 2;                // Ch <= 128: the weight slice must fit in shared memory
constexpr int kWTileBytes = NG * BK * 2;  // 6144: one (tap, k-chunk) B operand
constexpr int kSlabBytes = kMaxSlabRows * BK * 2;   // 20480
constexpr int kTmemCols = 64;
constexpr long long kSpinLimit = 4000000000LL;      // clock64 ticks (~2 s) before a dependency wait gives up

// single-MUFU gates as everywhere on the bf16 engine (rel. error ~2^-11, below the 2^-9 rounding of the bf16 operands): the
// accurate forms cost ~1700 instructions per thread and step, on the critical path of every step
__device__ __forceinline__ float tanh_fast(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigm(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }
// generic-proxy stores to GLOBAL memory -> async-proxy (TMA) reads: the .global form orders only that state space
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void ld8(const float* p, float* v) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p));
}
__device__ __forceinline__ void st8(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
               "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
               : "memory");
}
__device__ __forceinline__ void stbf16x16(__nv_bfloat16* p, const float v[16]) {
  uint32_t u[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    __nv_bfloat162 b = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    u[j] = *reinterpret_cast<uint32_t*>(&b);
  }
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]),
               "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7])
               : "memory");
}

// Phase timestamps (globaltimer ns) of CTA 0 for its first kTsSteps steps - rgp_convlstm_seq_profile().  Events:
// 0 dependencies satisfied (producer) | 1 slab TMA issued | 2 last slab box landed (MMA warp) | 3 MMAs issued |
// 4 accumulator seen by the epilogue | 5 gates done, operand rows stored | 6 fences + CTA barrier done (counter bumped)
constexpr int kTsSteps = 32, kTsEvents = 8;
__device__ unsigned long long g_lstm_ts[kTsSteps * kTsEvents];
__device__ __forceinline__ void stamp(bool on, int t, int ev) {
  if (on && t < kTsSteps) {
    unsigned long long v;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v));
    g_lstm_ts[t * kTsEvents + ev] = v;
  }
}

constexpr int kSmemW = 9 * kMaxKch * kWTileBytes;           // 110592
constexpr int kSmemSlab = kMaxKch * kSlabBytes;             // 40960
constexpr int kSmemPeep = 3 * NC * BM * (int)sizeof(float); // 24576
constexpr int kSmemBytes = kSmemW + kSmemSlab + kSmemPeep + 256 + 1024;

__global__ void __launch_bounds__(NTHREADS, 1)
lstm_seq_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmW,
                const rgp_lstm_seq_args a, const int slab_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                                   // [tap][kc] tiles of 48 rows x 128 B (i | f | o)
  uint8_t* slab = smem + kSmemW;                         // [kc] slab_rows x 128 B
  float* peep = reinterpret_cast<float*>(smem + kSmemW + kSmemSlab);   // [gate][channel][row]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemW + kSmemSlab + kSmemPeep);
  uint64_t* wfull = bars;
  uint64_t* fullA = bars + 1;                            // [kMaxKch]
  uint64_t* acc_full = fullA + kMaxKch;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ch = a.Ch, G = a.G, Wp = G + 1, S = Wp * Wp;
  const int nsplit = Ch / NC;
  const int tile = (int)(blockIdx.x / nsplit), split = (int)(blockIdx.x % nsplit);
  const int ch0 = split * NC;
  const int M = a.B * S;
  const int mtiles = (M + BM - 1) / BM;
  const int m0 = tile * BM;
  const int kch = Ch / BK;
  int* const err = a.flags + mtiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmH) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
    mbar_init(wfull, 1);
    for (int i = 0; i < kMaxKch; ++i) mbar_init(&fullA[i], 1);
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== dependency wait + TMA producer =====================
    if (lane == 0) {
      // the weight slice, once: 9 taps x kch k-chunks x 3 gates, boxes of 16 rows x 64 columns
      mbar_expect_tx(wfull, (uint32_t)(9 * kch * kWTileBytes));
      for (int tap = 0; tap < 9; ++tap)
        for (int kc = 0; kc < kch; ++kc)
          for (int g = 0; g < 3; ++g)
            tma_load_2d(wsm + (tap * kch + kc) * kWTileBytes + g * (NC * BK * 2), &tmW, wfull, kc * BK,
                        tap * 3 * Ch + g * Ch + ch0);
    }
    // lanes 0..2 watch the counters of tiles (tile-1, tile, tile+1) in parallel: one L2 round trip per poll, not three
    const int nb = tile - 1 + lane;
    const bool watch = lane < 3 && nb >= 0 && nb < mtiles;
    bool dead = false;
    for (int t = 0; t < a.T; ++t) {
      if (t > 0 && !dead) {
        const int need = nsplit * t;
        const long long t0 = clock64();
        for (;;) {
          const bool ok = !watch || ld_acquire(a.flags + nb) >= need;
          if (__all_sync(0xffffffffu, ok)) break;
          bool bad = false;
          if (lane == 0) {
            bad = ld_acquire(err) != 0;
            if (!bad && clock64() - t0 > kSpinLimit) { atomicExch(err, 1); bad = true; }
          }
          if (__any_sync(0xffffffffu, bad)) { dead = true; break; }
        }
      }
      if (lane == 0) {
        stamp(blockIdx.x == 0, t, 0);
        fence_proxy_async();        // the neighbours' generic-proxy stores of h_t, observed above, before the async-proxy read
        for (int kc = 0; kc < kch; ++kc) {
          mbar_expect_tx(&fullA[kc], (uint32_t)(slab_rows * BK * 2));
          tma_load_2d(slab + kc * kSlabBytes, &tmH, &fullA[kc], kc * BK, t * M + m0);
        }
        stamp(blockIdx.x == 0, t, 1);
      }
      __syncwarp();
      // the slab of step t+1 may only be overwritten after this step's MMAs: they complete before acc_full, which the
      // epilogue awaits before it bumps this tile's counter, which the wait above includes (nb == tile)
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NG >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mbar_wait(wfull, 0);
    for (int t = 0; t < a.T; ++t) {
      const uint32_t par = (uint32_t)(t & 1);
      for (int kc = 0; kc < kch; ++kc) {
        mbar_wait(&fullA[kc], par);
        tc_fence_after();
        if (lane == 0) {
          if (kc == kch - 1) stamp(blockIdx.x == 0, t, 2);
          for (int tap = 0; tap < 9; ++tap) {
            const int off = (tap / 3) * Wp + (tap % 3);
            const uint64_t adesc = make_sdesc(smem_u32(slab + kc * kSlabBytes) + (uint32_t)off * (BK * 2));
            const uint64_t bdesc = make_sdesc(smem_u32(wsm + (tap * kch + kc) * kWTileBytes));
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              tc_mma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kc | tap | k) ? 1u : 0u);
          }
          if (kc == kch - 1) { tc_commit(acc_full); stamp(blockIdx.x == 0, t, 3); }
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue warps 2..5: one thread = one q-row, NC channels =====================
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int lrow = lg * 32 + lane;
    const int m = m0 + lrow;
    const bool inrange = m < M;
    const int f = m / S, q = m - f * S;
    const int y = q / Wp, x = q - y * Wp;
    const bool valid = inrange && y < G && x < G;
    const long long prow = (long long)m + Wp + 1;
    const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16);
    // peepholes of this row -> shared memory [gate][channel][row] (read back by the same thread: no barrier needed)
    {
      const int p = y * G + x;
      const float* pw[3] = {a.wci, a.wcf, a.wco};
#pragma unroll
      for (int g = 0; g < 3; ++g)
#pragma unroll
        for (int j = 0; j < NC; ++j) peep[(g * NC + j) * BM + lrow] = valid ? pw[g][(long long)p * Ch + ch0 + j] : 0.f;
    }
    float c[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) c[j] = 0.f;
    if (valid) { ld8(a.Cf + prow * Ch + ch0, c); ld8(a.Cf + prow * Ch + ch0 + 8, c + 8); }
    const int fourCh = 4 * Ch;
    for (int t = 0; t < a.T; ++t) {
      const long long mt = (long long)t * M + m;
      // input-half pre-activations of this row: requested before the accumulator is awaited
      float xi[NC], xf[NC], xc[NC], xo[NC];
      if (valid) {
        const float* xr = a.Xg + mt * fourCh + ch0;
        ld8(xr, xi); ld8(xr + 8, xi + 8);
        ld8(xr + Ch, xf); ld8(xr + Ch + 8, xf + 8);
        ld8(xr + 2 * Ch, xc); ld8(xr + 2 * Ch + 8, xc + 8);
        ld8(xr + 3 * Ch, xo); ld8(xr + 3 * Ch + 8, xo + 8);
      }
      mbar_wait(acc_full, (uint32_t)(t & 1));
      tc_fence_after();
      const bool ts_on = blockIdx.x == 0 && warp == 2 && lane == 0;
      stamp(ts_on, t, 4);
      float hi[NC], hf[NC], ho[NC];
      tc_ld16(tlane, hi);
      tc_ld16(tlane + 16u, hf);
      tc_ld16(tlane + 32u, ho);
      tc_fence_before();
      float gi[NC], gf[NC], gg[NC], go[NC], hn[NC], cp[NC];
      long long rn = 0;
      if (valid) {
#pragma unroll
        for (int j = 0; j < NC; ++j) {           // gaze_lstm.py:112-133 as written: W_hi shared by i and the candidate,
          cp[j] = c[j];                          // the output gate peeps at the OLD cell
          gi[j] = sigm(xi[j] + hi[j] + peep[(0 * NC + j) * BM + lrow] * cp[j]);
          gf[j] = sigm(xf[j] + hf[j] + peep[(1 * NC + j) * BM + lrow] * cp[j]);
          gg[j] = tanh_fast(xc[j] + hi[j]);
          c[j] = gf[j] * cp[j] + gi[j] * gg[j];
          go[j] = sigm(xo[j] + ho[j] + peep[(2 * NC + j) * BM + lrow] * cp[j]);
          hn[j] = tanh_fast(c[j]) * go[j];
        }
        rn = ((long long)(t + 1) * M + prow) * Ch + ch0;                     // slab t+1 = h_{t+1}
        stbf16x16((TA*)a.Hop + rn, hn);      // the only rows the next step's MMAs read: published first
      }
      stamp(ts_on, t, 5);
      // publish: every thread hands its operand rows of h_{t+1} to the async proxy, the CTA barrier orders all 128
      // threads' stores before ONE release at gpu scope (cumulative over what the barrier made visible to thread 0) -
      // one fence per CTA and step instead of 128 (measured: 1.8 us of the 7.4 us step were the per-thread fences)
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(a.flags + tile, 1); }
      stamp(ts_on, t, 6);
      // tail, off the critical path: fp32 master, head rows, training saves
      if (valid) {
        st8(a.Hf + rn, hn); st8(a.Hf + rn + 8, hn + 8);
        if (a.HS) {
          const long long orow = ((long long)t * a.B + f) * a.So + (long long)(y + a.pado) * a.Wo + (x + a.pado);
          stbf16x16((TA*)a.HS + orow * Ch + ch0, hn);
        }
        if (a.save) {
          float* sv = a.save + mt * 5 * Ch + ch0;
          st8(sv, gi); st8(sv + 8, gi + 8);
          st8(sv + Ch, gf); st8(sv + Ch + 8, gf + 8);
          st8(sv + 2 * Ch, gg); st8(sv + 2 * Ch + 8, gg + 8);
          st8(sv + 3 * Ch, go); st8(sv + 3 * Ch + 8, go + 8);
          st8(sv + 4 * Ch, cp); st8(sv + 4 * Ch + 8, cp + 8);
        }
      }
    }
    if (valid) { st8(a.Cf + prow * Ch + ch0, c); st8(a.Cf + prow * Ch + ch0 + 8, c + 8); }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Multi-tile variant: more tiles than co-resident (tile, split) CTAs (14x14 at 64 clips: 113 tiles x 8 splits).  The grid is
// R row groups x nsplit splits with R = SMs / nsplit; CTA (g, split) keeps ITS split's weight slice resident and walks the
// tiles g, g+R, g+2R, ... of every step in that order.  Item (t, k) = tile k of step t; dependencies of an item are the
// step-(t-1) items of its tile and the two neighbours, all earlier in every CTA's program order, so the counter waits
// cannot cycle.  Pipelining inside a CTA: two slab buffers and two TMEM accumulators - the slab of item i+1 loads while
// the MMAs of item i run, the MMAs of item i+1 run while the epilogue of item i does the gate maths.  The cell state
// lives in global memory (fp32, own rows / channels only: no hazard), the peepholes of the CTA's 16 channels are a
// [gate][channel][position] table in shared memory.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kMaxP = 196;                                     // G <= 14
constexpr int kSmemPeepTab = 3 * NC * kMaxP * (int)sizeof(float);   // 37632
constexpr int kSmemBytesMulti = kSmemW + 2 * kSmemSlab + kSmemPeepTab + 256 + 1024;

__global__ void __launch_bounds__(NTHREADS_MULTI, 1)
lstm_seq_multi_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmW,
                      const rgp_lstm_seq_args a, const int slab_rows, const int R) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;
  uint8_t* slab = smem + kSmemW;                          // [buf][kc]
  float* peep = reinterpret_cast<float*>(smem + kSmemW + 2 * kSmemSlab);   // [gate][channel][position]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemW + 2 * kSmemSlab + kSmemPeepTab);
  uint64_t* wfull = bars;
  uint64_t* fullA = bars + 1;                             // [buf][kMaxKch]
  uint64_t* slab_free = fullA + 2 * kMaxKch;              // [buf]: the MMAs that read the buffer have completed
  uint64_t* acc_full = slab_free + 2;                     // [buf]
  uint64_t* acc_free = acc_full + 2;                      // [buf]: the epilogue has read the accumulator out of TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ch = a.Ch, G = a.G, Wp = G + 1, S = Wp * Wp, P = G * G;
  const int nsplit = Ch / NC;
  const int grp = (int)(blockIdx.x / nsplit), split = (int)(blockIdx.x % nsplit);
  const int ch0 = split * NC;
  const int M = a.B * S;
  const int mtiles = (M + BM - 1) / BM;
  const int ntl = (mtiles - grp + R - 1) / R;             // tiles of this CTA: grp, grp + R, ...
  const int kch = Ch / BK;
  int* const err = a.flags + mtiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmH) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
    mbar_init(wfull, 1);
    for (int i = 0; i < 2 * kMaxKch; ++i) mbar_init(&fullA[i], 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&slab_free[i], 1); mbar_init(&acc_full[i], 1); mbar_init(&acc_free[i], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(128u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // peephole table of this CTA's channels
  for (int i = threadIdx.x; i < 3 * NC * P; i += NTHREADS_MULTI) {
    const int g = i / (NC * P), r = i - g * (NC * P), j = r / P, p = r - j * P;
    const float* pw = g == 0 ? a.wci : (g == 1 ? a.wcf : a.wco);
    peep[(g * NC + j) * kMaxP + p] = pw[(long long)p * Ch + ch0 + j];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== dependency wait + TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(wfull, (uint32_t)(9 * kch * kWTileBytes));
      for (int tap = 0; tap < 9; ++tap)
        for (int kc = 0; kc < kch; ++kc)
          for (int g = 0; g < 3; ++g)
            tma_load_2d(wsm + (tap * kch + kc) * kWTileBytes + g * (NC * BK * 2), &tmW, wfull, kc * BK,
                        tap * 3 * Ch + g * Ch + ch0);
    }
    bool dead = false;
    long long it = 0;
    for (int t = 0; t < a.T; ++t) {
      for (int k = 0; k < ntl; ++k, ++it) {
        const int tile = grp + k * R;
        const int buf = (int)(it & 1);
        const int nb = tile - 1 + lane;
        const bool watch = lane < 3 && nb >= 0 && nb < mtiles;
        if (t > 0 && !dead) {
          const int need = nsplit * t;
          const long long t0 = clock64();
          for (;;) {
            const bool ok = !watch || ld_acquire(a.flags + nb) >= need;
            if (__all_sync(0xffffffffu, ok)) break;
            bool bad = false;
            if (lane == 0) {
              bad = ld_acquire(err) != 0;
              if (!bad && clock64() - t0 > kSpinLimit) { atomicExch(err, 1); bad = true; }
            }
            if (__any_sync(0xffffffffu, bad)) { dead = true; break; }
          }
        }
        if (lane == 0) {
          if (it >= 2) mbar_wait(&slab_free[buf], (uint32_t)(((it >> 1) - 1) & 1));   // MMAs of item it-2 are done
          fence_proxy_async();
          for (int kc = 0; kc < kch; ++kc) {
            mbar_expect_tx(&fullA[buf * kMaxKch + kc], (uint32_t)(slab_rows * BK * 2));
            tma_load_2d(slab + (buf * kMaxKch + kc) * kSlabBytes, &tmH, &fullA[buf * kMaxKch + kc], kc * BK,
                        t * M + tile * BM);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NG >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mbar_wait(wfull, 0);
    const long long nitems = (long long)a.T * ntl;
    for (long long it = 0; it < nitems; ++it) {
      const int buf = (int)(it & 1);
      const uint32_t par = (uint32_t)((it >> 1) & 1);
      if (it >= 2) mbar_wait(&acc_free[buf], par ^ 1);        // the epilogue of item it-2 has read this accumulator
      tc_fence_after();
      const uint32_t tacc = tmem_base + (uint32_t)(buf * 64);
      for (int kc = 0; kc < kch; ++kc) {
        mbar_wait(&fullA[buf * kMaxKch + kc], par);
        tc_fence_after();
        if (lane == 0) {
          for (int tap = 0; tap < 9; ++tap) {
            const int off = (tap / 3) * Wp + (tap % 3);
            const uint64_t adesc = make_sdesc(smem_u32(slab + (buf * kMaxKch + kc) * kSlabBytes) + (uint32_t)off * (BK * 2));
            const uint64_t bdesc = make_sdesc(smem_u32(wsm + (tap * kch + kc) * kWTileBytes));
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk)
              tc_mma_bf16(tacc, adesc + (uint64_t)(2 * kk), bdesc + (uint64_t)(2 * kk), idesc, (kc | tap | kk) ? 1u : 0u);
          }
          if (kc == kch - 1) { tc_commit(&slab_free[buf]); tc_commit(&acc_full[buf]); }
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue warps 2..5: items in order, accumulator buffer = it & 1 =====================
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int lrow = lg * 32 + lane;
    const int fourCh = 4 * Ch;
    const long long nitems = (long long)a.T * ntl;
    for (long long it = 0; it < nitems; ++it) {
      const int t = (int)(it / ntl), k = (int)(it - (long long)t * ntl);
      const int tile = grp + k * R;
      const int buf = (int)(it & 1);
      const int m = tile * BM + lrow;
      const bool inrange = m < M;
      const int f = m / S, q = m - f * S;
      const int y = q / Wp, x = q - y * Wp;
      const bool valid = inrange && y < G && x < G;
      const long long prow = (long long)m + Wp + 1;
      const long long mt = (long long)t * M + m;
      const int p = y * G + x;
      float xi[NC], xf[NC], xc[NC], xo[NC], c[NC];
      if (valid) {
        const float* xr = a.Xg + mt * fourCh + ch0;
        ld8(xr, xi); ld8(xr + 8, xi + 8);
        ld8(xr + Ch, xf); ld8(xr + Ch + 8, xf + 8);
        ld8(xr + 2 * Ch, xc); ld8(xr + 2 * Ch + 8, xc + 8);
        ld8(xr + 3 * Ch, xo); ld8(xr + 3 * Ch + 8, xo + 8);
      }
      // the cell of this row: stored by this same thread one step ago (program order)
      if (valid) { ld8(a.Cf + prow * Ch + ch0, c); ld8(a.Cf + prow * Ch + ch0 + 8, c + 8); }
      {   // the NEXT item: pull its pre-activation row from HBM into L2 now
        const long long it2 = it + 1;
        if (it2 < nitems) {
          const int t2 = (int)(it2 / ntl), k2 = (int)(it2 - (long long)t2 * ntl);
          const int m2 = (grp + k2 * R) * BM + lrow;
          if (m2 < M) {
            const float* x2 = a.Xg + ((long long)t2 * M + m2) * fourCh + ch0;
#pragma unroll
            for (int g = 0; g < 4; ++g) asm volatile("prefetch.global.L2 [%0];" ::"l"(x2 + g * Ch));
          }
        }
      }
      mbar_wait(&acc_full[buf], (uint32_t)((it >> 1) & 1));
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * 64);
      float hi[NC], hf[NC], ho[NC];
      tc_ld16(tlane, hi);
      tc_ld16(tlane + 16u, hf);
      tc_ld16(tlane + 32u, ho);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_free[buf]);
      float gi[NC], gf[NC], gg[NC], go[NC], hn[NC], cp[NC];
      long long rn = 0;
      if (valid) {
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          cp[j] = c[j];
          gi[j] = sigm(xi[j] + hi[j] + peep[(0 * NC + j) * kMaxP + p] * cp[j]);
          gf[j] = sigm(xf[j] + hf[j] + peep[(1 * NC + j) * kMaxP + p] * cp[j]);
          gg[j] = tanh_fast(xc[j] + hi[j]);
          c[j] = gf[j] * cp[j] + gi[j] * gg[j];
          go[j] = sigm(xo[j] + ho[j] + peep[(2 * NC + j) * kMaxP + p] * cp[j]);
          hn[j] = tanh_fast(c[j]) * go[j];
        }
        rn = ((long long)(t + 1) * M + prow) * Ch + ch0;
        stbf16x16((TA*)a.Hop + rn, hn);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(a.flags + tile, 1); }
      if (valid) {
        st8(a.Cf + prow * Ch + ch0, c); st8(a.Cf + prow * Ch + ch0 + 8, c + 8);
        st8(a.Hf + rn, hn); st8(a.Hf + rn + 8, hn + 8);
        if (a.HS) {
          const long long orow = ((long long)t * a.B + f) * a.So + (long long)(y + a.pado) * a.Wo + (x + a.pado);
          stbf16x16((TA*)a.HS + orow * Ch + ch0, hn);
        }
        if (a.save) {
          float* sv = a.save + mt * 5 * Ch + ch0;
          st8(sv, gi); st8(sv + 8, gi + 8);
          st8(sv + Ch, gf); st8(sv + Ch + 8, gf + 8);
          st8(sv + 2 * Ch, gg); st8(sv + 2 * Ch + 8, gg + 8);
          st8(sv + 3 * Ch, go); st8(sv + 3 * Ch + 8, go + 8);
          st8(sv + 4 * Ch, cp); st8(sv + 4 * Ch + 8, cp + 8);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Small-batch persistent ConvGRU (gaze_grcn.py:95-129 unrolled by :259-288) with the same structure: CTA (tile, split) owns
// 128 q-rows x 16 channels, its slices of U_z | U_r (9 taps x Ch/64 x 32 rows) and U (x 16 rows) - 110 KB at hidden 128, the
// reference's own width (gaze_grcn.py:211) - stay in shared memory, the state h and the update gate z of its rows live in
// registers.  A GRU step has TWO exchanges (r.h must be complete on the halo before the candidate's conv): phase ZR
// (N = 32) -> r.h rows published through flags[0..mtiles) ; phase C (N = 16) -> h_t rows published through
// flags[mtiles+1..2*mtiles+1).  Serves the batches the tile-independent gru_seq kernel is too coarse for (the reference
// trains at 28 clips: 14 tiles x 8 splits = 112 CTAs; there the per-step path is two launches of ~12 us per step).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kZRTileBytes = 2 * NC * BK * 2;     // 4096: z rows | r rows of one (tap, k-chunk)
constexpr int kCTileBytes = NC * BK * 2;          // 2048
constexpr int kSmemGruW = 9 * kMaxKch * (kZRTileBytes + kCTileBytes);   // 110592
constexpr int kSmemBytesGru = kSmemGruW + kSmemSlab + 256 + 1024;

__device__ __forceinline__ bool wait_tiles(const int* flags, int tile, int mtiles, int need, int* err, int lane, bool& dead) {
  // lanes 0..2 watch tiles (tile-1, tile, tile+1); returns when all show `need` (or the launch has been declared dead)
  const int nb = tile - 1 + lane;
  const bool watch = lane < 3 && nb >= 0 && nb < mtiles;
  if (dead) return false;
  const long long t0 = clock64();
  for (;;) {
    const bool ok = !watch || ld_acquire(flags + nb) >= need;
    if (__all_sync(0xffffffffu, ok)) return true;
    bool bad = false;
    if (lane == 0) {
      bad = ld_acquire(err) != 0;
      if (!bad && clock64() - t0 > kSpinLimit) { atomicExch(err, 1); bad = true; }
    }
    if (__any_sync(0xffffffffu, bad)) { dead = true; return false; }
  }
}

__global__ void __launch_bounds__(NTHREADS, 1)
gru_small_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmRH,
                 const __grid_constant__ CUtensorMap tmUzr, const __grid_constant__ CUtensorMap tmUc,
                 const rgp_gru_seq_args a, int* const flags, const int slab_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* wzr = smem;                                            // [tap][kc] 32 rows x 128 B
  uint8_t* wc = smem + 9 * kMaxKch * kZRTileBytes;                // [tap][kc] 16 rows x 128 B
  uint8_t* slab = smem + kSmemGruW;                               // [kc] slab_rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemGruW + kSmemSlab);
  uint64_t* wfull = bars;
  uint64_t* fullH = bars + 1;                                     // [kMaxKch] h_{t-1} slab
  uint64_t* fullR = fullH + kMaxKch;                              // [kMaxKch] r.h slab
  uint64_t* zr_full = fullR + kMaxKch;
  uint64_t* c_full = zr_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ch = a.Ch, G = a.G, Wp = G + 1, S = Wp * Wp;
  const int nsplit = Ch / NC;
  const int tile = (int)(blockIdx.x / nsplit), split = (int)(blockIdx.x % nsplit);
  const int ch0 = split * NC;
  const int M = a.B * S;
  const int mtiles = (M + BM - 1) / BM;
  const int m0 = tile * BM;
  const int kch = Ch / BK;
  int* const flagsR = flags;                   // [mtiles]  r.h of step t published
  int* const err = flags + mtiles;
  int* const flagsH = flags + mtiles + 1;      // [mtiles]  h_t published
  const bool train = a.Rs != nullptr;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmH) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmRH) : "memory");
    mbar_init(wfull, 1);
    for (int i = 0; i < kMaxKch; ++i) { mbar_init(&fullH[i], 1); mbar_init(&fullR[i], 1); }
    mbar_init(zr_full, 1); mbar_init(c_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== dependency waits + TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(wfull, (uint32_t)(9 * kch * (kZRTileBytes + kCTileBytes)));
      for (int tap = 0; tap < 9; ++tap)
        for (int kc = 0; kc < kch; ++kc) {
          for (int g = 0; g < 2; ++g)
            tma_load_2d(wzr + (tap * kch + kc) * kZRTileBytes + g * (NC * BK * 2), &tmUzr, wfull, kc * BK,
                        tap * 2 * Ch + g * Ch + ch0);
          tma_load_2d(wc + (tap * kch + kc) * kCTileBytes, &tmUc, wfull, kc * BK, tap * Ch + ch0);
        }
    }
    bool dead = false;
    for (int t = 0; t < a.T; ++t) {
      if (t > 0) wait_tiles(flagsH, tile, mtiles, nsplit * t, err, lane, dead);
      if (lane == 0) {
        fence_proxy_async();
        for (int kc = 0; kc < kch; ++kc) {
          mbar_expect_tx(&fullH[kc], (uint32_t)(slab_rows * BK * 2));
          tma_load_2d(slab + kc * kSlabBytes, &tmH, &fullH[kc], kc * BK, t * M + m0);
        }
      }
      __syncwarp();
      // r.h of this step: complete on the tile and both neighbours (their ZR MMAs, which read the h slab, are done too)
      wait_tiles(flagsR, tile, mtiles, nsplit * (t + 1), err, lane, dead);
      if (lane == 0) {
        fence_proxy_async();
        for (int kc = 0; kc < kch; ++kc) {
          mbar_expect_tx(&fullR[kc], (uint32_t)(slab_rows * BK * 2));
          tma_load_2d(slab + kc * kSlabBytes, &tmRH, &fullR[kc], kc * BK, (train ? t : 0) * M + m0);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_zr = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)((2 * NC) >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    constexpr uint32_t idesc_c = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NC >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mbar_wait(wfull, 0);
    for (int t = 0; t < a.T; ++t) {
      const uint32_t par = (uint32_t)(t & 1);
      for (int ph = 0; ph < 2; ++ph) {
        for (int kc = 0; kc < kch; ++kc) {
          mbar_wait(ph == 0 ? &fullH[kc] : &fullR[kc], par);
          tc_fence_after();
          if (lane == 0) {
            for (int tap = 0; tap < 9; ++tap) {
              const int off = (tap / 3) * Wp + (tap % 3);
              const uint64_t adesc = make_sdesc(smem_u32(slab + kc * kSlabBytes) + (uint32_t)off * (BK * 2));
              const uint64_t bdesc = make_sdesc(smem_u32(ph == 0 ? wzr + (tap * kch + kc) * kZRTileBytes
                                                                 : wc + (tap * kch + kc) * kCTileBytes));
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                tc_mma_bf16(tmem_base + (uint32_t)(ph == 0 ? 0 : 32), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k),
                            ph == 0 ? idesc_zr : idesc_c, (kc | tap | k) ? 1u : 0u);
            }
            if (kc == kch - 1) tc_commit(ph == 0 ? zr_full : c_full);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ===================== epilogue warps 2..5: one thread = one q-row, NC channels; h and z in registers ==========
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int lrow = lg * 32 + lane;
    const int m = m0 + lrow;
    const bool inrange = m < M;
    const int f = m / S, q = m - f * S;
    const int y = q / Wp, x = q - y * Wp;
    const bool valid = inrange && y < G && x < G;
    const long long prow = (long long)m + Wp + 1;
    const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16);
    const int threeCh = 3 * Ch;
    float h[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) h[j] = 0.f;
    if (valid) { ld8(a.Hf + prow * Ch + ch0, h); ld8(a.Hf + prow * Ch + ch0 + 8, h + 8); }      // h_0 (slab 0)
    for (int t = 0; t < a.T; ++t) {
      const long long mt = (long long)t * M + m;
      float xz[NC], xr[NC], xc[NC];
      if (valid) {
        const float* xp = a.Xg + mt * threeCh + ch0;
        ld8(xp, xz); ld8(xp + 8, xz + 8);
        ld8(xp + Ch, xr); ld8(xp + Ch + 8, xr + 8);
        ld8(xp + 2 * Ch, xc); ld8(xp + 2 * Ch + 8, xc + 8);
      }
      // ---------------- phase ZR: z, r ; r.h rows -> the candidate conv's operand ----------------
      mbar_wait(zr_full, (uint32_t)(t & 1));
      tc_fence_after();
      float z[NC], r[NC];
      tc_ld16(tlane, z);
      tc_ld16(tlane + 16u, r);
      tc_fence_before();
      if (valid) {
        float rh[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          z[j] = sigm(z[j] + xz[j]);
          r[j] = sigm(r[j] + xr[j]);
          rh[j] = r[j] * h[j];
        }
        stbf16x16((TA*)a.RH + ((long long)(train ? t : 0) * M + prow) * Ch + ch0, rh);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flagsR + tile, 1); }
      if (train && valid) {
        st8(a.Zs + mt * Ch + ch0, z); st8(a.Zs + mt * Ch + ch0 + 8, z + 8);
        st8(a.Rs + mt * Ch + ch0, r); st8(a.Rs + mt * Ch + ch0 + 8, r + 8);
      }
      // ---------------- phase C: candidate, state update ----------------
      mbar_wait(c_full, (uint32_t)(t & 1));
      tc_fence_after();
      float c[NC];
      tc_ld16(tlane + 32u, c);
      tc_fence_before();
      long long rn = 0;
      if (valid) {
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          c[j] = tanh_fast(c[j] + xc[j]);
          h[j] = z[j] * h[j] + (1.0f - z[j]) * c[j];
        }
        rn = ((long long)(t + 1) * M + prow) * Ch + ch0;
        stbf16x16((TA*)a.Hop + rn, h);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flagsH + tile, 1); }
      if (valid) {        // tail: fp32 master, saved candidate, BN-scaled head rows (gaze_grcn.py:325)
        st8(a.Hf + rn, h); st8(a.Hf + rn + 8, h + 8);
        if (train) { st8(a.Cs + mt * Ch + ch0, c); st8(a.Cs + mt * Ch + ch0 + 8, c + 8); }
        if (a.HS) {
          float v[NC];
          if (a.bn_gamma) {
            float g8[NC], b8[NC];
            ld8(a.bn_gamma + (long long)t * Ch + ch0, g8); ld8(a.bn_gamma + (long long)t * Ch + ch0 + 8, g8 + 8);
            ld8(a.bn_beta + (long long)t * Ch + ch0, b8); ld8(a.bn_beta + (long long)t * Ch + ch0 + 8, b8 + 8);
#pragma unroll
            for (int j = 0; j < NC; ++j) v[j] = g8[j] * (h[j] * a.alpha) + b8[j];
          } else {
#pragma unroll
            for (int j = 0; j < NC; ++j) v[j] = h[j];
          }
          const long long orow = ((long long)t * a.B + f) * a.So + (long long)(y + a.pado) * a.Wo + (x + a.pado);
          stbf16x16((TA*)a.HS + orow * Ch + ch0, v);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Small-batch persistent ConvGRU BPTT (tf.gradients of gaze_grcn.py:108-127, base.py:278-281): the reverse-time mirror of
// gru_small_kernel, all T steps in one launch.  CTA (tile, split) owns 128 q-rows x 16 INPUT channels of the two
// transposed convolutions; its slices of U^T (9 taps x Ch/64 x 16 rows) and [U_z|U_r]^T (9 x 2Ch/64 x 16 rows) - 110 KB at
// hidden 128 - stay in shared memory, dL/dh's non-conv part lives in registers.  Step t (t = T-1 .. 0):
//   phase RH : acc = conv^T(dpre_c_t ; U)       -> dpre_r_t = acc h_{t-1} r (1-r)  (dG_t, r columns) ; dhpart += acc r
//   phase H  : acc = conv^T([dpre_z_t, dpre_r_t] ; U_z|U_r) -> dh = acc + dhpart + head term ; gate backward of step t-1:
//              dpre_z_{t-1}, dpre_c_{t-1} (dG_{t-1}) ; dhpart = dh z
// Both operands are columns of the q-layout gradient buffer dG [T*M, 3Ch] (z | r | c), written 16 channels at a time by the
// CTAs and exchanged through per-tile counters exactly like the forward kernel: flags[0..mtiles) = "dpre_z / dpre_c of the
// step this iteration needs are complete", flags[mtiles+1..) = "dpre_r complete".  The gate backward of the last step
// (rgp_gru_bwd_last) is the kernel's prologue.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kBwdTileBytes = NC * BK * 2;                               // 2048: 16 rows of one (tap, k-chunk)
constexpr int kSmemBwdW = 9 * 3 * kMaxKch * kBwdTileBytes;               // U^T: kch chunks, [Uz|Ur]^T: 2*kch chunks -> 110592
constexpr int kSmemBytesGruBwd = kSmemBwdW + 2 * kSmemSlab + 256 + 1024;  // slab: up to 2*kch k-chunks

__global__ void __launch_bounds__(NTHREADS, 1)
gru_small_bwd_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmUcT,
                     const __grid_constant__ CUtensorMap tmUzrT, const rgp_gru_bwd_args a, int* const flags,
                     const int slab_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int Ch = a.Ch, G = a.G, Wp = G + 1, S = Wp * Wp;
  const int kch = Ch / BK;
  uint8_t* wc = smem;                                             // [tap][kc < kch]    16 rows x 128 B
  uint8_t* wzr = smem + 9 * kch * kBwdTileBytes;                  // [tap][kc < 2*kch]
  uint8_t* slab = smem + kSmemBwdW;                               // [kc < 2*kch] slab_rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemBwdW + 2 * kSmemSlab);
  uint64_t* wfull = bars;
  uint64_t* fullC = bars + 1;                                     // [kMaxKch]     dpre_c slab
  uint64_t* fullZR = fullC + kMaxKch;                             // [2*kMaxKch]   dpre_z | dpre_r slab
  uint64_t* rh_full = fullZR + 2 * kMaxKch;
  uint64_t* h_full = rh_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nsplit = Ch / NC;
  const int tile = (int)(blockIdx.x / nsplit), split = (int)(blockIdx.x % nsplit);
  const int ch0 = split * NC;
  const int M = a.B * S;
  const int mtiles = (M + BM - 1) / BM;
  const int m0 = tile * BM;
  int* const flagsC = flags;
  int* const err = flags + mtiles;
  int* const flagsR = flags + mtiles + 1;
  const int T = a.T;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmG) : "memory");
    mbar_init(wfull, 1);
    for (int i = 0; i < kMaxKch; ++i) mbar_init(&fullC[i], 1);
    for (int i = 0; i < 2 * kMaxKch; ++i) mbar_init(&fullZR[i], 1);
    mbar_init(rh_full, 1); mbar_init(h_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== dependency waits + TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(wfull, (uint32_t)(9 * 3 * kch * kBwdTileBytes));
      for (int tap = 0; tap < 9; ++tap) {
        for (int kc = 0; kc < kch; ++kc)
          tma_load_2d(wc + (tap * kch + kc) * kBwdTileBytes, &tmUcT, wfull, kc * BK, tap * Ch + ch0);
        for (int kc = 0; kc < 2 * kch; ++kc)
          tma_load_2d(wzr + (tap * 2 * kch + kc) * kBwdTileBytes, &tmUzrT, wfull, kc * BK, tap * Ch + ch0);
      }
    }
    bool dead = false;
    for (int it = 0; it < T; ++it) {
      const int t = T - 1 - it;
      const int row0 = t * M + m0 - (Wp + 1);          // the transposed taps reach Wp+1 rows back (negative: zero filled)
      wait_tiles(flagsC, tile, mtiles, nsplit * (it + 1), err, lane, dead);
      if (lane == 0) {
        fence_proxy_async();
        for (int kc = 0; kc < kch; ++kc) {
          mbar_expect_tx(&fullC[kc], (uint32_t)(slab_rows * BK * 2));
          tma_load_2d(slab + kc * kSlabBytes, &tmG, &fullC[kc], 2 * Ch + kc * BK, row0);
        }
      }
      __syncwarp();
      if (t == 0) break;                                // dL/dh_0 is not needed
      wait_tiles(flagsR, tile, mtiles, nsplit * (it + 1), err, lane, dead);
      if (lane == 0) {
        fence_proxy_async();
        for (int kc = 0; kc < 2 * kch; ++kc) {
          mbar_expect_tx(&fullZR[kc], (uint32_t)(slab_rows * BK * 2));
          tma_load_2d(slab + kc * kSlabBytes, &tmG, &fullZR[kc], kc * BK, row0);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NC >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mbar_wait(wfull, 0);
    for (int it = 0; it < T; ++it) {
      const int t = T - 1 - it;
      const uint32_t par = (uint32_t)(it & 1);
      for (int ph = 0; ph < 2; ++ph) {
        if (ph == 1 && t == 0) break;
        const int nk = ph == 0 ? kch : 2 * kch;
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(ph == 0 ? &fullC[kc] : &fullZR[kc], par);
          tc_fence_after();
          if (lane == 0) {
            for (int tap = 0; tap < 9; ++tap) {
              const int off = 2 * (Wp + 1) - ((tap / 3) * Wp + (tap % 3));      // btap + (Wp+1), btap = Wp+1 - o
              const uint64_t adesc = make_sdesc(smem_u32(slab + kc * kSlabBytes) + (uint32_t)off * (BK * 2));
              const uint64_t bdesc = make_sdesc(smem_u32(ph == 0 ? wc + (tap * kch + kc) * kBwdTileBytes
                                                                 : wzr + (tap * 2 * kch + kc) * kBwdTileBytes));
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                tc_mma_bf16(tmem_base + (uint32_t)(ph * 16), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                            (kc | tap | k) ? 1u : 0u);
            }
            if (kc == nk - 1) tc_commit(ph == 0 ? rh_full : h_full);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ===================== epilogue warps 2..5: one thread = one q-row, 16 channels =====================
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int lrow = lg * 32 + lane;
    const int m = m0 + lrow;
    const bool inrange = m < M;
    const int f = m / S, q = m - f * S;
    const int y = q / Wp, x = q - y * Wp;
    const bool valid = inrange && y < G && x < G;
    const long long prow = (long long)m + Wp + 1;
    const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16);
    const int threeCh = 3 * Ch;
    (void)f;
    auto ld16 = [&](const float* p, float* v) { ld8(p, v); ld8(p + 8, v + 8); };
    float dhp[NC], hp[NC];          // dhpart ; h_{t-1} of the step being differentiated (Hf slab t)
#pragma unroll
    for (int j = 0; j < NC; ++j) { dhp[j] = 0.f; hp[j] = 0.f; }
    // gate backward of one step given dh: writes dpre_z / dpre_c of step `ts` (its input state in hp_), dhpart = dh z
    auto gates = [&](int ts, const float* dh, const float* hp_) {
      const long long mt = (long long)ts * M + m;
      float zz[NC], cc[NC], dz[NC], dc[NC];
      ld16(a.Zs + mt * Ch + ch0, zz);
      ld16(a.Cs + mt * Ch + ch0, cc);
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        dz[j] = dh[j] * (hp_[j] - cc[j]) * zz[j] * (1.0f - zz[j]);
        dc[j] = dh[j] * (1.0f - zz[j]) * (1.0f - cc[j] * cc[j]);
        dhp[j] = dh[j] * zz[j];
      }
      stbf16x16((TA*)a.dG + mt * threeCh + ch0, dz);
      stbf16x16((TA*)a.dG + mt * threeCh + 2 * Ch + ch0, dc);
    };
    auto head_term = [&](int ts, float* hd) {          // dL/dh_{ts+1} through the head: dhs (x gamma alpha with the frozen BN)
      ld16(a.dhs + ((long long)ts * M + m) * Ch + ch0, hd);
      if (a.bn_gamma) {
        float g16[NC];
        ld16(a.bn_gamma + (long long)ts * Ch + ch0, g16);
#pragma unroll
        for (int j = 0; j < NC; ++j) hd[j] *= g16[j] * a.alpha;
      }
    };
    // ---- prologue = rgp_gru_bwd_last: step T-1 has no recurrent gradient
    if (valid) {
      float dh[NC];
      head_term(T - 1, dh);
      ld16(a.Hf + ((long long)(T - 1) * M + prow) * Ch + ch0, hp);
      gates(T - 1, dh, hp);
    }
    fence_proxy_async();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flagsC + tile, 1); }
    for (int it = 0; it < T; ++it) {
      const int t = T - 1 - it;
      const long long mt = (long long)t * M + m;
      float rr[NC], hd[NC], hp1[NC];
      if (valid) {
        ld16(a.Rs + mt * Ch + ch0, rr);
        if (t > 0) {
          head_term(t - 1, hd);
          ld16(a.Hf + ((long long)(t - 1) * M + prow) * Ch + ch0, hp1);
        }
      }
      // ---------------- phase RH ----------------
      mbar_wait(rh_full, (uint32_t)(it & 1));
      tc_fence_after();
      float acc[NC];
      tc_ld16(tlane, acc);
      tc_fence_before();
      if (valid) {
        float dr[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          dr[j] = acc[j] * hp[j] * rr[j] * (1.0f - rr[j]);
          dhp[j] = fmaf(acc[j], rr[j], dhp[j]);
        }
        stbf16x16((TA*)a.dG + mt * threeCh + Ch + ch0, dr);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flagsR + tile, 1); }
      if (t == 0) break;
      // ---------------- phase H: full dL/dh_t, then the gates of step t-1 ----------------
      mbar_wait(h_full, (uint32_t)(it & 1));
      tc_fence_after();
      tc_ld16(tlane + 16u, acc);
      tc_fence_before();
      if (valid) {
        float dh[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) { dh[j] = acc[j] + dhp[j] + hd[j]; hp[j] = hp1[j]; }
        gates(t - 1, dh, hp);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flagsC + tile, 1); }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Small-batch persistent ConvLSTM BPTT (tf.gradients of gaze_lstm.py:112-133 as written), the reverse-time mirror of
// lstm_seq_kernel: all T steps in one launch.  CTA (tile, split) owns 128 q-rows x 16 channels: the gate backward of its
// channels (rgp_lstm_gates_bwd's arithmetic) with dL/dc in registers and the peephole gradients accumulated per row in
// shared memory for the whole sequence (3 x 16 atomics per row at the END instead of per step), and 16 INPUT channels of
// the transposed hidden convolution dh_{t-1} = conv^T(dHc_t ; [W_hi|W_hf|W_ho]), K = 3Ch, whose weight slice (9 taps x
// 3Ch/64 x 16 rows = 110 KB at hidden 128) is resident.  dHc_t [M, 3Ch] is exchanged through the per-tile counters; its
// slab is loaded in halves of three k-chunks (shared memory: 110 KB weights + 60 KB slab + 48 KB peepholes / gradients).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kLbHalfKc = 3;
constexpr int kSmemLbW = 9 * 3 * kMaxKch * kBwdTileBytes;                 // 110592
constexpr int kSmemLbSlab = kLbHalfKc * kSlabBytes;                       // 61440
constexpr int kSmemBytesLstmBwd = kSmemLbW + kSmemLbSlab + 2 * kSmemPeep + 256 + 1024;

__global__ void __launch_bounds__(NTHREADS, 1)
lstm_small_bwd_kernel(const __grid_constant__ CUtensorMap tmHc, const __grid_constant__ CUtensorMap tmWhT,
                      const rgp_lstm_bwd_args a, int* const flags, const int slab_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int Ch = a.Ch, G = a.G, Wp = G + 1, S = Wp * Wp, T = a.T;
  const int kch3 = 3 * Ch / BK;                                  // k-chunks of K = 3Ch
  const int nh = (kch3 + kLbHalfKc - 1) / kLbHalfKc;             // slab halves per step
  uint8_t* wsm = smem;                                           // [tap][kc < kch3] 16 rows x 128 B
  uint8_t* slab = smem + kSmemLbW;                               // [kc < 3]
  float* peep = reinterpret_cast<float*>(smem + kSmemLbW + kSmemLbSlab);   // [gate][channel][row]
  float* gacc = peep + 3 * NC * BM;                              // peephole gradients of this row, same layout
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemLbW + kSmemLbSlab + 2 * kSmemPeep);
  uint64_t* wfull = bars;
  uint64_t* fullA = bars + 1;                                    // [3]
  uint64_t* half_free = fullA + kLbHalfKc;
  uint64_t* acc_full = half_free + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nsplit = Ch / NC;
  const int tile = (int)(blockIdx.x / nsplit), split = (int)(blockIdx.x % nsplit);
  const int ch0 = split * NC;
  const int M = a.B * S;
  const int mtiles = (M + BM - 1) / BM;
  const int m0 = tile * BM;
  int* const err = flags + mtiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmHc) : "memory");
    mbar_init(wfull, 1);
    for (int i = 0; i < kLbHalfKc; ++i) mbar_init(&fullA[i], 1);
    mbar_init(half_free, 1); mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== dependency wait + TMA producer (steps T-1 .. 1 need the transposed conv) =====================
    if (lane == 0) {
      mbar_expect_tx(wfull, (uint32_t)(9 * kch3 * kBwdTileBytes));
      for (int tap = 0; tap < 9; ++tap)
        for (int kc = 0; kc < kch3; ++kc)
          tma_load_2d(wsm + (tap * kch3 + kc) * kBwdTileBytes, &tmWhT, wfull, kc * BK, tap * Ch + ch0);
    }
    bool dead = false;
    long long u = 0;                                   // slab fills so far
    for (int it = 0; it + 1 < T; ++it) {
      const int t = T - 1 - it;
      const int row0 = t * M + m0 - (Wp + 1);
      wait_tiles(flags, tile, mtiles, nsplit * (it + 1), err, lane, dead);
      for (int hh = 0; hh < nh; ++hh, ++u) {
        if (lane == 0) {
          if (u > 0) mbar_wait(half_free, (uint32_t)((u - 1) & 1));       // the MMAs that read the previous fill are done
          fence_proxy_async();
          const int nk = min(kLbHalfKc, kch3 - hh * kLbHalfKc);
          for (int kc = 0; kc < nk; ++kc) {
            mbar_expect_tx(&fullA[kc], (uint32_t)(slab_rows * BK * 2));
            tma_load_2d(slab + kc * kSlabBytes, &tmHc, &fullA[kc], (hh * kLbHalfKc + kc) * BK, row0);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NC >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mbar_wait(wfull, 0);
    long long u = 0;
    for (int it = 0; it + 1 < T; ++it) {
      for (int hh = 0; hh < nh; ++hh, ++u) {
        const int nk = min(kLbHalfKc, kch3 - hh * kLbHalfKc);
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(&fullA[kc], (uint32_t)(u & 1));
          tc_fence_after();
          if (lane == 0) {
            for (int tap = 0; tap < 9; ++tap) {
              const int off = 2 * (Wp + 1) - ((tap / 3) * Wp + (tap % 3));
              const uint64_t adesc = make_sdesc(smem_u32(slab + kc * kSlabBytes) + (uint32_t)off * (BK * 2));
              const uint64_t bdesc = make_sdesc(smem_u32(wsm + (tap * kch3 + hh * kLbHalfKc + kc) * kBwdTileBytes));
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                tc_mma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (hh | kc | tap | k) ? 1u : 0u);
            }
            if (kc == nk - 1) {
              tc_commit(half_free);
              if (hh == nh - 1) tc_commit(acc_full);
            }
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ===================== epilogue warps 2..5: gate backward, one thread = one q-row x 16 channels =====================
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int lrow = lg * 32 + lane;
    const int m = m0 + lrow;
    const bool inrange = m < M;
    const int f = m / S, q = m - f * S;
    const int y = q / Wp, x = q - y * Wp;
    const bool valid = inrange && y < G && x < G;
    const long long prow = (long long)m + Wp + 1;
    const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16);
    (void)f;
    auto ld16 = [&](const float* p, float* v) { ld8(p, v); ld8(p + 8, v + 8); };
    const int pidx = y * G + x;
    {
      const float* pw[3] = {a.wci, a.wcf, a.wco};
#pragma unroll
      for (int g = 0; g < 3; ++g)
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          peep[(g * NC + j) * BM + lrow] = valid ? pw[g][(long long)pidx * Ch + ch0 + j] : 0.f;
          gacc[(g * NC + j) * BM + lrow] = 0.f;
        }
    }
    float dc[NC], cn[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) { dc[j] = 0.f; cn[j] = 0.f; }
    if (valid) ld16(a.Cf + prow * Ch + ch0, cn);                    // c_T: the cell the forward pass left behind
    for (int it = 0; it < T; ++it) {
      const int t = T - 1 - it;
      const long long mt = (long long)t * M + m;
      float gi[NC], gf[NC], gg[NC], go[NC], cp[NC], dh[NC];
      if (valid) {
        const float* sv = a.save + mt * 5 * Ch + ch0;
        ld16(sv, gi); ld16(sv + Ch, gf); ld16(sv + 2 * Ch, gg); ld16(sv + 3 * Ch, go); ld16(sv + 4 * Ch, cp);
        ld16(a.dhs + mt * Ch + ch0, dh);
      }
      if (it > 0) {                                                 // dL/dh_t through the hidden convolutions of step t+1
        mbar_wait(acc_full, (uint32_t)((it - 1) & 1));
        tc_fence_after();
        float acc[NC];
        tc_ld16(tlane, acc);
        tc_fence_before();
#pragma unroll
        for (int j = 0; j < NC; ++j) dh[j] += acc[j];
      }
      if (valid) {
        float dpi[NC], dpf[NC], dpg[NC], dpo[NC], dic[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) {
          const float tc = tanhf(cn[j]);
          const float dcn = dc[j] + dh[j] * go[j] * (1.0f - tc * tc);
          dpo[j] = dh[j] * tc * go[j] * (1.0f - go[j]);
          dpi[j] = dcn * gg[j] * gi[j] * (1.0f - gi[j]);
          dpg[j] = dcn * gi[j] * (1.0f - gg[j] * gg[j]);
          dpf[j] = dcn * cp[j] * gf[j] * (1.0f - gf[j]);
          dc[j] = dcn * gf[j] + dpi[j] * peep[(0 * NC + j) * BM + lrow] + dpf[j] * peep[(1 * NC + j) * BM + lrow] +
                  dpo[j] * peep[(2 * NC + j) * BM + lrow];
          gacc[(0 * NC + j) * BM + lrow] += dpi[j] * cp[j];
          gacc[(1 * NC + j) * BM + lrow] += dpf[j] * cp[j];
          gacc[(2 * NC + j) * BM + lrow] += dpo[j] * cp[j];
          dic[j] = dpi[j] + dpg[j];
          cn[j] = cp[j];                                            // c_{t-1} is the next iteration's c_t
        }
        TA* hc = (TA*)a.dHc + mt * 3 * Ch + ch0;
        stbf16x16(hc, dic); stbf16x16(hc + Ch, dpf); stbf16x16(hc + 2 * Ch, dpo);
        TA* g4 = (TA*)a.dG4 + mt * 4 * Ch + ch0;
        stbf16x16(g4, dpi); stbf16x16(g4 + Ch, dpf); stbf16x16(g4 + 2 * Ch, dpg); stbf16x16(g4 + 3 * Ch, dpo);
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) { __threadfence(); red_release_add(flags + tile, 1); }
    }
    if (valid) {
      float* gw[3] = {a.gwci, a.gwcf, a.gwco};
#pragma unroll
      for (int g = 0; g < 3; ++g)
#pragma unroll
        for (int j = 0; j < NC; ++j) atomicAdd(gw[g] + (long long)pidx * Ch + ch0 + j, gacc[(g * NC + j) * BM + lrow]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                 : "memory");
  }
}

int slab_rows_of(int G) { return (BM + 2 * (G + 1) + 2 + 7) & ~7; }

bool shape_ok(int B, int G, int Ch) {
  if (B <= 0 || G <= 0 || (Ch != 64 && Ch != 128)) return false;
  if (slab_rows_of(G) > kMaxSlabRows) return false;
  const int S = (G + 1) * (G + 1);
  const long long mtiles = ((long long)B * S + BM - 1) / BM;
  if (mtiles * (Ch / NC) <= num_sms()) return true;            // one (tile, split) CTA each
  // multi-tile variant: R = SMs / nsplit row groups walk the tiles; bounded so that a step stays short (<= 16 tiles per CTA)
  const int R = num_sms() / (Ch / NC);
  return R >= 1 && G <= 14 && mtiles <= 16LL * R;
}

bool gru_small_ok(int B, int G, int Ch) {
  if (B <= 0 || G <= 0 || (Ch != 64 && Ch != 128)) return false;
  if (slab_rows_of(G) > kMaxSlabRows) return false;
  const int S = (G + 1) * (G + 1);
  const long long mtiles = ((long long)B * S + BM - 1) / BM;
  return mtiles * (Ch / NC) <= num_sms();
}

}  // namespace lstmseq
using namespace lstmseq;

int convgru_seq_small_fwd(const rgp_gru_seq_args& a, int* flags, cudaStream_t s) {
  RGP_CHECK_ARG(a.dtype == 1, "rgp_convgru_seq_small_fwd: bf16 operands only (dtype=%d)", a.dtype);
  RGP_CHECK_ARG(a.T > 0 && gru_small_ok(a.B, a.G, a.Ch), "rgp_convgru_seq_small_fwd: B=%d G=%d Ch=%d T=%d does not fit (Ch 64 or "
                "128, G <= 14, tiles * Ch/16 <= SMs)", a.B, a.G, a.Ch, a.T);
  RGP_CHECK_ARG(a.Hop && a.Hf && a.RH && a.Uzr && a.Uc && a.Xg && flags, "rgp_convgru_seq_small_fwd: null pointer%s", "");
  RGP_CHECK_ARG((a.Rs == nullptr) == (a.Cs == nullptr) && (a.Rs == nullptr || a.Zs != nullptr),
                "rgp_convgru_seq_small_fwd: training needs Zs, Rs and Cs%s", "");
  RGP_CHECK_ARG((a.bn_gamma == nullptr) == (a.bn_beta == nullptr), "rgp_convgru_seq_small_fwd: bn_gamma / bn_beta%s", "");
  RGP_CHECK_ARG(((uintptr_t)a.Hop & 31) == 0 && ((uintptr_t)a.Hf & 31) == 0 && ((uintptr_t)a.RH & 31) == 0 &&
                ((uintptr_t)a.Xg & 31) == 0 && ((uintptr_t)a.HS & 31) == 0 && ((uintptr_t)a.Zs & 31) == 0 &&
                ((uintptr_t)a.Rs & 31) == 0 && ((uintptr_t)a.Cs & 31) == 0 && ((uintptr_t)a.bn_gamma & 31) == 0 &&
                ((uintptr_t)a.bn_beta & 31) == 0, "rgp_convgru_seq_small_fwd: buffers must be 32-byte aligned%s", "");
  const int Ch = a.Ch, S = (a.G + 1) * (a.G + 1);
  const int srows = slab_rows_of(a.G);
  CUtensorMap tmH, tmRH, tmUzr, tmUc;
  int rc = tc::encode_map_bf16(a.Hop, Ch, a.rowsH, (long long)Ch * 2, BK, srows, &tmH);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.RH, Ch, a.rowsRH, (long long)Ch * 2, BK, srows, &tmRH);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.Uzr, Ch, 18LL * Ch, (long long)Ch * 2, BK, NC, &tmUzr);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.Uc, Ch, 9LL * Ch, (long long)Ch * 2, BK, NC, &tmUc);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(gru_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesGru);
    if (e != cudaSuccess) { set_error("gru_small: cudaFuncSetAttribute(smem=%d): %s", kSmemBytesGru, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.B * S + BM - 1) / BM;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(mtiles * (Ch / NC)), 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = kSmemBytesGru;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gru_small_kernel, tmH, tmRH, tmUzr, tmUc, a, flags, srows);
  if (e != cudaSuccess) { set_error("gru_small launch: %s", cudaGetErrorString(e)); return -5; }
  return check_launch("gru_small");
}

int convlstm_seq_fwd(const rgp_lstm_seq_args& a, cudaStream_t s) {
  RGP_CHECK_ARG(a.dtype == 1, "rgp_convlstm_seq_fwd: bf16 operands only (dtype=%d)", a.dtype);
  RGP_CHECK_ARG(a.T > 0, "rgp_convlstm_seq_fwd: T=%d", a.T);
  RGP_CHECK_ARG(shape_ok(a.B, a.G, a.Ch), "rgp_convlstm_seq_fwd: B=%d G=%d Ch=%d does not fit (Ch 64 or 128, G <= 14, "
                "tiles <= 16 * SMs / (Ch/16)): use the per-step path", a.B, a.G, a.Ch);
  RGP_CHECK_ARG(a.Hop && a.Hf && a.Cf && a.Wh && a.Xg && a.wci && a.wcf && a.wco && a.flags, "rgp_convlstm_seq_fwd: null pointer%s", "");
  RGP_CHECK_ARG(((uintptr_t)a.Hop & 31) == 0 && ((uintptr_t)a.Hf & 31) == 0 && ((uintptr_t)a.Cf & 31) == 0 &&
                ((uintptr_t)a.Xg & 31) == 0 && ((uintptr_t)a.Wh & 15) == 0 && ((uintptr_t)a.HS & 31) == 0 &&
                ((uintptr_t)a.save & 31) == 0, "rgp_convlstm_seq_fwd: buffers must be 32-byte aligned%s", "");
  const int Ch = a.Ch, S = (a.G + 1) * (a.G + 1);
  const int srows = slab_rows_of(a.G);
  CUtensorMap tmH, tmW;
  int rc = tc::encode_map_bf16(a.Hop, Ch, a.rowsH, (long long)Ch * 2, BK, srows, &tmH);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.Wh, Ch, 27LL * Ch, (long long)Ch * 2, BK, NC, &tmW);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(lstm_seq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) { set_error("lstm_seq: cudaFuncSetAttribute(smem=%d): %s", kSmemBytes, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.B * S + BM - 1) / BM;
  const bool multi = mtiles * (Ch / NC) > num_sms();
  const int R = num_sms() / (Ch / NC);
  if (multi) {
    static PerDeviceFlag attr2 = {};
    if (attr2.need()) {
      cudaError_t e = cudaFuncSetAttribute(lstm_seq_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesMulti);
      if (e != cudaSuccess) { set_error("lstm_seq_multi: cudaFuncSetAttribute(smem=%d): %s", kSmemBytesMulti, cudaGetErrorString(e)); return -5; }
      attr2.set();
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((multi ? R : mtiles) * (Ch / NC)), 1, 1);
  cfg.blockDim = dim3(multi ? NTHREADS_MULTI : NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = multi ? kSmemBytesMulti : kSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;     // co-residency of all CTAs is what makes the counter waits safe
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = multi ? cudaLaunchKernelEx(&cfg, lstm_seq_multi_kernel, tmH, tmW, a, srows, R)
                        : cudaLaunchKernelEx(&cfg, lstm_seq_kernel, tmH, tmW, a, srows);
  if (e != cudaSuccess) { set_error("lstm_seq launch: %s", cudaGetErrorString(e)); return -5; }
  return check_launch("lstm_seq");
}

}  // namespace rgp

namespace rgp {
int convgru_seq_small_bwd(const rgp_gru_bwd_args& a, int* flags, cudaStream_t s) {
  RGP_CHECK_ARG(a.dtype == 1, "rgp_convgru_seq_small_bwd: bf16 operands only (dtype=%d)", a.dtype);
  RGP_CHECK_ARG(a.T > 0 && gru_small_ok(a.B, a.G, a.Ch), "rgp_convgru_seq_small_bwd: B=%d G=%d Ch=%d T=%d does not fit", a.B, a.G,
                a.Ch, a.T);
  RGP_CHECK_ARG(a.dG && a.UcT && a.UzrT && a.Zs && a.Rs && a.Cs && a.Hf && a.dhs && flags, "rgp_convgru_seq_small_bwd: null pointer%s", "");
  RGP_CHECK_ARG(((uintptr_t)a.dG & 31) == 0 && ((uintptr_t)a.Zs & 31) == 0 && ((uintptr_t)a.Rs & 31) == 0 &&
                ((uintptr_t)a.Cs & 31) == 0 && ((uintptr_t)a.Hf & 31) == 0 && ((uintptr_t)a.dhs & 31) == 0 &&
                ((uintptr_t)a.bn_gamma & 31) == 0 && ((uintptr_t)a.UcT & 15) == 0 && ((uintptr_t)a.UzrT & 15) == 0,
                "rgp_convgru_seq_small_bwd: buffers must be 32-byte aligned%s", "");
  const int Ch = a.Ch, S = (a.G + 1) * (a.G + 1);
  const int srows = slab_rows_of(a.G);
  CUtensorMap tmG, tmUcT, tmUzrT;
  int rc = tc::encode_map_bf16(a.dG, 3 * Ch, a.rowsG, (long long)3 * Ch * 2, BK, srows, &tmG);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.UcT, Ch, 9LL * Ch, (long long)Ch * 2, BK, NC, &tmUcT);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.UzrT, 2 * Ch, 9LL * Ch, (long long)2 * Ch * 2, BK, NC, &tmUzrT);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(gru_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesGruBwd);
    if (e != cudaSuccess) { set_error("gru_small_bwd: cudaFuncSetAttribute(smem=%d): %s", kSmemBytesGruBwd, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.B * S + BM - 1) / BM;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(mtiles * (Ch / NC)), 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = kSmemBytesGruBwd;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gru_small_bwd_kernel, tmG, tmUcT, tmUzrT, a, flags, srows);
  if (e != cudaSuccess) { set_error("gru_small_bwd launch: %s", cudaGetErrorString(e)); return -5; }
  return check_launch("gru_small_bwd");
}
}  // namespace rgp

namespace rgp {
int convlstm_seq_small_bwd(const rgp_lstm_bwd_args& a, int* flags, cudaStream_t s) {
  RGP_CHECK_ARG(a.dtype == 1, "rgp_convlstm_seq_small_bwd: bf16 operands only (dtype=%d)", a.dtype);
  RGP_CHECK_ARG(a.T > 0 && gru_small_ok(a.B, a.G, a.Ch), "rgp_convlstm_seq_small_bwd: B=%d G=%d Ch=%d T=%d does not fit", a.B, a.G,
                a.Ch, a.T);
  RGP_CHECK_ARG(a.dHc && a.dG4 && a.WhT && a.save && a.Cf && a.dhs && a.wci && a.wcf && a.wco && a.gwci && a.gwcf && a.gwco && flags,
                "rgp_convlstm_seq_small_bwd: null pointer%s", "");
  RGP_CHECK_ARG(((uintptr_t)a.dHc & 31) == 0 && ((uintptr_t)a.dG4 & 31) == 0 && ((uintptr_t)a.save & 31) == 0 &&
                ((uintptr_t)a.Cf & 31) == 0 && ((uintptr_t)a.dhs & 31) == 0 && ((uintptr_t)a.WhT & 15) == 0,
                "rgp_convlstm_seq_small_bwd: buffers must be 32-byte aligned%s", "");
  const int Ch = a.Ch, S = (a.G + 1) * (a.G + 1);
  const int srows = slab_rows_of(a.G);
  CUtensorMap tmHc, tmWhT;
  int rc = tc::encode_map_bf16(a.dHc, 3 * Ch, a.rowsHc, (long long)3 * Ch * 2, BK, srows, &tmHc);
  if (rc) return rc;
  rc = tc::encode_map_bf16(a.WhT, 3 * Ch, 9LL * Ch, (long long)3 * Ch * 2, BK, NC, &tmWhT);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(lstm_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesLstmBwd);
    if (e != cudaSuccess) { set_error("lstm_small_bwd: cudaFuncSetAttribute(smem=%d): %s", kSmemBytesLstmBwd, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.B * S + BM - 1) / BM;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(mtiles * (Ch / NC)), 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = kSmemBytesLstmBwd;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, lstm_small_bwd_kernel, tmHc, tmWhT, a, flags, srows);
  if (e != cudaSuccess) { set_error("lstm_small_bwd launch: %s", cudaGetErrorString(e)); return -5; }
  return check_launch("lstm_small_bwd");
}
}  // namespace rgp

extern "C" {

int rgp_convlstm_seq_small_bwd(const rgp_lstm_bwd_args* a, int* flags, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_convlstm_seq_small_bwd: null args%s", "");
  return rgp::convlstm_seq_small_bwd(*a, flags, (cudaStream_t)stream);
}

int rgp_convgru_seq_small_bwd(const rgp_gru_bwd_args* a, int* flags, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_convgru_seq_small_bwd: null args%s", "");
  return rgp::convgru_seq_small_bwd(*a, flags, (cudaStream_t)stream);
}

int rgp_convlstm_seq_fwd(const rgp_lstm_seq_args* a, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_convlstm_seq_fwd: null args%s", "");
  return rgp::convlstm_seq_fwd(*a, (cudaStream_t)stream);
}

int rgp_convgru_seq_small_fwd(const rgp_gru_seq_args* a, int* flags, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_convgru_seq_small_fwd: null args%s", "");
  return rgp::convgru_seq_small_fwd(*a, flags, (cudaStream_t)stream);
}

long long rgp_convgru_seq_small_workspace_bytes(int B, int G) {
  const long long mtiles = ((long long)B * (G + 1) * (G + 1) + 127) / 128;
  return (2 * mtiles + 1) * (long long)sizeof(int);
}

long long rgp_convlstm_seq_workspace_bytes(int B, int G) {
  const long long mtiles = ((long long)B * (G + 1) * (G + 1) + 127) / 128;
  return (mtiles + 1) * (long long)sizeof(int);
}

int rgp_convgru_seq_small_fits(int B, int G, int Ch) { return rgp::lstmseq::gru_small_ok(B, G, Ch) ? 1 : 0; }

int rgp_convlstm_seq_fits(int B, int G, int Ch) { return rgp::lstmseq::shape_ok(B, G, Ch) ? 1 : 0; }

int rgp_convlstm_seq_profile(unsigned long long* host_out, int n) {
  RGP_CHECK_ARG(host_out != nullptr && n > 0, "rgp_convlstm_seq_profile: bad arguments%s", "");
  if (n > rgp::lstmseq::kTsSteps * rgp::lstmseq::kTsEvents) n = rgp::lstmseq::kTsSteps * rgp::lstmseq::kTsEvents;
  cudaError_t e = cudaMemcpyFromSymbol(host_out, rgp::lstmseq::g_lstm_ts, sizeof(unsigned long long) * n);
  if (e != cudaSuccess) { rgp::set_error("rgp_convlstm_seq_profile: %s", cudaGetErrorString(e)); return -5; }
  return 0;
}

}  // extern "C"

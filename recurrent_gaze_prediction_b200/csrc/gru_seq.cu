// gru_seq.cu — persistent ConvGRU recurrence: all T steps of gaze_grcn.py:95-129 (unrolled by :259-288) in ONE launch.
//
// Why this shape.  With G = 7 a frame is S = 64 q-rows, so a 128-row MMA tile is exactly two clips and the 3x3 SAME
// halo of a tile only ever touches the zero pad rows of its neighbours: tiles are independent
// for the whole sequence.  One CTA (or a 2-CTA cluster splitting the channels) therefore owns its clips for all T
// steps and needs no grid barrier and no per-step launch.  Per step and CTA, in TMEM columns [0,NH) = z, [NH,2NH) = r
// and later c:
//     r-phase : acc_r  = sum_tap h_{t-1}[m+off] . U_r[tap]          (tcgen05, A = bf16 state rows via TMA)
//     z-phase : acc_z  = sum_tap h_{t-1}[m+off] . U_z[tap]          || r-epilogue: r = sig(acc_r+Xr), write r.h (bf16)
//     c-phase : acc_c  = sum_tap (r.h)[m+off]   . U[tap]            || z-epilogue: z = sig(acc_z+Xz) -> back into TMEM
//     c-epilogue: c = tanh(acc_c+Xc); h_t = z h_{t-1} + (1-z) c -> fp32 master, bf16 operand rows, BN-scaled head rows
// Warp roles: warp 0 = TMA producer (separate A and B rings; the weight ring runs ahead across the step boundary),
// warp 1 = TMEM allocator + single-thread MMA issuer, warps 2..9 = epilogue (TMEM lane group = warp % 4, two column
// halves).  Hand-offs are mbarriers; the state written with ordinary stores is made visible to the async proxy
// (TMA) with __threadfence + fence.proxy.async before the arrive.
#include "tc_common.cuh"

namespace rgp {
namespace {

using namespace tc;

constexpr int BM = 128, BK = 64;
constexpr int EPI_WARPS = 8;
constexpr int NTHREADS = 64 + 32 * EPI_WARPS;

__device__ __forceinline__ float tanh_fast(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }

// 32 consecutive values of one row per thread, as 256-bit accesses (sm_100: LDG/STG.E.ENL2.256).  A thread-per-row warp
// access touches 32 different lines per instruction, so the cost of an epilogue array is its instruction count x 32 L1 tag
// lookups: 256-bit words halve it against float4.  Every row start here is 32-byte aligned (Ch % 64 == 0, slabs and
// parameter tensors 256-byte aligned).
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

// registers -> TMEM (32 lanes x 32 columns), the inverse of tc_ld32
__device__ __forceinline__ void tc_st32(uint32_t taddr, const float v[32]) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// ---- cluster-scope mbarrier helpers (nsplit = 2: the two CTAs of a cluster signal each other) ----
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta_rank) {
  uint32_t raddr;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(bar)), "r"(cta_rank));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// non-blocking probe of a barrier phase
__device__ __forceinline__ bool mbar_test_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "mbarrier.test_wait.parity.acquire.cluster.shared::cta.b64 P1, [%1], %2;\n"
      "selp.u32 %0, 1, 0, P1;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// TMA tile load delivered to the same shared-memory offset (and mbarrier offset) of every CTA in cta_mask
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c0, int c1,
                                               uint16_t cta_mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(cta_mask)
      : "memory");
}
// tcgen05.commit arriving on the mbarrier at the same offset in every CTA of cta_mask
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(cta_mask)
               : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Phase timestamps (globaltimer, ns) of CTA 0 for its first kTsSteps steps — read back by rgp_convgru_seq_profile().
// Events: 0 r-phase gate open (MMA warp) | 1 r MMAs issued | 2 z MMAs issued | 3 c-phase gate open | 4 c MMAs issued |
//         5 r_full seen (epilogue warp 2) | 6 r-epilogue done | 7 z_full seen | 8 z-epilogue done | 9 c_full seen | 10 c-epilogue done
constexpr int kTsSteps = 32, kTsEvents = 12;
__device__ unsigned long long g_seq_ts[kTsSteps * kTsEvents];
__device__ __forceinline__ void stamp(bool on, long long it, int ev) {
  if (on && it < kTsSteps) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_seq_ts[it * kTsEvents + ev] = t;
  }
}

// State operand = a resident slab (same idea as the slab mode of tapgemm_tc.cu).  The 9 taps of the 3x3 conv read the
// tile's rows shifted by ky*Wp + kx <= 2*Wp + 2, so per k-chunk ONE box of kSlabRows rows holds the operand of all
// taps; the MMA's A descriptor starts at slab + off*128 (any row: the 128-byte swizzle acts on absolute smem address
// bits, probed in csrc/debug_umma.cu).  The slab holds ALL k-chunks of the state (Ch/64 boxes) and is filled twice per
// step: h_{t-1} (used by the r- and z-phase) and r.h (c-phase); box kc of r.h is fetched as soon as the z-phase has
// issued its last MMA on box kc.  L2->SMEM traffic per step: 3*36 state boxes of 16 KB (1.7 MB) -> 2*Ch/64 boxes of
// 19 KB (0.15 MB); only the weights (3.5 MB at Ch = 256) still stream.  k-steps run k-chunk-major (kc, tap).
constexpr int kSlabRows = 152;                       // 128 + 2*Wp + 2 = 146 at G = 7, rounded to the 8-row atom
constexpr int kSlabBoxBytes = kSlabRows * BK * 2;    // 19456 = 19 * 1024
constexpr int kMaxKChunks = 4;                       // Ch <= 256

// NT = tiles per CTA.  NT = 2 (with the 2-CTA channel split, NH = 128): a CTA walks TWO 128-row tiles for its half of the
// channels, so every weight box it streams (16 KB) feeds the MMAs of both tiles - the bytes the SM pulls through its
// L2->SMEM port per MMA are halved (what cta_group::2 would give), which is what bounded the MMA phases (16-18 us against
// 9.7 us of tensor work).  TMEM: tile tl uses columns [tl*2NH, tl*2NH + 2NH) (z | r/c), 512 columns in all.
template <int NH, int NT>
struct SeqCfg {
  static constexpr int kStages = (NH >= 256 || NT == 2) ? 4 : 6;
  static constexpr int kABytes = NT * kMaxKChunks * kSlabBoxBytes;  // the slab(s): box index = kc * NT + tl
  static constexpr int kBBytes = NH * BK * 2;
  static constexpr int kTmemCols = NT * 2 * NH;
  static constexpr int kBarBytes = 512;
  static constexpr int kSmemBytes = kABytes + kStages * kBBytes + 1024 + kBarBytes;
};

template <int NH, int NT>
__global__ void __launch_bounds__(NTHREADS, 1)
gru_seq_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmRH,
               const __grid_constant__ CUtensorMap tmUzr, const __grid_constant__ CUtensorMap tmUc,
               const rgp_gru_seq_args a) {
  using C = SeqCfg<NH, NT>;
  constexpr int kBoxes = NT * kMaxKChunks;
  constexpr int TM = NT * BM;                          // q-rows per CTA
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* slab = smem;                                // NT * kchunks boxes of kSlabRows x 64 bf16 (box = kc * NT + tl)
  uint8_t* ringB = smem + C::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kABytes + C::kStages * C::kBBytes);
  uint64_t* fullA = bars;                              // per slab box
  uint64_t* emptyA = fullA + kBoxes;
  uint64_t* fullB = emptyA + kBoxes;
  uint64_t* emptyB = fullB + C::kStages;
  uint64_t* r_full = emptyB + C::kStages;
  uint64_t* z_full = r_full + 1;
  uint64_t* c_full = z_full + 1;
  uint64_t* rh_ready = c_full + 1;     // r.h rows of this step are in global memory (all CTAs of the cluster)
  uint64_t* h_ready = rh_ready + 1;    // h_t rows are in global memory (all CTAs of the cluster)
  uint64_t* z_free = h_ready + 1;      // the z columns of TMEM (they carry h_t through the c-epilogue's tail) are read out
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(z_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // cluster = mcast tiles x nsplit channel halves; CTA rank = mrank * nsplit + nrank
  const int nsplit = a.nsplit, cs = a.mcast, csize = nsplit * cs;
  const int crank = (int)(blockIdx.x % csize);
  const int rank = crank % nsplit, mrank = crank / nsplit;
  const int cluster_id = (int)(blockIdx.x / csize) * cs + mrank;      // first tile of this CTA
  const int nclusters = (int)(gridDim.x / csize) * cs;                 // tile stride
  uint16_t bmask = 0;                                                   // CTAs that share this CTA's weight slice
  for (int j = 0; j < cs; ++j) bmask |= (uint16_t)(1u << (j * nsplit + rank));
  const int peer0 = mrank * nsplit;                                     // first CTA of this tile's channel-split group
  const int Ch = a.Ch, Wp = a.G + 1, S = Wp * Wp;
  const int M = a.B * S;                          // q-rows of one step
  const int mtiles_real = (M + TM - 1) / TM;
  // every CTA of a cluster walks the same number of tiles (weight slices are exchanged in lock-step); tiles past the
  // end are ghosts: all rows out of range, nothing stored
  const int mtiles = ((mtiles_real + cs - 1) / cs) * cs;
  const int col0 = rank * NH;                     // this CTA's channel slice [col0, col0+NH)
  const int kchunks = Ch / BK;
  const int ksteps = 9 * kchunks;                 // k-steps of one phase
  const bool train = a.Rs != nullptr;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmH) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmRH) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmUzr) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmUc) : "memory");
    for (int i = 0; i < kBoxes; ++i) { mbar_init(&fullA[i], 1); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < C::kStages; ++i) { mbar_init(&fullB[i], 1); mbar_init(&emptyB[i], (uint32_t)cs); }
    mbar_init(r_full, 1); mbar_init(z_full, 1); mbar_init(c_full, 1);
    mbar_init(rh_ready, EPI_WARPS * nsplit); mbar_init(h_ready, EPI_WARPS * nsplit);
    mbar_init(z_free, EPI_WARPS);
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
  if (csize > 1) cluster_sync_all();   // the peers' barriers exist before anyone arrives on them remotely
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer: two rings, the weight ring runs up to kStages k-steps ahead =====================
    if (lane == 0) {
      int sB = 0; uint32_t phB = 0;
      long long it = 0;                                   // running step counter of this CTA (barrier phases)
      for (int tile = cluster_id; tile < mtiles; tile += nclusters) {
        const int m0 = tile * TM;
        const int nboxes = NT * kchunks;
        for (int t = 0; t < a.T; ++t, ++it) {
          const long long rowH = (long long)t * M + m0;                 // slab t = h_{t-1} ... (slab 0 = h_0)
          const long long rowRH = (long long)(train ? t : 0) * M + m0;
          int iB = 0;                                                   // weight-ring cursor (k-step inside this step)
          auto issueB = [&]() {
            const int ph = iB / ksteps, ks = iB - ph * ksteps;
            const int kc = ks / 9, tap = ks - kc * 9;                   // k-chunk-major
            mbar_wait_cluster(&emptyB[sB], phB ^ 1);                    // every CTA of the group has released this stage
            mbar_expect_tx(&fullB[sB], C::kBBytes);                     // the whole box: 1/cs from each CTA of the group
            const int brow = ((ph == 0) ? tap * 2 * Ch + Ch : ((ph == 1) ? tap * 2 * Ch : tap * Ch)) + col0;
            const CUtensorMap* tm = (ph == 2) ? &tmUc : &tmUzr;
            if (cs == 1) {
              tma_load_2d(ringB + sB * C::kBBytes, tm, &fullB[sB], kc * BK, brow);
            } else {
              const int srows = NH / cs;                                // this CTA's slice of the box, multicast to the group
              tma_load_2d_mc(ringB + sB * C::kBBytes + mrank * srows * (BK * 2), tm, &fullB[sB], kc * BK,
                             brow + mrank * srows, bmask);
            }
            if (++sB == C::kStages) { sB = 0; phB ^= 1; }
            ++iB;
          };
          // (1) while the c-epilogue of the previous step still holds the state gate: r-phase weights into the slots the
          //     previous c-phase frees (their release never depends on the gate)
          while (iB < C::kStages) issueB();
          if (it > 0) { mbar_wait_cluster(h_ready, (uint32_t)((it - 1) & 1)); fence_proxy_async(); }
          // (2) h_{t-1}: all k-chunks of the slab (box kc was released by the previous c-phase: release parity 1)
          for (int bx = 0; bx < nboxes; ++bx) {
            mbar_wait(&emptyA[bx], 1);
            mbar_expect_tx(&fullA[bx], kSlabBoxBytes);
            tma_load_2d(slab + bx * kSlabBoxBytes, &tmH, &fullA[bx], (bx / NT) * BK, (int)rowH + (bx % NT) * BM);
          }
          // (3) the weights of the three phases in ring order, and r.h box kc as soon as (a) the r-epilogue has published
          //     the rows (rh_ready) and (b) the z-phase has issued its last MMA on box kc (release parity 0).  Both are
          //     polled between weight issues; only when the weight ring is full of c-phase stages that cannot drain
          //     without r.h (kStages ahead) does the producer block on them.
          int nextRH = 0;
          bool rh_ok = false;
          auto loadRH = [&]() {
            mbar_expect_tx(&fullA[nextRH], kSlabBoxBytes);
            tma_load_2d(slab + nextRH * kSlabBoxBytes, &tmRH, &fullA[nextRH], (nextRH / NT) * BK, (int)rowRH + (nextRH % NT) * BM);
            ++nextRH;
          };
          while (iB < 3 * ksteps) {
            if (nextRH < nboxes) {
              if (iB >= 2 * ksteps + C::kStages) {                    // must have r.h before more c-phase weights
                if (!rh_ok) { mbar_wait_cluster(rh_ready, (uint32_t)(it & 1)); fence_proxy_async(); rh_ok = true; }
                mbar_wait(&emptyA[nextRH], 0);
                loadRH();
                continue;
              }
              if (iB >= ksteps + 9 * (nextRH / NT + 1)) {             // the z-phase weights that free box nextRH are issued
                if (!rh_ok && mbar_test_cluster(rh_ready, (uint32_t)(it & 1))) { fence_proxy_async(); rh_ok = true; }
                if (rh_ok && mbar_test_cluster(&emptyA[nextRH], 0)) { loadRH(); continue; }
              }
            }
            issueB();
          }
          while (nextRH < nboxes) {                                     // (never taken: the ring forces r.h before the end)
            if (!rh_ok) { mbar_wait_cluster(rh_ready, (uint32_t)(it & 1)); fence_proxy_async(); rh_ok = true; }
            mbar_wait(&emptyA[nextRH], 0);
            loadRH();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NH >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    int sB = 0; uint32_t phB = 0;
    long long it = 0;
    for (int tile = cluster_id; tile < mtiles; tile += nclusters) {
      for (int t = 0; t < a.T; ++t, ++it) {
        for (int ph = 0; ph < 3; ++ph) {
          // TMEM hazards: r/c columns are free once the c-epilogue of the previous step (ph 0) resp. the r-epilogue
          // of this step (ph 2) has read them; z columns once the previous c-epilogue is done (implied by ph 0's wait)
          if (ph == 0 && it > 0) mbar_wait_cluster(h_ready, (uint32_t)((it - 1) & 1));
          if (ph == 2) mbar_wait_cluster(rh_ready, (uint32_t)(it & 1));
          if (ph == 1 && it > 0) mbar_wait(z_free, (uint32_t)((it - 1) & 1));   // tail of the previous c-epilogue
          tc_fence_after();
          if (ph != 1) stamp(blockIdx.x == 0 && lane == 0, it, ph == 0 ? 0 : 3);
          const uint32_t tacc = tmem_base + (uint32_t)(ph == 1 ? 0 : NH);
          uint64_t* done = (ph == 0) ? r_full : ((ph == 1) ? z_full : c_full);
          for (int ks = 0; ks < ksteps; ++ks) {
            const int kc = ks / 9, tap = ks - kc * 9;
            // slab box (kc, tl): fill 2*it (h_{t-1}, completes barrier phase parity 0) serves ph 0 and 1, fill 2*it+1 (r.h,
            // parity 1) serves ph 2
            if (tap == 0 && ph != 1)
              for (int tl = 0; tl < NT; ++tl) mbar_wait(&fullA[kc * NT + tl], ph == 0 ? 0u : 1u);
            mbar_wait(&fullB[sB], phB);
            tc_fence_after();
            if (lane == 0) {
              const int off = (tap / 3) * Wp + (tap % 3);
              const uint64_t bdesc = make_sdesc(smem_u32(ringB + sB * C::kBBytes));
#pragma unroll
              for (int tl = 0; tl < NT; ++tl) {                    // one weight box, NT tiles
                const uint64_t adesc = make_sdesc(smem_u32(slab + (kc * NT + tl) * kSlabBoxBytes) + (uint32_t)off * (BK * 2));
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)
                  tc_mma_bf16(tacc + (uint32_t)(tl * 2 * NH), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                              (ks | k) ? 1u : 0u);
              }
              if (tap == 8 && ph != 0)                             // last use of these boxes by the z-phase / the c-phase
                for (int tl = 0; tl < NT; ++tl) tc_commit(&emptyA[kc * NT + tl]);
              if (cs == 1) tc_commit(&emptyB[sB]); else tc_commit_mc(&emptyB[sB], bmask);
              if (ks == ksteps - 1) tc_commit(done);
            }
            __syncwarp();
            if (++sB == C::kStages) { sB = 0; phB ^= 1; }
          }
          stamp(blockIdx.x == 0 && lane == 0, it, ph == 0 ? 1 : (ph == 1 ? 2 : 4));
        }
      }
    }
  } else {
    // ===================== epilogue warps 2..9: TMEM lane group = warp % 4, column half = (warp-2)/4 =====================
    using TA = __nv_bfloat16;
    const int lg = warp & 3;
    const int cgi = (warp - 2) >> 2;
    constexpr int NCG = EPI_WARPS / 4;
    // NT = 1: the two warps of a lane group split the columns; NT = 2: they take one tile each, all NH columns
    constexpr int CW = (NT == 2) ? NH : NH / NCG;      // columns this warp handles per gate
    const int tl_e = (NT == 2) ? cgi : 0;              // this warp's tile
    const int cb = (NT == 2) ? 0 : cgi * CW;           // first local column
    const uint32_t tlane = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(tl_e * 2 * NH);
    const int threeCh = 3 * Ch;
    long long it = 0;
    for (int tile = cluster_id; tile < mtiles; tile += nclusters) {
      const int m = tile * TM + tl_e * BM + lg * 32 + lane;        // q-row inside a step slab
      const bool inrange = m < M;
      const int f = m / S, q = m - f * S;
      const int y = q / Wp, x = q - y * Wp;
      const bool valid = inrange && y < a.G && x < a.G;
      const long long prow = (long long)m + Wp + 1;
      for (int t = 0; t < a.T; ++t, ++it) {
        const uint32_t par = (uint32_t)(it & 1);
        const long long mt = (long long)t * M + m;                       // row in the [T*M, .] buffers
        const float* xg = a.Xg + mt * threeCh;
        const float* hprev = a.Hf + ((long long)t * M + prow) * Ch;      // slab t = h_{t-1}
        float* hnext = a.Hf + ((long long)(t + 1) * M + prow) * Ch;
        TA* hop_next = (TA*)a.Hop + ((long long)(t + 1) * M + prow) * Ch;
        TA* rh = (TA*)a.RH + ((long long)(train ? t : 0) * M + prow) * Ch;
        // pull this row's input-half pre-activations (HBM) into L2 while the r-phase MMAs run
        if (inrange) {
#pragma unroll
          for (int g = 0; g < 3; ++g)
            for (int c = 0; c < CW; c += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(xg + g * Ch + col0 + cb + c));
        }
        // ---------------- r-epilogue: r = sig(acc_r + Xr); r.h -> operand rows of the c-phase ----------------
        const bool ts_on = blockIdx.x == 0 && warp == 2 && lane == 0;
        mbar_wait(r_full, par);
        tc_fence_after();
        stamp(ts_on, it, 5);
#pragma unroll 1
        for (int c = 0; c < CW; c += 32) {
          const int n = col0 + cb + c;                                   // global channel
          float v[32], xx[32];
          tc_ld32(tlane + (uint32_t)(NH + cb + c), v);
          if (inrange) {
            ldf32(xg + Ch + n, xx);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = sigmoid_fast(v[i] + xx[i]);
            if (train) stf32(a.Rs + mt * Ch + n, v);
            if (valid) {
              ldf32(hprev + n, xx);
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] *= xx[i];
              stbf32(rh + n, v);
            }
          }
        }
        tc_fence_before();
        __threadfence();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (nsplit > 1) { for (int p = 0; p < nsplit; ++p) mbar_arrive_remote(rh_ready, (uint32_t)(peer0 + p)); }
          else mbar_arrive(rh_ready);
        }
        stamp(ts_on, it, 6);
        // ---------------- z-epilogue: z = sig(acc_z + Xz), kept in TMEM for the c-epilogue ----------------
        mbar_wait(z_full, par);
        tc_fence_after();
        stamp(ts_on, it, 7);
#pragma unroll 1
        for (int c = 0; c < CW; c += 32) {
          const int n = col0 + cb + c;
          float v[32], xx[32];
          tc_ld32(tlane + (uint32_t)(cb + c), v);
          if (inrange) {
            ldf32(xg + n, xx);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = sigmoid_fast(v[i] + xx[i]);
            if (train) stf32(a.Zs + mt * Ch + n, v);
          }
          tc_st32(tlane + (uint32_t)(cb + c), v);
        }
        stamp(ts_on, it, 8);
        // ---------------- c-epilogue: candidate, state update, BN-scaled head rows ----------------
        const long long orow = ((long long)t * a.B + f) * a.So + (long long)(y + a.pado) * a.Wo + (x + a.pado);
        mbar_wait(c_full, par);
        tc_fence_after();
        stamp(ts_on, it, 9);
        // The next step's MMAs need only the bf16 operand rows of h_t.  Critical part: candidate, state update, operand
        // rows (and the saved candidate when training); h_t itself is parked in the z columns of TMEM (z is dead once h_t
        // is formed).  The tail - fp32 master rows and the BN-scaled head rows - is written from there after h_ready has
        // been signalled, i.e. under the r-phase MMAs of step t+1; z_free orders it before that step's z-phase.
#pragma unroll 1
        for (int c = 0; c < CW; c += 32) {
          const int n = col0 + cb + c;
          float v[32], xx[32], zz[32];
          tc_ld32(tlane + (uint32_t)(NH + cb + c), v);
          tc_ld32(tlane + (uint32_t)(cb + c), zz);         // warp-collective: never under a divergent branch
          if (inrange) {
            ldf32(xg + 2 * Ch + n, xx);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = tanh_fast(v[i] + xx[i]);
            if (train) stf32(a.Cs + mt * Ch + n, v);
            if (valid) {
              ldf32(hprev + n, xx);
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = zz[i] * xx[i] + (1.0f - zz[i]) * v[i];
              stbf32(hop_next + n, v);
            }
          }
          tc_st32(tlane + (uint32_t)(cb + c), v);
        }
        tc_fence_before();
        __threadfence();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (nsplit > 1) { for (int p = 0; p < nsplit; ++p) mbar_arrive_remote(h_ready, (uint32_t)(peer0 + p)); }
          else mbar_arrive(h_ready);
        }
        stamp(ts_on, it, 10);
        // ---------------- tail of the c-epilogue (off the critical path) ----------------
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < CW; c += 32) {
          const int n = col0 + cb + c;
          float v[32], xx[32];
          tc_ld32(tlane + (uint32_t)(cb + c), v);
          if (valid) {
            stf32(hnext + n, v);
            if (a.HS) {
              if (a.bn_gamma) {
                ldf32(a.bn_gamma + (long long)t * Ch + n, xx);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = xx[i] * (v[i] * a.alpha);
                ldf32(a.bn_beta + (long long)t * Ch + n, xx);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] += xx[i];
              }
              stbf32((TA*)a.HS + orow * Ch + n, v);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(z_free);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (csize > 1) cluster_sync_all();   // nobody exits while a peer may still signal its barriers or write its smem
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::kTmemCols)
                 : "memory");
  }
}

template <int NH, int NT>
int launch_seq(const rgp_gru_seq_args& a, cudaStream_t s) {
  using C = SeqCfg<NH, NT>;
  const int Ch = a.Ch, S = (a.G + 1) * (a.G + 1);
  CUtensorMap tmH, tmRH, tmUzr, tmUc;
  int rc = encode_map_bf16(a.Hop, Ch, a.rowsH, (long long)Ch * 2, BK, kSlabRows, &tmH);
  if (rc) return rc;
  rc = encode_map_bf16(a.RH, Ch, a.rowsRH, (long long)Ch * 2, BK, kSlabRows, &tmRH);
  if (rc) return rc;
  rc = encode_map_bf16(a.Uzr, Ch, 9LL * 2 * Ch, (long long)Ch * 2, BK, NH / a.mcast, &tmUzr);
  if (rc) return rc;
  rc = encode_map_bf16(a.Uc, Ch, 9LL * Ch, (long long)Ch * 2, BK, NH / a.mcast, &tmUc);
  if (rc) return rc;
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(gru_seq_kernel<NH, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    if (e != cudaSuccess) { set_error("gru_seq: cudaFuncSetAttribute(smem=%d): %s", C::kSmemBytes, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  const int mtiles = (a.B * S + NT * BM - 1) / (NT * BM);
  const int csize = a.nsplit * a.mcast;
  int nclusters = (mtiles + a.mcast - 1) / a.mcast;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // all clusters must be co-resident?  No: a cluster is self-contained, later clusters simply start when SMs free up.
  // But a CTA walks several tiles only if the grid is capped, so cap it at what fits at once.
  static int max_clusters_dev[kMaxDevices][17] = {{0}};
  int* max_clusters = max_clusters_dev[current_device()];
  if (!max_clusters[csize]) {
    cfg.gridDim = dim3((unsigned)(csize * 64), 1, 1);
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, gru_seq_kernel<NH, NT>, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); n = num_sms() / csize; }
    max_clusters[csize] = n;
  }
  if (nclusters > max_clusters[csize]) nclusters = max_clusters[csize];
  cfg.gridDim = dim3((unsigned)(nclusters * csize), 1, 1);
  cudaError_t e = cudaLaunchKernelEx(&cfg, gru_seq_kernel<NH, NT>, tmH, tmRH, tmUzr, tmUc, a);
  if (e != cudaSuccess) { set_error("gru_seq launch: %s", cudaGetErrorString(e)); return -5; }
  return check_launch("gru_seq");
}

}  // namespace

int convgru_seq_profile(unsigned long long* host_out, int n) {
  if (n > kTsSteps * kTsEvents) n = kTsSteps * kTsEvents;
  cudaError_t e = cudaMemcpyFromSymbol(host_out, g_seq_ts, sizeof(unsigned long long) * n);
  if (e != cudaSuccess) { set_error("rgp_convgru_seq_profile: %s", cudaGetErrorString(e)); return -5; }
  return 0;
}

int convgru_seq_fwd(const rgp_gru_seq_args& a, cudaStream_t s) {
  const int S = (a.G + 1) * (a.G + 1);
  RGP_CHECK_ARG(a.dtype == 1, "rgp_convgru_seq_fwd: bf16 operands only (dtype=%d); fp32 uses the per-step rgp_tapgemm path", a.dtype);
  RGP_CHECK_ARG(a.B > 0 && a.T > 0, "rgp_convgru_seq_fwd: bad shape B=%d T=%d", a.B, a.T);
  RGP_CHECK_ARG(BM % S == 0, "rgp_convgru_seq_fwd: a 128-row tile must hold whole frames (S=%d; G must be 7, 3 or 1)", S);
  RGP_CHECK_ARG(a.nsplit == 1 || a.nsplit == 2 || a.nsplit == 4, "rgp_convgru_seq_fwd: nsplit=%d", a.nsplit);
  RGP_CHECK_ARG(a.mcast == 1 || a.mcast == 2 || a.mcast == 4 || a.mcast == 8, "rgp_convgru_seq_fwd: mcast=%d", a.mcast);
  RGP_CHECK_ARG(a.nsplit * a.mcast <= 8, "rgp_convgru_seq_fwd: cluster of %d CTAs (max 8)", a.nsplit * a.mcast);
  RGP_CHECK_ARG(a.Ch % 64 == 0 && a.Ch / 64 <= kMaxKChunks, "rgp_convgru_seq_fwd: Ch=%d must be a multiple of 64, at most %d",
                a.Ch, 64 * kMaxKChunks);
  RGP_CHECK_ARG(128 + 2 * (a.G + 1) + 2 <= kSlabRows, "rgp_convgru_seq_fwd: G=%d does not fit the state slab", a.G);
  RGP_CHECK_ARG((a.Rs == nullptr) == (a.Cs == nullptr), "rgp_convgru_seq_fwd: Rs and Cs must be both set (train) or both NULL");
  RGP_CHECK_ARG(a.Rs == nullptr || a.Zs != nullptr, "rgp_convgru_seq_fwd: training needs Zs, Rs and Cs");
  RGP_CHECK_ARG((a.bn_gamma == nullptr) == (a.bn_beta == nullptr), "rgp_convgru_seq_fwd: bn_gamma / bn_beta");
  RGP_CHECK_ARG(((uintptr_t)a.Hop & 15) == 0 && ((uintptr_t)a.RH & 15) == 0 && ((uintptr_t)a.Uzr & 15) == 0 &&
                ((uintptr_t)a.Uc & 15) == 0, "rgp_convgru_seq_fwd: operands must be 16-byte aligned");
  const int NH = a.Ch / a.nsplit;
  RGP_CHECK_ARG(a.ntile == 0 || a.ntile == 1 || (a.ntile == 2 && NH == 128 && a.mcast == 1),
                "rgp_convgru_seq_fwd: ntile=%d needs Ch/nsplit = 128 and mcast = 1", a.ntile);
  if (NH == 256) return launch_seq<256, 1>(a, s);
  if (NH == 128) return a.ntile == 2 ? launch_seq<128, 2>(a, s) : launch_seq<128, 1>(a, s);
  if (NH == 64) return launch_seq<64, 1>(a, s);          // four CTAs per tile at Ch = 256 (two at Ch = 128): small batches
  set_error("rgp_convgru_seq_fwd: Ch/nsplit = %d must be 64, 128 or 256", NH);
  return -22;
}

}  // namespace rgp

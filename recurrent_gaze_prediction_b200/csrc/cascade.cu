// cascade.cu — top level of the cascade model: models/gaze_grcn_cascade.py:343-424.
//
//   topgru_seq_kernel : the 5x5 ConvGRU with 3 units on the 49x49 grid (:346-351,379; cell :95-129), all T steps in
//                       one launch, state and r.h resident in shared memory for the whole sequence.  One CTA per clip, or
//                       (NC = 2, taken while two CTAs per clip still fit the SMs) a cluster of two CTAs per clip: each
//                       computes half of the rows, keeps whole planes, and writes the two rows next to the split into the
//                       partner's planes through distributed shared memory; two cluster barriers per step.
//   maxout_kernel     : bias + relu + max of the two halves of a maxout FC (:388-420) after its GEMM.
//
// The two GEMM-shaped pieces around them run on the tap-GEMM engines: the 11x11 / stride-7 transposed conv as a 4-tap
// GEMM (RGP_EPI_UP7), the input half of the top cell as a 25-tap GEMM, the two FCs as plain GEMMs.
#include <cooperative_groups.h>
#include <stdlib.h>
#include "common.cuh"

namespace cg = cooperative_groups;

namespace rgp {
namespace {

constexpr int GT = 49;              // top grid
constexpr int NU = 3;               // units of the top cell
constexpr int KT = 5;               // its kernel
constexpr int PW = GT + KT - 1;     // 53: zero-bordered plane width in shared memory
constexpr int PLANE = PW * PW;      // 2809
constexpr int NPIX = GT * GT;       // 2401
constexpr int TOP_THREADS = 512;
constexpr int PPT = (NPIX + TOP_THREADS - 1) / TOP_THREADS;   // 5 pixels per thread

__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

template <typename TA, int NC>
__global__ void __cluster_dims__(NC, 1, 1) __launch_bounds__(TOP_THREADS, 1)
topgru_seq_kernel(const float* __restrict__ xg, int ldx, int So, int Wo, const float* __restrict__ uzr,
                  const float* __restrict__ uc, const float* __restrict__ sal, const float* __restrict__ wsal,
                  float* __restrict__ state, int carry, TA* __restrict__ og_op, int ldo, float* __restrict__ h_f32,
                  int B, int T) {
  extern __shared__ float sm[];
  float* hs = sm;                          // [NU][PLANE]   h_{t-1}, zero border of 2
  float* rhs = hs + NU * PLANE;            // [NU][PLANE]   r . h_{t-1}
  float* ss = rhs + NU * PLANE;            // [PLANE]       frame saliency (reading B), zero border
  float* wzr = ss + PLANE + 1;             // [25][NU][6]  (+1: 7 * 2809 floats is odd, the float2 reads below need 8-byte alignment)
  float* wc = wzr + 25 * NU * 6;           // [25][NU][NU]
  float* ws = wc + 25 * NU * NU;           // [25][9]
  cg::cluster_group cl = cg::this_cluster();
  const int rank = NC > 1 ? (int)cl.block_rank() : 0;
  const int b = blockIdx.x / NC, tid = threadIdx.x;
  // rows [y0, y1) of the grid are this CTA's; it holds whole planes and receives the partner's two rows next to the split
  constexpr int RH = (GT + NC - 1) / NC;                 // 49 or 25
  constexpr int PPTN = (RH * GT + TOP_THREADS - 1) / TOP_THREADS;   // 5 or 3 pixels per thread
  const int y0 = rank * RH, y1 = min(GT, y0 + RH);
  const int npx = (y1 - y0) * GT;
  float* phs = NC > 1 ? cl.map_shared_rank(hs, rank ^ 1) : nullptr;       // the partner's planes
  float* prhs = NC > 1 ? cl.map_shared_rank(rhs, rank ^ 1) : nullptr;
  auto sync_all = [&]() { if (NC > 1) cl.sync(); else __syncthreads(); };
  // a row of mine the partner's 5x5 window reaches: the two rows at the split
  auto halo_row = [&](int y) { return NC > 1 && (rank == 0 ? y >= y1 - 2 : y < y0 + 2); };

  for (int i = tid; i < (2 * NU + 1) * PLANE; i += TOP_THREADS) sm[i] = 0.f;
  for (int i = tid; i < 25 * NU * 6; i += TOP_THREADS) wzr[i] = uzr[i];
  for (int i = tid; i < 25 * NU * NU; i += TOP_THREADS) wc[i] = uc[i];
  if (sal)
    for (int i = tid; i < 25 * 9; i += TOP_THREADS) ws[i] = wsal[i];
  __syncthreads();
  if (carry && state)
    for (int i = tid; i < NPIX * NU; i += TOP_THREADS) {
      const int p = i / NU, u = i - p * NU, y = p / GT, x = p - y * GT;
      hs[u * PLANE + (y + 2) * PW + x + 2] = state[(long long)b * NPIX * NU + i];
    }
  sync_all();                 // (cluster: nobody writes into a partner's planes before they are initialised)

  for (int t = 0; t < T; ++t) {
    const long long f = (long long)t * B + b;
    const float* xf = xg + f * So * ldx;
    if (sal) {
      const float* sf = sal + ((long long)b * T + t) * NPIX;
      for (int p = tid; p < NPIX; p += TOP_THREADS) {
        const int y = p / GT, x = p - y * GT;
        ss[(y + 2) * PW + x + 2] = sf[p];
      }
      __syncthreads();
    }
    float z[PPTN][NU], xc[PPTN][NU];
    // Tap-outer loops: the 18 (9) weights of a tap are read from shared memory ONCE per thread and applied to its ~5
    // pixels from registers.  (Pixel-outer, 21 LDS per pixel and tap, the kernel was bound by the LDS pipe: 33 us per step.)
    int base[PPTN];
    bool live[PPTN];
    // ---- phase A: z, r and r.h  (gaze_grcn_cascade.py:108-119)
    float acc[PPTN][6];
#pragma unroll
    for (int k = 0; k < PPTN; ++k) {
      const int pl = tid + k * TOP_THREADS;
      live[k] = pl < npx;
      const int pc = live[k] ? pl : 0;
      const int y = y0 + pc / GT, x = pc - (pc / GT) * GT;
      base[k] = y * PW + x;
      const float* xp = xf + (long long)(y * Wo + x) * ldx;
      const float4 x0 = *reinterpret_cast<const float4*>(xp);
      const float4 x1 = *reinterpret_cast<const float4*>(xp + 4);
      acc[k][0] = x0.x; acc[k][1] = x0.y; acc[k][2] = x0.z; acc[k][3] = x0.w; acc[k][4] = x1.x; acc[k][5] = x1.y;
      xc[k][0] = x1.z; xc[k][1] = x1.w; xc[k][2] = xp[8];
    }
#pragma unroll 1
    for (int tap = 0; tap < 25; ++tap) {
      const int to = (tap / KT) * PW + (tap % KT);
      float w[NU * 6];
#pragma unroll
      for (int i = 0; i < NU * 6; i += 2) {
        const float2 v = *reinterpret_cast<const float2*>(wzr + tap * NU * 6 + i);
        w[i] = v.x; w[i + 1] = v.y;
      }
#pragma unroll
      for (int k = 0; k < PPTN; ++k) {
        const int o = base[k] + to;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const float hv = hs[u * PLANE + o];
#pragma unroll
          for (int j = 0; j < 6; ++j) acc[k][j] = fmaf(hv, w[u * 6 + j], acc[k][j]);
        }
      }
      if (sal) {
        float q[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) q[i] = ws[tap * 9 + i];
#pragma unroll
        for (int k = 0; k < PPTN; ++k) {
          const float sv = ss[base[k] + to];
#pragma unroll
          for (int j = 0; j < 6; ++j) acc[k][j] = fmaf(sv, q[j], acc[k][j]);
#pragma unroll
          for (int j = 0; j < NU; ++j) xc[k][j] = fmaf(sv, q[6 + j], xc[k][j]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < PPTN; ++k) {
      if (live[k]) {
        const int c = base[k] + 2 * PW + 2;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          z[k][u] = sigm(acc[k][u]);
          const float rh = sigm(acc[k][NU + u]) * hs[u * PLANE + c];
          rhs[u * PLANE + c] = rh;
          if (halo_row(c / PW - 2)) prhs[u * PLANE + c] = rh;
        }
      }
    }
    sync_all();
    // ---- phase B: candidate and state update  (:122-127)
    float hn[PPTN][NU];
#pragma unroll 1
    for (int tap = 0; tap < 25; ++tap) {
      const int to = (tap / KT) * PW + (tap % KT);
      float w[NU * NU];
#pragma unroll
      for (int i = 0; i < NU * NU; ++i) w[i] = wc[tap * NU * NU + i];
#pragma unroll
      for (int k = 0; k < PPTN; ++k) {
        const int o = base[k] + to;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const float rv = rhs[u * PLANE + o];
#pragma unroll
          for (int j = 0; j < NU; ++j) xc[k][j] = fmaf(rv, w[u * NU + j], xc[k][j]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < PPTN; ++k) {
      if (live[k]) {
        const int c = base[k] + 2 * PW + 2;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const float cv = tanhf(xc[k][u]);
          hn[k][u] = z[k][u] * hs[u * PLANE + c] + (1.0f - z[k][u]) * cv;
        }
      }
    }
    // h_{t-1} is read by other threads (and by the partner, from its own planes) only in phase A, which every thread of
    // both CTAs has left at the barrier above: a thread may overwrite its own pixels now
#pragma unroll
    for (int k = 0; k < PPTN; ++k) {
      if (live[k]) {
        const int c = base[k] + 2 * PW + 2;
        const int y = c / PW - 2, x = c - (y + 2) * PW - 2;
        const int p = y * GT + x;
        const bool halo = halo_row(y);
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          hs[u * PLANE + c] = hn[k][u];
          if (halo) phs[u * PLANE + c] = hn[k][u];
          OpT<TA>::st(og_op + f * ldo + p * NU + u, hn[k][u]);       // tf.reshape(output_g, [B,-1]): (y, x, unit)
          if (h_f32) h_f32[f * (NPIX * NU) + p * NU + u] = hn[k][u];
        }
      }
    }
    sync_all();
  }
  if (state)
    for (int i = y0 * GT * NU + tid; i < y1 * GT * NU; i += TOP_THREADS) {
      const int p = i / NU, u = i - p * NU, y = p / GT, x = p - y * GT;
      state[(long long)b * NPIX * NU + i] = hs[u * PLANE + (y + 2) * PW + x + 2];
    }
}

template <typename TA>
__global__ void __launch_bounds__(256) maxout_kernel(const float* __restrict__ y, int ldy, const float* __restrict__ bias,
                                                     int half, int n, TA* __restrict__ out_op, int ldo,
                                                     float* __restrict__ out_maps, int B, int T) {
  const long long total = (long long)B * T * n;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long r = i / n;
    const int j = (int)(i - r * n);
    const float a = fmaxf(y[r * ldy + j] + bias[j], 0.f);
    const float c = fmaxf(y[r * ldy + half + j] + bias[half + j], 0.f);
    const float v = fmaxf(a, c);
    if (out_op) OpT<TA>::st(out_op + r * ldo + j, v);
    if (out_maps) {
      const long long t = r / B, b = r - t * B;      // rows are time-major
      out_maps[(b * T + t) * n + j] = v;
    }
  }
}

constexpr int kTopSmem = ((2 * NU + 1) * PLANE + 1 + 25 * NU * 6 + 25 * NU * NU + 25 * 9) * (int)sizeof(float);

template <typename TA>
int launch_topgru(const float* xg, int ldx, int So, int Wo, const float* uzr, const float* uc, const float* sal,
                  const float* wsal, float* state, int carry, void* og_op, int ldo, float* h_f32, int B, int T,
                  cudaStream_t s) {
  static PerDeviceFlag attr_set = {};
  if (attr_set.need()) {
    cudaError_t e = cudaFuncSetAttribute(topgru_seq_kernel<TA, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(topgru_seq_kernel<TA, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopSmem);
    if (e != cudaSuccess) { set_error("rgp_topgru_seq_fwd: cudaFuncSetAttribute(%d): %s", kTopSmem, cudaGetErrorString(e)); return -5; }
    attr_set.set();
  }
  // two CTAs per clip while they fit the SMs at once (the kernel is latency-bound: one CTA per clip leaves 84 of 148 SMs
  // idle at 64 clips); RGP_TOPGRU_SPLIT=1 forces one CTA per clip
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, current_device());
  static const bool one = [] { const char* e = getenv("RGP_TOPGRU_SPLIT"); return e && e[0] == '1'; }();
  if (!one && 2 * B <= nsm)
    topgru_seq_kernel<TA, 2><<<2 * B, TOP_THREADS, kTopSmem, s>>>(xg, ldx, So, Wo, uzr, uc, sal, wsal, state, carry, (TA*)og_op,
                                                                  ldo, h_f32, B, T);
  else
    topgru_seq_kernel<TA, 1><<<B, TOP_THREADS, kTopSmem, s>>>(xg, ldx, So, Wo, uzr, uc, sal, wsal, state, carry, (TA*)og_op,
                                                              ldo, h_f32, B, T);
  return check_launch("rgp_topgru_seq_fwd");
}

}  // namespace
}  // namespace rgp

using namespace rgp;

// Composed weights of the fused up-sampling + top input conv (rgp.h, rgp_cascade_compose).  One block per
// (class, tap, phase) and 64-wide chunk of k; a thread owns one k and the 9 gates.
template <typename TA>
__global__ void __launch_bounds__(64) cascade_compose_kernel(const float* __restrict__ up, int upn, const float* __restrict__ wtx,
                                                             TA* __restrict__ wc, int Ch) {
  const int k = blockIdx.y * 64 + threadIdx.x;
  int b = blockIdx.x;
  const int ph = b % 64; b /= 64;
  const int tap = b % 9, cls = b / 9;
  TA* dst = wc + ((long long)(cls * 9 + tap) * 1024 + ph * 16) * Ch + k;
  if (k >= Ch) return;
  float acc[9];
#pragma unroll
  for (int g = 0; g < 9; ++g) acc[g] = 0.f;
  if (ph < 49) {
    const int cy = cls / 3, cx = cls % 3, ky = tap / 3, kx = tap % 3, py = ph / 7, px = ph % 7;
    for (int dy = -2; dy <= 2; ++dy) {
      const int iy = 7 * (1 - ky) + py + dy + 2;
      if (iy < 0 || iy > 10 || (cy == 0 && py + dy < 0) || (cy == 2 && py + dy > 6)) continue;
      for (int dx = -2; dx <= 2; ++dx) {
        const int ix = 7 * (1 - kx) + px + dx + 2;
        if (ix < 0 || ix > 10 || (cx == 0 && px + dx < 0) || (cx == 2 && px + dx > 6)) continue;
        const int di = iy / 7, ry = iy - 7 * di, dj = ix / 7, rx = ix - 7 * dj;
        const float* u = up + ((long long)(di * 2 + dj) * upn + (ry * 7 + rx) * 64) * Ch + k;
        const float* w = wtx + ((dy + 2) * 5 + dx + 2) * 16 * 64;
        for (int c = 0; c < 64; ++c) {
          const float uv = u[(long long)c * Ch];
#pragma unroll
          for (int g = 0; g < 9; ++g) acc[g] = fmaf(uv, __ldg(w + g * 64 + c), acc[g]);
        }
      }
    }
  }
#pragma unroll
  for (int g = 0; g < 16; ++g) OpT<TA>::st(dst + (long long)g * Ch, g < 9 ? acc[g] : 0.f);
}

extern "C" {

int rgp_cascade_compose(const float* up, int upn, const float* wtx, void* wc, int Ch, int dtype, void* stream) {
  RGP_CHECK_ARG(up && wtx && wc, "rgp_cascade_compose: null pointer%s", "");
  RGP_CHECK_ARG(Ch > 0 && upn >= 49 * 64, "rgp_cascade_compose: Ch=%d upn=%d", Ch, upn);
  RGP_CHECK_ARG(dtype == 0 || dtype == 1, "rgp_cascade_compose: dtype %d", dtype);
  dim3 grid(9 * 9 * 64, (Ch + 63) / 64);
  if (dtype == 0)
    cascade_compose_kernel<float><<<grid, 64, 0, (cudaStream_t)stream>>>(up, upn, wtx, (float*)wc, Ch);
  else
    cascade_compose_kernel<__nv_bfloat16><<<grid, 64, 0, (cudaStream_t)stream>>>(up, upn, wtx, (__nv_bfloat16*)wc, Ch);
  return check_launch("cascade_compose");
}

int rgp_topgru_seq_fwd(const float* xg, int ldx, int So, int Wo, const float* uzr, const float* uc, const float* sal,
                       const float* wsal, float* state, int carry, void* og_op, int ldo, float* h_f32, int B, int T,
                       int dtype, void* stream) {
  RGP_CHECK_ARG(xg && uzr && uc && og_op, "rgp_topgru_seq_fwd: null pointer%s", "");
  RGP_CHECK_ARG(B > 0 && T > 0, "rgp_topgru_seq_fwd: B=%d T=%d", B, T);
  RGP_CHECK_ARG(ldx >= 9 && ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(xg) & 15) == 0,
                "rgp_topgru_seq_fwd: ldx=%d must be a multiple of 4 (>= 9) and xg 16-byte aligned", ldx);
  RGP_CHECK_ARG(Wo >= GT + 2 && So >= (GT - 1) * Wo + GT, "rgp_topgru_seq_fwd: So=%d Wo=%d too small for a 49x49 grid", So, Wo);
  RGP_CHECK_ARG(ldo >= NPIX * NU, "rgp_topgru_seq_fwd: ldo=%d < 7203", ldo);
  RGP_CHECK_ARG((sal == nullptr) == (wsal == nullptr), "rgp_topgru_seq_fwd: sal and wsal go together%s", "");
  RGP_CHECK_ARG(dtype == 0 || dtype == 1, "rgp_topgru_seq_fwd: dtype %d", dtype);
  if (dtype == 0)
    return launch_topgru<float>(xg, ldx, So, Wo, uzr, uc, sal, wsal, state, carry, og_op, ldo, h_f32, B, T,
                                (cudaStream_t)stream);
  return launch_topgru<__nv_bfloat16>(xg, ldx, So, Wo, uzr, uc, sal, wsal, state, carry, og_op, ldo, h_f32, B, T,
                                      (cudaStream_t)stream);
}

int rgp_maxout(const float* y, int ldy, const float* bias, int half, int n, void* out_op, int ldo, float* out_maps,
               int B, int T, int dtype, void* stream) {
  RGP_CHECK_ARG(y && bias && (out_op || out_maps), "rgp_maxout: null pointer%s", "");
  RGP_CHECK_ARG(n > 0 && n <= half && half + n <= ldy, "rgp_maxout: n=%d half=%d ldy=%d", n, half, ldy);
  RGP_CHECK_ARG(B > 0 && T > 0, "rgp_maxout: B=%d T=%d", B, T);
  RGP_CHECK_ARG(!out_op || ldo >= n, "rgp_maxout: ldo=%d < n=%d", ldo, n);
  const long long total = (long long)B * T * n;
  int grid = (int)((total + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  if (dtype == 0)
    maxout_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(y, ldy, bias, half, n, (float*)out_op, ldo, out_maps, B, T);
  else
    maxout_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(y, ldy, bias, half, n, (__nv_bfloat16*)out_op,
                                                                        ldo, out_maps, B, T);
  return check_launch("rgp_maxout");
}

}  // extern "C"

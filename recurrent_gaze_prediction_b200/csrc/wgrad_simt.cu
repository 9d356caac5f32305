// wgrad_simt.cu — fp32 (CUDA-core) engine of the weight-gradient contraction over rows:
//   D[tap][i, j] += sum_m Y[m, ycol0+i] * X[m + tap_off[tap], xcol0+j]
// (dW of every conv / projection of the path: base.py:278-281 applied to gaze_grcn.py:108-125,239-242).
// 64x64 output tile per CTA, 16-row K chunks, split over the row range; partials are added atomically.
#include "common.cuh"

namespace rgp {
namespace {
constexpr int TI = 64, TJ = 64, TK = 16;

template <typename TA>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(const rgp_wgrad_args a, int splits, int chunk) {
  __shared__ float Ys[TK][TI + 4];
  __shared__ float Xs[TK][TJ + 4];
  const int tilesj = (a.Nx + TJ - 1) / TJ;
  const int ti = blockIdx.x / tilesj, tj = blockIdx.x - ti * tilesj;
  const int tap = blockIdx.y;
  const int i0 = ti * TI, j0 = tj * TJ;
  const int mbeg = blockIdx.z * chunk;
  const int mend = min(a.M, mbeg + chunk);
  const int off = a.tap_off[tap];
  const TA* __restrict__ Y = (const TA*)a.Y;
  const TA* __restrict__ X = (const TA*)a.X;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int lr = tid >> 4, lc = (tid & 15) * 4;   // loader: row lr (0..15), 4 consecutive columns
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int m0 = mbeg; m0 < mend; m0 += TK) {
    const int m = m0 + lr;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float yv = 0.f, xv = 0.f;
      if (m < mend) {
        int ci = i0 + lc + e, cj = j0 + lc + e;
        if (ci < a.Ny) yv = OpT<TA>::ld(Y + (long long)m * a.ldy + a.ycol0 + ci);
        long long xr = (long long)m + off;
        if (cj < a.Nx && xr >= 0 && xr < a.rowsX) xv = OpT<TA>::ld(X + xr * a.ldx + a.xcol0 + cj);
      }
      Ys[lr][lc + e] = yv;
      Xs[lr][lc + e] = xv;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float4 y4 = *reinterpret_cast<const float4*>(&Ys[k][ty * 4]);
      float4 x4 = *reinterpret_cast<const float4*>(&Xs[k][tx * 4]);
      float yv[4] = {y4.x, y4.y, y4.z, y4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(yv[i], xv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* D = a.D + (long long)tap * a.Ny * a.Nx;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gi = i0 + ty * 4 + i;
    if (gi >= a.Ny) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gj = j0 + tx * 4 + j;
      if (gj < a.Nx) atomicAdd(D + (long long)gi * a.Nx + gj, acc[i][j]);
    }
  }
}
}  // namespace

int wgrad_simt(const rgp_wgrad_args& a, cudaStream_t s) {
  const int tiles = ((a.Ny + TI - 1) / TI) * ((a.Nx + TJ - 1) / TJ);
  int splits = (148 * 4 + tiles * a.ntaps - 1) / (tiles * a.ntaps);
  int maxs = (a.M + 255) / 256;
  if (splits > maxs) splits = maxs;
  if (splits < 1) splits = 1;
  int chunk = ((a.M + splits - 1) / splits + TK - 1) / TK * TK;
  splits = (a.M + chunk - 1) / chunk;
  dim3 grid(tiles, a.ntaps, splits);
  if (a.dtype == 0) wgrad_simt_kernel<float><<<grid, 256, 0, s>>>(a, splits, chunk);
  else wgrad_simt_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(a, splits, chunk);
  return check_launch("wgrad_simt");
}


}  // namespace rgp

// tapgemm_simt.cu — the fp32 (CUDA-core FMA) engine of the tap GEMM.
// It is the fp32 parity path of the north star ("fp32 gaze maps within 1e-4") and the anchor the
// tcgen05 engine is validated against.  Classic shared-memory tiling: 128x64 CTA tile, 16-wide K
// chunks, 256 threads x (8x4) register micro-tiles, double-buffered through registers.
#include "common.cuh"

namespace rgp {

namespace {
constexpr int BM = 128, BN = 64, BK = 16, NT = 256;
constexpr int TM = 8, TN = 4;

template <typename TA>
__device__ __forceinline__ void load4(const TA* p, float v[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float v[4]) {
  float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float v[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
  v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
}

template <typename TA>
__global__ void __launch_bounds__(NT) tapgemm_simt_kernel(const rgp_tapgemm_args a) {
  __shared__ float As[2][BK][BM + 4];
  __shared__ float Bs[2][BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const TA* __restrict__ A = (const TA*)a.A;
  const TA* __restrict__ Bw = (const TA*)a.Bw;

  // loader mapping: A tile 128 rows x 16 k = 512 quads -> 2 per thread; B tile 64 x 16 = 256 quads -> 1
  const int lrow = tid >> 2, lk = (tid & 3) * 4;
  const int kchunks = a.K / BK;
  const int total = a.ntaps * kchunks;

  float ra[2][4], rb[4];
  auto gload = [&](int it) {
    int tap = it / kchunks, kc = it - tap * kchunks;
    int k = kc * BK + lk;
    int off = a.tap_off[tap];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      long long row = (long long)m0 + lrow + h * 64 + off;
      if (row >= 0 && row < a.rowsA && (m0 + lrow + h * 64) < a.M) load4<TA>(A + row * a.lda + k, ra[h]);
      else { ra[h][0] = ra[h][1] = ra[h][2] = ra[h][3] = 0.f; }
    }
    int n = n0 + lrow;
    if (n < a.N) load4<TA>(Bw + ((long long)tap * a.N + n) * a.K + k, rb);
    else { rb[0] = rb[1] = rb[2] = rb[3] = 0.f; }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int j = 0; j < 4; ++j) As[buf][lk + j][lrow + h * 64] = ra[h][j];
#pragma unroll
    for (int j = 0; j < 4; ++j) Bs[buf][lk + j][lrow] = rb[j];
  };

  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 thread grid: rows ty*8.., cols tx*4..
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  gload(0);
  sstore(0);
  __syncthreads();
  for (int it = 0; it < total; ++it) {
    int buf = it & 1;
    if (it + 1 < total) gload(it + 1);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float av[TM], bv[TN];
      float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8 + 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
      av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
      bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (it + 1 < total) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < TM; ++i) {
    RowInfo r = make_row(a, m0 + ty * 8 + i);
#pragma unroll
    for (int j = 0; j < TN; ++j) epi_elem<TA>(a, r, n0 + tx * 4 + j, acc[i][j]);
  }
}
}  // namespace

int tapgemm_simt(const rgp_tapgemm_args& a, cudaStream_t s) {
  RGP_CHECK_ARG(a.K % BK == 0, "tapgemm_simt: K=%d must be a multiple of %d", a.K, BK);
  RGP_CHECK_ARG(a.lda % 4 == 0, "tapgemm_simt: lda=%d must be a multiple of 4", a.lda);
  dim3 grid((a.M + BM - 1) / BM, (a.N + BN - 1) / BN);
  if (a.dtype == 0) tapgemm_simt_kernel<float><<<grid, NT, 0, s>>>(a);
  else tapgemm_simt_kernel<__nv_bfloat16><<<grid, NT, 0, s>>>(a);
  return check_launch("tapgemm_simt");
}

}  // namespace rgp

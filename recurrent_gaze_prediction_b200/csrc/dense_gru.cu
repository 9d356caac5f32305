// dense_gru.cu — the dense-GRU variant of the gaze model (reference models/gaze_rnn.py:283-359).
//
// The contractions of this variant are plain GEMMs and run on the tap-GEMM engines with ntaps = 1 (rgp_tapgemm):
//   projection      [T*B*49, 1024] x [1024, 32]                       (gaze_rnn.py:294-301)
//   input half      [T*B, 1568]    x [1568, 3*1617]  for all T at once (the x rows of the GRUCell kernels)
//   per step        [B, 1617] x [1617, 2*1617]  (gates)  and  [B, 1617] x [1617, 1617]  (candidate, on r*h)
//   output          [T*B, 1617]    x [1617, 2401]    for all T at once (gaze_rnn.py:346-348)
// with every width padded to the engine's granularity (Ip = 1600, Hp = 1664, Np = 2432; padded weights are zero and the
// padded state columns stay exactly zero: sigmoid(0) = 0.5, tanh(0) = 0, h' = 0.5*0 + 0.5*0).
// What is left for this file is the element-wise glue, one pass over [M, Hp] each:
//   rgp_dense_pack   : projection accumulators + bias -> the GRU's input rows (flatten order y, x, c), operand dtype
//   rgp_dense_gru_ru : r, u = sigmoid(x-half + h-half + bias) ; r*h as the candidate GEMM's operand ; u kept
//   rgp_dense_gru_c  : c = tanh(x-half + (r*h)-half + bias) ; h' = u*h + (1-u)*c ; fp32 master + operand copy
//   rgp_dense_logits : output accumulators + bias -> logits [B, T, GH*GW] (time-major rows -> [B, T])
// Cell arithmetic: tf.nn.rnn_cell.GRUCell [TF-lib, un-vendored; parity unpinned, see oracle/model_oracle.py].
#include <cuda_bf16.h>
#include "common.cuh"

namespace rgp {
namespace {

template <class T> __device__ __forceinline__ T to_op(float v);
template <> __device__ __forceinline__ float to_op<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 to_op<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

template <class T>
__global__ void dense_pack_kernel(const float* __restrict__ acc, int ldacc, const float* __restrict__ bias,
                                  T* __restrict__ X, int ldx, long long rows, int P, int Cp) {
  const long long total = rows * ldx;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / ldx;
    const int j = (int)(i - r * ldx);
    float v = 0.0f;
    if (j < P * Cp) {
      const int pos = j / Cp, c = j - pos * Cp;
      v = acc[(r * P + pos) * ldacc + c] + bias[c];
    }
    X[i] = to_op<T>(v);
  }
}

template <class T>
__global__ void dense_gru_ru_kernel(const float* __restrict__ xg, int ldx, const float* __restrict__ hg,
                                    const float* __restrict__ bias_g, const float* __restrict__ h,
                                    T* __restrict__ rh, float* __restrict__ u_out, long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float rhv = 0.0f, uv = 0.0f;
    if (j < H) {
      const float pr = xg[m * ldx + j] + hg[m * 2 * Hp + j] + bias_g[j];
      const float pu = xg[m * ldx + Hp + j] + hg[m * 2 * Hp + Hp + j] + bias_g[Hp + j];
      rhv = sigmoidf_(pr) * h[i];
      uv = sigmoidf_(pu);
    }
    rh[i] = to_op<T>(rhv);
    u_out[i] = uv;
  }
}

template <class T>
__global__ void dense_gru_c_kernel(const float* __restrict__ xg, int ldx, const float* __restrict__ hc,
                                   const float* __restrict__ bias_c, const float* __restrict__ u,
                                   const float* __restrict__ h, float* __restrict__ hn, T* __restrict__ hn_op,
                                   long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float v = 0.0f;
    if (j < H) {
      const float c = tanhf(xg[m * ldx + 2 * Hp + j] + hc[i] + bias_c[j]);
      const float uu = u[i];
      v = uu * h[i] + (1.0f - uu) * c;
    }
    hn[i] = v;
    hn_op[i] = to_op<T>(v);
  }
}

__global__ void dense_logits_kernel(const float* __restrict__ acc, int ldacc, const float* __restrict__ bias,
                                    float* __restrict__ logits, int B, int T, int N) {
  const long long total = (long long)B * T * N;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long bt = i / N;
    const int j = (int)(i - bt * N);
    const int b = (int)(bt / T), t = (int)(bt - (long long)b * T);
    logits[i] = acc[((long long)t * B + b) * ldacc + j] + bias[j];
  }
}

int grid_for(long long total) {
  long long g = (total + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace
}  // namespace rgp

using namespace rgp;

extern "C" {

int rgp_dense_pack(const float* acc, int ldacc, const float* bias, void* X, int ldx, long long rows, int P, int Cp,
                   int dtype, void* stream) {
  RGP_CHECK_ARG(rows >= 0 && P > 0 && Cp > 0 && ldacc >= Cp && ldx >= P * Cp, "rgp_dense_pack: rows=%lld P=%d Cp=%d ldacc=%d ldx=%d",
                rows, P, Cp, ldacc, ldx);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == 0) dense_pack_kernel<float><<<grid_for(rows * ldx), 256, 0, s>>>(acc, ldacc, bias, (float*)X, ldx, rows, P, Cp);
  else dense_pack_kernel<__nv_bfloat16><<<grid_for(rows * ldx), 256, 0, s>>>(acc, ldacc, bias, (__nv_bfloat16*)X, ldx, rows, P, Cp);
  return check_launch("rgp_dense_pack");
}

int rgp_dense_gru_ru(const float* xg, int ldx, const float* hg, const float* bias_g, const float* h_f32, void* rh_op,
                     float* u_out, int M, int H, int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M >= 0 && H > 0 && Hp >= H && ldx >= 3 * Hp, "rgp_dense_gru_ru: M=%d H=%d Hp=%d ldx=%d", M, H, Hp, ldx);
  if (M == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_ru_kernel<float><<<grid_for(total), 256, 0, s>>>(xg, ldx, hg, bias_g, h_f32, (float*)rh_op, u_out, M, H, Hp);
  else dense_gru_ru_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(xg, ldx, hg, bias_g, h_f32, (__nv_bfloat16*)rh_op, u_out, M, H, Hp);
  return check_launch("rgp_dense_gru_ru");
}

int rgp_dense_gru_c(const float* xg, int ldx, const float* hc, const float* bias_c, const float* u, const float* h_f32,
                    float* hnext_f32, void* hnext_op, int M, int H, int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M >= 0 && H > 0 && Hp >= H && ldx >= 3 * Hp, "rgp_dense_gru_c: M=%d H=%d Hp=%d ldx=%d", M, H, Hp, ldx);
  if (M == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_c_kernel<float><<<grid_for(total), 256, 0, s>>>(xg, ldx, hc, bias_c, u, h_f32, hnext_f32, (float*)hnext_op, M, H, Hp);
  else dense_gru_c_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(xg, ldx, hc, bias_c, u, h_f32, hnext_f32, (__nv_bfloat16*)hnext_op, M, H, Hp);
  return check_launch("rgp_dense_gru_c");
}

int rgp_dense_logits(const float* acc, int ldacc, const float* bias, float* logits, int B, int T, int N, void* stream) {
  RGP_CHECK_ARG(B >= 0 && T > 0 && N > 0 && ldacc >= N, "rgp_dense_logits: B=%d T=%d N=%d ldacc=%d", B, T, N, ldacc);
  if (B == 0) return 0;
  dense_logits_kernel<<<grid_for((long long)B * T * N), 256, 0, (cudaStream_t)stream>>>(acc, ldacc, bias, logits, B, T, N);
  return check_launch("rgp_dense_logits");
}

}  // extern "C"

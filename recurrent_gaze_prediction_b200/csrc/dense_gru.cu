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

// tf.nn.dropout on the embedded features (gaze_rnn.py:302-303; live only while the dense variant trains, :529): element
// (row r, column j) is kept with probability keep_prob and scaled by 1/keep_prob.  The keep decision is a counter-based
// hash of (seed, r*ldx + j), so the backward pass (dense_unpack_kernel) re-derives the mask instead of storing it.
__device__ __forceinline__ bool dropout_keep(unsigned long long seed, long long idx, float keep_prob) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (unsigned long long)(idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f) < keep_prob;
}

template <class T>
__global__ void dense_pack_kernel(const float* __restrict__ acc, int ldacc, const float* __restrict__ bias,
                                  T* __restrict__ X, int ldx, long long rows, int P, int Cp, float keep_prob,
                                  unsigned long long seed) {
  const long long total = rows * ldx;
  const float inv = keep_prob < 1.0f ? 1.0f / keep_prob : 1.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / ldx;
    const int j = (int)(i - r * ldx);
    float v = 0.0f;
    if (j < P * Cp) {
      const int pos = j / Cp, c = j - pos * Cp;
      v = acc[(r * P + pos) * ldacc + c] + bias[c];
      if (keep_prob < 1.0f) v = dropout_keep(seed, i, keep_prob) ? v * inv : 0.0f;
    }
    X[i] = to_op<T>(v);
  }
}

// backward of dense_pack_kernel: dacc[(r*P + pos), c] = dX[r, pos*Cp + c] (x the dropout mask / keep_prob), operand dtype
template <class T>
__global__ void dense_unpack_kernel(const float* __restrict__ dX, int ldx, T* __restrict__ dacc, long long rows, int P,
                                    int Cp, float keep_prob, unsigned long long seed) {
  const long long total = rows * P * Cp;
  const float inv = keep_prob < 1.0f ? 1.0f / keep_prob : 1.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / (P * Cp);
    const int j = (int)(i - r * (P * Cp));
    float v = dX[r * ldx + j];
    if (keep_prob < 1.0f) v = dropout_keep(seed, r * ldx + j, keep_prob) ? v * inv : 0.0f;
    dacc[i] = to_op<T>(v);        // i = (r*P + pos)*Cp + c with j = pos*Cp + c
  }
}

// dlogits [B, T, N] fp32 -> time-major operand rows dL[(t*B + b), 0..Np) (zero padded): the Y operand of dW_out and the
// A operand of dL . W_out^T
template <class T>
__global__ void dense_dlogits_kernel(const float* __restrict__ dl, T* __restrict__ out, int B, int Tn, int N, int Np) {
  const long long total = (long long)B * Tn * Np;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Np;
    const int j = (int)(i - m * Np);
    const int t = (int)(m / B), b = (int)(m - (long long)t * B);
    out[i] = to_op<T>(j < N ? dl[((long long)b * Tn + t) * N + j] : 0.0f);
  }
}

// BPTT of one dense-GRU step (tf.gradients of  h' = u h + (1-u) c,  c = tanh(.. + (r h) Wc),  [r,u] = sig(..)):
// given dh = dL/dh' (= head term + recurrent term), writes d(pre-activations) of u and c into dG[:, Hp:2Hp] and
// dG[:, 2Hp:3Hp] (operand dtype) and the direct term dh*u of dL/dh into dhpart.
template <class T>
__global__ void dense_gru_bwd_c_kernel(const float* __restrict__ dh_head, const float* __restrict__ dh_rec,
                                       const float* __restrict__ u, const float* __restrict__ c,
                                       const float* __restrict__ hprev, T* __restrict__ dG, float* __restrict__ dhpart,
                                       long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float du = 0.f, dc = 0.f, dp = 0.f;
    if (j < H) {
      const float dh = dh_head[i] + (dh_rec ? dh_rec[i] : 0.f);
      const float uu = u[i], cc = c[i];
      du = dh * (hprev[i] - cc) * uu * (1.0f - uu);
      dc = dh * (1.0f - uu) * (1.0f - cc * cc);
      dp = dh * uu;
    }
    dG[m * 3 * Hp + Hp + j] = to_op<T>(du);
    dG[m * 3 * Hp + 2 * Hp + j] = to_op<T>(dc);
    dhpart[i] = dp;
  }
}

// ... then, given drh = dL/d(r h) = dpre_c . Wc_h^T: dpre_r = drh * h * r (1-r) into dG[:, 0:Hp]; dhpart += drh * r
template <class T>
__global__ void dense_gru_bwd_r_kernel(const float* __restrict__ drh, const float* __restrict__ r,
                                       const float* __restrict__ hprev, T* __restrict__ dG, float* __restrict__ dhpart,
                                       long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float dr = 0.f;
    if (j < H) {
      const float rr = r[i], d = drh[i];
      dr = d * hprev[i] * rr * (1.0f - rr);
      dhpart[i] += d * rr;
    }
    dG[m * 3 * Hp + j] = to_op<T>(dr);
  }
}

template <class T>
__global__ void dense_gru_ru_kernel(const float* __restrict__ xg, int ldx, const float* __restrict__ hg,
                                    const float* __restrict__ bias_g, const float* __restrict__ h,
                                    T* __restrict__ rh, float* __restrict__ u_out, float* __restrict__ r_save,
                                    long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float rhv = 0.0f, uv = 0.0f, rv = 0.0f;
    if (j < H) {
      const float pr = xg[m * ldx + j] + hg[m * 2 * Hp + j] + bias_g[j];
      const float pu = xg[m * ldx + Hp + j] + hg[m * 2 * Hp + Hp + j] + bias_g[Hp + j];
      rv = sigmoidf_(pr);
      rhv = rv * h[i];
      uv = sigmoidf_(pu);
    }
    rh[i] = to_op<T>(rhv);
    u_out[i] = uv;
    if (r_save) r_save[i] = rv;
  }
}

template <class T>
__global__ void dense_gru_c_kernel(const float* __restrict__ xg, int ldx, const float* __restrict__ hc,
                                   const float* __restrict__ bias_c, const float* __restrict__ u,
                                   const float* __restrict__ h, float* __restrict__ hn, T* __restrict__ hn_op,
                                   float* __restrict__ c_save, long long M, int H, int Hp) {
  const long long total = M * Hp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / Hp;
    const int j = (int)(i - m * Hp);
    float v = 0.0f, c = 0.0f;
    if (j < H) {
      c = tanhf(xg[m * ldx + 2 * Hp + j] + hc[i] + bias_c[j]);
      const float uu = u[i];
      v = uu * h[i] + (1.0f - uu) * c;
    }
    hn[i] = v;
    hn_op[i] = to_op<T>(v);
    if (c_save) c_save[i] = c;
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

int rgp_dense_pack_dropout(const float* acc, int ldacc, const float* bias, void* X, int ldx, long long rows, int P, int Cp,
                           float keep_prob, unsigned long long seed, int dtype, void* stream) {
  RGP_CHECK_ARG(rows >= 0 && P > 0 && Cp > 0 && ldacc >= Cp && ldx >= P * Cp, "rgp_dense_pack: rows=%lld P=%d Cp=%d ldacc=%d ldx=%d",
                rows, P, Cp, ldacc, ldx);
  RGP_CHECK_ARG(keep_prob > 0.f && keep_prob <= 1.f, "rgp_dense_pack_dropout: keep_prob=%f", (double)keep_prob);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == 0) dense_pack_kernel<float><<<grid_for(rows * ldx), 256, 0, s>>>(acc, ldacc, bias, (float*)X, ldx, rows, P, Cp, keep_prob, seed);
  else dense_pack_kernel<__nv_bfloat16><<<grid_for(rows * ldx), 256, 0, s>>>(acc, ldacc, bias, (__nv_bfloat16*)X, ldx, rows, P, Cp, keep_prob, seed);
  return check_launch("rgp_dense_pack");
}

int rgp_dense_unpack(const float* dX, int ldx, void* dacc, long long rows, int P, int Cp, float keep_prob,
                     unsigned long long seed, int dtype, void* stream) {
  RGP_CHECK_ARG(rows >= 0 && P > 0 && Cp > 0 && ldx >= P * Cp, "rgp_dense_unpack: rows=%lld P=%d Cp=%d ldx=%d", rows, P, Cp, ldx);
  RGP_CHECK_ARG(keep_prob > 0.f && keep_prob <= 1.f, "rgp_dense_unpack: keep_prob=%f", (double)keep_prob);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = rows * P * Cp;
  if (dtype == 0) dense_unpack_kernel<float><<<grid_for(total), 256, 0, s>>>(dX, ldx, (float*)dacc, rows, P, Cp, keep_prob, seed);
  else dense_unpack_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(dX, ldx, (__nv_bfloat16*)dacc, rows, P, Cp, keep_prob, seed);
  return check_launch("rgp_dense_unpack");
}

int rgp_dense_dlogits(const float* dlogits, void* out, int B, int T, int N, int Np, int dtype, void* stream) {
  RGP_CHECK_ARG(B > 0 && T > 0 && N > 0 && Np >= N, "rgp_dense_dlogits: B=%d T=%d N=%d Np=%d", B, T, N, Np);
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)B * T * Np;
  if (dtype == 0) dense_dlogits_kernel<float><<<grid_for(total), 256, 0, s>>>(dlogits, (float*)out, B, T, N, Np);
  else dense_dlogits_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(dlogits, (__nv_bfloat16*)out, B, T, N, Np);
  return check_launch("rgp_dense_dlogits");
}

int rgp_dense_gru_bwd_c(const float* dh_head, const float* dh_rec, const float* u, const float* c, const float* hprev,
                        void* dG, float* dhpart, int M, int H, int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M > 0 && H > 0 && Hp >= H && dh_head && u && c && hprev && dG && dhpart, "rgp_dense_gru_bwd_c: M=%d H=%d Hp=%d", M, H, Hp);
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_bwd_c_kernel<float><<<grid_for(total), 256, 0, s>>>(dh_head, dh_rec, u, c, hprev, (float*)dG, dhpart, M, H, Hp);
  else dense_gru_bwd_c_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(dh_head, dh_rec, u, c, hprev, (__nv_bfloat16*)dG, dhpart, M, H, Hp);
  return check_launch("rgp_dense_gru_bwd_c");
}

int rgp_dense_gru_bwd_r(const float* drh, const float* r, const float* hprev, void* dG, float* dhpart, int M, int H,
                        int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M > 0 && H > 0 && Hp >= H && drh && r && hprev && dG && dhpart, "rgp_dense_gru_bwd_r: M=%d H=%d Hp=%d", M, H, Hp);
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_bwd_r_kernel<float><<<grid_for(total), 256, 0, s>>>(drh, r, hprev, (float*)dG, dhpart, M, H, Hp);
  else dense_gru_bwd_r_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(drh, r, hprev, (__nv_bfloat16*)dG, dhpart, M, H, Hp);
  return check_launch("rgp_dense_gru_bwd_r");
}

int rgp_dense_pack(const float* acc, int ldacc, const float* bias, void* X, int ldx, long long rows, int P, int Cp,
                   int dtype, void* stream) {
  return rgp_dense_pack_dropout(acc, ldacc, bias, X, ldx, rows, P, Cp, 1.0f, 0ULL, dtype, stream);
}

int rgp_dense_gru_ru(const float* xg, int ldx, const float* hg, const float* bias_g, const float* h_f32, void* rh_op,
                     float* u_out, float* r_save, int M, int H, int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M >= 0 && H > 0 && Hp >= H && ldx >= 3 * Hp, "rgp_dense_gru_ru: M=%d H=%d Hp=%d ldx=%d", M, H, Hp, ldx);
  if (M == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_ru_kernel<float><<<grid_for(total), 256, 0, s>>>(xg, ldx, hg, bias_g, h_f32, (float*)rh_op, u_out, r_save, M, H, Hp);
  else dense_gru_ru_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(xg, ldx, hg, bias_g, h_f32, (__nv_bfloat16*)rh_op, u_out, r_save, M, H, Hp);
  return check_launch("rgp_dense_gru_ru");
}

int rgp_dense_gru_c(const float* xg, int ldx, const float* hc, const float* bias_c, const float* u, const float* h_f32,
                    float* hnext_f32, void* hnext_op, float* c_save, int M, int H, int Hp, int dtype, void* stream) {
  RGP_CHECK_ARG(M >= 0 && H > 0 && Hp >= H && ldx >= 3 * Hp, "rgp_dense_gru_c: M=%d H=%d Hp=%d ldx=%d", M, H, Hp, ldx);
  if (M == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = (long long)M * Hp;
  if (dtype == 0) dense_gru_c_kernel<float><<<grid_for(total), 256, 0, s>>>(xg, ldx, hc, bias_c, u, h_f32, hnext_f32, (float*)hnext_op, c_save, M, H, Hp);
  else dense_gru_c_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, s>>>(xg, ldx, hc, bias_c, u, h_f32, hnext_f32, (__nv_bfloat16*)hnext_op, c_save, M, H, Hp);
  return check_launch("rgp_dense_gru_c");
}

int rgp_dense_logits(const float* acc, int ldacc, const float* bias, float* logits, int B, int T, int N, void* stream) {
  RGP_CHECK_ARG(B >= 0 && T > 0 && N > 0 && ldacc >= N, "rgp_dense_logits: B=%d T=%d N=%d ldacc=%d", B, T, N, ldacc);
  if (B == 0) return 0;
  dense_logits_kernel<<<grid_for((long long)B * T * N), 256, 0, (cudaStream_t)stream>>>(acc, ldacc, bias, logits, B, T, N);
  return check_launch("rgp_dense_logits");
}

}  // extern "C"

// bwd_ops.cu — bandwidth-bound kernels of the hand-written BPTT (the reference leaves this to
// tf.gradients, base.py:278-281): last-step gate backward, per-tap weight transposes, im2col of dlogits for
// the composed head, weight-space chain rule of the head composition, frozen-BN and bias gradients.
#include "common.cuh"

namespace rgp {

template <typename TA>
__global__ void __launch_bounds__(256) gru_bwd_last_kernel(const float* __restrict__ dhs, const float* __restrict__ gamma,
                                                            float alpha, const float* __restrict__ z,
                                                            const float* __restrict__ c, const float* __restrict__ hprev,
                                                            TA* __restrict__ dg, int ldg, float* __restrict__ dhpart,
                                                            int B, int G, int Ch) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * P * Ch) return;
  int n = (int)(idx % Ch);
  int pos = (int)(idx / Ch);
  int b = pos / P, p = pos - b * P;
  int y = p / G, x = p - y * G;
  long long m = (long long)b * S + y * Wp + x;
  float dh = dhs[m * Ch + n];
  if (gamma) dh *= gamma[n] * alpha;
  gru_gate_bwd<TA>(z, c, hprev, dg, ldg, dhpart, Ch, m, m + Wp + 1, n, dh);
}

template <typename TA>
__global__ void __launch_bounds__(256) transpose_taps_kernel(const float* __restrict__ src, TA* __restrict__ dst, int R, int Cc) {
  __shared__ float tile[32][33];
  const int tap = blockIdx.z;
  const float* s = src + (long long)tap * R * Cc;
  TA* d = dst + (long long)tap * R * Cc;
  int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int i = ty; i < 32; i += 8) {
    int r = r0 + i, c = c0 + tx;
    tile[i][tx] = (r < R && c < Cc) ? s[(long long)r * Cc + c] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    int c = c0 + i, r = r0 + tx;
    if (c < Cc && r < R) OpT<TA>::st(d + (long long)c * R + r, tile[tx][i]);
  }
}

__device__ __forceinline__ void st4_ta(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
__device__ __forceinline__ void st4_ta(__nv_bfloat16* p, const float* v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ void ld4f(const float* p, float* v) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
// patches[(t*B+b)*100 + (i+3)*10 + (j+3), u*19+v] = dlogits[b,t,6i+u-3,6j+v-3]
// One CTA per frame: the 2401-float gradient map is staged in shared memory once, every thread then builds 4 consecutive
// patch elements and stores them as one 8-byte (bf16) / 16-byte (fp32) word.
template <typename TA>
__global__ void __launch_bounds__(256) head_patches_kernel(const float* __restrict__ dl, TA* __restrict__ patches, int B, int T) {
  __shared__ float src[2404];
  const int f = blockIdx.x;   // f = t*B + b
  const int t = f / B, b = f - t * B;
  const float* g = dl + ((long long)b * T + t) * 2401;
  for (int k = threadIdx.x; k < 2401; k += 256) src[k] = g[k];
  __syncthreads();
  for (int it = threadIdx.x; it < 49 * 96; it += 256) {
    const int pos = it / 96, k0 = (it - pos * 96) * 4;
    const int i = pos / 7, j = pos - i * 7;
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + q;
      const int u = k / 19, w = k - u * 19;
      const int Y = 6 * i + u - 3, X = 6 * j + w - 3;
      v[q] = (k < 361 && Y >= 0 && Y < 49 && X >= 0 && X < 49) ? src[Y * 49 + X] : 0.f;
    }
    st4_ta(patches + ((long long)f * 100 + (i + 3) * 10 + (j + 3)) * 384 + k0, v);
  }
}

template <typename TA>
__global__ void head_w123t_kernel(const float* __restrict__ w123, TA* __restrict__ out, int Ch) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (c, k): Ch*384
  if (idx >= Ch * 384) return;
  int k = idx % 384, c = idx / 384;
  OpT<TA>::st(out + idx, k < 361 ? w123[(long long)k * Ch + c] : 0.f);
}

// ---- weight-space chain rule of the head composition (see compose_* in ops.cu) ----
__global__ void cb_dw1_kernel(const float* __restrict__ dw123, const float* __restrict__ w23, float* __restrict__ gw1, int Ch) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (k1,l1,m,c)
  if (idx >= 25 * 64 * Ch) return;
  int c = idx % Ch, r = idx / Ch;
  int m = r & 63, kl = r >> 6;
  int k1 = kl / 5, l1 = kl - k1 * 5;
  float acc = 0.f;
  for (int s = 0; s < 11; ++s)
    for (int t = 0; t < 11; ++t)
      acc = fmaf(dw123[((long long)(2 * k1 + s) * 19 + (2 * l1 + t)) * Ch + c], w23[(s * 11 + t) * 64 + m], acc);
  gw1[idx] += acc;
}
__global__ void cb_dw23_kernel(const float* __restrict__ dw123, const float* __restrict__ w1, float* __restrict__ dw23, int Ch) {
  // one warp per (s,t,m): reduce over k1,l1,c
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= 121 * 64) return;
  int m = warp & 63, st = warp >> 6;
  int s = st / 11, t = st - s * 11;
  float acc = 0.f;
  for (int kl = 0; kl < 25; ++kl) {
    int k1 = kl / 5, l1 = kl - k1 * 5;
    const float* a = dw123 + ((long long)(2 * k1 + s) * 19 + (2 * l1 + t)) * Ch;
    const float* b = w1 + ((long long)kl * 64 + m) * Ch;
    for (int c = lane; c < Ch; c += 32) acc = fmaf(a[c], b[c], acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) dw23[warp] = acc;
}
__global__ void cb_dw2_kernel(const float* __restrict__ dw23, const float* __restrict__ w3f, float* __restrict__ gw2) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (k2,l2,e,m): 25*32*64
  if (idx >= 25 * 32 * 64) return;
  int m = idx & 63, r = idx >> 6;
  int e = r & 31, kl = r >> 5;
  int k2 = kl / 5, l2 = kl - k2 * 5;
  float acc = 0.f;
  for (int k3 = 0; k3 < 7; ++k3)
    for (int l3 = 0; l3 < 7; ++l3)
      acc = fmaf(dw23[((k2 + k3) * 11 + (l2 + l3)) * 64 + m], w3f[(k3 * 7 + l3) * 32 + e], acc);
  gw2[idx] += acc;
}
// one warp per (k3,l3,e): lanes run over the 64 mid channels (coalesced), 25 taps, shuffle reduction.  The earlier
// thread-per-output form ran 1600 serial FMAs on 1568 threads with 256-byte lane strides (222 us per pass).
__global__ void __launch_bounds__(256) cb_dw3f_kernel(const float* __restrict__ dw23, const float* __restrict__ w2, float* __restrict__ dw3f) {
  const int idx = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;   // (k3,l3,e): 49*32
  if (idx >= 49 * 32) return;
  const int e = idx & 31, kl = idx >> 5;
  const int k3 = kl / 7, l3 = kl - k3 * 7;
  float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 5
  for (int t = 0; t < 25; ++t) {
    const int k2 = t / 5, l2 = t - k2 * 5;
    const float* a = dw23 + ((k2 + k3) * 11 + (l2 + l3)) * 64;
    const float* b = w2 + (t * 32 + e) * 64;
    acc0 = fmaf(a[lane], b[lane], acc0);
    acc1 = fmaf(a[lane + 32], b[lane + 32], acc1);
  }
  float acc = acc0 + acc1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) dw3f[idx] = acc;
}
__global__ void cb_dw3_kernel(const float* __restrict__ dw3f, const float* __restrict__ w3, const float* __restrict__ ow,
                              float* __restrict__ gw3, float* __restrict__ gow) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (k3,l3,o,e): 49*12*32
  if (idx < 49 * 12 * 32) {
    int e = idx & 31, r = idx >> 5;
    int o = r % 12, kl = r / 12;
    gw3[idx] += dw3f[kl * 32 + e] * ow[o];
  }
  if (idx < 12) {   // d out_W[o] = sum_{kl,e} dw3f[kl,e] * w3[kl,o,e]
    float acc = 0.f;
    for (int kl = 0; kl < 49; ++kl)
      for (int e = 0; e < 32; ++e) acc = fmaf(dw3f[kl * 32 + e], w3[(kl * 12 + idx) * 32 + e], acc);
    gow[idx] += acc;
  }
}

// dgamma[c] += alpha * sum_{b,pos} dhs[m,c]*h[prow,c]; dbeta[c] += sum dhs[m,c]   (one step; block per 32 channels)
__global__ void __launch_bounds__(256) bn_bwd_kernel(const float* __restrict__ dhs, const float* __restrict__ h,
                                                      float* __restrict__ dgamma, float* __restrict__ dbeta, float alpha,
                                                      int B, int G, int Ch) {
  __shared__ float sg[8][32], sb[8][32];
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  // blockIdx.z = time step (rgp_bn_bwd_all): consecutive B*S-row slabs of dhs / h, consecutive [Ch] rows of dgamma / dbeta
  dhs += (long long)blockIdx.z * B * S * Ch;
  h += (long long)blockIdx.z * B * S * Ch;
  dgamma += (long long)blockIdx.z * Ch;
  dbeta += (long long)blockIdx.z * Ch;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int w = threadIdx.x >> 5;
  float ag = 0.f, ab = 0.f;
  if (c < Ch) {
    for (int pos = blockIdx.y * 8 + w; pos < B * P; pos += 8 * gridDim.y) {
      int b = pos / P, p = pos - b * P;
      int y = p / G, x = p - y * G;
      long long m = (long long)b * S + y * Wp + x;
      float d = dhs[m * Ch + c];
      ag = fmaf(d, h[(m + Wp + 1) * Ch + c], ag);
      ab += d;
    }
  }
  sg[w][threadIdx.x & 31] = ag; sb[w][threadIdx.x & 31] = ab;
  __syncthreads();
  if (w == 0 && c < Ch) {
    for (int k = 1; k < 8; ++k) { ag += sg[k][threadIdx.x]; ab += sb[k][threadIdx.x]; }
    atomicAdd(dgamma + c, ag * alpha);
    atomicAdd(dbeta + c, ab);
  }
}

// 16-byte form: a lane owns 4 consecutive channels, a warp 128 channels of a row, two rows in flight per warp (Ch % 4 == 0)
__global__ void __launch_bounds__(256) bn_bwd_v4_kernel(const float* __restrict__ dhs, const float* __restrict__ h,
                                                         float* __restrict__ dgamma, float* __restrict__ dbeta, float alpha,
                                                         int B, int G, int Ch) {
  __shared__ float sg[8][128], sb[8][128];
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  dhs += (long long)blockIdx.z * B * S * Ch;
  h += (long long)blockIdx.z * B * S * Ch;
  dgamma += (long long)blockIdx.z * Ch;
  dbeta += (long long)blockIdx.z * Ch;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 128 + lane * 4;
  float ag[4] = {0.f, 0.f, 0.f, 0.f}, ab[4] = {0.f, 0.f, 0.f, 0.f};
  if (c < Ch) {
    const int step = 8 * gridDim.y, n = B * P;
    for (int pos = blockIdx.y * 8 + w; pos < n; pos += 2 * step) {
      const int pos2 = pos + step < n ? pos + step : pos;
      const float k2 = pos + step < n ? 1.f : 0.f;
      const int b = pos / P, p = pos - b * P, y = p / G, x = p - y * G;
      const int b2 = pos2 / P, p2 = pos2 - b2 * P, y2 = p2 / G, x2 = p2 - y2 * G;
      const long long m = (long long)b * S + y * Wp + x, m2 = (long long)b2 * S + y2 * Wp + x2;
      float d[4], d2[4], hh[4], hh2[4];
      ld4f(dhs + m * Ch + c, d); ld4f(dhs + m2 * Ch + c, d2);
      ld4f(h + (m + Wp + 1) * Ch + c, hh); ld4f(h + (m2 + Wp + 1) * Ch + c, hh2);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        ag[j] = fmaf(d[j], hh[j], ag[j]); ab[j] += d[j];
        ag[j] = fmaf(k2 * d2[j], hh2[j], ag[j]); ab[j] += k2 * d2[j];
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) { sg[w][lane * 4 + j] = ag[j]; sb[w][lane * 4 + j] = ab[j]; }
  __syncthreads();
  if (threadIdx.x < 128) {
    const int cc = blockIdx.x * 128 + threadIdx.x;
    if (cc < Ch) {
      float a = 0.f, bsum = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) { a += sg[k][threadIdx.x]; bsum += sb[k][threadIdx.x]; }
      atomicAdd(dgamma + cc, a * alpha);
      atomicAdd(dbeta + cc, bsum);
    }
  }
}

template <typename TA>
__global__ void __launch_bounds__(256) colsum_kernel(const TA* __restrict__ X, int ld, int M, int N, float* __restrict__ out) {
  __shared__ float sm[8][32];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int w = threadIdx.x >> 5;
  float acc = 0.f;
  if (c < N)
    for (int m = blockIdx.y * 8 + w; m < M; m += 8 * gridDim.y) acc += OpT<TA>::ld(X + (long long)m * ld + c);
  sm[w][threadIdx.x & 31] = acc;
  __syncthreads();
  if (w == 0 && c < N) {
    for (int k = 1; k < 8; ++k) acc += sm[k][threadIdx.x];
    atomicAdd(out + c, acc);
  }
}

// 16-byte loads: a lane owns 16/sizeof(TA) consecutive columns, a warp a 512-byte row segment, the 8 warps of a CTA stride rows.
template <typename TA>
__global__ void __launch_bounds__(256) colsum_vec_kernel(const TA* __restrict__ X, int ld, int M, int N, float* __restrict__ out) {
  constexpr int V = 16 / (int)sizeof(TA);
  __shared__ float sm[8][32 * V];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + lane) * V;
  float acc[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc[j] = 0.f;
  if (c < N) {
    int m = blockIdx.y * 8 + w;
    const int step = 8 * gridDim.y;
    for (; m + step < M; m += 2 * step) {      // two rows in flight
      const uint4 u0 = *reinterpret_cast<const uint4*>(X + (long long)m * ld + c);
      const uint4 u1 = *reinterpret_cast<const uint4*>(X + (long long)(m + step) * ld + c);
      const TA* e0 = reinterpret_cast<const TA*>(&u0);
      const TA* e1 = reinterpret_cast<const TA*>(&u1);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += OpT<TA>::ld(e0 + j) + OpT<TA>::ld(e1 + j);
    }
    if (m < M) {
      const uint4 u0 = *reinterpret_cast<const uint4*>(X + (long long)m * ld + c);
      const TA* e0 = reinterpret_cast<const TA*>(&u0);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += OpT<TA>::ld(e0 + j);
    }
  }
#pragma unroll
  for (int j = 0; j < V; ++j) sm[w][lane * V + j] = acc[j];
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * V; i += 256) {
    const int cc = blockIdx.x * 32 * V + i;
    if (cc < N) {
      float a = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) a += sm[k][i];
      atomicAdd(out + cc, a);
    }
  }
}

// gaze_grcn77.py:206-211 backward; one warp per position
__global__ void __launch_bounds__(256) fc_head_bwd_kernel(const float* __restrict__ dl, const float* __restrict__ h,
                                                           const float* __restrict__ w, float* __restrict__ dhs,
                                                           float* __restrict__ gw, float* __restrict__ gb, int B, int T,
                                                           int t, int G, int Ch) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B * P) return;
  int b = warp / P, p = warp - b * P;
  int y = p / G, x = p - y * G;
  long long m = (long long)b * S + y * Wp + x;
  float d = dl[((long long)b * T + t) * P + p];
  const float* hr = h + (m + Wp + 1) * Ch;
  for (int c = lane; c < Ch; c += 32) {
    dhs[m * Ch + c] = d * w[c];
    atomicAdd(gw + c, d * hr[c]);
  }
  if (lane == 0) atomicAdd(gb, d);
}

// The same for all T steps in one launch.  A thread owns 4 consecutive channels and walks positions (t, b, p): the dhs row is
// written as it goes, the out_W gradient is accumulated in registers and leaves as 4 atomics per thread - the per-step kernel
// issues Ch atomics per position onto the same Ch addresses (measured 36 us per launch at 32 clips x 14x14, 4.6 ms per pass).
__global__ void __launch_bounds__(256) fc_head_bwd_all_kernel(const float* __restrict__ dl, const float* __restrict__ h1,
                                                               const float* __restrict__ w, float* __restrict__ dhs,
                                                               float* __restrict__ gw, float* __restrict__ gb, int B, int T,
                                                               int G, int Ch) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G, CQ = Ch >> 2;   // a thread owns 4 consecutive channels (16-byte accesses)
  const long long M = (long long)B * S, BP = (long long)B * P, N = (long long)T * BP;
  const int c = (threadIdx.x % CQ) * 4, sub = threadIdx.x / CQ, nsub = 256 / CQ;
  float wc[4], acc[4] = {0.f, 0.f, 0.f, 0.f}, accb = 0.0f;
  ld4f(w + c, wc);
  for (long long pos = (long long)blockIdx.x * nsub + sub; pos < N; pos += (long long)gridDim.x * nsub) {
    const int t = (int)(pos / BP);
    const int r = (int)(pos - (long long)t * BP);
    const int b = r / P, p = r - b * P;
    const int y = p / G, x = p - y * G;
    const long long m = (long long)t * M + (long long)b * S + y * Wp + x;
    const float d = dl[((long long)b * T + t) * P + p];
    float hh[4], o[4];
    ld4f(h1 + (m + Wp + 1) * Ch + c, hh);
#pragma unroll
    for (int j = 0; j < 4; ++j) { o[j] = d * wc[j]; acc[j] = fmaf(d, hh[j], acc[j]); }
    st4_ta(dhs + m * Ch + c, o);
    if (c == 0) accb += d;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) atomicAdd(gw + c + j, acc[j]);
  if (c == 0) atomicAdd(gb, accb);
}


// ConvLSTM gate backward of one step (tf.gradients of gaze_lstm.py:112-133 as written: the candidate shares the
// W_hi convolution with the input gate, the output gate peeps at the OLD cell).  One thread per (clip, position, channel).
//   save = (i, f, g, o, c_{t-1}) of this step; cn = c_t; dh = dhs (head) + dhrec (conv^T of step t+1);
//   dc (in: dL/dc_t from step t+1, out: dL/dc_{t-1}).
template <typename TA>
__global__ void __launch_bounds__(256) lstm_gates_bwd_kernel(
    const float* __restrict__ save, const float* __restrict__ cnext_save, const float* __restrict__ c_final,
    const float* __restrict__ dhs, const float* __restrict__ dhrec, float* __restrict__ dc,
    const float* __restrict__ wci, const float* __restrict__ wcf, const float* __restrict__ wco,
    TA* __restrict__ dg4, TA* __restrict__ dhc, float* __restrict__ gwci, float* __restrict__ gwcf,
    float* __restrict__ gwco, int B, int G, int Ch, int last) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * P * Ch) return;
  // clips are the fastest-varying index across blocks of the same (position, channel) range only through atomics
  int n = (int)(idx % Ch);
  int pos = (int)(idx / Ch);
  int b = pos / P, p = pos - b * P;
  int y = p / G, x = p - y * G;
  long long m = (long long)b * S + y * Wp + x;
  const float* sv = save + m * 5 * Ch;
  const float gi = sv[n], gf = sv[Ch + n], gg = sv[2 * Ch + n], go = sv[3 * Ch + n], cp = sv[4 * Ch + n];
  const float cn = cnext_save ? cnext_save[m * 5 * Ch + 4 * Ch + n] : c_final[(m + Wp + 1) * Ch + n];
  float dh = dhs[m * Ch + n];
  float dcn = 0.f;
  if (!last) { dh += dhrec[m * Ch + n]; dcn = dc[m * Ch + n]; }
  const float tc = tanhf_(cn);
  dcn += dh * go * (1.0f - tc * tc);
  const float dpo = dh * tc * go * (1.0f - go);
  const float dpi = dcn * gg * gi * (1.0f - gi);
  const float dpg = dcn * gi * (1.0f - gg * gg);
  const float dpf = dcn * cp * gf * (1.0f - gf);
  const int pe = p * Ch + n;
  dc[m * Ch + n] = dcn * gf + dpi * wci[pe] + dpf * wcf[pe] + dpo * wco[pe];
  TA* g4 = dg4 + m * 4 * Ch;
  OpT<TA>::st(g4 + n, dpi); OpT<TA>::st(g4 + Ch + n, dpf); OpT<TA>::st(g4 + 2 * Ch + n, dpg); OpT<TA>::st(g4 + 3 * Ch + n, dpo);
  TA* hc = dhc + m * 3 * Ch;
  OpT<TA>::st(hc + n, dpi + dpg); OpT<TA>::st(hc + Ch + n, dpf); OpT<TA>::st(hc + 2 * Ch + n, dpo);
  atomicAdd(gwci + pe, dpi * cp);
  atomicAdd(gwcf + pe, dpf * cp);
  atomicAdd(gwco + pe, dpo * cp);
}

// The same with one thread per (group of 4 clips, position, 4 consecutive channels): 16-byte accesses, and the peephole
// gradients of the 4 clips are summed in registers before they leave as 12 atomics per thread (the scalar kernel issues 3 per
// element, B-way contended).  Element arithmetic unchanged.
template <typename TA>
__global__ void __launch_bounds__(256) lstm_gates_bwd_v4_kernel(
    const float* __restrict__ save, const float* __restrict__ cnext_save, const float* __restrict__ c_final,
    const float* __restrict__ dhs, const float* __restrict__ dhrec, float* __restrict__ dc,
    const float* __restrict__ wci, const float* __restrict__ wcf, const float* __restrict__ wco,
    TA* __restrict__ dg4, TA* __restrict__ dhc, float* __restrict__ gwci, float* __restrict__ gwcf,
    float* __restrict__ gwco, int B, int G, int Ch, int last) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G, C4 = Ch >> 2, BG = (B + 3) >> 2;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)BG * P * C4) return;
  const int n = (int)(idx % C4) * 4;
  const int pos = (int)(idx / C4);
  const int p = pos % P, b0 = (pos / P) * 4;
  const int y = p / G, x = p - y * G;
  const int pe = p * Ch + n;
  float wi[4], wf[4], wo[4], ai[4] = {0.f, 0.f, 0.f, 0.f}, af[4] = {0.f, 0.f, 0.f, 0.f}, ao[4] = {0.f, 0.f, 0.f, 0.f};
  ld4f(wci + pe, wi); ld4f(wcf + pe, wf); ld4f(wco + pe, wo);
#pragma unroll
  for (int bb = 0; bb < 4; ++bb) {
    const int b = b0 + bb;
    if (b >= B) break;
    const long long m = (long long)b * S + y * Wp + x;
    const float* sv = save + m * 5 * Ch + n;
    float gi[4], gf[4], gg[4], go[4], cp[4], cn[4], dh[4], dcn[4] = {0.f, 0.f, 0.f, 0.f};
    ld4f(sv, gi); ld4f(sv + Ch, gf); ld4f(sv + 2 * Ch, gg); ld4f(sv + 3 * Ch, go); ld4f(sv + 4 * Ch, cp);
    if (cnext_save) ld4f(cnext_save + m * 5 * Ch + 4 * Ch + n, cn); else ld4f(c_final + (m + Wp + 1) * Ch + n, cn);
    ld4f(dhs + m * Ch + n, dh);
    if (!last) {
      float dr[4];
      ld4f(dhrec + m * Ch + n, dr); ld4f(dc + m * Ch + n, dcn);
#pragma unroll
      for (int j = 0; j < 4; ++j) dh[j] += dr[j];
    }
    float dpi[4], dpf[4], dpg[4], dpo[4], dco[4], dig[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float tc = tanhf_(cn[j]);
      dcn[j] += dh[j] * go[j] * (1.0f - tc * tc);
      dpo[j] = dh[j] * tc * go[j] * (1.0f - go[j]);
      dpi[j] = dcn[j] * gg[j] * gi[j] * (1.0f - gi[j]);
      dpg[j] = dcn[j] * gi[j] * (1.0f - gg[j] * gg[j]);
      dpf[j] = dcn[j] * cp[j] * gf[j] * (1.0f - gf[j]);
      dco[j] = dcn[j] * gf[j] + dpi[j] * wi[j] + dpf[j] * wf[j] + dpo[j] * wo[j];
      dig[j] = dpi[j] + dpg[j];
      ai[j] += dpi[j] * cp[j]; af[j] += dpf[j] * cp[j]; ao[j] += dpo[j] * cp[j];
    }
    st4_ta(dc + m * Ch + n, dco);
    TA* g4 = dg4 + m * 4 * Ch + n;
    st4_ta(g4, dpi); st4_ta(g4 + Ch, dpf); st4_ta(g4 + 2 * Ch, dpg); st4_ta(g4 + 3 * Ch, dpo);
    TA* hc = dhc + m * 3 * Ch + n;
    st4_ta(hc, dig); st4_ta(hc + Ch, dpf); st4_ta(hc + 2 * Ch, dpo);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    atomicAdd(gwci + pe + j, ai[j]);
    atomicAdd(gwcf + pe + j, af[j]);
    atomicAdd(gwco + pe + j, ao[j]);
  }
}

}  // namespace rgp

using namespace rgp;
#define DISPATCH_TA(dtype, CALL_F32, CALL_BF16) \
  do { if ((dtype) == 0) { CALL_F32; } else { CALL_BF16; } } while (0)

extern "C" {

int rgp_gru_bwd_last(const float* dhs, const float* gamma, float alpha, const float* z, const float* c,
                     const float* hprev_f32, void* dg, int ldg, float* dhpart, int B, int G, int Ch, int dtype,
                     void* stream) {
  long long total = (long long)B * G * G * Ch;
  int grid = (int)((total + 255) / 256);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_TA(dtype,
              (gru_bwd_last_kernel<float><<<grid, 256, 0, s>>>(dhs, gamma, alpha, z, c, hprev_f32, (float*)dg, ldg, dhpart, B, G, Ch)),
              (gru_bwd_last_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(dhs, gamma, alpha, z, c, hprev_f32, (__nv_bfloat16*)dg, ldg, dhpart, B, G, Ch)));
  return check_launch("rgp_gru_bwd_last");
}

int rgp_lstm_gates_bwd(const float* save, const float* cnext_save, const float* c_final, const float* dhs,
                       const float* dhrec, float* dc, const float* wci, const float* wcf, const float* wco, void* dg4,
                       void* dhc, float* gwci, float* gwcf, float* gwco, int B, int G, int Ch, int last, int dtype,
                       void* stream) {
  RGP_CHECK_ARG(cnext_save != nullptr || c_final != nullptr, "rgp_lstm_gates_bwd: need c_t (cnext_save or c_final)");
  RGP_CHECK_ARG(last || dhrec != nullptr, "rgp_lstm_gates_bwd: dhrec is required unless this is the last time step");
  long long total = (long long)B * G * G * Ch;
  cudaStream_t s = (cudaStream_t)stream;
  const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (Ch % 4 == 0 && al16(save) && al16(cnext_save) && al16(c_final) && al16(dhs) && al16(dhrec) && al16(dc) && al16(wci) &&
      al16(wcf) && al16(wco) && al16(dg4) && al16(dhc) && al16(gwci) && al16(gwcf) && al16(gwco)) {
    long long t4 = (long long)((B + 3) / 4) * G * G * (Ch / 4);
    int grid4 = (int)((t4 + 255) / 256);
    DISPATCH_TA(dtype,
                (lstm_gates_bwd_v4_kernel<float><<<grid4, 256, 0, s>>>(save, cnext_save, c_final, dhs, dhrec, dc, wci, wcf, wco, (float*)dg4, (float*)dhc, gwci, gwcf, gwco, B, G, Ch, last)),
                (lstm_gates_bwd_v4_kernel<__nv_bfloat16><<<grid4, 256, 0, s>>>(save, cnext_save, c_final, dhs, dhrec, dc, wci, wcf, wco, (__nv_bfloat16*)dg4, (__nv_bfloat16*)dhc, gwci, gwcf, gwco, B, G, Ch, last)));
    return check_launch("rgp_lstm_gates_bwd");
  }
  int grid = (int)((total + 255) / 256);
  DISPATCH_TA(dtype,
              (lstm_gates_bwd_kernel<float><<<grid, 256, 0, s>>>(save, cnext_save, c_final, dhs, dhrec, dc, wci, wcf, wco, (float*)dg4, (float*)dhc, gwci, gwcf, gwco, B, G, Ch, last)),
              (lstm_gates_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(save, cnext_save, c_final, dhs, dhrec, dc, wci, wcf, wco, (__nv_bfloat16*)dg4, (__nv_bfloat16*)dhc, gwci, gwcf, gwco, B, G, Ch, last)));
  return check_launch("rgp_lstm_gates_bwd");
}

int rgp_transpose_taps(const float* src, void* dst, int ntaps, int R, int Cc, int dtype, void* stream) {
  dim3 grid((Cc + 31) / 32, (R + 31) / 32, ntaps);
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_TA(dtype, (transpose_taps_kernel<float><<<grid, 256, 0, s>>>(src, (float*)dst, R, Cc)),
              (transpose_taps_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(src, (__nv_bfloat16*)dst, R, Cc)));
  return check_launch("rgp_transpose_taps");
}

int rgp_head_patches(const float* dlogits, void* patches, int B, int T, int dtype, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int grid = B * T;
  DISPATCH_TA(dtype, (head_patches_kernel<float><<<grid, 256, 0, s>>>(dlogits, (float*)patches, B, T)),
              (head_patches_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(dlogits, (__nv_bfloat16*)patches, B, T)));
  return check_launch("rgp_head_patches");
}

int rgp_head_w123t(const float* w123, void* out, int Ch, int dtype, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int grid = (Ch * 384 + 255) / 256;
  DISPATCH_TA(dtype, (head_w123t_kernel<float><<<grid, 256, 0, s>>>(w123, (float*)out, Ch)),
              (head_w123t_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(w123, (__nv_bfloat16*)out, Ch)));
  return check_launch("rgp_head_w123t");
}

int rgp_head_compose_bwd(const float* dw123, const float* w1, const float* w2, const float* w3, const float* out_w,
                         const float* w3f, const float* w23, float* dw23, float* dw3f, float* gw1, float* gw2,
                         float* gw3, float* gout_w, int Ch, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  cb_dw1_kernel<<<(25 * 64 * Ch + 255) / 256, 256, 0, s>>>(dw123, w23, gw1, Ch);
  cb_dw23_kernel<<<(121 * 64 * 32 + 255) / 256, 256, 0, s>>>(dw123, w1, dw23, Ch);
  cb_dw2_kernel<<<(25 * 32 * 64 + 255) / 256, 256, 0, s>>>(dw23, w3f, gw2);
  cb_dw3f_kernel<<<(49 * 32 * 32 + 255) / 256, 256, 0, s>>>(dw23, w2, dw3f);
  cb_dw3_kernel<<<(49 * 12 * 32 + 255) / 256, 256, 0, s>>>(dw3f, w3, out_w, gw3, gout_w);
  return check_launch("rgp_head_compose_bwd");
}

int rgp_bn_bwd(const float* dhs, const float* h_f32, float* dgamma, float* dbeta, float alpha, int B, int G, int Ch,
               void* stream) {
  int gy = (B * G * G + 63) / 64;
  if (gy > 64) gy = 64;
  dim3 grid((Ch + 31) / 32, gy);
  bn_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dhs, h_f32, dgamma, dbeta, alpha, B, G, Ch);
  return check_launch("rgp_bn_bwd");
}

int rgp_bn_bwd_all(const float* dhs, const float* h_f32, float* dgamma, float* dbeta, float alpha, int B, int T, int G,
                   int Ch, void* stream) {
  if (Ch % 4 == 0 && ((reinterpret_cast<uintptr_t>(dhs) | reinterpret_cast<uintptr_t>(h_f32)) & 15) == 0) {
    int gyv = (B * G * G + 31) / 32;
    if (gyv > 32) gyv = 32;
    if (gyv < 1) gyv = 1;
    dim3 gridv((Ch + 127) / 128, gyv, T);
    bn_bwd_v4_kernel<<<gridv, 256, 0, (cudaStream_t)stream>>>(dhs, h_f32, dgamma, dbeta, alpha, B, G, Ch);
    return check_launch("rgp_bn_bwd_all");
  }
  int gy = (B * G * G + 63) / 64;
  if (gy > 16) gy = 16;
  if (gy < 1) gy = 1;
  dim3 grid((Ch + 31) / 32, gy, T);
  bn_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dhs, h_f32, dgamma, dbeta, alpha, B, G, Ch);
  return check_launch("rgp_bn_bwd_all");
}

int rgp_colsum(const void* X, int ld, int M, int N, float* out, int dtype, void* stream) {
  int gy = (M + 511) / 512;
  if (gy > 296) gy = 296;
  if (gy < 1) gy = 1;
  cudaStream_t s = (cudaStream_t)stream;
  const int V = dtype == 0 ? 4 : 8;
  if (N % V == 0 && ld % V == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0) {
    int gyv = (M + 255) / 256;
    if (gyv > 592) gyv = 592;
    if (gyv < 1) gyv = 1;
    dim3 gridv((N + 32 * V - 1) / (32 * V), gyv);
    DISPATCH_TA(dtype, (colsum_vec_kernel<float><<<gridv, 256, 0, s>>>((const float*)X, ld, M, N, out)),
                (colsum_vec_kernel<__nv_bfloat16><<<gridv, 256, 0, s>>>((const __nv_bfloat16*)X, ld, M, N, out)));
    return check_launch("rgp_colsum");
  }
  dim3 grid((N + 31) / 32, gy);
  DISPATCH_TA(dtype, (colsum_kernel<float><<<grid, 256, 0, s>>>((const float*)X, ld, M, N, out)),
              (colsum_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)X, ld, M, N, out)));
  return check_launch("rgp_colsum");
}

int rgp_fc_head_bwd(const float* dlogits, const float* h_f32, const float* out_W, float* dhs, float* gout_W,
                    float* gout_b, int B, int T, int t, int G, int Ch, void* stream) {
  int warps = B * G * G;
  int grid = (warps * 32 + 255) / 256;
  fc_head_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dlogits, h_f32, out_W, dhs, gout_W, gout_b, B, T, t, G, Ch);
  return check_launch("rgp_fc_head_bwd");
}

int rgp_fc_head_bwd_all(const float* dlogits, const float* h1_f32, const float* out_W, float* dhs, float* gout_W,
                        float* gout_b, int B, int T, int G, int Ch, void* stream) {
  RGP_CHECK_ARG(Ch >= 4 && Ch % 4 == 0 && 256 % (Ch / 4) == 0, "rgp_fc_head_bwd_all: Ch/4 (Ch=%d) must divide 256", Ch);
  RGP_CHECK_ARG(((reinterpret_cast<uintptr_t>(h1_f32) | reinterpret_cast<uintptr_t>(out_W) | reinterpret_cast<uintptr_t>(dhs)) & 15) == 0,
                "rgp_fc_head_bwd_all: h1_f32, out_W and dhs must be 16-byte aligned%s", "");
  RGP_CHECK_ARG(B > 0 && T > 0 && G > 0, "rgp_fc_head_bwd_all: B=%d T=%d G=%d", B, T, G);
  const long long N = (long long)T * B * G * G;
  const int nsub = 256 / (Ch / 4);
  long long grid = (N + nsub - 1) / nsub;
  if (grid > 148 * 8) grid = 148 * 8;
  fc_head_bwd_all_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(dlogits, h1_f32, out_W, dhs, gout_W, gout_b, B, T, G, Ch);
  return check_launch("rgp_fc_head_bwd_all");
}

}  // extern "C"

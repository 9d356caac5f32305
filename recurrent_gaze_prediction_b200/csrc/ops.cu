// ops.cu — bandwidth-bound kernels of the RGP path: layout packing, gate maths of the ConvLSTM,
// head weight composition, spatial softmax + loss, gradient norm + clip + TF-Adam.
#include <mutex>
#include "common.cuh"

namespace rgp {

// ---------------------------------------------------------------------------------------------
// pack: gaze_grcn.py:225-227 (NCHW -> NHWC transpose of the C3D blob) + cast + time-major reorder
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void st4_op(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void st4_op(__nv_bfloat16* p, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}

// One CTA = one frame x 64 channels.  In: 64 x P floats, contiguous (16-byte loads, copied linearly into shared memory:
// conflict-free).  Out: P rows x 64 channels; a thread converts 4 consecutive channels of one position (4 shared reads at
// stride P, 2-way bank conflicts) and stores them as one 8-byte (bf16) / 16-byte (fp32) word, so a warp writes two full
// 128-byte row segments per instruction instead of 64 bytes of 2-byte stores.
//
// TS = type of the fed features: float (the reference feeds fp32, gaze_rnn.py:520-531) or bf16 (half the host->device
// bytes; the bf16 engine rounds the features to bf16 here anyway, so the packed operand is bit-identical).
template <typename TA, typename TS>
__global__ void __launch_bounds__(256) pack_c3d_kernel(const TS* __restrict__ c3d, TA* __restrict__ out,
                                                        int B, int T, int C, int P) {
  extern __shared__ float tile[];  // [64][P]
  const int f_in = blockIdx.x;     // b*T + t
  const int b = f_in / T, t = f_in - b * T;
  const int c0 = blockIdx.y * 64;
  const TS* src = c3d + ((long long)f_in * C + c0) * P;
  const int n = 64 * P;
  if (sizeof(TS) == 4 && ((n | (int)(((long long)f_in * C + c0) * P)) & 3) == 0) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* t4 = reinterpret_cast<float4*>(tile);
    for (int i = threadIdx.x; i < n / 4; i += 256) t4[i] = s4[i];
  } else if (sizeof(TS) == 2 && ((n | (int)(((long long)f_in * C + c0) * P)) & 7) == 0) {
    const uint4* s8 = reinterpret_cast<const uint4*>(src);       // 8 bf16 per 16-byte load
    for (int i = threadIdx.x; i < n / 8; i += 256) {
      const uint4 u = s8[i];
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        tile[8 * i + 2 * j] = __uint_as_float(w[j] << 16);
        tile[8 * i + 2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
      }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += 256) tile[i] = OpT<TS>::ld(src + i);
  }
  __syncthreads();
  TA* dst = out + ((long long)(t * B + b) * P) * C + c0;
  for (int i = threadIdx.x; i < 16 * P; i += 256) {
    const int p = i >> 4, c = (i & 15) * 4;
    st4_op(dst + (long long)p * C + c, tile[c * P + p], tile[(c + 1) * P + p], tile[(c + 2) * P + p], tile[(c + 3) * P + p]);
  }
}

// emb [B,T,G,G,C] fp32 -> p-layout [(t*B+b)*S + (y+1)*Wp + x+1, C]
template <typename TA>
__global__ void __launch_bounds__(256) pack_emb_kernel(const float* __restrict__ emb, TA* __restrict__ out,
                                                        int B, int T, int G, int C) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;  // one thread per 4 channels
  long long total = (long long)B * T * P * (C / 4);
  if (idx >= total) return;
  int c4 = (int)(idx % (C / 4));
  long long pos = idx / (C / 4);
  int p = (int)(pos % P);
  long long f_in = pos / P;
  int b = (int)(f_in / T), t = (int)(f_in - (long long)b * T);
  int y = p / G, x = p - y * G;
  float4 v = *reinterpret_cast<const float4*>(emb + pos * C + c4 * 4);
  TA* d = out + (((long long)(t * B + b)) * S + (y + 1) * Wp + (x + 1)) * C + c4 * 4;
  OpT<TA>::st(d, v.x); OpT<TA>::st(d + 1, v.y); OpT<TA>::st(d + 2, v.z); OpT<TA>::st(d + 3, v.w);
}

template <typename TA>
__global__ void cast_kernel(const float* __restrict__ s, TA* __restrict__ d, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) OpT<TA>::st(d + i, s[i]);
}

// ---------------------------------------------------------------------------------------------
// gaze_grcn77.py:206-211 : per-position FC Ch -> 1 on the fp32 state (p-layout), one warp per position
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fc_head_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                                       const float* __restrict__ bias, float* __restrict__ logits,
                                                       int B, int T, int t, int G, int Ch) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B * P) return;
  int b = warp / P, p = warp - b * P;
  int y = p / G, x = p - y * G;
  const float* hr = h + ((long long)b * S + (y + 1) * Wp + (x + 1)) * Ch;
  float s = 0.f;
  for (int c = lane; c < Ch; c += 32) s = fmaf(hr[c], w[c], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) logits[((long long)b * T + t) * P + p] = s + bias[0];
}

// The same for all T steps in one launch (h1 = slab of h_1; slab t holds h_{t+1}); 16-byte loads, Ch % 4 == 0.
__global__ void __launch_bounds__(256) fc_head_all_kernel(const float* __restrict__ h1, const float* __restrict__ w,
                                                           const float* __restrict__ bias, float* __restrict__ logits,
                                                           int B, int T, int G, int Ch) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  const long long BP = (long long)B * P;
  const long long warp = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= (long long)T * BP) return;
  const int t = (int)(warp / BP);
  const int r = (int)(warp - (long long)t * BP);
  const int b = r / P, p = r - b * P;
  const int y = p / G, x = p - y * G;
  const float* hr = h1 + (((long long)t * B + b) * S + (y + 1) * Wp + (x + 1)) * Ch;
  float s = 0.f;
  for (int c = lane * 4; c < Ch; c += 128) {
    const float4 a = *reinterpret_cast<const float4*>(hr + c), ww = *reinterpret_cast<const float4*>(w + c);
    s = fmaf(a.x, ww.x, s); s = fmaf(a.y, ww.y, s); s = fmaf(a.z, ww.z, s); s = fmaf(a.w, ww.w, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) logits[((long long)b * T + t) * P + p] = s + bias[0];
}

// ---------------------------------------------------------------------------------------------
// ConvLSTM gates — gaze_lstm.py:112-133, literally (candidate reuses the W_hi conv; output gate
// peeps at the OLD cell; peepholes are full [G,G,Ch] tensors).
// xg [M,4Ch] = conv(x,{W_xi,W_xf,W_xc,W_xo}); hconv [M,3Ch] = conv(h,{W_hi,W_hf,W_ho}); q-layout rows.
// save (optional, training) [M,5Ch] = i, f, g(tanh candidate), o, c_prev.
// ---------------------------------------------------------------------------------------------
template <typename TA>
__global__ void __launch_bounds__(256) lstm_gates_kernel(const float* __restrict__ xg, int ldx,
                                                          const float* __restrict__ hconv,
                                                          const float* __restrict__ wci, const float* __restrict__ wcf,
                                                          const float* __restrict__ wco, float* __restrict__ c_f32,
                                                          float* __restrict__ h_f32, TA* __restrict__ h_op_next,
                                                          TA* __restrict__ hs_op, int B, int G, int Ch, int So, int Wo,
                                                          int pado, float* __restrict__ save) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * P * Ch) return;
  int n = (int)(idx % Ch);
  int pos = (int)(idx / Ch);
  int b = pos / P, p = pos - b * P;
  int y = p / G, x = p - y * G;
  long long m = (long long)b * S + y * Wp + x;
  long long prow = m + Wp + 1;
  const float* xr = xg + m * ldx;
  const float* hr = hconv + m * 3 * Ch;
  float cp = c_f32[prow * Ch + n];
  int pe = (y * G + x) * Ch + n;
  float hi = hr[n];
  float gi = sigmoidf_(xr[n] + hi + wci[pe] * cp);
  float gf = sigmoidf_(xr[Ch + n] + hr[Ch + n] + wcf[pe] * cp);
  float gg = tanhf_(xr[2 * Ch + n] + hi);
  float cn = gf * cp + gi * gg;
  float go = sigmoidf_(xr[3 * Ch + n] + hr[2 * Ch + n] + wco[pe] * cp);
  float hn = tanhf_(cn) * go;
  c_f32[prow * Ch + n] = cn;
  h_f32[prow * Ch + n] = hn;
  OpT<TA>::st(h_op_next + prow * Ch + n, hn);
  if (hs_op) {
    long long orow = (long long)b * So + (y + pado) * Wo + (x + pado);
    OpT<TA>::st(hs_op + orow * Ch + n, hn);
  }
  if (save) {
    float* sv = save + m * 5 * Ch;
    sv[n] = gi; sv[Ch + n] = gf; sv[2 * Ch + n] = gg; sv[3 * Ch + n] = go; sv[4 * Ch + n] = cp;
  }
}

// The same, one thread per 4 consecutive channels (Ch % 4 == 0, ldx % 4 == 0): 16-byte loads and stores, element arithmetic
// unchanged.  The scalar kernel moved ~70 MB per step at 2.8 TB/s (24.8 us per step at 64 clips x 14x14 x 128).
template <typename TA>
__global__ void __launch_bounds__(256) lstm_gates_v4_kernel(const float* __restrict__ xg, int ldx,
                                                             const float* __restrict__ hconv,
                                                             const float* __restrict__ wci, const float* __restrict__ wcf,
                                                             const float* __restrict__ wco, float* __restrict__ c_f32,
                                                             float* __restrict__ h_f32, TA* __restrict__ h_op_next,
                                                             TA* __restrict__ hs_op, int B, int G, int Ch, int So, int Wo,
                                                             int pado, float* __restrict__ save) {
  const int Wp = G + 1, S = Wp * Wp, P = G * G, C4 = Ch >> 2;
  long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * P * C4) return;
  const int n = (int)(idx % C4) * 4;
  const int pos = (int)(idx / C4);
  const int b = pos / P, p = pos - b * P;
  const int y = p / G, x = p - y * G;
  const long long m = (long long)b * S + y * Wp + x;
  const long long prow = m + Wp + 1;
  const float* xr = xg + m * ldx + n;
  const float* hr = hconv + m * 3 * Ch + n;
  const int pe = p * Ch + n;
  const float4 cp4 = *reinterpret_cast<const float4*>(c_f32 + prow * Ch + n);
  const float4 xi4 = *reinterpret_cast<const float4*>(xr), xf4 = *reinterpret_cast<const float4*>(xr + Ch);
  const float4 xc4 = *reinterpret_cast<const float4*>(xr + 2 * Ch), xo4 = *reinterpret_cast<const float4*>(xr + 3 * Ch);
  const float4 hi4 = *reinterpret_cast<const float4*>(hr), hf4 = *reinterpret_cast<const float4*>(hr + Ch);
  const float4 ho4 = *reinterpret_cast<const float4*>(hr + 2 * Ch);
  const float4 wi4 = *reinterpret_cast<const float4*>(wci + pe), wf4 = *reinterpret_cast<const float4*>(wcf + pe);
  const float4 wo4 = *reinterpret_cast<const float4*>(wco + pe);
  const float cp[4] = {cp4.x, cp4.y, cp4.z, cp4.w}, xi[4] = {xi4.x, xi4.y, xi4.z, xi4.w}, xf[4] = {xf4.x, xf4.y, xf4.z, xf4.w};
  const float xc[4] = {xc4.x, xc4.y, xc4.z, xc4.w}, xo[4] = {xo4.x, xo4.y, xo4.z, xo4.w}, hi[4] = {hi4.x, hi4.y, hi4.z, hi4.w};
  const float hf[4] = {hf4.x, hf4.y, hf4.z, hf4.w}, ho[4] = {ho4.x, ho4.y, ho4.z, ho4.w}, wi[4] = {wi4.x, wi4.y, wi4.z, wi4.w};
  const float wf[4] = {wf4.x, wf4.y, wf4.z, wf4.w}, wo[4] = {wo4.x, wo4.y, wo4.z, wo4.w};
  float gi[4], gf[4], gg[4], go[4], cn[4], hn[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    gi[j] = sigmoidf_(xi[j] + hi[j] + wi[j] * cp[j]);
    gf[j] = sigmoidf_(xf[j] + hf[j] + wf[j] * cp[j]);
    gg[j] = tanhf_(xc[j] + hi[j]);
    cn[j] = gf[j] * cp[j] + gi[j] * gg[j];
    go[j] = sigmoidf_(xo[j] + ho[j] + wo[j] * cp[j]);
    hn[j] = tanhf_(cn[j]) * go[j];
  }
  *reinterpret_cast<float4*>(c_f32 + prow * Ch + n) = make_float4(cn[0], cn[1], cn[2], cn[3]);
  *reinterpret_cast<float4*>(h_f32 + prow * Ch + n) = make_float4(hn[0], hn[1], hn[2], hn[3]);
  st4_op(h_op_next + prow * Ch + n, hn[0], hn[1], hn[2], hn[3]);
  if (hs_op) {
    const long long orow = (long long)b * So + (y + pado) * Wo + (x + pado);
    st4_op(hs_op + orow * Ch + n, hn[0], hn[1], hn[2], hn[3]);
  }
  if (save) {
    float* sv = save + m * 5 * Ch + n;
    *reinterpret_cast<float4*>(sv) = make_float4(gi[0], gi[1], gi[2], gi[3]);
    *reinterpret_cast<float4*>(sv + Ch) = make_float4(gf[0], gf[1], gf[2], gf[3]);
    *reinterpret_cast<float4*>(sv + 2 * Ch) = make_float4(gg[0], gg[1], gg[2], gg[3]);
    *reinterpret_cast<float4*>(sv + 3 * Ch) = make_float4(go[0], go[1], go[2], go[3]);
    *reinterpret_cast<float4*>(sv + 4 * Ch) = cp4;
  }
}

// ---------------------------------------------------------------------------------------------
// Head weight composition (weight space).  The head of gaze_grcn.py:326-366 is linear:
// deconv 5x5/s3 (Ch->64) o deconv 5x5/s2 (64->32) o deconv 7x7/s1 SAME (32->12) o FC 12->1.
// Composite = one transposed conv 19x19 / stride 6 (Ch -> 1), VALID 7->55, cropped by 3 to 49:
//   logit[Y,X] = sum_{i,j,c} hs[i,j,c] * W123[Y+3-6i, X+3-6j, c] + out_b
// ---------------------------------------------------------------------------------------------
__global__ void compose_w3f_kernel(const float* __restrict__ w3, const float* __restrict__ ow, float* __restrict__ w3f) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;  // (k3,l3,e): 49*32
  if (i >= 49 * 32) return;
  int e = i & 31, kl = i >> 5;
  float s = 0.f;
  for (int o = 0; o < 12; ++o) s = fmaf(w3[(kl * 12 + o) * 32 + e], ow[o], s);
  w3f[i] = s;
}
__global__ void compose_w23_kernel(const float* __restrict__ w2, const float* __restrict__ w3f, float* __restrict__ w23) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;  // (s,t,m): 121*64
  if (i >= 121 * 64) return;
  int m = i & 63, st = i >> 6;
  int s = st / 11, t = st - s * 11;
  float acc = 0.f;
  for (int k2 = 0; k2 < 5; ++k2) {
    int k3 = s - k2;
    if (k3 < 0 || k3 > 6) continue;
    for (int l2 = 0; l2 < 5; ++l2) {
      int l3 = t - l2;
      if (l3 < 0 || l3 > 6) continue;
      const float* a = w2 + ((k2 * 5 + l2) * 32) * 64 + m;  // w2[k2,l2,e,m]
      const float* b = w3f + (k3 * 7 + l3) * 32;
      for (int e = 0; e < 32; ++e) acc = fmaf(a[e * 64], b[e], acc);
    }
  }
  w23[i] = acc;
}
__global__ void compose_w123_kernel(const float* __restrict__ w1, const float* __restrict__ w23, float* __restrict__ w123, int Ch) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;  // (u,v,c): 361*Ch
  if (i >= 361 * Ch) return;
  int c = i % Ch, uv = i / Ch;
  int u = uv / 19, v = uv - u * 19;
  float acc = 0.f;
  for (int k1 = 0; k1 < 5; ++k1) {
    int s = u - 2 * k1;
    if (s < 0 || s > 10) continue;
    for (int l1 = 0; l1 < 5; ++l1) {
      int t = v - 2 * l1;
      if (t < 0 || t > 10) continue;
      const float* a = w1 + ((long long)(k1 * 5 + l1) * 64) * Ch + c;  // w1[k1,l1,m,c]
      const float* b = w23 + (s * 11 + t) * 64;
      for (int m = 0; m < 64; ++m) acc = fmaf(a[(long long)m * Ch], b[m], acc);
    }
  }
  w123[i] = acc;
}
template <typename TA>
__global__ void compose_bhead_kernel(const float* __restrict__ w123, TA* __restrict__ bhead, int Ch) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;  // (tap, n, c): 16*48*Ch
  if (i >= 16 * 48 * Ch) return;
  int c = i % Ch, tn = i / Ch;
  int n = tn % 48, tap = tn / 48;
  int dy = tap >> 2, dx = tap & 3;
  float v = 0.f;
  if (n < 36) {
    int ry = n / 6, rx = n - ry * 6;
    int u = 6 * dy + ry, w = 6 * dx + rx;
    if (u < 19 && w < 19) v = w123[(u * 19 + w) * Ch + c];
  }
  OpT<TA>::st(bhead + i, v);
}

// ---------------------------------------------------------------------------------------------
// Spatial softmax + loss + dlogits: model_util.py:61-72, gaze_rnn.py:149-159,381-408.
// One CTA per frame; the frame (<= 4096 positions) lives in shared memory; one pass over HBM:
// read logits + gt, write probs and/or dlogits.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float t = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, t) : v + t;
  }
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < (int)(blockDim.x >> 5); ++i) r = is_max ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

__global__ void __launch_bounds__(256) softmax_xent_kernel(const float* __restrict__ logits, const float* __restrict__ gt,
                                                            float* __restrict__ probs, float* __restrict__ dlogits,
                                                            float* __restrict__ frame_loss, int hw, float inv_bt,
                                                            int loss_type) {
  extern __shared__ float row[];  // hw logits
  __shared__ float red[8];
  const long long base = (long long)blockIdx.x * hw;
  const float* lg = logits + base;
  const float* g = gt + base;
  if (loss_type == 1) {  // l2: gaze_rnn.py:387-389 — no softmax, model output is the raw map
    float s = 0.f;
    for (int i = threadIdx.x; i < hw; i += 256) {
      float d = lg[i] - g[i];
      s = fmaf(0.5f * d, d, s);
      if (probs) probs[base + i] = lg[i];
      if (dlogits) dlogits[base + i] = d * inv_bt;
    }
    s = block_reduce(s, red, false);
    if (threadIdx.x == 0) frame_loss[blockIdx.x] = s * inv_bt;
    return;
  }
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < hw; i += 256) {
    float v = lg[i];
    row[i] = v;
    mx = fmaxf(mx, v);
  }
  mx = block_reduce(mx, red, true);
  float se = 0.f;
  for (int i = threadIdx.x; i < hw; i += 256) se += expf(row[i] - mx);
  se = block_reduce(se, red, false);
  const float lse = logf(se);
  float loss = 0.f;
  for (int i = threadIdx.x; i < hw; i += 256) {
    float gv = g[i];
    float x = row[i] - mx - lse;  // log_softmax (own elements of row[]: no hazard)
    loss = fmaf(-gv, x, loss);
    if (loss_type == 2 && gv > 0.f) loss = fmaf(gv, logf(gv), loss);
    float p = expf(x);
    if (probs) probs[base + i] = p;
    if (dlogits) dlogits[base + i] = (p - gv) * inv_bt;  // TF gradient of softmax_cross_entropy_with_logits
  }
  loss = block_reduce(loss, red, false);
  if (threadIdx.x == 0) frame_loss[blockIdx.x] = loss * inv_bt;
}

__global__ void __launch_bounds__(1024) reduce_sum_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (long long i = threadIdx.x; i < n; i += 1024) s += (double)x[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 32; ++i) t += red[i];
    out[0] = (float)t;
  }
}

// Large n: kRedBlocks CTAs write fixed-order partials; the last CTA to arrive adds them in index order, so the
// result does not depend on scheduling.  The scratch is library-static: calls must not overlap on different
// streams of one process (include/rgp.h: one stream per model instance).
constexpr int kRedBlocks = 296;
constexpr int kRedSlots = 16;            // concurrent streams served without sharing scratch (host side: red_slot())
__device__ double g_red_part_all[kRedSlots][kRedBlocks];
__device__ unsigned int g_red_count_all[kRedSlots] = {0};

__global__ void __launch_bounds__(512) reduce_sum_big_kernel(const float* __restrict__ x, long long n, float* __restrict__ out,
                                                             int slot) {
  double* g_red_part = g_red_part_all[slot];
  unsigned int& g_red_count = g_red_count_all[slot];
  __shared__ double red[16];
  __shared__ bool last;
  const long long per = ((n / 4 + gridDim.x - 1) / gridDim.x) * 4;         // float4-aligned contiguous share per CTA
  const long long b = (long long)blockIdx.x * per;
  const long long e = (b + per < n) ? b + per : n;
  double s = 0.0;
  if (((uintptr_t)x & 15) == 0) {
    const float4* x4 = reinterpret_cast<const float4*>(x + b);
    const long long n4 = (e > b) ? (e - b) / 4 : 0;
    for (long long i = threadIdx.x; i < n4; i += 512) { const float4 v = x4[i]; s += (double)v.x + (double)v.y + (double)v.z + (double)v.w; }
    for (long long i = b + n4 * 4 + threadIdx.x; i < e; i += 512) s += (double)x[i];
  } else {
    for (long long i = b + threadIdx.x; i < e; i += 512) s += (double)x[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 16; ++i) t += red[i];
    g_red_part[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(&g_red_count, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double t = 0.0;
    for (unsigned i = 0; i < gridDim.x; ++i) t += *(volatile double*)&g_red_part[i];
    out[0] = (float)t;
    g_red_count = 0;
  }
}

// ---------------------------------------------------------------------------------------------
// base.py:287-297: clip_by_global_norm + tf.train.AdamOptimizer (epsilon outside the bias correction)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, long long n, double* __restrict__ out) {
  __shared__ double red[8];
  double s = 0.0;
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  long long stride = (long long)gridDim.x * 256;
  for (; i < n; i += stride) { double v = g[i]; s += v * v; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += red[k];
    atomicAdd(out, t);
  }
}

// gaze_rnn.py:439-443 + the python wrapper of tf.train.AdamOptimizer [TF-lib], evaluated ON THE DEVICE so that a captured
// CUDA graph of the training step replays a correct schedule: state[0] = global_step (number of updates applied).
//   lr   = lr0 * decay^floor(step / decay_steps)        (tf.train.exponential_decay, staircase)
//   lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t),  t = step + 1
// sched[0] = lr_t (consumed by clip_adam_kernel), sched[1] = lr; then global_step <- t.  One thread, double arithmetic.
__global__ void adam_schedule_kernel(long long* __restrict__ step, const double* __restrict__ lr0, double decay,
                                     int decay_steps, double beta1, double beta2, float* __restrict__ sched) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const long long s0 = step[0];
  const double t = (double)(s0 + 1);
  const double lr = lr0[0] * pow(decay, (double)(s0 / (decay_steps > 0 ? decay_steps : 1)));
  const double lr_t = lr * sqrt(1.0 - pow(beta2, t)) / (1.0 - pow(beta1, t));
  sched[0] = (float)lr_t;
  sched[1] = (float)lr;
  step[0] = s0 + 1;
}

template <typename TA>
__global__ void __launch_bounds__(256) clip_adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                         float* __restrict__ m, float* __restrict__ v, long long n,
                                                         const double* __restrict__ sumsq, float clip, float lr_host,
                                                         const float* __restrict__ lr_dev,
                                                         float b1, float b2, float eps, TA* __restrict__ p_op) {
  const float lr_t = lr_dev ? lr_dev[0] : lr_host;
  const float norm = (float)sqrt(sumsq[0]);
  const float scale = clip > 0.f ? clip / fmaxf(norm, clip) : 1.0f;
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  long long stride = (long long)gridDim.x * 256;
  for (; i < n; i += stride) {
    float gi = g[i] * scale;
    float mi = b1 * m[i] + (1.0f - b1) * gi;
    float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    float pi = p[i] - lr_t * mi / (sqrtf(vi) + eps);
    m[i] = mi; v[i] = vi; p[i] = pi;
    if (p_op) OpT<TA>::st(p_op + i, pi);
  }
}

// base.py:270-273 — the reference's other two optimisers, in TF1's update rules [TF-lib]:
//   kind 0  tf.train.MomentumOptimizer(lr, momentum=0.9) (ApplyMomentum, no Nesterov):
//             accum = momentum * accum + g;  theta -= lr * accum
//   kind 1  tf.train.RMSPropOptimizer(lr, decay=0.9, momentum=0.9, epsilon=1e-10) (ApplyRMSProp; the `rms` slot starts at ONE):
//             ms = decay * ms + (1 - decay) * g^2;  mom = momentum * mom + lr * g / sqrt(ms + eps);  theta -= mom
// g is the globally clipped gradient (base.py:287-291); lr = sched[1], the staircase learning rate rgp_adam_schedule leaves
// on the device (the same schedule kernel advances global_step for every optimiser).
template <typename TA>
__global__ void __launch_bounds__(256) clip_momentum_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                             float* __restrict__ acc, float* __restrict__ ms, long long n,
                                                             const double* __restrict__ sumsq, float clip,
                                                             const float* __restrict__ sched, int kind, float momentum,
                                                             float rms_decay, float eps, TA* __restrict__ p_op) {
  const float lr = sched[1];
  const float norm = (float)sqrt(sumsq[0]);
  const float scale = clip > 0.f ? clip / fmaxf(norm, clip) : 1.0f;
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long stride = (long long)gridDim.x * 256;
  for (; i < n; i += stride) {
    const float gi = g[i] * scale;
    float pi, ai;
    if (kind == 0) {
      ai = momentum * acc[i] + gi;
      pi = p[i] - lr * ai;
    } else {
      const float mi = rms_decay * ms[i] + (1.0f - rms_decay) * gi * gi;
      ms[i] = mi;
      ai = momentum * acc[i] + lr * gi * rsqrtf(mi + eps);
      pi = p[i] - ai;
    }
    acc[i] = ai; p[i] = pi;
    if (p_op) OpT<TA>::st(p_op + i, pi);
  }
}

}  // namespace rgp

using namespace rgp;

// gaze_rnn.py:504-510 — the random horizontal flip of half the batch, applied to the fed device buffers instead of the
// host arrays: for every selected clip, each row of W consecutive floats (the last axis of [clip, rows, W]) is
// reversed in place.  One thread per row; a row of 7 (C3D) or 49 (gaze map) floats is read and written by its owner only.
namespace rgp {
namespace {
__global__ void flip_w_kernel(float* __restrict__ x, const int* __restrict__ clips, int nsel, long long rows_per_clip, int W) {
  const long long total = (long long)nsel * rows_per_clip;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / rows_per_clip);
    const long long r = i - (long long)k * rows_per_clip;
    float* row = x + ((long long)clips[k] * rows_per_clip + r) * W;
    for (int a = 0, b = W - 1; a < b; ++a, --b) { const float t = row[a]; row[a] = row[b]; row[b] = t; }
  }
}
}  // namespace
}  // namespace rgp

// ---- host helpers of the entry points below (C++ linkage)
template <typename TA, typename TS>
static int launch_pack_c3d(const void* c3d, void* out, int B, int T, int C, int P, void* stream) {
  dim3 grid(B * T, C / 64);
  size_t sm = (size_t)64 * P * sizeof(float);
  if (sm > 48 * 1024) cudaFuncSetAttribute(pack_c3d_kernel<TA, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  pack_c3d_kernel<TA, TS><<<grid, 256, sm, (cudaStream_t)stream>>>((const TS*)c3d, (TA*)out, B, T, C, P);
  return check_launch("rgp_pack_c3d");
}

// The partial-sum scratch of the large reduction is keyed by (device, stream): two engines on different streams (or
// devices: __device__ arrays are per device already) never share a slot.  More than kRedSlots live streams on one
// device fall back to the single-CTA kernel, which needs no scratch.
static int red_slot(void* stream) {
  static std::mutex mu;
  static void* owner[64][kRedSlots] = {};
  static int used[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  std::lock_guard<std::mutex> g(mu);
  for (int i = 0; i < used[dev]; ++i)
    if (owner[dev][i] == stream) return i;
  if (used[dev] < kRedSlots) { owner[dev][used[dev]] = stream; return used[dev]++; }
  return -1;
}

static int launch_clip_adam(float* p, const float* g, float* m, float* v, long long n, const double* sumsq, float clip,
                            float lr_t, const float* lr_dev, float beta1, float beta2, float eps, void* p_op, int dtype,
                            void* stream) {
  int grid = (int)((n + 1023) / 1024);
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  if (dtype == 0 || p_op == nullptr)
    clip_adam_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, sumsq, clip, lr_t, lr_dev, beta1, beta2,
                                                                     eps, (float*)p_op);
  else
    clip_adam_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, sumsq, clip, lr_t, lr_dev, beta1,
                                                                             beta2, eps, (__nv_bfloat16*)p_op);
  return check_launch("rgp_gradnorm_clip_adam");
}

extern "C" {

int rgp_pack_c3d_src(const void* c3d, int src_dtype, void* out, int B, int T, int C, int P, int dtype, void* stream) {
  RGP_CHECK_ARG(C % 64 == 0, "rgp_pack_c3d: C=%d must be a multiple of 64", C);
  RGP_CHECK_ARG(src_dtype == 0 || src_dtype == 1, "rgp_pack_c3d_src: src_dtype=%d (0 fp32, 1 bf16)", src_dtype);
  RGP_CHECK_ARG((size_t)64 * P * sizeof(float) <= 200 * 1024, "rgp_pack_c3d: P=%d positions per frame do not fit in shared memory", P);
  if (src_dtype == 0)
    return dtype == 0 ? launch_pack_c3d<float, float>(c3d, out, B, T, C, P, stream)
                      : launch_pack_c3d<__nv_bfloat16, float>(c3d, out, B, T, C, P, stream);
  return dtype == 0 ? launch_pack_c3d<float, __nv_bfloat16>(c3d, out, B, T, C, P, stream)
                    : launch_pack_c3d<__nv_bfloat16, __nv_bfloat16>(c3d, out, B, T, C, P, stream);
}

int rgp_pack_c3d(const float* c3d, void* out, int B, int T, int C, int P, int dtype, void* stream) {
  return rgp_pack_c3d_src(c3d, 0, out, B, T, C, P, dtype, stream);
}

int rgp_pack_emb(const float* emb, void* out, int B, int T, int G, int C, int dtype, void* stream) {
  RGP_CHECK_ARG(C % 4 == 0, "rgp_pack_emb: C=%d must be a multiple of 4", C);
  long long total = (long long)B * T * G * G * (C / 4);
  int grid = (int)((total + 255) / 256);
  if (dtype == 0) pack_emb_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(emb, (float*)out, B, T, G, C);
  else pack_emb_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(emb, (__nv_bfloat16*)out, B, T, G, C);
  return check_launch("rgp_pack_emb");
}

int rgp_cast(const float* src, void* dst, long long n, int dtype, void* stream) {
  int grid = (int)((n + 1023) / 1024);
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  if (dtype == 0) cast_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(src, (float*)dst, n);
  else cast_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n);
  return check_launch("rgp_cast");
}

int rgp_fc_head(const float* h_f32, const float* out_W, const float* out_b, float* logits, int B, int T, int t,
                int G, int Ch, void* stream) {
  int warps = B * G * G;
  int grid = (warps * 32 + 255) / 256;
  fc_head_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(h_f32, out_W, out_b, logits, B, T, t, G, Ch);
  return check_launch("rgp_fc_head");
}

int rgp_fc_head_all(const float* h1_f32, const float* out_W, const float* out_b, float* logits, int B, int T, int G,
                    int Ch, void* stream) {
  RGP_CHECK_ARG(Ch > 0 && Ch % 4 == 0, "rgp_fc_head_all: Ch=%d must be a multiple of 4", Ch);
  RGP_CHECK_ARG(B > 0 && T > 0 && G > 0, "rgp_fc_head_all: B=%d T=%d G=%d", B, T, G);
  RGP_CHECK_ARG(((reinterpret_cast<uintptr_t>(h1_f32) | reinterpret_cast<uintptr_t>(out_W)) & 15) == 0,
                "rgp_fc_head_all: h1_f32 and out_W must be 16-byte aligned%s", "");
  const long long warps = (long long)T * B * G * G;
  const long long grid = (warps * 32 + 255) / 256;
  fc_head_all_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(h1_f32, out_W, out_b, logits, B, T, G, Ch);
  return check_launch("rgp_fc_head_all");
}

int rgp_lstm_gates(const float* xg, int ldx, const float* hconv, const float* wci, const float* wcf,
                   const float* wco, float* c_f32, float* h_f32, void* h_op_next, void* hs_op, int B, int G,
                   int Ch, int So, int Wo, int pado, int dtype, float* save, void* stream) {
  long long total = (long long)B * G * G * Ch;
  const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (Ch % 4 == 0 && ldx % 4 == 0 && al16(xg) && al16(hconv) && al16(wci) && al16(wcf) && al16(wco) && al16(c_f32) &&
      al16(h_f32) && al16(h_op_next) && al16(hs_op) && al16(save)) {
    int grid4 = (int)((total / 4 + 255) / 256);
    if (dtype == 0)
      lstm_gates_v4_kernel<float><<<grid4, 256, 0, (cudaStream_t)stream>>>(xg, ldx, hconv, wci, wcf, wco, c_f32, h_f32,
                                                                           (float*)h_op_next, (float*)hs_op, B, G, Ch, So,
                                                                           Wo, pado, save);
    else
      lstm_gates_v4_kernel<__nv_bfloat16><<<grid4, 256, 0, (cudaStream_t)stream>>>(
          xg, ldx, hconv, wci, wcf, wco, c_f32, h_f32, (__nv_bfloat16*)h_op_next, (__nv_bfloat16*)hs_op, B, G, Ch, So,
          Wo, pado, save);
    return check_launch("rgp_lstm_gates");
  }
  int grid = (int)((total + 255) / 256);
  if (dtype == 0)
    lstm_gates_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(xg, ldx, hconv, wci, wcf, wco, c_f32, h_f32,
                                                                      (float*)h_op_next, (float*)hs_op, B, G, Ch, So,
                                                                      Wo, pado, save);
  else
    lstm_gates_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(
        xg, ldx, hconv, wci, wcf, wco, c_f32, h_f32, (__nv_bfloat16*)h_op_next, (__nv_bfloat16*)hs_op, B, G, Ch, So,
        Wo, pado, save);
  return check_launch("rgp_lstm_gates");
}

int rgp_head_compose(const float* w1, const float* w2, const float* w3, const float* out_w, float* w3f, float* w23,
                     float* w123, void* bhead, int Ch, int dtype, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  compose_w3f_kernel<<<(49 * 32 + 255) / 256, 256, 0, s>>>(w3, out_w, w3f);
  compose_w23_kernel<<<(121 * 64 + 255) / 256, 256, 0, s>>>(w2, w3f, w23);
  compose_w123_kernel<<<(361 * Ch + 255) / 256, 256, 0, s>>>(w1, w23, w123, Ch);
  int n = 16 * 48 * Ch;
  if (dtype == 0) compose_bhead_kernel<float><<<(n + 255) / 256, 256, 0, s>>>(w123, (float*)bhead, Ch);
  else compose_bhead_kernel<__nv_bfloat16><<<(n + 255) / 256, 256, 0, s>>>(w123, (__nv_bfloat16*)bhead, Ch);
  return check_launch("rgp_head_compose");
}

// rgp_head_compose_static: the same four kernels with the three intermediates (w3f, w23, w123) in a file-scope device
// array instead of caller buffers - for callers that cannot allocate (rgp.h).  In host code `g_head_scratch` is only the
// host-side stand-in of the symbol: its device address is looked up once per device (that device current) and cached.
constexpr int kHeadStaticMaxCh = 512;
constexpr int kHeadW3fOff = 0;                       // [7,7,32]   1568 floats
constexpr int kHeadW23Off = 1600;                    // [11,11,64] 7744 floats, 256-byte aligned offset
constexpr int kHeadW123Off = 1600 + 7808;            // [19,19,Ch] 361 * Ch floats, 256-byte aligned offset
constexpr int kHeadScratchFloats = kHeadW123Off + 361 * kHeadStaticMaxCh;      // 194 240 floats = 776 960 bytes
static __device__ __align__(256) float g_head_scratch[kHeadScratchFloats];

static float* head_scratch_ptr() {
  static float* ptr[rgp::kMaxDevices] = {nullptr};
  static std::mutex mu;
  const int dev = rgp::current_device();
  std::lock_guard<std::mutex> g(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_head_scratch);
    if (e != cudaSuccess || !p) {
      rgp::set_error("rgp_head_compose_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptr[dev] = static_cast<float*>(p);
  }
  return ptr[dev];
}

int rgp_head_compose_static(const float* w1, const float* w2, const float* w3, const float* out_w, void* bhead, int Ch,
                            int dtype, void* stream) {
  RGP_CHECK_ARG(w1 && w2 && w3 && out_w && bhead, "rgp_head_compose_static: null pointer%s", "");
  RGP_CHECK_ARG(dtype == 0 || dtype == 1, "rgp_head_compose_static: dtype %d", dtype);
  RGP_CHECK_ARG(Ch > 0 && Ch <= kHeadStaticMaxCh,
                "rgp_head_compose_static: Ch=%d is beyond the built-in scratch (at most %d channels = %d bytes); use "
                "rgp_head_compose with caller-owned workspaces", Ch, kHeadStaticMaxCh, kHeadScratchFloats * 4);
  float* sc = head_scratch_ptr();
  if (!sc) return -5;
  // every scratch element that is read below is written first by the kernel before it in stream order: nothing to clear
  return rgp_head_compose(w1, w2, w3, out_w, sc + kHeadW3fOff, sc + kHeadW23Off, sc + kHeadW123Off, bhead, Ch, dtype, stream);
}

int rgp_softmax_xent_fwd_bwd(const float* logits, const float* gt, float* probs, float* dlogits, float* frame_loss,
                             int n_frames, int hw, float inv_bt, int loss_type, void* stream) {
  RGP_CHECK_ARG(hw > 0 && hw <= 12000, "rgp_softmax_xent_fwd_bwd: hw=%d unsupported", hw);
  RGP_CHECK_ARG(loss_type >= 0 && loss_type <= 2, "rgp_softmax_xent_fwd_bwd: unknown loss_type %d", loss_type);
  softmax_xent_kernel<<<n_frames, 256, hw * sizeof(float), (cudaStream_t)stream>>>(logits, gt, probs, dlogits,
                                                                                     frame_loss, hw, inv_bt, loss_type);
  return check_launch("rgp_softmax_xent_fwd_bwd");
}

int rgp_reduce_sum(const float* x, long long n, float* out, void* stream) {
  const int slot = (n > (1 << 16)) ? red_slot(stream) : -1;
  if (slot >= 0) reduce_sum_big_kernel<<<kRedBlocks, 512, 0, (cudaStream_t)stream>>>(x, n, out, slot);
  else reduce_sum_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(x, n, out);
  return check_launch("rgp_reduce_sum");
}

int rgp_sumsq(const float* g, long long n, double* sumsq, void* stream) {
  int grid = (int)((n + 2047) / 2048);
  if (grid > 148 * 4) grid = 148 * 4;
  if (grid < 1) grid = 1;
  sumsq_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g, n, sumsq);
  return check_launch("rgp_sumsq");
}

int rgp_gradnorm_clip_adam(float* p, const float* g, float* m, float* v, long long n, const double* sumsq, float clip,
                           float lr_t, float beta1, float beta2, float eps, void* p_op, int dtype, void* stream) {
  return launch_clip_adam(p, g, m, v, n, sumsq, clip, lr_t, nullptr, beta1, beta2, eps, p_op, dtype, stream);
}

int rgp_gradnorm_clip_adam_dev(float* p, const float* g, float* m, float* v, long long n, const double* sumsq, float clip,
                               const float* lr_t_dev, float beta1, float beta2, float eps, void* p_op, int dtype,
                               void* stream) {
  RGP_CHECK_ARG(lr_t_dev != nullptr, "rgp_gradnorm_clip_adam_dev: lr_t_dev is NULL%s", "");
  return launch_clip_adam(p, g, m, v, n, sumsq, clip, 0.f, lr_t_dev, beta1, beta2, eps, p_op, dtype, stream);
}

int rgp_gradnorm_clip_momentum_dev(float* p, const float* g, float* acc, float* ms, long long n, const double* sumsq,
                                   float clip, const float* sched, int kind, float momentum, float rms_decay, float eps,
                                   void* p_op, int dtype, void* stream) {
  RGP_CHECK_ARG(p && g && acc && sumsq && sched, "rgp_gradnorm_clip_momentum_dev: null pointer%s", "");
  RGP_CHECK_ARG(kind == 0 || kind == 1, "rgp_gradnorm_clip_momentum_dev: kind=%d (0 momentum, 1 rmsprop)", kind);
  RGP_CHECK_ARG(kind == 0 || ms != nullptr, "rgp_gradnorm_clip_momentum_dev: rmsprop needs the mean-square slot%s", "");
  RGP_CHECK_ARG(n >= 0, "rgp_gradnorm_clip_momentum_dev: n=%lld", n);
  if (n == 0) return 0;
  int grid = (int)((n + 1023) / 1024);
  if (grid > 148 * 8) grid = 148 * 8;
  if (dtype == 0 || p_op == nullptr)
    clip_momentum_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(p, g, acc, ms, n, sumsq, clip, sched, kind, momentum,
                                                                         rms_decay, eps, (float*)p_op);
  else
    clip_momentum_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(p, g, acc, ms, n, sumsq, clip, sched, kind,
                                                                                 momentum, rms_decay, eps, (__nv_bfloat16*)p_op);
  return check_launch("rgp_gradnorm_clip_momentum_dev");
}

int rgp_adam_schedule(long long* step, const double* lr0, double decay, int decay_steps, double beta1, double beta2,
                      float* sched, void* stream) {
  RGP_CHECK_ARG(step && lr0 && sched, "rgp_adam_schedule: null pointer%s", "");
  RGP_CHECK_ARG(decay_steps > 0, "rgp_adam_schedule: decay_steps=%d", decay_steps);
  adam_schedule_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(step, lr0, decay, decay_steps, beta1, beta2, sched);
  return check_launch("rgp_adam_schedule");
}

int rgp_flip_w(float* x, const int* clips, int nsel, long long rows_per_clip, int W, void* stream) {
  RGP_CHECK_ARG(nsel >= 0 && rows_per_clip > 0 && W > 0, "rgp_flip_w: nsel=%d rows=%lld W=%d", nsel, rows_per_clip, W);
  if (nsel == 0) return 0;
  const long long total = (long long)nsel * rows_per_clip;
  int grid = (int)((total + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  rgp::flip_w_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, clips, nsel, rows_per_clip, W);
  return check_launch("rgp_flip_w");
}

int rgp_zero(void* p, long long bytes, void* stream) {
  cudaError_t e = cudaMemsetAsync(p, 0, (size_t)bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) { set_error("rgp_zero: %s", cudaGetErrorString(e)); return -5; }
  return 0;
}

}  // extern "C"

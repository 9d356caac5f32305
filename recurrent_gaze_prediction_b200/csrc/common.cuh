// common.cuh — shared device helpers and the fused epilogues of the tap-GEMM engines.
// Reference arithmetic cited per epilogue (paths relative to the reference root).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/rgp.h"

namespace rgp {

void set_error(const char* fmt, ...);
int check_launch(const char* what);

// Per-DEVICE one-time setup (cudaFuncSetAttribute is a per-device attribute; the SM count and the occupancy of a
// cluster shape differ per device): a flag / value slot per device ordinal of the calling thread's current device.
constexpr int kMaxDevices = 64;
inline int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return (dev >= 0 && dev < kMaxDevices) ? dev : 0;
}
struct PerDeviceFlag {
  volatile char done[kMaxDevices];
  // true when this device still needs its setup (idempotent setup: a racing second thread merely repeats it)
  bool need() const { return !done[current_device()]; }
  void set() { done[current_device()] = 1; }
};

#define RGP_CHECK_ARG(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      rgp::set_error(__VA_ARGS__);               \
      return -22;                                \
    }                                            \
  } while (0)

template <typename T> struct OpT;
template <> struct OpT<float> {
  static __device__ __forceinline__ float ld(const float* p) { return *p; }
  static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
};
template <> struct OpT<__nv_bfloat16> {
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

// accurate libm forms: the gates sit in the epilogue, their cost is noise next to the GEMM, and
// the fp32 parity bar (1e-4 relative on the gaze maps) leaves no room for ex2.approx on tiny pre-activations
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float tanhf_(float x) { return tanhf(x); }

// Per-row geometry computed once per output row of a tile.
struct RowInfo {
  int m;
  bool inrange;   // m < M
  bool valid;     // row is a real (y<G, x<G) position of its frame
  long long prow; // p-layout row of this position in the primary geometry (m + Wp + 1)
  long long orow; // row in the secondary output layout (PROJ: p-row; GRU_C: head-input row)
  long long obase;// HEAD: element offset of logits[b,t,0,0]; hy, hx = 6a-3, 6b-3
  int hy, hx;
};

__device__ __forceinline__ RowInfo make_row(const rgp_tapgemm_args& a, int m) {
  RowInfo r;
  r.m = m;
  r.inrange = m < a.M;
  r.valid = false;
  r.prow = 0; r.orow = 0; r.obase = 0; r.hy = 0; r.hx = 0;
  if (!r.inrange) return r;
  switch (a.epi) {
    case RGP_EPI_PROJ: {
      int f = m / a.P, p = m - f * a.P;
      int y = p / a.G, x = p - y * a.G;
      r.valid = true;
      r.orow = (long long)f * a.So + (y + a.pado) * a.Wo + (x + a.pado);
      break;
    }
    case RGP_EPI_HEAD: {
      int f = m / a.S, q = m - f * a.S;
      int ya = q / a.Wp, xb = q - ya * a.Wp;
      r.valid = (ya < 9) && (xb < 9);
      int t = f / a.B, b = f - t * a.B;   // frames are time-major
      r.obase = ((long long)b * a.T + t) * 2401;
      r.hy = 6 * ya - 3; r.hx = 6 * xb - 3;
      break;
    }
    case RGP_EPI_STORE_F32:
    case RGP_EPI_STORE_OP:
      r.valid = true;
      if (a.im2col) {        // dense-M conv: row m = f*G*G + y*G + x is stored at the q-row f*S + y*Wp + x
        const int P = a.G * a.G;
        const int f = m / P, p = m - f * P;
        const int y = p / a.G, x = p - y * a.G;
        r.m = f * a.S + y * a.Wp + x;
      }
      break;
    case RGP_EPI_UP7X: {  // position-major only: dense row m = f*49 + y*7 + x -> block origin (7y, 7x) of frame f's So rows
      const int P = a.G * a.G;
      const int f = m / P, p = m - f * P;
      const int y = p / a.G, x = p - y * a.G;
      r.valid = true;
      r.obase = (long long)f * a.So;
      r.hy = 7 * y; r.hx = 7 * x;
      break;
    }
    case RGP_EPI_UP7: {   // rows = the 8x8 p-layout blocks (a,b) of the bottom state; hy, hx = 7a, 7b (operand coordinates)
      int f = m / a.S, q = m - f * a.S;
      int ya = q / a.Wp, xb = q - ya * a.Wp;
      r.valid = true;
      r.obase = (long long)f * a.So;
      r.hy = 7 * ya; r.hx = 7 * xb;
      break;
    }
    case RGP_EPI_DHS: {   // M rows in the head-input layout (S=100, Wp=10, data at +3,+3) -> recurrence q-row
      int f = m / a.S, q = m - f * a.S;
      int ya = q / a.Wp, xb = q - ya * a.Wp;
      int i = ya - a.pado, j = xb - a.pado;
      r.valid = (i >= 0) && (i < a.G) && (j >= 0) && (j < a.G);
      r.orow = (long long)f * a.So + i * a.Wo + j;
      break;
    }
    default: {
      if (a.im2col && a.epi == RGP_EPI_DX_DENSE) {   // dense-M transposed conv: row m IS the dense row f*P + y*G + x
        r.valid = true;
        r.orow = m;
        break;
      }
      int f = m / a.S, q = m - f * a.S;
      int y = q / a.Wp, x = q - y * a.Wp;
      r.valid = (y < a.G) && (x < a.G);
      r.prow = (long long)m + a.Wp + 1;
      if (a.epi == RGP_EPI_GRU_C && a.out_op2)
        r.orow = (long long)f * a.So + (y + a.pado) * a.Wo + (x + a.pado);
      if (a.epi == RGP_EPI_DX_DENSE) r.orow = (long long)f * a.P + y * a.G + x;
      break;
    }
  }
  return r;
}

// Gate backward of one GRU step at one (position m, channel n), given dh = dL/dh_t (gaze_grcn.py:118-127):
//   h_t = z h_{t-1} + (1-z) c  =>  dz = dh (h_{t-1} - c), dc = dh (1-z), dh_{t-1} += dh z
//   dpre_z = dz z (1-z) ; dpre_c = dc (1-c^2)      (dpre_r needs d(r.h), i.e. the transposed conv: BWD_RH)
template <typename TA>
__device__ __forceinline__ void gru_gate_bwd(const float* __restrict__ save_z, const float* __restrict__ save_c,
                                             const float* __restrict__ hprev, TA* __restrict__ dg, int ldg,
                                             float* __restrict__ dhpart, int Ch, long long m, long long prow, int n,
                                             float dh) {
  float z = save_z[m * Ch + n], c = save_c[m * Ch + n], hp = hprev[prow * Ch + n];
  float dz = dh * (hp - c), dc = dh * (1.0f - z);
  OpT<TA>::st(dg + m * ldg + n, dz * z * (1.0f - z));
  OpT<TA>::st(dg + m * ldg + 2 * Ch + n, dc * (1.0f - c * c));
  dhpart[m * Ch + n] = dh * z;
}

// One output element. TA = operand dtype.
template <typename TA>
__device__ __forceinline__ void epi_elem(const rgp_tapgemm_args& a, const RowInfo& r, int n, float acc) {
  if (!r.inrange || n >= a.N) return;
  switch (a.epi) {
    case RGP_EPI_STORE_F32: {
      float* o = a.out_f32 + (long long)r.m * a.ldo + n;
      *o = a.accumulate ? (*o + acc) : acc;
      break;
    }
    case RGP_EPI_STORE_OP: {
      OpT<TA>::st((TA*)a.out_op + (long long)r.m * a.ldo + n, acc);
      break;
    }
    case RGP_EPI_STORE_OP_P: {
      if (r.valid) OpT<TA>::st((TA*)a.out_op + r.prow * a.ldo + n, acc);
      break;
    }
    case RGP_EPI_PROJ: {  // gaze_grcn.py:239-242 xw_plus_b, written straight into the conv operand layout
      OpT<TA>::st((TA*)a.out_op + r.orow * a.ldo + n, acc + a.bias[n]);
      break;
    }
    case RGP_EPI_GRU_ZR: {  // gaze_grcn.py:108-119
      float pre = acc + a.xg[(long long)r.m * a.ldx + n];
      float s = sigmoidf_(pre);
      if (n < a.Ch) {
        a.z_f32[(long long)r.m * a.Ch + n] = s;
      } else {
        int j = n - a.Ch;
        if (a.save_r) a.save_r[(long long)r.m * a.Ch + j] = s;
        if (r.valid) {
          float hp = a.hprev_f32[r.prow * a.Ch + j];
          OpT<TA>::st((TA*)a.out_op + r.prow * a.Ch + j, s * hp);   // r (.) h, the operand of U
        }
      }
      break;
    }
    case RGP_EPI_GRU_C: {  // gaze_grcn.py:122-127 ; BN-scale gaze_grcn.py:325
      float pre = acc + a.xg[(long long)r.m * a.ldx + 2 * a.Ch + n];
      float c = tanhf_(pre);
      if (a.save_c) a.save_c[(long long)r.m * a.Ch + n] = c;
      if (r.valid) {
        float z = a.z_f32[(long long)r.m * a.Ch + n];
        float hp = a.hprev_f32[r.prow * a.Ch + n];
        float hn = z * hp + (1.0f - z) * c;
        a.hnext_f32[r.prow * a.Ch + n] = hn;
        OpT<TA>::st((TA*)a.out_op + r.prow * a.Ch + n, hn);
        if (a.out_op2) {
          float v = hn;
          if (a.bias) v = a.bias[n] * (hn * a.alpha) + a.bias2[n];
          OpT<TA>::st((TA*)a.out_op2 + r.orow * a.Ch + n, v);
        }
      }
      break;
    }
    case RGP_EPI_BWD_RH: {  // acc = dL/d(r.h) at this position
      if (r.valid) {
        long long m = r.m;
        float rr = a.save_r[m * a.Ch + n], hp = a.hprev_f32[r.prow * a.Ch + n];
        OpT<TA>::st((TA*)a.dg + m * a.ldg + a.Ch + n, acc * hp * rr * (1.0f - rr));
        a.dhpart[m * a.Ch + n] += acc * rr;
      }
      break;
    }
    case RGP_EPI_BWD_H: {  // acc = convT([dpre_z,dpre_r]) of step t -> full dL/dh_{t-1}, then gates of step t-1
      if (r.valid) {
        long long m = r.m;
        float hd = a.dhs[m * a.Ch + n];
        if (a.bias) hd *= a.bias[n] * a.alpha;
        float dh = acc + a.dhpart[m * a.Ch + n] + hd;
        gru_gate_bwd<TA>(a.save_z, a.save_c, a.hprev_f32, (TA*)a.dg, a.ldg, a.dhpart, a.Ch, m, r.prow, n, dh);
      }
      break;
    }
    case RGP_EPI_DHS: {
      if (r.valid) a.out_f32[r.orow * a.ldo + n] = acc;
      break;
    }
    case RGP_EPI_DX_DENSE: {
      if (r.valid) OpT<TA>::st((TA*)a.out_op + r.orow * a.ldo + n, acc);
      break;
    }
    case RGP_EPI_UP7: {   // gaze_grcn_cascade.py:326-331: column n = (ry*7+rx)*P + co -> pixel (7a+ry-2, 7b+rx-2)
      const int pix = n / a.P;
      if (pix < 49) {
        const int co = n - pix * a.P;
        const int ry = pix / 7, rx = pix - ry * 7;
        const int Y = r.hy + ry, X = r.hx + rx;             // operand coordinates = pixel + 2
        if (Y >= 2 && Y < 51 && X >= 2 && X < 51)
          OpT<TA>::st((TA*)a.out_op + (r.obase + (long long)Y * a.Wo + X) * a.ldo + co, acc);
      }
      break;
    }
    case RGP_EPI_HEAD: {  // composed head: class n = ry*6+rx -> logits[b,t,6a+ry-3,6b+rx-3]
      if (r.valid && n < 36) {
        int ry = n / 6, rx = n - ry * 6;
        int Y = r.hy + ry, X = r.hx + rx;
        if (Y >= 0 && Y < 49 && X >= 0 && X < 49) a.out_f32[r.obase + Y * 49 + X] = acc + a.bias[0];
      }
      break;
    }
  }
}

int tapgemm_simt(const rgp_tapgemm_args& a, cudaStream_t s);
int wgrad_simt(const rgp_wgrad_args& a, cudaStream_t s);
int wgrad_tc(const rgp_wgrad_args& a, cudaStream_t s);
int tapgemm_tc(const rgp_tapgemm_args& a, cudaStream_t s);
int convgru_seq_fwd(const rgp_gru_seq_args& a, cudaStream_t s);
int convgru_seq_profile(unsigned long long* host_out, int n);

}  // namespace rgp

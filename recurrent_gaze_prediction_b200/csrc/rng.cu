// rng.cu — counter-based device random numbers for the OPT-IN fast path of the saliency evaluation.
//
// The reference draws AUC-Judd's tie-breaking jitter (evaluation_metrics.py:78-79, rand(H,W)*1e-7 per frame) and the
// AUC-Borji negative pixels (:148, randint(0, n_pix, [n_fix, n_rep])) from NumPy's global Mersenne-Twister stream; the
// default path of evaluate_batch reproduces that stream on the host and ships the arrays (1.4 GB of doubles per 1024 frames
// at 360x480 - 99 % of the call's wall time).  With rng="device" the same quantities are generated here instead:
// statistically equivalent, not stream-identical to NumPy.  splitmix64 of (seed, element index); 53-bit doubles in [0,1).
#include "common.cuh"

namespace rgp {
namespace {

__device__ __forceinline__ unsigned long long mix64(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256) uniform_f64_kernel(double* __restrict__ out, long long n, unsigned long long seed) {
  // two consecutive doubles per thread: 16-byte stores
  const long long pairs = (n + 1) >> 1;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < pairs; i += (long long)gridDim.x * 256) {
    const double a = (double)(mix64(seed, 2 * i) >> 11) * (1.0 / 9007199254740992.0);
    const double b = (double)(mix64(seed, 2 * i + 1) >> 11) * (1.0 / 9007199254740992.0);
    if (2 * i + 1 < n) *reinterpret_cast<double2*>(out + 2 * i) = make_double2(a, b);
    else out[2 * i] = a;
  }
}

// idx[f, r, c] for r < nfix[f] : uniform pixel in [0, npix) ; rows r >= nfix[f] : -1 (absent, as _pad_index writes them)
__global__ void __launch_bounds__(256) randint_rows_kernel(int* __restrict__ out, const int* __restrict__ nfix, int nframes,
                                                          int nrows, int nrep, int npix, unsigned long long seed) {
  const long long total = (long long)nframes * nrows * nrep;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long fr = i / nrep;
    const int f = (int)(fr / nrows), r = (int)(fr - (long long)f * nrows);
    // 64-bit multiply-high maps the hash to [0, npix) without modulo bias beyond 2^-64
    out[i] = r < nfix[f] ? (int)__umul64hi(mix64(seed, (unsigned long long)i), (unsigned long long)npix) : -1;
  }
}

int grid_for(long long work) {
  long long g = (work + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  return g < 1 ? 1 : (int)g;
}

}  // namespace
}  // namespace rgp

using namespace rgp;

extern "C" {

int rgp_fill_uniform_f64(double* out, long long n, unsigned long long seed, void* stream) {
  RGP_CHECK_ARG(out != nullptr && n >= 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
                "rgp_fill_uniform_f64: out must be 16-byte aligned, n=%lld", n);
  if (n == 0) return 0;
  uniform_f64_kernel<<<grid_for((n + 1) / 2), 256, 0, (cudaStream_t)stream>>>(out, n, seed);
  return check_launch("rgp_fill_uniform_f64");
}

int rgp_fill_randint_rows(int* out, const int* nfix, int nframes, int nrows, int nrep, int npix, unsigned long long seed,
                          void* stream) {
  RGP_CHECK_ARG(out && nfix && nframes > 0 && nrows > 0 && nrep > 0 && npix > 0,
                "rgp_fill_randint_rows: nframes=%d nrows=%d nrep=%d npix=%d", nframes, nrows, nrep, npix);
  randint_rows_kernel<<<grid_for((long long)nframes * nrows * nrep), 256, 0, (cudaStream_t)stream>>>(out, nfix, nframes, nrows,
                                                                                                     nrep, npix, seed);
  return check_launch("rgp_fill_randint_rows");
}

}  // extern "C"

// api.cu — C-ABI plumbing: error slot, version, engine dispatch of rgp_tapgemm.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace rgp {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return -5;
  }
  return 0;
}
}  // namespace rgp

extern "C" {

const char* rgp_last_error(void) { return rgp::g_err; }
int rgp_version(void) { return 100; }

int rgp_has_tcgen05(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaDeviceProp p;
  if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 0;
  return p.major == 10 ? 1 : 0;
}

int rgp_tapgemm(const rgp_tapgemm_args* a, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_tapgemm: null args");
  RGP_CHECK_ARG(a->ntaps >= 1 && a->ntaps <= RGP_MAX_TAPS, "rgp_tapgemm: ntaps=%d out of range", a->ntaps);
  RGP_CHECK_ARG(a->M > 0 && a->N > 0 && a->K > 0, "rgp_tapgemm: bad shape M=%d N=%d K=%d", a->M, a->N, a->K);
  RGP_CHECK_ARG(a->dtype == 0 || a->dtype == 1, "rgp_tapgemm: dtype %d", a->dtype);
  RGP_CHECK_ARG(a->epi >= 0 && a->epi <= RGP_EPI_UP7X, "rgp_tapgemm: unknown epilogue %d", a->epi);
  RGP_CHECK_ARG(!a->im2col || (a->dtype == 1 && a->engine != 1), "rgp_tapgemm: im2col mode needs the tcgen05 engine (dtype %d)", a->dtype);
  RGP_CHECK_ARG(a->epi != RGP_EPI_UP7X || a->im2col == 3, "rgp_tapgemm: RGP_EPI_UP7X needs im2col = 3 (tcgen05 engine), got %d", a->im2col);
  if (a->dtype == 1 && a->engine != 1) return rgp::tapgemm_tc(*a, (cudaStream_t)stream);
  return rgp::tapgemm_simt(*a, (cudaStream_t)stream);
}

int rgp_convgru_seq_fwd(const rgp_gru_seq_args* a, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_convgru_seq_fwd: null args");
  return rgp::convgru_seq_fwd(*a, (cudaStream_t)stream);
}

int rgp_convgru_seq_profile(unsigned long long* host_out, int n) {
  RGP_CHECK_ARG(host_out != nullptr && n > 0, "rgp_convgru_seq_profile: bad arguments");
  return rgp::convgru_seq_profile(host_out, n);
}

int rgp_wgrad(const rgp_wgrad_args* a, void* stream) {
  RGP_CHECK_ARG(a != nullptr, "rgp_wgrad: null args");
  RGP_CHECK_ARG(a->ntaps >= 1 && a->ntaps <= RGP_MAX_TAPS, "rgp_wgrad: ntaps=%d out of range", a->ntaps);
  RGP_CHECK_ARG(a->M > 0 && a->Ny > 0 && a->Nx > 0, "rgp_wgrad: bad shape M=%d Ny=%d Nx=%d", a->M, a->Ny, a->Nx);
  if (a->dtype == 1 && a->engine != 1) return rgp::wgrad_tc(*a, (cudaStream_t)stream);
  return rgp::wgrad_simt(*a, (cudaStream_t)stream);
}

}  // extern "C"

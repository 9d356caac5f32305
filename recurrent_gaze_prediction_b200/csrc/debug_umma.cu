// debug_umma.cu — hardware probe behind the "resident A slab" tap-GEMM: can a K-major
// SWIZZLE_128B UMMA operand start at a row that is NOT a multiple of the 8-row swizzle atom?
//
// A slab of `rows` x 64 bf16 is TMA-loaded once (SWIZZLE_128B, 1024-byte aligned); D[128, N] = A[off : off+128, :] . B^T
// is then issued with the A descriptor's start address = slab + off*128 (+ k*32 per K16 step) and the descriptor's
// base-offset field (bits 49-51) = mode ? ((start >> 7) & 7) : 0.  The host compares D with a CPU product
// (tools/umma_rowoff_probe.py).  Not part of the product path; not declared in include/rgp.h.
// Not part of the product library: compiled only when the build defines RGP_BUILD_PROBES
// (RGP_BUILD_PROBES=1 python -c "import __graft_entry__ as g; g.build(force=True)").
#ifdef RGP_BUILD_PROBES
#include <cuda_bf16.h>
#include "tc_common.cuh"

namespace rgp {
namespace {
using namespace tc;

constexpr int kN = 128;

__global__ void __launch_bounds__(128, 1)
umma_rowoff_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* __restrict__ D,
                   int rows, int off, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sa = smem;                                   // rows x 128 B
  uint8_t* sb = smem + 256 * 128;                       // kN x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(sb + kN * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bars[0], rows * 128 + kN * 128);
    tma_load_2d(sa, &tmA, &bars[0], 0, 0);
    tma_load_2d(sb, &tmB, &bars[0], 0, 0);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int k = 0; k < 4; ++k) {
      const uint32_t a_addr = smem_u32(sa) + off * 128 + k * 32;
      uint64_t ad = make_sdesc(a_addr);
      if (mode) ad |= (uint64_t)((a_addr >> 7) & 7) << 49;
      const uint64_t bd = make_sdesc(smem_u32(sb) + k * 32);
      tc_mma_bf16(tmem, ad, bd, idesc, k > 0);
    }
    tc_commit(&bars[1]);
  }
  __syncthreads();
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  for (int c = 0; c < kN; c += 32) {
    float v[32];
    tc_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    for (int i = 0; i < 32; ++i) D[(warp * 32 + lane) * kN + c + i] = v[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
}

}  // namespace
}  // namespace rgp

extern "C" int rgp_debug_umma_rowoff(const void* A, int rows, const void* B, float* D, int off, int mode, void* stream) {
  using namespace rgp;
  RGP_CHECK_ARG(rows >= 128 && rows <= 256 && rows % 8 == 0 && off >= 0 && off + 128 <= rows, "rgp_debug_umma_rowoff: rows=%d off=%d", rows, off);
  CUtensorMap tmA, tmB;
  int rc = tc::encode_map_bf16(A, 64, rows, 128, 64, rows, &tmA);
  if (rc) return rc;
  rc = tc::encode_map_bf16(B, 64, kN, 128, 64, kN, &tmB);
  if (rc) return rc;
  const int smem = 256 * 128 + kN * 128 + 1024 + 64;
  cudaFuncSetAttribute(umma_rowoff_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  umma_rowoff_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tmA, tmB, D, rows, off, mode);
  return check_launch("rgp_debug_umma_rowoff");
}

#endif  // RGP_BUILD_PROBES

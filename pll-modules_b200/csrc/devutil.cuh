/*
 * devutil.cuh — small device helpers shared by the kernels: 256-bit global accesses
 * (LDG.E.ENL2.256 / STG.E.ENL2.256 on sm_100a: one request per (site, category) element of a
 * 4-state CLV) and the 2^+-256 scaling constants (pll.h:96-100).
 */
#ifndef PLLB200_DEVUTIL_CUH_
#define PLLB200_DEVUTIL_CUH_

#include <cuda_runtime.h>

struct __align__(32) d4
{
  double x, y, z, w;
};
struct __align__(16) d2
{
  double x, y;
};

__device__ __forceinline__ d4 ldg256(const double * p)
{
  d4 r;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg256(double * p, const d4 & r)
{
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(r.x), "d"(r.y), "d"(r.z),
               "d"(r.w)
               : "memory");
}

/* ---- mbarrier / TMA bulk copy (PTX ISA: cp.async.bulk, mbarrier) ------------------------------ */
__device__ __forceinline__ unsigned int smem_u32(const void * p)
{
  return (unsigned int)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long * bar, unsigned int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long * bar, unsigned int bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long * bar, unsigned int parity)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_1d(void * dst_smem, const void * src_gmem, unsigned int bytes,
                                            unsigned long long * bar)
{
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

#define SCALE_FACTOR_D 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define SCALE_THRESHOLD_D (1.0 / SCALE_FACTOR_D)
#define LN_SCALE_THRESHOLD (-177.445678223345993274) /* ln(2^-256) */

#endif

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

#define SCALE_FACTOR_D 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define SCALE_THRESHOLD_D (1.0 / SCALE_FACTOR_D)
#define LN_SCALE_THRESHOLD (-177.445678223345993274) /* ln(2^-256) */

#endif

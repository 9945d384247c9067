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

/* same store with the streaming hint: the line is not expected to be read again (a parent CLV whose only reader takes
 * it from registers), so L2 may write it back first */
__device__ __forceinline__ void stg256_cs(double * p, const d4 & r)
{
  asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(r.x), "d"(r.y), "d"(r.z),
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

/* ---- 20-state DMMA operand order (k_traversal_aa.cu, k_pmatrix.cu) -------------------------------
 * X[site][i] = sum_j C[site][j] P[i][j] runs as m8n8k4 tiles: M = 8 sites, N = 8 parent states i
 * (3 n-tiles), K = 4 child states j (5 k-steps).  Which state sits in which k-slot and which n-column
 * is free as long as the A fragment (child CLV), the B fragment (P^T image) and the reader of the D
 * fragment agree.  The order is chosen so that fragment lane column c (= lane & 3) owns the SAME states
 * on both sides, and owns them contiguously in memory:
 *   k-slot  (ks, c):  state 4c + ks for ks < 4,           state 16 + c for ks = 4;
 *   D pair  (nt, c):  states 4c + 2nt, 4c + 2nt + 1 for nt < 2,   16 + 2c, 16 + 2c + 1 for nt = 2 (c < 2).
 * A parent written by one operation is then the A operand of the next one without leaving the registers
 * (one shuffle for k-slot 4), a CLV row is read with one 32-byte and one 8-byte load per lane and category,
 * and written with one 32-byte store (+ one 16-byte store from half the lanes). */
__host__ __device__ __forceinline__ unsigned int aa_state_of_slot(unsigned int ks, unsigned int c)
{
  return ks < 4 ? 4 * c + ks : 16 + c;
}
__host__ __device__ __forceinline__ unsigned int aa_kstep_of_state(unsigned int j) { return j < 16 ? (j & 3u) : 4; }
__host__ __device__ __forceinline__ unsigned int aa_col_of_state(unsigned int j) { return j < 16 ? j >> 2 : j - 16; }
/* first state of the D pair (nt, c) */
__host__ __device__ __forceinline__ unsigned int aa_dstate(unsigned int nt, unsigned int c)
{
  return nt < 2 ? 4 * c + 2 * nt : 16 + 2 * c;
}
/* n-tile and column inside the tile (0..7) of parent state i */
__host__ __device__ __forceinline__ unsigned int aa_ntile_of_state(unsigned int i) { return i < 16 ? (i & 3u) >> 1 : 2; }
__host__ __device__ __forceinline__ unsigned int aa_ncol_of_state(unsigned int i)
{
  return i < 16 ? 2 * (i >> 2) + (i & 1u) : i - 16;
}

#define SCALE_FACTOR_D 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define SCALE_THRESHOLD_D (1.0 / SCALE_FACTOR_D)
#define LN_SCALE_THRESHOLD (-177.445678223345993274) /* ln(2^-256) */

#endif

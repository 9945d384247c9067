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

/* ---- 20-state DMMA operand order (k_traversal_aa.cu, k_pmatrix.cu) -------------------------------
 * The contraction index j of X[site][i] = sum_j C[site][j] P[i][j] may be fed to the five m8n8k4
 * k-steps in any order as long as the A fragment (child CLV) and the B fragment (P^T image) agree.
 * The order is chosen so that lane column c of k-step ks holds the state the SAME lane holds in the
 * D fragment of the previous update: D of n-tile nt = states nt*8 + 2c, 2c+1, hence
 *   (ks, c) <-> state (ks>>1)*8 + 2c + (ks&1)   for ks < 4,      (4, c) <-> state 16 + c.
 * A parent written by one operation is then the A operand of the next one without leaving the
 * registers (one shuffle for k-step 4), and a CLV read from memory is two 16-byte loads and one
 * 8-byte load per category instead of five 8-byte loads. */
__host__ __device__ __forceinline__ unsigned int aa_state_of_slot(unsigned int ks, unsigned int c)
{
  return ks < 4 ? (ks >> 1) * 8 + 2 * c + (ks & 1u) : 16 + c;
}
__host__ __device__ __forceinline__ unsigned int aa_kstep_of_state(unsigned int j)
{
  return j < 16 ? (j >> 3) * 2 + (j & 1u) : 4;
}
__host__ __device__ __forceinline__ unsigned int aa_col_of_state(unsigned int j)
{
  return j < 16 ? (j & 7u) >> 1 : j - 16;
}

#define SCALE_FACTOR_D 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define SCALE_THRESHOLD_D (1.0 / SCALE_FACTOR_D)
#define LN_SCALE_THRESHOLD (-177.445678223345993274) /* ln(2^-256) */

#endif

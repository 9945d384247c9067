/*
 * k_traversal_aa.cu — 20-state (protein) site-column traversal kernel on the FP64 tensor cores.
 *
 * Replaces pll_core_update_partial_{ii,ti,tt} for states == 20 (libs/libpll/src/core_partials.c:82-200,
 * 354-610,612-764; AVX2 20x20 kernels core_partials_avx2.c:49-1009).  For one rate category the child
 * contraction X[site][i] = sum_j V[site][j] * P[i][j] is a dense [sites x 20] x [20 x 20] product, so it
 * runs as DMMA tiles (mma.sync.m8n8k4.f64): M = 8 sites, N = 8 parent states (3 tiles, 20 -> 24), K = 4
 * child states (5 steps) -> 15 DMMA per side, tile and category.  The B operand (P^T) is pre-arranged by
 * pmatrix_kernel as a fragment image [k][ntile][kstep][lane] so a lane fetches its fragment with one
 * conflict-free LDS.64; both matrices of the NEXT operation (2 x K x 3840 B) arrive by TMA bulk copy into
 * a double-buffered shared-memory stage while the current operation computes.
 *
 * Like the 4-state kernel, ONE launch walks the whole operation list for a chunk of sites (a CTA of 4
 * warps owns 4 x 8T consecutive sites), so children written a few ops earlier come back from L1/L2.
 * Here the lanes of a warp exchange data through global memory (D-fragment layout != A-fragment layout),
 * which the per-operation __syncthreads orders.
 *
 * Tip children (PLL_ATTRIB_PATTERN_TIP): the A fragment is bit (j) of tipmap[code] as 0.0/1.0.
 * Scaling (per-site mode): max high word over all K*20 entries of a site < 2^-256 -> the site is
 * rescaled in a (rare) second pass and its scaler incremented (core_partials.c:729-763).
 * Per-rate scalers with 20 states use the generic kernel (k_partials.cu).
 * Algorithmic bytes per site (Gamma4): ii 1932, ti 1289, tt 646; 6480 flop -> 3.35 flop/B.
 */
#include "internal.h"
#include "devutil.cuh"

#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include <vector>

struct __align__(16) AaOp
{
  unsigned long long p_off, c1_off, c2_off; /* doubles into CLV arena / bytes into tipchars */
  unsigned long long ps_off, s1_off, s2_off;
  unsigned int m1_off, m2_off;              /* doubles into the fragment image */
  unsigned int flags, pad;
};
#define AOPF_TIP1 1u
#define AOPF_TIP2 2u
#define AOPF_PS 4u
#define AOPF_S1 8u
#define AOPF_S2 16u

struct AaArgs
{
  const AaOp * ops;
  unsigned int nops, sites, K;
  double * clv;
  double * pbase; /* where parents are written: the CLV arena, or a sumtable buffer (launch_sumtable_aa) */
  unsigned int * scaler;
  const double * frag;
  const double * tiptab; /* per matrix [code 24][k][20] = P x tip vector (k_pmatrix.cu), or null */
  const unsigned char * tipchars;
  const unsigned long long * tipmap;
  int l1_prefetch; /* Gamma4 kernel: pull the next operation's memory children into L1 while this one computes */
};

/* defaults: 4 warps per CTA, 4 tiles of 8 sites per warp and chunk (template parameters below) */
#define AA_SEG 128 /* ops staged per CTA */
#define AA_FRAG 480 /* doubles per (matrix, category): 3 ntiles x 5 ksteps x 32 lanes */
/* signed compare on the high word: x < 2^-256 for every non-NaN x, negative round-off values included,
 * as the reference's `x < PLL_SCALE_THRESHOLD` (core_partials.c:729-763) */
#define AA_THR_HI 0x2FF00000
#define AA_HI_MIN (-2147483647 - 1)
struct PmatIdx
{
  unsigned int v[8];
};

__device__ __forceinline__ unsigned int aa_smem_u32(const void * p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void aa_mbar_init(unsigned long long * b, unsigned int n)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(aa_smem_u32(b)), "r"(n) : "memory");
}
__device__ __forceinline__ void aa_mbar_expect(unsigned long long * b, unsigned int bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(aa_smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void aa_mbar_wait(unsigned long long * b, unsigned int parity)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(
          aa_smem_u32(b)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void aa_tma(void * dst, const void * src, unsigned int bytes, unsigned long long * b)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   aa_smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(aa_smem_u32(b))
               : "memory");
}
__device__ __forceinline__ void dmma(double & c0, double & c1, double a, double b)
{
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

/* A fragment of one (site, category) from memory in the slot order of devutil.cuh: states 4c .. 4c+3 (k-slots
 * 0-3, one 32-byte load) and 16+c (k-slot 4) */
__device__ __forceinline__ void aa_load_slots(const double * p, int col, bool act, double (&A)[5])
{
  if (act)
  {
    const d4 v = ldg256(p + 4 * col);
    A[0] = v.x;
    A[1] = v.y;
    A[2] = v.z;
    A[3] = v.w;
    A[4] = p[16 + col];
  }
  else
    A[0] = A[1] = A[2] = A[3] = A[4] = 0.0;
}

/* dynamic shared memory: AaOp ops[AA_SEG] | u64 tipmap[256] | u64 bar[2] | u32 cnt[2] | double mat[2 stages][2][K*480] */
template <int AA_T, int AA_WARPS, int MINB>
__device__ __forceinline__ void trav_aa_body(const AaArgs & a)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  AaOp * s_ops = reinterpret_cast<AaOp *>(smem_raw);
  unsigned long long * s_tipmap = reinterpret_cast<unsigned long long *>(smem_raw + sizeof(AaOp) * AA_SEG);
  unsigned long long * s_bar = s_tipmap + 256;
  unsigned int * s_cnt = reinterpret_cast<unsigned int *>(s_bar + 2); /* warps done with stage 0 / 1 */
  double * s_mat = reinterpret_cast<double *>(s_bar + 4); /* keeps the TMA destination 16-byte aligned */

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = lane >> 2, col = lane & 3; /* fragment coordinates */
  const unsigned int K = a.K;
  const unsigned int mat_doubles = K * AA_FRAG;
  const unsigned int mat_bytes = mat_doubles * sizeof(double);
  const unsigned long long span = (unsigned long long)K * 20; /* doubles per site */
  const unsigned long long sites_per_cta = (unsigned long long)AA_WARPS * 8 * AA_T;
  const unsigned long long nchunks = (a.sites + sites_per_cta - 1) / sites_per_cta;
  if (blockIdx.x >= nchunks) return; /* batched launch: the grid is sized for the largest partition */

  for (int i = threadIdx.x; i < 256; i += blockDim.x) s_tipmap[i] = a.tipmap ? a.tipmap[i] : 0ull;
  if (threadIdx.x == 0)
  {
    aa_mbar_init(&s_bar[0], 1);
    aa_mbar_init(&s_bar[1], 1);
    s_cnt[0] = s_cnt[1] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  unsigned int ph = 0;

  for (unsigned int seg = 0; seg < a.nops; seg += AA_SEG)
  {
    const unsigned int nseg = min(a.nops - seg, (unsigned int)AA_SEG);
    __syncthreads();
    {
      const uint4 * src = reinterpret_cast<const uint4 *>(a.ops + seg);
      uint4 * dst = reinterpret_cast<uint4 *>(s_ops);
      for (unsigned int i = threadIdx.x; i < nseg * 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    for (unsigned long long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x)
    {
      const unsigned long long wsite0 = chunk * sites_per_cta + (unsigned long long)warp * 8 * AA_T;
      __syncthreads(); /* both stages free, counters zero */
      /* a launch of at most two operations (the single CLV updates and the sumtable of branch-length
       * optimisation) keeps its matrices in the stages: loaded for the CTA's first chunk only */
      const bool fresh = a.nops > 2 || chunk == blockIdx.x;
      if (threadIdx.x == 0 && fresh)
        for (unsigned int o = 0; o < 2 && o < nseg; ++o)
        {
          double * dstm = s_mat + (size_t)o * 2 * mat_doubles;
          aa_mbar_expect(&s_bar[o], 2 * mat_bytes);
          aa_tma(dstm, ((s_ops[o].flags & AOPF_TIP1) && a.tiptab ? a.tiptab : a.frag) + s_ops[o].m1_off, mat_bytes, &s_bar[o]);
          aa_tma(dstm + mat_doubles, ((s_ops[o].flags & AOPF_TIP2) && a.tiptab ? a.tiptab : a.frag) + s_ops[o].m2_off,
                 mat_bytes, &s_bar[o]);
        }
      for (unsigned int o = 0; o < nseg; ++o)
      {
        const AaOp & d = s_ops[o];
        const unsigned int flags = d.flags;
        const bool tip1 = flags & AOPF_TIP1, tip2 = flags & AOPF_TIP2;
        /* a tip child with a table: its slot holds T[code][k][20] = P x (tip vector of the code); the D
         * fragment of that child is two table entries per n-tile instead of 5 k-steps of DMMA */
        const bool look1 = tip1 && a.tiptab, look2 = tip2 && a.tiptab;
        const unsigned int st = o & 1u;
        /* A warp only ever touches its own sites, so the warps of a CTA are not kept in step: the CLV this
         * warp wrote in op o-1 only has to be visible to its own lanes */
        __syncwarp();
        if (fresh)
        {
          aa_mbar_wait(&s_bar[st], (ph >> st) & 1u);
          ph ^= 1u << st;
        }
        const double * sL = s_mat + (size_t)st * 2 * mat_doubles;
        const double * sR = sL + mat_doubles;
        const bool check = (flags & AOPF_PS) && !(tip1 && tip2);
        int smax[AA_T]; /* per tile: max (signed) hi word of this lane's site (row) over k, i */
#pragma unroll
        for (int t = 0; t < AA_T; ++t) smax[t] = AA_HI_MIN;

        /* A fragments (child values of 8 sites x 4 child states per k-step) are the only operands that
         * come from global memory; the ones of step (k, t) + 1 are requested before the 30 DMMAs of step
         * (k, t) run (ncu: the first DMMA of every tile was waiting on them, 34 % of all stall samples).
         * The op body is instantiated per (LOOK1, LOOK2): a tip-tip op has no DMMA at all, a tip-inner op
         * one accumulator family, and neither pays registers for the other's operands. */
        auto run = [&](auto look1_tag, auto look2_tag) {
          constexpr bool LOOK1 = decltype(look1_tag)::value, LOOK2 = decltype(look2_tag)::value;
          auto load_a = [&](unsigned int k, int t, double (&aL)[5], double (&aR)[5]) {
            const unsigned long long site = wsite0 + 8 * t + row;
            const bool act = site < a.sites;
            if (LOOK1)
              aL[0] = __hiloint2double(0, act ? (int)a.tipchars[d.c1_off + site] : 0); /* the code rides in the fragment */
            else if (tip1)
            {
              const unsigned long long m = act ? s_tipmap[a.tipchars[d.c1_off + site]] : 0ull;
#pragma unroll
              for (int ks = 0; ks < 5; ++ks) aL[ks] = (double)((m >> aa_state_of_slot(ks, col)) & 1ull);
            }
            else
            {
              aa_load_slots(a.clv + d.c1_off + site * span + k * 20, col, act, aL);
            }
            if (LOOK2)
              aR[0] = __hiloint2double(0, act ? (int)a.tipchars[d.c2_off + site] : 0);
            else if (tip2)
            {
              const unsigned long long m = act ? s_tipmap[a.tipchars[d.c2_off + site]] : 0ull;
#pragma unroll
              for (int ks = 0; ks < 5; ++ks) aR[ks] = (double)((m >> aa_state_of_slot(ks, col)) & 1ull);
            }
            else
            {
              aa_load_slots(a.clv + d.c2_off + site * span + k * 20, col, act, aR);
            }
          };
          constexpr int NA1 = LOOK1 ? 1 : 5, NA2 = LOOK2 ? 1 : 5; /* live operand registers per child */
          double aL[5], aR[5];
          load_a(0, 0, aL, aR);
          for (unsigned int k = 0; k < K; ++k)
          {
            double bL[LOOK1 ? 1 : 15], bR[LOOK2 ? 1 : 15];
            if (!LOOK1)
#pragma unroll
              for (int f = 0; f < 15; ++f) bL[LOOK1 ? 0 : f] = sL[(k * 15 + f) * 32 + lane];
            if (!LOOK2)
#pragma unroll
              for (int f = 0; f < 15; ++f) bR[LOOK2 ? 0 : f] = sR[(k * 15 + f) * 32 + lane];
#pragma unroll
            for (int t = 0; t < AA_T; ++t)
            {
              const unsigned long long site = wsite0 + 8 * t + row;
              const bool act = site < a.sites;
              double nL[5], nR[5];
              if (t + 1 < AA_T)
                load_a(k, t + 1, nL, nR);
              else if (k + 1 < K)
                load_a(k + 1, 0, nL, nR);
              double xl[3][2], xr[3][2];
#pragma unroll
              for (int nt = 0; nt < 3; ++nt) xl[nt][0] = xl[nt][1] = xr[nt][0] = xr[nt][1] = 0.0;
              if (LOOK1)
              {
                const double * row_l = sL + ((unsigned int)__double2loint(aL[0]) * K + k) * 20;
#pragma unroll
                for (int nt = 0; nt < 3; ++nt)
                  if (nt < 2 || col < 2)
                  {
                    const double2 v = *reinterpret_cast<const double2 *>(row_l + aa_dstate(nt, col));
                    xl[nt][0] = v.x;
                    xl[nt][1] = v.y;
                  }
              }
              if (LOOK2)
              {
                const double * row_r = sR + ((unsigned int)__double2loint(aR[0]) * K + k) * 20;
#pragma unroll
                for (int nt = 0; nt < 3; ++nt)
                  if (nt < 2 || col < 2)
                  {
                    const double2 v = *reinterpret_cast<const double2 *>(row_r + aa_dstate(nt, col));
                    xr[nt][0] = v.x;
                    xr[nt][1] = v.y;
                  }
              }
              /* k-step outermost: the independent accumulator chains are interleaved, so a DMMA never
               * waits on the one issued just before it */
#pragma unroll
              for (int ks = 0; ks < 5; ++ks)
#pragma unroll
                for (int nt = 0; nt < 3; ++nt)
                {
                  if (!LOOK1) dmma(xl[nt][0], xl[nt][1], aL[ks], bL[LOOK1 ? 0 : nt * 5 + ks]);
                  if (!LOOK2) dmma(xr[nt][0], xr[nt][1], aR[ks], bR[LOOK2 ? 0 : nt * 5 + ks]);
                }
              /* D fragment: this lane holds parent[site][k][aa_dstate(nt, col), +1] */
              double * pd = a.pbase + d.p_off + site * span + k * 20;
#pragma unroll
              for (int nt = 0; nt < 3; ++nt)
              {
                const double o0 = xl[nt][0] * xr[nt][0], o1 = xl[nt][1] * xr[nt][1];
                if (nt < 2 || col < 2)
                {
                  if (check) smax[t] = max(smax[t], max(__double2hiint(o0), __double2hiint(o1)));
                  if (act) *reinterpret_cast<double2 *>(pd + aa_dstate(nt, col)) = make_double2(o0, o1);
                }
              }
#pragma unroll
              for (int ks = 0; ks < NA1; ++ks) aL[ks] = nL[ks];
#pragma unroll
              for (int ks = 0; ks < NA2; ++ks) aR[ks] = nR[ks];
            }
          }
        };
        if (look1 && look2)
          run(std::true_type{}, std::true_type{});
        else if (look1)
          run(std::true_type{}, std::false_type{});
        else if (look2)
          run(std::false_type{}, std::true_type{});
        else
          run(std::false_type{}, std::false_type{});

        /* ---- scalers + (rare) rescale pass ---- */
        if (flags & AOPF_PS)
        {
#pragma unroll
          for (int t = 0; t < AA_T; ++t)
          {
            int m = smax[t];
            m = max(m, __shfl_xor_sync(0xffffffffu, m, 1));
            m = max(m, __shfl_xor_sync(0xffffffffu, m, 2));
            const bool scale = check && m < AA_THR_HI;
            const unsigned long long site = wsite0 + 8 * t + row;
            if (site < a.sites)
            {
              if (col == 0)
              {
                unsigned int v = scale ? 1u : 0u;
                if (flags & AOPF_S1) v += a.scaler[d.s1_off + site];
                if (flags & AOPF_S2) v += a.scaler[d.s2_off + site];
                a.scaler[d.ps_off + site] = v;
              }
              if (scale)
              {
                /* each lane rescales exactly the entries it wrote: all k, cols 2*col(+1) of the n-tiles */
                for (unsigned int k = 0; k < K; ++k)
                {
                  double * pd = a.pbase + d.p_off + site * span + k * 20;
                  for (int nt = 0; nt < 3; ++nt)
                    if (nt < 2 || col < 2)
                    {
                      pd[aa_dstate(nt, col)] *= SCALE_FACTOR_D;
                      pd[aa_dstate(nt, col) + 1] *= SCALE_FACTOR_D;
                    }
                }
              }
            }
          }
        }
        /* the last warp to leave op o refills its stage with the matrices of op o+2 */
        __syncwarp();
        if (lane == 0)
        {
          const unsigned int done = atomicAdd(&s_cnt[st], 1u);
          if (done == AA_WARPS - 1)
          {
            s_cnt[st] = 0;
            if (o + 2 < nseg)
            {
              double * dstm = s_mat + (size_t)st * 2 * mat_doubles;
              aa_mbar_expect(&s_bar[st], 2 * mat_bytes);
              aa_tma(dstm, ((s_ops[o + 2].flags & AOPF_TIP1) && a.tiptab ? a.tiptab : a.frag) + s_ops[o + 2].m1_off,
                     mat_bytes, &s_bar[st]);
              aa_tma(dstm + mat_doubles,
                     ((s_ops[o + 2].flags & AOPF_TIP2) && a.tiptab ? a.tiptab : a.frag) + s_ops[o + 2].m2_off, mat_bytes,
                     &s_bar[st]);
            }
          }
        }
      }
    }
  }
}

template <int AA_T, int AA_WARPS, int MINB>
__global__ void __launch_bounds__(AA_WARPS * 32, MINB) trav_aa_kernel(AaArgs a)
{
  trav_aa_body<AA_T, AA_WARPS, MINB>(a);
}

/* many partitions in one launch (blockIdx.y = partition; batch mode, internal.h) */
template <int AA_T, int AA_WARPS, int MINB>
__global__ void __launch_bounds__(AA_WARPS * 32, MINB) trav_aa_batch_kernel(const AaArgs * __restrict__ arr)
{
  __shared__ AaArgs sa;
  for (unsigned int i = threadIdx.x; i < sizeof(AaArgs) / 4; i += blockDim.x)
    reinterpret_cast<unsigned int *>(&sa)[i] = reinterpret_cast<const unsigned int *>(arr + blockIdx.y)[i];
  __syncthreads();
  trav_aa_body<AA_T, AA_WARPS, MINB>(sa);
}

/* ---- Gamma4 kernel with register forwarding ------------------------------------------------------
 * In a post-order operation list the parent written by operation o-1 is a child of operation o whenever
 * o has an inner child at all that was not finished earlier: always for a tip-inner operation, and for
 * one of the two children of an inner-inner operation.  With the slot order of devutil.cuh the D
 * fragments a lane holds after o-1 ARE its A fragments for o (k-steps 0-3; one shuffle for k-step 4), so
 * that child is never read back: of 2 x 129 + 238 child CLVs of the C3 tree only 129 come from memory.
 * The host puts the children of every operation in the order (lookup | load, lookup | load | forward)
 * — the product of the two child terms commutes bit for bit — so five op bodies cover all cases.
 * Rate categories are a template constant (4): the forwarded fragments live in registers
 * fw[k][tile][slot], which needs every index to be known at compile time. */
#define AOPF_FWD2 32u
/* ops staged per CTA.  64 x 64 B + 32 B + 2 stages x 30 KB = 65,568 B: two CTAs fit the 132 KB shared-memory
 * carve-out and leave 124 KB of L1 (with 128 ops and the unused tip map the next step, 164 KB, was taken) */
#define AA4_SEG 64
#define AOPF_SYNC0 64u /* a child read from memory is the previous operation's parent (both children are that node) */
enum { AK_LOOK = 0, AK_LOAD = 2, AK_FWD = 3 };

/* TS: the parent tile (8 sites x 640 B, contiguous in the CLV arena) is assembled in a per-warp shared-memory
 * buffer `stg` and leaves with one TMA bulk store per operation instead of 16-/32-byte global stores from the
 * fragment layout (which cost twice the L1TEX wavefronts of the bytes they carry) */
__device__ __forceinline__ void aa_bulk_wait_read()
{
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void aa_bulk_wait()
{
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void aa_bulk_store(double * gdst, const double * ssrc, unsigned int bytes)
{
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(aa_smem_u32(ssrc)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

template <int C1, int C2, int AA_T, bool TS>
__device__ __forceinline__ void aa4_op(const AaArgs & a, const AaOp & d, const double * __restrict__ sL,
                                       const double * __restrict__ sR, unsigned long long wsite0, int lane, bool check,
                                       const unsigned int (&code1)[AA_T], const unsigned int (&code2)[AA_T],
                                       double (&fw)[4][AA_T][5], int (&smax)[AA_T], double * stg,
                                       const double (&cur)[AA_T][5])
{
  constexpr int K = 4;
  constexpr unsigned long long span = K * 20;
  const int row = lane >> 2, col = lane & 3;
  const int src4 = (lane & ~3) | (col >> 1); /* lane that holds state 16 + col of this row in its D fragment */
  bool act[AA_T];
  const double * c1p[AA_T];
  const double * c2p[AA_T];
  double * pp[AA_T];
#pragma unroll
  for (int t = 0; t < AA_T; ++t)
  {
    const unsigned long long site = wsite0 + 8 * t + row;
    act[t] = site < a.sites;
    if (C1 == AK_LOAD) c1p[t] = a.clv + d.c1_off + site * span;
    if (C2 == AK_LOAD) c2p[t] = a.clv + d.c2_off + site * span;
    pp[t] = a.pbase + d.p_off + site * span;
  }
  /* children that come from memory: the fragments of category k+1 are requested before the DMMAs of k run */
  constexpr bool PF = AA_T == 1; /* two tiles per warp: no registers left for the look-ahead copy */
  double aL[AA_T][5], aR[AA_T][5], nL[AA_T][5], nR[AA_T][5];
  /* category 0 of the first child that comes from memory was requested during the previous operation (`cur`) */
  if (PF)
#pragma unroll
    for (int t = 0; t < AA_T; ++t)
    {
      if (C1 == AK_LOAD)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) aL[t][ks] = cur[t][ks];
      if (C2 == AK_LOAD && C1 != AK_LOAD)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) aR[t][ks] = cur[t][ks];
      if (C2 == AK_LOAD && C1 == AK_LOAD) aa_load_slots(c2p[t], col, act[t], aR[t]);
    }
#pragma unroll
  for (int k = 0; k < K; ++k)
  {
    if (PF && k + 1 < K)
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
      {
        if (C1 == AK_LOAD) aa_load_slots(c1p[t] + (k + 1) * 20, col, act[t], nL[t]);
        if (C2 == AK_LOAD) aa_load_slots(c2p[t] + (k + 1) * 20, col, act[t], nR[t]);
      }
    if (!PF)
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
      {
        if (C1 == AK_LOAD) aa_load_slots(c1p[t] + k * 20, col, act[t], aL[t]);
        if (C2 == AK_LOAD) aa_load_slots(c2p[t] + k * 20, col, act[t], aR[t]);
      }
    if (C2 == AK_FWD)
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) aR[t][ks] = fw[k][t][ks];
    double o[AA_T][3][2];
    /* tip children: x = T[code][k][.] from the table in this child's matrix slot.  A 16-byte shared-memory load is
     * served a quarter warp (two site rows) at a time and every table row starts at the same bank, so two rows with
     * different codes would collide on every access: even rows therefore fetch their n-tile 0 pair while odd rows
     * fetch their n-tile 1 pair (the other 16 bytes of the same 32), and the second load does the opposite */
    double lk1[AA_T][3][2], lk2[AA_T][3][2];
    const int odd = row & 1;
#pragma unroll
    for (int t = 0; t < AA_T; ++t)
    {
      if (C1 == AK_LOOK)
      {
        const double * rp = sL + (code1[t] * K + k) * 20;
        const double2 x = *reinterpret_cast<const double2 *>(rp + 4 * col + 2 * odd);
        const double2 y = *reinterpret_cast<const double2 *>(rp + 4 * col + 2 * (1 - odd));
        lk1[t][0][0] = odd ? y.x : x.x;
        lk1[t][0][1] = odd ? y.y : x.y;
        lk1[t][1][0] = odd ? x.x : y.x;
        lk1[t][1][1] = odd ? x.y : y.y;
        lk1[t][2][0] = lk1[t][2][1] = 0.0;
        if (col < 2)
        {
          const double2 z = *reinterpret_cast<const double2 *>(rp + 16 + 2 * col);
          lk1[t][2][0] = z.x;
          lk1[t][2][1] = z.y;
        }
      }
      if (C2 == AK_LOOK)
      {
        const double * rp = sR + (code2[t] * K + k) * 20;
        const double2 x = *reinterpret_cast<const double2 *>(rp + 4 * col + 2 * odd);
        const double2 y = *reinterpret_cast<const double2 *>(rp + 4 * col + 2 * (1 - odd));
        lk2[t][0][0] = odd ? y.x : x.x;
        lk2[t][0][1] = odd ? y.y : x.y;
        lk2[t][1][0] = odd ? x.x : y.x;
        lk2[t][1][1] = odd ? x.y : y.y;
        lk2[t][2][0] = lk2[t][2][1] = 0.0;
        if (col < 2)
        {
          const double2 z = *reinterpret_cast<const double2 *>(rp + 16 + 2 * col);
          lk2[t][2][0] = z.x;
          lk2[t][2][1] = z.y;
        }
      }
    }
    /* n-tile outermost: only the five B fragments of one n-tile (per child) are live, and they serve all tiles */
#pragma unroll
    for (int nt = 0; nt < 3; ++nt)
    {
      double bL[5], bR[5];
      if (C1 != AK_LOOK)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) bL[ks] = sL[(k * 15 + nt * 5 + ks) * 32 + lane];
      if (C2 != AK_LOOK)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) bR[ks] = sR[(k * 15 + nt * 5 + ks) * 32 + lane];
      double xl[AA_T][2], xr[AA_T][2];
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
      {
        xl[t][0] = xl[t][1] = xr[t][0] = xr[t][1] = 0.0;
        if (C1 == AK_LOOK)
        {
          xl[t][0] = lk1[t][nt][0];
          xl[t][1] = lk1[t][nt][1];
        }
        if (C2 == AK_LOOK)
        {
          xr[t][0] = lk2[t][nt][0];
          xr[t][1] = lk2[t][nt][1];
        }
      }
      /* k-step outermost: the 2 x AA_T accumulator chains are interleaved */
#pragma unroll
      for (int ks = 0; ks < 5; ++ks)
#pragma unroll
        for (int t = 0; t < AA_T; ++t)
        {
          if (C1 != AK_LOOK) dmma(xl[t][0], xl[t][1], aL[t][ks], bL[ks]);
          if (C2 != AK_LOOK) dmma(xr[t][0], xr[t][1], aR[t][ks], bR[ks]);
        }
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
      {
        o[t][nt][0] = xl[t][0] * xr[t][0];
        o[t][nt][1] = xl[t][1] * xr[t][1];
        if (check && (nt < 2 || col < 2))
          smax[t] = max(smax[t], max(__double2hiint(o[t][nt][0]), __double2hiint(o[t][nt][1])));
      }
    }
    /* D fragments of this lane: parent[site][k][4 col .. 4 col + 3] and, for col < 2, [16 + 2 col, + 1] */
#pragma unroll
    for (int t = 0; t < AA_T; ++t)
    {
      if (TS)
      {
        if (k == 0 && t == 0)
        {
          /* the bulk store of the previous operation has finished reading the buffer */
          if (lane == 0) aa_bulk_wait_read();
          __syncwarp();
        }
        /* every row of the buffer starts at the same bank: odd rows write the two halves of their 32 bytes in the
         * opposite order (as for the table lookups) */
        double * sp = stg + (8 * t + row) * 80 + k * 20;
        const int odd = row & 1;
        *reinterpret_cast<double2 *>(sp + 4 * col + 2 * odd) = odd ? make_double2(o[t][1][0], o[t][1][1]) : make_double2(o[t][0][0], o[t][0][1]);
        *reinterpret_cast<double2 *>(sp + 4 * col + 2 * (1 - odd)) = odd ? make_double2(o[t][0][0], o[t][0][1]) : make_double2(o[t][1][0], o[t][1][1]);
        if (col < 2) *reinterpret_cast<double2 *>(sp + 16 + 2 * col) = make_double2(o[t][2][0], o[t][2][1]);
      }
      else if (act[t])
      {
        d4 v;
        v.x = o[t][0][0];
        v.y = o[t][0][1];
        v.z = o[t][1][0];
        v.w = o[t][1][1];
        stg256(pp[t] + k * 20 + 4 * col, v);
        if (col < 2) *reinterpret_cast<double2 *>(pp[t] + k * 20 + 16 + 2 * col) = make_double2(o[t][2][0], o[t][2][1]);
      }
      /* ... and, re-read as an A fragment, the operand of the next operation */
      fw[k][t][0] = o[t][0][0];
      fw[k][t][1] = o[t][0][1];
      fw[k][t][2] = o[t][1][0];
      fw[k][t][3] = o[t][1][1];
      const double v0 = __shfl_sync(0xffffffffu, o[t][2][0], src4), v1 = __shfl_sync(0xffffffffu, o[t][2][1], src4);
      fw[k][t][4] = (col & 1) ? v1 : v0;
    }
    if (PF && k + 1 < K)
#pragma unroll
      for (int t = 0; t < AA_T; ++t)
#pragma unroll
        for (int ks = 0; ks < 5; ++ks)
        {
          if (C1 == AK_LOAD) aL[t][ks] = nL[t][ks];
          if (C2 == AK_LOAD) aR[t][ks] = nR[t][ks];
        }
  }
}

/* dynamic shared memory: AaOp ops[AA4_SEG] | u64 bar[NS] | u32 cnt[NS] (+ pad) | double mat[NS stages][2][4*480] | staging */
template <int AA_T, int AA_WARPS, int MINB, bool TS, int NS = 2> /* NS: matrix stages (operations in flight) */
__device__ __forceinline__ void trav_aa4_body(const AaArgs & a)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  AaOp * s_ops = reinterpret_cast<AaOp *>(smem_raw);
  unsigned long long * s_bar = reinterpret_cast<unsigned long long *>(smem_raw + sizeof(AaOp) * AA4_SEG);
  unsigned int * s_cnt = reinterpret_cast<unsigned int *>(s_bar + NS);
  double * s_mat = reinterpret_cast<double *>(s_bar + 2 * NS); /* NS = 2: 32 bytes as in trav_aa_body, NS = 3: 48 */

  constexpr unsigned int K = 4;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = lane >> 2, col = lane & 3;
  constexpr unsigned int mat_doubles = K * AA_FRAG;
  /* TS: per-warp parent tile buffers behind the matrix stages, [AA_T x 8 sites][K][20] */
  double * const stg = TS ? s_mat + (size_t)2 * NS * mat_doubles + (size_t)warp * AA_T * 8 * 80 : nullptr;
  constexpr unsigned int mat_bytes = mat_doubles * sizeof(double);
  constexpr unsigned long long span = K * 20;
  const unsigned long long sites_per_cta = (unsigned long long)AA_WARPS * 8 * AA_T;
  const unsigned long long nchunks = (a.sites + sites_per_cta - 1) / sites_per_cta;
  if (blockIdx.x >= nchunks) return;

  if (threadIdx.x == 0)
  {
    for (int i = 0; i < NS; ++i)
    {
      aa_mbar_init(&s_bar[i], 1);
      s_cnt[i] = 0;
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  unsigned int ph = 0;
  double fw[4][AA_T][5];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int t = 0; t < AA_T; ++t)
#pragma unroll
      for (int ks = 0; ks < 5; ++ks) fw[k][t][ks] = 0.0;

  for (unsigned int seg = 0; seg < a.nops; seg += AA4_SEG)
  {
    const unsigned int nseg = min(a.nops - seg, (unsigned int)AA4_SEG);
    __syncthreads();
    {
      const uint4 * src = reinterpret_cast<const uint4 *>(a.ops + seg);
      uint4 * dst = reinterpret_cast<uint4 *>(s_ops);
      for (unsigned int i = threadIdx.x; i < nseg * 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    for (unsigned long long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x)
    {
      const unsigned long long wsite0 = chunk * sites_per_cta + (unsigned long long)warp * 8 * AA_T;
      __syncthreads(); /* both stages free, counters zero */
      /* a launch of at most NS operations (the single CLV updates of branch-length optimisation) keeps its
       * matrices in the stages: loaded for the CTA's first chunk only */
      const bool fresh = a.nops > NS || chunk == blockIdx.x;
      if (threadIdx.x == 0 && fresh)
        for (unsigned int o = 0; o < NS && o < nseg; ++o)
        {
          double * dstm = s_mat + (size_t)o * 2 * mat_doubles;
          aa_mbar_expect(&s_bar[o], 2 * mat_bytes);
          aa_tma(dstm, ((s_ops[o].flags & AOPF_TIP1) ? a.tiptab : a.frag) + s_ops[o].m1_off, mat_bytes, &s_bar[o]);
          aa_tma(dstm + mat_doubles, ((s_ops[o].flags & AOPF_TIP2) ? a.tiptab : a.frag) + s_ops[o].m2_off, mat_bytes,
                 &s_bar[o]);
        }
      unsigned int st = 0; /* stage of operation o: o mod NS */
      /* tip codes of the sites of this warp: the ones of op o+1 are requested while op o computes */
      unsigned int code1[AA_T], code2[AA_T], ncode1[AA_T], ncode2[AA_T];
      auto load_codes = [&](const AaOp & q, unsigned int (&q1)[AA_T], unsigned int (&q2)[AA_T]) {
#pragma unroll
        for (int t = 0; t < AA_T; ++t)
        {
          const unsigned long long site = wsite0 + 8 * t + row;
          q1[t] = ((q.flags & AOPF_TIP1) && site < a.sites) ? a.tipchars[q.c1_off + site] : 0u;
          q2[t] = ((q.flags & AOPF_TIP2) && site < a.sites) ? a.tipchars[q.c2_off + site] : 0u;
        }
      };
      load_codes(s_ops[0], code1, code2);
      /* category 0 of the first child of an operation that is read from memory (not a tip, not forwarded): requested
       * one operation ahead, like the tip codes.  That child was written at least two operations earlier. */
      double cur[AA_T][5], nxt[AA_T][5];
      auto load_first = [&](const AaOp & q, bool fwd_ok, double (&dst)[AA_T][5]) {
        const bool l1 = !(q.flags & AOPF_TIP1);
        const bool l2 = !(q.flags & AOPF_TIP2) && !((q.flags & AOPF_FWD2) && fwd_ok);
        if (AA_T == 1 && (l1 || l2))
        {
          const unsigned long long off = l1 ? q.c1_off : q.c2_off;
#pragma unroll
          for (int t = 0; t < AA_T; ++t)
          {
            const unsigned long long site = wsite0 + 8 * t + row;
            aa_load_slots(a.clv + off + site * span, col, site < a.sites, dst[t]);
          }
        }
      };
      load_first(s_ops[0], false, nxt);
      for (unsigned int o = 0; o < nseg; ++o)
      {
        const bool prefetched = o == 0 || (!TS && !(s_ops[o].flags & AOPF_SYNC0));
        if (prefetched)
#pragma unroll
          for (int t = 0; t < AA_T; ++t)
#pragma unroll
            for (int ks = 0; ks < 5; ++ks) cur[t][ks] = nxt[t][ks];
        const AaOp & d = s_ops[o];
        const unsigned int flags = d.flags;
        const bool look1 = flags & AOPF_TIP1, look2 = flags & AOPF_TIP2;
        const bool fwd2 = (flags & AOPF_FWD2) && o > 0; /* the first operation of a chunk has no predecessor */
        if (o + 1 < nseg) load_codes(s_ops[o + 1], ncode1, ncode2);
        if (!TS && o + 1 < nseg && !(s_ops[o + 1].flags & AOPF_SYNC0)) load_first(s_ops[o + 1], true, nxt);
        if (!TS && a.l1_prefetch && o + 1 < nseg && wsite0 + 8 * AA_T <= a.sites)
        {
          /* the warp's tile of a child is AA_T x 5 KB of consecutive memory: 40 lines per 8 sites */
          const AaOp & q = s_ops[o + 1];
          const bool l1 = !(q.flags & AOPF_TIP1), l2 = !(q.flags & (AOPF_TIP2 | AOPF_FWD2));
#pragma unroll
          for (int i = 0; i < (40 * AA_T + 31) / 32; ++i)
          {
            const int line = i * 32 + lane;
            if (line < 40 * AA_T)
            {
              if (l1) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.clv + q.c1_off + wsite0 * span + line * 16));
              if (l2) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.clv + q.c2_off + wsite0 * span + line * 16));
            }
          }
        }
        /* child scaler counts: requested now, added after the operation */
        unsigned int sv1[AA_T], sv2[AA_T]; /* two registers: an add here would wait for both loads right away */
#pragma unroll
        for (int t = 0; t < AA_T; ++t)
        {
          const unsigned long long site = wsite0 + 8 * t + row;
          sv1[t] = sv2[t] = 0;
          if (col == 0 && site < a.sites)
          {
            if (flags & AOPF_S1) sv1[t] = a.scaler[d.s1_off + site];
            if (flags & AOPF_S2) sv2[t] = a.scaler[d.s2_off + site];
          }
        }
        __syncwarp(); /* children written by earlier operations of this warp are visible to all its lanes */
        if (fresh)
        {
          aa_mbar_wait(&s_bar[st], (ph >> st) & 1u);
          ph ^= 1u << st;
        }
        const double * sL = s_mat + (size_t)st * 2 * mat_doubles;
        const double * sR = sL + mat_doubles;
        const bool check = (flags & AOPF_PS) && !(look1 && look2);
        int smax[AA_T];
#pragma unroll
        for (int t = 0; t < AA_T; ++t) smax[t] = AA_HI_MIN;
        if (TS && !(look1 && (look2 || fwd2)))
        {
          /* a child read from memory was written at least two operations ago: all bulk stores but the newest are
           * complete (and their data visible to this thread, then to the warp) */
          if (lane == 0)
          {
            if (flags & AOPF_SYNC0)
              aa_bulk_wait<0>();
            else
              aa_bulk_wait<1>();
          }
          __syncwarp();
        }
        if (!prefetched) load_first(d, true, cur);

        if (look1 && look2)
          aa4_op<AK_LOOK, AK_LOOK, AA_T, TS>(a, d, sL, sR, wsite0, lane, check, code1, code2, fw, smax, stg, cur);
        else if (look1 && fwd2)
          aa4_op<AK_LOOK, AK_FWD, AA_T, TS>(a, d, sL, sR, wsite0, lane, check, code1, code2, fw, smax, stg, cur);
        else if (look1)
          aa4_op<AK_LOOK, AK_LOAD, AA_T, TS>(a, d, sL, sR, wsite0, lane, check, code1, code2, fw, smax, stg, cur);
        else if (fwd2)
          aa4_op<AK_LOAD, AK_FWD, AA_T, TS>(a, d, sL, sR, wsite0, lane, check, code1, code2, fw, smax, stg, cur);
        else
          aa4_op<AK_LOAD, AK_LOAD, AA_T, TS>(a, d, sL, sR, wsite0, lane, check, code1, code2, fw, smax, stg, cur);

        /* ---- scalers + (rare) rescale pass ---- */
        if (flags & AOPF_PS)
        {
#pragma unroll
          for (int t = 0; t < AA_T; ++t)
          {
            int m = smax[t];
            m = max(m, __shfl_xor_sync(0xffffffffu, m, 1));
            m = max(m, __shfl_xor_sync(0xffffffffu, m, 2));
            const bool scale = check && m < AA_THR_HI;
            const unsigned long long site = wsite0 + 8 * t + row;
            if (site < a.sites)
            {
              if (col == 0) a.scaler[d.ps_off + site] = sv1[t] + sv2[t] + (scale ? 1u : 0u);
              if (scale)
              {
                /* each lane rescales exactly the entries it wrote, in memory and in the forwarded copy */
                for (unsigned int k = 0; k < K; ++k)
                {
                  double * pd = TS ? stg + (8 * t + row) * 80 + k * 20 : a.pbase + d.p_off + site * span + k * 20;
                  for (int nt = 0; nt < 3; ++nt)
                    if (nt < 2 || col < 2)
                    {
                      pd[aa_dstate(nt, col)] *= SCALE_FACTOR_D;
                      pd[aa_dstate(nt, col) + 1] *= SCALE_FACTOR_D;
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                  for (int ks = 0; ks < 5; ++ks) fw[k][t][ks] *= SCALE_FACTOR_D;
              }
            }
          }
        }
#pragma unroll
        for (int t = 0; t < AA_T; ++t)
        {
          code1[t] = ncode1[t];
          code2[t] = ncode2[t];
        }
        if (TS)
        {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* this lane's buffer writes -> async proxy */
          __syncwarp();
          if (lane == 0 && wsite0 < a.sites)
          {
            const unsigned long long n = min((unsigned long long)(8 * AA_T), (unsigned long long)a.sites - wsite0);
            aa_bulk_store(a.pbase + d.p_off + wsite0 * span, stg, (unsigned int)(n * span * sizeof(double)));
          }
        }
        /* the last warp to leave op o refills its stage with the matrices of op o+2 */
        __syncwarp();
        if (lane == 0)
        {
          const unsigned int done = atomicAdd(&s_cnt[st], 1u);
          if (done == AA_WARPS - 1)
          {
            s_cnt[st] = 0;
            if (o + NS < nseg)
            {
              double * dstm = s_mat + (size_t)st * 2 * mat_doubles;
              aa_mbar_expect(&s_bar[st], 2 * mat_bytes);
              aa_tma(dstm, ((s_ops[o + NS].flags & AOPF_TIP1) ? a.tiptab : a.frag) + s_ops[o + NS].m1_off, mat_bytes,
                     &s_bar[st]);
              aa_tma(dstm + mat_doubles, ((s_ops[o + NS].flags & AOPF_TIP2) ? a.tiptab : a.frag) + s_ops[o + NS].m2_off,
                     mat_bytes, &s_bar[st]);
            }
          }
        }
        st = st + 1 == NS ? 0 : st + 1;
      }
      if (TS)
      {
        /* segment / chunk boundary: every store of this warp is complete before its sites are revisited */
        if (lane == 0) aa_bulk_wait<0>();
        __syncwarp();
      }
    }
  }
}

template <int AA_T, int AA_WARPS, int MINB, bool TS, int NS>
__global__ void __launch_bounds__(AA_WARPS * 32, MINB) trav_aa4_kernel(AaArgs a)
{
  trav_aa4_body<AA_T, AA_WARPS, MINB, TS, NS>(a);
}

template <int AA_T, int AA_WARPS, int MINB, bool TS>
__global__ void __launch_bounds__(AA_WARPS * 32, MINB) trav_aa4_batch_kernel(const AaArgs * __restrict__ arr)
{
  __shared__ AaArgs sa;
  for (unsigned int i = threadIdx.x; i < sizeof(AaArgs) / 4; i += blockDim.x)
    reinterpret_cast<unsigned int *>(&sa)[i] = reinterpret_cast<const unsigned int *>(arr + blockIdx.y)[i];
  __syncthreads();
  trav_aa4_body<AA_T, AA_WARPS, MINB, TS>(sa);
}

/* ---- host ------------------------------------------------------------------------------------ */

static int aa_env_int(const char * name, int dflt)
{
  const char * v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}
int g_opt_aa_variant = aa_env_int("PLLB200_AA_VARIANT", 0); /* tuning knob: tile / warp / occupancy variants below */
int g_opt_aa_tiptab = 1; /* tip children by table lookup instead of DMMA */

template <int T, int W, int MINB>
static bool launch_aa_variant(PartCtx * c, const AaArgs & a)
{
  const unsigned int K = a.K;
  const size_t smem = sizeof(AaOp) * AA_SEG + 256 * 8 + 4 * 8 + (size_t)2 * 2 * K * AA_FRAG * sizeof(double);
  /* function attributes belong to a device: one slot per device (a process may drive several) */
  static size_t smem_set_dev[64] = {0};
  static int per_sm_dev[64] = {0};
  size_t & smem_set = smem_set_dev[c->device & 63];
  int & per_sm = per_sm_dev[c->device & 63];
  if (smem > smem_set)
  {
    if (!pllb200_cuda_ok(cudaFuncSetAttribute(trav_aa_kernel<T, W, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem),
                         "trav_aa smem attribute"))
      return false;
    smem_set = smem;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trav_aa_kernel<T, W, MINB>, W * 32, smem) != cudaSuccess ||
        per_sm < 1)
      per_sm = 1;
  }
  const unsigned long long spc = (unsigned long long)W * 8 * T;
  unsigned long long blocks = (a.sites + spc - 1) / spc;
  blocks = std::min<unsigned long long>(blocks, (unsigned long long)c->sm_count * per_sm);
  if (blocks < 1) blocks = 1;
  if (batch_collects(BK_TRAV_AA) && T == 2 && W == 4 && MINB == 3)
  {
    batch_collect(BK_TRAV_AA, &a, sizeof(a), (unsigned int)blocks, smem);
    return true;
  }
  trav_aa_kernel<T, W, MINB><<<(unsigned int)blocks, W * 32, smem, xstream(c)>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "trav_aa_kernel launch");
}

template <int T, int W, int MINB, bool TS = false, int NS = 2>
static bool launch_aa4_variant(PartCtx * c, const AaArgs & a)
{
  const size_t smem = sizeof(AaOp) * AA4_SEG + 2 * NS * 8 + (size_t)NS * 2 * 4 * AA_FRAG * sizeof(double) +
                      (TS ? (size_t)W * T * 8 * 80 * sizeof(double) : 0);
  static bool attr_set_dev[64] = {false};
  static int per_sm_dev[64] = {0};
  bool & attr_set = attr_set_dev[c->device & 63];
  int & per_sm = per_sm_dev[c->device & 63];
  if (!attr_set)
  {
    if (!pllb200_cuda_ok(cudaFuncSetAttribute(trav_aa4_kernel<T, W, MINB, TS, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem),
                         "trav_aa4 smem attribute"))
      return false;
    attr_set = true;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trav_aa4_kernel<T, W, MINB, TS, NS>, W * 32, smem) != cudaSuccess ||
        per_sm < 1)
      per_sm = 1;
  }
  const unsigned long long spc = (unsigned long long)W * 8 * T;
  unsigned long long blocks = (a.sites + spc - 1) / spc;
  blocks = std::min<unsigned long long>(blocks, (unsigned long long)c->sm_count * per_sm);
  if (blocks < 1) blocks = 1;
  trav_aa4_kernel<T, W, MINB, TS, NS><<<(unsigned int)blocks, W * 32, smem, xstream(c)>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "trav_aa4_kernel launch");
}

/* tile / warp / occupancy of the default Gamma4 kernel (measured best of the variants below at C3) */
#define AA4_T 1
#define AA4_W 8
#define AA4_MINB 2
#define AA4_TS false
static const size_t AA4_SMEM = sizeof(AaOp) * AA4_SEG + 4 * 8 + (size_t)2 * 2 * 4 * AA_FRAG * sizeof(double);

/* children in the order the Gamma4 kernel expects (lookup | load, lookup | load | forward) */
static void aa4_canonical_children(AaOp * dops, unsigned int count)
{
  for (unsigned int i = 0; i < count; ++i)
  {
    AaOp & d = dops[i];
    const bool t1 = d.flags & AOPF_TIP1, t2 = d.flags & AOPF_TIP2;
    const bool f1 = i > 0 && !t1 && d.c1_off == dops[i - 1].p_off;
    const bool f2 = i > 0 && !t2 && d.c2_off == dops[i - 1].p_off;
    const bool swap = f1 ? !f2 : (!f2 && !t1 && t2);
    if (swap)
    {
      std::swap(d.c1_off, d.c2_off);
      std::swap(d.s1_off, d.s2_off);
      std::swap(d.m1_off, d.m2_off);
      const unsigned int f = d.flags;
      d.flags = (f & ~(AOPF_TIP1 | AOPF_TIP2 | AOPF_S1 | AOPF_S2)) | ((f & AOPF_TIP1) ? AOPF_TIP2 : 0u) |
                ((f & AOPF_TIP2) ? AOPF_TIP1 : 0u) | ((f & AOPF_S1) ? AOPF_S2 : 0u) | ((f & AOPF_S2) ? AOPF_S1 : 0u);
    }
    if (f1 || f2) d.flags |= AOPF_FWD2;
    if (f1 && f2) d.flags |= AOPF_SYNC0;
  }
}

bool launch_trav_aa(PartCtx * c, const pll_operation_t * ops, unsigned int count)
{
  pll_partition_t * p = &c->pub;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  const unsigned int K = p->rate_cats;
  static_assert(sizeof(AaOp) == 64, "AaOp is 8 x u64");
  c->devop_scratch.resize((size_t)count * 8);
  AaOp * dops = reinterpret_cast<AaOp *>(c->devop_scratch.data());
  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops[i];
    AaOp & d = dops[i];
    memset(&d, 0, sizeof(d));
    d.p_off = (unsigned long long)(o.parent_clv_index - c->clv_first) * c->clv_stride;
    if (pattern_tip && o.child1_clv_index < p->tips)
    {
      d.flags |= AOPF_TIP1;
      d.c1_off = (unsigned long long)o.child1_clv_index * c->tip_stride;
    }
    else
      d.c1_off = (unsigned long long)(o.child1_clv_index - c->clv_first) * c->clv_stride;
    if (pattern_tip && o.child2_clv_index < p->tips)
    {
      d.flags |= AOPF_TIP2;
      d.c2_off = (unsigned long long)o.child2_clv_index * c->tip_stride;
    }
    else
      d.c2_off = (unsigned long long)(o.child2_clv_index - c->clv_first) * c->clv_stride;
    if (o.parent_scaler_index >= 0)
    {
      d.flags |= AOPF_PS;
      d.ps_off = (unsigned long long)o.parent_scaler_index * c->scaler_stride;
    }
    if (o.child1_scaler_index >= 0)
    {
      d.flags |= AOPF_S1;
      d.s1_off = (unsigned long long)o.child1_scaler_index * c->scaler_stride;
    }
    if (o.child2_scaler_index >= 0)
    {
      d.flags |= AOPF_S2;
      d.s2_off = (unsigned long long)o.child2_scaler_index * c->scaler_stride;
    }
    d.m1_off = o.child1_matrix_index * K * AA_FRAG;
    d.m2_off = o.child2_matrix_index * K * AA_FRAG;
  }
  const bool collect = batch_collects(BK_TRAV_AA) && g_opt_aa_variant == 0; /* as in launch_aa_variant */
  const bool tiptab_ok = g_opt_aa_tiptab && c->d_tiptab && c->d_tipmap && p->maxstates <= 24;
  /* Gamma4 kernel with register forwarding: every tip child must be a table lookup */
  const bool fwd_kernel = K == 4 && (g_opt_aa_variant == 0 || g_opt_aa_variant >= 20) && (tiptab_ok || !pattern_tip);
  if (fwd_kernel) aa4_canonical_children(dops, count);
  const AaOp * d_ops = collect ? (const AaOp *)batch_pool_add(dops, sizeof(AaOp) * count)
                               : (const AaOp *)dev_stage_small(c, dops, sizeof(AaOp) * count);
  if (!d_ops) return false;
  AaArgs a;
  a.ops = d_ops;
  a.nops = count;
  a.sites = c->sites_alloc;
  a.K = K;
  a.clv = c->d_clv;
  a.pbase = c->d_clv;
  a.scaler = c->d_scaler;
  a.frag = c->d_pmat_il;
  a.tiptab = nullptr;
  if (g_opt_aa_tiptab && c->d_tiptab && c->d_tipmap && p->maxstates <= 24)
  {
    if (c->tiptab_stale && !launch_pmatrix_reinterleave(c)) return false; /* tip map changed since the tables were built */
    a.tiptab = c->d_tiptab;
  }
  a.tipchars = c->d_tipchars;
  a.tipmap = c->d_tipmap;
  a.l1_prefetch = g_opt_trav_prefetch;
  if (fwd_kernel && collect)
  {
    const unsigned long long spc = (unsigned long long)AA4_W * 8 * AA4_T;
    batch_collect(BK_TRAV_AA4, &a, sizeof(a), (unsigned int)((a.sites + spc - 1) / spc), AA4_SMEM);
    return true;
  }
  if (fwd_kernel)
    switch (g_opt_aa_variant)
    {
      case 21: return launch_aa4_variant<1, 4, 3>(c, a);
      case 22: return launch_aa4_variant<2, 4, 2>(c, a);
      case 23: return launch_aa4_variant<1, 8, 1>(c, a);
      case 24: return launch_aa4_variant<2, 4, 3>(c, a);
      case 25: return launch_aa4_variant<1, 6, 2>(c, a);
      case 26: return launch_aa4_variant<1, 5, 3>(c, a);
      case 27: return launch_aa4_variant<1, 8, 2>(c, a);
      case 28: return launch_aa4_variant<2, 5, 2>(c, a);
      case 29: return launch_aa4_variant<2, 6, 2>(c, a);
      case 30: return launch_aa4_variant<1, 6, 3>(c, a);
      case 31: return launch_aa4_variant<1, 9, 2>(c, a);
      case 32: return launch_aa4_variant<1, 10, 2>(c, a);
      case 33: return launch_aa4_variant<1, 7, 3>(c, a);
      case 40: return launch_aa4_variant<1, 8, 2, true>(c, a);
      case 50: return launch_aa4_variant<1, 8, 2, false, 3>(c, a);
      case 51: return launch_aa4_variant<1, 6, 2, false, 3>(c, a);
      case 41: return launch_aa4_variant<1, 6, 3, true>(c, a);
      case 42: return launch_aa4_variant<1, 9, 2, true>(c, a);
      default: return launch_aa4_variant<AA4_T, AA4_W, AA4_MINB, AA4_TS>(c, a);
    }
  switch (g_opt_aa_variant)
  {
    case 1: return launch_aa_variant<8, 4, 3>(c, a);
    case 2: return launch_aa_variant<4, 8, 2>(c, a);
    case 3: return launch_aa_variant<2, 4, 3>(c, a);
    case 4: return launch_aa_variant<4, 6, 2>(c, a);
    case 5: return launch_aa_variant<3, 4, 3>(c, a);
    case 6: return launch_aa_variant<2, 8, 2>(c, a);
    case 7: return launch_aa_variant<2, 6, 3>(c, a);
    case 8: return launch_aa_variant<1, 4, 3>(c, a);
    case 9: return launch_aa_variant<2, 6, 2>(c, a);
    case 10: return launch_aa_variant<1, 8, 2>(c, a);
    case 11: return launch_aa_variant<1, 8, 3>(c, a);
    case 12: return launch_aa_variant<4, 4, 3>(c, a);
    default: return launch_aa_variant<2, 4, 3>(c, a);
  }
}

bool flush_trav_aa(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  static size_t smem_set_dev[64] = {0};
  size_t & smem_set = smem_set_dev[lead->device & 63];
  if (k.smem > smem_set)
  {
    if (!pllb200_cuda_ok(cudaFuncSetAttribute(trav_aa_batch_kernel<2, 4, 3>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k.smem),
                         "trav_aa batch smem attribute"))
      return false;
    smem_set = k.smem;
  }
  /* measured at 50 x 2500 sites: one CTA per 64-site chunk 80 us, one resident wave of CTAs looping over chunks 98 us */
  trav_aa_batch_kernel<2, 4, 3><<<dim3(k.max_blocks, k.count), 4 * 32, k.smem, lead->stream>>>(
      static_cast<const AaArgs *>(d_args));
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "batched trav_aa kernel launch");
}

bool flush_trav_aa4(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  static bool attr_set_dev[64] = {false};
  bool & attr_set = attr_set_dev[lead->device & 63];
  if (!attr_set)
  {
    if (!pllb200_cuda_ok(cudaFuncSetAttribute(trav_aa4_batch_kernel<AA4_T, AA4_W, AA4_MINB, AA4_TS>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AA4_SMEM),
                         "trav_aa4 batch smem attribute"))
      return false;
    attr_set = true;
  }
  trav_aa4_batch_kernel<AA4_T, AA4_W, AA4_MINB, AA4_TS><<<dim3(k.max_blocks, k.count), AA4_W * 32, AA4_SMEM, lead->stream>>>(
      static_cast<const AaArgs *>(d_args));
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "batched trav_aa4 kernel launch");
}

/* ---- 20-state sumtable on the same kernel ------------------------------------------------------
 * sum[n,k,j] = (sum_i P[n,k,i] pi_i Vinv[i][j]) * (sum_i V[j][i] C[n,k,i])   (core_derivatives.c:116-640)
 * has the shape of one CLV update: two [sites x 20] x [20 x 20] contractions and an element-wise
 * product.  So the "P-matrices" are replaced by fragment images of  L[j][i] = pi_i Vinv[i][j]  and
 * R[j][i] = V[j][i]  (rebuilt only when the eigen data of a category changes), the parent buffer is the
 * sumtable, and there is no scaling.  The generic kernel took 1.56 ms for 200 k sites (1.2 TB/s). */
__global__ void sumfrag_kernel(double * __restrict__ frag, const double * __restrict__ eigen,
                               unsigned long long eigen_stride, unsigned int K, PmatIdx idx,
                               const unsigned long long * __restrict__ tipmap)
{
  const unsigned int k = blockIdx.x;
  const double * blk = eigen + (unsigned long long)idx.v[k] * eigen_stride;
  const double * ev = blk, * iev = blk + 400, * fr = blk + 800 + 20;
  /* third block: what a tip code contributes through L, [code 24][k][20] like the P-matrix tip tables of
   * k_pmatrix.cu (states added in order, like the masked sums of core_derivatives.c:300-340) */
  if (tipmap)
    for (unsigned int t = threadIdx.x; t < 480; t += blockDim.x)
    {
      const unsigned int code = t / 20, o = t % 20;
      const unsigned long long mask = tipmap[code];
      double v = 0.0;
      for (unsigned int n = 0; n < 20; ++n)
        if ((mask >> n) & 1ull) v += fr[n] * iev[n * 20 + o];
      frag[(size_t)2 * K * AA_FRAG + ((size_t)code * K + k) * 20 + o] = v;
    }
  for (unsigned int t = threadIdx.x; t < 400; t += blockDim.x)
  {
    const unsigned int o = t / 20, n = t % 20; /* output state j, summed state i */
    const unsigned int nt = aa_ntile_of_state(o), ks = aa_kstep_of_state(n), lane = (aa_ncol_of_state(o) << 2) | aa_col_of_state(n);
    const size_t at = (size_t)nt * 160 + ks * 32 + lane;
    frag[(size_t)k * AA_FRAG + at] = fr[n] * iev[n * 20 + o];
    frag[((size_t)K + k) * AA_FRAG + at] = ev[o * 20 + n];
  }
}

bool launch_sumtable_aa(PartCtx * c, bool first_is_tip, unsigned int first, unsigned int second,
                        const unsigned int * params_indices, double * d_sumtable)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats;
  const size_t frag_doubles = (size_t)3 * K * AA_FRAG; /* L, R, tip table of L */
  if (!c->d_sumfrag)
  {
    if (!pllb200_cuda_ok(cudaMalloc(&c->d_sumfrag, frag_doubles * sizeof(double)), "sumtable fragments")) return false;
    if (!pllb200_cuda_ok(cudaMemsetAsync(c->d_sumfrag, 0, frag_doubles * sizeof(double), xstream(c)), "sumtable fragments"))
      return false;
    c->sumfrag_key.clear();
  }
  /* Gamma-4 edges run on the forwarding kernel (one operation: 82 us against 108 us at C3); its tip children are
   * table lookups, so the table of L rides along */
  const bool tip_lookup = c->d_tipmap && p->tipmap && p->maxstates <= 24;
  const bool aa4 = K == 4 && g_opt_aa_variant == 0 && !batch_collects(BK_TRAV_AA) && (!first_is_tip || tip_lookup);
  std::vector<unsigned long long> key(K + 1);
  for (unsigned int k = 0; k < K; ++k)
    key[k] = ((unsigned long long)params_indices[k] << 32) | c->eigen_version[params_indices[k]];
  key[K] = tip_lookup ? c->tipmap_gen + 1 : 0;
  if (key != c->sumfrag_key)
  {
    PmatIdx idx;
    for (unsigned int k = 0; k < K; ++k) idx.v[k] = params_indices[k];
    sumfrag_kernel<<<K, 128, 0, xstream(c)>>>(c->d_sumfrag, c->d_eigen, c->eigen_stride, K, idx,
                                              tip_lookup ? c->d_tipmap : nullptr);
    ++g_kernel_launches;
    if (!pllb200_cuda_ok(cudaGetLastError(), "sumfrag_kernel launch")) return false;
    c->sumfrag_key = key;
  }
  AaOp op;
  memset(&op, 0, sizeof(op));
  if (first_is_tip)
  {
    op.flags |= AOPF_TIP1;
    op.c1_off = (unsigned long long)first * c->tip_stride;
  }
  else
    op.c1_off = (unsigned long long)(first - c->clv_first) * c->clv_stride;
  op.c2_off = (unsigned long long)(second - c->clv_first) * c->clv_stride;
  op.m1_off = 0;
  op.m2_off = K * AA_FRAG;
  const AaOp * d_ops = batch_collects(BK_TRAV_AA) ? (const AaOp *)batch_pool_add(&op, sizeof(op))
                                                   : (const AaOp *)dev_stage_small(c, &op, sizeof(op));
  if (!d_ops) return false;
  AaArgs a;
  a.ops = d_ops;
  a.nops = 1;
  a.sites = c->sites_alloc;
  a.K = K;
  a.clv = c->d_clv;
  a.pbase = d_sumtable;
  a.scaler = c->d_scaler;
  a.frag = c->d_sumfrag;
  a.tiptab = nullptr; /* the "matrices" here are eigenvector images, not P-matrices */
  a.tipchars = c->d_tipchars;
  a.tipmap = c->d_tipmap;
  a.l1_prefetch = 0;
  if (aa4)
  {
    a.tiptab = c->d_sumfrag + (size_t)2 * K * AA_FRAG;
    return launch_aa4_variant<AA4_T, AA4_W, AA4_MINB, AA4_TS>(c, a);
  }
  return launch_aa_variant<2, 4, 3>(c, a);
}

/*
 * k_traversal_dna.cu — the 4-state "site-column" traversal kernel (see k_partials.cu header for the
 * design rationale and the reference functions it replaces: core_partials.c:48-200,202-351,612-764,
 * 1013-1210 and their AVX/AVX2 twins).
 *
 * One launch evaluates a whole operation list.  A warp owns a chunk of 32*R consecutive
 * (site, rate-category) elements and walks the list for it; lane l handles elements
 * e0+l, e0+l+32, ..., all of the same category k = l mod K, so one pass over the two staged
 * P-matrices (16 broadcast LDS.128) serves R elements.  Per element and op: one 256-bit load per
 * inner child (or one byte per tip child, expanded in registers to a 0/1 vector so that tip-tip,
 * tip-inner and inner-inner share one arithmetic path: sum_j M[i][j]*{0,1} is bit-identical to the
 * reference's masked row sums), 32 FP64 FMA/MUL, one 256-bit store.
 *
 * Latency structure (what ncu showed to matter — the kernel is not DRAM-bound because child CLVs
 * are re-read from L1/L2, so dependent-load latency per op is the limiter):
 *   - operations are pre-resolved on the host into 64-byte DevOp records (absolute offsets, flags)
 *     and staged per CTA in shared memory, 256 ops per segment: reading op o+1 ahead costs nothing;
 *   - the two P-matrices of op o+1 are fetched by TMA (cp.async.bulk global->shared, completion on a
 *     per-warp mbarrier) while op o computes: no registers, no LSU instructions, no L1 pollution;
 *   - the tip codes of op o+1 (the only loads that always come from DRAM) are issued during op o;
 *   - full chunks run a predicate-free path; only the last partial chunk takes the guarded path.
 */
#include "internal.h"
#include "devutil.cuh"

#include <algorithm>
#include <type_traits>

struct __align__(16) DevOp
{
  unsigned long long p_off;  /* parent CLV: offset in doubles from the CLV arena */
  unsigned long long c1_off; /* child 1: CLV offset in doubles, or byte offset into tipchars */
  unsigned long long c2_off;
  unsigned long long ps_off; /* scaler offsets in u32 elements from the scaler arena */
  unsigned long long s1_off;
  unsigned long long s2_off;
  unsigned int m1_off;       /* P-matrix offsets in doubles into the interleaved image */
  unsigned int m2_off;
  unsigned int flags;
  unsigned int pad;
};
#define OPF_TIP1 1u
#define OPF_TIP2 2u
#define OPF_PS 4u
#define OPF_S1 8u
#define OPF_S2 16u
#define OPF_FWD1 32u  /* child 1 is the parent of the previous op: take it from registers */
#define OPF_FWD2 64u
#define OPF_FS1 128u  /* child-1 scaler is the previous op's parent scaler */
#define OPF_FS2 256u
#define OPF_STREAM 512u /* the parent is read by the next op only, out of registers: store it with the streaming hint
                           (trav_prefetch bit 2; measured: no change, 29.2-29.3 ms either way) */

struct TravArgs
{
  const DevOp * ops;
  unsigned int nops;
  unsigned int sites; /* sites_alloc */
  double * clv;
  unsigned int * scaler;
  const double * pmat_il;
  const unsigned char * tipchars;
  int l2_prefetch; /* ask L2 for the next op's children that are neither tips nor forwarded in registers */
};

#define TRAV_WARPS 8
#define TRAV_SEG(K) ((K) == 8 ? 128 : 256) /* DevOps staged per CTA at a time (<= 16 KB) */

/* 4-bit tip code -> 0.0/1.0 vector (hi word of 1.0 is 0x3FF00000) */
__device__ __forceinline__ d4 tip_vector(unsigned int code)
{
  d4 v;
  v.x = __hiloint2double((int)((code & 1u) * 0x3FF00000u), 0);
  v.y = __hiloint2double((int)(((code >> 1) & 1u) * 0x3FF00000u), 0);
  v.z = __hiloint2double((int)(((code >> 2) & 1u) * 0x3FF00000u), 0);
  v.w = __hiloint2double((int)(((code >> 3) & 1u) * 0x3FF00000u), 0);
  return v;
}

template <int K>
struct TravSmem
{
  DevOp ops[TRAV_SEG(K)];
  double mat[TRAV_WARPS][2][2][16 * K]; /* [warp][stage][L/R][row i][cat k][col j] */
  unsigned long long bar[TRAV_WARPS][2];
};

/* signed hi word: x < 2^-256  <=>  hi(x) < 0x2FF00000 (lo word of 2^-256 is 0) for every non-NaN x, negative
 * round-off values included, as the reference's `x < PLL_SCALE_THRESHOLD` (core_partials.c:729-763) */
__device__ __forceinline__ int hi_word(double x) { return __double2hiint(x); }
#define SCALE_THRESHOLD_HI 0x2FF00000

/* arithmetic of one op for the R elements of a lane; T1/T2: child is a pattern tip */
template <int K, int R, bool T1, bool T2, bool FULL>
__device__ __forceinline__ void op_body(const double * __restrict__ p1, const double * __restrict__ p2,
                                        unsigned long long t1, unsigned long long t2, const double * sL,
                                        const double * sR, int k, unsigned int actmask,
                                        unsigned long long * bar, unsigned int parity,
                                        const d4 (&fw)[R], unsigned int flags, double (&out)[R][4])
{
  constexpr int ESTEP = K * 4;
  d4 v1[R], v2[R];
  /* register forwarding: with the depth-first schedule the op right before a parent is (almost
   * always) one of its inner children, so that operand never makes the round trip through L1/L2 */
  const bool f1 = !T1 && (flags & OPF_FWD1), f2 = !T2 && (flags & OPF_FWD2);
#pragma unroll
  for (int r = 0; r < R; ++r)
  {
    if (FULL || ((actmask >> r) & 1u))
    {
      if (!T1)
      {
        if (f1)
          v1[r] = fw[r];
        else
          v1[r] = ldg256(p1 + ESTEP * r);
      }
      if (!T2)
      {
        if (f2)
          v2[r] = fw[r];
        else
          v2[r] = ldg256(p2 + ESTEP * r);
      }
    }
    else
      v1[r] = v2[r] = d4{0.0, 0.0, 0.0, 0.0};
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
  {
    if (T1) v1[r] = tip_vector((unsigned int)(t1 >> (8 * r)));
    if (T2) v2[r] = tip_vector((unsigned int)(t2 >> (8 * r)));
  }
  mbar_wait(bar, parity); /* matrices of this op have landed */
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    const double2 la = *reinterpret_cast<const double2 *>(sL + (i * K + k) * 4);
    const double2 lb = *reinterpret_cast<const double2 *>(sL + (i * K + k) * 4 + 2);
    const double2 ra = *reinterpret_cast<const double2 *>(sR + (i * K + k) * 4);
    const double2 rb = *reinterpret_cast<const double2 *>(sR + (i * K + k) * 4 + 2);
#pragma unroll
    for (int r = 0; r < R; ++r)
    {
      const double xl = la.x * v1[r].x + la.y * v1[r].y + lb.x * v1[r].z + lb.y * v1[r].w;
      const double xr = ra.x * v2[r].x + ra.y * v2[r].y + rb.x * v2[r].z + rb.y * v2[r].w;
      out[r][i] = xl * xr;
    }
  }
}

/* One chunk (32*R elements = 32*R/K consecutive sites) through ops [0, nseg) of the staged segment.
 * Lane l owns category k = l mod K of the R CONSECUTIVE sites s0 .. s0+R-1, s0 = chunk_site0 + (l/K)*R:
 * a warp-level 256-bit access still covers whole 128-byte lines (K lanes x 32 B per site), the R tip
 * codes of a lane are one packed load, and its R per-site scalers are contiguous.
 * FULL: every site of the chunk exists (no guards).  ph: mbarrier phase bits of the two stages. */
template <int K, int R, bool RATE_SCALERS, bool FULL>
__device__ __forceinline__ void run_segment(const TravArgs & a, TravSmem<K> & sm, unsigned int nseg,
                                            unsigned int s0, int lane, int warp,
                                            unsigned int & ph)
{
  const int k = lane % K;
  constexpr unsigned int MAT_BYTES = 16 * K * sizeof(double);
  constexpr int ESTEP = K * 4;                               /* doubles between consecutive sites */
  const unsigned long long eoff = ((unsigned long long)s0 * K + k) * 4; /* doubles */
  const unsigned long long sidx0 = RATE_SCALERS ? (unsigned long long)s0 * K + k : s0; /* scaler element of r = 0 */
  constexpr int SC_STEP = RATE_SCALERS ? K : 1;
  const bool sc_lane = RATE_SCALERS || k == 0;
  unsigned int actmask = 0;
#pragma unroll
  for (int r = 0; r < R; ++r)
    if (FULL || s0 + r < a.sites) actmask |= 1u << r;

  /* R tip codes packed little-endian into one word (R = 4: one aligned 32-bit load, R = 8: one 64-bit load) */
  using tip_t = typename std::conditional<(R > 4), unsigned long long, unsigned int>::type;
  auto load_tips = [&](const unsigned char * base) -> tip_t {
    if (FULL && R == 8) return (tip_t)__ldg(reinterpret_cast<const unsigned long long *>(base));
    if (FULL && R == 4) return __ldg(reinterpret_cast<const unsigned int *>(base));
    if (FULL && R == 2) return __ldg(reinterpret_cast<const unsigned short *>(base));
    tip_t t = 0;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if ((actmask >> r) & 1u) t |= (tip_t)__ldg(base + r) << (8 * r);
    return t;
  };
  /* R scaler counts of one child (per-site mode with R = 4: one 128-bit load) */
  auto add_scalers = [&](const unsigned int * base, unsigned int (&sc)[R]) {
    if (FULL && (R == 4 || R == 8) && !RATE_SCALERS)
    {
#pragma unroll
      for (int h = 0; h < R / 4; ++h)
      {
        const uint4 q = *reinterpret_cast<const uint4 *>(base + 4 * h);
        sc[4 * h + 0] += q.x;
        sc[4 * h + 1] += q.y;
        sc[4 * h + 2] += q.z;
        sc[4 * h + 3] += q.w;
      }
    }
    else
    {
#pragma unroll
      for (int r = 0; r < R; ++r)
        if (FULL || ((actmask >> r) & 1u)) sc[r] += base[SC_STEP * r];
    }
  };

  /* prologue: matrices + tip codes of op 0 */
  __syncwarp();
  if (lane == 0)
  {
    mbar_expect_tx(&sm.bar[warp][0], 2 * MAT_BYTES);
    tma_load_1d(sm.mat[warp][0][0], a.pmat_il + sm.ops[0].m1_off, MAT_BYTES, &sm.bar[warp][0]);
    tma_load_1d(sm.mat[warp][0][1], a.pmat_il + sm.ops[0].m2_off, MAT_BYTES, &sm.bar[warp][0]);
  }
  d4 fw[R];                /* parent CLV elements of the previous op (register forwarding) */
  unsigned int fsc[R];     /* ... and its scaler counts */
#pragma unroll
  for (int r = 0; r < R; ++r)
  {
    fw[r] = d4{0.0, 0.0, 0.0, 0.0};
    fsc[r] = 0;
  }
  tip_t t1 = 0, t2 = 0; /* packed tip codes of the CURRENT op */
  {
    const unsigned int f0 = sm.ops[0].flags;
    if (f0 & OPF_TIP1) t1 = load_tips(a.tipchars + sm.ops[0].c1_off + s0);
    if (f0 & OPF_TIP2) t2 = load_tips(a.tipchars + sm.ops[0].c2_off + s0);
  }

  for (unsigned int o = 0; o < nseg; ++o)
  {
    const DevOp & d = sm.ops[o];
    const unsigned int flags = d.flags;
    const unsigned int st = o & 1u;

    /* all lanes are done with stage st^1 (op o-1): refill it with the matrices of op o+1 */
    __syncwarp();
    if (lane == 0 && o + 1 < nseg)
    {
      mbar_expect_tx(&sm.bar[warp][st ^ 1u], 2 * MAT_BYTES);
      tma_load_1d(sm.mat[warp][st ^ 1u][0], a.pmat_il + sm.ops[o + 1].m1_off, MAT_BYTES, &sm.bar[warp][st ^ 1u]);
      tma_load_1d(sm.mat[warp][st ^ 1u][1], a.pmat_il + sm.ops[o + 1].m2_off, MAT_BYTES, &sm.bar[warp][st ^ 1u]);
    }
    /* tip codes of the NEXT op (always a DRAM access: issued a whole op early) */
    tip_t n1 = 0, n2 = 0;
    if (o + 1 < nseg)
    {
      const unsigned int fn = sm.ops[o + 1].flags;
      if (fn & OPF_TIP1) n1 = load_tips(a.tipchars + sm.ops[o + 1].c1_off + s0);
      if (fn & OPF_TIP2) n2 = load_tips(a.tipchars + sm.ops[o + 1].c2_off + s0);
    }
    /* tip codes are the one operand that always comes from DRAM, and under the kernel's write load a DRAM read
     * takes about as long as a whole operation: ask L2 for the codes of op o + (l2_prefetch >> 4) now, so that the
     * register load one op ahead (above) finds them there.  32*R/K sites = one 32-byte sector (R = 4) per child. */
    {
      const unsigned int tipd = (unsigned int)a.l2_prefetch >> 4;
      if (tipd && o + tipd < nseg && lane < 2)
      {
        const DevOp & q = sm.ops[o + tipd];
        const unsigned int cs0 = s0 - (unsigned int)(lane / K) * R;
        if (lane == 0 && (q.flags & OPF_TIP1)) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.tipchars + q.c1_off + cs0));
        if (lane == 1 && (q.flags & OPF_TIP2)) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.tipchars + q.c2_off + cs0));
      }
    }
    /* inner children of the NEXT op that will not come out of registers: one 128-byte line per lane
     * covers the warp's 4 KB of each child.  Costs no register; turns the DRAM miss that op o+1 would
     * wait for (ncu: 15 % of the stall samples sit on the first use of those loads) into an L2 hit. */
    if (FULL && R == 4 && K == 4 && (a.l2_prefetch & 3) && o + 1 < nseg)
    {
      const DevOp & nx = sm.ops[o + 1];
      const unsigned long long woff = (unsigned long long)(s0 - (unsigned int)(lane / K) * R) * ESTEP + lane * 16;
      /* mode 2: into L1 — the child was written by this warp (L2 holds it), and a CLV is written once per launch,
       * so the line cannot go stale before its single reader uses it */
      if ((a.l2_prefetch & 3) == 2)
      {
        if (!(nx.flags & (OPF_TIP1 | OPF_FWD1)))
          asm volatile("prefetch.global.L1 [%0];" ::"l"(a.clv + nx.c1_off + woff));
        if (!(nx.flags & (OPF_TIP2 | OPF_FWD2)))
          asm volatile("prefetch.global.L1 [%0];" ::"l"(a.clv + nx.c2_off + woff));
      }
      else
      {
        if (!(nx.flags & (OPF_TIP1 | OPF_FWD1)))
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a.clv + nx.c1_off + woff));
        if (!(nx.flags & (OPF_TIP2 | OPF_FWD2)))
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a.clv + nx.c2_off + woff));
      }
    }
    /* child scaler counts (the lanes that own a scaler entry only): the loads are issued here, the
     * counts are only combined when the parent scaler is written, after the arithmetic (ncu: adding
     * them right away exposed the full L2 latency, 22 % of all stall samples) */
    unsigned int q1[R], q2[R];
#pragma unroll
    for (int r = 0; r < R; ++r) q1[r] = q2[r] = 0;
    if ((flags & OPF_PS) && sc_lane)
    {
      if (flags & OPF_FS1)
      {
#pragma unroll
        for (int r = 0; r < R; ++r) q1[r] = fsc[r];
      }
      else if (flags & OPF_S1)
        add_scalers(a.scaler + d.s1_off + sidx0, q1);
      if (flags & OPF_FS2)
      {
#pragma unroll
        for (int r = 0; r < R; ++r) q2[r] = fsc[r];
      }
      else if (flags & OPF_S2)
        add_scalers(a.scaler + d.s2_off + sidx0, q2);
    }

    /* ---- parent[i] = (L_k[i,:] . v1) * (R_k[i,:] . v2), one specialised body per tip pattern ---- */
    const double * p1 = a.clv + d.c1_off + eoff;
    const double * p2 = a.clv + d.c2_off + eoff;
    const double * sL = sm.mat[warp][st][0];
    const double * sR = sm.mat[warp][st][1];
    unsigned long long * bar = &sm.bar[warp][st];
    const unsigned int parity = (ph >> st) & 1u;
    ph ^= 1u << st;
    double out[R][4];
    switch (flags & (OPF_TIP1 | OPF_TIP2))
    {
      case 0: op_body<K, R, false, false, FULL>(p1, p2, t1, t2, sL, sR, k, actmask, bar, parity, fw, flags, out); break;
      case OPF_TIP1: op_body<K, R, true, false, FULL>(p1, p2, t1, t2, sL, sR, k, actmask, bar, parity, fw, flags, out); break;
      case OPF_TIP2: op_body<K, R, false, true, FULL>(p1, p2, t1, t2, sL, sR, k, actmask, bar, parity, fw, flags, out); break;
      default: op_body<K, R, true, true, FULL>(p1, p2, t1, t2, sL, sR, k, actmask, bar, parity, fw, flags, out); break;
    }

    /* ---- underflow scaling vote + scaler update ---- */
    unsigned int scalemask = 0;
    unsigned int sc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) sc[r] = q1[r] + q2[r];
    if (flags & OPF_PS)
    {
      if ((flags & (OPF_TIP1 | OPF_TIP2)) != (OPF_TIP1 | OPF_TIP2))
      {
#pragma unroll
        for (int r = 0; r < R; ++r)
        {
          const int m = max(max(hi_word(out[r][0]), hi_word(out[r][1])),
                                     max(hi_word(out[r][2]), hi_word(out[r][3])));
          if (m < SCALE_THRESHOLD_HI) scalemask |= 1u << r;
        }
        if (!RATE_SCALERS)
        {
#pragma unroll
          for (int off = 1; off < K; off <<= 1) scalemask &= __shfl_xor_sync(0xffffffffu, scalemask, off);
        }
      }
      if (sc_lane)
      {
        unsigned int * ps = a.scaler + d.ps_off + sidx0;
        if (FULL && (R == 4 || R == 8) && !RATE_SCALERS)
        {
#pragma unroll
          for (int h = 0; h < R / 4; ++h)
            *reinterpret_cast<uint4 *>(ps + 4 * h) =
                make_uint4(sc[4 * h] + ((scalemask >> (4 * h)) & 1u), sc[4 * h + 1] + ((scalemask >> (4 * h + 1)) & 1u),
                           sc[4 * h + 2] + ((scalemask >> (4 * h + 2)) & 1u), sc[4 * h + 3] + ((scalemask >> (4 * h + 3)) & 1u));
        }
        else
        {
#pragma unroll
          for (int r = 0; r < R; ++r)
            if (FULL || ((actmask >> r) & 1u)) ps[SC_STEP * r] = sc[r] + ((scalemask >> r) & 1u);
        }
      }
      if (scalemask)
      {
#pragma unroll
        for (int r = 0; r < R; ++r)
          if ((scalemask >> r) & 1u)
          {
            out[r][0] *= SCALE_FACTOR_D;
            out[r][1] *= SCALE_FACTOR_D;
            out[r][2] *= SCALE_FACTOR_D;
            out[r][3] *= SCALE_FACTOR_D;
          }
      }
    }

    /* ---- store ---- */
    double * pdst = a.clv + d.p_off + eoff;
#pragma unroll
    for (int r = 0; r < R; ++r)
    {
      fw[r] = d4{out[r][0], out[r][1], out[r][2], out[r][3]};
      if (FULL || ((actmask >> r) & 1u))
      {
        if (flags & OPF_STREAM)
          stg256_cs(pdst + ESTEP * r, fw[r]);
        else
          stg256(pdst + ESTEP * r, fw[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) fsc[r] = (flags & OPF_PS) ? sc[r] + ((scalemask >> r) & 1u) : 0u;
    t1 = n1;
    t2 = n2;
  }
}

/* the last, partial chunk: same code with guards, kept out of line so that it does not inflate the
 * register allocation of the hot path.  MINB is a template parameter on purpose: one callee per kernel,
 * so that ptxas sizes it under that kernel's register cap. */
template <int K, int R, bool RATE_SCALERS, int MINB, int TAG>
__device__ __noinline__ void run_segment_tail(const TravArgs & a, TravSmem<K> & sm, unsigned int nseg,
                                              unsigned int s0, int lane, int warp,
                                              unsigned int & ph)
{
  run_segment<K, R, RATE_SCALERS, false>(a, sm, nseg, s0, lane, warp, ph);
}

template <int K, int R, bool RATE_SCALERS, int MINB, int TAG> /* TAG: one out-of-line tail per kernel */
__device__ __forceinline__ void trav_dna_body(const TravArgs & a)
{
  __shared__ __align__(128) TravSmem<K> sm;

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  constexpr unsigned long long SPC = 32ull * R / K; /* sites per chunk */
  const unsigned long long nchunks = ((unsigned long long)a.sites + SPC - 1) / SPC;
  const unsigned long long gwarp = (unsigned long long)blockIdx.x * TRAV_WARPS + warp;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * TRAV_WARPS;

  if (lane == 0)
  {
    mbar_init(&sm.bar[warp][0], 1);
    mbar_init(&sm.bar[warp][1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  unsigned int ph = 0;

  for (unsigned int seg = 0; seg < a.nops; seg += TRAV_SEG(K))
  {
    const unsigned int nseg = min(a.nops - seg, (unsigned int)TRAV_SEG(K));
    __syncthreads(); /* previous segment fully consumed by every warp of the CTA */
    {
      const uint4 * src = reinterpret_cast<const uint4 *>(a.ops + seg);
      uint4 * dst = reinterpret_cast<uint4 *>(sm.ops);
      for (unsigned int i = threadIdx.x; i < nseg * 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    for (unsigned long long chunk = gwarp; chunk < nchunks; chunk += nwarps)
    {
      const unsigned int s0 = (unsigned int)(chunk * SPC) + (unsigned int)(lane / K) * R;
      if ((chunk + 1) * SPC <= a.sites)
        run_segment<K, R, RATE_SCALERS, true>(a, sm, nseg, s0, lane, warp, ph);
      else
        run_segment_tail<K, R, RATE_SCALERS, MINB, TAG>(a, sm, nseg, s0, lane, warp, ph);
    }
  }
}

template <int K, int R, bool RATE_SCALERS, int MINB>
__global__ void __launch_bounds__(TRAV_WARPS * 32, MINB) trav_dna_kernel(TravArgs a)
{
  trav_dna_body<K, R, RATE_SCALERS, MINB, 0>(a);
}

/* ---- host ------------------------------------------------------------------------------------ */

int g_opt_trav_variant = 0;
int g_opt_trav_short = 3; /* lists of at most this many operations run on the kernel below: they are pure streams of
                             loads, and the stash kernel leaves only ~24 KB of L1 beside its shared memory — too little
                             to keep a stream in flight (measured: 102 vs 60 us per single operation at 1 M sites) */
int g_opt_trav_forward = 1;
int g_opt_trav_prefetch = 0; /* measured: 28.3 ms with, 27.9 ms without at C2 — the kernel is not waiting on DRAM latency */

template <int K, int R, bool RS, int MINB>
static bool launch_variant(PartCtx * c, const TravArgs & a)
{
  static int per_sm_cache = 0;
  if (!per_sm_cache)
  {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, trav_dna_kernel<K, R, RS, MINB>,
                                                      TRAV_WARPS * 32, 0) != cudaSuccess || nb < 1)
      nb = MINB;
    per_sm_cache = nb;
  }
  int per_sm = per_sm_cache;
  if (g_opt_trav_ctas_per_sm > 0) per_sm = std::min(per_sm, g_opt_trav_ctas_per_sm);
  const unsigned long long spc = 32ull * R / K;
  const unsigned long long nchunks = ((unsigned long long)a.sites + spc - 1) / spc;
  unsigned long long blocks = (nchunks + TRAV_WARPS - 1) / TRAV_WARPS;
  const unsigned long long cap = (unsigned long long)c->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  trav_dna_kernel<K, R, RS, MINB><<<(unsigned int)blocks, TRAV_WARPS * 32, 0, xstream(c)>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "trav_dna_kernel launch");
}

template <int K>
static bool launch_k(PartCtx * c, const TravArgs & a, bool rate_scalers)
{
  if (rate_scalers) return launch_variant<K, 2, true, 3>(c, a);
  if (K == 4)
  {
    switch (g_opt_trav_variant)
    {
      case 1: return launch_variant<K, 1, false, 4>(c, a);
      case 2: return launch_variant<K, 2, false, 3>(c, a);
      case 3: return launch_variant<K, 4, false, 2>(c, a);
      case 4: return launch_variant<K, 2, false, 2>(c, a);
      case 5: return launch_variant<K, 8, false, 1>(c, a);
      default: break;
    }
    return launch_variant<K, 4, false, 2>(c, a);
  }
  return launch_variant<K, 2, false, 3>(c, a);
}

bool launch_trav_dna(PartCtx * c, const pll_operation_t * ops, unsigned int count)
{
  pll_partition_t * p = &c->pub;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  const bool rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;
  const unsigned int K = p->rate_cats;

  /* Gamma-4 with per-site scalers: the forwarding + stash kernel (variant 0, the default); the kernel
   * below stays for other category counts, per-rate scalers and as the bit-for-bit cross-check */
  if (g_opt_trav_variant == 0 && lean_supported(c) && (count > (unsigned int)g_opt_trav_short || g_batch))
    return launch_trav_dna4(c, ops, count);

  static_assert(sizeof(DevOp) == 64, "DevOp is 8 x u64");
  c->devop_scratch.resize((size_t)count * 8);
  DevOp * dops = reinterpret_cast<DevOp *>(c->devop_scratch.data());
  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops[i];
    DevOp & d = dops[i];
    memset(&d, 0, sizeof(d));
    d.p_off = (unsigned long long)(o.parent_clv_index - c->clv_first) * c->clv_stride;
    if (pattern_tip && o.child1_clv_index < p->tips)
    {
      d.flags |= OPF_TIP1;
      d.c1_off = (unsigned long long)o.child1_clv_index * c->tip_stride;
    }
    else
      d.c1_off = (unsigned long long)(o.child1_clv_index - c->clv_first) * c->clv_stride;
    if (pattern_tip && o.child2_clv_index < p->tips)
    {
      d.flags |= OPF_TIP2;
      d.c2_off = (unsigned long long)o.child2_clv_index * c->tip_stride;
    }
    else
      d.c2_off = (unsigned long long)(o.child2_clv_index - c->clv_first) * c->clv_stride;
    if (o.parent_scaler_index >= 0)
    {
      d.flags |= OPF_PS;
      d.ps_off = (unsigned long long)o.parent_scaler_index * c->scaler_stride;
    }
    if (o.child1_scaler_index >= 0)
    {
      d.flags |= OPF_S1;
      d.s1_off = (unsigned long long)o.child1_scaler_index * c->scaler_stride;
    }
    if (o.child2_scaler_index >= 0)
    {
      d.flags |= OPF_S2;
      d.s2_off = (unsigned long long)o.child2_scaler_index * c->scaler_stride;
    }
    d.m1_off = o.child1_matrix_index * 16 * K;
    d.m2_off = o.child2_matrix_index * 16 * K;
    /* forwarding from the previous op of the same staged segment (segments restart per chunk) */
    const unsigned int seg_len = K == 8 ? 128 : 256;
    if (g_opt_trav_forward && i % seg_len != 0)
    {
      const pll_operation_t & q = ops[i - 1];
      if (!(d.flags & OPF_TIP1) && o.child1_clv_index == q.parent_clv_index) d.flags |= OPF_FWD1;
      if (!(d.flags & OPF_TIP2) && o.child2_clv_index == q.parent_clv_index &&
          !(d.flags & OPF_FWD1))
        d.flags |= OPF_FWD2;
      if (q.parent_scaler_index >= 0 && (d.flags & OPF_PS))
      {
        if (o.child1_scaler_index == q.parent_scaler_index) d.flags |= OPF_FS1;
        if (o.child2_scaler_index == q.parent_scaler_index) d.flags |= OPF_FS2;
      }
      if ((d.flags & (OPF_FWD1 | OPF_FWD2)) && (g_opt_trav_prefetch & 4)) dops[i - 1].flags |= OPF_STREAM;
    }
  }
  const DevOp * d_ops = (const DevOp *)dev_stage_small(c, dops, sizeof(DevOp) * count);
  if (!d_ops) return false;
  TravArgs a;
  a.ops = d_ops;
  a.nops = count;
  a.sites = c->sites_alloc;
  a.clv = c->d_clv;
  a.scaler = c->d_scaler;
  a.pmat_il = c->d_pmat_il;
  a.tipchars = c->d_tipchars;
  a.l2_prefetch = g_opt_trav_prefetch;
  switch (K)
  {
    case 1: return launch_k<1>(c, a, rate_scalers);
    case 2: return launch_k<2>(c, a, rate_scalers);
    case 4: return launch_k<4>(c, a, rate_scalers);
    default: return launch_k<8>(c, a, rate_scalers);
  }
}

/* batch mode collects Gamma-4 partitions on the lean kernel only (k_traversal_dna4.cu) */
bool flush_trav_dna4(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  return flush_trav_dna4_lean(lead, d_args, k);
}

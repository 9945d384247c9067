/*
 * k_traversal_dna4.cu — the 4-state, Gamma-4, per-site-scaler traversal kernel ("lean" site-column kernel).
 *
 * Same contract as k_traversal_dna.cu (pll_update_partials, partials.c:237-291; core kernels
 * core_partials.c:48-200,202-351,612-764,1013-1210): one launch evaluates a whole operation list, a warp
 * owns 32 consecutive sites (lane = (group of 4 consecutive sites, rate category)) and walks the list
 * for them.  What differs is where operands come from — the round-1 kernel was bound by the latency
 * chain of an operation (ncu: issue slots 50 % busy at 16 warps/SM, long_scoreboard 32 %), not by DRAM:
 *
 *   - NO inner child is read back from L1/L2 in the common case.  The host orders the list Sethi-Ullman
 *     style (trav_schedule.cpp) and tells every operation where each child lives: in the registers that
 *     still hold the previous operation's parent, or in one of two per-warp shared-memory stash slots
 *     (4 KB each) that its producer filled next to the global store.  Subtrees of Strahler number <= 3
 *     (all but a handful of operations of a 1000-taxon tree) never load a CLV; the rest, and children
 *     computed by earlier launches, come from global memory as before.  Scaler counts travel the same way.
 *   - tip children never touch the FP64 pipe: the P-matrix kernel also writes, per matrix, the five
 *     column sums a DNA tip code can select ([A, C, G, T, N][cat][row], 640 B), the tip arena has a
 *     twin that holds the column number of every code, and a tip term is two 128-bit shared-memory reads
 *     (rare ambiguity codes add their columns in state order, which is what the masked row sum of
 *     core_partials.c:290-351 does).  The table of a tip child takes the place of its matrix in the
 *     TMA-staged ring.
 *   - operations are 32-byte records of ROW indices; every address is one IMAD.WIDE from a per-lane base.
 *   - the accumulator is updated in place: acc = R x child2 (child2 is loaded INTO acc, or is already there
 *     when forwarded), then acc *= L x child1: no copy between operations, ~100 live registers.
 *   - the scaling vote leaves early when no lane of the warp saw an element below 2^-256.
 * Arithmetic per element is the same expression as the round-1 kernel (bit-identical results).
 *
 * HBM traffic per site and operation: 128 B CLV + 4 B scaler written, 0-1 B tip codes read; everything
 * else is on chip.  Bound: DRAM writes.
 */
#include "internal.h"
#include "devutil.cuh"
#include "trav_schedule.h"

#include <algorithm>
#include <atomic>

struct LeanArgs
{
  const LeanOp * ops; /* MUST stay first (batch pool token, device_ctx.cu) */
  unsigned int nops;
  unsigned int sites; /* sites_alloc */
  unsigned int seg;   /* operations per staged segment */
  unsigned int tip_stride;
  char * clv;
  char * scaler;
  const double * pmat_il;   /* [m][row i][cat k][col j] : 512 B per matrix */
  const double * pmat_cols; /* [m][column c 0..4][cat k][row i] : 640 B per matrix */
  const unsigned char * tipcols;
  unsigned int clv_stride_b, scaler_stride_b; /* row strides in bytes (< 4 GB: lean_supported) */
  /* regraft scoring (lean_op_special) */
  double * score_out;        /* [candidate][score_stride] per-warp partial sums */
  const unsigned int * weights;
  unsigned int score_stride;
  double wpi[16];            /* [category][state]: rate weight x frequency */
  unsigned int dbg; /* timing ablations only (option trav_debug): 1 no tip loads, 2 no CLV stores, 4 no scaler stores, 8 plain chunk stride in the last round */
};

#define L4_WARPS 8
#ifndef L4_TIP_AHEAD
/* tip codes are requested this many operations before their use (1 or 2).  2 was measured: the two extra live
 * registers make ptxas park an in-flight load result in local memory (STL waits for the load): 32.1 vs 24.9 ms at C2 */
#define L4_TIP_AHEAD 1
#endif
#define L4_SEG_MAX 512
#define L4_STASH_B 4096u
#define L4_WARP_SM (2 * L4_STASH_B + 2 * 128 + 2 * 2 * 640 + 16 + 112) /* stash | scaler stash | matrices | barriers | pad -> 11136 */
#define L4_OFF_SST (2 * L4_STASH_B)
#define L4_OFF_MAT (L4_OFF_SST + 256)
#define L4_OFF_BAR (L4_OFF_MAT + 2560)

static_assert(L4_WARP_SM % 128 == 0, "per-warp shared memory block keeps 128-byte alignment");

__device__ __forceinline__ double2 lds_d2(unsigned int addr)
{
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_d2(unsigned int addr, double x, double y)
{
  asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(addr), "d"(x), "d"(y) : "memory");
}
__device__ __forceinline__ uint4 lds_u4(unsigned int addr)
{
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds_u2(unsigned int addr)
{
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ unsigned int lds_u1(unsigned int addr)
{
  unsigned int v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u4(unsigned int addr, const uint4 & v)
{
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cp_async16(unsigned int dst, const void * src)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait_u32(unsigned int bar, unsigned int parity)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_u32(unsigned int dst, const void * src, unsigned int bytes, unsigned int bar)
{
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void mbar_expect_u32(unsigned int bar, unsigned int bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

/* signed compare on the high word: x < 2^-256 for every non-NaN x, negative values (round-off in a
 * P-matrix) included, as the reference's `x < PLL_SCALE_THRESHOLD` (core_partials.c:729-763) */
#define L4_THRESHOLD_HI 0x2FF00000
__device__ __forceinline__ int hi_s(double x) { return __double2hiint(x); }

/* term of a tip child for one element.  `b` is the tip's column code (tipcols_encode_kernel): 16-byte offset of
 * the code's column in the staged table, or — ambiguity codes other than N — 0x80 | code: the sum of the columns
 * of the code's bits in state order.
 * Bank conflicts: every column starts at the same bank, and a 16-byte shared-memory load is served a quarter warp
 * (two site groups) at a time, so two groups with different codes collided on the banks of their category.  The
 * odd columns (C, T) therefore hold their two halves in the opposite order (pmatrix_dna_cols) and the code's bit 0
 * says so: rows 0-1 of an even column sit in banks 0-3 (+ 8 k), of an odd column in banks 4-7 — two groups only
 * collide when their codes differ but have the same parity */
template <bool SLOW>
__device__ __forceinline__ void tip_term(unsigned int tab_k, unsigned int b, double (&t)[4])
{
  if (SLOW && (b & 0x80u))
  {
    t[0] = t[1] = t[2] = t[3] = 0.0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if ((b >> j) & 1u)
      {
        const double2 a = lds_d2(tab_k + j * 128 + (j & 1) * 16), c = lds_d2(tab_k + j * 128 + 16 - (j & 1) * 16);
        t[0] += a.x;
        t[1] += a.y;
        t[2] += c.x;
        t[3] += c.y;
      }
    return;
  }
  const unsigned int p = tab_k + b * 16;
  const double2 a = lds_d2(p), c = lds_d2(p ^ 16u); /* tab_k is a multiple of 32 */
  t[0] = a.x;
  t[1] = a.y;
  t[2] = c.x;
  t[3] = c.y;
}

/* acc[r] <- M_k x acc[r] for the 4 elements of the lane; M staged as [row i][cat k][col j] */
__device__ __forceinline__ void transform_inplace(unsigned int mat_k, double (&acc)[4][4])
{
  double2 ma[4], mb[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    ma[i] = lds_d2(mat_k + i * 128);
    mb[i] = lds_d2(mat_k + i * 128 + 16);
  }
#pragma unroll
  for (int r = 0; r < 4; ++r)
  {
    double t[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      t[i] = ma[i].x * acc[r][0] + ma[i].y * acc[r][1] + mb[i].x * acc[r][2] + mb[i].y * acc[r][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[r][i] = t[i];
  }
}

/* acc[r][i] <- (M_k[i,:] . v[r]) * acc[r][i] */
__device__ __forceinline__ void multiply_by_term(unsigned int mat_k, const double (&v)[4][4], double (&acc)[4][4])
{
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    const double2 la = lds_d2(mat_k + i * 128), lb = lds_d2(mat_k + i * 128 + 16);
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      const double xl = la.x * v[r][0] + la.y * v[r][1] + lb.x * v[r][2] + lb.y * v[r][3];
      acc[r][i] = xl * acc[r][i];
    }
  }
}

/* per-lane constants of one chunk: as few registers as possible, everything else is derived where it is used */
struct LaneCtx
{
  unsigned int s0;      /* first of the lane's 4 consecutive sites */
  unsigned int clv_off; /* byte offset of element (s0, k) inside a CLV: (s0 * 4 + k) * 32 (< 4 GB: lean_supported) */
  unsigned int wsm;     /* this warp's shared-memory block (shared-window address) */
  unsigned int lane, actmask;
  __device__ __forceinline__ unsigned int k() const { return lane & 3u; }
  __device__ __forceinline__ bool sc_lane() const { return (lane & 3u) == 0; }
  __device__ __forceinline__ unsigned int stash_lane() const { return wsm + lane * 16; }
  __device__ __forceinline__ unsigned int sst_lane() const { return wsm + L4_OFF_SST + (lane >> 2) * 16; }
  __device__ __forceinline__ unsigned int mat_lane() const { return wsm + L4_OFF_MAT + lane * 16; }
  __device__ __forceinline__ unsigned int mat_k(unsigned int st) const { return wsm + L4_OFF_MAT + st * 1280 + (lane & 3u) * 32; }
};
__device__ __forceinline__ char * lean_clv_ptr(const LeanArgs & a, const LaneCtx & L, unsigned int row)
{
  return a.clv + ((unsigned long long)row * a.clv_stride_b + L.clv_off);
}
__device__ __forceinline__ unsigned int * lean_scaler_ptr(const LeanArgs & a, const LaneCtx & L, unsigned int row)
{
  return reinterpret_cast<unsigned int *>(a.scaler + ((unsigned long long)row * a.scaler_stride_b + L.s0 * 4u));
}

template <bool FULL>
__device__ __forceinline__ unsigned int lean_load_tips(const LeanArgs & a, const LaneCtx & L, unsigned int tip)
{
  const unsigned char * p = a.tipcols + ((unsigned long long)tip * a.tip_stride + L.s0);
  if (a.dbg & 1u) return 0u;
  if (FULL) return __ldg(reinterpret_cast<const unsigned int *>(p));
  unsigned int t = 0;
#pragma unroll
  for (int r = 0; r < 4; ++r)
    if ((L.actmask >> r) & 1u) t |= (unsigned int)__ldg(p + r) << (8 * r);
  return t;
}

template <bool FULL>
__device__ __forceinline__ uint4 lean_load_scalers(const LeanArgs & a, const LaneCtx & L, unsigned int row)
{
  const unsigned int * p = lean_scaler_ptr(a, L, row);
  if (FULL) return *reinterpret_cast<const uint4 *>(p);
  uint4 q = make_uint4(0, 0, 0, 0);
  if (L.actmask & 1u) q.x = p[0];
  if (L.actmask & 2u) q.y = p[1];
  if (L.actmask & 4u) q.z = p[2];
  if (L.actmask & 8u) q.w = p[3];
  return q;
}

/* Matrices (or tip column tables) and tip codes of the operation whose record sits at `rec`, into stage `st`:
 * 16-byte cp.async by all lanes (512 B = one per lane; the fifth column of a tip table is 128 B more from
 * lanes 0-7), one commit group per call.  Straight-line (selects and predicated copies, no branch), so that
 * it can be scheduled into the arithmetic of the current operation.  [A per-warp TMA bulk copy cost ~80 issue
 * slots per operation: address set-up plus the elect / R2UR sequence around each UBLKCP.] */
template <bool FULL>
__device__ __forceinline__ void lean_prefetch(const LeanArgs & a, const LaneCtx & L, unsigned int rec, unsigned int rec_tip,
                                              unsigned int st, unsigned int & n1, unsigned int & n2)
{
  const uint4 x1 = lds_u4(rec + 16);
  const unsigned int fn = x1.z >> 24;
  const bool t1 = LOPA_SRC1(fn) == SRC_TIP, t2 = LOPA_SRC2(fn) == SRC_TIP;
  const char * cols = reinterpret_cast<const char *>(a.pmat_cols) + L.lane * 16;
  const char * rows = reinterpret_cast<const char *>(a.pmat_il) + L.lane * 16;
  const char * g1 = (t1 ? cols : rows) + (unsigned long long)(x1.z & LOP_MATRIX_MASK) * (t1 ? 640u : 512u);
  const char * g2 = (t2 ? cols : rows) + (unsigned long long)(x1.w & LOP_MATRIX_MASK) * (t2 ? 640u : 512u);
  const unsigned int dst = L.mat_lane() + st * 1280;
  cp_async16(dst, g1);
  cp_async16(dst + 640, g2);
  if (t1 && L.lane < 8) cp_async16(dst + 512, g1 + 512);
  if (t2 && L.lane < 8) cp_async16(dst + 640 + 512, g2 + 512);
  cp_async_commit();
  /* tip codes: of the operation at rec_tip (L4_TIP_AHEAD operations ahead) */
  const uint4 y0 = lds_u4(rec_tip);
  const unsigned int ft = L4_TIP_AHEAD == 1 ? fn : lds_u1(rec_tip + 24) >> 24;
  n1 = 0;
  n2 = 0;
  if (LOPA_SRC1(ft) == SRC_TIP) n1 = lean_load_tips<FULL>(a, L, y0.y);
  if (LOPA_SRC2(ft) == SRC_TIP) n2 = lean_load_tips<FULL>(a, L, y0.z);
}

/* tip column codes in flight: of the current operation (t) and, with L4_TIP_AHEAD == 2, of the next one (u) */
struct TipQ
{
  unsigned int t1, t2, u1, u2;
  __device__ __forceinline__ void advance(unsigned int n1, unsigned int n2)
  {
    if (L4_TIP_AHEAD == 1)
    {
      t1 = n1;
      t2 = n2;
    }
    else
    {
      t1 = u1;
      t2 = u2;
      u1 = n1;
      u2 = n2;
    }
  }
};

/* underflow scaling vote (core_partials.c:729-763; never for tip-tip: :185-186), scaler counts, global store,
 * stash.  sc: sum of the children's counts (sc lanes).  w0 = {p, c1, c2, ps}. */
template <bool FULL, bool TIPTIP>
__device__ __forceinline__ void lean_finish(const LeanArgs & a, const LaneCtx & L, const uint4 & w0, unsigned int fa,
                                            unsigned int fb, double (&acc)[4][4], uint4 sc, uint4 & fsc)
{
  if (fa & LOPA_PS)
  {
    if (!TIPTIP)
    {
      unsigned int scalemask = 0;
#pragma unroll
      for (int r = 0; r < 4; ++r)
      {
        const int m = max(max(hi_s(acc[r][0]), hi_s(acc[r][1])), max(hi_s(acc[r][2]), hi_s(acc[r][3])));
        if (m < L4_THRESHOLD_HI) scalemask |= 1u << r;
      }
      if (__any_sync(0xffffffffu, scalemask != 0u))
      {
        scalemask &= __shfl_xor_sync(0xffffffffu, scalemask, 1);
        scalemask &= __shfl_xor_sync(0xffffffffu, scalemask, 2);
        if (scalemask)
        {
#pragma unroll
          for (int r = 0; r < 4; ++r)
            if ((scalemask >> r) & 1u)
            {
#pragma unroll
              for (int i = 0; i < 4; ++i) acc[r][i] *= SCALE_FACTOR_D;
            }
          sc.x += scalemask & 1u;
          sc.y += (scalemask >> 1) & 1u;
          sc.z += (scalemask >> 2) & 1u;
          sc.w += (scalemask >> 3) & 1u;
        }
      }
    }
    if (L.sc_lane() && !(a.dbg & 4u))
    {
      unsigned int * ps = lean_scaler_ptr(a, L, w0.w);
      if (FULL)
        *reinterpret_cast<uint4 *>(ps) = sc;
      else
      {
        if (L.actmask & 1u) ps[0] = sc.x;
        if (L.actmask & 2u) ps[1] = sc.y;
        if (L.actmask & 4u) ps[2] = sc.z;
        if (L.actmask & 8u) ps[3] = sc.w;
      }
    }
    fsc = sc;
  }
  else
    fsc = make_uint4(0, 0, 0, 0);

  double * pdst = reinterpret_cast<double *>(lean_clv_ptr(a, L, w0.x));
#pragma unroll
  for (int r = 0; r < 4; ++r)
    if ((FULL || ((L.actmask >> r) & 1u)) && !(a.dbg & 2u)) stg256(pdst + 16 * r, d4{acc[r][0], acc[r][1], acc[r][2], acc[r][3]});
  if (fa & LOPA_STASH_W)
  {
    const unsigned int ws = LOPB_WSLOT(fb);
    const unsigned int sl = L.stash_lane() + ws * L4_STASH_B;
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      sts_d2(sl + r * 1024, acc[r][0], acc[r][1]);
      sts_d2(sl + r * 1024 + 512, acc[r][2], acc[r][3]);
    }
    if (L.sc_lane()) sts_u4(L.sst_lane() + ws * 128, fsc);
  }
}

__device__ __forceinline__ void lean_stash_read(unsigned int sl, double (&v)[4][4])
{
#pragma unroll
  for (int r = 0; r < 4; ++r)
  {
    const double2 x = lds_d2(sl + r * 1024), y = lds_d2(sl + r * 1024 + 512);
    v[r][0] = x.x;
    v[r][1] = x.y;
    v[r][2] = y.x;
    v[r][3] = y.y;
  }
}

__device__ __forceinline__ void lean_tip_assign(unsigned int tab, unsigned int t, bool slow, double (&acc)[4][4])
{
#pragma unroll
  for (int r = 0; r < 4; ++r)
  {
    if (slow)
      tip_term<true>(tab, (t >> (8 * r)) & 0xffu, acc[r]);
    else
      tip_term<false>(tab, (t >> (8 * r)) & 0xffu, acc[r]);
  }
}
__device__ __forceinline__ void lean_tip_multiply(unsigned int tab, unsigned int t, bool slow, double (&acc)[4][4])
{
#pragma unroll
  for (int r = 0; r < 4; ++r)
  {
    double x[4];
    if (slow)
      tip_term<true>(tab, (t >> (8 * r)) & 0xffu, x);
    else
      tip_term<false>(tab, (t >> (8 * r)) & 0xffu, x);
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[r][i] = x[i] * acc[r][i];
  }
}

/* One operation whose operand sources are compile-time constants: one basic block from the record read to the
 * scaling vote, so that ptxas overlaps the next operation's prefetch, the scaler traffic and the shared-memory
 * reads with the FP64 chains (the round-2 ncu capture of the all-runtime version: 49 branches per operation,
 * issue slots 34 % busy, a ~2700-cycle dependent chain per operation).  Preconditions checked by the host
 * (LOPB_TYPE != 0): the parent has a scaler row; scaler counts of a child come from where its CLV comes from;
 * full chunk; no ambiguity code other than N in the tip codes (checked here: `slow` runs the generic body). */
template <int SRC1, int SRC2>
__device__ __forceinline__ void lean_op_hot(const LeanArgs & a, const LaneCtx & L, unsigned int rec, unsigned int st,
                                            double (&acc)[4][4], uint4 & fsc, TipQ & q)
{
  const uint4 w0 = lds_u4(rec), w1 = lds_u4(rec + 16);
  const unsigned int fa = w1.z >> 24, fb = w1.w >> 24;
  unsigned int n1, n2;
  lean_prefetch<true>(a, L, rec + 32, rec + 32 * L4_TIP_AHEAD, st ^ 1u, n1, n2);
  const unsigned int t1 = q.t1, t2 = q.t2;

  uint4 q1 = make_uint4(0, 0, 0, 0), q2 = make_uint4(0, 0, 0, 0);
  if (L.sc_lane())
  {
    if (SRC1 == SRC_STASH) q1 = lds_u4(L.sst_lane() + LOPA_SLOT1(fa) * 128);
    if (SRC1 == SRC_GLOBAL) q1 = lean_load_scalers<true>(a, L, w1.x);
    if (SRC2 == SRC_FWD) q2 = fsc;
    if (SRC2 == SRC_STASH) q2 = lds_u4(L.sst_lane() + LOPA_SLOT2(fa) * 128);
    if (SRC2 == SRC_GLOBAL) q2 = lean_load_scalers<true>(a, L, w1.y);
  }
  /* child 1 from global memory: requested now, next to child 2, so that both are in flight together — a
   * short list (branch-length optimisation, a regraft candidate) is a pure stream and needs the bytes in flight */
  double v1[4][4];
  if (SRC1 == SRC_GLOBAL)
  {
    const double * p1 = reinterpret_cast<const double *>(lean_clv_ptr(a, L, w0.y));
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      const d4 v = ldg256(p1 + 16 * r);
      v1[r][0] = v.x;
      v1[r][1] = v.y;
      v1[r][2] = v.z;
      v1[r][3] = v.w;
    }
  }
  if (SRC2 == SRC_GLOBAL)
  {
    const double * p2 = reinterpret_cast<const double *>(lean_clv_ptr(a, L, w0.z));
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      const d4 v = ldg256(p2 + 16 * r);
      acc[r][0] = v.x;
      acc[r][1] = v.y;
      acc[r][2] = v.z;
      acc[r][3] = v.w;
    }
  }
  if (SRC2 == SRC_STASH) lean_stash_read(L.stash_lane() + LOPA_SLOT2(fa) * L4_STASH_B, acc);

  const unsigned int matL = L.mat_k(st), matR = matL + 640;
  cp_async_wait1(); /* this operation's matrices have landed (the next one's may be in flight) ... */
  __syncwarp();     /* ... for every lane's part of them */
  if (SRC2 == SRC_TIP)
    lean_tip_assign(matR, t2, false, acc);
  else
    transform_inplace(matR, acc);
  if (SRC1 == SRC_TIP)
    lean_tip_multiply(matL, t1, false, acc);
  else
  {
    if (SRC1 == SRC_STASH) lean_stash_read(L.stash_lane() + LOPA_SLOT1(fa) * L4_STASH_B, v1);
    multiply_by_term(matL, v1, acc);
  }

  const uint4 sc = make_uint4(q1.x + q2.x, q1.y + q2.y, q1.z + q2.z, q1.w + q2.w);
  lean_finish<true, SRC1 == SRC_TIP && SRC2 == SRC_TIP>(a, L, w0, fa | LOPA_PS, fb, acc, sc, fsc);
  q.advance(n1, n2);
}

/* both child terms of an operation whose sources are read at run time: acc <- (M2 x child2) * (M1 x child1),
 * sc <- sum of the children's scaler counts (valid in the lanes with want_scalers).  one_child: child 2 only. */
template <bool FULL>
__device__ __forceinline__ void lean_children_generic(const LeanArgs & a, const LaneCtx & L, const uint4 & w0,
                                                      const uint4 & w1, unsigned int st, bool want_scalers, bool one_child,
                                                      double (&acc)[4][4], const uint4 & fsc, unsigned int t1,
                                                      unsigned int t2, uint4 & sc)
{
  const unsigned int fa = w1.z >> 24, fb = w1.w >> 24;
  const unsigned int src1 = LOPA_SRC1(fa), src2 = LOPA_SRC2(fa);
  uint4 q1 = make_uint4(0, 0, 0, 0), q2 = make_uint4(0, 0, 0, 0);
  if (want_scalers)
  {
    const unsigned int ss1 = LOPB_SS1(fb), ss2 = LOPB_SS2(fb);
    if (ss1 == SSRC_GLOBAL)
      q1 = lean_load_scalers<FULL>(a, L, w1.x);
    else if (ss1 == SSRC_SAME)
      q1 = lds_u4(L.sst_lane() + LOPA_SLOT1(fa) * 128); /* child 1 is never forwarded */
    if (ss2 == SSRC_GLOBAL)
      q2 = lean_load_scalers<FULL>(a, L, w1.y);
    else if (ss2 == SSRC_SAME)
      q2 = src2 == SRC_FWD ? fsc : lds_u4(L.sst_lane() + LOPA_SLOT2(fa) * 128);
  }
  const unsigned int matL = L.mat_k(st), matR = matL + 640;
  const bool slow = __any_sync(0xffffffffu, ((t1 | t2) & 0x80808080u) != 0u);

  /* ---- acc <- term of child 2 ---- */
  if (src2 == SRC_GLOBAL)
  {
    const double * p2 = reinterpret_cast<const double *>(lean_clv_ptr(a, L, w0.z));
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (FULL || ((L.actmask >> r) & 1u))
      {
        const d4 v = ldg256(p2 + 16 * r);
        acc[r][0] = v.x;
        acc[r][1] = v.y;
        acc[r][2] = v.z;
        acc[r][3] = v.w;
      }
  }
  else if (src2 == SRC_STASH)
    lean_stash_read(L.stash_lane() + LOPA_SLOT2(fa) * L4_STASH_B, acc);
  cp_async_wait1();
  __syncwarp();
  if (src2 == SRC_TIP)
    lean_tip_assign(matR, t2, slow, acc);
  else
    transform_inplace(matR, acc);

  sc = make_uint4(q1.x + q2.x, q1.y + q2.y, q1.z + q2.z, q1.w + q2.w);
  if (one_child) return;

  /* ---- acc *= term of child 1 ---- */
  if (src1 == SRC_TIP)
    lean_tip_multiply(matL, t1, slow, acc);
  else
  {
    double v1[4][4];
    if (src1 == SRC_STASH)
      lean_stash_read(L.stash_lane() + LOPA_SLOT1(fa) * L4_STASH_B, v1);
    else
    {
      const double * p1 = reinterpret_cast<const double *>(lean_clv_ptr(a, L, w0.y));
#pragma unroll
      for (int r = 0; r < 4; ++r)
      {
        d4 v = d4{0.0, 0.0, 0.0, 0.0};
        if (FULL || ((L.actmask >> r) & 1u)) v = ldg256(p1 + 16 * r);
        v1[r][0] = v.x;
        v1[r][1] = v.y;
        v1[r][2] = v.z;
        v1[r][3] = v.w;
      }
    }
    multiply_by_term(matL, v1, acc);
  }
}

/* any operation: sources, scaler sources and flags read at run time */
template <bool FULL>
__device__ __forceinline__ void lean_op_generic(const LeanArgs & a, const LaneCtx & L, unsigned int rec, unsigned int st,
                                                double (&acc)[4][4], uint4 & fsc, TipQ & q)
{
  const uint4 w0 = lds_u4(rec), w1 = lds_u4(rec + 16);
  const unsigned int fa = w1.z >> 24, fb = w1.w >> 24;
  unsigned int n1, n2;
  lean_prefetch<FULL>(a, L, rec + 32, rec + 32 * L4_TIP_AHEAD, st ^ 1u, n1, n2);
  const unsigned int t1 = q.t1, t2 = q.t2;
  uint4 sc;
  lean_children_generic<FULL>(a, L, w0, w1, st, (fa & LOPA_PS) && L.sc_lane(), false, acc, fsc, t1, t2, sc);
  if (LOPA_SRC1(fa) == SRC_TIP && LOPA_SRC2(fa) == SRC_TIP)
    lean_finish<FULL, true>(a, L, w0, fa, fb, acc, sc, fsc);
  else
    lean_finish<FULL, false>(a, L, w0, fa, fb, acc, sc, fsc);
  q.advance(n1, n2);
}

/* The two operations of a batched regraft scoring list (algo_search.cpp, pllmod_algo_spr_round: every candidate
 * branch of one pruned subtree in ONE launch; reference loop: src/algorithm/algo_search.c:727-860).
 *   SETUP (once, first): U[k][i] = w_k pi_i sum_j P_z[i][j] S[k][j] for the pruned subtree's CLV S over the prune
 *     branch, into stash slot 1 (with S's scaler counts): the part of the edge log-likelihood
 *     (core_likelihood.c:1388-1490) that is the same for every candidate.
 *   SCORE: X = (M a)(M b), the CLV the new inner node would get between the two halves of the candidate branch
 *     (a: the CLV pointing away from the pruning point, b: the one pointing towards it), is never stored: the site
 *     likelihood sum_k sum_i X[k][i] U[k][i] goes through log, scaler term and pattern weight, the warp adds its
 *     32 sites and accumulates into score_out[candidate][warp].  X is not rescaled (its children's counts only
 *     enter the log term; 2^-768 is still a normal double), which changes nothing mathematically. */
template <bool FULL>
__device__ __forceinline__ void lean_op_special(const LeanArgs & a, const LaneCtx & L, unsigned int rec, unsigned int st,
                                                double (&acc)[4][4], uint4 & fsc, TipQ & q)
{
  const uint4 w0 = lds_u4(rec), w1 = lds_u4(rec + 16);
  unsigned int n1, n2;
  lean_prefetch<FULL>(a, L, rec + 32, rec + 32 * L4_TIP_AHEAD, st ^ 1u, n1, n2);
  const unsigned int t1 = q.t1, t2 = q.t2;
  const bool setup = w0.x == LSPECIAL_SETUP;
  uint4 sc;
  lean_children_generic<FULL>(a, L, w0, w1, st, L.sc_lane(), setup, acc, fsc, t1, t2, sc);
  const unsigned int k = L.k();
  const unsigned int ul = L.stash_lane() + L4_STASH_B; /* stash slot 1 */
  if (setup)
  {
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[r][i] *= a.wpi[k * 4 + i];
      sts_d2(ul + r * 1024, acc[r][0], acc[r][1]);
      sts_d2(ul + r * 1024 + 512, acc[r][2], acc[r][3]);
    }
    if (L.sc_lane()) sts_u4(L.sst_lane() + 128, sc);
  }
  else
  {
    double lh[4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      const double2 x = lds_d2(ul + r * 1024), y = lds_d2(ul + r * 1024 + 512);
      lh[r] = acc[r][0] * x.x + acc[r][1] * x.y + acc[r][2] * y.x + acc[r][3] * y.y;
    }
    /* sum over the 4 categories (the 4 lanes of a site group), then lane k finishes site s0 + k */
#pragma unroll
    for (int r = 0; r < 4; ++r)
    {
      lh[r] += __shfl_xor_sync(0xffffffffu, lh[r], 1);
      lh[r] += __shfl_xor_sync(0xffffffffu, lh[r], 2);
    }
    const uint4 su = lds_u4(L.sst_lane() + 128);
    /* the children's counts live in the group's first lane */
    const int lead = (int)(L.lane & ~3u);
    const unsigned int cx = __shfl_sync(0xffffffffu, sc.x, lead), cy = __shfl_sync(0xffffffffu, sc.y, lead);
    const unsigned int cz = __shfl_sync(0xffffffffu, sc.z, lead), cw = __shfl_sync(0xffffffffu, sc.w, lead);
    const double mine = k == 0 ? lh[0] : (k == 1 ? lh[1] : (k == 2 ? lh[2] : lh[3]));
    const unsigned int cnt = k == 0 ? cx + su.x : (k == 1 ? cy + su.y : (k == 2 ? cz + su.z : cw + su.w));
    double v = 0.0;
    if (FULL || ((L.actmask >> k) & 1u)) v = (log(mine) + (double)cnt * LN_SCALE_THRESHOLD) * (double)a.weights[L.s0 + k];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (L.lane == 0) a.score_out[(size_t)w0.w * a.score_stride + (blockIdx.x * L4_WARPS + (threadIdx.x >> 5))] += v;
  }
  fsc = make_uint4(0, 0, 0, 0);
  q.advance(n1, n2);
}

/* One chunk (32 consecutive sites) through the staged operations [0, nseg); the record after the last one is
 * all zero (a harmless prefetch target).  wsm: this warp's shared-memory block (shared-window address), s0:
 * first of the lane's 4 consecutive sites.  FULL: all 32 sites exist — only then the specialised bodies run. */
template <bool FULL, bool SCORING>
__device__ __forceinline__ void lean_chunk(const LeanArgs & a, unsigned int sops, unsigned int nseg, unsigned int s0,
                                           int lane, unsigned int wsm)
{
  LaneCtx L;
  L.lane = (unsigned int)lane;
  L.s0 = s0;
  L.clv_off = (s0 * 4u + (lane & 3u)) * 32u;
  L.wsm = wsm;
  L.actmask = 0xfu;
  if (!FULL)
  {
    L.actmask = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (s0 + r < a.sites) L.actmask |= 1u << r;
  }

  double acc[4][4]; /* [element r][state i]: the parent of the previous operation when an operation starts */
#pragma unroll
  for (int r = 0; r < 4; ++r) acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.0;
  uint4 fsc = make_uint4(0, 0, 0, 0); /* ... and its scaler counts (sc lanes) */
  TipQ q;                             /* packed tip column codes of the CURRENT operation (and the next) */

  __syncwarp();
  lean_prefetch<FULL>(a, L, sops, sops, 0, q.t1, q.t2);
  q.u1 = q.u2 = 0;
  if (L4_TIP_AHEAD == 2)
  {
    const uint4 y0 = lds_u4(sops + 32);
    const unsigned int ft = lds_u1(sops + 32 + 24) >> 24;
    if (LOPA_SRC1(ft) == SRC_TIP) q.u1 = lean_load_tips<FULL>(a, L, y0.y);
    if (LOPA_SRC2(ft) == SRC_TIP) q.u2 = lean_load_tips<FULL>(a, L, y0.z);
  }

  /* The loop variable is the record's shared-memory address (records are 32 bytes: bit 5 = stage).  There is no
   * trip count: the record behind the last operation is all zero (no real operation has parent row = child rows =
   * 0) and decodes to the generic type, whose branch leaves the loop.  [With `rec != rec_end` ptxas evaluated the
   * test at the top of an iteration and kept the predicate in LOCAL memory across the body — every predicate
   * register is in use —, and its reload behind the body's stores took 19 % of the kernel's stall samples
   * (profiles/r02u_trav_dna4_stalls_by_sass.txt); a test pinned behind the body reloaded `nseg` instead.] */
  (void)nseg;
  for (unsigned int rec = sops;; rec += 32)
  {
    const unsigned int st = ((rec - sops) >> 5) & 1u;
    __syncwarp(); /* every lane is done with stage st^1 (previous operation): the bodies refill it for the next one */
    const unsigned int m2f = lds_u1(rec + 28);
    unsigned int type = LOPB_TYPE(m2f >> 24);
    if (!(SCORING && type == LTYPE_SPECIAL) && (!FULL || __any_sync(0xffffffffu, ((q.t1 | q.t2) & 0x80808080u) != 0u)))
      type = 0; /* partial chunk, or ambiguity codes: generic body */
    switch (type)
    {
      case LTYPE_TIP_TIP: lean_op_hot<SRC_TIP, SRC_TIP>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_TIP_FWD: lean_op_hot<SRC_TIP, SRC_FWD>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_STASH_FWD: lean_op_hot<SRC_STASH, SRC_FWD>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_GLOBAL_FWD: lean_op_hot<SRC_GLOBAL, SRC_FWD>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_TIP_GLOBAL: lean_op_hot<SRC_TIP, SRC_GLOBAL>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_GLOBAL_GLOBAL: lean_op_hot<SRC_GLOBAL, SRC_GLOBAL>(a, L, rec, st, acc, fsc, q); break;
      case LTYPE_SPECIAL:
        if (SCORING) lean_op_special<FULL>(a, L, rec, st, acc, fsc, q); /* only the scoring kernel carries this body */
        break;
      default:
        if (m2f == 0u)
        {
          const uint4 z = lds_u4(rec);
          if ((z.x | z.y | z.z) == 0u) goto list_done;
        }
        lean_op_generic<FULL>(a, L, rec, st, acc, fsc, q);
        break;
    }
  }
list_done:
  /* the last operation prefetched the zero record behind the segment: that copy must have landed before the
   * next chunk's first copies go into the same stage (copies of different groups may land in any order) */
  cp_async_wait0();
  __syncwarp();
}

template <int TAG, bool SCORING>
__device__ __noinline__ void lean_chunk_tail(const LeanArgs & a, unsigned int sops, unsigned int nseg, unsigned int s0,
                                             int lane, unsigned int wsm)
{
  lean_chunk<false, SCORING>(a, sops, nseg, s0, lane, wsm);
}

/* bid / nblk: this CTA's number among the CTAs that share the chunks of one list (blockIdx.x / gridDim.x everywhere
 * but in the sub-list kernel below) */
template <int TAG, bool SCORING>
__device__ __forceinline__ void trav_dna4_body(const LeanArgs & a, const unsigned int bid, const unsigned int nblk)
{
  extern __shared__ __align__(128) unsigned char l4_smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned int sm0 = smem_u32(l4_smem);
  const unsigned int sops = sm0;
  const unsigned int wsm = sm0 + (a.seg + 4) * 32 + warp * L4_WARP_SM; /* + the zero record, 128-byte aligned */
  const unsigned int nchunks = (unsigned int)(((unsigned long long)a.sites + 31) / 32);
  const unsigned int nwarps = nblk * L4_WARPS;


  for (unsigned int seg = 0; seg < a.nops; seg += a.seg)
  {
    const unsigned int nseg = min(a.nops - seg, a.seg);
    __syncthreads(); /* previous segment fully consumed by every warp of the CTA */
    {
      const uint4 * src = reinterpret_cast<const uint4 *>(a.ops + seg);
      for (unsigned int i = threadIdx.x; i < nseg * 2; i += blockDim.x) sts_u4(sops + i * 16, __ldg(src + i));
      if (threadIdx.x < 8) sts_u4(sops + (nseg * 2 + threadIdx.x) * 16, make_uint4(0, 0, 0, 0)); /* prefetch targets behind the last operation */
    }
    __syncthreads();
    /* full rounds: a CTA takes L4_WARPS consecutive chunks; the last, partial round is dealt out warp-major over the
     * CTAs (warp 0 of every CTA, then warp 1 ...) so that every SM keeps the same number of warps busy: with the
     * plain stride the first CTAs ran a whole extra round at full occupancy while the other SMs idled (125 k sites
     * per GPU = 1.65 rounds cost 2.0) */
    const unsigned int full_end = (a.dbg & 8u) ? nchunks : nchunks - nchunks % nwarps;
    for (unsigned int round0 = 0; round0 < nchunks; round0 += nwarps)
    {
      const unsigned int chunk =
        round0 + (round0 < full_end ? bid * L4_WARPS + warp : warp * nblk + bid);
      if (chunk >= nchunks) break;
      const unsigned int s0 = chunk * 32u + (unsigned int)(lane >> 2) * 4;
      if ((unsigned long long)(chunk + 1) * 32 <= a.sites)
        lean_chunk<true, SCORING>(a, sops, nseg, s0, lane, wsm);
      else
        lean_chunk_tail<TAG, SCORING>(a, sops, nseg, s0, lane, wsm);
    }
  }
}

__global__ void __launch_bounds__(L4_WARPS * 32, 2) trav_dna4_kernel(LeanArgs a) { trav_dna4_body<0, false>(a, blockIdx.x, gridDim.x); }
/* the same kernel with the SETUP / SCORE bodies of the batched regraft scoring (lean_op_special) */
__global__ void __launch_bounds__(L4_WARPS * 32, 2) trav_dna4_score_kernel(LeanArgs a) { trav_dna4_body<2, true>(a, blockIdx.x, gridDim.x); }

/* many partitions in one launch (blockIdx.y = partition; batch mode, internal.h) */
__global__ void __launch_bounds__(L4_WARPS * 32, 2) trav_dna4_batch_kernel(const LeanArgs * __restrict__ arr)
{
  __shared__ LeanArgs sa;
  for (unsigned int i = threadIdx.x; i < sizeof(LeanArgs) / 4; i += blockDim.x)
    reinterpret_cast<unsigned int *>(&sa)[i] = reinterpret_cast<const unsigned int *>(arr + blockIdx.y)[i];
  __syncthreads();
  const LeanArgs a = sa;
  trav_dna4_body<1, false>(a, blockIdx.x, gridDim.x);
}

/* One full traversal as independent sub-lists + the operations above them, in ONE launch (launch_trav_dna4_split):
 * blockIdx.x = sub-list (arr[0 .. nsub-1]), blockIdx.y = group of L4_WARPS chunks.  A CTA walks its sub-list for its
 * group and takes a ticket of the group; the CTA that takes the last one finds every subtree root of the group in
 * global memory and walks the top list (arr[nsub]) for it right away, while other CTAs still work on sub-lists of other
 * groups — the short, latency-bound top chains hide between them instead of forming a sparse second launch.
 * tickets[gridDim.y] are zero at launch and zero again at exit. */
__global__ void __launch_bounds__(L4_WARPS * 32, 2) trav_dna4_split_kernel(const LeanArgs * __restrict__ arr, unsigned int nsub,
                                                                          unsigned int * __restrict__ tickets)
{
  __shared__ LeanArgs sa;
  __shared__ unsigned int s_last;
  unsigned int which = blockIdx.x;
  for (int phase = 0; phase < 2; ++phase)
  {
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < sizeof(LeanArgs) / 4; i += blockDim.x)
      reinterpret_cast<unsigned int *>(&sa)[i] = reinterpret_cast<const unsigned int *>(arr + which)[i];
    __syncthreads();
    {
      const LeanArgs a = sa;
      trav_dna4_body<3, false>(a, blockIdx.y, gridDim.y);
    }
    if (phase) break;
    __threadfence(); /* this thread's CLV and scaler stores are visible device-wide before the ticket is */
    __syncthreads();
    if (threadIdx.x == 0)
    {
      const unsigned int t = atomicAdd(tickets + blockIdx.y, 1u);
      s_last = t + 1 == nsub;
      if (s_last) tickets[blockIdx.y] = 0;
    }
    __syncthreads();
    if (!s_last) break;
    __threadfence(); /* ... and the other CTAs' stores before this CTA's loads */
    which = nsub;
  }
}

/* ---- tip column codes ---------------------------------------------------------------------------- */

/* 4-bit DNA code -> 16-byte offset of its column in the table (column c at 128 c, odd columns with their halves
 * exchanged: + 1): A C G T -> 0, 9, 16, 25, N (all four) -> 32, anything else -> 0x80 | code */
__global__ void tipcols_encode_kernel(const unsigned char * __restrict__ src, unsigned char * __restrict__ dst, size_t n)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
  {
    const unsigned int c = src[i] & 15u;
    unsigned char v;
    switch (c)
    {
      case 1: v = 0; break;
      case 2: v = 9; break;
      case 4: v = 16; break;
      case 8: v = 25; break;
      case 15: v = L4_TIPCOL_N; break;
      default: v = (unsigned char)(0x80u | c); break;
    }
    dst[i] = v;
  }
}

bool launch_tipcols_encode(PartCtx * c, unsigned int tip)
{
  if (!c->d_tipcols) return true;
  const size_t n = c->tip_stride;
  const size_t off = (size_t)tip * c->tip_stride;
  unsigned int blocks = (unsigned int)std::min<size_t>((n + 255) / 256, 1184);
  tipcols_encode_kernel<<<blocks, 256, 0, xstream(c)>>>(c->d_tipchars + off, c->d_tipcols + off, n);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "tipcols_encode_kernel launch");
}

/* ---- host ------------------------------------------------------------------------------------ */

int g_opt_trav_stash = 2; /* stash slots per warp: 0, 1 or 2 */
int g_opt_trav_debug = 0; /* LeanArgs::dbg */

static size_t lean_smem_bytes(unsigned int seg) { return (size_t)(seg + 4) * 32 + (size_t)L4_WARPS * L4_WARP_SM; }

static void lean_fill_args(PartCtx * c, LeanArgs & a, const LeanOp * d_ops, unsigned int count, unsigned int seg);
static bool lean_set_smem_attr();

bool lean_supported(const PartCtx * c)
{
  const pll_partition_t * p = &c->pub;
  return p->states == 4 && p->rate_cats == 4 && !(p->attributes & PLL_ATTRIB_RATE_SCALERS) && c->d_pmat_cols &&
         (!(p->attributes & PLL_ATTRIB_PATTERN_TIP) || c->d_tipcols) &&
         c->clv_stride * sizeof(double) < (1ull << 32) && c->scaler_stride * sizeof(unsigned int) < (1ull << 32);
}

/* A full traversal as independent sub-lists (trav_split_ops): a warp's walk over a whole list is one dependent chain
 * of `count` operations (~1 us each), so a launch cannot finish faster than (rounds of chunks per warp) x that chain,
 * however few warps the last round keeps busy: an 8-GPU shard of C2 (125 k sites = 1.65 rounds) paid 2.3 rounds.
 * Here a CTA takes ONE sub-list (subtrees of at most `target` operations, nothing shared between sub-lists) for 8
 * chunks, the block scheduler deals CTAs out as others finish, and the CTA that completes a group of chunks walks the
 * operations above the subtrees for it (trav_dna4_split_kernel). */
/* measured on one B200 (profiles/r02q_split_sweep.txt, r02v_split_fused.txt), kernel time of a full C2 traversal, single
 * list -> sub-lists of 128: 125 k sites 4.49 -> 3.44 ms, 250 k 7.57 -> 6.77, 500 k 12.8 -> 13.05 and 1 M 25.4 -> 26.6
 * (worse: many rounds hide the chain, and the top's operations read their children from DRAM) */
int g_opt_trav_split = 128;      /* sub-list length; 0 = never split */
int g_opt_trav_split_rounds = 4; /* split only when a resident warp would take fewer chunks than this (~300 k sites) */

static std::atomic<unsigned long long> g_split_launches{0};
/* tests: how many traversals ran as sub-lists so far */
extern "C" unsigned long long pllb200_debug_split_launches(void) { return g_split_launches.load(); }

static bool launch_trav_dna4_split(PartCtx * c, const pll_operation_t * ops, unsigned int count, bool * done)
{
  pll_partition_t * p = &c->pub;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  static thread_local std::vector<pll_operation_t> ordered;
  static thread_local std::vector<unsigned int> bounds;
  static thread_local std::vector<LeanOp> lean, part;
  static thread_local std::vector<LeanArgs> args;
  *done = false;
  const unsigned int target = std::min<unsigned int>((unsigned int)g_opt_trav_split, L4_SEG_MAX);
  if (!trav_split_ops(ops, count, p->nodes, target, ordered, bounds)) return true;
  const unsigned int nlists = (unsigned int)bounds.size() - 1; /* the last one is the top */
  if (nlists < 2 || !c->d_split_tickets) return true;
  lean.resize(count);
  for (unsigned int l = 0; l < nlists; ++l)
  {
    const unsigned int b = bounds[l], n = bounds[l + 1] - b;
    if (!trav_lean_resolve(ordered.data() + b, n, pattern_tip ? p->tips : 0, c->clv_first, std::min<unsigned int>(L4_SEG_MAX, n),
                           g_opt_trav_stash, g_opt_trav_forward != 0, part, nullptr))
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Operation list cannot be packed (matrix index >= 2^24).");
      return false;
    }
    std::copy(part.begin(), part.end(), lean.begin() + b);
  }
  if (!lean_set_smem_attr()) return false;
  const LeanOp * d_ops = (const LeanOp *)dev_stage_small(c, lean.data(), sizeof(LeanOp) * count);
  if (!d_ops) return false;
  const unsigned long long nchunks = ((unsigned long long)c->sites_alloc + 31) / 32;
  const unsigned long long groups = (nchunks + L4_WARPS - 1) / L4_WARPS;
  if (groups > 65535ull) return true; /* gridDim.y */
  args.resize(nlists);
  unsigned int maxseg = 1;
  for (unsigned int l = 0; l < nlists; ++l)
  {
    const unsigned int n = bounds[l + 1] - bounds[l];
    const unsigned int seg = std::min<unsigned int>(L4_SEG_MAX, (n + 3u) & ~3u); /* keeps the per-warp blocks behind the records 128-byte aligned */
    lean_fill_args(c, args[l], d_ops + bounds[l], n, seg);
    maxseg = std::max(maxseg, seg);
  }
  /* read by every CTA: device memory, never the mapped host ring */
  const LeanArgs * d_args = (const LeanArgs *)dev_stage(c, args.data(), sizeof(LeanArgs) * args.size());
  if (!d_args) return false;
  if (!pllb200_cuda_ok(cudaMemsetAsync(c->d_split_tickets, 0, sizeof(unsigned int) * groups, xstream(c)), "split tickets"))
    return false;
  trav_dna4_split_kernel<<<dim3(nlists - 1, (unsigned int)groups), L4_WARPS * 32, lean_smem_bytes(maxseg), xstream(c)>>>(
      d_args, nlists - 1, c->d_split_tickets);
  ++g_kernel_launches;
  ++g_split_launches;
  if (!pllb200_cuda_ok(cudaGetLastError(), "trav_dna4_split_kernel launch")) return false;
  *done = true;
  return true;
}

bool launch_trav_dna4(PartCtx * c, const pll_operation_t * ops, unsigned int count)
{
  pll_partition_t * p = &c->pub;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  static thread_local std::vector<LeanOp> lean;
  const bool collect = batch_collects(BK_TRAV_DNA4);
  if (!collect && g_opt_trav_split > 1 && count > 2u * (unsigned int)g_opt_trav_split)
  {
    const unsigned long long nwarps = (unsigned long long)c->sm_count * 2 * L4_WARPS;
    const unsigned long long nchunks = ((unsigned long long)c->sites_alloc + 31) / 32;
    if (nchunks < (unsigned long long)g_opt_trav_split_rounds * nwarps)
    {
      bool done = false;
      if (!launch_trav_dna4_split(c, ops, count, &done)) return false;
      if (done) return true;
    }
  }
  const unsigned int seg = std::min<unsigned int>(L4_SEG_MAX, count);
  if (!trav_lean_resolve(ops, count, pattern_tip ? p->tips : 0, c->clv_first, seg, g_opt_trav_stash,
                         g_opt_trav_forward != 0, lean, nullptr))
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Operation list cannot be packed (matrix index >= 2^24).");
    return false;
  }
  if (!lean_set_smem_attr()) return false;
  const LeanOp * d_ops = collect ? (const LeanOp *)batch_pool_add(lean.data(), sizeof(LeanOp) * count)
                                 : (const LeanOp *)dev_stage_small(c, lean.data(), sizeof(LeanOp) * count);
  if (!d_ops) return false;
  LeanArgs a;
  lean_fill_args(c, a, d_ops, count, seg);

  const unsigned long long nchunks = ((unsigned long long)a.sites + 31) / 32;
  unsigned long long blocks = (nchunks + L4_WARPS - 1) / L4_WARPS;
  int per_sm = 2;
  if (g_opt_trav_ctas_per_sm > 0) per_sm = std::min(per_sm, g_opt_trav_ctas_per_sm);
  const unsigned long long cap = (unsigned long long)c->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (collect)
  {
    batch_collect(BK_TRAV_DNA4, &a, sizeof(a), (unsigned int)blocks, lean_smem_bytes(L4_SEG_MAX));
    return true;
  }
  trav_dna4_kernel<<<(unsigned int)blocks, L4_WARPS * 32, lean_smem_bytes(seg), xstream(c)>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "trav_dna4_kernel launch");
}

/* ---- batched regraft scoring (algo_search.cpp) ------------------------------------------------------ */

/* score[c] = sum over the launch's warps of score_out[c][warp], in a fixed order (deterministic) */
__global__ void __launch_bounds__(256) spr_score_reduce_kernel(const double * __restrict__ part, unsigned int stride,
                                                              double * __restrict__ out)
{
  __shared__ double s_red[256];
  const double * row = part + (size_t)blockIdx.x * stride;
  double x = 0.0;
  for (unsigned int i = threadIdx.x; i < stride; i += 256) x += row[i];
  s_red[threadIdx.x] = x;
  __syncthreads();
  for (unsigned int h = 128; h > 0; h >>= 1)
  {
    if (threadIdx.x < h) s_red[threadIdx.x] += s_red[threadIdx.x + h];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = s_red[0];
}

static void lean_fill_args(PartCtx * c, LeanArgs & a, const LeanOp * d_ops, unsigned int count, unsigned int seg)
{
  memset(&a, 0, sizeof(a));
  a.ops = d_ops;
  a.nops = count;
  a.sites = c->sites_alloc;
  a.seg = seg;
  a.tip_stride = (unsigned int)c->tip_stride;
  a.clv = reinterpret_cast<char *>(c->d_clv);
  a.scaler = reinterpret_cast<char *>(c->d_scaler);
  a.pmat_il = c->d_pmat_il;
  a.pmat_cols = c->d_pmat_cols;
  a.tipcols = c->d_tipcols;
  a.dbg = (unsigned int)g_opt_trav_debug;
  a.clv_stride_b = (unsigned int)(c->clv_stride * sizeof(double));
  a.scaler_stride_b = (unsigned int)(c->scaler_stride * sizeof(unsigned int));
}

static bool lean_set_smem_attr()
{
  static bool attr_set = false;
  if (attr_set) return true;
  const int maxb = (int)lean_smem_bytes(L4_SEG_MAX);
  if (!pllb200_cuda_ok(cudaFuncSetAttribute(trav_dna4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb),
                       "trav_dna4_kernel smem") ||
      !pllb200_cuda_ok(cudaFuncSetAttribute(trav_dna4_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb),
                       "trav_dna4_batch_kernel smem") ||
      !pllb200_cuda_ok(cudaFuncSetAttribute(trav_dna4_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb),
                       "trav_dna4_score_kernel smem") ||
      !pllb200_cuda_ok(cudaFuncSetAttribute(trav_dna4_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, maxb),
                       "trav_dna4_split_kernel smem"))
    return false;
  attr_set = true;
  return true;
}

/* One launch over a list that mixes ordinary operations (kind 0; they may write scratch CLV rows: clv index
 * nodes + r, scaler index scale_buffers + r), one SETUP (kind LSPECIAL_SETUP, first) and SCORE operations (kind
 * LSPECIAL_SCORE, cand[i] = candidate number).  Special operations name their operands like ordinary ones
 * (SETUP: child 2 = the pruned subtree over the prune branch; SCORE: the two CLVs either side of the candidate
 * branch, both over the candidate's half-length matrix) and a parent row nobody reads.  host_scores[ncand] comes
 * back after a stream synchronisation.  Preconditions (checked by the caller): lean_supported, no invariant
 * sites, no ascertainment bias. */
bool launch_trav_dna4_score(PartCtx * c, const pll_operation_t * ops, const unsigned char * kind,
                            const unsigned int * cand, unsigned int count, unsigned int ncand,
                            const unsigned int * params_indices, double * host_scores)
{
  pll_partition_t * p = &c->pub;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  if (!dev_sync_inputs(c)) return false;
  static thread_local std::vector<LeanOp> lean;
  const unsigned int seg = std::min<unsigned int>(L4_SEG_MAX, count);
  /* stash slot 1 holds the SETUP result for the whole list (one slot is left for ordinary values) and does not
   * survive a segment boundary: the list must fit one segment */
  if (count > L4_SEG_MAX || ncand > PLLB200_SCORE_MAX ||
      !trav_lean_resolve(ops, count, pattern_tip ? p->tips : 0, c->clv_first, seg, 1, true, lean, nullptr))
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Regraft scoring list cannot be packed.");
    return false;
  }
  for (unsigned int i = 0; i < count; ++i)
    if (kind[i])
    {
      LeanOp & d = lean[i];
      unsigned int fa = d.m1f >> 24, fb = d.m2f >> 24;
      fa &= ~(unsigned int)(LOPA_PS | LOPA_STASH_W);
      fb = (fb & 0x0fu) | ((unsigned int)LTYPE_SPECIAL << 5);
      d.m1f = (d.m1f & LOP_MATRIX_MASK) | (fa << 24);
      d.m2f = (d.m2f & LOP_MATRIX_MASK) | (fb << 24);
      d.p = kind[i];
      d.ps = cand[i];
    }
  if (!lean_set_smem_attr()) return false;
  const LeanOp * d_ops = (const LeanOp *)dev_stage(c, lean.data(), sizeof(LeanOp) * count);
  if (!d_ops) return false;
  LeanArgs a;
  lean_fill_args(c, a, d_ops, count, seg);
  const unsigned long long nchunks = ((unsigned long long)a.sites + 31) / 32;
  unsigned long long blocks = std::min<unsigned long long>((nchunks + L4_WARPS - 1) / L4_WARPS, (unsigned long long)c->sm_count * 2);
  if (blocks < 1) blocks = 1;
  const unsigned int stride = (unsigned int)blocks * L4_WARPS;
  const size_t need = (size_t)ncand * stride + ncand;
  if (c->score_cap < need)
  {
    CUDA_TRY(cudaStreamSynchronize(c->stream), false);
    cudaFree(c->d_score);
    c->d_score = nullptr;
    c->score_cap = 0;
    CUDA_TRY(cudaMalloc(&c->d_score, need * 2 * sizeof(double)), false);
    c->score_cap = need * 2;
  }
  if (!c->h_score) CUDA_TRY(cudaHostAlloc(&c->h_score, PLLB200_SCORE_MAX * sizeof(double), cudaHostAllocDefault), false);
  CUDA_TRY(cudaMemsetAsync(c->d_score, 0, (size_t)ncand * stride * sizeof(double), c->stream), false);
  a.score_out = c->d_score;
  a.score_stride = stride;
  a.weights = c->d_weights;
  for (unsigned int k = 0; k < 4; ++k)
    for (unsigned int i = 0; i < 4; ++i) a.wpi[k * 4 + i] = p->rate_weights[k] * p->frequencies[params_indices[k]][i];
  trav_dna4_score_kernel<<<(unsigned int)blocks, L4_WARPS * 32, lean_smem_bytes(seg), c->stream>>>(a);
  ++g_kernel_launches;
  if (!pllb200_cuda_ok(cudaGetLastError(), "trav_dna4_score_kernel launch")) return false;
  double * d_out = c->d_score + (size_t)ncand * stride;
  spr_score_reduce_kernel<<<ncand, 256, 0, c->stream>>>(c->d_score, stride, d_out);
  ++g_kernel_launches;
  if (!pllb200_cuda_ok(cudaGetLastError(), "spr_score_reduce_kernel launch")) return false;
  CUDA_TRY(cudaMemcpyAsync(c->h_score, d_out, sizeof(double) * ncand, cudaMemcpyDeviceToHost, c->stream), false);
  g_d2h_bytes += sizeof(double) * ncand;
  if (!dev_flush(c)) return false;
  memcpy(host_scores, c->h_score, sizeof(double) * ncand);
  return true;
}

bool flush_trav_dna4_lean(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  trav_dna4_batch_kernel<<<dim3(batch_grid_x(lead, k, 2), k.count), L4_WARPS * 32, k.smem, lead->stream>>>(
      static_cast<const LeanArgs *>(d_args));
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "batched trav_dna4 kernel launch");
}

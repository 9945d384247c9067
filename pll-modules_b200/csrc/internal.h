/*
 * internal.h — private state behind pll_partition_t for the B200 library.
 *
 * A partition handed out by pll_partition_create is the first member of a PartCtx, so the public
 * pointer doubles as the handle to its HBM arenas, stream and staging buffers
 * (SURVEY.md §8b "Device-state attachment": over-allocation instead of a side registry).
 */
#ifndef PLLB200_INTERNAL_H_
#define PLLB200_INTERNAL_H_

#include "../../include/pllb200.h"

#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <unordered_map>
#include <vector>

#define PLLB200_MAGIC 0xB200C1F5u
#define PLLB200_RESULT_SLOTS 64   /* async reduction results in flight per partition */
#define PLLB200_MAX_RED_BLOCKS 2048 /* upper bound on reduction grid size */
#define PLLB200_SCORE_MAX 1024     /* regraft candidates per scoring launch */
#define PLLB200_MAX_SUMTABLES 64   /* device shadows of caller-owned sumtable buffers per partition */

struct PendingResult
{
  double * host_dst;
  int slot;
  int comp; /* component inside the slot */
};

struct PartCtx
{
  pll_partition_t pub; /* MUST stay first */
  unsigned int magic;
  int device;
  int sm_count;

  cudaStream_t own_stream;
  cudaStream_t stream;

  /* geometry */
  unsigned int sites_alloc;   /* sites + asc_additional_sites */
  unsigned int clv_first;     /* first node index that owns a CLV buffer */
  size_t clv_elems;           /* doubles per CLV: sites_alloc*K*sp */
  size_t clv_stride;          /* doubles between consecutive CLVs (padded to 256 B) */
  size_t scaler_elems;        /* u32 per scale buffer */
  size_t scaler_stride;
  size_t pmat_elems;          /* doubles per P-matrix: K*s*sp */
  size_t tip_stride;          /* bytes between tipchar rows */
  size_t eigen_stride;        /* doubles per rate matrix in the eigen block */

  /* HBM arenas */
  double * d_clv;
  unsigned int * d_scaler;
  double * d_pmat;
  double * d_pmat_il;         /* 4-state interleaved copy [i][k][j] for the DNA traversal kernel */
  /* 4 states: per matrix the five column sums a DNA tip code selects, [m][A C G T N][cat k][row i] — what a tip
   * child contributes to a parent, read by table lookup in k_traversal_dna4.cu */
  double * d_pmat_cols;
  unsigned int * d_split_tickets; /* one counter per group of 8 chunks (trav_dna4_split_kernel), zero between launches */
  unsigned char * d_tipcols;  /* twin of d_tipchars (4 states + tip patterns): column number of every code */
  /* 20 states + tip patterns: per matrix the products P x (tip state vector) of every tip code,
   * [prob_matrices][code 24][k][20] — the same K*480 doubles per matrix as the fragment image, so a tip
   * child's slot in the traversal kernel's shared memory holds this table instead (k_traversal_aa.cu) */
  double * d_tiptab;
  bool tiptab_stale;          /* tip map changed after tables were built: rebuild before the next use */
  unsigned char * d_tipchars;
  unsigned int * d_weights;
  int * d_invariant;
  double * d_eigen;           /* [rate_matrix][ evecs s*sp | inv_evecs s*sp | evals sp | freqs sp ] */
  unsigned long long * d_tipmap;
  double * d_persite;         /* lazily allocated, sites doubles */
  double * d_ancestral;       /* lazily allocated, sites*states doubles (pll_compute_node_ancestral) */
  /* multi-GPU exchange for the fused Newton kernel (k_newton.cu): own buffer + peers' IPC mappings */
  double * xchg_peer[8];
  int xchg_rank, xchg_nranks;
  double * h_xred;            /* mapped pinned result of pllb200_exchange_reduce_cb (+ completion tag) */
  double * h_xred_dev;
  unsigned long long xred_seq;
  double * d_sumfrag;         /* lazily allocated: 20-state sumtable matrices as DMMA fragment images */
  std::vector<unsigned long long> sumfrag_key;
  double * d_sumgen = nullptr; /* lazily allocated: 17..64-state sumtable matrices [L | R][K][s][sp] */
  std::vector<unsigned long long> sumgen_key;
  unsigned long long tipmap_gen = 0; /* bumped whenever the tip map is uploaded again */
  /* batched regraft scoring (k_traversal_dna4.cu): per-warp partial sums + results, pinned landing buffer */
  double * d_score;
  size_t score_cap;
  double * h_score;
  unsigned int scratch_rows;  /* extra CLV / scaler rows behind the arenas (clv index nodes + r, scaler index scale_buffers + r) */
  unsigned int scratch_mats;  /* extra P-matrix slots (matrix index prob_matrices + j) */
  double * d_brlen;           /* lazily allocated, [prob_matrices][4]: queued Newton results (k_newton.cu) */
  /* multi-partition Newton (k_newton.cu): descriptor array on the device + the host image it mirrors;
   * used on the first partition of a set; ev_batch orders this partition's stream before a batched launch */
  void * d_multi;
  size_t multi_cap;
  std::vector<unsigned char> multi_shadow, multi_shadow_next;
  cudaEvent_t ev_batch;

  /* staging ring: pinned host + device mirror, stream-ordered */
  char * h_stage;
  char * d_stage;
  size_t stage_size;
  size_t stage_head;
  std::vector<void *> retired_dev, retired_host; /* rings replaced by a bigger one: freed at the next dev_flush */
  char * h_small;            /* mapped pinned ring for payloads of a few bytes (dev_stage_small) */
  char * h_small_dev;
  size_t small_head;

  /* reductions */
  double * d_partials;        /* PLLB200_MAX_RED_BLOCKS*3 doubles */
  unsigned int * d_ticket;
  double * d_result;          /* PLLB200_RESULT_SLOTS*4 doubles on device */
  double * h_result;          /* same, pinned + mapped: kernels write results straight to the host */
  double * h_result_dev;      /* device alias of h_result */
  /* completion tags, one per slot, behind the results in the same mapped allocation: a kernel writes
   * slot_tag[slot] there after its result, the host polls (dev_wait_slot) instead of synchronising */
  volatile unsigned long long * h_tag;
  unsigned long long * h_tag_dev;
  unsigned long long slot_tag[PLLB200_RESULT_SLOTS];
  unsigned long long tag_counter;
  int next_slot;
  std::vector<PendingResult> pending;

  /* derivative passes: table of the last call and its direction (k_derivatives.cu) */
  const double * deriv_last_table;
  int deriv_reverse;

  /* sumtable shadows keyed by the caller's host pointer */
  std::unordered_map<const void *, double *> sumtables;
  /* last use of every shadow: at most PLLB200_MAX_SUMTABLES are kept, least recently used first out */
  std::unordered_map<const void *, unsigned long long> sumtable_stamp;
  unsigned long long sumtable_clock;

  /* host->device mirroring state */
  std::vector<unsigned char> tip_dirty;
  bool any_tip_dirty;
  bool weights_dirty;
  bool invariant_dirty;
  bool tipmap_dirty;
  std::vector<unsigned int> eigen_version;     /* bumped by pll_update_eigen */
  std::vector<unsigned int> eigen_version_dev; /* version mirrored on the device */

  /* scratch for op-list scheduling */
  std::vector<pll_operation_t> op_scratch;
  std::vector<int> idx_scratch;
  std::vector<unsigned long long> devop_scratch; /* resolved ops for the 4-state traversal kernel */
};

static inline PartCtx * ctx_of(pll_partition_t * p) { return reinterpret_cast<PartCtx *>(p); }
static inline const PartCtx * ctx_of(const pll_partition_t * p)
{
  return reinterpret_cast<const PartCtx *>(p);
}

/* error plumbing (errors.cpp) */
void pllb200_set_error(int code, const char * fmt, ...);
bool pllb200_cuda_ok(cudaError_t e, const char * what);
#define CUDA_TRY(expr, failret)                        \
  do                                                   \
  {                                                    \
    if (!pllb200_cuda_ok((expr), #expr)) return failret; \
  } while (0)

/* global counters (device_ctx.cu) */
/* statistics only, but partitions may be driven from several threads: atomic */
extern std::atomic<unsigned long long> g_kernel_launches;
extern std::atomic<unsigned long long> g_newton_iterations;
extern std::atomic<unsigned long long> g_h2d_bytes;
extern std::atomic<unsigned long long> g_d2h_bytes;
extern int g_opt_trav_ctas_per_sm;
extern int g_opt_trav_reorder;
extern int g_opt_fuse_root;
extern int g_opt_trav_variant;
extern int g_opt_trav_short;
extern int g_opt_trav_split;
extern int g_opt_trav_split_rounds;
extern int g_opt_trav_forward;
extern int g_opt_trav_prefetch;
extern int g_opt_fused_newton;
extern int g_opt_queued_newton;
extern int g_opt_aa_variant;
extern int g_opt_aa_tiptab;

/* ---- batch mode (device_ctx.cu) ------------------------------------------------------------------
 * Many small partitions sharing one tree: between batch_begin and batch_end the per-partition launchers
 * of the kinds below do not launch; they append their argument struct to the collector, and batch_flush
 * issues ONE launch per kind with blockIdx.y = partition, on the lead partition's stream (staging goes
 * through the lead's ring too).  begin makes the lead stream wait for every member stream, end makes the
 * member streams wait for the lead.  Launchers that are not batch-aware must not run in between. */
/* flush order: P-matrices, CLV updates, sumtables.  Calls collected between two flushes must otherwise be
 * independent of each other (a flush is one phase of the caller's loop over partitions). */
/* BK_TRAV_AA4: Gamma4 20-state CLV updates on the register-forwarding kernel; flushed before BK_TRAV_AA */
enum { BK_PMAT = 0, BK_TRAV_DNA4, BK_TRAV_AA, BK_SUM_DNA4, BK_TRAV_AA4, BK_COUNT };
struct BatchKind
{
  std::vector<unsigned char> bytes; /* argument structs back to back */
  unsigned int count, max_blocks;
  size_t smem;
};
struct BatchCtx
{
  PartCtx * lead;
  std::vector<PartCtx *> members;
  BatchKind kind[BK_COUNT];
  std::vector<unsigned char> pool; /* resolved ops of the collected launches: staged once per flush */
};
/* ops of a launch that WILL be collected: returns a token that batch_flush turns into a device pointer
 * (the `ops` pointer is the first member of TravArgs and AaArgs) */
#define BATCH_POOL_TAG 0x7000000000000000ull
const void * batch_pool_add(const void * src, size_t bytes);
static inline bool batch_collects(int kind);
extern thread_local BatchCtx * g_batch;
extern int g_opt_batch;
void batch_collect(int kind, const void * args, size_t bytes, unsigned int blocks, size_t smem);
bool batch_flush();
/* per kind, in the kernel's own file: launch `k.count` argument structs staged at `d_args` */
bool flush_sum_dna4(PartCtx * lead, const void * d_args, const BatchKind & k);
bool flush_trav_dna4(PartCtx * lead, const void * d_args, const BatchKind & k);
bool flush_trav_aa(PartCtx * lead, const void * d_args, const BatchKind & k);
bool flush_trav_aa4(PartCtx * lead, const void * d_args, const BatchKind & k);
bool flush_pmat(PartCtx * lead, const void * d_args, const BatchKind & k);
bool batch_eligible(const PartCtx * c);
/* the stream a launcher uses: the lead's in batch mode */
static inline cudaStream_t xstream(const PartCtx * c) { return g_batch ? g_batch->lead->stream : c->stream; }
/* grid.x of a batched launch: enough CTAs per partition for the largest one, but not more than one
 * resident wave over all partitions (every CTA then pays its start-up — matrices by TMA, tables — once) */
static inline unsigned int batch_grid_x(const PartCtx * lead, const BatchKind & k, unsigned int ctas_per_sm)
{
  const unsigned int wave = (unsigned int)lead->sm_count * ctas_per_sm;
  unsigned int gx = (wave + k.count - 1) / k.count;
  if (gx > k.max_blocks) gx = k.max_blocks;
  return gx < 1 ? 1 : gx;
}
static inline bool batch_collects(int kind) { return g_batch && !(g_opt_batch & (2 << kind)); }

/* device context (device_ctx.cu) */
bool dev_ctx_init(PartCtx * c);      /* allocates arenas for c->pub geometry */
void dev_ctx_destroy(PartCtx * c);
void * dev_stage(PartCtx * c, const void * src, size_t bytes); /* returns device ptr, stream ordered */
#define PLLB200_SMALL_RING (256u << 10)
#define PLLB200_SMALL_MAX 256u
void * dev_stage_small(PartCtx * c, const void * src, size_t bytes); /* <= 256 B: mapped host memory, no copy call */
extern int g_opt_small_stage;
bool dev_sync_inputs(PartCtx * c);   /* push dirty tipchars / weights / invariant / tipmap */
bool dev_sync_eigen(PartCtx * c, const unsigned int * params_indices); /* ensure eigen block current */
bool dev_flush(PartCtx * c);            /* stream sync + deliver pending async results */
void dev_slot_tag(PartCtx * c, int slot, unsigned long long ** tag_host, unsigned long long * tag);
bool dev_wait_slot(PartCtx * c, int slot);
bool dev_wait_results(PartCtx * c);  /* wait (by tag) for the last queued result, deliver all pending ones */
int dev_alloc_slot(PartCtx * c);
/* sumtable shadows: global key registry so pll_aligned_free(key) releases the device buffer */
void sumtable_forget_key(const void * key);
void sumtable_forget_partition(PartCtx * c);

static inline double * clv_dev(const PartCtx * c, unsigned int clv_index)
{
  return c->d_clv + (size_t)(clv_index - c->clv_first) * c->clv_stride;
}
static inline unsigned int * scaler_dev(const PartCtx * c, int scaler_index)
{
  return scaler_index < 0 ? nullptr : c->d_scaler + (size_t)scaler_index * c->scaler_stride;
}
static inline bool is_tip_pattern(const PartCtx * c, unsigned int clv_index)
{
  return (c->pub.attributes & PLL_ATTRIB_PATTERN_TIP) && clv_index < c->pub.tips;
}

/* kernel launchers (k_*.cu). All are asynchronous on c->stream and return false on launch error. */
bool launch_pmatrix(PartCtx * c, const unsigned int * params_indices,
                    const unsigned int * matrix_indices, const double * branch_lengths,
                    unsigned int count);
bool launch_pmatrix_reinterleave(PartCtx * c);
bool launch_partials(PartCtx * c, const pll_operation_t * ops, unsigned int count);
bool launch_trav_dna(PartCtx * c, const pll_operation_t * ops, unsigned int count); /* k_traversal_dna.cu */
/* k_traversal_dna4.cu: Gamma-4, per-site scalers — forwarding + shared-memory stash + tip column tables */
bool lean_supported(const PartCtx * c);
bool launch_trav_dna4(PartCtx * c, const pll_operation_t * ops, unsigned int count);
bool flush_trav_dna4_lean(PartCtx * lead, const void * d_args, const BatchKind & k);
#define L4_TIPCOL_N 32 /* column code of N in d_tipcols (tipcols_encode_kernel) */
bool launch_tipcols_encode(PartCtx * c, unsigned int tip);
bool launch_trav_dna4_score(PartCtx * c, const pll_operation_t * ops, const unsigned char * kind, const unsigned int * cand,
                            unsigned int count, unsigned int ncand, const unsigned int * params_indices, double * host_scores);
extern int g_opt_generic_dmma;
extern int g_opt_eval_timing;
void eval_timing_begin(cudaStream_t s);
void eval_timing_end(cudaStream_t s);
extern int g_opt_fast_evals; /* DNA Gamma-4 logL / derivative kernels on the 32-site tile mapping */
extern int g_opt_scratch_rows;
extern int g_opt_spr_batch;
extern int g_opt_trav_stash;
extern int g_opt_trav_debug;
bool launch_trav_aa(PartCtx * c, const pll_operation_t * ops, unsigned int count);  /* k_traversal_aa.cu */
bool launch_ancestral(PartCtx * c, unsigned int node_clv, int node_scaler, unsigned int other_clv,
                      int other_scaler, unsigned int matrix_index, const unsigned int * freqs_indices,
                      double * d_out);
bool launch_edge_logl(PartCtx * c, unsigned int parent_clv, int parent_scaler,
                      unsigned int child_clv, int child_scaler, unsigned int matrix_index,
                      const unsigned int * freqs_indices, double * d_persite, int slot);
bool launch_root_logl(PartCtx * c, unsigned int clv_index, int scaler_index,
                      const unsigned int * freqs_indices, double * d_persite, int slot);
bool launch_sumtable(PartCtx * c, unsigned int parent_clv, unsigned int child_clv,
                     int parent_scaler, int child_scaler, const unsigned int * params_indices,
                     double * d_sumtable);
/* k_newton.cu: whole Newton-Raphson loop on one branch in one cooperative launch; result slot =
 * {x, status (0 ok / 1 iteration limit / 2 non-finite derivative), iterations, last f} */
bool newton_fused_supported(const PartCtx * c);
bool launch_newton(PartCtx * c, const unsigned int * params_indices, const double * d_sumtable,
                   double xmin, double x0, double xmax, double tol, unsigned int max_iters,
                   double scale, int slot, double * brlen_io = nullptr, int x0_from_device = 0,
                   int revert_on_limit = 0, int eval_only = 0);
bool launch_newton_multi(PartCtx ** cs, unsigned int count, unsigned int * const * params_indices,
                         double * const * d_sumtables, const double * scales, double xmin, double x0,
                         double xmax, double tol, unsigned int max_iters, int slot);
bool generic_dmma_usable(const PartCtx * c);
bool launch_sumtable_dmma(PartCtx * c, bool first_is_tip, unsigned int first, unsigned int second,
                          const unsigned int * params_indices, double * d_sumtable);
bool launch_sumtable_aa(PartCtx * c, bool first_is_tip, unsigned int first, unsigned int second,
                        const unsigned int * params_indices, double * d_sumtable);
bool launch_pmatrix_devlen(PartCtx * c, const unsigned int * params_indices, unsigned int matrix_index,
                           const double * d_length);
bool launch_derivatives(PartCtx * c, int parent_scaler, int child_scaler, double branch_length,
                        const unsigned int * params_indices, const double * d_sumtable, int slot);

#endif

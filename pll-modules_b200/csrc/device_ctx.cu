/*
 * device_ctx.cu — HBM residency for a partition: arenas, stream, staging ring, host->device
 * mirroring of the small host-visible arrays, result delivery.
 *
 * Layout in HBM (one cudaMalloc per arena, sized once at pll_partition_create; nothing is
 * allocated on the hot path):
 *   d_clv      [clv buffers][sites_alloc][rate_cats][states_padded] double   site-major, rate
 *              interleaved, state fastest — the reference's own order (pll.c:546-581) so a
 *              downloaded CLV is byte-compatible
 *   d_scaler   [scale_buffers][sites_alloc (x rate_cats)] u32
 *   d_pmat     [prob_matrices][rate_cats][states][states_padded] double (pll.c:593-617)
 *   d_pmat_il  4-state only: [prob_matrices][row i][cat k][col j], the shared-memory image the DNA
 *              traversal kernel stages with one bulk copy per matrix (bank-conflict free)
 *   d_tipchars [tips][tip_stride] u8, d_weights u32, d_invariant i32, d_eigen model block
 */
#include "internal.h"

#include <algorithm>

std::atomic<unsigned long long> g_kernel_launches{0};
std::atomic<unsigned long long> g_newton_iterations{0};
std::atomic<unsigned long long> g_h2d_bytes{0};
std::atomic<unsigned long long> g_d2h_bytes{0};
int g_opt_trav_ctas_per_sm = 0; /* 0 = occupancy query */
int g_opt_trav_reorder = 1;
int g_opt_fuse_root = 1;
/* scratch CLV / scaler rows behind the arenas of Gamma-4 DNA partitions: the batched regraft scoring of
 * pllmod_algo_spr_round keeps one outward-pointing CLV per search depth there (radius_max + 2 rows) */
int g_opt_scratch_rows = 12;
/* measurement aid (bench.py): device time of the last logL / derivative kernel, by events on its stream */
int g_opt_eval_timing = 0;
static cudaEvent_t g_et0, g_et1;
static bool g_et_ready = false, g_et_pending = false;
void eval_timing_begin(cudaStream_t s)
{
  if (!g_opt_eval_timing) return;
  if (!g_et_ready)
  {
    if (cudaEventCreate(&g_et0) != cudaSuccess || cudaEventCreate(&g_et1) != cudaSuccess) return;
    g_et_ready = true;
  }
  cudaEventRecord(g_et0, s);
}
void eval_timing_end(cudaStream_t s)
{
  if (!g_opt_eval_timing || !g_et_ready) return;
  cudaEventRecord(g_et1, s);
  g_et_pending = true;
}
extern "C" double pllb200_last_eval_kernel_us(void)
{
  if (!g_et_pending) return -1.0;
  float ms = 0.f;
  if (cudaEventSynchronize(g_et1) != cudaSuccess || cudaEventElapsedTime(&ms, g_et0, g_et1) != cudaSuccess) return -1.0;
  return 1000.0 * (double)ms;
}
int g_opt_spr_batch = 1; /* pllmod_algo_spr_round, fast mode: score all candidates of a pruned subtree in one launch */

static __thread int t_device = -1;

extern "C" int pllb200_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess)
  {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int pllb200_set_device(int device)
{
  int n = pllb200_device_count();
  if (device < 0 || device >= n)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid device %d (have %d).", device, n);
    return PLL_FAILURE;
  }
  t_device = device;
  return PLL_SUCCESS;
}

extern "C" const char * pllb200_version(void) { return "pllb200 0.1 sm_100a"; }

extern "C" void pllb200_get_stats(pllb200_stats_t * out)
{
  out->kernel_launches = g_kernel_launches;
  out->h2d_bytes = g_h2d_bytes;
  out->d2h_bytes = g_d2h_bytes;
  out->newton_iterations = g_newton_iterations;
}

extern "C" void pllb200_reset_stats(void)
{
  g_kernel_launches = 0;
  g_h2d_bytes = 0;
  g_d2h_bytes = 0;
  g_newton_iterations = 0;
}

extern "C" int pllb200_set_option(const char * name, int value)
{
  int * slot = nullptr;
  if (!strcmp(name, "trav_ctas_per_sm")) slot = &g_opt_trav_ctas_per_sm;
  else if (!strcmp(name, "trav_reorder")) slot = &g_opt_trav_reorder;
  else if (!strcmp(name, "fuse_root")) slot = &g_opt_fuse_root;
  else if (!strcmp(name, "trav_variant")) slot = &g_opt_trav_variant;
  else if (!strcmp(name, "trav_forward")) slot = &g_opt_trav_forward;
  else if (!strcmp(name, "trav_stash")) slot = &g_opt_trav_stash;
  else if (!strcmp(name, "trav_short")) slot = &g_opt_trav_short;
  else if (!strcmp(name, "trav_split")) slot = &g_opt_trav_split;
  else if (!strcmp(name, "trav_split_rounds")) slot = &g_opt_trav_split_rounds;
  else if (!strcmp(name, "scratch_rows")) slot = &g_opt_scratch_rows;
  else if (!strcmp(name, "generic_dmma")) slot = &g_opt_generic_dmma;
  else if (!strcmp(name, "fast_evals")) slot = &g_opt_fast_evals;
  else if (!strcmp(name, "eval_timing")) slot = &g_opt_eval_timing;
  else if (!strcmp(name, "spr_batch")) slot = &g_opt_spr_batch;
  else if (!strcmp(name, "trav_debug")) slot = &g_opt_trav_debug;
  else if (!strcmp(name, "fused_newton")) slot = &g_opt_fused_newton;
  else if (!strcmp(name, "queued_newton")) slot = &g_opt_queued_newton;
  else if (!strcmp(name, "trav_prefetch")) slot = &g_opt_trav_prefetch;
  else if (!strcmp(name, "aa_variant")) slot = &g_opt_aa_variant;
  else if (!strcmp(name, "small_stage")) slot = &g_opt_small_stage;
  else if (!strcmp(name, "aa_tiptab")) slot = &g_opt_aa_tiptab;
  else if (!strcmp(name, "batch")) slot = &g_opt_batch;
  if (!slot) return -1;
  int old = *slot;
  *slot = value;
  return old;
}

static size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

bool dev_ctx_init(PartCtx * c)
{
  pll_partition_t * p = &c->pub;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
  {
    cudaGetLastError();
    pllb200_set_error(PLLB200_ERROR_DEVICE,
                      "No usable CUDA device (%s): libpll_b200 has no CPU fallback.",
                      e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return false;
  }
  int dev = t_device;
  if (dev < 0) CUDA_TRY(cudaGetDevice(&dev), false);
  c->device = dev;
  CUDA_TRY(cudaSetDevice(dev), false);
  CUDA_TRY(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, dev), false);

  const unsigned int K = p->rate_cats, s = p->states, sp = p->states_padded;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  const bool rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;

  c->clv_first = pattern_tip ? p->tips : 0;
  c->clv_elems = (size_t)c->sites_alloc * K * sp;
  c->clv_stride = round_up(c->clv_elems, 32);           /* 256 B */
  c->scaler_elems = (size_t)c->sites_alloc * (rate_scalers ? K : 1);
  c->scaler_stride = round_up(c->scaler_elems, 64);     /* 256 B */
  c->pmat_elems = (size_t)K * s * sp;
  c->tip_stride = round_up(c->sites_alloc, 256);
  c->eigen_stride = (size_t)2 * s * sp + 2 * sp;

  c->d_clv = nullptr; c->d_scaler = nullptr; c->d_pmat = nullptr; c->d_pmat_il = nullptr; c->d_tiptab = nullptr; c->tiptab_stale = false;
  c->d_pmat_cols = nullptr; c->d_tipcols = nullptr; c->d_split_tickets = nullptr;
  c->d_tipchars = nullptr; c->d_weights = nullptr; c->d_invariant = nullptr; c->d_eigen = nullptr;
  c->d_tipmap = nullptr; c->d_persite = nullptr; c->d_ancestral = nullptr; c->d_brlen = nullptr; c->d_sumfrag = nullptr;
  c->d_multi = nullptr; c->multi_cap = 0; c->ev_batch = nullptr;
  for (int r = 0; r < 8; ++r) c->xchg_peer[r] = nullptr;
  c->xchg_rank = 0; c->xchg_nranks = 0; c->h_stage = nullptr; c->d_stage = nullptr;
  c->h_xred = nullptr; c->h_xred_dev = nullptr; c->xred_seq = 0;
  c->h_small = nullptr; c->h_small_dev = nullptr; c->small_head = 0;
  c->d_partials = nullptr; c->d_ticket = nullptr; c->d_result = nullptr; c->h_result = nullptr;
  c->own_stream = nullptr;

  CUDA_TRY(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking), false);
  c->stream = c->own_stream;

  const bool lean_shape = s == 4 && K == 4 && !rate_scalers;
  c->scratch_rows = lean_shape && g_opt_scratch_rows > 0 ? (unsigned int)g_opt_scratch_rows : 0;
  c->scratch_mats = c->scratch_rows ? 512 : 0;
  c->d_score = nullptr; c->score_cap = 0; c->h_score = nullptr;
  c->deriv_last_table = nullptr; c->deriv_reverse = 0;
  const unsigned int nclv = p->nodes - c->clv_first + c->scratch_rows;
  if (nclv)
  {
    const size_t bytes = (size_t)nclv * c->clv_stride * sizeof(double);
    e = cudaMalloc(&c->d_clv, bytes);
    if (e != cudaSuccess)
    {
      cudaGetLastError();
      pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Unable to allocate %.2f GB of HBM for CLVs (%s).",
                        bytes / 1e9, cudaGetErrorString(e));
      dev_ctx_destroy(c);
      return false;
    }
    /* padded state columns must read as 0 forever; inner CLVs are fully overwritten otherwise */
    if (sp != s || !pattern_tip) CUDA_TRY(cudaMemsetAsync(c->d_clv, 0, bytes, c->stream), false);
    for (unsigned int i = c->clv_first; i < p->nodes; ++i) p->clv[i] = clv_dev(c, i);
  }
  if (p->scale_buffers)
  {
    const size_t bytes = (size_t)(p->scale_buffers + c->scratch_rows) * c->scaler_stride * sizeof(unsigned int);
    CUDA_TRY(cudaMalloc(&c->d_scaler, bytes), false);
    CUDA_TRY(cudaMemsetAsync(c->d_scaler, 0, bytes, c->stream), false);
    for (unsigned int i = 0; i < p->scale_buffers; ++i) p->scale_buffer[i] = scaler_dev(c, (int)i);
  }
  if (p->prob_matrices)
  {
    const size_t bytes = (size_t)(p->prob_matrices + c->scratch_mats) * c->pmat_elems * sizeof(double);
    CUDA_TRY(cudaMalloc(&c->d_pmat, bytes), false);
    CUDA_TRY(cudaMemsetAsync(c->d_pmat, 0, bytes, c->stream), false);
    for (unsigned int i = 0; i < p->prob_matrices; ++i) p->pmatrix[i] = c->d_pmat + i * c->pmat_elems;
    if (s == 4)
    {
      CUDA_TRY(cudaMalloc(&c->d_pmat_il, bytes), false);
      CUDA_TRY(cudaMemsetAsync(c->d_pmat_il, 0, bytes, c->stream), false);
      const size_t cbytes = (size_t)(p->prob_matrices + c->scratch_mats) * 5 * K * 4 * sizeof(double);
      CUDA_TRY(cudaMalloc(&c->d_pmat_cols, cbytes), false);
      CUDA_TRY(cudaMemsetAsync(c->d_pmat_cols, 0, cbytes, c->stream), false);
      const size_t tbytes = ((size_t)c->sites_alloc / 256 + 1) * sizeof(unsigned int);
      CUDA_TRY(cudaMalloc(&c->d_split_tickets, tbytes), false);
      CUDA_TRY(cudaMemsetAsync(c->d_split_tickets, 0, tbytes, c->stream), false);
    }
    else if (s == 20)
    {
      /* DMMA B-fragment image for the 20-state traversal kernel: [m][k][ntile 3][kstep 5][lane 32] */
      const size_t fbytes = (size_t)p->prob_matrices * K * 480 * sizeof(double);
      CUDA_TRY(cudaMalloc(&c->d_pmat_il, fbytes), false);
      CUDA_TRY(cudaMemsetAsync(c->d_pmat_il, 0, fbytes, c->stream), false);
      if (pattern_tip && sp == 20 && !rate_scalers)
      {
        CUDA_TRY(cudaMalloc(&c->d_tiptab, fbytes), false);
        CUDA_TRY(cudaMemsetAsync(c->d_tiptab, 0, fbytes, c->stream), false);
      }
    }
  }
  if (pattern_tip && p->tips)
  {
    CUDA_TRY(cudaMalloc(&c->d_tipchars, (size_t)p->tips * c->tip_stride), false);
    CUDA_TRY(cudaMemsetAsync(c->d_tipchars, s == 4 ? 15 : 0, (size_t)p->tips * c->tip_stride,
                             c->stream), false);
    if (s == 4)
    {
      CUDA_TRY(cudaMalloc(&c->d_tipcols, (size_t)p->tips * c->tip_stride), false);
      CUDA_TRY(cudaMemsetAsync(c->d_tipcols, L4_TIPCOL_N, (size_t)p->tips * c->tip_stride, c->stream), false); /* code 15 = N */
    }
    CUDA_TRY(cudaMalloc(&c->d_tipmap, 256 * sizeof(unsigned long long)), false);
    CUDA_TRY(cudaMemsetAsync(c->d_tipmap, 0, 256 * sizeof(unsigned long long), c->stream), false);
  }
  /* padded to whole 256-site tiles: the Newton kernel fetches weights by 16-byte-granular bulk copies */
  CUDA_TRY(cudaMalloc(&c->d_weights, sizeof(unsigned int) * round_up(c->sites_alloc + 4, 256)), false);
  CUDA_TRY(cudaMemsetAsync(c->d_weights, 0, sizeof(unsigned int) * round_up(c->sites_alloc + 4, 256),
                           c->stream), false);
  CUDA_TRY(cudaMalloc(&c->d_eigen, sizeof(double) * c->eigen_stride * p->rate_matrices), false);
  CUDA_TRY(cudaMemsetAsync(c->d_eigen, 0, sizeof(double) * c->eigen_stride * p->rate_matrices,
                           c->stream), false);

  c->stage_size = 4u << 20;
  c->stage_head = 0;
  CUDA_TRY(cudaHostAlloc(&c->h_stage, c->stage_size, cudaHostAllocDefault), false);
  CUDA_TRY(cudaHostAlloc(&c->h_small, PLLB200_SMALL_RING, cudaHostAllocMapped), false);
  CUDA_TRY(cudaHostGetDevicePointer((void **)&c->h_small_dev, c->h_small, 0), false);
  CUDA_TRY(cudaMalloc(&c->d_stage, c->stage_size), false);

  CUDA_TRY(cudaMalloc(&c->d_partials, sizeof(double) * 4 * PLLB200_MAX_RED_BLOCKS), false);
  CUDA_TRY(cudaMalloc(&c->d_ticket, sizeof(unsigned int) * 4), false);
  CUDA_TRY(cudaMemsetAsync(c->d_ticket, 0, sizeof(unsigned int) * 4, c->stream), false);
  CUDA_TRY(cudaMalloc(&c->d_result, sizeof(double) * 4 * PLLB200_RESULT_SLOTS), false);
  CUDA_TRY(cudaMemsetAsync(c->d_result, 0, sizeof(double) * 4 * PLLB200_RESULT_SLOTS, c->stream), false);
  /* results and, behind them, one completion tag per slot */
  CUDA_TRY(cudaHostAlloc(&c->h_result, (sizeof(double) * 4 + sizeof(unsigned long long)) * PLLB200_RESULT_SLOTS,
                         cudaHostAllocMapped), false);
  memset(c->h_result, 0, (sizeof(double) * 4 + sizeof(unsigned long long)) * PLLB200_RESULT_SLOTS);
  CUDA_TRY(cudaHostGetDevicePointer(&c->h_result_dev, c->h_result, 0), false);
  c->h_tag = reinterpret_cast<volatile unsigned long long *>(c->h_result + 4 * PLLB200_RESULT_SLOTS);
  c->h_tag_dev = reinterpret_cast<unsigned long long *>(c->h_result_dev + 4 * PLLB200_RESULT_SLOTS);
  memset(c->slot_tag, 0, sizeof(c->slot_tag));
  c->tag_counter = 0;
  c->next_slot = 0;
  CUDA_TRY(cudaStreamSynchronize(c->stream), false);
  return true;
}

void dev_ctx_destroy(PartCtx * c)
{
  if (c->device >= 0) cudaSetDevice(c->device);
  if (c->own_stream) cudaStreamSynchronize(c->own_stream);
  if (c->stream && c->stream != c->own_stream) cudaStreamSynchronize(c->stream);
  for (auto & kv : c->sumtables) cudaFree(kv.second);
  c->sumtables.clear();
  cudaFree(c->d_clv); cudaFree(c->d_scaler); cudaFree(c->d_pmat); cudaFree(c->d_pmat_il); cudaFree(c->d_tiptab);
  cudaFree(c->d_pmat_cols); cudaFree(c->d_tipcols); cudaFree(c->d_split_tickets); cudaFree(c->d_score);
  if (c->h_score) cudaFreeHost(c->h_score);
  cudaFree(c->d_tipchars); cudaFree(c->d_weights); cudaFree(c->d_invariant); cudaFree(c->d_eigen);
  cudaFree(c->d_tipmap); cudaFree(c->d_persite); cudaFree(c->d_ancestral); cudaFree(c->d_brlen); cudaFree(c->d_sumfrag); cudaFree(c->d_sumgen);
  cudaFree(c->d_multi);
  if (c->ev_batch) cudaEventDestroy(c->ev_batch);
  for (int r = 0; r < 8; ++r)
    if (c->xchg_peer[r])
    {
      if (c->xchg_nranks > 0 && r != c->xchg_rank) cudaIpcCloseMemHandle(c->xchg_peer[r]); else cudaFree(c->xchg_peer[r]);
    } cudaFree(c->d_stage); cudaFree(c->d_partials);
  cudaFree(c->d_ticket); cudaFree(c->d_result);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->h_small) cudaFreeHost(c->h_small);
  if (c->h_xred) cudaFreeHost(c->h_xred);
  if (c->h_result) cudaFreeHost(c->h_result);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  cudaGetLastError();
  c->d_clv = nullptr;
  c->own_stream = nullptr;
}

/* Copy `bytes` of host data into the stream-ordered staging ring and return the device address the
 * kernels queued after this call may read.  The pinned half is reused lap by lap; a lap boundary
 * waits for the stream so earlier async copies have consumed their source. */
void * dev_stage(PartCtx * c, const void * src, size_t bytes)
{
  if (g_batch && g_batch->lead != c) return dev_stage(g_batch->lead, src, bytes);
  const size_t need = round_up(bytes, 256);
  /* a launch stages a handful of arrays and holds their addresses until it is queued: keep one array below
   * an eighth of the ring, so that a lap can never land on an address handed out for the same launch */
  if (need * 8 > c->stage_size)
  {
    /* grow: rare (op lists beyond 16k ops).  Blocks handed out from the old ring stay valid until the next
     * stream synchronisation of this partition (dev_flush). */
    c->retired_host.push_back(c->h_stage);
    c->retired_dev.push_back(c->d_stage);
    c->stage_size = round_up(need * 16, 1 << 20);
    if (!pllb200_cuda_ok(cudaHostAlloc(&c->h_stage, c->stage_size, cudaHostAllocDefault), "stage grow") ||
        !pllb200_cuda_ok(cudaMalloc(&c->d_stage, c->stage_size), "stage grow"))
      return nullptr;
    c->stage_head = 0;
  }
  if (c->stage_head + need > c->stage_size)
  {
    if (!pllb200_cuda_ok(cudaStreamSynchronize(c->stream), "stage lap sync")) return nullptr;
    c->stage_head = 0;
  }
  char * h = c->h_stage + c->stage_head;
  char * d = c->d_stage + c->stage_head;
  memcpy(h, src, bytes);
  if (!pllb200_cuda_ok(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, c->stream), "stage copy"))
    return nullptr;
  c->stage_head += need;
  g_h2d_bytes += bytes;
  return d;
}

/* Few bytes (one or two resolved ops, a matrix index and its length): written into a ring of MAPPED
 * pinned memory that the kernel reads directly — no copy call.  A branch of a 100-partition analysis
 * queued 250 such copies (profiles/README.md); the read costs the kernel one PCIe round trip. */
int g_opt_small_stage = 1; /* 0 never, 1 for partitions whose kernels are shorter than the round trip saved, 2 always */
void * dev_stage_small(PartCtx * c, const void * src, size_t bytes)
{
  if (g_batch) return dev_stage(g_batch->lead, src, bytes); /* batched kernels read argument arrays from HBM */
  /* measured (BLO pass, 200 taxa): 1 k sites 45.3 -> 40.5 us per branch, 1 M sites 259.7 -> 269.2 us — a
   * long kernel would rather have the copy overlap its predecessor than wait ~2.7 us at its start */
  if (bytes > PLLB200_SMALL_MAX || !g_opt_small_stage || (g_opt_small_stage == 1 && c->clv_elems > (1u << 20)))
    return dev_stage(c, src, bytes);
  const size_t need = round_up(bytes, 64);
  if (c->small_head + need > PLLB200_SMALL_RING)
  {
    /* one lap = thousands of launches; everything that read the old entries must be done */
    if (!pllb200_cuda_ok(cudaStreamSynchronize(c->stream), "small stage lap sync")) return nullptr;
    c->small_head = 0;
  }
  memcpy(c->h_small + c->small_head, src, bytes);
  void * d = c->h_small_dev + c->small_head;
  c->small_head += need;
  g_h2d_bytes += bytes;
  return d;
}

/* big host arrays: chunk through the staging ring (pageable sources, async-safe) */
static bool upload_big(PartCtx * c, void * dst, const void * src, size_t bytes)
{
  const size_t chunk = c->stage_size / 2;
  const char * s = (const char *)src;
  char * d = (char *)dst;
  while (bytes)
  {
    const size_t n = std::min(bytes, chunk);
    void * st = dev_stage(c, s, n);
    if (!st) return false;
    if (!pllb200_cuda_ok(cudaMemcpyAsync(d, st, n, cudaMemcpyDeviceToDevice, xstream(c)), "upload_big"))
      return false;
    s += n; d += n; bytes -= n;
  }
  return true;
}

bool dev_sync_inputs(PartCtx * c)
{
  pll_partition_t * p = &c->pub;
  if (c->any_tip_dirty && c->d_tipchars && p->tipchars)
  {
    for (unsigned int t = 0; t < p->tips; ++t)
      if (c->tip_dirty[t])
      {
        if (!upload_big(c, c->d_tipchars + (size_t)t * c->tip_stride, p->tipchars[t], c->sites_alloc))
          return false;
        if (!launch_tipcols_encode(c, t)) return false;
        c->tip_dirty[t] = 0;
      }
    c->any_tip_dirty = false;
  }
  if (c->tipmap_dirty && c->d_tipmap && p->tipmap)
  {
    void * st = dev_stage(c, p->tipmap, 256 * sizeof(unsigned long long));
    if (!st) return false;
    CUDA_TRY(cudaMemcpyAsync(c->d_tipmap, st, 256 * sizeof(unsigned long long),
                             cudaMemcpyDeviceToDevice, xstream(c)), false);
    c->tipmap_dirty = false;
    ++c->tipmap_gen;
    if (c->d_tiptab) c->tiptab_stale = true; /* tables of already computed matrices lack the new codes */
  }
  if (c->weights_dirty)
  {
    if (!upload_big(c, c->d_weights, p->pattern_weights, sizeof(unsigned int) * c->sites_alloc))
      return false;
    c->weights_dirty = false;
  }
  if (c->invariant_dirty && p->invariant)
  {
    if (!c->d_invariant) CUDA_TRY(cudaMalloc(&c->d_invariant, sizeof(int) * p->sites), false);
    if (!upload_big(c, c->d_invariant, p->invariant, sizeof(int) * p->sites)) return false;
    c->invariant_dirty = false;
  }
  return true;
}

/* Model block of every rate matrix referenced by params_indices: recompute the decomposition if
 * the caller invalidated it (models.c:421-428) and mirror it when its version moved. */
bool dev_sync_eigen(PartCtx * c, const unsigned int * params_indices)
{
  pll_partition_t * p = &c->pub;
  const unsigned int s = p->states, sp = p->states_padded;
  for (unsigned int k = 0; k < p->rate_cats; ++k)
  {
    const unsigned int m = params_indices[k];
    if (m >= p->rate_matrices)
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "params index %u out of range", m);
      return false;
    }
    if (!p->eigen_decomp_valid[m] && !pll_update_eigen(p, m)) return false;
    if (c->eigen_version_dev[m] == c->eigen_version[m]) continue;
    std::vector<double> blk(c->eigen_stride, 0.0);
    memcpy(blk.data(), p->eigenvecs[m], sizeof(double) * s * sp);
    memcpy(blk.data() + s * sp, p->inv_eigenvecs[m], sizeof(double) * s * sp);
    memcpy(blk.data() + 2 * s * sp, p->eigenvals[m], sizeof(double) * s);
    memcpy(blk.data() + 2 * s * sp + sp, p->frequencies[m], sizeof(double) * s);
    void * st = dev_stage(c, blk.data(), sizeof(double) * c->eigen_stride);
    if (!st) return false;
    CUDA_TRY(cudaMemcpyAsync(c->d_eigen + (size_t)m * c->eigen_stride, st,
                             sizeof(double) * c->eigen_stride, cudaMemcpyDeviceToDevice, xstream(c)),
             false);
    c->eigen_version_dev[m] = c->eigen_version[m];
  }
  return true;
}

int dev_alloc_slot(PartCtx * c)
{
  /* all slots in flight -> drain */
  if (c->pending.size() >= PLLB200_RESULT_SLOTS - 2) dev_flush(c);
  int s = c->next_slot;
  c->next_slot = (c->next_slot + 1) % PLLB200_RESULT_SLOTS;
  c->slot_tag[s] = ++c->tag_counter;
  return s;
}

void dev_slot_tag(PartCtx * c, int slot, unsigned long long ** tag_host, unsigned long long * tag)
{
  *tag_host = c->h_tag_dev + slot;
  *tag = c->slot_tag[slot];
}

/* Poll the slot's completion tag in mapped host memory.  Every so often the stream is queried: a failed
 * launch or a kernel fault would otherwise never show up, and an idle stream without the tag means the
 * slot was never written. */
bool dev_wait_slot(PartCtx * c, int slot)
{
  const unsigned long long want = c->slot_tag[slot];
  /* in batch mode the kernels of every member run on the lead partition's stream */
  const cudaStream_t stream = xstream(c);
  for (unsigned long long spins = 0;; ++spins)
  {
    if (c->h_tag[slot] == want) return true;
    if ((spins & 0x3fffull) == 0x3fffull)
    {
      const cudaError_t e = cudaStreamQuery(stream);
      if (e == cudaErrorNotReady) continue;
      if (e != cudaSuccess) return pllb200_cuda_ok(e, "waiting for a device result");
      /* stream idle: the tag is there now, or nobody is going to write it */
      if (c->h_tag[slot] == want) return true;
      CUDA_TRY(cudaStreamSynchronize(stream), false);
      if (c->h_tag[slot] == want) return true;
      pllb200_set_error(PLLB200_ERROR_DEVICE, "A device result was awaited that no queued kernel writes (slot %d).", slot);
      return false;
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
}

bool dev_wait_results(PartCtx * c)
{
  if (c->pending.empty()) return true;
  /* one stream, in-order: the last queued result is the last to arrive */
  if (!dev_wait_slot(c, c->pending.back().slot)) return false;
  for (const PendingResult & r : c->pending) *r.host_dst = c->h_result[r.slot * 4 + r.comp];
  g_d2h_bytes += sizeof(double) * c->pending.size();
  c->pending.clear();
  return true;
}

bool dev_flush(PartCtx * c)
{
  CUDA_TRY(cudaStreamSynchronize(c->stream), false);
  if (g_batch && g_batch->lead != c) CUDA_TRY(cudaStreamSynchronize(g_batch->lead->stream), false);
  for (void * p : c->retired_dev) cudaFree(p);
  for (void * p : c->retired_host) cudaFreeHost(p);
  c->retired_dev.clear();
  c->retired_host.clear();
  for (const PendingResult & r : c->pending) *r.host_dst = c->h_result[r.slot * 4 + r.comp];
  g_d2h_bytes += sizeof(double) * c->pending.size(); /* written by the kernel into mapped host memory */
  c->pending.clear();
  return true;
}

/* ---- additive C ABI: transfers, stream control ------------------------------------------- */

static bool check_ctx(pll_partition_t * p)
{
  if (!p || ctx_of(p)->magic != PLLB200_MAGIC)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Not a libpll_b200 partition.");
    return false;
  }
  cudaSetDevice(ctx_of(p)->device);
  return true;
}

extern "C" int pllb200_download_clv(pll_partition_t * p, unsigned int clv_index, double * dst)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (clv_index >= p->nodes || clv_index < c->clv_first)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "CLV %u has no device buffer.", clv_index);
    return PLL_FAILURE;
  }
  CUDA_TRY(cudaMemcpyAsync(dst, clv_dev(c, clv_index), sizeof(double) * c->clv_elems,
                           cudaMemcpyDeviceToHost, c->stream), PLL_FAILURE);
  g_d2h_bytes += sizeof(double) * c->clv_elems;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_upload_clv(pll_partition_t * p, unsigned int clv_index, const double * src)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (clv_index >= p->nodes || clv_index < c->clv_first)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "CLV %u has no device buffer.", clv_index);
    return PLL_FAILURE;
  }
  return upload_big(c, clv_dev(c, clv_index), src, sizeof(double) * c->clv_elems) ? PLL_SUCCESS
                                                                                  : PLL_FAILURE;
}

extern "C" int pllb200_download_scaler(pll_partition_t * p, unsigned int scaler_index,
                                       unsigned int * dst)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (scaler_index >= p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid scaler index %u.", scaler_index);
    return PLL_FAILURE;
  }
  CUDA_TRY(cudaMemcpyAsync(dst, scaler_dev(c, (int)scaler_index),
                           sizeof(unsigned int) * c->scaler_elems, cudaMemcpyDeviceToHost,
                           c->stream), PLL_FAILURE);
  g_d2h_bytes += sizeof(unsigned int) * c->scaler_elems;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_download_pmatrix(pll_partition_t * p, unsigned int matrix_index, double * dst)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (matrix_index >= p->prob_matrices)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid matrix index %u.", matrix_index);
    return PLL_FAILURE;
  }
  CUDA_TRY(cudaMemcpyAsync(dst, c->d_pmat + (size_t)matrix_index * c->pmat_elems,
                           sizeof(double) * c->pmat_elems, cudaMemcpyDeviceToHost, c->stream),
           PLL_FAILURE);
  g_d2h_bytes += sizeof(double) * c->pmat_elems;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_download_pmatrices(pll_partition_t * p, double * dst)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  const size_t bytes = sizeof(double) * c->pmat_elems * p->prob_matrices;
  CUDA_TRY(cudaMemcpyAsync(dst, c->d_pmat, bytes, cudaMemcpyDeviceToHost, c->stream), PLL_FAILURE);
  g_d2h_bytes += bytes;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_upload_pmatrices(pll_partition_t * p, const double * src)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (!upload_big(c, c->d_pmat, src, sizeof(double) * c->pmat_elems * p->prob_matrices)) return PLL_FAILURE;
  return launch_pmatrix_reinterleave(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_upload_scaler(pll_partition_t * p, unsigned int scaler_index, const unsigned int * src)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (scaler_index >= p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid scaler index %u.", scaler_index);
    return PLL_FAILURE;
  }
  return upload_big(c, scaler_dev(c, (int)scaler_index), src, sizeof(unsigned int) * c->scaler_elems)
             ? PLL_SUCCESS
             : PLL_FAILURE;
}

/* ---- multi-GPU exchange buffers (one process per GPU, CUDA IPC over NVLink/NVSwitch) ---------- */

#define XCHG_BYTES 32768 /* Newton exchange (k_newton.cu: NW_XCHG_DOUBLES * 8 = 640 B) + the reduce inboxes below */
/* all-reduce inboxes: [2 parities][8 ranks]{values[XRED_MAX], seq, -} from this double offset on */
#define XRED_BASE 128
#define XRED_MAX 128
#define XRED_SLOT (XRED_MAX + 2)
#define XRED_SPIN_LIMIT 40000000ull

struct XredArgs
{
  double * inbox[8];
  double * d_in;             /* this rank's values (staged) */
  double * out;              /* mapped host result */
  unsigned long long * tag_host;
  unsigned long long tag, seq;
  int rank, nranks, n, op;
};

/* All-reduce of n <= XRED_MAX doubles over NVLink peer memory (the collective step of
 * pllmod_treeinfo_compute_loglh, treeinfo.c:1061-1067, and of the derivative callback, pll_optimize.c:1270-1286):
 * every rank pushes its values into every rank's inbox (st.release.sys), waits for all ranks in its own inbox
 * (ld.acquire.sys) and combines them in rank order — the result is bit-identical on all GPUs.  One CTA.
 * `seq` comes from the host (same on all ranks: the call is collective), so a timed-out call cannot leave a
 * stale counter behind; slots alternate with its parity. */
__global__ void xchg_reduce_kernel(XredArgs a)
{
  const int i = threadIdx.x;
  const unsigned int par = (unsigned int)(a.seq & 1ull);
  __shared__ int s_fail;
  if (i == 0) s_fail = 0;
  if (i < a.n)
  {
    const double v = a.d_in[i];
    for (int r = 0; r < a.nranks; ++r) a.inbox[r][XRED_BASE + (par * 8 + a.rank) * XRED_SLOT + i] = v;
  }
  __threadfence_system();
  __syncthreads();
  if (i == 0)
  {
    for (int r = 0; r < a.nranks; ++r)
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(a.inbox[r] + XRED_BASE + (par * 8 + a.rank) * XRED_SLOT + XRED_MAX),
                   "l"(a.seq)
                   : "memory");
    unsigned long long spins = 0;
    for (int r = 0; r < a.nranks && !s_fail; ++r)
    {
      const double * flag = a.inbox[a.rank] + XRED_BASE + (par * 8 + r) * XRED_SLOT + XRED_MAX;
      unsigned long long v;
      do
      {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
        if (v != a.seq && ++spins > XRED_SPIN_LIMIT) s_fail = 1;
      } while (v != a.seq && !s_fail);
    }
  }
  __syncthreads();
  if (i < a.n)
  {
    double acc = 0.0;
    for (int r = 0; r < a.nranks; ++r)
    {
      const double v = __ldcg(a.inbox[a.rank] + XRED_BASE + (par * 8 + r) * XRED_SLOT + i);
      acc = r == 0 ? v : (a.op == 0 ? acc + v : (a.op == 1 ? fmax(acc, v) : fmin(acc, v)));
    }
    a.out[i] = s_fail ? __longlong_as_double(0x7ff8000000000000ll) : acc;
  }
  __threadfence_system();
  __syncthreads();
  if (i == 0) *a.tag_host = a.tag;
}

extern "C" int pllb200_exchange_create(pll_partition_t * p, void * ipc_handle_out)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (c->xchg_nranks != 0)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "exchange_create: the partition is already connected.");
    return PLL_FAILURE;
  }
  if (!c->xchg_peer[0])
  {
    double * buf = nullptr;
    CUDA_TRY(cudaMalloc(&buf, XCHG_BYTES), PLL_FAILURE);
    CUDA_TRY(cudaMemset(buf, 0, XCHG_BYTES), PLL_FAILURE);
    c->xchg_peer[0] = buf; /* parked in slot 0 until the rank is known */
    c->xchg_rank = 0;
    c->xchg_nranks = 0;
  }
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaIpcGetMemHandle(&h, c->xchg_peer[0]), PLL_FAILURE);
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(ipc_handle_out, &h, sizeof(h));
  return PLL_SUCCESS;
}

extern "C" int pllb200_exchange_connect(pll_partition_t * p, int rank, int nranks, const void * all_handles)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks || !c->xchg_peer[0] || c->xchg_nranks != 0)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "exchange_connect: need 1..8 ranks and a buffer from exchange_create.");
    return PLL_FAILURE;
  }
  double * own = c->xchg_peer[0];
  c->xchg_peer[0] = nullptr;
  for (int r = 0; r < nranks; ++r)
  {
    if (r == rank)
    {
      c->xchg_peer[r] = own;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char *)all_handles + (size_t)r * sizeof(h), sizeof(h));
    void * ptr = nullptr;
    if (!pllb200_cuda_ok(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle"))
    {
      /* undo: close what was opened, park the own buffer in slot 0 again (destroy then frees it) */
      for (int q = 0; q < r; ++q)
        if (q != rank && c->xchg_peer[q]) cudaIpcCloseMemHandle(c->xchg_peer[q]);
      for (int q = 0; q < 8; ++q) c->xchg_peer[q] = nullptr;
      c->xchg_peer[0] = own;
      return PLL_FAILURE;
    }
    c->xchg_peer[r] = (double *)ptr;
  }
  c->xchg_rank = rank;
  c->xchg_nranks = nranks;
  return PLL_SUCCESS;
}

/* parallel_reduce_cb-compatible collective (pll_tree.h: void (*)(void *, double *, size_t, int)); context = a
 * partition connected with pllb200_exchange_connect.  op: 0 sum, 1 max, 2 min (PLLMOD_COMMON_REDUCE_*). */
extern "C" void pllb200_exchange_reduce_cb(void * context, double * data, size_t count, int op)
{
  pll_partition_t * p = static_cast<pll_partition_t *>(context);
  if (!check_ctx(p)) return;
  PartCtx * c = ctx_of(p);
  if (c->xchg_nranks < 2) return;
  if (!c->h_xred)
  {
    if (!pllb200_cuda_ok(cudaHostAlloc(&c->h_xred, (XRED_MAX + 1) * sizeof(double), cudaHostAllocMapped), "reduce buffer") ||
        !pllb200_cuda_ok(cudaHostGetDevicePointer((void **)&c->h_xred_dev, c->h_xred, 0), "reduce buffer"))
      return;
    c->h_xred[XRED_MAX] = 0.0;
  }
  for (size_t off = 0; off < count; off += XRED_MAX)
  {
    const int n = (int)std::min<size_t>(XRED_MAX, count - off);
    double * d_in = static_cast<double *>(dev_stage_small(c, data + off, sizeof(double) * n));
    if (!d_in) return;
    XredArgs a;
    for (int r = 0; r < 8; ++r) a.inbox[r] = c->xchg_peer[r];
    a.d_in = d_in;
    a.out = c->h_xred_dev;
    a.tag_host = reinterpret_cast<unsigned long long *>(c->h_xred_dev + XRED_MAX);
    a.seq = ++c->xred_seq;
    a.tag = a.seq;
    a.rank = c->xchg_rank;
    a.nranks = c->xchg_nranks;
    a.n = n;
    a.op = op;
    xchg_reduce_kernel<<<1, XRED_MAX, 0, c->stream>>>(a);
    ++g_kernel_launches;
    if (!pllb200_cuda_ok(cudaGetLastError(), "xchg_reduce_kernel launch")) return;
    volatile unsigned long long * tag = reinterpret_cast<volatile unsigned long long *>(c->h_xred + XRED_MAX);
    for (unsigned long long spins = 0; *tag != a.tag; ++spins)
      if ((spins & 0xfffffull) == 0xfffffull && cudaStreamQuery(c->stream) != cudaErrorNotReady)
      {
        if (*tag == a.tag) break;
        if (!pllb200_cuda_ok(cudaStreamSynchronize(c->stream), "exchange reduce")) return;
        break;
      }
    for (int i = 0; i < n; ++i) data[off + i] = c->h_xred[i];
    g_d2h_bytes += sizeof(double) * n;
    if (data[off] != data[off]) pllb200_set_error(PLLB200_ERROR_DEVICE, "exchange all-reduce: a peer did not answer");
  }
}

extern "C" int pllb200_exchange_ranks(pll_partition_t * p)
{
  if (!p || ctx_of(p)->magic != PLLB200_MAGIC) return 0;
  return ctx_of(p)->xchg_nranks;
}

extern "C" int pllb200_download_sumtable(pll_partition_t * p, const double * key, double * dst)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  auto it = c->sumtables.find(key);
  if (it == c->sumtables.end())
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Unknown sumtable buffer.");
    return PLL_FAILURE;
  }
  CUDA_TRY(cudaMemcpyAsync(dst, it->second, sizeof(double) * c->clv_elems, cudaMemcpyDeviceToHost,
                           c->stream), PLL_FAILURE);
  g_d2h_bytes += sizeof(double) * c->clv_elems;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_sync_host_arrays(pll_partition_t * p)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  std::fill(c->tip_dirty.begin(), c->tip_dirty.end(), 1);
  c->any_tip_dirty = c->weights_dirty = c->tipmap_dirty = true;
  c->invariant_dirty = p->invariant != nullptr;
  unsigned int sum = 0;
  for (unsigned int i = 0; i < p->sites; ++i) sum += p->pattern_weights[i];
  p->pattern_weight_sum = sum;
  return dev_sync_inputs(c) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_partition_synchronize(pll_partition_t * p)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  return dev_flush(ctx_of(p)) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_partition_wait_results(pll_partition_t * p)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  return dev_wait_results(ctx_of(p)) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_partition_set_stream(pll_partition_t * p, void * cuda_stream)
{
  if (!check_ctx(p)) return PLL_FAILURE;
  PartCtx * c = ctx_of(p);
  if (!dev_flush(c)) return PLL_FAILURE;
  c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
  return PLL_SUCCESS;
}

extern "C" uint64_t pllb200_result_device_ptr(pll_partition_t * p)
{
  if (!check_ctx(p)) return 0;
  return (uint64_t)(uintptr_t)ctx_of(p)->d_result;
}


/* ---- batch mode --------------------------------------------------------------------------------- */
thread_local BatchCtx * g_batch = nullptr;
int g_opt_batch = 1; /* bit 0: on; bits 1.. : keep kind (bit - 1) out of the collector (debugging) */

bool batch_eligible(const PartCtx * c)
{
  const pll_partition_t * p = &c->pub;
  if (p->attributes & (PLL_ATTRIB_RATE_SCALERS | PLL_ATTRIB_AB_MASK)) return false;
  if (p->states == 4) return p->rate_cats == 4;
  if (p->states == 20)
    return p->states_padded == 20 && (p->rate_cats == 1 || p->rate_cats == 2 || p->rate_cats == 4 || p->rate_cats == 8) &&
           c->d_pmat_il != nullptr;
  return false;
}

void batch_collect(int kind, const void * args, size_t bytes, unsigned int blocks, size_t smem)
{
  BatchKind & k = g_batch->kind[kind];
  const unsigned char * b = static_cast<const unsigned char *>(args);
  k.bytes.insert(k.bytes.end(), b, b + bytes);
  ++k.count;
  if (blocks > k.max_blocks) k.max_blocks = blocks;
  if (smem > k.smem) k.smem = smem;
}

const void * batch_pool_add(const void * src, size_t bytes)
{
  std::vector<unsigned char> & pool = g_batch->pool;
  const size_t off = round_up(pool.size(), 64);
  pool.resize(off + bytes);
  memcpy(pool.data() + off, src, bytes);
  g_h2d_bytes += bytes;
  return reinterpret_cast<const void *>(BATCH_POOL_TAG + off);
}

bool batch_flush()
{
  BatchCtx * b = g_batch;
  if (!b) return true;
  bool ok = true;
  if (!b->pool.empty())
  {
    const unsigned long long d_pool = (unsigned long long)dev_stage(b->lead, b->pool.data(), b->pool.size());
    if (!d_pool) ok = false;
    g_h2d_bytes -= b->pool.size(); /* counted when added */
    /* pointer members that may hold pool tokens: the first one of TravArgs / AaArgs (ops), the first two
     * of PmatBatchPart (indices, lengths) */
    static const int kinds[4] = {BK_TRAV_DNA4, BK_TRAV_AA, BK_TRAV_AA4, BK_PMAT};
    static const int nptr[4] = {1, 1, 1, 2};
    for (int q = 0; q < 4; ++q)
    {
      BatchKind & k = b->kind[kinds[q]];
      if (!k.count) continue;
      const size_t stride = k.bytes.size() / k.count;
      for (unsigned int i = 0; i < k.count; ++i)
        for (int f = 0; f < nptr[q]; ++f)
        {
          unsigned long long * ptr = reinterpret_cast<unsigned long long *>(k.bytes.data() + i * stride) + f;
          if ((*ptr >> 60) == 7ull) *ptr = d_pool + (*ptr - BATCH_POOL_TAG);
        }
    }
    b->pool.clear();
  }
  static const int order[BK_COUNT] = {BK_PMAT, BK_TRAV_DNA4, BK_TRAV_AA4, BK_TRAV_AA, BK_SUM_DNA4};
  for (int q = 0; q < BK_COUNT && ok; ++q)
  {
    const int kind = order[q];
    BatchKind & k = b->kind[kind];
    if (!k.count) continue;
    const void * d_args = dev_stage(b->lead, k.bytes.data(), k.bytes.size());
    if (!d_args)
      ok = false;
    else
      switch (kind)
      {
        case BK_PMAT: ok = flush_pmat(b->lead, d_args, k); break;
        case BK_TRAV_DNA4: ok = flush_trav_dna4(b->lead, d_args, k); break;
        case BK_TRAV_AA: ok = flush_trav_aa(b->lead, d_args, k); break;
        case BK_TRAV_AA4: ok = flush_trav_aa4(b->lead, d_args, k); break;
        default: ok = flush_sum_dna4(b->lead, d_args, k); break;
      }
    k.bytes.clear();
    k.count = k.max_blocks = 0;
    k.smem = 0;
  }
  return ok;
}

/* Returns 1 when batch mode is on, 0 when these partitions cannot be batched (callers carry on with the
 * per-partition path), -1 on error. */
extern "C" int pllb200_batch_begin(pll_partition_t ** partitions, unsigned int count)
{
  if (g_batch || !(g_opt_batch & 1)) return 0;
  std::vector<PartCtx *> m;
  for (unsigned int i = 0; i < count; ++i)
  {
    if (!partitions[i]) continue;
    PartCtx * c = ctx_of(partitions[i]);
    if (c->magic != PLLB200_MAGIC || !batch_eligible(c)) return 0;
    if (!m.empty() && c->device != m[0]->device) return 0;
    m.push_back(c);
  }
  if (m.size() < 2) return 0;
  PartCtx * lead = m[0];
  if (!pllb200_cuda_ok(cudaSetDevice(lead->device), "cudaSetDevice")) return -1;
  for (size_t i = 1; i < m.size(); ++i)
  {
    PartCtx * c = m[i];
    if (c->stream == lead->stream) continue;
    if (!c->ev_batch &&
        !pllb200_cuda_ok(cudaEventCreateWithFlags(&c->ev_batch, cudaEventDisableTiming), "batch event"))
      return -1;
    if (!pllb200_cuda_ok(cudaEventRecord(c->ev_batch, c->stream), "batch event record") ||
        !pllb200_cuda_ok(cudaStreamWaitEvent(lead->stream, c->ev_batch, 0), "batch event wait"))
      return -1;
  }
  BatchCtx * b = new BatchCtx();
  b->lead = lead;
  b->members = m;
  for (int k = 0; k < BK_COUNT; ++k)
  {
    b->kind[k].count = b->kind[k].max_blocks = 0;
    b->kind[k].smem = 0;
  }
  g_batch = b;
  return 1;
}

extern "C" int pllb200_batch_flush(void) { return batch_flush() ? PLL_SUCCESS : PLL_FAILURE; }

extern "C" int pllb200_batch_active(void) { return g_batch != nullptr; }

extern "C" int pllb200_batch_end(void)
{
  BatchCtx * b = g_batch;
  if (!b) return PLL_SUCCESS;
  bool ok = batch_flush();
  g_batch = nullptr;
  PartCtx * lead = b->lead;
  if (!lead->ev_batch)
    ok = pllb200_cuda_ok(cudaEventCreateWithFlags(&lead->ev_batch, cudaEventDisableTiming), "batch event") && ok;
  if (lead->ev_batch)
  {
    ok = pllb200_cuda_ok(cudaEventRecord(lead->ev_batch, lead->stream), "batch event record") && ok;
    for (size_t i = 1; i < b->members.size(); ++i)
      if (b->members[i]->stream != lead->stream)
        ok = pllb200_cuda_ok(cudaStreamWaitEvent(b->members[i]->stream, lead->ev_batch, 0), "batch event wait") && ok;
  }
  delete b;
  return ok ? PLL_SUCCESS : PLL_FAILURE;
}

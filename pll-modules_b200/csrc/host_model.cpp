/*
 * host_model.cpp — host side of the libpll partition API: object lifecycle, tip encoding, model
 * setters, eigen-decomposition and discrete-Gamma rates.
 *
 * Everything here is O(states^3) or O(tips*sites) byte shuffling that feeds the device; the
 * likelihood arithmetic itself lives in the k_*.cu kernels.  Reference behaviour followed:
 *   pll_partition_create / destroy      libs/libpll/src/pll.c:424-868, 31-132, 870
 *   pll_set_tip_states / set_tip_clv    pll.c:875-1129 (charmap/tipmap: pll.c:156-422)
 *   pll_set_pattern_weights             pll.c:1131
 *   model setters                       models.c:445-495,546
 *   pll_update_eigen                    models.c:182-410 (symmetrised Q, zero-frequency elimination)
 *   invariant sites                     models.c:495-750
 *   pll_compute_gamma_cats              gamma.c:24-292 (AS32 / AS91 / AS70 / ACM291)
 */
#include "internal.h"

#include <cassert>
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <new>

/* ------------------------------------------------------------------------------------------ */
/* errors                                                                                     */
/* ------------------------------------------------------------------------------------------ */

extern "C" {
__thread int pll_errno = 0;
__thread char pll_errmsg[200] = {0};
}

void pllb200_set_error(int code, const char * fmt, ...)
{
  pll_errno = code;
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(pll_errmsg, sizeof(pll_errmsg), fmt, ap);
  va_end(ap);
}

bool pllb200_cuda_ok(cudaError_t e, const char * what)
{
  if (e == cudaSuccess) return true;
  pllb200_set_error(PLLB200_ERROR_DEVICE, "CUDA error %d (%s) in %.100s", (int)e,
                    cudaGetErrorString(e), what);
  return false;
}

extern "C" void pllmod_set_error(int errno_, const char * errmsg_fmt, ...)
{
  /* src/pllmod_common.c:42-51 */
  pll_errno = errno_;
  va_list ap;
  va_start(ap, errmsg_fmt);
  vsnprintf(pll_errmsg, sizeof(pll_errmsg), errmsg_fmt, ap);
  va_end(ap);
}

extern "C" void pllmod_reset_error(void)
{
  pll_errno = 0;
  pll_errmsg[0] = 0;
}

/* ------------------------------------------------------------------------------------------ */
/* memory + state maps                                                                         */
/* ------------------------------------------------------------------------------------------ */

extern "C" void * pll_aligned_alloc(size_t size, size_t alignment)
{
  void * mem = nullptr;
  if (alignment < sizeof(void *)) alignment = sizeof(void *);
  if (posix_memalign(&mem, alignment, size ? size : alignment)) mem = nullptr;
  return mem;
}

extern "C" void pll_aligned_free(void * ptr)
{
  if (ptr) sumtable_forget_key(ptr); /* a freed sumtable key releases its device shadow */
  free(ptr);
}

/* pll_map_bin / pll_map_nt / pll_map_aa live in state_maps.c (C99 designated initialisers). */

/* ------------------------------------------------------------------------------------------ */
/* partition lifecycle                                                                         */
/* ------------------------------------------------------------------------------------------ */

static void free_host_arrays(pll_partition_t * p)
{
  unsigned int i;
  free(p->eigen_decomp_valid);
  free(p->clv);
  free(p->pmatrix);
  free(p->scale_buffer);
  if (p->subst_params)
    for (i = 0; i < p->rate_matrices; ++i) pll_aligned_free(p->subst_params[i]);
  free(p->subst_params);
  if (p->frequencies)
    for (i = 0; i < p->rate_matrices; ++i) pll_aligned_free(p->frequencies[i]);
  free(p->frequencies);
  if (p->eigenvecs)
    for (i = 0; i < p->rate_matrices; ++i) pll_aligned_free(p->eigenvecs[i]);
  free(p->eigenvecs);
  if (p->inv_eigenvecs)
    for (i = 0; i < p->rate_matrices; ++i) pll_aligned_free(p->inv_eigenvecs[i]);
  free(p->inv_eigenvecs);
  if (p->eigenvals)
    for (i = 0; i < p->rate_matrices; ++i) pll_aligned_free(p->eigenvals[i]);
  free(p->eigenvals);
  free(p->rates);
  free(p->rate_weights);
  free(p->prop_invar);
  free(p->invariant);
  free(p->pattern_weights);
  if (p->tipchars)
    for (i = 0; i < p->tips; ++i) free(p->tipchars[i]);
  free(p->tipchars);
  free(p->charmap);
  free(p->tipmap);
}

extern "C" pll_partition_t * pll_partition_create(unsigned int tips, unsigned int clv_buffers,
                                                  unsigned int states, unsigned int sites,
                                                  unsigned int rate_matrices,
                                                  unsigned int prob_matrices,
                                                  unsigned int rate_cats,
                                                  unsigned int scale_buffers,
                                                  unsigned int attributes)
{
  if (__builtin_popcount(attributes & PLL_ATTRIB_ARCH_MASK) > 1)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Multiple architecture flags specified.");
    return nullptr;
  }
  if (attributes & PLL_ATTRIB_SITE_REPEATS)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "PLL_ATTRIB_SITE_REPEATS is not supported by the B200 library "
                      "(dense site-major streaming only).");
    return nullptr;
  }
  if (!states || !sites || !rate_cats || !rate_matrices || states > 64)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "Invalid partition geometry (states=%u sites=%u cats=%u).", states, sites,
                      rate_cats);
    return nullptr;
  }

  PartCtx * c = new (std::nothrow) PartCtx();
  if (!c)
  {
    pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for partition.");
    return nullptr;
  }
  pll_partition_t * p = &c->pub;
  memset(p, 0, sizeof(*p));
  c->magic = PLLB200_MAGIC;

  /* vector width of the device layout: 256-bit accesses need states_padded % 4 == 0 */
  p->alignment = PLL_ALIGNMENT_AVX;
  p->states_padded = (states + 3) & ~3u;
  p->attributes = attributes;
  p->tips = tips;
  p->clv_buffers = clv_buffers;
  p->nodes = tips + clv_buffers;
  p->states = states;
  p->sites = sites;
  p->pattern_weight_sum = sites;
  p->rate_matrices = rate_matrices;
  p->prob_matrices = prob_matrices;
  p->rate_cats = rate_cats;
  p->scale_buffers = scale_buffers;
  p->asc_bias_alloc = (attributes & (PLL_ATTRIB_AB_MASK | PLL_ATTRIB_AB_FLAG)) != 0;
  p->asc_additional_sites = p->asc_bias_alloc ? (int)states : 0;
  const unsigned int sites_alloc = sites + (unsigned int)p->asc_additional_sites;
  const unsigned int sp = p->states_padded;

  bool ok = true;
  p->eigen_decomp_valid = (int *)calloc(rate_matrices, sizeof(int));
  p->clv = (double **)calloc(p->nodes ? p->nodes : 1, sizeof(double *));
  p->pmatrix = (double **)calloc(prob_matrices ? prob_matrices : 1, sizeof(double *));
  p->scale_buffer = (unsigned int **)calloc(scale_buffers ? scale_buffers : 1, sizeof(unsigned int *));
  p->eigenvecs = (double **)calloc(rate_matrices, sizeof(double *));
  p->inv_eigenvecs = (double **)calloc(rate_matrices, sizeof(double *));
  p->eigenvals = (double **)calloc(rate_matrices, sizeof(double *));
  p->subst_params = (double **)calloc(rate_matrices, sizeof(double *));
  p->frequencies = (double **)calloc(rate_matrices, sizeof(double *));
  p->rates = (double *)calloc(rate_cats, sizeof(double));
  p->rate_weights = (double *)calloc(rate_cats, sizeof(double));
  p->prop_invar = (double *)calloc(rate_matrices, sizeof(double));
  p->pattern_weights = (unsigned int *)malloc(sizeof(unsigned int) * sites_alloc);
  ok = p->eigen_decomp_valid && p->clv && p->pmatrix && p->scale_buffer && p->eigenvecs &&
       p->inv_eigenvecs && p->eigenvals && p->subst_params && p->frequencies && p->rates &&
       p->rate_weights && p->prop_invar && p->pattern_weights;
  if (ok)
  {
    const unsigned int nsubst = states * (states - 1) / 2;
    for (unsigned int i = 0; i < rate_matrices && ok; ++i)
    {
      p->eigenvecs[i] = (double *)pll_aligned_alloc(sizeof(double) * states * sp, 32);
      p->inv_eigenvecs[i] = (double *)pll_aligned_alloc(sizeof(double) * states * sp, 32);
      p->eigenvals[i] = (double *)pll_aligned_alloc(sizeof(double) * sp, 32);
      p->subst_params[i] = (double *)pll_aligned_alloc(sizeof(double) * (nsubst ? nsubst : 1), 32);
      p->frequencies[i] = (double *)pll_aligned_alloc(sizeof(double) * sp, 32);
      ok = p->eigenvecs[i] && p->inv_eigenvecs[i] && p->eigenvals[i] && p->subst_params[i] &&
           p->frequencies[i];
      if (ok)
      {
        memset(p->eigenvecs[i], 0, sizeof(double) * states * sp);
        memset(p->inv_eigenvecs[i], 0, sizeof(double) * states * sp);
        memset(p->eigenvals[i], 0, sizeof(double) * sp);
        memset(p->frequencies[i], 0, sizeof(double) * sp);
        memset(p->subst_params[i], 0, sizeof(double) * (nsubst ? nsubst : 1));
      }
    }
  }
  if (!ok)
  {
    free_host_arrays(p);
    delete c;
    pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Unable to allocate enough memory.");
    return nullptr;
  }

  /* default: equal category weights (pll.c:746-752), pattern weights 1 (pll.c:806-812) */
  for (unsigned int i = 0; i < rate_cats; ++i) p->rate_weights[i] = 1.0 / rate_cats;
  for (unsigned int i = 0; i < sites_alloc; ++i) p->pattern_weights[i] = 1;
  for (unsigned int i = sites; i < sites_alloc; ++i) p->pattern_weights[i] = 0;

  c->sites_alloc = sites_alloc;
  c->eigen_version.assign(rate_matrices, 0);
  c->eigen_version_dev.assign(rate_matrices, 0xffffffffu);
  c->tip_dirty.assign(tips ? tips : 1, 0);
  c->any_tip_dirty = false;
  c->weights_dirty = true;
  c->invariant_dirty = false;
  c->tipmap_dirty = false;

  /* HBM arenas, stream, staging.  No device -> hard failure, never a CPU fallback. */
  if (!dev_ctx_init(c))
  {
    free_host_arrays(p);
    delete c;
    return nullptr;
  }
  return p;
}

extern "C" void pll_partition_destroy(pll_partition_t * partition)
{
  if (!partition) return;
  PartCtx * c = ctx_of(partition);
  if (c->magic != PLLB200_MAGIC) return;
  sumtable_forget_partition(c);
  dev_ctx_destroy(c);
  free_host_arrays(partition);
  c->magic = 0;
  delete c;
}

/* ------------------------------------------------------------------------------------------ */
/* tips                                                                                        */
/* ------------------------------------------------------------------------------------------ */

/* Build (first call) or extend (later calls with a different map) the code <-> state tables used
 * under PLL_ATTRIB_PATTERN_TIP: charmap[ascii] -> code, tipmap[code] -> state bitmask.
 * 4-state partitions store the raw mask as the code (pll.c:875-909). */
static int ensure_charmap(pll_partition_t * p, const pll_state_t * map)
{
  PartCtx * c = ctx_of(p);
  if (!p->tipchars)
  {
    p->charmap = (unsigned char *)calloc(PLL_ASCII_SIZE, 1);
    p->tipmap = (pll_state_t *)calloc(PLL_ASCII_SIZE, sizeof(pll_state_t));
    p->tipchars = (unsigned char **)calloc(p->tips, sizeof(unsigned char *));
    if (!p->charmap || !p->tipmap || !p->tipchars)
    {
      pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate tip pattern tables.");
      return PLL_FAILURE;
    }
    for (unsigned int i = 0; i < p->tips; ++i)
    {
      p->tipchars[i] = (unsigned char *)malloc(c->sites_alloc);
      if (!p->tipchars[i])
      {
        pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate space for storing tip characters.");
        return PLL_FAILURE;
      }
      /* a tip never set reads as "all states" = harmless; 4-state: mask 15, else code 0 fixed below */
      memset(p->tipchars[i], p->states == 4 ? 15 : 0, c->sites_alloc);
    }
    p->maxstates = 0;
  }

  /* merge the states of this map into tipmap, keep codes of already known states stable */
  unsigned int known = 0;
  while (known < PLL_ASCII_SIZE && p->tipmap[known]) ++known;
  unsigned int added = 0;
  memset(p->charmap, 0, PLL_ASCII_SIZE);
  for (unsigned int ch = 0; ch < PLL_ASCII_SIZE; ++ch)
  {
    const pll_state_t st = map[ch];
    if (!st) continue;
    unsigned int code = 0;
    while (code < known && p->tipmap[code] != st) ++code;
    if (code == known)
    {
      if (known >= PLL_ASCII_SIZE - 1)
      {
        pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                          "Cannot specify 256 or more states with PLL_ATTRIB_PATTERN_TIP.");
        return PLL_FAILURE;
      }
      p->tipmap[known++] = st;
      ++added;
    }
    p->charmap[ch] = (unsigned char)code;
  }
  if (added)
  {
    if (p->states == 4)
    {
      pll_state_t m = 0;
      for (unsigned int i = 0; i < known; ++i)
        if (p->tipmap[i] > m) m = p->tipmap[i];
      p->maxstates = (unsigned int)m + 1;
    }
    else
      p->maxstates = known;
    c->tipmap_dirty = true;
  }
  return PLL_SUCCESS;
}

extern "C" int pll_set_tip_states(pll_partition_t * partition, unsigned int tip_index,
                                  const pll_state_t * map, const char * sequence)
{
  PartCtx * c = ctx_of(partition);
  pll_partition_t * p = partition;
  if (tip_index >= p->tips)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid tip index %u.", tip_index);
    return PLL_FAILURE;
  }
  const unsigned int states = p->states, sites = p->sites;

  if (p->attributes & PLL_ATTRIB_PATTERN_TIP)
  {
    if (!ensure_charmap(p, map)) return PLL_FAILURE;
    unsigned char * row = p->tipchars[tip_index];
    for (unsigned int i = 0; i < sites; ++i)
    {
      const pll_state_t st = map[(unsigned char)sequence[i]];
      if (!st)
      {
        pllb200_set_error(PLL_ERROR_TIPDATA_ILLEGALSTATE, "Illegal state code in tip \"%c\"",
                          sequence[i]);
        return PLL_FAILURE;
      }
      row[i] = states == 4 ? (unsigned char)st : p->charmap[(unsigned char)sequence[i]];
    }
    if (p->asc_bias_alloc)
    {
      /* one pseudo-site per state, in state order (pll.c:896-907, 937-955) */
      if (states == 4)
        for (unsigned int i = 0; i < states; ++i) row[sites + i] = (unsigned char)(1u << i);
      else
      {
        memset(row + sites, 0, states);
        for (unsigned int code = 0; code < p->maxstates; ++code)
        {
          const pll_state_t st = p->tipmap[code];
          if (__builtin_popcountll(st) == 1)
          {
            const unsigned int pos = (unsigned int)__builtin_ctzll(st);
            if (pos < states) row[sites + pos] = (unsigned char)code;
          }
        }
      }
    }
    c->tip_dirty[tip_index] = 1;
    c->any_tip_dirty = true;
    return PLL_SUCCESS;
  }

  /* tip CLV mode: expand the bitmask to 0/1 doubles, replicated over rate categories, and push
   * it to the tip's device CLV (pll.c:958-1024) */
  const unsigned int sp = p->states_padded, K = p->rate_cats;
  std::vector<double> buf(c->clv_elems, 0.0);
  double * out = buf.data();
  for (unsigned int i = 0; i < sites; ++i)
  {
    pll_state_t st = map[(unsigned char)sequence[i]];
    if (!st)
    {
      pllb200_set_error(PLL_ERROR_TIPDATA_ILLEGALSTATE, "Illegal state code in tip \"%c\"",
                        sequence[i]);
      return PLL_FAILURE;
    }
    for (unsigned int k = 0; k < K; ++k)
    {
      pll_state_t s2 = st;
      for (unsigned int j = 0; j < states; ++j, s2 >>= 1) out[j] = (double)(s2 & 1);
      out += sp;
    }
  }
  if (p->asc_bias_alloc)
    for (unsigned int i = 0; i < states; ++i)
      for (unsigned int k = 0; k < K; ++k)
      {
        out[i] = 1.0;
        out += sp;
      }
  return pllb200_upload_clv(p, tip_index, buf.data());
}

extern "C" int pll_set_tip_clv(pll_partition_t * partition, unsigned int tip_index,
                               const double * clv, int padding)
{
  pll_partition_t * p = partition;
  PartCtx * c = ctx_of(p);
  if (p->attributes & PLL_ATTRIB_PATTERN_TIP)
  {
    pllb200_set_error(PLL_ERROR_TIPDATA_ILLEGALFUNCTION,
                      "Cannot use pll_set_tip_clv with PLL_ATTRIB_PATTERN_TIP.");
    return PLL_FAILURE;
  }
  const unsigned int sp = p->states_padded, K = p->rate_cats, states = p->states;
  const unsigned int in_stride = padding ? sp : states;
  std::vector<double> buf(c->clv_elems, 0.0);
  double * out = buf.data();
  for (unsigned int i = 0; i < p->sites; ++i, clv += in_stride)
    for (unsigned int k = 0; k < K; ++k, out += sp) memcpy(out, clv, sizeof(double) * states);
  if (p->asc_bias_alloc)
    for (unsigned int i = 0; i < states; ++i)
      for (unsigned int k = 0; k < K; ++k, out += sp) out[i] = 1.0;
  return pllb200_upload_clv(p, tip_index, buf.data());
}

extern "C" void pll_set_pattern_weights(pll_partition_t * partition,
                                        const unsigned int * pattern_weights)
{
  pll_partition_t * p = partition;
  memcpy(p->pattern_weights, pattern_weights, sizeof(unsigned int) * p->sites);
  unsigned int sum = 0;
  for (unsigned int i = 0; i < p->sites; ++i) sum += pattern_weights[i];
  p->pattern_weight_sum = sum;
  ctx_of(p)->weights_dirty = true;
}

extern "C" int pll_set_asc_bias_type(pll_partition_t * p, int asc_bias_type)
{
  /* pll.c:1145-1191 */
  const int attr = asc_bias_type & PLL_ATTRIB_AB_MASK;
  if (!p->asc_bias_alloc)
  {
    pllb200_set_error(PLL_ERROR_AB_NOSUPPORT,
                      "Partition was not created with ascertainment bias support");
    return PLL_FAILURE;
  }
  double pinv = 0.0;
  for (unsigned int i = 0; i < p->rate_matrices; ++i) pinv += p->prop_invar[i];
  if (asc_bias_type != 0 && pinv != 0.0)
  {
    pllb200_set_error(PLL_ERROR_INVAR_INCOMPAT,
                      "Invariant sites are not compatible with asc bias correction");
    return PLL_FAILURE;
  }
  if (attr != asc_bias_type || attr > PLL_ATTRIB_AB_STAMATAKIS)
  {
    pllb200_set_error(PLL_ERROR_AB_INVALIDMETHOD, "Illegal ascertainment bias algorithm \"%d\"",
                      asc_bias_type);
    return PLL_FAILURE;
  }
  p->attributes &= ~(unsigned int)PLL_ATTRIB_AB_MASK;
  p->attributes |= (unsigned int)attr;
  return PLL_SUCCESS;
}

extern "C" void pll_set_asc_state_weights(pll_partition_t * p, const unsigned int * state_weights)
{
  /* pll.c:1193-1200: weights of the `states` dummy sites that follow the real ones */
  memcpy(p->pattern_weights + p->sites, state_weights, sizeof(unsigned int) * p->states);
  ctx_of(p)->weights_dirty = true;
}

extern "C" void pll_fill_parent_scaler(unsigned int scaler_size, unsigned int * parent_scaler,
                                       const unsigned int * left_scaler,
                                       const unsigned int * right_scaler)
{
  /* host utility twin of the device rule in k_partials.cu (pll.c:1202-1224) */
  for (unsigned int i = 0; i < scaler_size; ++i)
    parent_scaler[i] = (left_scaler ? left_scaler[i] : 0) + (right_scaler ? right_scaler[i] : 0);
}

/* ------------------------------------------------------------------------------------------ */
/* model setters                                                                               */
/* ------------------------------------------------------------------------------------------ */

extern "C" void pll_set_frequencies(pll_partition_t * p, unsigned int idx, const double * f)
{
  double sum = 0;
  memcpy(p->frequencies[idx], f, sizeof(double) * p->states);
  for (unsigned int i = 0; i < p->states; ++i) sum += p->frequencies[idx][i];
  if (fabs(sum - 1.0) > PLL_MISC_EPSILON)
    for (unsigned int i = 0; i < p->states; ++i) p->frequencies[idx][i] /= sum;
  p->eigen_decomp_valid[idx] = 0;
}

extern "C" void pll_set_category_rates(pll_partition_t * p, const double * rates)
{
  memcpy(p->rates, rates, sizeof(double) * p->rate_cats);
}

extern "C" void pll_set_category_weights(pll_partition_t * p, const double * w)
{
  memcpy(p->rate_weights, w, sizeof(double) * p->rate_cats);
}

extern "C" void pll_set_subst_params(pll_partition_t * p, unsigned int idx, const double * params)
{
  memcpy(p->subst_params[idx], params, sizeof(double) * (p->states * (p->states - 1) / 2));
  p->eigen_decomp_valid[idx] = 0;
}

/* ------------------------------------------------------------------------------------------ */
/* eigen-decomposition (host; O(states^3), microseconds)                                       */
/* ------------------------------------------------------------------------------------------ */

/* Cyclic Jacobi for a dense symmetric n x n matrix.  On return d[m] are the eigenvalues and
 * row m of U is the matching unit eigenvector.  Chosen over Householder+QL for its uniformly
 * small backward error (every rotation is orthogonal to machine precision), which matters
 * because P-matrices are rebuilt from V, V^-1 thousands of times. */
static void jacobi_sym(std::vector<double> & A, int n, std::vector<double> & d,
                       std::vector<double> & U)
{
  U.assign((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i) U[(size_t)i * n + i] = 1.0;
  for (int sweep = 0; sweep < 100; ++sweep)
  {
    double off = 0, diag = 0;
    for (int i = 0; i < n; ++i)
    {
      diag += A[(size_t)i * n + i] * A[(size_t)i * n + i];
      for (int j = i + 1; j < n; ++j) off += A[(size_t)i * n + j] * A[(size_t)i * n + j];
    }
    if (off <= 1e-40 * (diag > 0 ? diag : 1.0) || off == 0.0) break;
    for (int p = 0; p < n - 1; ++p)
      for (int q = p + 1; q < n; ++q)
      {
        const double apq = A[(size_t)p * n + q];
        if (apq == 0.0) continue;
        const double app = A[(size_t)p * n + p], aqq = A[(size_t)q * n + q];
        const double theta = (aqq - app) / (2.0 * apq);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int k = 0; k < n; ++k)
        {
          const double akp = A[(size_t)k * n + p], akq = A[(size_t)k * n + q];
          A[(size_t)k * n + p] = cs * akp - sn * akq;
          A[(size_t)k * n + q] = sn * akp + cs * akq;
        }
        for (int k = 0; k < n; ++k)
        {
          const double apk = A[(size_t)p * n + k], aqk = A[(size_t)q * n + k];
          A[(size_t)p * n + k] = cs * apk - sn * aqk;
          A[(size_t)q * n + k] = sn * apk + cs * aqk;
        }
        for (int k = 0; k < n; ++k)
        {
          const double upk = U[(size_t)p * n + k], uqk = U[(size_t)q * n + k];
          U[(size_t)p * n + k] = cs * upk - sn * uqk;
          U[(size_t)q * n + k] = sn * upk + cs * uqk;
        }
      }
  }
  d.resize(n);
  for (int i = 0; i < n; ++i) d[i] = A[(size_t)i * n + i];
}

extern "C" int pll_update_eigen(pll_partition_t * p, unsigned int idx)
{
  PartCtx * c = ctx_of(p);
  const unsigned int s = p->states, sp = p->states_padded;
  const double * freqs = p->frequencies[idx];
  const double * rates = p->subst_params[idx];
  const unsigned int nsubst = s * (s - 1) / 2;

  /* exchangeabilities relative to the last one (models.c:195-203) */
  std::vector<double> r(rates, rates + nsubst);
  if (nsubst && r[nsubst - 1] > 0.0)
    for (unsigned int i = 0; i < nsubst; ++i) r[i] /= rates[nsubst - 1];

  /* B = D^1/2 Q D^-1/2 restricted to states with non-negligible frequency (models.c:227-291) */
  std::vector<int> keep;
  for (unsigned int i = 0; i < s; ++i)
    if (freqs[i] > PLL_EIGEN_MINFREQ) keep.push_back((int)i);
  std::vector<double> Q((size_t)s * s, 0.0);
  unsigned int k = 0;
  for (unsigned int i = 0; i < s; ++i)
    for (unsigned int j = i + 1; j < s; ++j, ++k)
    {
      const double f =
          (freqs[i] <= PLL_EIGEN_MINFREQ || freqs[j] <= PLL_EIGEN_MINFREQ) ? 0.0 : r[k];
      Q[(size_t)i * s + j] = Q[(size_t)j * s + i] = f * sqrt(freqs[i] * freqs[j]);
      Q[(size_t)i * s + i] -= f * freqs[j];
      Q[(size_t)j * s + j] -= f * freqs[i];
    }
  double mean = 0;
  for (unsigned int i = 0; i < s; ++i) mean += freqs[i] * (-Q[(size_t)i * s + i]);
  if (!(mean > 0))
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Degenerate rate matrix (mean rate %g).", mean);
    return PLL_FAILURE;
  }
  for (auto & v : Q) v /= mean;

  const int n = (int)keep.size();
  std::vector<double> B((size_t)n * n);
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b) B[(size_t)a * n + b] = Q[(size_t)keep[a] * s + keep[b]];
  std::vector<double> d, U;
  jacobi_sym(B, n, d, U);

  double * ev = p->eigenvecs[idx];
  double * iev = p->inv_eigenvecs[idx];
  double * val = p->eigenvals[idx];
  memset(ev, 0, sizeof(double) * s * sp);
  memset(iev, 0, sizeof(double) * s * sp);
  memset(val, 0, sizeof(double) * sp);
  /* eliminated states: eigenvalue 0 and identity rows (models.c:346-355) */
  for (unsigned int i = 0; i < s; ++i) ev[(size_t)i * sp + i] = iev[(size_t)i * sp + i] = 1.0;
  for (int a = 0; a < n; ++a)
  {
    val[keep[a]] = d[a];
    ev[(size_t)keep[a] * sp + keep[a]] = iev[(size_t)keep[a] * sp + keep[a]] = 0.0;
  }
  /* V[m][j] = U[m][j]*sqrt(pi_j),  V^-1[i][m] = U[m][i]/sqrt(pi_i)   (models.c:388-397) */
  for (int a = 0; a < n; ++a)
    for (int b = 0; b < n; ++b)
    {
      ev[(size_t)keep[a] * sp + keep[b]] = U[(size_t)a * n + b] * sqrt(freqs[keep[b]]);
      iev[(size_t)keep[a] * sp + keep[b]] = U[(size_t)b * n + a] / sqrt(freqs[keep[a]]);
    }
  p->eigen_decomp_valid[idx] = 1;
  c->eigen_version[idx]++;
  return PLL_SUCCESS;
}

/* ------------------------------------------------------------------------------------------ */
/* invariant sites                                                                             */
/* ------------------------------------------------------------------------------------------ */

static pll_state_t gap_state_of(unsigned int states)
{
  return states >= 64 ? ~0ull : ((1ull << states) - 1ull);
}

/* AND of all tip state masks per site; needs host-visible tip data, i.e. PATTERN_TIP tipchars or
 * a download of tip CLVs */
static bool site_state_and(pll_partition_t * p, std::vector<pll_state_t> & acc)
{
  PartCtx * c = ctx_of(p);
  acc.assign(p->sites, gap_state_of(p->states));
  if (p->attributes & PLL_ATTRIB_PATTERN_TIP)
  {
    if (!p->tipchars) return true;
    for (unsigned int t = 0; t < p->tips; ++t)
      for (unsigned int j = 0; j < p->sites; ++j)
        acc[j] &= p->states == 4 ? (pll_state_t)p->tipchars[t][j] : p->tipmap[p->tipchars[t][j]];
    return true;
  }
  std::vector<double> buf(c->clv_elems);
  const size_t span = (size_t)p->rate_cats * p->states_padded;
  for (unsigned int t = 0; t < p->tips; ++t)
  {
    if (!pllb200_download_clv(p, t, buf.data())) return false;
    for (unsigned int j = 0; j < p->sites; ++j)
    {
      pll_state_t st = 0;
      for (unsigned int k = 0; k < p->states; ++k)
        st |= ((pll_state_t)(buf[j * span + k] != 0.0)) << k;
      acc[j] &= st;
    }
  }
  return true;
}

extern "C" int pll_update_invariant_sites(pll_partition_t * p)
{
  if (!p->invariant) p->invariant = (int *)malloc(sizeof(int) * p->sites);
  if (!p->invariant)
  {
    pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate invariant sites array.");
    return PLL_FAILURE;
  }
  std::vector<pll_state_t> acc;
  if (!site_state_and(p, acc)) return PLL_FAILURE;
  for (unsigned int i = 0; i < p->sites; ++i)
    p->invariant[i] = (__builtin_popcountll(acc[i]) == 1) ? __builtin_ctzll(acc[i]) : -1;
  ctx_of(p)->invariant_dirty = true;
  return PLL_SUCCESS;
}

extern "C" unsigned int pll_count_invariant_sites(pll_partition_t * p,
                                                  unsigned int * state_inv_count)
{
  unsigned int count = 0;
  if (state_inv_count) memset(state_inv_count, 0, sizeof(unsigned int) * p->states);
  if (p->invariant)
  {
    for (unsigned int i = 0; i < p->sites; ++i)
      if (p->invariant[i] > -1)
      {
        count += p->pattern_weights[i];
        if (state_inv_count) state_inv_count[p->invariant[i]]++;
      }
    return count;
  }
  std::vector<pll_state_t> acc;
  if (!site_state_and(p, acc)) return 0;
  for (unsigned int i = 0; i < p->sites; ++i)
    if (__builtin_popcountll(acc[i]) == 1)
    {
      count += p->pattern_weights[i];
      if (state_inv_count) state_inv_count[__builtin_ctzll(acc[i])]++;
    }
  return count;
}

extern "C" int pll_update_invariant_sites_proportion(pll_partition_t * p, unsigned int idx,
                                                     double prop_invar)
{
  if (prop_invar != 0.0 && (p->attributes & PLL_ATTRIB_AB_MASK))
  {
    pllb200_set_error(PLL_ERROR_INVAR_INCOMPAT,
                      "Invariant sites are not compatible with asc bias correction");
    return PLL_FAILURE;
  }
  if (prop_invar < 0 || prop_invar >= 1)
  {
    pllb200_set_error(PLL_ERROR_INVAR_PROPORTION, "Invalid proportion of invariant sites (%f)",
                      prop_invar);
    return PLL_FAILURE;
  }
  if (idx > p->rate_matrices)
  {
    pllb200_set_error(PLL_ERROR_INVAR_PARAMINDEX, "Invalid params index (%u)", idx);
    return PLL_FAILURE;
  }
  if (prop_invar > 0.0 && !p->invariant)
    if (!pll_update_invariant_sites(p))
    {
      pllb200_set_error(PLL_ERROR_INVAR_NONEFOUND, "No invariant sites found");
      return PLL_FAILURE;
    }
  p->prop_invar[idx] = prop_invar;
  return PLL_SUCCESS;
}

/* ------------------------------------------------------------------------------------------ */
/* discrete Gamma rates                                                                        */
/* ------------------------------------------------------------------------------------------ */

namespace
{
/* ln Gamma(x), Pike & Hill (1966) ACM alg. 291: shift to x >= 7 then Stirling (gamma.c:101-136) */
double ln_gamma(double alpha)
{
  double x = alpha, shift = 0.0;
  if (x < 7.0)
  {
    double f = 1.0, z = alpha;
    while (z < 7.0)
    {
      f *= z;
      z += 1.0;
    }
    x = z;
    shift = -log(f);
  }
  const double z = 1.0 / (x * x);
  return shift + (x - 0.5) * log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z +
          .083333333333333) / x;
}

/* regularised lower incomplete gamma, Bhattacharjee (1970) AS32 (gamma.c:27-99): power series for
 * x <= 1 or x < p, else a continued fraction evaluated by forward recurrence with rescaling */
double inc_gamma(double x, double p, double ln_gamma_p)
{
  const double accurate = 1e-8, overflow = 1e30;
  if (x == 0) return 0;
  if (x < 0 || p <= 0) return -1;
  const double factor = exp(p * log(x) - x - ln_gamma_p);
  if (!(x > 1 && x >= p))
  {
    double gin = 1, term = 1, rn = p;
    do
    {
      rn += 1;
      term *= x / rn;
      gin += term;
    } while (term > accurate);
    return gin * factor / p;
  }
  double a = 1 - p, b = a + x + 1, term = 0;
  double pn[6] = {1, x, x + 1, x * b, 0, 0};
  double gin = pn[2] / pn[3];
  for (;;)
  {
    a += 1;
    b += 2;
    term += 1;
    const double an = a * term;
    pn[4] = b * pn[2] - an * pn[0];
    pn[5] = b * pn[3] - an * pn[1];
    if (pn[5] != 0)
    {
      const double rn = pn[4] / pn[5];
      const double dif = fabs(gin - rn);
      if (dif <= accurate && dif <= accurate * rn) return 1 - factor * gin;
      gin = rn;
    }
    for (int i = 0; i < 4; ++i) pn[i] = pn[i + 2];
    if (fabs(pn[4]) >= overflow)
      for (int i = 0; i < 4; ++i) pn[i] /= overflow;
  }
}

/* standard normal quantile, Odeh & Evans (1974) AS70 (gamma.c:138-164) */
double point_normal(double prob)
{
  const double a0 = -.322232431088, a1 = -1, a2 = -.342242088547, a3 = -.0204231210245,
               a4 = -.453642210148e-4, b0 = .0993484626060, b1 = .588581570495,
               b2 = .531103462366, b3 = .103537752850, b4 = .0038560700634;
  const double p1 = prob < 0.5 ? prob : 1 - prob;
  if (p1 < 1e-20) return -9999;
  const double y = sqrt(log(1 / (p1 * p1)));
  const double z = y + ((((y * a4 + a3) * y + a2) * y + a1) * y + a0) /
                           ((((y * b4 + b3) * y + b2) * y + b1) * y + b0);
  return prob < 0.5 ? -z : z;
}

/* chi-square quantile, Best & Roberts (1975) AS91 (gamma.c:166-219): starting value by regime,
 * then the seven-term Taylor refinement until relative change < 5e-7 */
double point_chi2(double prob, double v)
{
  const double e = .5e-6, aa = .6931471805;
  if (prob < .000002 || prob > .999998 || v <= 0) return -1;
  const double g = ln_gamma(v / 2), xx = v / 2, c = xx - 1;
  double ch;
  if (v < -1.24 * log(prob))
  {
    ch = pow(prob * xx * exp(g + xx * aa), 1 / xx);
    if (ch - e < 0) return ch;
  }
  else if (v <= .32)
  {
    ch = 0.4;
    const double a = log(1 - prob);
    double q;
    do
    {
      q = ch;
      const double p1 = 1 + ch * (4.67 + ch), p2 = ch * (6.73 + ch * (6.66 + ch));
      const double t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - exp(a + g + .5 * ch + c * aa) * p2 / p1) / t;
    } while (fabs(q / ch - 1) - .01 > 0);
  }
  else
  {
    const double x = point_normal(prob), p1 = 0.222222 / v;
    ch = v * pow(x * sqrt(p1) + 1 - p1, 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (log(1 - prob) - c * log(.5 * ch) + g);
  }
  double q;
  do
  {
    q = ch;
    const double p1 = .5 * ch;
    const double ig = inc_gamma(p1, xx, g);
    if (ig < 0) return -1;
    const double p2 = prob - ig;
    const double t = p2 * exp(xx * aa + g + p1 - c * log(ch));
    const double b = t / ch, a = 0.5 * t - b * c;
    const double s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    const double s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    const double s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    const double s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    const double s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    const double s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 - b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (fabs(q / ch - 1) > e);
  return ch;
}

double point_gamma(double prob, double alpha, double beta)
{
  return point_chi2(prob, 2.0 * alpha) / (2.0 * beta);
}
}  // namespace

extern "C" int pll_compute_gamma_cats(double alpha, unsigned int categories,
                                      double * output_rates, int rates_mode)
{
  if (alpha < 0.02 || categories < 1)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid alpha value (%f)", alpha);
    return PLL_FAILURE;
  }
  const double factor = (double)categories;
  if (categories == 1)
  {
    output_rates[0] = 1.0;
    return PLL_SUCCESS;
  }
  if (rates_mode == PLL_GAMMA_RATES_MEDIAN)
  {
    const double middle = 1.0 / (2.0 * categories);
    double t = 0;
    for (unsigned int i = 0; i < categories; ++i)
    {
      output_rates[i] = point_gamma((double)(i * 2 + 1) * middle, alpha, alpha);
      t += output_rates[i];
    }
    for (unsigned int i = 0; i < categories; ++i) output_rates[i] *= factor / t;
    return PLL_SUCCESS;
  }
  if (rates_mode == PLL_GAMMA_RATES_MEAN)
  {
    std::vector<double> cut(categories);
    const double lnga1 = ln_gamma(alpha + 1);
    for (unsigned int i = 0; i + 1 < categories; ++i)
      cut[i] = inc_gamma(point_gamma((i + 1.0) / categories, alpha, alpha) * alpha, alpha + 1, lnga1);
    output_rates[0] = cut[0] * factor;
    output_rates[categories - 1] = (1 - cut[categories - 2]) * factor;
    for (unsigned int i = 1; i + 1 < categories; ++i)
      output_rates[i] = (cut[i] - cut[i - 1]) * factor;
    return PLL_SUCCESS;
  }
  pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid GAMMA discretization mode (%d)", rates_mode);
  return PLL_FAILURE;
}

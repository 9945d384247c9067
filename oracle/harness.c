/*
 * harness.c — TEST INFRASTRUCTURE.  One source, two builds:
 *   oracle/_ref/libharness_ref.so            gcc -DHARNESS_REFERENCE against the reference headers
 *                                            + oracle/_ref/libpll_ref.so (the unmodified reference)
 *   pll-modules_b200/lib/libharness_b200.so  against include/ + libpll_b200.so (the product)
 * It only uses the public C API both libraries share (pll_partition_*, pll_update_*,
 * pll_compute_*, pllmod_treeinfo_*, pllmod_opt_*), so every parity test drives the reference and
 * the product through literally the same caller code — the way the reference's own tests
 * (test/src/optimize/blopt-minimal.c, libs/libpll/test/src/derivatives.c) drive libpll.
 * The only #ifdef is how a CLV / scaler / P-matrix is read back (host memory vs device download).
 *
 * Synthetic inputs follow SURVEY.md §8d: random-insertion topology, Exp(mean) branch lengths
 * clipped to [1e-4, 1] * scale, iid alignment with 5 % gaps from a counter-based hash (so any site
 * slice of the same alignment can be generated independently: multi-GPU site sharding).
 */
#include "pll.h"
#include "pll_tree.h"
#include "pll_optimize.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef struct
{
  unsigned int tips;
  unsigned int nparts;
  pll_utree_t * tree;
  pllmod_treeinfo_t * ti;
  pll_partition_t ** parts;
  unsigned int ** param_idx;
  unsigned int * site_begin; /* slice of the global alignment this handle owns */
  unsigned int * sites;      /* local sites per partition */
  double ** sumtables;
  int brlen_linkage;
} harness_t;

/* ---- deterministic generators -------------------------------------------------------------- */

static unsigned long long xs64(unsigned long long * s)
{
  unsigned long long x = *s;
  x ^= x << 13;
  x ^= x >> 7;
  x ^= x << 17;
  return *s = x;
}
static double u01(unsigned long long * s) { return (double)(xs64(s) >> 11) / 9007199254740992.0; }

static unsigned long long mix64(unsigned long long z)
{
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static const char * ALPHABET_NT = "ACGT";
static const char * ALPHABET_AA = "ARNDCQEGHILKMFPSTWYV";

/* character of (partition, tip, global site): 5 % gap, else uniform over the states */
static char cell_char(unsigned long long seed, unsigned int part, unsigned int tip,
                      unsigned long long site, unsigned int states)
{
  const unsigned long long h =
      mix64(seed ^ mix64(((unsigned long long)part << 48) ^ ((unsigned long long)tip << 32) ^ site));
  const unsigned int r = (unsigned int)(h % 10000u);
  if (r < 500u) return '-';
  /* seeds with bit 40 set: 3 % IUPAC ambiguity codes in DNA (multi-bit tip masks other than the gap) */
  if (states == 4 && ((seed >> 40) & 1ull) && r < 800u) return "RYMKSWBDHV"[(h >> 24) % 10u];
  const unsigned int st = (unsigned int)((h >> 20) % states);
  if (states == 4) return ALPHABET_NT[st];
  if (states == 20) return ALPHABET_AA[st];
  return (char)('0' + st); /* generic alphabets use a custom map built in make_partition */
}

static pll_unode_t * new_tip(unsigned int idx)
{
  pll_unode_t * t = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
  t->clv_index = t->node_index = t->pmatrix_index = idx;
  t->scaler_index = PLL_SCALE_BUFFER_NONE;
  return t;
}

static pll_unode_t * new_inner(unsigned int clv, int scaler, unsigned int * node_id)
{
  pll_unode_t * n[3];
  for (int i = 0; i < 3; ++i)
  {
    n[i] = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
    n[i]->clv_index = clv;
    n[i]->scaler_index = scaler;
    n[i]->node_index = (*node_id)++;
  }
  n[0]->next = n[1];
  n[1]->next = n[2];
  n[2]->next = n[0];
  return n[0];
}

static void join(pll_unode_t * a, pll_unode_t * b, double len, unsigned int pm)
{
  a->back = b;
  b->back = a;
  a->length = b->length = len;
  a->pmatrix_index = b->pmatrix_index = pm;
}

/* random-insertion topology; branch e of the growing tree is remembered by one of its ends */
static pll_unode_t * build_tree(unsigned int tips, unsigned long long seed, double brlen_mean,
                                double brlen_scale)
{
  unsigned long long s = seed ? seed : 88172645463325252ull;
  unsigned int node_id = tips;
  pll_unode_t ** ends = (pll_unode_t **)malloc(sizeof(pll_unode_t *) * (2 * tips - 3));
  unsigned int nb = 0;
  pll_unode_t * c0 = new_inner(tips, 0, &node_id);
  pll_unode_t * t0 = new_tip(0), *t1 = new_tip(1), *t2 = new_tip(2);
  join(t0, c0, 0.1, 0);
  join(t1, c0->next, 0.1, 1);
  join(t2, c0->next->next, 0.1, 2);
  ends[nb++] = t0;
  ends[nb++] = t1;
  ends[nb++] = t2;
  for (unsigned int t = 3; t < tips; ++t)
  {
    pll_unode_t * in = new_inner(tips + t - 2, (int)(t - 2), &node_id);
    pll_unode_t * tip = new_tip(t);
    pll_unode_t * e = ends[xs64(&s) % nb];
    pll_unode_t * eb = e->back;
    const unsigned int pm = e->pmatrix_index;
    join(e, in, 0.1, pm);
    join(eb, in->next, 0.1, nb);
    join(tip, in->next->next, 0.1, nb + 1);
    ends[nb++] = in->next;
    ends[nb++] = tip;
  }
  for (unsigned int i = 0; i < nb; ++i)
  {
    double l = -brlen_mean * log(1.0 - u01(&s));
    if (l < 1e-4) l = 1e-4;
    if (l > 1.0) l = 1.0;
    l *= brlen_scale;
    ends[i]->length = ends[i]->back->length = l;
  }
  free(ends);
  return c0;
}

/* ---- construction -------------------------------------------------------------------------- */

static pll_partition_t * make_partition(unsigned int part, unsigned int tips, unsigned int states,
                                        unsigned int site_begin, unsigned int sites,
                                        unsigned int rate_cats, unsigned int attributes,
                                        unsigned long long seed, double alpha, double pinv,
                                        const double * subst, const double * freqs)
{
  pll_partition_t * p = pll_partition_create(tips, tips - 2, states, sites, 1, 2 * tips - 3,
                                             rate_cats, tips - 2, attributes);
  if (!p) return NULL;
  double rates[64];
  if (!pll_compute_gamma_cats(alpha, rate_cats, rates, PLL_GAMMA_RATES_MEAN)) return NULL;
  pll_set_category_rates(p, rates);
  pll_set_frequencies(p, 0, freqs);
  pll_set_subst_params(p, 0, subst);

  pll_state_t gmap[256];
  const pll_state_t * map = states == 4 ? pll_map_nt : (states == 20 ? pll_map_aa : gmap);
  if (map == gmap)
  {
    memset(gmap, 0, sizeof(gmap));
    for (unsigned int i = 0; i < states; ++i) gmap['0' + i] = 1ull << i;
    gmap['-'] = (states >= 64) ? ~0ull : ((1ull << states) - 1);
  }
  char * seq = (char *)malloc((size_t)sites + 1);
  for (unsigned int t = 0; t < tips; ++t)
  {
    for (unsigned int i = 0; i < sites; ++i)
      seq[i] = cell_char(seed, part, t, (unsigned long long)site_begin + i, states);
    seq[sites] = 0;
    if (!pll_set_tip_states(p, t, map, seq))
    {
      free(seq);
      return NULL;
    }
  }
  free(seq);
  if (attributes & PLL_ATTRIB_AB_MASK)
  {
    /* ascertainment bias: weights of the per-state dummy sites (used by Felsenstein / Stamatakis) */
    unsigned int aw[64];
    for (unsigned int i = 0; i < states; ++i) aw[i] = 1 + (i % 3);
    pll_set_asc_state_weights(p, aw);
  }
  if (pinv > 0.0)
    if (!pll_update_invariant_sites_proportion(p, 0, pinv)) return NULL;
  return p;
}

/* Create tree + nparts partitions + treeinfo.
 * states[p], sites_total[p] describe the GLOBAL alignment; this handle owns the site slice
 * [floor(rank*S/nranks), floor((rank+1)*S/nranks)) of every partition.
 * subst/freqs: concatenated per partition (states*(states-1)/2 and states doubles). */
void * h_create(unsigned int tips, unsigned int nparts, const unsigned int * states,
                const unsigned int * sites_total, unsigned int rate_cats, unsigned int attributes,
                unsigned long long seed, double brlen_mean, double brlen_scale,
                const double * alphas, const double * pinvs, const double * subst,
                const double * freqs, int brlen_linkage, unsigned int rank, unsigned int nranks)
{
  harness_t * h = (harness_t *)calloc(1, sizeof(harness_t));
  h->tips = tips;
  h->nparts = nparts;
  h->brlen_linkage = brlen_linkage;
  h->parts = (pll_partition_t **)calloc(nparts, sizeof(pll_partition_t *));
  h->param_idx = (unsigned int **)calloc(nparts, sizeof(unsigned int *));
  h->site_begin = (unsigned int *)calloc(nparts, sizeof(unsigned int));
  h->sites = (unsigned int *)calloc(nparts, sizeof(unsigned int));
  h->sumtables = (double **)calloc(nparts, sizeof(double *));
  pll_unode_t * root = build_tree(tips, seed, brlen_mean, brlen_scale);
  h->ti = pllmod_treeinfo_create(root, tips, nparts, brlen_linkage);
  if (!h->ti)
  {
    fprintf(stderr, "harness: treeinfo_create failed: %s\n", pll_errmsg);
    return NULL;
  }
  h->tree = h->ti->tree;
  for (unsigned int p = 0; p < nparts; ++p)
  {
    const unsigned long long S = sites_total[p];
    const unsigned int b = (unsigned int)(S * rank / nranks), e = (unsigned int)(S * (rank + 1) / nranks);
    h->site_begin[p] = b;
    h->sites[p] = e - b;
    h->parts[p] = make_partition(p, tips, states[p], b, e - b, rate_cats, attributes, seed,
                                 alphas[p], pinvs ? pinvs[p] : 0.0, subst, freqs);
    if (!h->parts[p])
    {
      fprintf(stderr, "harness: partition %u failed: %s\n", p, pll_errmsg);
      return NULL;
    }
    subst += states[p] * (states[p] - 1) / 2;
    freqs += states[p];
    h->param_idx[p] = (unsigned int *)calloc(rate_cats, sizeof(unsigned int));
    if (!pllmod_treeinfo_init_partition(h->ti, p, h->parts[p], 0, PLL_GAMMA_RATES_MEAN, alphas[p],
                                        h->param_idx[p], NULL))
    {
      fprintf(stderr, "harness: init_partition failed: %s\n", pll_errmsg);
      return NULL;
    }
    h->sumtables[p] = h->ti->deriv_precomp[p];
  }
  return h;
}

void h_destroy(void * hv)
{
  harness_t * h = (harness_t *)hv;
  if (!h) return;
  pll_utree_t * tree = h->ti->tree;
  /* treeinfo frees its wrapper but not the nodes: collect them first */
  const unsigned int n = tree->tip_count + tree->inner_count;
  pll_unode_t ** nodes = (pll_unode_t **)malloc(sizeof(pll_unode_t *) * n);
  memcpy(nodes, tree->nodes, sizeof(pll_unode_t *) * n);
  const unsigned int tipc = tree->tip_count;
  for (unsigned int p = 0; p < h->nparts; ++p) pll_partition_destroy(h->parts[p]);
  pllmod_treeinfo_destroy(h->ti);
  for (unsigned int i = 0; i < n; ++i)
  {
    if (i < tipc)
      free(nodes[i]);
    else
    {
      pll_unode_t * a = nodes[i], *b = a->next, *c = b->next;
      free(a);
      free(b);
      free(c);
    }
  }
  free(nodes);
  for (unsigned int p = 0; p < h->nparts; ++p) free(h->param_idx[p]);
  free(h->parts);
  free(h->param_idx);
  free(h->site_begin);
  free(h->sites);
  free(h->sumtables);
  free(h);
}

/* ---- accessors ----------------------------------------------------------------------------- */

unsigned int h_local_sites(void * hv, unsigned int p) { return ((harness_t *)hv)->sites[p]; }
unsigned int h_states_padded(void * hv, unsigned int p) { return ((harness_t *)hv)->parts[p]->states_padded; }
void * h_partition(void * hv, unsigned int p) { return ((harness_t *)hv)->parts[p]; }
void * h_treeinfo(void * hv) { return ((harness_t *)hv)->ti; }
int h_errno(void) { return pll_errno; }
const char * h_errmsg(void) { return pll_errmsg; }

void h_set_reduce_cb(void * hv, void * ctx, void (*cb)(void *, double *, size_t, int))
{
  pllmod_treeinfo_set_parallel_context(((harness_t *)hv)->ti, ctx, cb);
}

/* out: sites * rate_cats * states (UNPADDED, so the two libraries compare independent of padding) */
int h_get_clv(void * hv, unsigned int p, unsigned int clv_index, double * out)
{
  harness_t * h = (harness_t *)hv;
  pll_partition_t * part = h->parts[p];
  const unsigned int sp = part->states_padded, s = part->states, K = part->rate_cats;
  const size_t n = (size_t)part->sites * K;
  double * buf;
#ifdef HARNESS_REFERENCE
  buf = part->clv[clv_index];
  if (!buf) return 0;
#else
  buf = (double *)malloc(sizeof(double) * ((size_t)part->sites + part->asc_additional_sites) * K * sp);
  if (!pllb200_download_clv(part, clv_index, buf))
  {
    free(buf);
    return 0;
  }
#endif
  for (size_t e = 0; e < n; ++e) memcpy(out + e * s, buf + e * sp, sizeof(double) * s);
#ifndef HARNESS_REFERENCE
  free(buf);
#endif
  return 1;
}

int h_get_scaler(void * hv, unsigned int p, unsigned int scaler_index, unsigned int * out)
{
  harness_t * h = (harness_t *)hv;
  pll_partition_t * part = h->parts[p];
  const size_t n = (size_t)part->sites * ((part->attributes & PLL_ATTRIB_RATE_SCALERS) ? part->rate_cats : 1);
#ifdef HARNESS_REFERENCE
  memcpy(out, part->scale_buffer[scaler_index], sizeof(unsigned int) * n);
  return 1;
#else
  const size_t na = ((size_t)part->sites + part->asc_additional_sites) *
                    ((part->attributes & PLL_ATTRIB_RATE_SCALERS) ? part->rate_cats : 1);
  unsigned int * buf = (unsigned int *)malloc(sizeof(unsigned int) * na);
  const int rc = pllb200_download_scaler(part, scaler_index, buf);
  if (rc) memcpy(out, buf, sizeof(unsigned int) * n);
  free(buf);
  return rc;
#endif
}

/* out: rate_cats * states * states (unpadded) */
int h_get_pmatrix(void * hv, unsigned int p, unsigned int matrix_index, double * out)
{
  harness_t * h = (harness_t *)hv;
  pll_partition_t * part = h->parts[p];
  const unsigned int sp = part->states_padded, s = part->states, K = part->rate_cats;
  double * buf;
#ifdef HARNESS_REFERENCE
  buf = part->pmatrix[matrix_index];
#else
  buf = (double *)malloc(sizeof(double) * K * s * sp);
  if (!pllb200_download_pmatrix(part, matrix_index, buf))
  {
    free(buf);
    return 0;
  }
#endif
  for (unsigned int r = 0; r < K * s; ++r) memcpy(out + (size_t)r * s, buf + (size_t)r * sp, sizeof(double) * s);
#ifndef HARNESS_REFERENCE
  free(buf);
#endif
  return 1;
}

void h_get_rates(void * hv, unsigned int p, double * out)
{
  pll_partition_t * part = ((harness_t *)hv)->parts[p];
  memcpy(out, part->rates, sizeof(double) * part->rate_cats);
}

/* branch lengths by pmatrix index (2*tips-3 doubles) */
void h_get_branch_lengths(void * hv, double * out)
{
  harness_t * h = (harness_t *)hv;
  for (unsigned int i = 0; i < h->ti->subnode_count; ++i)
    out[h->ti->subnodes[i]->pmatrix_index] = h->ti->subnodes[i]->length;
}

void h_set_branch_lengths(void * hv, const double * in)
{
  harness_t * h = (harness_t *)hv;
  for (unsigned int i = 0; i < h->ti->subnode_count; ++i)
    h->ti->subnodes[i]->length = in[h->ti->subnodes[i]->pmatrix_index];
  pllmod_treeinfo_invalidate_all(h->ti);
}

unsigned int h_root_node_index(void * hv) { return ((harness_t *)hv)->ti->root->node_index; }
int h_set_root(void * hv, unsigned int node_index)
{
  harness_t * h = (harness_t *)hv;
  return pllmod_treeinfo_set_root(h->ti, h->ti->subnodes[node_index]);
}
/* describe subnode i: {clv_index, scaler_index, pmatrix_index, back node_index, is_tip} */
void h_node_info(void * hv, unsigned int node_index, int * out)
{
  pll_unode_t * n = ((harness_t *)hv)->ti->subnodes[node_index];
  out[0] = (int)n->clv_index;
  out[1] = n->scaler_index;
  out[2] = (int)n->pmatrix_index;
  out[3] = (int)n->back->node_index;
  out[4] = n->next == NULL;
}

/* ---- the hot path ---------------------------------------------------------------------------- */

double h_compute_loglh(void * hv, int incremental)
{
  return pllmod_treeinfo_compute_loglh(((harness_t *)hv)->ti, incremental);
}

void h_partition_loglh(void * hv, double * out)
{
  harness_t * h = (harness_t *)hv;
  memcpy(out, h->ti->partition_loglh, sizeof(double) * h->nparts);
}

double h_compute_loglh_persite(void * hv, double * persite_p0)
{
  harness_t * h = (harness_t *)hv;
  double ** ps = (double **)calloc(h->nparts, sizeof(double *));
  ps[0] = persite_p0;
  const double r = pllmod_treeinfo_compute_loglh_persite(h->ti, 0, ps);
  free(ps);
  return r;
}

/* only the CLV updates of a full/partial traversal (what BASELINE's "CLV site-updates/s" times):
 * the op list is built once, then pll_update_partials is called reps times. Returns seconds/rep. */
double h_time_update_partials(void * hv, int reps)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  (void)pllmod_treeinfo_compute_loglh(ti, 0); /* P-matrices + op list in ti->operations */
  const unsigned int ops = ti->tip_count - 2;
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int r = 0; r < reps; ++r)
    for (unsigned int p = 0; p < h->nparts; ++p) pll_update_partials(h->parts[p], ti->operations, ops);
#ifndef HARNESS_REFERENCE
  for (unsigned int p = 0; p < h->nparts; ++p) pllb200_partition_synchronize(h->parts[p]);
#endif
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return ((t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec)) / reps;
}

/* queue ONE full-traversal CLV update per partition without waiting (device timing is done by the
 * caller with CUDA events on the partition stream) */
void h_queue_update_partials(void * hv)
{
  harness_t * h = (harness_t *)hv;
  const unsigned int ops = h->ti->tip_count - 2;
  for (unsigned int p = 0; p < h->nparts; ++p) pll_update_partials(h->parts[p], h->ti->operations, ops);
}

int h_update_partials_only(void * hv, int incremental)
{
#ifdef HARNESS_REFERENCE
  (void)hv;
  (void)incremental;
  return 0; /* additive symbol, absent from the reference */
#else
  return pllmod_treeinfo_update_partials(((harness_t *)hv)->ti, incremental);
#endif
}

double h_edge_loglh(void * hv, unsigned int p, unsigned int node_index)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * n = h->ti->subnodes[node_index];
  return pll_compute_edge_loglikelihood(h->parts[p], n->clv_index, n->scaler_index,
                                        n->back->clv_index, n->back->scaler_index,
                                        n->pmatrix_index, h->param_idx[p], NULL);
}

double h_root_loglh(void * hv, unsigned int p, unsigned int node_index)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * n = h->ti->subnodes[node_index];
  return pll_compute_root_loglikelihood(h->parts[p], n->clv_index, n->scaler_index,
                                        h->param_idx[p], NULL);
}

/* sumtable on the edge at subnode `node_index` (CLVs must be oriented towards it, e.g. the root
 * edge after h_compute_loglh), then derivatives at each of the n branch lengths */
int h_derivatives(void * hv, unsigned int p, unsigned int node_index, const double * brlens, int n,
                  double * df, double * ddf)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * e = h->ti->subnodes[node_index];
  if (!pll_update_sumtable(h->parts[p], e->clv_index, e->back->clv_index, e->scaler_index,
                           e->back->scaler_index, h->param_idx[p], h->sumtables[p]))
    return 0;
  for (int i = 0; i < n; ++i)
    if (!pll_compute_likelihood_derivatives(h->parts[p], e->scaler_index, e->back->scaler_index,
                                            brlens[i], h->param_idx[p], h->sumtables[p], &df[i], &ddf[i]))
      return 0;
  return 1;
}

/* seconds per call of pll_compute_edge_loglikelihood / pll_compute_likelihood_derivatives in a C loop (no
 * interpreter between the calls); the derivative loop varies the branch length like a Newton iteration does */
double h_time_edge_loglh(void * hv, unsigned int p, unsigned int node_index, int reps)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * n = h->ti->subnodes[node_index];
  struct timespec t0, t1;
  volatile double sink = 0.0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int r = 0; r < reps; ++r)
    sink += pll_compute_edge_loglikelihood(h->parts[p], n->clv_index, n->scaler_index, n->back->clv_index,
                                           n->back->scaler_index, n->pmatrix_index, h->param_idx[p], NULL);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  (void)sink;
  return ((double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec)) / (reps > 0 ? reps : 1);
}

double h_time_derivatives(void * hv, unsigned int p, unsigned int node_index, int reps)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * e = h->ti->subnodes[node_index];
  if (!pll_update_sumtable(h->parts[p], e->clv_index, e->back->clv_index, e->scaler_index,
                           e->back->scaler_index, h->param_idx[p], h->sumtables[p]))
    return -1.0;
  struct timespec t0, t1;
  double df, ddf;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int r = 0; r < reps; ++r)
    if (!pll_compute_likelihood_derivatives(h->parts[p], e->scaler_index, e->back->scaler_index,
                                            0.01 + 0.49 * (double)r / (double)(reps > 1 ? reps - 1 : 1),
                                            h->param_idx[p], h->sumtables[p], &df, &ddf))
      return -1.0;
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return ((double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec)) / (reps > 0 ? reps : 1);
}

/* out: sites*rate_cats*states (unpadded) */
int h_get_sumtable(void * hv, unsigned int p, double * out)
{
  harness_t * h = (harness_t *)hv;
  pll_partition_t * part = h->parts[p];
  const unsigned int sp = part->states_padded, s = part->states, K = part->rate_cats;
  const size_t n = (size_t)part->sites * K;
#ifdef HARNESS_REFERENCE
  const double * buf = h->sumtables[p];
#else
  double * buf = (double *)malloc(sizeof(double) * ((size_t)part->sites + part->asc_additional_sites) * K * sp);
  if (!pllb200_download_sumtable(part, h->sumtables[p], buf))
  {
    free(buf);
    return 0;
  }
#endif
  for (size_t e = 0; e < n; ++e) memcpy(out + e * s, buf + e * sp, sizeof(double) * s);
#ifndef HARNESS_REFERENCE
  free((void *)buf);
#endif
  return 1;
}

/* Newton-Raphson branch-length optimisation around the current root (all branches if radius -1) */
double h_optimize_branch_lengths(void * hv, int method, int radius, int max_iters, double lh_epsilon,
                                 double brlen_min, double brlen_max)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  return pllmod_opt_optimize_branch_lengths_local_multi(
      ti->partitions, ti->partition_count, ti->root, ti->param_indices, ti->deriv_precomp,
      ti->branch_lengths, ti->brlen_scalers, brlen_min, brlen_max, lh_epsilon, max_iters, radius, 1,
      method, ti->brlen_linkage, ti->parallel_context, ti->parallel_reduce_cb);
}

/* legacy single-partition driver (pllmod_opt_optimize_branch_lengths_local) on partition 0 */
double h_optimize_branch_lengths_single(void * hv, int radius, int smoothings, double tolerance)
{
  harness_t * h = (harness_t *)hv;
  return pllmod_opt_optimize_branch_lengths_local(h->parts[0], h->ti->root, h->param_idx[0], 0, 0,
                                                  tolerance, smoothings, radius, 1);
}

/* change the Gamma shape of partition p (model-parameter optimisers do this between evaluations) */
int h_set_alpha(void * hv, unsigned int p, double alpha)
{
  harness_t * h = (harness_t *)hv;
  double rates[64];
  if (!pll_compute_gamma_cats(alpha, h->parts[p]->rate_cats, rates, PLL_GAMMA_RATES_MEAN)) return 0;
  pll_set_category_rates(h->parts[p], rates);
  const int old = h->ti->active_partition;
  pllmod_treeinfo_set_active_partition(h->ti, (int)p);
  pllmod_treeinfo_invalidate_all(h->ti);
  pllmod_treeinfo_set_active_partition(h->ti, old);
  return 1;
}

int h_gamma_cats(double alpha, unsigned int cats, int mode, double * out)
{
  return pll_compute_gamma_cats(alpha, cats, out, mode);
}

/* invalidate the CLVs/P-matrix around an edge as an SPR regraft evaluation would (algo_search.c:
 * 753-790), then change its length: exercises small partial traversals */
void h_perturb_edge(void * hv, unsigned int node_index, double new_length)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * e = h->ti->subnodes[node_index];
  pllmod_treeinfo_set_branch_length(h->ti, e, new_length);
  pllmod_treeinfo_invalidate_pmatrix(h->ti, e);
  /* every CLV that "sees" this edge from the root side is stale: walk towards the root is not
   * needed — the reference's callers invalidate the two end nodes' other orientations */
  if (e->next)
  {
    pllmod_treeinfo_invalidate_clv(h->ti, e->next);
    pllmod_treeinfo_invalidate_clv(h->ti, e->next->next);
  }
  if (e->back->next)
  {
    pllmod_treeinfo_invalidate_clv(h->ti, e->back->next);
    pllmod_treeinfo_invalidate_clv(h->ti, e->back->next->next);
  }
}

unsigned int h_ops_counter(void * hv) { return ((harness_t *)hv)->ti->counter; }

/* {tip-tip, tip-inner, inner-inner} operation counts of the last traversal of partition p and the
 * number of ops; used by bench.py to turn a launch into algorithmic bytes */
void h_op_type_counts(void * hv, unsigned int p, unsigned int nops, unsigned int * out)
{
  harness_t * h = (harness_t *)hv;
  const pll_partition_t * part = h->parts[p];
  const int pattern = (part->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  out[0] = out[1] = out[2] = 0;
  for (unsigned int i = 0; i < nops; ++i)
  {
    const pll_operation_t * o = &h->ti->operations[i];
    const int t1 = pattern && o->child1_clv_index < part->tips;
    const int t2 = pattern && o->child2_clv_index < part->tips;
    out[t1 && t2 ? 0 : ((t1 || t2) ? 1 : 2)]++;
  }
}

/* the operation list of the last traversal (8 ints per operation, pll_operation_t order); returns what fits */
unsigned int h_get_operations(void * hv, unsigned int nops, int * out)
{
  harness_t * h = (harness_t *)hv;
  memcpy(out, h->ti->operations, (size_t)nops * sizeof(pll_operation_t));
  return nops;
}

/* ---- scalar optimisers --------------------------------------------------- */

/* analytic objectives so that Brent / L-BFGS-B / EM can be compared between the two libraries
 * without a device: the algorithms are pure host code */
typedef struct
{
  int kind;
  int n;
  int evals;
} analytic_t;

static double analytic_value(const analytic_t * a, const double * x)
{
  double f = 0.0;
  switch (a->kind)
  {
    case 0: /* extended Rosenbrock */
      for (int i = 0; i + 1 < a->n; ++i)
        f += 100.0 * (x[i + 1] - x[i] * x[i]) * (x[i + 1] - x[i] * x[i]) + (1.0 - x[i]) * (1.0 - x[i]);
      return f + 3.0;
    case 1: /* ill-conditioned quadratic with minimum outside typical boxes */
      for (int i = 0; i < a->n; ++i) f += (double)(i + 1) * (i + 1) * (x[i] - 3.0 + i) * (x[i] - 3.0 + i);
      return f + 10.0;
    case 2: /* smooth non-quadratic: sum of cosh + coupling */
      for (int i = 0; i < a->n; ++i) f += cosh(0.5 * (x[i] - 1.5)) + 0.1 * x[i] * x[(i + 1) % a->n];
      return f + 100.0;
    default: /* -log of a Gamma-like density in every coordinate (likelihood-shaped) */
      for (int i = 0; i < a->n; ++i) f += -((2.0 + i) * log(x[i]) - 1.7 * x[i]);
      return f + 50.0;
  }
}
static double analytic_cb(void * p, double * x)
{
  analytic_t * a = (analytic_t *)p;
  a->evals++;
  const double v = analytic_value(a, x);
  if (getenv("H_TRACE"))
  {
    fprintf(stderr, "eval %d f=%.17g x=", a->evals, v);
    for (int i = 0; i < a->n; ++i) fprintf(stderr, " %.17g", x[i]);
    fprintf(stderr, "\n");
  }
  return v;
}
static double analytic_cb1(void * p, double x)
{
  analytic_t * a = (analytic_t *)p;
  a->evals++;
  switch (a->kind)
  {
    case 0: return (x - 2.0) * (x - 2.0) + 1.0;
    case 1: return cosh(x - 0.3) + 0.01 * x;
    case 2: return -(3.0 * log(x) - 4.0 * x) + 20.0;
    default: return fabs(x - 0.7) + 0.5 * x * x; /* kink */
  }
}

/* out = {xopt, fx, f2x, evals} */
void h_test_brent(int kind, double xmin, double xguess, double xmax, double xtol, double * out)
{
  analytic_t a = {kind, 1, 0};
  double fx = 0, f2x = 0;
  out[0] = pllmod_opt_minimize_brent(xmin, xguess, xmax, xtol, &fx, &f2x, &a, analytic_cb1);
  out[1] = fx;
  out[2] = f2x;
  out[3] = a.evals;
}

typedef struct
{
  int n;
  int kinds[8];
  int evals;
} analytic_multi_t;
static double analytic_multi_cb(void * p, double * x, double * fx, int * converged)
{
  analytic_multi_t * a = (analytic_multi_t *)p;
  a->evals++;
  double tot = 0;
  for (int i = 0; i < a->n; ++i)
  {
    analytic_t one = {a->kinds[i], 1, 0};
    const double v = analytic_cb1(&one, x[i]);
    if (fx) fx[i] = v;
    tot += v;
  }
  if (converged)
  {
    int all = 1;
    for (int i = 0; i < a->n; ++i) all &= converged[i];
    converged[a->n] = all;
  }
  return tot;
}
/* xguess[n] in; out = xopt[n] then fx[n] then {evals} */
int h_test_brent_multi(int n, const int * kinds, const int * mask, const double * xmin,
                       double * xguess, const double * xmax, double xtol, int global_range, double * out)
{
  analytic_multi_t a;
  a.n = n;
  a.evals = 0;
  for (int i = 0; i < n; ++i) a.kinds[i] = kinds[i];
  double f2x[8] = {0};
  const int rc = pllmod_opt_minimize_brent_multi((unsigned)n, (int *)mask, (double *)xmin, xguess,
                                                 (double *)xmax, xtol, out, out + n, f2x, &a,
                                                 analytic_multi_cb, global_range);
  out[2 * n] = a.evals;
  return rc;
}

/* x[n] in/out; returns score; evals[0] = number of objective calls */
double h_test_lbfgsb(int kind, int n, double * x, double * xmin, double * xmax, int * bound,
                     double factr, double pgtol, int * evals)
{
  analytic_t a = {kind, n, 0};
  const double score = pllmod_opt_minimize_lbfgsb(x, xmin, xmax, bound, (unsigned)n, factr, pgtol, &a,
                                                  analytic_cb);
  *evals = a.evals;
  return score;
}

typedef struct
{
  int parts;
  int kinds[8];
  int n[8];
  int evals;
} lbfgsb_multi_t;
static double lbfgsb_multi_cb(void * p, double ** x, double * fx, int * flags)
{
  lbfgsb_multi_t * a = (lbfgsb_multi_t *)p;
  if (!x)
  {
    /* flag reduction (what the parallel callback does across ranks) */
    int all = 1;
    for (int i = 0; i < a->parts; ++i) all &= flags[i];
    flags[a->parts] = all;
    return 0.0;
  }
  a->evals++;
  double tot = 0;
  for (int i = 0; i < a->parts; ++i)
  {
    if (flags && flags[i]) continue;
    analytic_t one = {a->kinds[i], a->n[i], 0};
    const double v = analytic_value(&one, x[i]);
    if (fx) fx[i] = v;
    tot += v;
  }
  return tot;
}
/* xflat: concatenated x per partition (in/out), bounds likewise */
double h_test_lbfgsb_multi(int parts, const int * kinds, const int * n, double * xflat, double * lo,
                           double * hi, int * bound, double factr, double pgtol, int * evals)
{
  lbfgsb_multi_t a;
  a.parts = parts;
  a.evals = 0;
  double * xs[8], *los[8], *his[8];
  int * bs[8];
  unsigned int ns[8], nmax = 0;
  int off = 0;
  for (int i = 0; i < parts; ++i)
  {
    a.kinds[i] = kinds[i];
    a.n[i] = n[i];
    xs[i] = xflat + off, los[i] = lo + off, his[i] = hi + off, bs[i] = bound + off;
    ns[i] = (unsigned)n[i];
    if (ns[i] > nmax) nmax = ns[i];
    off += n[i];
  }
  const double s = pllmod_opt_minimize_lbfgsb_multi((unsigned)parts, xs, los, his, bs, ns, nmax, factr,
                                                    pgtol, &a, lbfgsb_multi_cb);
  *evals = a.evals;
  return s;
}

typedef struct
{
  unsigned int l, c;
  const double * base; /* l*c per-site, per-category conditional likelihoods */
  double * w;
} em_t;
static double em_cb(void * p, double * sitecat)
{
  em_t * e = (em_t *)p;
  for (unsigned int i = 0; i < e->l; ++i)
    for (unsigned int k = 0; k < e->c; ++k) sitecat[i * e->c + k] = e->base[i * e->c + k] * e->w[k];
  return 0.0;
}
void h_test_em(unsigned int l, unsigned int c, const double * base, unsigned int * site_w, double * w)
{
  em_t e = {l, c, base, w};
  double * buf = (double *)calloc((size_t)l * c, sizeof(double));
  pllmod_opt_minimize_em(w, c, buf, site_w, l, &e, em_cb);
  free(buf);
}

/* model-parameter optimisation of partition p at the current root edge through
 * pllmod_opt_optimize_onedim (Brent) or _multidim (L-BFGS-B); `which` = PLLMOD_OPT_PARAM_* mask.
 * out = {score, alpha, pinv, rates[6 or n], freqs[states], ...} filled by h_get_model */
static int cb_trav_all(pll_unode_t * n)
{
  (void)n;
  return 1;
}
double h_optimize_params(void * hv, unsigned int p, unsigned int which, double factr, double pgtol,
                         double alpha_guess, const int * symmetries, int use_onedim)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  pll_partition_t * part = h->parts[p];
  const unsigned int tips = h->tips, nodes = 2 * tips - 2;
  pll_unode_t ** trav = (pll_unode_t **)malloc(sizeof(pll_unode_t *) * nodes);
  unsigned int trav_size = 0, mcount = 0, ocount = 0;
  pll_unode_t * root = ti->root;
  if (!pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, cb_trav_all, trav, &trav_size)) return NAN;
  double * brlens = (double *)malloc(sizeof(double) * (2 * tips - 3));
  unsigned int * midx = (unsigned int *)malloc(sizeof(unsigned int) * (2 * tips - 3));
  pll_operation_t * ops = (pll_operation_t *)malloc(sizeof(pll_operation_t) * (tips - 2));
  pll_utree_create_operations(trav, trav_size, brlens, midx, ops, &mcount, &ocount);

  pll_optimize_options_t o;
  memset(&o, 0, sizeof(o));
  o.lk_params.partition = part;
  o.lk_params.operations = ops;
  o.lk_params.branch_lengths = brlens;
  o.lk_params.matrix_indices = midx;
  o.lk_params.rooted = 0;
  o.lk_params.params_indices = h->param_idx[p];
  o.lk_params.alpha_value = alpha_guess;
  o.lk_params.where.unrooted_t.parent_clv_index = root->clv_index;
  o.lk_params.where.unrooted_t.parent_scaler_index = root->scaler_index;
  o.lk_params.where.unrooted_t.child_clv_index = root->back->clv_index;
  o.lk_params.where.unrooted_t.child_scaler_index = root->back->scaler_index;
  o.lk_params.where.unrooted_t.edge_pmatrix_index = root->pmatrix_index;
  o.params_index = 0;
  o.which_parameters = which;
  o.subst_params_symmetries = (int *)symmetries;
  o.factr = factr;
  o.pgtol = pgtol;
  /* bring P-matrices and CLVs up to date first (the reference callers do the same) */
  pll_update_prob_matrices(part, h->param_idx[p], midx, brlens, mcount);
  pll_update_partials(part, ops, ocount);
  const double score = use_onedim ? pllmod_opt_optimize_onedim(&o, 0, 0) : pllmod_opt_optimize_multidim(&o, NULL, NULL);
  free(trav);
  free(brlens);
  free(midx);
  free(ops);
  pllmod_treeinfo_invalidate_all(ti);
  return score;
}

/* out: alpha-independent model read-back = rates[K], rate_weights[K], pinv, subst[s(s-1)/2], freqs[s] */
void h_get_model(void * hv, unsigned int p, double * out)
{
  pll_partition_t * part = ((harness_t *)hv)->parts[p];
  const unsigned int K = part->rate_cats, s = part->states, nr = s * (s - 1) / 2;
  memcpy(out, part->rates, sizeof(double) * K);
  memcpy(out + K, part->rate_weights, sizeof(double) * K);
  out[2 * K] = part->prop_invar[0];
  memcpy(out + 2 * K + 1, part->subst_params[0], sizeof(double) * nr);
  memcpy(out + 2 * K + 1 + nr, part->frequencies[0], sizeof(double) * s);
}

/* ---- SPR round (BASELINE configs[4]) -------------------------------------- */
#include "pllmod_algorithm.h"

double h_spr_round(void * hv, unsigned int radius_min, unsigned int radius_max, unsigned int ntopol_keep,
                   int thorough, int blo_method, double bl_min, double bl_max, int smoothings,
                   double epsilon, int use_cutoff, double subtree_cutoff, double lh_eps_triplet,
                   int fast_clv_updates, double * cutoff_out)
{
  harness_t * h = (harness_t *)hv;
  cutoff_info_t ci;
  memset(&ci, 0, sizeof(ci));
  ci.lh_cutoff = 1e30; /* first round: no cutoff yet (what raxml-ng passes) */
  const double lh = pllmod_algo_spr_round(h->ti, radius_min, radius_max, ntopol_keep, thorough, blo_method,
                                          bl_min, bl_max, smoothings, epsilon, use_cutoff ? &ci : NULL,
                                          subtree_cutoff, lh_eps_triplet, fast_clv_updates);
  if (cutoff_out)
  {
    cutoff_out[0] = ci.lh_cutoff;
    cutoff_out[1] = ci.lh_dec_sum;
    cutoff_out[2] = ci.lh_dec_count;
  }
  return lh;
}

/* order-independent description of the current topology: for every branch the hash of the tip set on
 * the side that does not contain tip 0, with the branch length; sorted by hash.
 * out_hash / out_len: 2*tips-3 entries */
static unsigned long long split_hash(pll_unode_t * n, int * has0)
{
  if (!n->next)
  {
    if (n->clv_index == 0) *has0 = 1;
    return mix64(0xABCDEFull + n->clv_index);
  }
  return split_hash(n->next->back, has0) + split_hash(n->next->next->back, has0);
}
void h_topology_signature(void * hv, unsigned long long * out_hash, double * out_len)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  const unsigned int nb = 2 * h->tips - 3;
  unsigned long long total = 0;
  for (unsigned int t = 0; t < h->tips; ++t) total += mix64(0xABCDEFull + t);
  unsigned int k = 0;
  char * seen = (char *)calloc(ti->subnode_count, 1);
  for (unsigned int i = 0; i < ti->subnode_count && k < nb; ++i)
  {
    pll_unode_t * n = ti->subnodes[i];
    if (seen[n->node_index] || seen[n->back->node_index]) continue;
    seen[n->node_index] = seen[n->back->node_index] = 1;
    int has0 = 0;
    unsigned long long hsh = split_hash(n->back, &has0); /* subtree behind n->back's far side */
    if (has0) hsh = total - hsh;
    out_hash[k] = hsh;
    out_len[k] = n->length;
    ++k;
  }
  free(seen);
  /* insertion sort by hash */
  for (unsigned int i = 1; i < k; ++i)
  {
    unsigned long long hh = out_hash[i];
    double ll = out_len[i];
    unsigned int j = i;
    while (j > 0 && out_hash[j - 1] > hh)
    {
      out_hash[j] = out_hash[j - 1];
      out_len[j] = out_len[j - 1];
      --j;
    }
    out_hash[j] = hh;
    out_len[j] = ll;
  }
}

/* ---- treeinfo-level model optimisers ---------------------------------------- */
/* which: 0 alpha (Brent)  1 pinv (Brent)  2 subst rates  3 frequencies  4 alpha+pinv  5 free rates+weights
 *        6 branch-length scalers (SCALED linkage)  7 branch lengths (NR, all branches)
 * params_to_optimize of every partition is set to the family under test first. Returns -logL. */
double h_algo_opt_treeinfo(void * hv, int which, double tolerance)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  static const int masks[8] = {PLLMOD_OPT_PARAM_ALPHA, PLLMOD_OPT_PARAM_PINV, PLLMOD_OPT_PARAM_SUBST_RATES,
                               PLLMOD_OPT_PARAM_FREQUENCIES, PLLMOD_OPT_PARAM_ALPHA | PLLMOD_OPT_PARAM_PINV,
                               PLLMOD_OPT_PARAM_FREE_RATES | PLLMOD_OPT_PARAM_RATE_WEIGHTS,
                               PLLMOD_OPT_PARAM_BRANCH_LEN_SCALER, 0};
  for (unsigned int p = 0; p < ti->partition_count; ++p) ti->params_to_optimize[p] = masks[which];
  pllmod_treeinfo_compute_loglh(ti, 0);
  switch (which)
  {
    case 0: return pllmod_algo_opt_onedim_treeinfo(ti, PLLMOD_OPT_PARAM_ALPHA, 0.02, 100.0, tolerance);
    case 1: return pllmod_algo_opt_onedim_treeinfo(ti, PLLMOD_OPT_PARAM_PINV, 1e-4, 0.99, tolerance);
    case 2: return pllmod_algo_opt_subst_rates_treeinfo(ti, 0, 1e-3, 1000.0, 1e7, tolerance);
    case 3: return pllmod_algo_opt_frequencies_treeinfo(ti, 0, 1e-3, 100.0, 1e7, tolerance);
    case 4: return pllmod_algo_opt_alpha_pinv_treeinfo(ti, 0, 0.02, 100.0, 1e-4, 0.99, 1e7, tolerance);
    case 5: return pllmod_algo_opt_rates_weights_treeinfo(ti, 0.02, 100.0, 1e-6, 100.0, 1e7, tolerance);
    case 6: return pllmod_algo_opt_brlen_scalers_treeinfo(ti, 0.01, 100.0, 1e-6, 100.0, tolerance);
    default: return pllmod_algo_opt_brlen_treeinfo(ti, 1e-6, 100.0, tolerance, 4, PLLMOD_OPT_BLO_NEWTON_FAST, -1);
  }
}

/* out: alphas[nparts] then brlen_scalers[nparts] (1.0 when the treeinfo has none) */
void h_get_treeinfo_params(void * hv, double * out)
{
  harness_t * h = (harness_t *)hv;
  for (unsigned int p = 0; p < h->nparts; ++p)
  {
    out[p] = h->ti->alphas[p];
    out[h->nparts + p] = h->ti->brlen_scalers ? h->ti->brlen_scalers[p] : 1.0;
  }
}

/* ---- binary checkpoint I/O (src/binary/pll_binary.h) ----------------------- */
#include "pll_binary.h"

#define H_BIN_PARTITION 10
#define H_BIN_TREE 11
#define H_BIN_CUSTOM 12
#define H_BIN_CLV 13

/* partition 0 (with CLVs, scalers, P-matrices, weights), the tree, the branch lengths as a custom
 * block and the root CLV as a block of its own, in a random-access file */
int h_binary_dump(void * hv, const char * path)
{
  harness_t * h = (harness_t *)hv;
  pll_binary_header_t hd;
  FILE * f = pllmod_binary_create(path, &hd, PLLMOD_BIN_ACCESS_RANDOM, 8);
  if (!f) return 0;
  pll_unode_t * root = h->ti->root;
  const unsigned int nbr = 2 * h->tips - 3;
  int ok = pllmod_binary_partition_dump(f, H_BIN_PARTITION, h->parts[0],
                                        PLLMOD_BIN_ATTRIB_UPDATE_MAP | PLLMOD_BIN_ATTRIB_PARTITION_DUMP_CLV |
                                            PLLMOD_BIN_ATTRIB_PARTITION_DUMP_WGT);
  ok = ok && pllmod_binary_utree_dump(f, H_BIN_TREE, root, h->tips, PLLMOD_BIN_ATTRIB_UPDATE_MAP);
  ok = ok && pllmod_binary_custom_dump(f, H_BIN_CUSTOM, h->ti->branch_lengths[0], sizeof(double) * nbr,
                                       PLLMOD_BIN_ATTRIB_UPDATE_MAP);
  ok = ok && pllmod_binary_clv_dump(f, H_BIN_CLV, h->parts[0], root->clv_index, PLLMOD_BIN_ATTRIB_UPDATE_MAP);
  pllmod_binary_close(f);
  return ok;
}

static double tree_length_rec(const pll_unode_t * n, unsigned int * tips)
{
  if (!n->next)
  {
    ++*tips;
    return 0.0;
  }
  return n->next->length + tree_length_rec(n->next->back, tips) + n->next->next->length +
         tree_length_rec(n->next->next->back, tips);
}

/* Open a file written by h_binary_dump (of either library) and use what is in it WITHOUT recomputing
 * anything: out[0] = edge logL at the stored root from the loaded CLVs / scalers / P-matrices / weights,
 * out[1] = the same after the root CLV was loaded again from its own block, out[2] = tips of the loaded
 * tree, out[3] = its total branch length, out[4] = sum of the custom block's doubles, out[5] = blocks in
 * the map, out[6] = attributes stored with the partition block.  Returns 1 on success. */
int h_binary_load(const char * path, double * out)
{
  pll_binary_header_t hd;
  FILE * f = pllmod_binary_open(path, &hd);
  if (!f) return 0;
  int ok = 0;
  unsigned int attr = 0, tattr = 0, cattr = 0, ctype = 0, vattr = 0, nblocks = 0;
  size_t csize = 0;
  pll_partition_t * part = NULL;
  pll_unode_t * root = NULL;
  double * custom = NULL;
  pll_block_map_t * map = pllmod_binary_get_map(f, &nblocks);
  part = pllmod_binary_partition_load(f, H_BIN_PARTITION, NULL, &attr, PLLMOD_BIN_ACCESS_SEEK);
  if (part) root = pllmod_binary_utree_load(f, H_BIN_TREE, &tattr, PLLMOD_BIN_ACCESS_SEEK);
  if (root) custom = (double *)pllmod_binary_custom_load(f, H_BIN_CUSTOM, &csize, &ctype, &cattr, PLLMOD_BIN_ACCESS_SEEK);
  if (custom && ctype == PLLMOD_BIN_BLOCK_CUSTOM)
  {
    unsigned int * pidx = (unsigned int *)calloc(part->rate_cats, sizeof(unsigned int));
    out[0] = pll_compute_edge_loglikelihood(part, root->clv_index, root->scaler_index, root->back->clv_index,
                                            root->back->scaler_index, root->pmatrix_index, pidx, NULL);
    ok = pllmod_binary_clv_load(f, H_BIN_CLV, part, root->clv_index, &vattr, PLLMOD_BIN_ACCESS_SEEK);
    out[1] = pll_compute_edge_loglikelihood(part, root->clv_index, root->scaler_index, root->back->clv_index,
                                            root->back->scaler_index, root->pmatrix_index, pidx, NULL);
    unsigned int tips = 0;
    out[3] = root->length + tree_length_rec(root, &tips) + tree_length_rec(root->back, &tips);
    out[2] = tips;
    out[4] = 0.0;
    for (size_t i = 0; i < csize / sizeof(double); ++i) out[4] += custom[i];
    out[5] = nblocks;
    out[6] = attr;
    free(pidx);
  }
  free(custom);
  free(map);
  if (root) pll_utree_graph_destroy(root, NULL);
  if (part) pll_partition_destroy(part);
  pllmod_binary_close(f);
  return ok;
}

/* tree + custom blocks only (no partition, so no device): sequential file.  h_binary_tree_load returns
 * tips, total length, label checksum and the custom payload sum in out[0..3]. */
int h_binary_tree_dump(const char * path, unsigned int tips, unsigned int seed)
{
  char ** names = (char **)malloc(sizeof(char *) * tips);
  for (unsigned int i = 0; i < tips; ++i)
  {
    names[i] = (char *)malloc(24);
    snprintf(names[i], 24, "taxon_%u_%u", i, (i * 2654435761u) % 97u);
  }
  pll_utree_t * t = pllmod_utree_create_random(tips, (const char * const *)names, seed);
  for (unsigned int i = 0; i < tips; ++i) free(names[i]);
  free(names);
  if (!t) return 0;
  unsigned long long s = 88172645463325252ull + seed;
  for (unsigned int i = 0; i < t->tip_count + 3 * t->inner_count; ++i)
  {
    pll_unode_t * n = t->nodes[i < t->tip_count ? i : t->tip_count + (i - t->tip_count) / 3];
    if (i >= t->tip_count)
      for (unsigned int r = 0; r < (i - t->tip_count) % 3; ++r) n = n->next;
    if (n->node_index < n->back->node_index) n->length = n->back->length = 0.001 + u01(&s);
  }
  pll_binary_header_t hd;
  FILE * f = pllmod_binary_create(path, &hd, PLLMOD_BIN_ACCESS_SEQUENTIAL, 0);
  if (!f) return 0;
  double payload[5] = {1.5, -2.25, 3.0, (double)tips, (double)seed};
  int ok = pllmod_binary_custom_dump(f, 1, payload, sizeof(payload), 0);
  ok = ok && pllmod_binary_utree_dump(f, 2, t->vroot, tips, 0);
  pllmod_binary_close(f);
  pll_utree_destroy(t, NULL);
  return ok;
}

static void label_sum_rec(const pll_unode_t * n, double * acc)
{
  if (!n->next)
  {
    if (n->label)
      for (const char * c = n->label; *c; ++c) *acc += (double)(unsigned char)*c * (1 + n->node_index % 7);
    return;
  }
  label_sum_rec(n->next->back, acc);
  label_sum_rec(n->next->next->back, acc);
}

int h_binary_tree_load(const char * path, double * out)
{
  pll_binary_header_t hd;
  FILE * f = pllmod_binary_open(path, &hd);
  if (!f) return 0;
  size_t csize = 0;
  unsigned int ctype = 0, cattr = 0, tattr = 0, tips = 0;
  double * custom = (double *)pllmod_binary_custom_load(f, 1, &csize, &ctype, &cattr, 0);
  pll_unode_t * root = custom ? pllmod_binary_utree_load(f, 2, &tattr, 0) : NULL;
  int ok = root != NULL;
  if (ok)
  {
    out[1] = root->length + tree_length_rec(root, &tips) + tree_length_rec(root->back, &tips);
    out[0] = tips;
    out[2] = 0.0;
    label_sum_rec(root, &out[2]);
    label_sum_rec(root->back, &out[2]);
    out[3] = 0.0;
    for (size_t i = 0; i < csize / sizeof(double); ++i) out[3] += custom[i];
    out[4] = hd.n_blocks;
    pll_utree_graph_destroy(root, NULL);
  }
  free(custom);
  pllmod_binary_close(f);
  return ok;
}

/* ---- marginal ancestral states (treeinfo.c:1611-1712) ----------------------- */

/* out: node_count * sum_p(sites_p * states_p) doubles, nodes in the order of the result; labels_sum gets a
 * checksum of the generated node labels.  Returns the node count (0 on failure). */
unsigned int h_ancestral(void * hv, double * out, double * labels_sum)
{
  harness_t * h = (harness_t *)hv;
  pllmod_ancestral_t * a = pllmod_treeinfo_compute_ancestral(h->ti);
  if (!a) return 0;
  size_t span = 0;
  for (unsigned int p = 0; p < h->nparts; ++p) span += (size_t)h->parts[p]->sites * h->parts[p]->states;
  *labels_sum = 0.0;
  /* results are keyed by clv_index so that both libraries report the nodes in one order */
  for (unsigned int i = 0; i < a->node_count; ++i)
  {
    const unsigned int slot = a->nodes[i]->clv_index - h->tips;
    memcpy(out + (size_t)slot * span, a->probs[i], sizeof(double) * span);
    for (const char * c = a->nodes[i]->label; c && *c; ++c) *labels_sum += (double)(unsigned char)*c * (1 + slot % 5);
  }
  const unsigned int n = a->node_count;
  pllmod_treeinfo_destroy_ancestral(a);
  return n;
}

/* ---- a13 entry points that the other drivers do not reach (pll_optimize.c:1115,1146,1182) ---------- */

/* pllmod_opt_optimize_branch_lengths_iterative on partition 0 (all branches, legacy driver) */
double h_optimize_branch_lengths_iterative(void * hv, int smoothings, double tolerance)
{
  harness_t * h = (harness_t *)hv;
  return pllmod_opt_optimize_branch_lengths_iterative(h->parts[0], h->ti->root, h->param_idx[0], 0, 0, tolerance,
                                                      smoothings, 1);
}

/* exported by the reference (pll_optimize.c:1182, PLL_EXPORT) but missing from its header pll_optimize.h */
double pllmod_opt_compute_edge_loglikelihood_multi(pll_partition_t ** partitions, size_t partition_count,
                                                   unsigned int parent_clv_index, int parent_scaler_index,
                                                   unsigned int child_clv_index, int child_scaler_index,
                                                   unsigned int matrix_index, unsigned int ** const params_indices,
                                                   double * persite_lnl, void * parallel_context,
                                                   void (*parallel_reduce_cb)(void *, double *, size_t, int));

/* pllmod_opt_compute_edge_loglikelihood_multi over all partitions at the edge of subnode node_index;
 * persite (may be NULL): sum of sites of all partitions doubles */
double h_edge_loglh_multi(void * hv, unsigned int node_index, double * persite)
{
  harness_t * h = (harness_t *)hv;
  pllmod_treeinfo_t * ti = h->ti;
  pll_unode_t * n = ti->subnodes[node_index];
  return pllmod_opt_compute_edge_loglikelihood_multi(ti->partitions, ti->partition_count, n->clv_index,
                                                     n->scaler_index, n->back->clv_index, n->back->scaler_index,
                                                     n->pmatrix_index, ti->param_indices, persite,
                                                     ti->parallel_context, ti->parallel_reduce_cb);
}

/* pllmod_opt_derivative_func: the Newton callback of the single-partition driver, fed by hand */
int h_derivative_func(void * hv, unsigned int p, unsigned int node_index, const double * brlens, int n,
                      double * df, double * ddf)
{
  harness_t * h = (harness_t *)hv;
  pll_unode_t * e = h->ti->subnodes[node_index];
  if (!pll_update_sumtable(h->parts[p], e->clv_index, e->back->clv_index, e->scaler_index,
                           e->back->scaler_index, h->param_idx[p], h->sumtables[p]))
    return 0;
  pll_optimize_options_t o;
  memset(&o, 0, sizeof(o));
  o.lk_params.partition = h->parts[p];
  o.lk_params.rooted = 0;
  o.lk_params.params_indices = h->param_idx[p];
  o.lk_params.where.unrooted_t.parent_clv_index = e->clv_index;
  o.lk_params.where.unrooted_t.parent_scaler_index = e->scaler_index;
  o.lk_params.where.unrooted_t.child_clv_index = e->back->clv_index;
  o.lk_params.where.unrooted_t.child_scaler_index = e->back->scaler_index;
  o.lk_params.where.unrooted_t.edge_pmatrix_index = e->pmatrix_index;
  o.sumtable = h->sumtables[p];
  for (int i = 0; i < n; ++i) pllmod_opt_derivative_func(&o, brlens[i], &df[i], &ddf[i]);
  return 1;
}

/* ---- thread safety across partitions (SURVEY.md 8b "Threading"): every handle is driven from its own
 *      pthread at the same time — full logL, one Newton-Raphson smoothing pass, logL again.
 *      out[3 * i + {0,1,2}] per handle. ---- */
#include <pthread.h>
typedef struct
{
  harness_t * h;
  double * out;
} thread_job_t;

static void * thread_job(void * arg)
{
  thread_job_t * j = (thread_job_t *)arg;
  pllmod_treeinfo_t * ti = j->h->ti;
  j->out[0] = pllmod_treeinfo_compute_loglh(ti, 0);
  j->out[1] = pllmod_opt_optimize_branch_lengths_local_multi(
      ti->partitions, ti->partition_count, ti->root, ti->param_indices, ti->deriv_precomp, ti->branch_lengths,
      ti->brlen_scalers, 0, 0, 0.1, 2, -1, 1, PLLMOD_OPT_BLO_NEWTON_FAST, ti->brlen_linkage, ti->parallel_context,
      ti->parallel_reduce_cb);
  j->out[2] = pllmod_treeinfo_compute_loglh(ti, 0);
  return NULL;
}

int h_run_concurrently(void ** handles, int n, int threaded, double * out)
{
  pthread_t th[16];
  thread_job_t jobs[16];
  if (n > 16) return 0;
  for (int i = 0; i < n; ++i)
  {
    jobs[i].h = (harness_t *)handles[i];
    jobs[i].out = out + 3 * i;
  }
  if (!threaded)
  {
    for (int i = 0; i < n; ++i) thread_job(&jobs[i]);
    return 1;
  }
  for (int i = 0; i < n; ++i)
    if (pthread_create(&th[i], NULL, thread_job, &jobs[i])) return 0;
  for (int i = 0; i < n; ++i) pthread_join(th[i], NULL);
  return 1;
}

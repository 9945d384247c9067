/*
 * pll_oracle.c — TEST INFRASTRUCTURE: a plain-C restatement of the reference's arithmetic for the
 * likelihood hot path.  Scalar loops, no SIMD, no GPU; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline leg of bench.py may load it, and only as the checker.  The product never links it.
 *
 * PINNING: this restatement is checked (tests/test_oracle_cpu.py) against
 *   (1) the unmodified reference compiled from /root/reference into oracle/_ref (same seeded inputs,
 *       logL to 1e-10, derivatives to 1e-8, scaler counts exact), and
 *   (2) the committed golden fixtures that reference produced (tests/golden/ref_cases.json).
 * Each function cites the reference code it follows (paths relative to the reference tree).
 *
 * The synthetic-input generators (tree, alignment) are the same ones oracle/harness.c uses, so a case is
 * fully described by (tips, sites, states, rate_cats, seed, brlen_mean, brlen_scale, alpha, pinv, model).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORC_SCALE_FACTOR 115792089237316195423570985008687907853269984665640564039457584007913129639936.0 /* 2^256 */
#define ORC_SCALE_THRESHOLD (1.0 / ORC_SCALE_FACTOR)
#define ORC_MAXDIFF 4
#define ORC_MISC_EPSILON 1e-8
#define ORC_EIGEN_MINFREQ 1e-6

/* ------------------------------------------------------------------------------------------------ */
/* Gamma rates — libs/libpll/src/gamma.c:24-292 (AS32, ACM291, AS70, AS91)                          */
/* ------------------------------------------------------------------------------------------------ */
static double orc_lngamma(double alpha)
{
  double x = alpha, f = 0.0, z;
  if (x < 7.0)
  {
    f = 1.0;
    z = alpha - 1.0;
    while ((z = z + 1.0) < 7.0) f *= z;
    x = z;
    f = -log(f);
  }
  z = 1 / (x * x);
  return f + (x - 0.5) * log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z + .083333333333333) / x;
}

static double orc_incgamma(double x, double alpha, double lnga)
{
  const double accurate = 1e-8, overflow = 1e30;
  double p = alpha, factor, gin, rn, a, b, an, dif, term, pn[6];
  int i;
  if (x == 0) return 0;
  if (x < 0 || p <= 0) return -1;
  factor = exp(p * log(x) - x - lnga);
  if (!(x > 1 && x >= p))
  {
    gin = 1; term = 1; rn = p;
    do { rn++; term *= x / rn; gin += term; } while (term > accurate);
    return gin * factor / p;
  }
  a = 1 - p; b = a + x + 1; term = 0;
  pn[0] = 1; pn[1] = x; pn[2] = x + 1; pn[3] = x * b;
  gin = pn[2] / pn[3];
  for (;;)
  {
    a++; b += 2; term++;
    an = a * term;
    for (i = 0; i < 2; i++) pn[i + 4] = b * pn[i + 2] - an * pn[i];
    if (pn[5] != 0)
    {
      rn = pn[4] / pn[5];
      dif = fabs(gin - rn);
      if (dif <= accurate && dif <= accurate * rn) return 1 - factor * gin;
      gin = rn;
    }
    for (i = 0; i < 4; i++) pn[i] = pn[i + 2];
    if (fabs(pn[4]) >= overflow)
      for (i = 0; i < 4; i++) pn[i] /= overflow;
  }
}

static double orc_pointnormal(double prob)
{
  const double a0 = -.322232431088, a1 = -1, a2 = -.342242088547, a3 = -.0204231210245, a4 = -.453642210148e-4;
  const double b0 = .0993484626060, b1 = .588581570495, b2 = .531103462366, b3 = .103537752850, b4 = .0038560700634;
  double y, z, p1 = prob < 0.5 ? prob : 1 - prob;
  if (p1 < 1e-20) return -9999;
  y = sqrt(log(1 / (p1 * p1)));
  z = y + ((((y * a4 + a3) * y + a2) * y + a1) * y + a0) / ((((y * b4 + b3) * y + b2) * y + b1) * y + b0);
  return prob < 0.5 ? -z : z;
}

static double orc_pointchi2(double prob, double v)
{
  const double e = .5e-6, aa = .6931471805;
  double p = prob, g, xx, c, ch, a, q, p1, p2, t, x, b, s1, s2, s3, s4, s5, s6;
  if (p < .000002 || p > .999998 || v <= 0) return -1;
  g = orc_lngamma(v / 2);
  xx = v / 2; c = xx - 1;
  if (v < -1.24 * log(p))
  {
    ch = pow((p * xx * exp(g + xx * aa)), 1 / xx);
    if (ch - e < 0) return ch;
  }
  else if (v <= .32)
  {
    ch = 0.4; a = log(1 - p);
    do
    {
      q = ch; p1 = 1 + ch * (4.67 + ch); p2 = ch * (6.73 + ch * (6.66 + ch));
      t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - exp(a + g + .5 * ch + c * aa) * p2 / p1) / t;
    } while (fabs(q / ch - 1) - .01 > 0);
  }
  else
  {
    x = orc_pointnormal(p);
    p1 = 0.222222 / v; ch = v * pow((x * sqrt(p1) + 1 - p1), 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (log(1 - p) - c * log(.5 * ch) + g);
  }
  do
  {
    q = ch; p1 = .5 * ch;
    if ((t = orc_incgamma(p1, xx, g)) < 0) return -1;
    p2 = p - t;
    t = p2 * exp(xx * aa + g + p1 - c * log(ch));
    b = t / ch; a = 0.5 * t - b * c;
    s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 - b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (fabs(q / ch - 1) > e);
  return ch;
}

/* gamma.c:220-292, PLL_GAMMA_RATES_MEAN (mode 0) / MEDIAN (mode 1) */
int orc_gamma_rates(double alpha, unsigned int cats, int mode, double * out)
{
  unsigned int i;
  const double factor = cats;
  if (alpha < 0.02 || cats < 1) return 0;
  if (cats == 1) { out[0] = 1.0; return 1; }
  if (mode == 1)
  {
    double t = 0, middle = 1.0 / (2.0 * cats);
    for (i = 0; i < cats; i++) { out[i] = orc_pointchi2((i * 2 + 1) * middle, 2 * alpha) / (2 * alpha); t += out[i]; }
    for (i = 0; i < cats; i++) out[i] *= factor / t;
    return 1;
  }
  {
    double * gp = (double *)malloc(sizeof(double) * cats);
    const double lnga1 = orc_lngamma(alpha + 1);
    for (i = 0; i < cats - 1; i++) gp[i] = orc_pointchi2((i + 1.0) / cats, 2 * alpha) / (2 * alpha);
    for (i = 0; i < cats - 1; i++) gp[i] = orc_incgamma(gp[i] * alpha, alpha + 1, lnga1);
    out[0] = gp[0] * factor;
    out[cats - 1] = (1 - gp[cats - 2]) * factor;
    for (i = 1; i < cats - 1; i++) out[i] = (gp[i] - gp[i - 1]) * factor;
    free(gp);
  }
  return 1;
}

/* ------------------------------------------------------------------------------------------------ */
/* eigen-decomposition — models.c:182-410.  The reference uses Householder + QL; any accurate         */
/* symmetric solver gives the same P-matrices, this one is a cyclic Jacobi.                         */
/* V[m][j] = U[m][j] sqrt(pi_j), Vinv[i][m] = U[m][i]/sqrt(pi_i), eliminated states: identity rows   */
/* ------------------------------------------------------------------------------------------------ */
int orc_eigen(unsigned int s, const double * subst, const double * freqs, double * evecs, double * inv_evecs,
              double * evals)
{
  const unsigned int np = s * (s - 1) / 2;
  double * r = (double *)malloc(sizeof(double) * (np ? np : 1));
  double * Q = (double *)calloc((size_t)s * s, sizeof(double));
  int * keep = (int *)malloc(sizeof(int) * s);
  unsigned int i, j, k = 0, n = 0;
  double mean = 0;
  memcpy(r, subst, sizeof(double) * np);
  if (np && r[np - 1] > 0) for (i = 0; i < np; ++i) r[i] /= subst[np - 1];
  for (i = 0; i < s; ++i)
    for (j = i + 1; j < s; ++j, ++k)
    {
      const double f = (freqs[i] <= ORC_EIGEN_MINFREQ || freqs[j] <= ORC_EIGEN_MINFREQ) ? 0 : r[k];
      Q[i * s + j] = Q[j * s + i] = f * sqrt(freqs[i] * freqs[j]);
      Q[i * s + i] -= f * freqs[j];
      Q[j * s + j] -= f * freqs[i];
    }
  for (i = 0; i < s; ++i) mean += freqs[i] * (-Q[i * s + i]);
  for (i = 0; i < s * s; ++i) Q[i] /= mean;
  for (i = 0; i < s; ++i) if (freqs[i] > ORC_EIGEN_MINFREQ) keep[n++] = (int)i;
  {
    double * A = (double *)malloc(sizeof(double) * n * n);
    double * U = (double *)calloc((size_t)n * n, sizeof(double));
    unsigned int a, b, p, q, sweep;
    for (a = 0; a < n; ++a) for (b = 0; b < n; ++b) A[a * n + b] = Q[keep[a] * s + keep[b]];
    for (a = 0; a < n; ++a) U[a * n + a] = 1;
    for (sweep = 0; sweep < 100; ++sweep)
    {
      double off = 0, dg = 0;
      for (a = 0; a < n; ++a) { dg += A[a * n + a] * A[a * n + a]; for (b = a + 1; b < n; ++b) off += A[a * n + b] * A[a * n + b]; }
      if (off <= 1e-40 * (dg > 0 ? dg : 1)) break;
      for (p = 0; p + 1 < n; ++p)
        for (q = p + 1; q < n; ++q)
        {
          const double apq = A[p * n + q];
          double theta, t, c, sn;
          if (apq == 0) continue;
          theta = (A[q * n + q] - A[p * n + p]) / (2 * apq);
          t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1));
          c = 1 / sqrt(t * t + 1); sn = t * c;
          for (k = 0; k < n; ++k) { const double x = A[k * n + p], y = A[k * n + q]; A[k * n + p] = c * x - sn * y; A[k * n + q] = sn * x + c * y; }
          for (k = 0; k < n; ++k) { const double x = A[p * n + k], y = A[q * n + k]; A[p * n + k] = c * x - sn * y; A[q * n + k] = sn * x + c * y; }
          for (k = 0; k < n; ++k) { const double x = U[p * n + k], y = U[q * n + k]; U[p * n + k] = c * x - sn * y; U[q * n + k] = sn * x + c * y; }
        }
    }
    memset(evecs, 0, sizeof(double) * s * s);
    memset(inv_evecs, 0, sizeof(double) * s * s);
    memset(evals, 0, sizeof(double) * s);
    for (i = 0; i < s; ++i) evecs[i * s + i] = inv_evecs[i * s + i] = 1;
    for (a = 0; a < n; ++a) { evals[keep[a]] = A[a * n + a]; evecs[keep[a] * s + keep[a]] = inv_evecs[keep[a] * s + keep[a]] = 0; }
    for (a = 0; a < n; ++a)
      for (b = 0; b < n; ++b)
      {
        evecs[keep[a] * s + keep[b]] = U[a * n + b] * sqrt(freqs[keep[b]]);
        inv_evecs[keep[a] * s + keep[b]] = U[b * n + a] / sqrt(freqs[keep[a]]);
      }
    free(A); free(U);
  }
  free(r); free(Q); free(keep);
  return 1;
}

/* core_pmatrix.c:187-253: P_k = I + Vinv diag(expm1(lambda r_k t / (1-pinv))) V ; t == 0 -> I */
void orc_pmatrix(unsigned int s, unsigned int K, const double * rates, double pinv, const double * evals,
                 const double * evecs, const double * inv_evecs, double t, double * out)
{
  unsigned int k, i, j, m;
  double * expd = (double *)malloc(sizeof(double) * s);
  for (k = 0; k < K; ++k)
  {
    double * P = out + (size_t)k * s * s;
    if (t > 0)
    {
      for (j = 0; j < s; ++j)
        expd[j] = (pinv > ORC_MISC_EPSILON) ? expm1(evals[j] * rates[k] * t / (1.0 - pinv)) : expm1(evals[j] * rates[k] * t);
      for (i = 0; i < s; ++i)
        for (j = 0; j < s; ++j)
        {
          double v = (i == j) ? 1.0 : 0.0;
          for (m = 0; m < s; ++m) v += (inv_evecs[i * s + m] * expd[m]) * evecs[m * s + j];
          P[i * s + j] = v;
        }
    }
    else
      for (i = 0; i < s; ++i) for (j = 0; j < s; ++j) P[i * s + j] = (i == j);
  }
  free(expd);
}

/* ------------------------------------------------------------------------------------------------ */
/* CLV update — core_partials.c:24-46 (scaler init), :709-764 (inner-inner), :290-351 / :1013-1210   */
/* (tip sides as masked row sums; tip-tip never rescales, :185-186).                                */
/* A child is either a CLV (clv != NULL) or a row of state bitmasks (mask != NULL).                 */
/* per_rate: PLL_ATTRIB_RATE_SCALERS semantics (one counter per site and category).                 */
/* ------------------------------------------------------------------------------------------------ */
void orc_update_partial(unsigned int s, unsigned int sites, unsigned int K, const double * lclv,
                        const unsigned long long * lmask, const double * rclv, const unsigned long long * rmask,
                        const double * lmat, const double * rmat, const unsigned int * lscaler,
                        const unsigned int * rscaler, double * parent, unsigned int * pscaler, int per_rate)
{
  const unsigned int span = K * s;
  const size_t nsc = (size_t)sites * (per_rate ? K : 1);
  unsigned int n, k, i, j;
  if (pscaler)
  {
    size_t q;
    for (q = 0; q < nsc; ++q) pscaler[q] = (lscaler ? lscaler[q] : 0) + (rscaler ? rscaler[q] : 0);
  }
  for (n = 0; n < sites; ++n)
  {
    int site_small = 1;
    for (k = 0; k < K; ++k)
    {
      int rate_small = 1;
      double * p = parent + (size_t)n * span + k * s;
      for (i = 0; i < s; ++i)
      {
        double x = 0, y = 0;
        const double * L = lmat + ((size_t)k * s + i) * s;
        const double * R = rmat + ((size_t)k * s + i) * s;
        if (lclv) for (j = 0; j < s; ++j) x += L[j] * lclv[(size_t)n * span + k * s + j];
        else for (j = 0; j < s; ++j) if ((lmask[n] >> j) & 1ull) x += L[j];
        if (rclv) for (j = 0; j < s; ++j) y += R[j] * rclv[(size_t)n * span + k * s + j];
        else for (j = 0; j < s; ++j) if ((rmask[n] >> j) & 1ull) y += R[j];
        p[i] = x * y;
        rate_small &= (p[i] < ORC_SCALE_THRESHOLD);
      }
      if (per_rate && pscaler && !(lmask && rmask) && rate_small)
      {
        for (i = 0; i < s; ++i) p[i] *= ORC_SCALE_FACTOR;
        pscaler[(size_t)n * K + k] += 1;
      }
      site_small &= rate_small;
    }
    if (!per_rate && pscaler && !(lmask && rmask) && site_small)
    {
      for (i = 0; i < span; ++i) parent[(size_t)n * span + i] *= ORC_SCALE_FACTOR;
      pscaler[n] += 1;
    }
  }
}

/* site log-likelihood with scaling (core_likelihood.c:1388-1490) */
static double orc_site_finish(double terma, double terminv, unsigned int sc)
{
  if (sc)
  {
    if (terminv > 0)
    {
      unsigned int c = sc < ORC_MAXDIFF ? sc : ORC_MAXDIFF, q;
      double f = 1.0;
      for (q = 0; q < c; ++q) f *= ORC_SCALE_THRESHOLD;
      return log(terma * f + terminv);
    }
    return log(terma) + sc * log(ORC_SCALE_THRESHOLD);
  }
  return log(terma + terminv);
}

/* edge log-likelihood (likelihood.c:586-636 + core_likelihood.c:468-572,1192-1490): the matrix is
 * applied to the child end; a tip child is a mask row.  persite may be NULL. */
double orc_edge_loglh(unsigned int s, unsigned int sites, unsigned int K, const double * pclv,
                      const unsigned int * pscaler, const double * cclv, const unsigned long long * cmask,
                      const unsigned int * cscaler, const double * pmat, const double * freqs,
                      const double * weights, double pinv, const int * invariant,
                      const unsigned int * pattern_weights, int per_rate, double * persite)
{
  const unsigned int span = K * s;
  double logl = 0;
  unsigned int n, k, i, j;
  unsigned int rs[64];
  for (n = 0; n < sites; ++n)
  {
    double terma = 0, terminv = 0, lk;
    unsigned int site_sc = 0;
    if (per_rate)
    {
      site_sc = 0xffffffffu;
      for (k = 0; k < K; ++k)
      {
        rs[k] = (pscaler ? pscaler[(size_t)n * K + k] : 0) + (cscaler ? cscaler[(size_t)n * K + k] : 0);
        if (rs[k] < site_sc) site_sc = rs[k];
      }
      for (k = 0; k < K; ++k) { rs[k] -= site_sc; if (rs[k] > ORC_MAXDIFF) rs[k] = ORC_MAXDIFF; }
    }
    else
      site_sc = (pscaler ? pscaler[n] : 0) + (cscaler ? cscaler[n] : 0);
    for (k = 0; k < K; ++k)
    {
      double tr = 0;
      for (i = 0; i < s; ++i)
      {
        double tb = 0;
        const double * P = pmat + ((size_t)k * s + i) * s;
        if (cclv) for (j = 0; j < s; ++j) tb += P[j] * cclv[(size_t)n * span + k * s + j];
        else for (j = 0; j < s; ++j) if ((cmask[n] >> j) & 1ull) tb += P[j];
        tr += pclv[(size_t)n * span + k * s + i] * freqs[i] * tb;
      }
      if (per_rate) { unsigned int q; for (q = 0; q < rs[k]; ++q) tr *= ORC_SCALE_THRESHOLD; }
      if (pinv > 0)
      {
        terma += weights[k] * tr * (1.0 - pinv);
        if (invariant && invariant[n] != -1) terminv += weights[k] * freqs[invariant[n]] * pinv;
      }
      else
        terma += tr * weights[k];
    }
    lk = orc_site_finish(terma, terminv, site_sc) * pattern_weights[n];
    if (persite) persite[n] = lk;
    logl += lk;
  }
  return logl;
}

/* sumtable (core_derivatives.c:418-465, tip side :615-627): the tip, if any, takes the pi*Vinv role.
 * per-rate scalers (pscaler/cscaler non-NULL): entries of category k are multiplied by
 * 2^(-256*min(c_k - min_k c_k, 4)) (core_derivatives.c:418-440,460-463) */
void orc_sumtable(unsigned int s, unsigned int sites, unsigned int K, const double * pclv,
                  const unsigned long long * pmask, const double * cclv, const double * evecs,
                  const double * inv_evecs, const double * freqs, const unsigned int * pscaler,
                  const unsigned int * cscaler, int per_rate, double * sum)
{
  const unsigned int span = K * s;
  unsigned int n, k, i, j;
  unsigned int rs[64];
  for (n = 0; n < sites; ++n)
  {
    if (per_rate)
    {
      unsigned int mn = 0xffffffffu;
      for (k = 0; k < K; ++k)
      {
        rs[k] = (pscaler ? pscaler[(size_t)n * K + k] : 0) + (cscaler ? cscaler[(size_t)n * K + k] : 0);
        if (rs[k] < mn) mn = rs[k];
      }
      for (k = 0; k < K; ++k) { rs[k] -= mn; if (rs[k] > ORC_MAXDIFF) rs[k] = ORC_MAXDIFF; }
    }
    for (k = 0; k < K; ++k)
      for (j = 0; j < s; ++j)
      {
        double l = 0, r = 0;
        for (i = 0; i < s; ++i)
        {
          const double pi = pclv ? pclv[(size_t)n * span + k * s + i] : (double)((pmask[n] >> i) & 1ull);
          l += pi * freqs[i] * inv_evecs[i * s + j];
          r += evecs[j * s + i] * cclv[(size_t)n * span + k * s + i];
        }
        sum[(size_t)n * span + k * s + j] = l * r;
        if (per_rate) { unsigned int q; for (q = 0; q < rs[k]; ++q) sum[(size_t)n * span + k * s + j] *= ORC_SCALE_THRESHOLD; }
      }
  }
}

/* first/second derivative of -logL (core_derivatives.c:643-849) */
void orc_derivatives(unsigned int s, unsigned int sites, unsigned int K, const double * sum,
                     const double * rates, const double * weights, const double * evals,
                     const double * freqs, double pinv, const int * invariant,
                     const unsigned int * pattern_weights, double t, double * d_f, double * dd_f)
{
  const unsigned int span = K * s;
  double * diag = (double *)malloc(sizeof(double) * span * 3);
  unsigned int n, k, j;
  *d_f = *dd_f = 0;
  for (k = 0; k < K; ++k)
  {
    const double ki = rates[k] / (1.0 - pinv);
    for (j = 0; j < s; ++j)
    {
      const double e = exp(evals[j] * ki * t);
      diag[(k * s + j) * 3 + 0] = e;
      diag[(k * s + j) * 3 + 1] = evals[j] * ki * e;
      diag[(k * s + j) * 3 + 2] = evals[j] * ki * evals[j] * ki * e;
    }
  }
  for (n = 0; n < sites; ++n)
  {
    double L0 = 0, L1 = 0, L2 = 0, d1;
    for (k = 0; k < K; ++k)
    {
      double c0 = 0, c1 = 0, c2 = 0;
      for (j = 0; j < s; ++j)
      {
        const double v = sum[(size_t)n * span + k * s + j];
        c0 += v * diag[(k * s + j) * 3 + 0];
        c1 += v * diag[(k * s + j) * 3 + 1];
        c2 += v * diag[(k * s + j) * 3 + 2];
      }
      if (pinv > 0)
      {
        const double inv = (invariant && invariant[n] != -1) ? freqs[invariant[n]] * pinv : 0;
        c0 = c0 * (1. - pinv) + inv;
        c1 *= (1. - pinv);
        c2 *= (1. - pinv);
      }
      L0 += c0 * weights[k]; L1 += c1 * weights[k]; L2 += c2 * weights[k];
    }
    d1 = -L1 / L0;
    *d_f += pattern_weights[n] * d1;
    *dd_f += pattern_weights[n] * (d1 * d1 - L2 / L0);
  }
  free(diag);
}

/* ------------------------------------------------------------------------------------------------ */
/* Whole synthetic case: same generators as oracle/harness.c, post-order traversal from the root edge */
/* (treeinfo.c:966-1058: CLVs towards root = first inner node, logL on the edge root <-> root->back). */
/* ------------------------------------------------------------------------------------------------ */
typedef struct onode
{
  struct onode * next, * back;
  double length;
  int tip;   /* tip index or -1 */
  int inner; /* inner index (scaler / CLV slot) or -1 */
} onode_t;

static unsigned long long xs64(unsigned long long * s) { unsigned long long x = *s; x ^= x << 13; x ^= x >> 7; x ^= x << 17; return *s = x; }
static double u01(unsigned long long * s) { return (double)(xs64(s) >> 11) / 9007199254740992.0; }
static unsigned long long mix64(unsigned long long z)
{
  z += 0x9E3779B97F4A7C15ull; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
/* state bitmask of (partition, tip, site): 5 % gaps (all states), else one uniform state */
static unsigned long long cell_mask(unsigned long long seed, unsigned int part, unsigned int tip,
                                    unsigned long long site, unsigned int states)
{
  const unsigned long long h = mix64(seed ^ mix64(((unsigned long long)part << 48) ^ ((unsigned long long)tip << 32) ^ site));
  if ((unsigned int)(h % 10000u) < 500u) return states >= 64 ? ~0ull : ((1ull << states) - 1);
  return 1ull << ((h >> 20) % states);
}
static onode_t * otip(int idx) { onode_t * t = (onode_t *)calloc(1, sizeof(onode_t)); t->tip = idx; t->inner = -1; return t; }
static onode_t * oinner(int idx)
{
  onode_t * n[3]; int i;
  for (i = 0; i < 3; ++i) { n[i] = (onode_t *)calloc(1, sizeof(onode_t)); n[i]->tip = -1; n[i]->inner = idx; }
  n[0]->next = n[1]; n[1]->next = n[2]; n[2]->next = n[0];
  return n[0];
}
static void ojoin(onode_t * a, onode_t * b, double l) { a->back = b; b->back = a; a->length = b->length = l; }

typedef struct
{
  unsigned int tips, sites, s, K;
  int per_rate;
  double ** clv;             /* per inner node */
  unsigned int ** scaler;    /* per inner node */
  unsigned long long ** masks; /* per tip */
  double * rates, * evals, * evecs, * inv_evecs;
  const double * freqs;
  double pinv;
  double * pm1, * pm2;
} ocase_t;

/* orient the CLV of inner node n towards n->back (post-order recursion) */
static void orc_orient(ocase_t * c, onode_t * n)
{
  onode_t * l, * r;
  if (n->tip >= 0) return;
  l = n->next->back; r = n->next->next->back;
  orc_orient(c, l);
  orc_orient(c, r);
  orc_pmatrix(c->s, c->K, c->rates, c->pinv, c->evals, c->evecs, c->inv_evecs, l->length, c->pm1);
  orc_pmatrix(c->s, c->K, c->rates, c->pinv, c->evals, c->evecs, c->inv_evecs, r->length, c->pm2);
  orc_update_partial(c->s, c->sites, c->K, l->tip >= 0 ? NULL : c->clv[l->inner], l->tip >= 0 ? c->masks[l->tip] : NULL,
                     r->tip >= 0 ? NULL : c->clv[r->inner], r->tip >= 0 ? c->masks[r->tip] : NULL, c->pm1, c->pm2,
                     l->tip >= 0 ? NULL : c->scaler[l->inner], r->tip >= 0 ? NULL : c->scaler[r->inner],
                     c->clv[n->inner], c->scaler[n->inner], c->per_rate);
}

/* Runs one synthetic single-partition case.  Outputs: out[0] = logL at the root edge; for each of the
 * nbr branch lengths brl[i]: df[i], ddf[i] on the root edge; scaler_sums[tips-2] (sum of every inner
 * node's scaler array, nodes in harness order).  Returns 1 on success. */
int orc_run_case(unsigned int tips, unsigned int sites, unsigned int states, unsigned int K, int per_rate,
                 unsigned long long seed, double brlen_mean, double brlen_scale, double alpha, double pinv,
                 const double * subst, const double * freqs, unsigned int site_begin, const double * brl,
                 int nbr, double * out, double * df, double * ddf, unsigned long long * scaler_sums)
{
  ocase_t c;
  unsigned long long s = seed ? seed : 88172645463325252ull;
  unsigned int t, i, nb = 0;
  onode_t ** ends = (onode_t **)malloc(sizeof(onode_t *) * (2 * tips - 3));
  onode_t * c0 = oinner(0), * t0 = otip(0), * t1 = otip(1), * t2 = otip(2), * root, * other;
  const size_t span = (size_t)K * states;
  double * weights = (double *)malloc(sizeof(double) * K);
  unsigned int * pw = (unsigned int *)malloc(sizeof(unsigned int) * sites);
  int * invariant = NULL;
  memset(&c, 0, sizeof(c));
  ojoin(t0, c0, 0.1); ojoin(t1, c0->next, 0.1); ojoin(t2, c0->next->next, 0.1);
  ends[nb++] = t0; ends[nb++] = t1; ends[nb++] = t2;
  for (t = 3; t < tips; ++t)
  {
    onode_t * in = oinner((int)(t - 2)), * tip = otip((int)t), * e = ends[xs64(&s) % nb], * eb = e->back;
    ojoin(e, in, 0.1); ojoin(eb, in->next, 0.1); ojoin(tip, in->next->next, 0.1);
    ends[nb++] = in->next; ends[nb++] = tip;
  }
  for (i = 0; i < nb; ++i)
  {
    double l = -brlen_mean * log(1.0 - u01(&s));
    if (l < 1e-4) l = 1e-4;
    if (l > 1.0) l = 1.0;
    ends[i]->length = ends[i]->back->length = l * brlen_scale;
  }
  c.tips = tips; c.sites = sites; c.s = states; c.K = K; c.per_rate = per_rate; c.freqs = freqs; c.pinv = pinv;
  c.rates = (double *)malloc(sizeof(double) * K);
  orc_gamma_rates(alpha, K, 0, c.rates);
  c.evals = (double *)malloc(sizeof(double) * states);
  c.evecs = (double *)malloc(sizeof(double) * states * states);
  c.inv_evecs = (double *)malloc(sizeof(double) * states * states);
  orc_eigen(states, subst, freqs, c.evecs, c.inv_evecs, c.evals);
  c.pm1 = (double *)malloc(sizeof(double) * span * states);
  c.pm2 = (double *)malloc(sizeof(double) * span * states);
  c.clv = (double **)malloc(sizeof(double *) * (tips - 2));
  c.scaler = (unsigned int **)malloc(sizeof(unsigned int *) * (tips - 2));
  for (i = 0; i < tips - 2; ++i)
  {
    c.clv[i] = (double *)malloc(sizeof(double) * sites * span);
    c.scaler[i] = (unsigned int *)calloc((size_t)sites * (per_rate ? K : 1), sizeof(unsigned int));
  }
  c.masks = (unsigned long long **)malloc(sizeof(unsigned long long *) * tips);
  for (t = 0; t < tips; ++t)
  {
    c.masks[t] = (unsigned long long *)malloc(sizeof(unsigned long long) * sites);
    for (i = 0; i < sites; ++i) c.masks[t][i] = cell_mask(seed, 0, t, (unsigned long long)site_begin + i, states);
  }
  for (i = 0; i < K; ++i) weights[i] = 1.0 / K;
  for (i = 0; i < sites; ++i) pw[i] = 1;
  if (pinv > 0)
  {
    invariant = (int *)malloc(sizeof(int) * sites); /* models.c:651-750 */
    for (i = 0; i < sites; ++i)
    {
      unsigned long long m = states >= 64 ? ~0ull : ((1ull << states) - 1);
      for (t = 0; t < tips; ++t) m &= c.masks[t][i];
      invariant[i] = (m && !(m & (m - 1))) ? __builtin_ctzll(m) : -1;
    }
  }
  /* root = first inner node; the root edge is (root, root->back) and both ends are oriented */
  root = c0; other = c0->back;
  orc_orient(&c, other);
  orc_orient(&c, root);
  orc_pmatrix(states, K, c.rates, pinv, c.evals, c.evecs, c.inv_evecs, root->length, c.pm1);
  out[0] = orc_edge_loglh(states, sites, K, c.clv[root->inner], c.scaler[root->inner],
                          other->tip >= 0 ? NULL : c.clv[other->inner], other->tip >= 0 ? c.masks[other->tip] : NULL,
                          other->tip >= 0 ? NULL : c.scaler[other->inner], c.pm1, freqs, weights, pinv, invariant, pw,
                          per_rate, NULL);
  if (nbr > 0)
  {
    double * sum = (double *)malloc(sizeof(double) * sites * span);
    if (other->tip >= 0)
      orc_sumtable(states, sites, K, NULL, c.masks[other->tip], c.clv[root->inner], c.evecs, c.inv_evecs, freqs,
                   c.scaler[root->inner], NULL, per_rate, sum);
    else
      orc_sumtable(states, sites, K, c.clv[root->inner], NULL, c.clv[other->inner], c.evecs, c.inv_evecs, freqs,
                   c.scaler[root->inner], c.scaler[other->inner], per_rate, sum);
    for (i = 0; i < (unsigned int)nbr; ++i)
      orc_derivatives(states, sites, K, sum, c.rates, weights, c.evals, freqs, pinv, invariant, pw, brl[i], &df[i], &ddf[i]);
    free(sum);
  }
  if (scaler_sums)
    for (i = 0; i < tips - 2; ++i)
    {
      size_t q, nq = (size_t)sites * (per_rate ? K : 1);
      scaler_sums[i] = 0;
      for (q = 0; q < nq; ++q) scaler_sums[i] += c.scaler[i][q];
    }
  /* teardown (nodes leak-free enough for a test helper: free the big arrays) */
  for (i = 0; i < tips - 2; ++i) { free(c.clv[i]); free(c.scaler[i]); }
  for (t = 0; t < tips; ++t) free(c.masks[t]);
  free(c.clv); free(c.scaler); free(c.masks); free(c.rates); free(c.evals); free(c.evecs); free(c.inv_evecs);
  free(c.pm1); free(c.pm2); free(weights); free(pw); free(invariant); free(ends);
  return 1;
}

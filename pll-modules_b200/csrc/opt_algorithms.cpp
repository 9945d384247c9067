/*
 * opt_algorithms.cpp — the derivative-free / quasi-Newton scalar optimisers of pll-modules
 * (SURVEY.md §8a row a14).  Pure host code: every objective evaluation they request is a device
 * traversal + logL through the C ABI, so the drivers are organised as resumable state machines that
 * advance ALL partitions with one objective call per round (the reference's lock-step protocol,
 * which is also what keeps multi-GPU ranks in step).
 *
 *   pllmod_opt_minimize_brent / _brent_multi    src/optimize/opt_algorithms.c:789-1460
 *   pllmod_opt_minimize_em                      opt_algorithms.c:1473-1546
 *   pllmod_opt_minimize_lbfgsb / _lbfgsb_multi  opt_algorithms.c:418-776 (driver around the
 *                                               vendored L-BFGS-B 3.0, src/optimize/lbfgsb/)
 *
 * L-BFGS-B here is written from the published algorithm (Byrd, Lu, Nocedal, Zhu 1995; Morales &
 * Nocedal 2011 for the projected subspace step; More' & Thuente 1994 for the line search) with the
 * reference's constants: m = 5 corrections, ftol 1e-3, gtol 0.9, xtol 0.1, eps 2.2e-16, forward
 * differences with step 1e-4*|x|.  The dimension is small (<= 190), so the compact representation
 * B = theta*I - W M W' is kept as dense matrices and the subspace system is solved by Cholesky.
 */
#include "internal.h"

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <vector>

#define LBFGSB_ERROR 1.0e-4 /* lbfgsb.h:23 (finite-difference step) */
#define LBFGSB_EPS 2.2e-16  /* lbfgsb.h:168 */

static inline bool bad_score(double v) { return v != v || fabs(v - (-INFINITY)) < 1e-10; }

/* =============================================================================================== */
/* Brent                                                                                           */
/* =============================================================================================== */
namespace
{
const int BRENT_ITMAX = 100;
const double BRENT_GOLD = 0.3819660;
const double BRENT_ZEPS = 1.0e-7;

/* One bracketed 1-D minimisation in "ask/tell" form: trial() is the abscissa to evaluate next,
 * tell(f) consumes the value and returns false once the bracket is below tolerance. */
struct BrentState
{
  double a, b;       /* bracket */
  double x, w, v;    /* best, second best, previous second best */
  double fx, fw, fv;
  double d, e;       /* last step, step before last */
  double u;          /* pending trial */
  double tol;
  double x0, fx0;    /* starting point (returned if nothing better is found) */
  double * fopt;
  double * f2opt;
  bool active;

  static double with_sign(double mag, double s) { return s >= 0.0 ? fabs(mag) : -fabs(mag); }

  /* bracket test + next trial point; false = converged (writes fopt / f2opt) */
  bool propose()
  {
    const double xm = 0.5 * (a + b);
    const double tol1 = tol * fabs(x) + BRENT_ZEPS, tol2 = 2.0 * tol1;
    if (fabs(x - xm) <= tol2 - 0.5 * (b - a))
    {
      if (fopt) *fopt = fx;
      if (f2opt)
      {
        const double xw = x - w, wv = w - v, vx = v - x;
        *f2opt = 2.0 * (fv * xw + fx * wv + fw * vx) / (v * v * xw + x * x * wv + w * w * vx);
      }
      return false;
    }
    bool golden = true;
    if (fabs(e) > tol1)
    {
      /* parabola through (x,fx) (w,fw) (v,fv) */
      const double r = (x - w) * (fx - fv);
      double q = (x - v) * (fx - fw);
      double p = (x - v) * q - (x - w) * r;
      q = 2.0 * (q - r);
      if (q > 0.0) p = -p;
      q = fabs(q);
      const double e_prev = e;
      e = d;
      if (!(fabs(p) >= fabs(0.5 * q * e_prev) || p <= q * (a - x) || p >= q * (b - x)))
      {
        d = p / q;
        const double t = x + d;
        if (t - a < tol2 || b - t < tol2) d = with_sign(tol1, xm - x);
        golden = false;
      }
    }
    if (golden)
    {
      e = (x >= xm) ? a - x : b - x;
      d = BRENT_GOLD * e;
    }
    u = (fabs(d) >= tol1) ? x + d : x + with_sign(tol1, d);
    return true;
  }

  bool start(double ax, double bx, double cx, double fa, double fb, double fc, double tol_,
             double * fopt_, double * f2opt_)
  {
    tol = tol_;
    fopt = fopt_;
    f2opt = f2opt_;
    a = std::min(ax, cx);
    b = std::max(ax, cx);
    x0 = x = bx;
    fx0 = fx = fb;
    if (fa < fc)
    {
      w = ax, fw = fa, v = cx, fv = fc;
    }
    else
    {
      w = cx, fw = fc, v = ax, fv = fa;
    }
    d = e = 0.0;
    u = 0.0;
    active = propose();
    return active;
  }

  bool tell(double fu)
  {
    if (fu <= fx)
    {
      if (u >= x)
        a = x;
      else
        b = x;
      v = w, w = x, x = u;
      fv = fw, fw = fx, fx = fu;
    }
    else
    {
      if (u < x)
        a = u;
      else
        b = u;
      if (fu <= fw || w == x)
      {
        v = w, w = u;
        fv = fw, fw = fu;
      }
      else if (fu <= fv || v == x || v == w)
      {
        v = u;
        fv = fu;
      }
    }
    active = propose();
    return active;
  }
};

typedef double (*brent_multi_fn)(void *, double *, double *, int *);

int brent_drive(unsigned int xnum, int * opt_mask, double * xmin, double * xguess, double * xmax,
                double xtol, double * xopt, double * fx, double * f2x, void * params,
                brent_multi_fn target, int global_range)
{
  std::vector<BrentState> st(xnum);
  std::vector<double> lo(xnum), hi(xnum), ax(xnum, 0.0), cx(xnum, 0.0);
  std::vector<double> fa(xnum, 0.0), fb(xnum, 0.0), fc(xnum, 0.0), flo(xnum, 0.0), fhi(xnum, 0.0);
  std::vector<int> converged(xnum + 1, 0);
  for (unsigned int i = 0; i < xnum; ++i)
  {
    lo[i] = global_range ? xmin[0] : xmin[i];
    hi[i] = global_range ? xmax[0] : xmax[i];
    st[i] = BrentState();
    st[i].active = false;
  }

  /* first bracket: a narrow window around the guess, opt_algorithms.c:1110-1135 */
  for (unsigned int i = 0; i < xnum; ++i)
  {
    if (opt_mask && !opt_mask[i]) continue;
    xguess[i] = std::min(std::max(xguess[i], lo[i]), hi[i]);
    const double eps = xguess[i] > 0 ? xguess[i] * xtol * 50.0 : 2.0 * xtol;
    ax[i] = std::max(xguess[i] - eps, lo[i]);
    cx[i] = std::min(xguess[i] + eps, hi[i]);
  }
  target(params, ax.data(), fa.data(), nullptr);
  target(params, xguess, fb.data(), nullptr);
  target(params, cx.data(), fc.data(), nullptr);
  target(params, lo.data(), flo.data(), nullptr);
  target(params, hi.data(), fhi.data(), nullptr);

  for (unsigned int i = 0; i < xnum; ++i)
  {
    if (opt_mask && !opt_mask[i]) continue;
    if (fa[i] < fb[i] || fc[i] < fb[i])
    {
      /* the guess is not the lowest of the three: fall back to the full range */
      ax[i] = lo[i], cx[i] = hi[i];
      fa[i] = flo[i], fc[i] = fhi[i];
    }
    if (!st[i].start(ax[i], xguess[i], cx[i], fa[i], fb[i], fc[i], xtol, fx, f2x)) converged[i] = 1;
  }

  std::vector<double> u(xnum), fu(xnum);
  int rounds = 0;
  bool go = true;
  while (go)
  {
    for (unsigned int i = 0; i < xnum; ++i) u[i] = st[i].u;
    target(params, u.data(), fu.data(), converged.data());
    go = !converged[xnum]; /* the callee reduces the flags over ranks into the last slot */
    for (unsigned int i = 0; i < xnum; ++i)
    {
      if (opt_mask && !opt_mask[i]) continue;
      if (!converged[i]) converged[i] = !st[i].tell(fu[i]);
    }
    ++rounds;
    go = go && rounds <= BRENT_ITMAX;
  }

  for (unsigned int i = 0; i < xnum; ++i) xopt[i] = (st[i].fx > st[i].fx0) ? st[i].x0 : st[i].x;
  target(params, xopt, fx, nullptr);
  return PLL_SUCCESS;
}

struct BrentSingle
{
  double (*fn)(void *, double);
  void * params;
};
double brent_single_cb(void * p, double * x, double * fx, int *)
{
  BrentSingle * s = (BrentSingle *)p;
  *fx = s->fn(s->params, *x);
  return *fx;
}
}  // namespace

extern "C" double pllmod_opt_minimize_brent(double xmin, double xguess, double xmax, double xtol,
                                            double * fx, double * f2x, void * params,
                                            double (*target_funk)(void *, double))
{
  BrentSingle s{target_funk, params};
  double optx = xguess;
  brent_drive(1, nullptr, &xmin, &xguess, &xmax, xtol, &optx, fx, f2x, &s, brent_single_cb, 1);
  return optx;
}

extern "C" int pllmod_opt_minimize_brent_multi(unsigned int xnum, int * opt_mask, double * xmin,
                                               double * xguess, double * xmax, double xtol,
                                               double * xopt, double * fx, double * f2x,
                                               void * params,
                                               double (*target_funk)(void *, double *, double *, int *),
                                               int global_range)
{
  return brent_drive(xnum, opt_mask, xmin, xguess, xmax, xtol, xopt, fx, f2x, params, target_funk,
                     global_range);
}

/* =============================================================================================== */
/* EM for mixture weights (Wang, Li, Susko, Roger 2008)                                           */
/* =============================================================================================== */
extern "C" void pllmod_opt_minimize_em(double * w, unsigned int w_count, double * sitecat_lh,
                                       unsigned int * site_w, unsigned int l, void * params,
                                       double (*update_sitecatlk_funk)(void *, double *))
{
  /* opt_algorithms.c:1473-1546: at most 10 rounds; posterior-weighted category shares, divided by
   * the pattern count l; converged when every weight moves < 1e-4.  From the second round on the
   * freshly computed per-category likelihoods are rescaled by the previous weight ratios. */
  std::vector<double> acc(w_count), ratio(w_count, 1.0);
  bool have_ratio = false, done = false;
  for (int round = 0; round < 10 && !done; ++round)
  {
    update_sitecatlk_funk(params, sitecat_lh);
    if (have_ratio)
      for (unsigned int i = 0; i < l; ++i)
        for (unsigned int c = 0; c < w_count; ++c) sitecat_lh[(size_t)i * w_count + c] *= ratio[c];
    have_ratio = true;
    std::fill(acc.begin(), acc.end(), 0.0);
    for (unsigned int i = 0; i < l; ++i)
    {
      const double * row = sitecat_lh + (size_t)i * w_count;
      double tot = 0.0;
      for (unsigned int c = 0; c < w_count; ++c) tot += row[c];
      const double scale = site_w[i] / tot;
      for (unsigned int c = 0; c < w_count; ++c) acc[c] += row[c] * scale;
    }
    done = true;
    for (unsigned int c = 0; c < w_count; ++c)
    {
      acc[c] /= l;
      done = done && fabs(w[c] - acc[c]) < 1e-4;
      ratio[c] = acc[c] / w[c];
      w[c] = acc[c];
    }
  }
}

/* =============================================================================================== */
/* L-BFGS-B                                                                                        */
/* =============================================================================================== */
namespace
{
/* More'-Thuente safeguarded cubic/quadratic step (MINPACK-2 "dcstep") */
void mt_step(double & stx, double & fx, double & dx, double & sty, double & fy, double & dy,
             double & stp, double fp, double dp, bool & brackt, double stpmin, double stpmax)
{
  const double sgnd = dp * (dx / fabs(dx));
  double stpf;
  if (fp > fx)
  {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = std::max(fabs(theta), std::max(fabs(dx), fabs(dp)));
    double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp < stx) gamma = -gamma;
    const double p = (gamma - dx) + theta, q = ((gamma - dx) + gamma) + dp, r = p / q;
    const double stpc = stx + r * (stp - stx);
    const double stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
    stpf = (fabs(stpc - stx) < fabs(stpq - stx)) ? stpc : stpc + (stpq - stpc) / 2.0;
    brackt = true;
  }
  else if (sgnd < 0.0)
  {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = std::max(fabs(theta), std::max(fabs(dx), fabs(dp)));
    double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
    if (stp > stx) gamma = -gamma;
    const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dx, r = p / q;
    const double stpc = stp + r * (stx - stp);
    const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
    stpf = (fabs(stpc - stp) > fabs(stpq - stp)) ? stpc : stpq;
    brackt = true;
  }
  else if (fabs(dp) < fabs(dx))
  {
    const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
    const double s = std::max(fabs(theta), std::max(fabs(dx), fabs(dp)));
    double gamma = s * sqrt(std::max(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
    if (stp > stx) gamma = -gamma;
    const double p = (gamma - dp) + theta, q = (gamma + (dx - dp)) + gamma, r = p / q;
    double stpc;
    if (r < 0.0 && gamma != 0.0)
      stpc = stp + r * (stx - stp);
    else
      stpc = (stp > stx) ? stpmax : stpmin;
    const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
    if (brackt)
    {
      stpf = (fabs(stpc - stp) < fabs(stpq - stp)) ? stpc : stpq;
      stpf = (stp > stx) ? std::min(stp + 0.66 * (sty - stp), stpf)
                         : std::max(stp + 0.66 * (sty - stp), stpf);
    }
    else
    {
      stpf = (fabs(stpc - stp) > fabs(stpq - stp)) ? stpc : stpq;
      stpf = std::max(stpmin, std::min(stpmax, stpf));
    }
  }
  else
  {
    if (brackt)
    {
      const double theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
      const double s = std::max(fabs(theta), std::max(fabs(dy), fabs(dp)));
      double gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
      if (stp > sty) gamma = -gamma;
      const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dy, r = p / q;
      stpf = stp + r * (sty - stp);
    }
    else
      stpf = (stp > stx) ? stpmax : stpmin;
  }
  if (fp > fx)
  {
    sty = stp, fy = fp, dy = dp;
  }
  else
  {
    if (sgnd < 0.0)
    {
      sty = stx, fy = fx, dy = dx;
    }
    stx = stp, fx = fp, dx = dp;
  }
  stp = stpf;
}

/* More'-Thuente line search in ask/tell form (MINPACK-2 "dcsrch") */
struct LineSearch
{
  enum Status { EVALUATE, CONVERGED, WARNING, ERROR };
  double ftol, gtol, xtol, stpmin, stpmax;
  bool brackt;
  int stage;
  double finit, ginit, gtest, width, width1;
  double stx, fx, gx, sty, fy, gy, stmin, stmax;

  Status start(double f, double g, double & stp, double stpmin_, double stpmax_)
  {
    ftol = 1e-3, gtol = 0.9, xtol = 0.1;
    stpmin = stpmin_, stpmax = stpmax_;
    if (stp < stpmin || stp > stpmax || g >= 0.0 || stpmax < stpmin) return ERROR;
    brackt = false;
    stage = 1;
    finit = f, ginit = g, gtest = ftol * ginit;
    width = stpmax - stpmin, width1 = 2.0 * width;
    stx = 0.0, fx = finit, gx = ginit;
    sty = 0.0, fy = finit, gy = ginit;
    stmin = 0.0, stmax = stp + 4.0 * stp;
    return EVALUATE;
  }

  Status next(double f, double g, double & stp)
  {
    const double ftest = finit + stp * gtest;
    if (stage == 1 && f <= ftest && g >= 0.0) stage = 2;
    Status st = EVALUATE;
    if (brackt && (stp <= stmin || stp >= stmax)) st = WARNING;
    if (brackt && stmax - stmin <= xtol * stmax) st = WARNING;
    if (stp == stpmax && f <= ftest && g <= gtest) st = WARNING;
    if (stp == stpmin && (f > ftest || g >= gtest)) st = WARNING;
    if (f <= ftest && fabs(g) <= gtol * (-ginit)) st = CONVERGED;
    if (st != EVALUATE) return st;

    if (stage == 1 && f <= fx && f > ftest)
    {
      /* work on the modified function psi(stp) = f - f0 - ftol*g0*stp */
      double fm = f - stp * gtest, fxm = fx - stx * gtest, fym = fy - sty * gtest;
      double gm = g - gtest, gxm = gx - gtest, gym = gy - gtest;
      mt_step(stx, fxm, gxm, sty, fym, gym, stp, fm, gm, brackt, stmin, stmax);
      fx = fxm + stx * gtest, fy = fym + sty * gtest;
      gx = gxm + gtest, gy = gym + gtest;
    }
    else
      mt_step(stx, fx, gx, sty, fy, gy, stp, f, g, brackt, stmin, stmax);

    if (brackt)
    {
      if (fabs(sty - stx) >= 0.66 * width1) stp = stx + 0.5 * (sty - stx);
      width1 = width;
      width = fabs(sty - stx);
      stmin = std::min(stx, sty), stmax = std::max(stx, sty);
    }
    else
    {
      stmin = stp + 1.1 * (stp - stx);
      stmax = stp + 4.0 * (stp - stx);
    }
    stp = std::max(stp, stpmin);
    stp = std::min(stp, stpmax);
    if ((brackt && (stp <= stmin || stp >= stmax)) || (brackt && stmax - stmin <= xtol * stmax))
      stp = stx;
    return EVALUATE;
  }
};

/* dense helpers for the (2*col)x(2*col) middle matrix and the reduced Hessian */
bool invert_small(std::vector<double> & A, int n)
{
  /* Gauss-Jordan with partial pivoting, in place */
  std::vector<double> inv((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i) inv[(size_t)i * n + i] = 1.0;
  for (int c = 0; c < n; ++c)
  {
    int piv = c;
    for (int r = c + 1; r < n; ++r)
      if (fabs(A[(size_t)r * n + c]) > fabs(A[(size_t)piv * n + c])) piv = r;
    const double pv = A[(size_t)piv * n + c];
    if (!(fabs(pv) > 0.0) || !std::isfinite(pv)) return false;
    if (piv != c)
      for (int k = 0; k < n; ++k)
      {
        std::swap(A[(size_t)piv * n + k], A[(size_t)c * n + k]);
        std::swap(inv[(size_t)piv * n + k], inv[(size_t)c * n + k]);
      }
    const double ip = 1.0 / pv;
    for (int k = 0; k < n; ++k)
    {
      A[(size_t)c * n + k] *= ip;
      inv[(size_t)c * n + k] *= ip;
    }
    for (int r = 0; r < n; ++r)
    {
      if (r == c) continue;
      const double fct = A[(size_t)r * n + c];
      if (fct == 0.0) continue;
      for (int k = 0; k < n; ++k)
      {
        A[(size_t)r * n + k] -= fct * A[(size_t)c * n + k];
        inv[(size_t)r * n + k] -= fct * inv[(size_t)c * n + k];
      }
    }
  }
  A.swap(inv);
  return true;
}

bool cholesky_solve(std::vector<double> & A, int n, std::vector<double> & b)
{
  for (int j = 0; j < n; ++j)
  {
    double s = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) s -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(s > 0.0) || !std::isfinite(s)) return false;
    const double ljj = sqrt(s);
    A[(size_t)j * n + j] = ljj;
    for (int i = j + 1; i < n; ++i)
    {
      double t = A[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) t -= A[(size_t)i * n + k] * A[(size_t)j * n + k];
      A[(size_t)i * n + j] = t / ljj;
    }
  }
  for (int i = 0; i < n; ++i)
  {
    double t = b[i];
    for (int k = 0; k < i; ++k) t -= A[(size_t)i * n + k] * b[k];
    b[i] = t / A[(size_t)i * n + i];
  }
  for (int i = n - 1; i >= 0; --i)
  {
    double t = b[i];
    for (int k = i + 1; k < n; ++k) t -= A[(size_t)k * n + i] * b[k];
    b[i] = t / A[(size_t)i * n + i];
  }
  return true;
}

/* L-BFGS-B solver in reverse-communication form: advance() returns what the caller must do next */
struct Lbfgsb
{
  enum Request { NEED_FG, NEW_X, STOP };
  enum Phase { PH_INIT, PH_FIRST_FG, PH_LINESEARCH, PH_NEW_X, PH_DONE };

  int n, m;
  double * x;
  const double * l;
  const double * u;
  const int * nbd;
  double factr, pgtol;
  double f;
  std::vector<double> g;

  Phase phase;
  bool constrained, boxed;
  int iter, col;
  double theta, sbgnrm, fold, stp, stpmx, gd, gdold, dnorm, dtd;
  int nfun_ls, iback;
  bool failed; /* abnormal termination */
  std::vector<int> iwhere;
  std::vector<double> S, Y;           /* n x m, column j = correction pair j (oldest first) */
  std::vector<double> z, d, xold, gold;
  std::vector<double> Mmid;           /* (2col)^2: inverse of [[-D, L'],[L, theta S'S]] */
  LineSearch ls;

  void init(int n_, double * x_, const double * l_, const double * u_, const int * nbd_,
            double factr_, double pgtol_)
  {
    n = n_, m = 5;
    x = x_, l = l_, u = u_, nbd = nbd_;
    factr = factr_, pgtol = pgtol_;
    f = 0.0;
    g.assign(n, 0.0);
    phase = PH_INIT;
    failed = false;
  }

  void reset_memory()
  {
    col = 0;
    theta = 1.0;
    Mmid.clear();
  }

  double projected_gradient_norm() const
  {
    double nrm = 0.0;
    for (int i = 0; i < n; ++i)
    {
      double gi = g[i];
      if (nbd[i] != 0)
      {
        if (gi < 0.0)
        {
          if (nbd[i] >= 2) gi = std::max(x[i] - u[i], gi);
        }
        else if (nbd[i] <= 2)
          gi = std::min(x[i] - l[i], gi);
      }
      nrm = std::max(nrm, fabs(gi));
    }
    return nrm;
  }

  /* rows of W = [Y  theta*S] */
  inline double Wrow(int i, int j) const
  {
    return j < col ? Y[(size_t)i * m + j] : theta * S[(size_t)i * m + (j - col)];
  }
  void mid_times(const std::vector<double> & v, std::vector<double> & out) const
  {
    const int c2 = 2 * col;
    out.assign(c2, 0.0);
    for (int r = 0; r < c2; ++r)
    {
      double s = 0.0;
      for (int k = 0; k < c2; ++k) s += Mmid[(size_t)r * c2 + k] * v[k];
      out[r] = s;
    }
  }

  bool rebuild_middle()
  {
    const int c2 = 2 * col;
    Mmid.assign((size_t)c2 * c2, 0.0);
    for (int i = 0; i < col; ++i)
      for (int j = 0; j < col; ++j)
      {
        double sy = 0.0, ss = 0.0;
        for (int k = 0; k < n; ++k)
        {
          sy += S[(size_t)k * m + i] * Y[(size_t)k * m + j];
          ss += S[(size_t)k * m + i] * S[(size_t)k * m + j];
        }
        if (i == j) Mmid[(size_t)i * c2 + j] = -sy;
        if (i > j)
        {
          Mmid[(size_t)(col + i) * c2 + j] = sy; /* L */
          Mmid[(size_t)j * c2 + (col + i)] = sy; /* L' */
        }
        Mmid[(size_t)(col + i) * c2 + (col + j)] = theta * ss;
      }
    return invert_small(Mmid, c2);
  }

  /* generalised Cauchy point along the projected steepest-descent path; fills z and c = W'(z-x) */
  bool cauchy(std::vector<double> & c)
  {
    const int c2 = 2 * col;
    z.assign(x, x + n);
    c.assign(c2, 0.0);
    if (sbgnrm <= 0.0) return true;
    std::vector<double> dd(n, 0.0), tbreak;
    std::vector<int> ibreak;
    std::vector<double> p(c2, 0.0), v, wbp(c2);
    bool bounded = true;
    int nfree_dirs = 0;
    double f1 = 0.0;
    for (int i = 0; i < n; ++i)
    {
      const double neggi = -g[i];
      double tl = 0.0, tu = 0.0;
      if (iwhere[i] != 3 && iwhere[i] != -1)
      {
        if (nbd[i] <= 2) tl = x[i] - l[i];
        if (nbd[i] >= 2) tu = u[i] - x[i];
        const bool xlower = nbd[i] <= 2 && tl <= 0.0, xupper = nbd[i] >= 2 && tu <= 0.0;
        iwhere[i] = 0;
        if (xlower)
        {
          if (neggi <= 0.0) iwhere[i] = 1;
        }
        else if (xupper)
        {
          if (neggi >= 0.0) iwhere[i] = 2;
        }
        else if (fabs(neggi) <= 0.0)
          iwhere[i] = -3;
      }
      if (iwhere[i] != 0 && iwhere[i] != -1) continue;
      dd[i] = neggi;
      f1 -= neggi * neggi;
      for (int j = 0; j < c2; ++j) p[j] += Wrow(i, j) * neggi;
      if (nbd[i] <= 2 && nbd[i] != 0 && neggi < 0.0)
      {
        ibreak.push_back(i);
        tbreak.push_back(tl / (-neggi));
      }
      else if (nbd[i] >= 2 && neggi > 0.0)
      {
        ibreak.push_back(i);
        tbreak.push_back(tu / neggi);
      }
      else
      {
        ++nfree_dirs;
        if (fabs(neggi) > 0.0) bounded = false;
      }
    }
    const int nbreak = (int)ibreak.size();
    if (nbreak == 0 && nfree_dirs == 0) return true;

    double f2 = -theta * f1;
    const double f2_org = f2;
    if (col > 0)
    {
      mid_times(p, v);
      for (int j = 0; j < c2; ++j) f2 -= v[j] * p[j];
    }
    double dtm = -f1 / f2, tsum = 0.0;
    bool skip_tail = false;
    if (nbreak > 0)
    {
      std::vector<int> order(nbreak);
      for (int k = 0; k < nbreak; ++k) order[k] = k;
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return tbreak[a] < tbreak[b]; });
      double tj = 0.0;
      int nleft = nbreak;
      for (int k = 0; k < nbreak; ++k)
      {
        const double tj0 = tj;
        const int ib = ibreak[order[k]];
        tj = tbreak[order[k]];
        const double dt = tj - tj0;
        if (dtm < dt) break;
        tsum += dt;
        --nleft;
        const double dibp = dd[ib];
        dd[ib] = 0.0;
        double zibp;
        if (dibp > 0.0)
        {
          zibp = u[ib] - x[ib];
          z[ib] = u[ib];
          iwhere[ib] = 2;
        }
        else
        {
          zibp = l[ib] - x[ib];
          z[ib] = l[ib];
          iwhere[ib] = 1;
        }
        if (nleft == 0 && nbreak == n)
        {
          dtm = dt;
          skip_tail = true;
          break;
        }
        const double dibp2 = dibp * dibp;
        f1 = f1 + dt * f2 + dibp2 - theta * dibp * zibp;
        f2 = f2 - theta * dibp2;
        if (col > 0)
        {
          for (int j = 0; j < c2; ++j) c[j] += dt * p[j];
          for (int j = 0; j < c2; ++j) wbp[j] = Wrow(ib, j);
          mid_times(wbp, v);
          double wmc = 0.0, wmp = 0.0, wmw = 0.0;
          for (int j = 0; j < c2; ++j)
          {
            wmc += c[j] * v[j];
            wmp += p[j] * v[j];
            wmw += wbp[j] * v[j];
          }
          for (int j = 0; j < c2; ++j) p[j] -= dibp * wbp[j];
          f1 += dibp * wmc;
          f2 += 2.0 * dibp * wmp - dibp2 * wmw;
        }
        f2 = std::max(LBFGSB_EPS * f2_org, f2);
        if (nleft > 0)
          dtm = -f1 / f2;
        else if (bounded)
        {
          f1 = f2 = 0.0;
          dtm = 0.0;
        }
        else
          dtm = -f1 / f2;
      }
    }
    if (!skip_tail)
    {
      if (dtm <= 0.0) dtm = 0.0;
      tsum += dtm;
      for (int i = 0; i < n; ++i) z[i] += tsum * dd[i];
    }
    if (col > 0)
      for (int j = 0; j < c2; ++j) c[j] += dtm * p[j];
    for (int i = 0; i < n; ++i)
      if (!std::isfinite(z[i])) return false;
    return true;
  }

  /* minimise the quadratic model over the variables free at the Cauchy point (direct primal
   * method), then project onto the box (Morales-Nocedal) */
  bool subspace_min(const std::vector<double> & c, const std::vector<int> & free_idx)
  {
    const int nf = (int)free_idx.size(), c2 = 2 * col;
    std::vector<double> r(nf), mc;
    mid_times(c, mc);
    for (int i = 0; i < nf; ++i)
    {
      const int k = free_idx[i];
      double wmc = 0.0;
      for (int j = 0; j < c2; ++j) wmc += Wrow(k, j) * mc[j];
      r[i] = constrained ? (-theta * (z[k] - x[k]) - g[k] + wmc) : (-g[k] + wmc);
    }
    /* B_FF = theta I - W_F M W_F' */
    std::vector<double> B((size_t)nf * nf), WM((size_t)nf * c2);
    for (int i = 0; i < nf; ++i)
      for (int j = 0; j < c2; ++j)
      {
        double s = 0.0;
        for (int k = 0; k < c2; ++k) s += Wrow(free_idx[i], k) * Mmid[(size_t)k * c2 + j];
        WM[(size_t)i * c2 + j] = s;
      }
    for (int i = 0; i < nf; ++i)
      for (int j = 0; j <= i; ++j)
      {
        double s = 0.0;
        for (int k = 0; k < c2; ++k) s += WM[(size_t)i * c2 + k] * Wrow(free_idx[j], k);
        const double val = (i == j ? theta : 0.0) - s;
        B[(size_t)i * nf + j] = B[(size_t)j * nf + i] = val;
      }
    if (!cholesky_solve(B, nf, r)) return false;

    std::vector<double> zc(z);
    bool hit = false;
    for (int i = 0; i < nf; ++i)
    {
      const int k = free_idx[i];
      const double t = zc[k] + r[i];
      switch (nbd[k])
      {
        case 0: z[k] = t; break;
        case 1:
          z[k] = std::max(l[k], t);
          if (z[k] == l[k]) hit = true;
          break;
        case 2:
          z[k] = std::min(u[k], std::max(l[k], t));
          if (z[k] == l[k] || z[k] == u[k]) hit = true;
          break;
        default:
          z[k] = std::min(u[k], t);
          if (z[k] == u[k]) hit = true;
          break;
      }
    }
    if (!hit) return true;
    double ddp = 0.0;
    for (int i = 0; i < n; ++i) ddp += (z[i] - x[i]) * g[i];
    if (ddp <= 0.0) return true;
    /* the projected point is not a descent direction: truncate the step at the first bound */
    z = zc;
    double alpha = 1.0;
    int ibd = -1;
    for (int i = 0; i < nf; ++i)
    {
      const int k = free_idx[i];
      const double dk = r[i];
      double t1 = alpha;
      if (nbd[k] != 0)
      {
        if (dk < 0.0 && nbd[k] <= 2)
        {
          const double t2 = l[k] - z[k];
          if (t2 >= 0.0)
            t1 = 0.0;
          else if (dk * alpha < t2)
            t1 = t2 / dk;
        }
        else if (dk > 0.0 && nbd[k] >= 2)
        {
          const double t2 = u[k] - z[k];
          if (t2 <= 0.0)
            t1 = 0.0;
          else if (dk * alpha > t2)
            t1 = t2 / dk;
        }
        if (t1 < alpha)
        {
          alpha = t1;
          ibd = i;
        }
      }
    }
    if (alpha < 1.0 && ibd >= 0)
    {
      const int k = free_idx[ibd];
      if (r[ibd] > 0.0)
      {
        z[k] = u[k];
        r[ibd] = 0.0;
      }
      else if (r[ibd] < 0.0)
      {
        z[k] = l[k];
        r[ibd] = 0.0;
      }
    }
    for (int i = 0; i < nf; ++i) z[free_idx[i]] += alpha * r[i];
    return true;
  }

  /* search direction for the current iterate; false = numerical trouble (caller resets memory) */
  bool compute_direction()
  {
    std::vector<double> c;
    std::vector<int> free_idx;
    if (!constrained && col > 0)
    {
      z.assign(x, x + n);
      c.assign(2 * col, 0.0);
      for (int i = 0; i < n; ++i) free_idx.push_back(i);
    }
    else
    {
      if (!cauchy(c)) return false;
      for (int i = 0; i < n; ++i)
        if (iwhere[i] <= 0) free_idx.push_back(i);
    }
    if (!free_idx.empty() && col > 0)
      if (!subspace_min(c, free_idx)) return false;
    d.resize(n);
    for (int i = 0; i < n; ++i) d[i] = z[i] - x[i];
    return true;
  }

  /* set up the line search along d; returns false when d is not a descent direction */
  bool begin_line_search()
  {
    dtd = 0.0;
    for (int i = 0; i < n; ++i) dtd += d[i] * d[i];
    dnorm = sqrt(dtd);
    stpmx = 1e10;
    if (constrained)
    {
      if (iter == 0)
        stpmx = 1.0;
      else
        for (int i = 0; i < n; ++i)
        {
          const double a1 = d[i];
          if (nbd[i] == 0) continue;
          if (a1 < 0.0 && nbd[i] <= 2)
          {
            const double a2 = l[i] - x[i];
            if (a2 >= 0.0)
              stpmx = 0.0;
            else if (a1 * stpmx < a2)
              stpmx = a2 / a1;
          }
          else if (a1 > 0.0 && nbd[i] >= 2)
          {
            const double a2 = u[i] - x[i];
            if (a2 <= 0.0)
              stpmx = 0.0;
            else if (a1 * stpmx > a2)
              stpmx = a2 / a1;
          }
        }
    }
    stp = (iter == 0 && !boxed) ? std::min(1.0 / dnorm, stpmx) : 1.0;
    xold.assign(x, x + n);
    gold = g;
    fold = f;
    nfun_ls = 0;
    iback = 0;
    gd = 0.0;
    for (int i = 0; i < n; ++i) gd += g[i] * d[i];
    gdold = gd;
    if (gd >= 0.0) return false;
    return ls.start(f, gd, stp, 0.0, stpmx) == LineSearch::EVALUATE;
  }

  void take_trial_step()
  {
    ++nfun_ls;
    iback = nfun_ls - 1;
    if (stp == 1.0)
      for (int i = 0; i < n; ++i) x[i] = z[i];
    else
      for (int i = 0; i < n; ++i) x[i] = stp * d[i] + xold[i];
  }

  void restore_iterate()
  {
    for (int i = 0; i < n; ++i) x[i] = xold[i];
    g = gold;
    f = fold;
  }

  /* one (re)start of an iteration: direction + line-search set-up, resetting the memory on trouble */
  Request start_iteration()
  {
    for (int attempt = 0; attempt < 3; ++attempt)
    {
      if (compute_direction() && begin_line_search())
      {
        take_trial_step();
        phase = PH_LINESEARCH;
        return NEED_FG;
      }
      if (col == 0) break;
      reset_memory();
    }
    failed = true;
    phase = PH_DONE;
    return STOP;
  }

  bool store_pair()
  {
    /* r = g - g_old, s = stp*d */
    std::vector<double> yv(n), sv(n);
    double rr = 0.0;
    for (int i = 0; i < n; ++i)
    {
      yv[i] = g[i] - gold[i];
      rr += yv[i] * yv[i];
    }
    double dr, ddum;
    if (stp == 1.0)
    {
      dr = gd - gdold;
      ddum = -gdold;
      sv = d;
    }
    else
    {
      dr = (gd - gdold) * stp;
      ddum = -gdold * stp;
      for (int i = 0; i < n; ++i) sv[i] = stp * d[i];
    }
    if (dr <= LBFGSB_EPS * ddum) return true; /* curvature too small: skip the update */
    if (col == m)
    {
      for (int i = 0; i < n; ++i)
        for (int j = 0; j + 1 < m; ++j)
        {
          S[(size_t)i * m + j] = S[(size_t)i * m + j + 1];
          Y[(size_t)i * m + j] = Y[(size_t)i * m + j + 1];
        }
      --col;
    }
    for (int i = 0; i < n; ++i)
    {
      S[(size_t)i * m + col] = sv[i];
      Y[(size_t)i * m + col] = yv[i];
    }
    ++col;
    theta = rr / dr;
    return rebuild_middle();
  }

  Request advance()
  {
    switch (phase)
    {
      case PH_INIT:
      {
        if (n <= 0 || factr < 0.0)
        {
          failed = true;
          phase = PH_DONE;
          return STOP;
        }
        constrained = false;
        boxed = true;
        iwhere.assign(n, 0);
        for (int i = 0; i < n; ++i)
        {
          if (nbd[i] == 2 && l[i] > u[i])
          {
            failed = true;
            phase = PH_DONE;
            return STOP;
          }
          if (nbd[i] > 0)
          {
            if (nbd[i] <= 2 && x[i] <= l[i])
              x[i] = l[i];
            else if (nbd[i] >= 2 && x[i] >= u[i])
              x[i] = u[i];
          }
          if (nbd[i] != 2) boxed = false;
          if (nbd[i] == 0)
            iwhere[i] = -1;
          else
          {
            constrained = true;
            iwhere[i] = (nbd[i] == 2 && u[i] - l[i] <= 0.0) ? 3 : 0;
          }
        }
        S.assign((size_t)n * m, 0.0);
        Y.assign((size_t)n * m, 0.0);
        reset_memory();
        iter = 0;
        phase = PH_FIRST_FG;
        return NEED_FG;
      }
      case PH_FIRST_FG:
        sbgnrm = projected_gradient_norm();
        if (sbgnrm <= pgtol)
        {
          phase = PH_DONE;
          return STOP;
        }
        return start_iteration();
      case PH_LINESEARCH:
      {
        gd = 0.0;
        for (int i = 0; i < n; ++i) gd += g[i] * d[i];
        const LineSearch::Status st = ls.next(f, gd, stp);
        if (st == LineSearch::EVALUATE)
        {
          if (iback >= 19)
          {
            /* 20 backtracks: give up on this direction */
            restore_iterate();
            if (col == 0)
            {
              failed = true;
              phase = PH_DONE;
              return STOP;
            }
            reset_memory();
            return start_iteration();
          }
          take_trial_step();
          return NEED_FG;
        }
        ++iter;
        sbgnrm = projected_gradient_norm();
        phase = PH_NEW_X;
        return NEW_X;
      }
      case PH_NEW_X:
      {
        if (sbgnrm <= pgtol)
        {
          phase = PH_DONE;
          return STOP;
        }
        const double scale = std::max(std::max(fabs(fold), fabs(f)), 1.0);
        if (fold - f <= factr * LBFGSB_EPS * scale)
        {
          phase = PH_DONE;
          return STOP;
        }
        if (!store_pair()) reset_memory();
        return start_iteration();
      }
      default: return STOP;
    }
  }
};

/* forward-difference step for coordinate value v (opt_algorithms.c:503-509) */
inline double fd_step(double v)
{
  const double h = LBFGSB_ERROR * fabs(v);
  return h < 1e-12 ? LBFGSB_ERROR : h;
}
}  // namespace

extern "C" double pllmod_opt_minimize_lbfgsb(double * x, double * xmin, double * xmax, int * bound,
                                             unsigned int n, double factr, double pgtol,
                                             void * params, double (*target_funk)(void *, double *))
{
  pll_errno = 0;
  Lbfgsb opt;
  opt.init((int)n, x, xmin, xmax, bound, factr, pgtol);
  double score = 0.0;
  for (;;)
  {
    const Lbfgsb::Request rq = opt.advance();
    if (rq == Lbfgsb::STOP) break;
    if (rq == Lbfgsb::NEW_X) continue;
    score = target_funk(params, x);
    if (bad_score(score)) break;
    opt.f = score;
    for (unsigned int i = 0; i < n; ++i)
    {
      const double keep = x[i];
      x[i] = keep + fd_step(keep);
      const double h = x[i] - keep;
      const double shifted = target_funk(params, x);
      opt.g[i] = (shifted - score) / h;
      x[i] = keep;
    }
  }
  /* leave the model at the optimum */
  score = target_funk(params, x);
  if (score != score)
  {
    score = -INFINITY;
    if (!pll_errno) pllmod_set_error(PLLMOD_OPT_ERROR_LBFGSB_UNKNOWN, "Unknown LBFGSB error");
  }
  return score;
}

extern "C" double pllmod_opt_minimize_lbfgsb_multi(
    unsigned int xnum, double ** x, double ** xmin, double ** xmax, int ** bound, unsigned int * n,
    unsigned int nmax, double factr, double pgtol, void * params,
    double (*target_funk)(void *, double **, double *, int *))
{
  /* opt_algorithms.c:628-776: one solver per LOCAL partition (x[p] == NULL marks a remote one);
   * all solvers advance in lock step; the objective callback doubles as the cross-rank AND
   * reduction of the flag arrays (x == NULL, result in flags[xnum]). */
  double score = -INFINITY;
  std::vector<double> lh_old(xnum, 0.0), lh_new(xnum, 0.0), keep(xnum, 0.0), h(xnum, 0.0);
  std::vector<int> converged(xnum + 1, 0), skip(xnum + 1, 0);
  std::vector<Lbfgsb> opts(xnum);
  for (unsigned int p = 0; p < xnum; ++p)
  {
    if (!x[p])
    {
      converged[p] = skip[p] = 1;
      continue;
    }
    opts[p].init((int)n[p], x[p], xmin[p], xmax[p], bound[p], factr, pgtol);
  }
  pll_errno = 0;
  bool go = true;
  while (go)
  {
    for (unsigned int p = 0; p < xnum; ++p)
    {
      if (converged[p])
      {
        skip[p] = 1;
        continue;
      }
      const Lbfgsb::Request rq = opts[p].advance();
      skip[p] = rq != Lbfgsb::NEED_FG;
      converged[p] = rq == Lbfgsb::STOP;
    }
    target_funk(params, nullptr, nullptr, converged.data());
    go = !converged[xnum];
    target_funk(params, nullptr, nullptr, skip.data());
    const bool all_skip = skip[xnum] != 0;
    if (all_skip || !go) continue;

    score = target_funk(params, x, lh_old.data(), skip.data());
    if (bad_score(score)) break;
    for (unsigned int p = 0; p < xnum; ++p)
      if (!skip[p]) opts[p].f = lh_old[p];
    for (unsigned int i = 0; i < nmax; ++i)
    {
      for (unsigned int p = 0; p < xnum; ++p)
      {
        if (!skip[p] && i >= n[p]) skip[p] = 1; /* fewer parameters than the widest partition */
        if (skip[p]) continue;
        keep[p] = x[p][i];
        x[p][i] = keep[p] + fd_step(keep[p]);
        h[p] = x[p][i] - keep[p];
      }
      score = target_funk(params, x, lh_new.data(), skip.data());
      for (unsigned int p = 0; p < xnum; ++p)
      {
        if (skip[p]) continue;
        opts[p].g[i] = (lh_new[p] - lh_old[p]) / h[p];
        x[p][i] = keep[p];
      }
    }
  }
  score = target_funk(params, x, nullptr, nullptr);
  if (score != score)
  {
    score = -INFINITY;
    if (!pll_errno) pllmod_set_error(PLLMOD_OPT_ERROR_LBFGSB_UNKNOWN, "Unknown LBFGSB error");
  }
  return score;
}

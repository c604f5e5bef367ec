// Host-side driver of the length-scale optimisation: L-BFGS-B specialised to ONE bounded variable.
//
// Reference call site: sklearn GaussianProcessRegressor._constrained_optimization (_gpr.py:656-676),
// i.e. scipy.optimize.minimize(method="L-BFGS-B", jac=True, bounds=log([1e-5, 1e5])) on
// theta = log(length_scale) -- reached from BOOptimizer._maximize (optimizer.py:21) once per restart.
// scipy (pinned 1.1.0, installed 1.18.1) is not vendored in the reference; this restates the
// published algorithm (Byrd, Lu, Nocedal, Zhu 1995; L-BFGS-B 3.0 with the More'-Thuente line search
// dcsrch/dcstep of MINPACK-2) for n = 1, where it collapses exactly:
//   * the limited-memory BFGS matrix of any number of pairs is the scalar B = y/s of the newest pair
//     (B = 1 while no pair is stored);
//   * the generalised Cauchy point is clip(x - g/B) and subspace minimisation leaves it unchanged;
//   * everything else -- first step length, stpmx, the Wolfe tests (ftol 1e-3, gtol 0.9, xtol 0.1),
//     20-evaluation line-search cap with memory reset, curvature-skip rule, and the two convergence
//     tests (projected gradient <= 1e-5, relative reduction <= 1e7 * eps) -- is followed step by step
//     so that the same local optimum is reached from the same start.
// It is written as a resumable state machine so that all restarts advance in lock step and every
// round of objective evaluations is ONE batched device launch sequence.
#pragma once
#include <cmath>
#include <algorithm>

#ifdef __CUDACC__
#define PLX_HD __host__ __device__
#else
#define PLX_HD
#endif

namespace plx {

PLX_HD inline double hd_abs(double a) { return a < 0.0 ? -a : a; }
PLX_HD inline double hd_max(double a, double b) { return a < b ? b : a; }   // exactly std::max / std::min,
PLX_HD inline double hd_min(double a, double b) { return b < a ? b : a; }   // including their NaN behaviour

struct Dcsrch {
  // persistent state of one line search
  bool brackt; int stage;
  double ginit, gtest, gx, gy, finit, fx, fy, stx, sty, stmin, stmax, width, width1;
  enum Task { START, FG, CONVERGENCE, WARNING, ERROR } task;

  PLX_HD static void dcstep(double& stx, double& fx, double& dx, double& sty, double& fy, double& dy, double& stp,
                     double fp, double dp, bool& brackt, double stpmin, double stpmax) {
    const double sgnd = dp * (dx / hd_abs(dx));
    double stpf;
    if (fp > fx) {                                   // case 1: higher value -> minimum bracketed
      const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
      const double s = hd_max(hd_abs(theta), hd_max(hd_abs(dx), hd_abs(dp)));
      double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
      if (stp < stx) gamma = -gamma;
      const double p = (gamma - dx) + theta, q = ((gamma - dx) + gamma) + dp, r = p / q;
      const double stpc = stx + r * (stp - stx);
      const double stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
      stpf = (hd_abs(stpc - stx) < hd_abs(stpq - stx)) ? stpc : stpc + (stpq - stpc) / 2.0;
      brackt = true;
    } else if (sgnd < 0.0) {                         // case 2: derivatives of opposite sign
      const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
      const double s = hd_max(hd_abs(theta), hd_max(hd_abs(dx), hd_abs(dp)));
      double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
      if (stp > stx) gamma = -gamma;
      const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dx, r = p / q;
      const double stpc = stp + r * (stx - stp);
      const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
      stpf = (hd_abs(stpc - stp) > hd_abs(stpq - stp)) ? stpc : stpq;
      brackt = true;
    } else if (hd_abs(dp) < hd_abs(dx)) {      // case 3: derivative magnitude decreases
      const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
      const double s = hd_max(hd_abs(theta), hd_max(hd_abs(dx), hd_abs(dp)));
      double gamma = s * sqrt(hd_max(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
      if (stp > stx) gamma = -gamma;
      const double p = (gamma - dp) + theta, q = (gamma + (dx - dp)) + gamma, r = p / q;
      double stpc;
      if (r < 0.0 && gamma != 0.0) stpc = stp + r * (stx - stp);
      else stpc = (stp > stx) ? stpmax : stpmin;
      const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
      if (brackt) {
        stpf = (hd_abs(stpc - stp) < hd_abs(stpq - stp)) ? stpc : stpq;
        if (stp > stx) stpf = hd_min(stp + 0.66 * (sty - stp), stpf);
        else stpf = hd_max(stp + 0.66 * (sty - stp), stpf);
      } else {
        stpf = (hd_abs(stpc - stp) > hd_abs(stpq - stp)) ? stpc : stpq;
        stpf = hd_min(stpmax, stpf);
        stpf = hd_max(stpmin, stpf);
      }
    } else {                                          // case 4: derivative does not decrease
      if (brackt) {
        const double theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
        const double s = hd_max(hd_abs(theta), hd_max(hd_abs(dy), hd_abs(dp)));
        double gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
        if (stp > sty) gamma = -gamma;
        const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dy, r = p / q;
        stpf = stp + r * (sty - stp);
      } else {
        stpf = (stp > stx) ? stpmax : stpmin;
      }
    }
    if (fp > fx) { sty = stp; fy = fp; dy = dp; }
    else {
      if (sgnd < 0.0) { sty = stx; fy = fx; dy = dx; }
      stx = stp; fx = fp; dx = dp;
    }
    stp = stpf;
  }

  // one call of dcsrch; `first` distinguishes the START entry
  PLX_HD void step(bool first, double f, double g, double& stp, double ftol, double gtol, double xtol,
            double stpmin, double stpmax) {
    const double xtrapl = 1.1, xtrapu = 4.0;
    if (first) {
      if (stp < stpmin || stp > stpmax || g >= 0.0 || stpmax < stpmin) { task = ERROR; return; }
      brackt = false; stage = 1; finit = f; ginit = g; gtest = ftol * ginit;
      width = stpmax - stpmin; width1 = width / 0.5;
      stx = 0.0; fx = finit; gx = ginit; sty = 0.0; fy = finit; gy = ginit;
      stmin = 0.0; stmax = stp + xtrapu * stp;
      task = FG; return;
    }
    const double ftest = finit + stp * gtest;
    if (stage == 1 && f <= ftest && g >= 0.0) stage = 2;
    task = FG;
    if (brackt && (stp <= stmin || stp >= stmax)) task = WARNING;
    if (brackt && stmax - stmin <= xtol * stmax) task = WARNING;
    if (stp == stpmax && f <= ftest && g <= gtest) task = WARNING;
    if (stp == stpmin && (f > ftest || g >= gtest)) task = WARNING;
    if (f <= ftest && hd_abs(g) <= gtol * (-ginit)) task = CONVERGENCE;
    if (task == WARNING || task == CONVERGENCE) return;
    if (stage == 1 && f <= fx && f > ftest) {
      const double fm = f - stp * gtest; double fxm = fx - stx * gtest, fym = fy - sty * gtest;
      const double gm = g - gtest; double gxm = gx - gtest, gym = gy - gtest;
      dcstep(stx, fxm, gxm, sty, fym, gym, stp, fm, gm, brackt, stmin, stmax);
      fx = fxm + stx * gtest; fy = fym + sty * gtest; gx = gxm + gtest; gy = gym + gtest;
    } else {
      dcstep(stx, fx, gx, sty, fy, gy, stp, f, g, brackt, stmin, stmax);
    }
    if (brackt) {
      if (hd_abs(sty - stx) >= 0.66 * width1) stp = stx + 0.5 * (sty - stx);
      width1 = width; width = hd_abs(sty - stx);
    }
    if (brackt) { stmin = hd_min(stx, sty); stmax = hd_max(stx, sty); }
    else { stmin = stp + xtrapl * (stp - stx); stmax = stp + xtrapu * (stp - stx); }
    stp = hd_max(stp, stpmin);
    stp = hd_min(stp, stpmax);
    if ((brackt && (stp <= stmin || stp >= stmax)) || (brackt && stmax - stmin <= xtol * stmax)) stp = stx;
    task = FG;
  }
};

struct Lbfgsb1D {
  // configuration (scipy defaults used by sklearn)
  double lo = -11.512925464970229, hi = 11.512925464970229;
  double factr = 1e7, pgtol = 1e-5;
  int maxiter = 15000, maxfun = 15000, maxls = 20;

  // iterate
  double x = 0, f = 0, g = 0;
  double B = 1.0; bool have_pair = false;
  int iter = 0, nfev = 0;
  // line search
  Dcsrch ls; double d = 0, z = 0, stp = 0, stpmx = 0, gd = 0, gdold = 0, xt = 0, ft = 0, gt = 0;
  int ifun = 0, iback = 0;
  // driver
  bool done = false, waiting_first = true;
  double x_eval = 0;
  int status = 0;            // 0 converged, 1 iteration/evaluation limit, 2 abnormal line search
  const char* message = "";

  static constexpr double epsmch = 2.220446049250313e-16;

  PLX_HD double projgr() const {
    double gi = g;
    if (gi < 0.0) gi = hd_max(x - hi, gi); else gi = hd_min(x - lo, gi);
    return hd_abs(gi);
  }

  PLX_HD void start(double x0) {
    x = hd_min(hd_max(x0, lo), hi);
    x_eval = x; waiting_first = true; done = false; iter = 0; nfev = 0; have_pair = false; B = 1.0;
  }

  PLX_HD void finish(int st, const char* msg) { done = true; status = st; message = msg; }

  PLX_HD void begin_iteration() {
    for (;;) {
      // generalised Cauchy point of the one-variable quadratic model
      const double neggi = -g;
      double tb = HUGE_VAL;
      if (neggi < 0.0) tb = (x - lo) / (-neggi); else if (neggi > 0.0) tb = (hi - x) / neggi;
      const double dtm = 1.0 / (have_pair ? B : 1.0);
      if (dtm < tb) z = x + dtm * neggi; else z = (neggi < 0.0) ? lo : hi;
      d = z - x;
      // line-search set-up (lnsrlb)
      stpmx = 1e10;
      if (iter == 0) stpmx = 1.0;
      else if (d < 0.0) { const double a2 = lo - x; if (a2 >= 0.0) stpmx = 0.0; else if (d * stpmx < a2) stpmx = a2 / d; }
      else if (d > 0.0) { const double a2 = hi - x; if (a2 <= 0.0) stpmx = 0.0; else if (d * stpmx > a2) stpmx = a2 / d; }
      stp = 1.0;
      xt = x; ft = f; gt = g; ifun = 0; iback = 0;
      if (line_search_step(true)) return;      // waiting for an evaluation, or finished
      // line search failed before any evaluation with memory to drop: loop with B reset
    }
  }

  // returns true when the machine is waiting for f/g at x_eval or has finished;
  // false when the caller must restart the iteration with a reset model
  PLX_HD bool line_search_step(bool first) {
    gd = g * d;
    bool fail = false;
    if (first) { gdold = gd; if (gd >= 0.0) fail = true; }
    if (!fail) {
      ls.step(first, f, gd, stp, 1e-3, 0.9, 0.1, 0.0, stpmx);
      if (ls.task != Dcsrch::CONVERGENCE && ls.task != Dcsrch::WARNING) {
        ++ifun; ++nfev; iback = ifun - 1;
        if (iback >= maxls) fail = true;
        else {
          x_eval = (stp == 1.0) ? z : stp * d + xt;
          x = x_eval;
          return true;
        }
      } else {
        finish_iteration();
        return true;
      }
    }
    // failure: restore the last iterate
    x = xt; f = ft; g = gt;
    if (!have_pair) {
      if (ifun > 0) { --nfev; }
      ++iter;
      finish(2, "ABNORMAL_TERMINATION_IN_LNSRCH");
      return true;
    }
    have_pair = false; B = 1.0;
    return false;
  }

  PLX_HD void finish_iteration() {
    ++iter;
    if (iter >= maxiter) { finish(1, "STOP: TOTAL NO. OF ITERATIONS REACHED LIMIT"); return; }
    if (nfev > maxfun) { finish(1, "STOP: TOTAL NO. OF F,G EVALUATIONS EXCEEDS LIMIT"); return; }
    if (projgr() <= pgtol) { finish(0, "CONVERGENCE: NORM OF PROJECTED GRADIENT <= PGTOL"); return; }
    const double ddum0 = hd_max(hd_abs(ft), hd_max(hd_abs(f), 1.0));
    if (ft - f <= epsmch * factr * ddum0) { finish(0, "CONVERGENCE: RELATIVE REDUCTION OF F <= FACTR*EPSMCH"); return; }
    const double r = g - gt, rr = r * r;
    double dr, ddum;
    if (stp == 1.0) { dr = gd - gdold; ddum = -gdold; }
    else { dr = (gd - gdold) * stp; ddum = -gdold * stp; }
    if (!(dr <= epsmch * ddum)) { have_pair = true; B = rr / dr; }
    begin_iteration();
  }

  // hand the objective and derivative at x_eval back to the machine
  PLX_HD void feed(double fval, double gval) {
    if (done) return;
    f = fval; g = gval;
    if (waiting_first) {
      waiting_first = false; nfev = 1;
      if (projgr() <= pgtol) { finish(0, "CONVERGENCE: NORM OF PROJECTED GRADIENT <= PGTOL"); return; }
      begin_iteration();
      return;
    }
    if (!line_search_step(false)) begin_iteration();
  }
};

}  // namespace plx

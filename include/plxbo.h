/*
 * plxbo.h -- C ABI of the B200-native Bayesian-optimisation suggestion path.
 *
 * The reference (Polyaxon 0.2.8, polyaxon/hpsearch/search_managers/bayesian_optimization/) is pure
 * Python and has no FFI of its own: all arithmetic happens inside scikit-learn / scipy calls.  Each
 * entry point below replaces one of those call sites; the reference file:line is given beside it.
 * INTEGRATION.md shows the ctypes stub a maintainer of the reference would bind these with.
 *
 * Conventions
 *   - plain C types; every matrix is float64, row-major (numpy C order), caller-owned;
 *   - pointers are HOST pointers unless the parameter name ends in _dev;
 *   - every function returns a status (0 = ok); plx_last_error() gives the message;
 *   - PLX_ERR_NOT_PD reports a Cholesky failure -- the Python layer raises numpy.linalg.LinAlgError,
 *     which is what sklearn does at _gpr.py:351-361;
 *   - one handle per process and device, not thread-safe; work is issued on the handle's stream
 *     (plx_bo_set_stream) and functions that return results to the host synchronise that stream;
 *   - there is no CPU fallback: without a CUDA device plx_bo_create fails.
 */
#ifndef PLXBO_H
#define PLXBO_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct plx_bo* plx_bo_handle;

enum {
  PLX_OK = 0,
  PLX_ERR_INVALID = 1,     /* bad argument */
  PLX_ERR_CUDA = 2,        /* CUDA runtime / driver error */
  PLX_ERR_NOT_PD = 3,      /* kernel matrix not positive definite */
  PLX_ERR_STATE = 4,       /* call order: data / theta not set */
  PLX_ERR_UNSUPPORTED = 5  /* e.g. Matern nu without a closed form */
};

/* kernels: acquisition_function.py:36-40 (sklearn RBF / Matern, scalar length scale) */
enum { PLX_KERNEL_RBF = 0, PLX_KERNEL_MATERN = 1 };
/* acquisition functions: acquisition_function.py:48-68 */
enum { PLX_ACQ_UCB = 0, PLX_ACQ_EI = 1, PLX_ACQ_POI = 2 };

int plx_bo_create(int device, plx_bo_handle* out);
void plx_bo_destroy(plx_bo_handle h);
const char* plx_last_error(plx_bo_handle h);   /* h may be NULL: error of the last failed create */
const char* plx_version(void);

/* Use the caller's CUDA stream (a cudaStream_t passed as void*); NULL restores the handle's own. */
int plx_bo_set_stream(plx_bo_handle h, void* stream);

/* Options: "sweep_impl" (3 [default] = a thread per candidate against the state in shared memory up to 64 observations
 * (sweep_small.cuh; 7 = the same rule, named), int8-sliced sweep on the tcgen05 tensor cores from 256 observations up
 * (fused form below 1536 observations -- 1024 where the tensor-path builders apply, see "i8_distance" --, cooperative form from there) and the FP64 DMMA kernel below 256; 5 = always the cooperative int8-sliced sweep (four CTAs per
 * candidate tile); 6 = the cooperative form with the four CTAs as two clusters of two that share the candidate planes'
 * stream by TMA multicast (same results, measured equal in speed); 4 = the fused int8-sliced sweep (one CTA per tile); 2 = int8-sliced as two kernels per chunk;
 * 1 = always FP64 DMMA with the TMA ring; 0 = plain-load cross-check); "i8_operand" (0 [default] = the digit planes of
 * the candidate tile reach the tensor core through tensor memory, 1 = both operands from shared memory; same
 * results, kept for A/B timing);
 * "fit_small" (1 [default]: with at most 64 observations plx_gp_fit runs every optimiser run from start to
 * convergence in ONE launch, optimiser included; 2: the same up to 128 observations (slower than the batched
 * sequence there, kept for tests); 0: always the batched launch sequence used for larger sets);
 * "i8_distance" (builders of the int8-sliced sweeps, at most 32 dimensions: 0 = squared distances as (c - x)^2 summed
 * directly, the arithmetic of scipy's cdist; 1 = as |u|^2 + |v|^2 - 2 u.v with u, v centred on the observations'
 * mean, u.v on the FP64 tensor path (mma.sync), pairs where that cancels recomputed directly; 2 [default] = 1 in the
 * cooperative form from 16 dimensions up (4 % faster there at the C3 shape), 0 otherwise (the fused form is slower with
 * it, and with few dimensions too many pairs need the recomputation); posterior equal to 1e-10);
 * "gen_table" (1 [default]: for a general Matern nu the kernel value and its derivative are read from Chebyshev
 * tables of k(t) e^t and g(t) e^t filled once per nu from the Bessel routine -- sweeps 27x faster at the C3 shape, fits
 * as fast as with a closed form, equal to the routine within 1e-14 --; 0: the routine for every value);
 * "blocking_sync" (0 [default]: calls wait for the GPU by spinning on the stream; 1: the host thread sleeps on an
 * event instead -- for processes whose worker threads outnumber their cores);
 * "potf_fused" (1 [default]: the diagonal 128-blocks of the fit are factorised and inverted in one pass of 128 column
 * steps; 0: in two passes, the round-1 kernel, kept for A/B runs);
 * "refine_persist" (1 [default]: with at most 256 observations plx_acq_refine runs every seed from start to
 * convergence in one launch; 0: four launches per round, as for larger sets);
 * "fit_apply" (1 [default]: plx_gp_fit ends by building the posterior state at the best theta; 0: it returns the
 * optimiser runs only -- a rank's share of a fit whose runs are dealt to several GPUs);
 * "lml_batch_limit" (k > 0: at most k thetas per pass of plx_gp_lml / plx_gp_fit; 0 = as many as fit
 * in the free device memory, which is also what bounds a batch too large for one pass). */
int plx_bo_set_option(plx_bo_handle h, const char* name, long long value);

/* Allocate, ahead of time, every workspace a problem of this shape needs (n observations of d encoded dimensions,
 * `runs` optimiser runs = 1 + n_restarts_optimizer, `seeds` refinement seeds = n_iter, `candidates` resident warm-up
 * rows = n_warmup; zeros skip a part), so that the first call of a group carries no cudaMalloc.  Workspaces are keyed
 * on n rounded up to 128.  Call it before plx_gp_set_data.  No counterpart in the reference (numpy allocates as it
 * goes); the caller is the worker start-up of cmd/celery_worker, once per process. */
int plx_bo_reserve(plx_bo_handle h, int n, int d, int runs, int seeds, long long candidates);

/* Observations.  Replaces the arguments of GaussianProcessRegressor.fit(X, y) at optimizer.py:21
 * and the kernel construction at acquisition_function.py:36-46.  nu must be 0.5, 1.5, 2.5 or inf
 * for PLX_KERNEL_MATERN; jitter is sklearn's `alpha` (1e-10 in the reference). */
int plx_gp_set_data(plx_bo_handle h, const double* X, const double* y, int n, int d, int kernel, double nu,
                    double jitter);

/* Batched log-marginal likelihood and d/dtheta at B values of theta = log(length_scale).
 * Replaces GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True)
 * (sklearn _gpr.py:583-654).  A non-PD member yields lml = -inf, grad = 0 as sklearn does (:592-593).
 * grad may be NULL. */
int plx_gp_lml(plx_bo_handle h, const double* thetas, int B, double* lml, double* grad);

/* Hyper-parameter fit.  Replaces the optimiser loop of GaussianProcessRegressor.fit
 * (sklearn _gpr.py:299-368): one bounded L-BFGS-B run from theta0 and one from each of
 * theta_restarts[0..n_restarts), all advanced together; the run with the smallest -LML wins
 * (first minimum, as numpy.argmin at :336-337) and the posterior state is built at its theta.
 * per_run (nullable) receives n_restarts+1 rows of {theta_start, theta_opt, lml_opt, n_evals}. */
int plx_gp_fit(plx_bo_handle h, double theta0, const double* theta_restarts, int n_restarts, double theta_lo,
               double theta_hi, double* theta_star, double* lml_star, double* per_run);

/* Posterior state at a fixed theta: L = chol(K + jitter I), alpha = K^-1 y, L^-1.
 * Replaces sklearn _gpr.py:349-367. */
int plx_gp_set_theta(plx_bo_handle h, double theta);

/* One more observation at the current theta, in O(n^2) instead of the O(n^3) refit: the factor, its
 * inverse and alpha are extended by one row (bordered Cholesky), giving the state that
 * plx_gp_set_data on the n+1 observations followed by plx_gp_set_theta(theta) would build, up to
 * rounding.  The reference has no such step -- it refits all observations on every iteration
 * (optimizer.py:21) -- so the host layer uses it only when asked to hold theta between refits
 * (PLX_BO_REFIT_EVERY > 1).  x_new is d doubles (HOST).  PLX_ERR_NOT_PD leaves the state untouched. */
int plx_gp_append(plx_bo_handle h, const double* x_new, double y_new);

/* Copy the state back for verification: L (n x n, lower), alpha (n), Linv (n x n); any may be NULL. */
int plx_gp_get_state(plx_bo_handle h, double* L, double* alpha, double* Linv);

/* Warm-up sweep.  Replaces UtilityFunction.max_compute's sampling stage
 * (acquisition_function.py:88-92) including compute() -> gp.predict(return_std=True)
 * (:48-68, sklearn _gpr.py:446-500).  Candidates are N x d; best_index is idx_base + the position of
 * the first maximum (numpy.argmax semantics, NaN wins); best_x receives that candidate (d doubles).
 * mean / std / acq (nullable, HOST, N doubles each) receive the per-candidate values. */
int plx_acq_sweep(plx_bo_handle h, const double* Xc, long long N, long long idx_base, int acq, double kappa,
                  double eps, double y_max, double* best_value, long long* best_index, double* best_x,
                  double* mean, double* std, double* acq_values);

/* Same, candidates already resident in device memory (no per-candidate dumps). */
int plx_acq_sweep_dev(plx_bo_handle h, const double* Xc_dev, long long N, long long idx_base, int acq,
                      double kappa, double eps, double y_max, double* best_value, long long* best_index,
                      double* best_x);

/* Resident candidates.  A seeded experiment group re-creates RandomState(seed) on every call
 * (manager.py:30, acquisition_function.py:27), so `x_tries` (acquisition_function.py:88-89) is the
 * same matrix on every BO iteration; plx_candidates_set keeps one N x d copy in device memory and
 * plx_acq_sweep_resident sweeps it without drawing or copying it again.  The host layer decides when
 * the copy is still valid (same generator state, bounds and sizes). */
int plx_candidates_set(plx_bo_handle h, const double* Xc, long long N, int d);
/* Draw the resident candidate matrix on the device, bit-identical to
 * numpy.random.RandomState.uniform(low, low + range, size=(N, d)) (acquisition_function.py:88-89):
 * key[624] / pos are the generator's MT19937 state (RandomState.get_state()[1:3]); key_out / pos_out
 * receive the state after the draw so the host generator can be moved on.  pos must be even. */
int plx_candidates_draw_mt19937(plx_bo_handle h, const unsigned* key, int pos, const double* low,
                                const double* range, long long N, int d, unsigned* key_out, int* pos_out);
/* The same draw restricted to rows [first, first + count) of that N x d matrix: the block a rank of a multi-GPU job
 * sweeps (each rank calls it with the SAME key / pos / N and its own range).  The generator is not run through the
 * rows before `first`: the device jumps to their end in O(log) steps (csrc/mt19937.cuh), so the cost is that of
 * `count` rows, spread over up to two CTAs per SM.  Only the block is held; plx_acq_sweep_resident_range and
 * plx_candidates_get take GLOBAL row numbers and refuse rows outside it.  key_out / pos_out are the state after the
 * WHOLE matrix (identical on every rank), in numpy's own representation. */
int plx_candidates_draw_mt19937_range(plx_bo_handle h, const unsigned* key, int pos, const double* low,
                                      const double* range, long long N, int d, long long first, long long count,
                                      unsigned* key_out, int* pos_out);
/* Copy rows [first, first + count) (global row numbers, inside the resident block) back to the host. */
int plx_candidates_get(plx_bo_handle h, double* out, long long first, long long count);
int plx_acq_sweep_resident(plx_bo_handle h, long long idx_base, int acq, double kappa, double eps, double y_max,
                           double* best_value, long long* best_index, double* best_x);

/* Verification hook: out[i] = k(r2[i]) exactly as the sweep's inner loop evaluates the closed-form
 * kernels (sklearn kernels.py:1563, :1722-1729 on squared scaled distances), i.e. with the branch-free
 * sqrt / exp of that loop.  Host arrays of length n. */
int plx_kernel_values(plx_bo_handle h, int kernel, double nu, const double* r2, int n, double* out);

/* Sweep rows [first, first + count) of the resident matrix; best_index is the row number in the whole
 * matrix.  This is the per-rank call of the multi-GPU path: every rank holds the same matrix, sweeps
 * its contiguous slice, and the (value, index) records are merged with numpy.argmax semantics. */
int plx_acq_sweep_resident_range(plx_bo_handle h, long long first, long long count, int acq, double kappa,
                                 double eps, double y_max, double* best_value, long long* best_index,
                                 double* best_x);

/* Device time in milliseconds of the main sweep kernel(s) of the last plx_acq_sweep* call, measured
 * with CUDA events on the launching stream, and how many kernels that call launched. */
int plx_last_sweep_stats(plx_bo_handle h, double* kernel_ms, long long* launches);
/* How many candidates of the last plx_acq_sweep* had a negative posterior variance set to zero: the event sklearn's
 * predict warns about (_gpr.py:485-491, reached from acquisition_function.py:49); the host layer raises the warning. */
int plx_last_sweep_clamped(plx_bo_handle h, long long* count);

/* Measured issue rate, in TOP/s, of the int8 tensor instruction the sliced sweep is built on
 * (tcgen05.mma kind::i8, m128 n128 k32, operands resident in shared memory, one CTA per SM): the roofline
 * denominator bench.py reports for that path.  No counterpart in the reference. */
int plx_measure_i8_peak(plx_bo_handle h, double* tops);

/* Total number of kernels this handle has launched since creation. */
long long plx_launch_count(plx_bo_handle h);

/* Refinement stage.  Replaces the loop of scipy.optimize.minimize(method="L-BFGS-B") calls in
 * UtilityFunction.max_compute (acquisition_function.py:95-111): every seed (S x d, clipped into the
 * box like scipy clips x0) is refined by a bounded limited-memory BFGS with analytic gradients, all
 * seeds in lock step on the device.  bounds is d x 2 (low, high) as SearchSpace.bounds.  Outputs per
 * seed: refined point (S x d), acquisition value there (the reference's -res.fun), success flag
 * (the reference's res.success: converged, not a line-search failure or an iteration limit) and the
 * number of evaluations (nullable).  max_iter <= 0 means scipy's default 15000.  d <= 128.
 * The caller applies the acceptance rule of :105-111 (skip failures; `>=`, so a later seed wins ties). */
int plx_acq_refine(plx_bo_handle h, const double* seeds, int S, const double* bounds, int acq, double kappa,
                   double eps, double y_max, int max_iter, double* x_out, double* acq_out, int* success_out,
                   int* nfev_out);

/* Acquisition value and analytic gradient at S points (S x d): what one compute() call of the
 * reference returns (acquisition_function.py:62-68) plus d/dx, which the reference obtains by finite
 * differences inside scipy.  grads is S x d. */
int plx_acq_value_grad(plx_bo_handle h, const double* Xp, int S, int acq, double kappa, double eps, double y_max,
                       double* values, double* grads);

/* Warm-up sweep that allocates nothing.  No counterpart in the reference: it exists for host
 * applications that own the GPU memory budget themselves (all memory handed to their own pool, or a
 * fixed MPS / container limit) and so cannot let a library call cudaMalloc, and for latency-critical
 * sections where an allocation is not acceptable.  Every input and output is a DEVICE pointer of the
 * caller: candidates Xc (N x d), observations X (n x d), alpha (n), Linv (n x n lower triangular,
 * leading dimension ldz >= n) -- the last two as plx_gp_get_state returns them -- and the two
 * result cells.  Scratch comes from a static device array (about 8.9 MB per GPU, held until the
 * process exits), which limits it to n <= 256 observations; larger n is refused with
 * PLX_ERR_INVALID.  The first call on a device does a one-time set-up and should come before any
 * stream capture.  All calls share that array: calls that overlap on the GPU race, serialise them.
 * The work is queued on `stream` (a cudaStream_t as void*, NULL = default stream) and the function
 * returns without synchronising; with N == 0 the result cells hold (-inf, INT64_MAX).
 * Errors: plx_last_error(NULL). */
int plx_acq_sweep_noalloc(int device, void* stream, const double* Xc_dev, long long N, long long idx_base,
                          const double* X_dev, const double* alpha_dev, const double* Linv_dev, int ldz, int n,
                          int d, int kernel, double nu, double theta, int acq, double kappa, double eps,
                          double y_max, double* best_value_dev, long long* best_index_dev);

/* Host-only: the one-variable L-BFGS-B used by plx_gp_fit, driven by a callback (for testing against
 * scipy.optimize.minimize(method="L-BFGS-B", jac=True), sklearn _gpr.py:660-666). */
typedef void (*plx_fg_callback)(double x, double* f, double* g, void* user);
int plx_lbfgsb1d_minimize(plx_fg_callback fg, void* user, double x0, double lo, double hi, double* x, double* f,
                          int* n_evals, int* n_iters, int* status);

#ifdef __cplusplus
}
#endif
#endif /* PLXBO_H */

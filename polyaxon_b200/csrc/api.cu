// C ABI of libplxbo.so: handle, workspace management and launch sequences.
// Declarations and the reference call site each entry point replaces: include/plxbo.h.
#include "../../include/plxbo.h"
#include "refine.cuh"
#include "append.cuh"
#include "sweep_i8p.cuh"
#include "mt19937.cuh"
#include "fit_small.cuh"
#include "sweep_small.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace plx;

namespace {
std::string g_create_error;
// Errors of the handle-less entry point plx_acq_sweep_noalloc: a pointer to a string literal (never freed, one
// word), so two host threads failing at once cannot corrupt anything; plx_last_error(NULL) returns it while it is
// newer than the last plx_bo_create.
const char* volatile g_noalloc_error = nullptr;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
}  // namespace

struct plx_bo {
  int device = 0, num_sms = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr, copy_stream = nullptr;
  cudaStream_t fact_stream = nullptr;      // high priority: diagonal-block factorisations of the fit
  std::string err;
  long long launches = 0;
  EncodeTiledFn encode = nullptr;

  // observations
  bool have_data = false;
  int n = 0, d = 0, npad = 0, nb = 0, kind = 0;
  MaternGen mg{};
  double jitter = 1e-10;
  double* dX = nullptr; size_t x_cap = 0;
  double* dy = nullptr; size_t y_cap = 0;
  int kernel_id = 0; double nu = 0.0;          // as given to plx_gp_set_data (plx_gp_append may replay it)
  double* d_app = nullptr; size_t app_cap = 0; // plx_gp_append scratch: k | l | w | pivot

  // batched fit workspace
  int bcap = 0, wnpad = 0;
  BatchPtrs bp{};
  FitMaps fit_maps{};          // 3-D tensor maps over A, L, U, T, W of the workspace
  double *d_thetas = nullptr, *d_logdet = nullptr, *d_w = nullptr, *d_alpha = nullptr, *d_partial = nullptr,
         *d_lml = nullptr, *d_grad = nullptr, *d_xs = nullptr;
  size_t xs_cap = 0;
  int* d_info = nullptr;
  int ntri = 0;

  // posterior state used by the sweep
  bool have_state = false, state_factor_valid = false;
  double theta = 0.0;
  double* dZ = nullptr; size_t z_cap = 0;
  double* d_alpha_state = nullptr; size_t as_cap = 0;
  double* dKs = nullptr; size_t ks_cap = 0;
  double* dXt = nullptr; size_t xt_cap = 0;
  int i8_distance = 2;                                // builders of the int8 sweeps: 0 direct differences, 1 FP64 tensor path, 2 by form
  double* d_mgtab = nullptr; size_t mgtab_cap = 0; int gen_table = 1;   // Chebyshev table of the general-nu kernel value (sweeps)
  double* d_xrow = nullptr; size_t xrow_cap = 0;      // centred scaled observation rows [npad][40], their squared norms and the centre:
  double* d_xnc = nullptr; size_t xnc_cap = 0;        // operands of the tensor-path builders (i8_distance 1 / 2, d <= 32)
  double* d_xmean = nullptr; size_t xmean_cap = 0;
  CUtensorMap tmZ, tmK;
  // int8-sliced variant of the sweep (sweep_impl 2, sweep_i8.cuh)
  signed char* d_kplanes = nullptr; size_t kplanes_cap = 0;     // [I8_S][i8_rows][npad]
  signed char* d_zplanes = nullptr; size_t zplanes_cap = 0;     // [I8_S][npad][npad]
  double* d_zscale = nullptr; size_t zscale_cap = 0;
  double* d_i8mean = nullptr; size_t i8mean_cap = 0;
  double* d_i8part = nullptr; size_t i8part_cap = 0;             // cooperative form: partial sums per group
  unsigned* d_i8sync = nullptr; size_t i8sync_cap = 0;           //                   counters per group
  CUtensorMap tmA8, tmB8, tmA8p;          // tmA8p: box of one plane (paired form, sweep_i8p.cuh)
  int coop_cluster_ok = 1;                // cooperative + cluster launch accepted by the driver (tried once)
  long long i8_rows = 0; int i8_npad = 0;
  bool i8_state_valid = false;

  // sweep buffers
  double* d_blk_val = nullptr; long long* d_blk_idx = nullptr;
  // result record of a sweep on the device: [0] best value, [1] best index, [2] clamped-variance count, [3..] winning row;
  // it comes back in ONE copy into pinned host memory (a pageable device->host copy blocks per call)
  double* d_rec = nullptr; size_t rec_cap = 0; double* h_pin = nullptr; size_t pin_cap = 0;
  double* d_best_val = nullptr; long long* d_best_idx = nullptr;           // = d_rec, d_rec + 1
  unsigned long long* d_clamped = nullptr; long long last_clamped = 0;     // = d_rec + 2: negative variances set to zero by the last sweep
  double* dcand[2] = {nullptr, nullptr}; size_t cand_cap = 0;
  double* d_out[3] = {nullptr, nullptr, nullptr}; size_t out_cap = 0;
  cudaEvent_t ev_copy[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
  std::vector<cudaEvent_t> ev_pool;
  int sweep_impl = 3, sweep_dbg = 0, lml_batch_limit = 0, i8_a_from_smem = 0, fit_apply = 1, fit_small = 1, refine_persist = 1, potf_fused = 1, blocking_sync = 0;
  cudaEvent_t ev_block = nullptr;
  double* d_fs = nullptr; size_t fs_cap = 0;      // one-launch fit (fit_small.cuh): starts | per-run results      // 3 = choose per GP state, see launch_sweep
  double last_ms = 0.0; long long last_launches = 0;

  // resident candidate matrix (iteration-invariant for a seeded group)
  double* d_resident = nullptr; size_t resident_cap = 0; long long resident_n = 0, resident_first = 0; int resident_d = 0;
  //   rows [resident_first, resident_first + resident_n) of the group's candidate matrix (a rank of a multi-GPU job holds its slice)

  // small persistent scratch (generator state + bounds of the device draw; verification probe)
  double* d_misc = nullptr; size_t misc_cap = 0;

  // refinement workspace
  int rf_cap_s = 0, rf_cap_npad = 0, rf_cap_d = 0;
  double *rf_kv = nullptr, *rf_kp = nullptr, *rf_v = nullptr, *rf_w = nullptr, *rf_wpart = nullptr, *rf_vec = nullptr,
         *rf_bounds = nullptr, *rf_val = nullptr, *rf_grad = nullptr;
  RefineScalars* rf_sc = nullptr;
  int* rf_active = nullptr;
  long long last_refine_rounds = 0;
};

// Wait for the handle's stream.  The default spins on the stream (lowest latency); with the "blocking_sync" option the
// host thread sleeps on an event instead, which is what a process wants whose worker threads outnumber its cores
// (polyaxon_b200.groups: 8 threads per rank, 8 ranks per box).
static cudaError_t sync_stream(plx_bo* h) {
  if (!h->blocking_sync) return cudaStreamSynchronize(h->stream);
  if (!h->ev_block) {
    const cudaError_t e = cudaEventCreateWithFlags(&h->ev_block, cudaEventBlockingSync | cudaEventDisableTiming);
    if (e != cudaSuccess) return e;
  }
  const cudaError_t e = cudaEventRecord(h->ev_block, h->stream);
  return e != cudaSuccess ? e : cudaEventSynchronize(h->ev_block);
}

#define CK(call)                                                                            \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                          \
      return PLX_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)

#define LAUNCH_CHECK(name)                                                                  \
  do {                                                                                      \
    ++h->launches;                                                                          \
    cudaError_t e_ = cudaGetLastError();                                                    \
    if (e_ != cudaSuccess) {                                                                \
      h->err = std::string("launch of " name ": ") + cudaGetErrorString(e_);                \
      return PLX_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)

static int fail(plx_bo* h, int code, const std::string& msg) { h->err = msg; return code; }

template <typename T>
static int ensure(plx_bo* h, T*& p, size_t& cap, size_t need) {
  if (need <= cap && p != nullptr) return PLX_OK;
  if (p) CK(cudaFree(p));
  p = nullptr; cap = 0;
  CK(cudaMalloc(reinterpret_cast<void**>(&p), need * sizeof(T)));
  cap = need;
  return PLX_OK;
}

static int kernel_kind(int kernel, double nu, int* kind) {
  if (kernel == PLX_KERNEL_RBF) { *kind = K_RBF; return PLX_OK; }
  if (kernel != PLX_KERNEL_MATERN) return PLX_ERR_INVALID;
  if (nu == 0.5) *kind = K_M05;
  else if (nu == 1.5) *kind = K_M15;
  else if (nu == 2.5) *kind = K_M25;
  else if (std::isinf(nu) && nu > 0) *kind = K_RBF;
  else if (nu > 0.0 && nu <= 50.0) *kind = K_GEN;
  else return PLX_ERR_UNSUPPORTED;
  return PLX_OK;
}

// nu-dependent constants of the general Matern kernel (Temme's Gamma_1 / Gamma_2 etc.)
static MaternGen make_matern_gen(double nu) {
  MaternGen m{};
  m.nu = nu;
  m.nl = static_cast<int>(nu + 0.5);
  m.mu = (m.nl == 0) ? -nu : nu - m.nl;
  m.coef = std::pow(2.0, 1.0 - nu) / std::tgamma(nu);
  m.s = std::sqrt(2.0 * nu);
  const long double mu = m.mu;
  const long double gampl = 1.0L / std::tgamma(1.0L + mu), gammi = 1.0L / std::tgamma(1.0L - mu);
  m.gampl = static_cast<double>(gampl);
  m.gammi = static_cast<double>(gammi);
  m.gam2 = static_cast<double>((gammi + gampl) / 2.0L);
  if (std::fabs(m.mu) < 1e-3) {
    // 1/Gamma(1+z) = 1 + g z + c3 z^2 + c4 z^3 + c5 z^4 + c6 z^5 ...; Gamma_1 is minus its odd part over z
    const double g = 0.5772156649015329, c4 = -0.0420026350340952, c6 = -0.0421977345555443;
    m.gam1 = -(g + c4 * m.mu * m.mu + c6 * m.mu * m.mu * m.mu * m.mu);
  } else {
    m.gam1 = static_cast<double>((gammi - gampl) / (2.0L * mu));
  }
  return m;
}

extern "C" const char* plx_version(void) { return "plxbo 0.1 (sm_100a)"; }

static const char* last_handleless_error();
extern "C" const char* plx_last_error(plx_bo_handle h) { return h ? h->err.c_str() : last_handleless_error(); }

extern "C" int plx_bo_create(int device, plx_bo_handle* out) {
  if (!out) return PLX_ERR_INVALID;
  *out = nullptr;
  g_noalloc_error = nullptr;               // plx_last_error(NULL) now reports this call
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                     " (this library has no CPU path)";
    return PLX_ERR_CUDA;
  }
  if (device < 0 || device >= count) { g_create_error = "device index out of range"; return PLX_ERR_INVALID; }
  plx_bo* h = new plx_bo();
  h->device = device;
  auto bail = [&](const std::string& m) { g_create_error = m; delete h; return PLX_ERR_CUDA; };
  if (cudaSetDevice(device) != cudaSuccess) return bail("cudaSetDevice failed");
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail("cudaGetDeviceProperties failed");
  if (prop.major != 10) return bail(std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                    "; this library is built for sm_100a only");
  h->num_sms = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) return bail("stream create failed");
  if (cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return bail("stream create failed");
  h->stream = h->own_stream;
  {
    int lo = 0, hi = 0;                                    // numerically lowest value = highest priority
    if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess ||
        cudaStreamCreateWithPriority(&h->fact_stream, cudaStreamNonBlocking, hi) != cudaSuccess)
      return bail("priority stream create failed");
  }
  for (int i = 0; i < 2; ++i) {
    if (cudaEventCreateWithFlags(&h->ev_copy[i], cudaEventDisableTiming) != cudaSuccess) return bail("event create failed");
    if (cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming) != cudaSuccess) return bail("event create failed");
  }
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn ||
      qres != cudaDriverEntryPointSuccess)
    return bail("cuTensorMapEncodeTiled not available from the driver");
  h->encode = reinterpret_cast<EncodeTiledFn>(fn);
  if (const char* e = std::getenv("PLX_BO_NO_COOP_CLUSTER"))     // profilers that cannot replay a cooperative cluster launch
    if (std::atoi(e) != 0) h->coop_cluster_ok = 0;
  if (const char* e = std::getenv("PLX_BO_SWEEP_IMPL")) {       // A/B switch for measurements; see plx_bo_set_option
    const int v = std::atoi(e);
    if (v >= 0 && v <= 7) h->sweep_impl = v;
  }
  // opt in to large dynamic shared memory once
  cudaFuncSetAttribute(potf2_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, POTF_SMEM);
  cudaFuncSetAttribute(potf2_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, POTF_SMEM);
  cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FIT_SMEM);
  cudaFuncSetAttribute(chol_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FIT_SMEM);
  cudaFuncSetAttribute(trtri_t_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FIT_SMEM);
  cudaFuncSetAttribute(trtri_u_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FIT_SMEM);
  cudaFuncSetAttribute(kinv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FIT_SMEM);
  cudaFuncSetAttribute(sweep_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SWEEP_SMEM);
  cudaFuncSetAttribute(sweep_simple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SWEEP_SIMPLE_SMEM);
  cudaFuncSetAttribute(sweep_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8_SMEM);
  cudaFuncSetAttribute(sweep_i8f_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8F_SMEM);
  cudaFuncSetAttribute(sweep_i8c_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8F_SMEM);
  cudaFuncSetAttribute(sweep_i8f_gen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8F_SMEM);
  cudaFuncSetAttribute(sweep_i8c_gen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8F_SMEM);
  cudaFuncSetAttribute(sweep_i8p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, I8F_SMEM);
  cudaFuncSetAttribute(kstar_slices_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, KSL_SMEM);
  cudaFuncSetAttribute(refine_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(double) * rf_persist_smem_doubles(RF_PERSIST_ZSMEM, RF_PERSIST_ZSMEM, RF_PERSIST_DSMEM)));
  cudaFuncSetAttribute(sweep_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sweep_small_smem(SS_NMAX, SS_XMAX / SS_NMAX)));
  cudaFuncSetAttribute(fit_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FS_SMEM);
  cudaFuncSetAttribute(state_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FS_SMEM);
  cudaFuncSetAttribute(mt19937_draw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(unsigned) * MT_SEQ + sizeof(double) * 2 * 2048));
  if (cudaGetLastError() != cudaSuccess) return bail("cudaFuncSetAttribute failed (is the library built for this GPU?)");
  if (cudaMalloc(&h->d_blk_val, sizeof(double) * 1024) != cudaSuccess ||
      cudaMalloc(&h->d_blk_idx, sizeof(long long) * 1024) != cudaSuccess ||
      cudaMalloc(&h->d_rec, sizeof(double) * (3 + 2048)) != cudaSuccess ||
      cudaHostAlloc(&h->h_pin, sizeof(double) * (3 + 2048), cudaHostAllocDefault) != cudaSuccess)
    return bail("cudaMalloc failed");
  h->rec_cap = h->pin_cap = 3 + 2048;
  h->d_best_val = h->d_rec; h->d_best_idx = reinterpret_cast<long long*>(h->d_rec + 1);
  h->d_clamped = reinterpret_cast<unsigned long long*>(h->d_rec + 2);
  *out = h;
  return PLX_OK;
}

extern "C" void plx_bo_destroy(plx_bo_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  void* ptrs[] = {h->dX, h->dy, h->bp.A, h->bp.L, h->bp.U, h->bp.W, h->bp.T, h->d_thetas, h->d_logdet, h->d_w,
                  h->d_alpha, h->d_partial, h->d_lml, h->d_grad, h->d_xs, h->d_info, h->dZ, h->d_alpha_state, h->dKs, h->dXt,
                  h->d_blk_val, h->d_blk_idx, h->d_rec, h->dcand[0], h->dcand[1],
                  h->d_out[0], h->d_out[1], h->d_out[2], h->rf_kv, h->rf_kp, h->rf_v, h->rf_w, h->rf_wpart, h->rf_vec,
                  h->rf_bounds, h->rf_val, h->rf_grad, h->rf_sc, h->rf_active, h->d_resident, h->d_misc, h->d_app,
                  h->d_kplanes, h->d_zplanes, h->d_zscale, h->d_i8mean, h->d_i8part, h->d_i8sync, h->d_fs, h->d_mgtab, h->d_xrow, h->d_xnc, h->d_xmean};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (h->h_pin) cudaFreeHost(h->h_pin);
  for (int i = 0; i < 2; ++i) { if (h->ev_copy[i]) cudaEventDestroy(h->ev_copy[i]); if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]); }
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  if (h->ev_block) cudaEventDestroy(h->ev_block);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->fact_stream) cudaStreamDestroy(h->fact_stream);
  delete h;
}

extern "C" int plx_bo_set_stream(plx_bo_handle h, void* stream) {
  if (!h) return PLX_ERR_INVALID;
  h->stream = stream ? reinterpret_cast<cudaStream_t>(stream) : h->own_stream;
  return PLX_OK;
}

extern "C" int plx_bo_set_option(plx_bo_handle h, const char* name, long long value) {
  if (!h || !name) return PLX_ERR_INVALID;
  if (std::strcmp(name, "sweep_impl") == 0 && value >= 0 && value <= 7) { h->sweep_impl = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "lml_batch_limit") == 0 && value >= 0) { h->lml_batch_limit = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "gen_table") == 0 && (value == 0 || value == 1)) {
    h->gen_table = static_cast<int>(value);
    if (h->kind == K_GEN && h->d_mgtab && h->have_data) h->mg.tab = value ? h->d_mgtab : nullptr;
    return PLX_OK;
  }
  if (std::strcmp(name, "i8_distance") == 0 && value >= 0 && value <= 2) { h->i8_distance = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "blocking_sync") == 0 && (value == 0 || value == 1)) { h->blocking_sync = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "potf_fused") == 0 && (value == 0 || value == 1)) { h->potf_fused = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "refine_persist") == 0 && (value == 0 || value == 1)) { h->refine_persist = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "fit_small") == 0 && value >= 0 && value <= 2) { h->fit_small = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "fit_apply") == 0 && (value == 0 || value == 1)) { h->fit_apply = static_cast<int>(value); return PLX_OK; }
  if (std::strcmp(name, "i8_operand") == 0 && (value == 0 || value == 1)) { h->i8_a_from_smem = static_cast<int>(value); return PLX_OK; }
#ifdef PLX_PROFILE
  if (std::strcmp(name, "sweep_dbg") == 0) { h->sweep_dbg = static_cast<int>(value); return PLX_OK; }
#endif
  return fail(h, PLX_ERR_INVALID, std::string("unknown option or value: ") + name);
}

extern "C" long long plx_launch_count(plx_bo_handle h) { return h ? h->launches : 0; }

extern "C" int plx_gp_set_data(plx_bo_handle h, const double* X, const double* y, int n, int d, int kernel,
                               double nu, double jitter) {
  if (!h || !X || !y || n <= 0 || d <= 0) return h ? fail(h, PLX_ERR_INVALID, "set_data: bad arguments") : PLX_ERR_INVALID;
  int kind;
  const int rc = kernel_kind(kernel, nu, &kind);
  if (rc == PLX_ERR_UNSUPPORTED)
    return fail(h, rc, "Matern nu must be positive and at most 50");
  if (rc != PLX_OK) return fail(h, rc, "unknown kernel id");
  CK(cudaSetDevice(h->device));
  h->n = n; h->d = d; h->kind = kind; h->jitter = jitter; h->kernel_id = kernel; h->nu = nu;
  h->mg = (kind == K_GEN) ? make_matern_gen(nu) : MaternGen{};
  if (kind == K_GEN) {
    // kernel value and derivative come from tables filled once per nu (gp_kernels.cuh)
    int rt;
    if ((rt = ensure(h, h->d_mgtab, h->mgtab_cap, static_cast<size_t>(2) * MG_NINT * MG_ROW)) != PLX_OK) return rt;
    matern_tab_kernel<<<MG_NINT, MG_DEG, 0, h->stream>>>(h->mg, h->d_mgtab);
    LAUNCH_CHECK("matern_tab_kernel");
    if (h->gen_table) h->mg.tab = h->d_mgtab;
  }
  h->npad = (n + 127) / 128 * 128; h->nb = h->npad / 128;
  { const int nt32 = (n + 31) / 32; h->ntri = nt32 * (nt32 + 1) / 2; }   // lower 32-tiles that hold observations
  h->have_data = false; h->have_state = false; h->state_factor_valid = false;
  int r;
  if ((r = ensure(h, h->dX, h->x_cap, static_cast<size_t>(n) * d)) != PLX_OK) return r;
  if ((r = ensure(h, h->dy, h->y_cap, static_cast<size_t>(h->npad))) != PLX_OK) return r;
  CK(cudaMemsetAsync(h->dy, 0, sizeof(double) * h->npad, h->stream));
  CK(cudaMemcpyAsync(h->dX, X, sizeof(double) * n * d, cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(h->dy, y, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
  CK(sync_stream(h));
  h->have_data = true;
  return PLX_OK;
}

// 3-D map over a batched buffer: `cols` doubles per row, `rows` rows per batch member, members
// `stride` doubles apart; box = one 16 x 128 chunk of one member, 128-byte swizzle.
static int make_map3(plx_bo* h, CUtensorMap* map, double* base, size_t cols, size_t rows, size_t batch, size_t stride) {
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(batch)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols) * 8, static_cast<cuuint64_t>(stride) * 8};
  cuuint32_t box[3] = {KCH, TILE, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, PLX_ERR_CUDA, "cuTensorMapEncodeTiled (3-D) failed with code " + std::to_string(r));
  return PLX_OK;
}

static int ensure_batch(plx_bo* h, int B) {
  if (B <= h->bcap && h->wnpad == h->npad) return PLX_OK;
  const int cap = std::max(B, h->wnpad == h->npad ? h->bcap : 0);
  void* old[] = {h->bp.A, h->bp.L, h->bp.U, h->bp.W, h->bp.T, h->d_thetas, h->d_logdet, h->d_w, h->d_alpha,
                 h->d_partial, h->d_lml, h->d_grad, h->d_info};
  for (void* p : old) if (p) CK(cudaFree(p));
  h->bp = BatchPtrs{}; h->d_thetas = h->d_logdet = h->d_w = h->d_alpha = h->d_partial = h->d_lml = h->d_grad = nullptr;
  h->d_info = nullptr; h->bcap = 0; h->state_factor_valid = false;
  const size_t mat = static_cast<size_t>(h->npad) * h->npad, blk = static_cast<size_t>(h->nb) * 128 * 128;
  const int nt32max = h->npad / 32;                   // partial sums: room for every 32-tile of npad
  const int ntri_cap = nt32max * (nt32max + 1) / 2;
  CK(cudaMalloc(&h->bp.A, sizeof(double) * mat * cap));
  CK(cudaMalloc(&h->bp.L, sizeof(double) * mat * cap));
  CK(cudaMalloc(&h->bp.U, sizeof(double) * mat * cap));
  CK(cudaMalloc(&h->bp.W, sizeof(double) * blk * cap));
  CK(cudaMalloc(&h->bp.T, sizeof(double) * mat * cap));
  CK(cudaMalloc(&h->d_thetas, sizeof(double) * cap));
  CK(cudaMalloc(&h->d_logdet, sizeof(double) * cap));
  CK(cudaMalloc(&h->d_w, sizeof(double) * h->npad * cap));
  CK(cudaMalloc(&h->d_alpha, sizeof(double) * h->npad * cap));
  CK(cudaMalloc(&h->d_partial, sizeof(double) * ntri_cap * cap));
  CK(cudaMalloc(&h->d_lml, sizeof(double) * cap));
  CK(cudaMalloc(&h->d_grad, sizeof(double) * cap));
  CK(cudaMalloc(&h->d_info, sizeof(int) * cap));
  h->bp.mat_stride = mat; h->bp.blk_stride = blk;
  h->bcap = cap; h->wnpad = h->npad;
  {
    int r;
    const size_t np = h->npad;
    if ((r = make_map3(h, &h->fit_maps.A, h->bp.A, np, np, cap, mat)) != PLX_OK) return r;
    if ((r = make_map3(h, &h->fit_maps.L, h->bp.L, np, np, cap, mat)) != PLX_OK) return r;
    if ((r = make_map3(h, &h->fit_maps.U, h->bp.U, np, np, cap, mat)) != PLX_OK) return r;
    if ((r = make_map3(h, &h->fit_maps.T, h->bp.T, np, np, cap, mat)) != PLX_OK) return r;
    if ((r = make_map3(h, &h->fit_maps.W, h->bp.W, 128, static_cast<size_t>(h->nb) * 128, cap, blk)) != PLX_OK) return r;
  }
  return PLX_OK;
}

// Launch sequence of one batched LML (+gradient) evaluation; results stay on the device.
static int launch_lml(plx_bo* h, int B, bool want_grad) {
  cudaStream_t st = h->stream;
  const int npad = h->npad, nb = h->nb, n = h->n;
  {
    int r;
    if ((r = ensure(h, h->d_xs, h->xs_cap, static_cast<size_t>(B) * n * h->d)) != PLX_OK) return r;
  }
  scale_x_batch_kernel<<<dim3((n * h->d + 255) / 256, B), 256, 0, st>>>(h->dX, n * h->d, h->d_thetas, h->d_xs, h->d_logdet,
                                                                       h->d_info);
  LAUNCH_CHECK("scale_x_batch_kernel");
  const int t32 = npad / 32;
  kmat_kernel<0><<<dim3(t32 * (t32 + 1) / 2, 1, B), 256, 0, st>>>(h->d_xs, n, h->d, npad, h->kind, h->mg, h->jitter,
                                                                 h->bp, nullptr, nullptr, 0);
  LAUNCH_CHECK("kmat_kernel");
  // Blocked right-looking factorisation in panels of PANEL block columns.  The diagonal-block kernel is a
  // serial chain of 128 column steps on one CTA per theta, so it runs on a second stream: the trailing
  // update that feeds it is issued in two parts, the next diagonal tile first, and potf2(k + 1) then
  // overlaps the rest of that update.
  const int PANEL = 4;
  cudaStream_t s2 = h->fact_stream;
  cudaEvent_t ev_diag = h->ev_copy[0], ev_fact = h->ev_done[0];
  auto syrk_tiles = [&](int jlo, int jhi) { int t = 0; for (int j = jlo; j < jhi; ++j) t += nb - j; return t; };
  auto potf2 = h->potf_fused ? potf2_fused_kernel : potf2_inv_kernel;     // factor + inverse in one pass of 128 steps, or in two
  potf2<<<B, 256, POTF_SMEM, st>>>(h->bp, 0, n, npad, h->d_logdet, h->d_info);
  LAUNCH_CHECK("potf2_inv_kernel");
  for (int p0 = 0; p0 < nb; p0 += PANEL) {
    const int p1 = std::min(p0 + PANEL, nb);
    for (int k = p0; k < p1; ++k) {
      if (k > 0) CK(cudaStreamWaitEvent(st, ev_fact, 0));          // potf2(k) ran on the second stream
      const int T = nb - k - 1;
      if (T == 0) break;
      chol_panel_kernel<<<dim3(T, B), 384, FIT_SMEM, st>>>(h->fit_maps.A, h->fit_maps.W, h->bp, k, npad);
      LAUNCH_CHECK("chol_panel_kernel");
      // inside the panel only its remaining columns need column k now; after the panel's last column
      // everything to its right is updated once with the whole panel as the contraction
      const bool inside = k + 1 < p1;
      const int kcol = inside ? k : p0, kw = inside ? 1 : p1 - p0;
      const int jlo = k + 1, jhi = inside ? p1 : nb;
      const int tiles = syrk_tiles(jlo, jhi);
      chol_syrk_kernel<<<dim3(1, B), 384, FIT_SMEM, st>>>(h->fit_maps.L, h->bp, kcol, kw, jlo, jhi, nb, npad, 0);
      LAUNCH_CHECK("chol_syrk_kernel");
      CK(cudaEventRecord(ev_diag, st));
      CK(cudaStreamWaitEvent(s2, ev_diag, 0));
      potf2<<<B, 256, POTF_SMEM, s2>>>(h->bp, k + 1, n, npad, h->d_logdet, h->d_info);
      LAUNCH_CHECK("potf2_inv_kernel");
      CK(cudaEventRecord(ev_fact, s2));
      if (tiles > 1) {
        chol_syrk_kernel<<<dim3(tiles - 1, B), 384, FIT_SMEM, st>>>(h->fit_maps.L, h->bp, kcol, kw, jlo, jhi, nb, npad, 1);
        LAUNCH_CHECK("chol_syrk_kernel");
      }
    }
  }
  for (int s = 1; s < nb; s *= 2) {               // recursive doubling: 2 launches per level
    const int tiles = ((nb + 2 * s - 1) / (2 * s)) * s * s;
    trtri_t_kernel<<<dim3(tiles, B), 384, FIT_SMEM, st>>>(h->fit_maps.U, h->fit_maps.L, h->bp, s, nb, npad);
    LAUNCH_CHECK("trtri_t_kernel");
    trtri_u_kernel<<<dim3(tiles, B), 384, FIT_SMEM, st>>>(h->fit_maps.T, h->fit_maps.A, h->bp, s, nb, npad);
    LAUNCH_CHECK("trtri_u_kernel");
  }
  uty_kernel<<<dim3(npad / 32, B), 256, 0, st>>>(h->bp, h->dy, npad, h->d_w);
  LAUNCH_CHECK("uty_kernel");
  uw_kernel<<<dim3(npad / 8, B), 256, 0, st>>>(h->bp, h->d_w, npad, h->d_alpha);
  LAUNCH_CHECK("uw_kernel");
  if (want_grad) {
    kinv_kernel<<<dim3(nb * (nb + 1) / 2, B), 384, FIT_SMEM, st>>>(h->fit_maps.U, h->bp, npad);
    LAUNCH_CHECK("kinv_kernel");
    kmat_kernel<1><<<dim3(h->ntri, 1, B), 256, 0, st>>>(h->d_xs, n, h->d, npad, h->kind, h->mg, h->jitter, h->bp,
                                                       h->d_alpha, h->d_partial, h->ntri);
    LAUNCH_CHECK("kmat_kernel<grad>");
  }
  lml_finalize_kernel<<<B, 256, 0, st>>>(h->dy, h->d_alpha, h->d_logdet, h->d_info,
                                         want_grad ? h->d_partial : nullptr, h->ntri, n, npad, h->d_lml,
                                         want_grad ? h->d_grad : nullptr);
  LAUNCH_CHECK("lml_finalize_kernel");
  return PLX_OK;
}

// Largest number of thetas evaluated in one pass: five n x n matrices per theta must fit in (most of)
// the memory that is free now plus what the current workspace already holds.
static int batch_limit(plx_bo* h) {
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return 1;
  const size_t mat = static_cast<size_t>(h->npad) * h->npad * sizeof(double);
  const size_t per_theta = 4 * mat + static_cast<size_t>(h->nb) * 128 * 128 * sizeof(double) + 4 * mat / 64;
  size_t avail = free_b;
  if (h->wnpad == h->npad) avail += static_cast<size_t>(h->bcap) * per_theta;
  const size_t usable = avail / 10 * 8;
  return static_cast<int>(std::max<size_t>(1, std::min<size_t>(usable / per_theta, 1 << 20)));
}

static int eval_lml(plx_bo* h, const double* thetas, int B, double* lml, double* grad) {
  int r;
  int limit = (B <= h->bcap && h->wnpad == h->npad) ? B : batch_limit(h);
  if (h->lml_batch_limit > 0) limit = std::min(limit, h->lml_batch_limit);
  for (int first = 0; first < B; first += limit) {
    const int b = std::min(limit, B - first);
    if ((r = ensure_batch(h, b)) != PLX_OK) return r;
    h->state_factor_valid = false;
    // thetas in and results out through pinned memory: the copies queue on the stream and one synchronisation
    // collects both results (pageable device->host copies block one by one)
    const bool pinned = static_cast<size_t>(3) * b <= h->pin_cap;
    if (pinned) {
      std::memcpy(h->h_pin, thetas + first, sizeof(double) * b);
      CK(cudaMemcpyAsync(h->d_thetas, h->h_pin, sizeof(double) * b, cudaMemcpyHostToDevice, h->stream));
    } else {
      CK(cudaMemcpyAsync(h->d_thetas, thetas + first, sizeof(double) * b, cudaMemcpyHostToDevice, h->stream));
    }
    if ((r = launch_lml(h, b, grad != nullptr)) != PLX_OK) return r;
    double* out_l = pinned ? h->h_pin + b : lml + first;
    double* out_g = pinned ? h->h_pin + 2 * b : (grad ? grad + first : nullptr);
    CK(cudaMemcpyAsync(out_l, h->d_lml, sizeof(double) * b, cudaMemcpyDeviceToHost, h->stream));
    if (grad) CK(cudaMemcpyAsync(out_g, h->d_grad, sizeof(double) * b, cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
    if (pinned) {
      std::memcpy(lml + first, out_l, sizeof(double) * b);
      if (grad) std::memcpy(grad + first, out_g, sizeof(double) * b);
    }
  }
  return PLX_OK;
}

extern "C" int plx_gp_lml(plx_bo_handle h, const double* thetas, int B, double* lml, double* grad) {
  if (!h || !thetas || !lml || B <= 0) return h ? fail(h, PLX_ERR_INVALID, "gp_lml: bad arguments") : PLX_ERR_INVALID;
  if (!h->have_data) return fail(h, PLX_ERR_STATE, "gp_lml: call plx_gp_set_data first");
  CK(cudaSetDevice(h->device));
  return eval_lml(h, thetas, B, lml, grad);
}

static int make_map(plx_bo* h, CUtensorMap* map, double* base, size_t rows, size_t cols, int box_rows = TILE) {
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * 8};
  cuuint32_t box[2] = {KCH, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, base, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, PLX_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string(r));
  return PLX_OK;
}

extern "C" int plx_gp_set_theta(plx_bo_handle h, double theta) {
  if (!h) return PLX_ERR_INVALID;
  if (!h->have_data) return fail(h, PLX_ERR_STATE, "gp_set_theta: call plx_gp_set_data first");
  CK(cudaSetDevice(h->device));
  h->have_state = false;
  int r;
  const size_t mat = static_cast<size_t>(h->npad) * h->npad;
  const bool small = h->fit_small && h->n <= (h->fit_small == 2 ? FS_NMAX : FS_NUSE);
  if (small) {
    if ((r = ensure_batch(h, 1)) != PLX_OK) return r;           // L lives in the fit workspace (plx_gp_get_state, plx_gp_append)
    h->state_factor_valid = false;
  } else {
    double lml;
    r = eval_lml(h, &theta, 1, &lml, nullptr);
    if (r != PLX_OK) return r;
    if (std::isinf(lml) && lml < 0)
      return fail(h, PLX_ERR_NOT_PD, "kernel matrix is not positive definite at theta=" + std::to_string(theta));
  }
  if ((r = ensure(h, h->dZ, h->z_cap, mat)) != PLX_OK) return r;
  if ((r = ensure(h, h->d_alpha_state, h->as_cap, static_cast<size_t>(h->npad))) != PLX_OK) return r;
  if ((r = ensure(h, h->dKs, h->ks_cap, static_cast<size_t>(h->num_sms) * 128 * h->npad)) != PLX_OK) return r;
  if ((r = ensure(h, h->dXt, h->xt_cap, static_cast<size_t>(h->d) * h->npad)) != PLX_OK) return r;
  scale_transpose_x_kernel<<<dim3((h->npad + 255) / 256, h->d), 256, 0, h->stream>>>(h->dX, h->n, h->d, h->npad,
                                                                                   std::exp(theta), h->dXt);
  LAUNCH_CHECK("scale_transpose_x_kernel");
  if (small) {
    // small observation sets: factor, inverse and alpha in one launch of one CTA (fit_small.cuh)
    if ((r = ensure(h, h->d_fs, h->fs_cap, 8)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_xs, h->xs_cap, static_cast<size_t>(h->n) * h->d)) != PLX_OK) return r;
    FitSmallParams fp{};
    fp.X = h->dX; fp.y = h->dy; fp.n = h->n; fp.d = h->d; fp.kind = h->kind; fp.R = 1; fp.mg = h->mg;
    fp.jitter = h->jitter; fp.xs = h->d_xs;
    state_small_kernel<<<1, FS_THREADS, FS_SMEM, h->stream>>>(fp, theta, h->npad, h->bp.L, h->dZ, h->d_alpha_state, h->d_fs);
    LAUNCH_CHECK("state_small_kernel");
    double* st = h->h_pin;
    CK(cudaMemcpyAsync(st, h->d_fs, sizeof(double) * 2, cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
    if (st[0] != 0.0)
      return fail(h, PLX_ERR_NOT_PD, "kernel matrix is not positive definite at theta=" + std::to_string(theta));
  } else {
    transpose_kernel<<<dim3(h->npad / 32, h->npad / 32), 256, 0, h->stream>>>(h->bp.U, h->dZ, h->npad);
    LAUNCH_CHECK("transpose_kernel");
    CK(cudaMemcpyAsync(h->d_alpha_state, h->d_alpha, sizeof(double) * h->npad, cudaMemcpyDeviceToDevice, h->stream));
    CK(sync_stream(h));
  }
  if ((r = make_map(h, &h->tmZ, h->dZ, h->npad, h->npad)) != PLX_OK) return r;
  if ((r = make_map(h, &h->tmK, h->dKs, static_cast<size_t>(h->num_sms) * 128, h->npad)) != PLX_OK) return r;
  h->theta = theta; h->have_state = true; h->state_factor_valid = true; h->i8_state_valid = false;
  return PLX_OK;
}

extern "C" int plx_gp_get_state(plx_bo_handle h, double* L, double* alpha, double* Linv) {
  if (!h) return PLX_ERR_INVALID;
  if (!h->have_state) return fail(h, PLX_ERR_STATE, "gp_get_state: no posterior state");
  CK(cudaSetDevice(h->device));
  const size_t n = h->n, npad = h->npad;
  if (L) {
    if (!h->state_factor_valid) return fail(h, PLX_ERR_STATE, "gp_get_state: L was overwritten by a later plx_gp_lml");
    CK(cudaMemcpy2DAsync(L, n * 8, h->bp.L, npad * 8, n * 8, n, cudaMemcpyDeviceToHost, h->stream));
  }
  if (Linv) CK(cudaMemcpy2DAsync(Linv, n * 8, h->dZ, npad * 8, n * 8, n, cudaMemcpyDeviceToHost, h->stream));
  if (alpha) CK(cudaMemcpyAsync(alpha, h->d_alpha_state, n * 8, cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  if (L)      // the factor's strict upper triangle is never written on the device
    for (size_t i = 0; i < n; ++i) for (size_t j = i + 1; j < n; ++j) L[i * n + j] = 0.0;
  if (Linv)
    for (size_t i = 0; i < n; ++i) for (size_t j = i + 1; j < n; ++j) Linv[i * n + j] = 0.0;
  return PLX_OK;
}

// One more observation at the current theta.  See append.cuh for the algebra.
extern "C" int plx_gp_append(plx_bo_handle h, const double* x_new, double y_new) {
  if (!h || !x_new) return h ? fail(h, PLX_ERR_INVALID, "gp_append: bad arguments") : PLX_ERR_INVALID;
  if (!h->have_state) return fail(h, PLX_ERR_STATE, "gp_append: no posterior state (call plx_gp_fit or plx_gp_set_theta)");
  CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  const int n0 = h->n, d = h->d, npad = h->npad;
  int r;
  if (h->x_cap < static_cast<size_t>(n0 + 1) * d) {       // grow the observation matrix, keeping its rows
    double* nx = nullptr;
    const size_t cap = static_cast<size_t>(npad + 128) * d;
    CK(cudaMalloc(reinterpret_cast<void**>(&nx), cap * sizeof(double)));
    CK(cudaMemcpyAsync(nx, h->dX, sizeof(double) * n0 * d, cudaMemcpyDeviceToDevice, st));
    CK(sync_stream(h));
    CK(cudaFree(h->dX));
    h->dX = nx; h->x_cap = cap;
  }
  if ((r = ensure(h, h->d_app, h->app_cap, static_cast<size_t>(3) * (npad + 128) + 8)) != PLX_OK) return r;
  double *kv = h->d_app, *lv = kv + npad, *wv = lv + npad, *piv = wv + npad;
  const double ell = std::exp(h->theta);
  // the bordered pivot first: k = k(X, x_new), l = L^-1 k, pivot^2 = 1 + jitter - l.l.  Only scratch (and the
  // unused row n0 of the observation matrix) is written before it is known to be positive.
  CK(cudaMemcpyAsync(h->dX + static_cast<size_t>(n0) * d, x_new, sizeof(double) * d, cudaMemcpyHostToDevice, st));
  append_kvec_kernel<<<(n0 + 255) / 256, 256, 0, st>>>(h->dX, n0, d, ell, h->kind, h->mg, kv);
  LAUNCH_CHECK("append_kvec_kernel");
  trmv_lower_kernel<<<(n0 + 7) / 8, 256, 0, st>>>(h->dZ, npad, n0, kv, lv);
  LAUNCH_CHECK("trmv_lower_kernel");
  append_pivot_kernel<<<1, 256, 0, st>>>(lv, n0, 1.0 + h->jitter, piv);
  LAUNCH_CHECK("append_pivot_kernel");
  double lam2 = 0.0;
  CK(cudaMemcpyAsync(&lam2, piv, sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(sync_stream(h));
  if (!(lam2 > 0.0))       // nothing of the state has been touched yet
    return fail(h, PLX_ERR_NOT_PD, "gp_append: the extended kernel matrix is not positive definite (pivot^2 = " +
                                       std::to_string(lam2) + ")");
  if (n0 == npad) {
    // the padded storage is full: rebuild at the next padding step from the device copy of the data (the pivot
    // above says the extended matrix is positive definite, so the rebuild does not fail on that account)
    std::vector<double> X(static_cast<size_t>(n0 + 1) * d), y(n0 + 1);
    CK(cudaMemcpyAsync(X.data(), h->dX, sizeof(double) * n0 * d, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(y.data(), h->dy, sizeof(double) * n0, cudaMemcpyDeviceToHost, st));
    CK(sync_stream(h));
    std::memcpy(X.data() + static_cast<size_t>(n0) * d, x_new, sizeof(double) * d);
    y[n0] = y_new;
    const double theta = h->theta;
    r = plx_gp_set_data(h, X.data(), y.data(), n0 + 1, d, h->kernel_id, h->nu, h->jitter);
    return r != PLX_OK ? r : plx_gp_set_theta(h, theta);
  }
  const double lam = std::sqrt(lam2);
  // Z row n0 <- -(l^T Z) / lambda, written in place (only rows < n0 are read), then the diagonal
  trmv_lower_t_kernel<<<(n0 + 31) / 32, 256, 0, st>>>(h->dZ, npad, n0, lv, -1.0 / lam,
                                                      h->dZ + static_cast<size_t>(n0) * npad);
  LAUNCH_CHECK("trmv_lower_t_kernel");
  append_row_kernel<<<(n0 + 256) / 256, 256, 0, st>>>(h->dZ, npad, n0, nullptr, 1.0 / lam);   // diagonal only
  LAUNCH_CHECK("append_row_kernel");
  if (h->state_factor_valid) {
    append_row_kernel<<<(n0 + 256) / 256, 256, 0, st>>>(h->bp.L, npad, n0, lv, lam);
    LAUNCH_CHECK("append_row_kernel");
  }
  CK(cudaMemcpyAsync(h->dy + n0, &y_new, sizeof(double), cudaMemcpyHostToDevice, st));
  const int n1 = n0 + 1;
  trmv_lower_kernel<<<(n1 + 7) / 8, 256, 0, st>>>(h->dZ, npad, n1, h->dy, wv);
  LAUNCH_CHECK("trmv_lower_kernel");
  trmv_lower_t_kernel<<<(n1 + 31) / 32, 256, 0, st>>>(h->dZ, npad, n1, wv, 1.0, h->d_alpha_state);
  LAUNCH_CHECK("trmv_lower_t_kernel");
  scale_transpose_x_kernel<<<dim3((npad + 255) / 256, d), 256, 0, st>>>(h->dX, n1, d, npad, ell, h->dXt);
  LAUNCH_CHECK("scale_transpose_x_kernel");
  CK(sync_stream(h));          // y_new / lam2 live on this frame
  h->n = n1; h->i8_state_valid = false;
  { const int nt32 = (n1 + 31) / 32; h->ntri = nt32 * (nt32 + 1) / 2; }
  return PLX_OK;
}

extern "C" int plx_gp_fit(plx_bo_handle h, double theta0, const double* theta_restarts, int n_restarts,
                          double theta_lo, double theta_hi, double* theta_star, double* lml_star, double* per_run) {
  if (!h || n_restarts < 0 || (n_restarts > 0 && !theta_restarts) || !(theta_lo < theta_hi))
    return h ? fail(h, PLX_ERR_INVALID, "gp_fit: bad arguments") : PLX_ERR_INVALID;
  if (!h->have_data) return fail(h, PLX_ERR_STATE, "gp_fit: call plx_gp_set_data first");
  CK(cudaSetDevice(h->device));
  const int R = n_restarts + 1;
  if (h->fit_small && h->n <= (h->fit_small == 2 ? FS_NMAX : FS_NUSE)) {
    // small observation sets: every run from start to convergence in one launch (fit_small.cuh)
    int r;
    if ((r = ensure(h, h->d_fs, h->fs_cap, static_cast<size_t>(5) * R)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_xs, h->xs_cap, static_cast<size_t>(R) * h->n * h->d)) != PLX_OK) return r;
    std::vector<double> st(R), out(static_cast<size_t>(4) * R);
    for (int i = 0; i < R; ++i) st[i] = (i == 0) ? theta0 : theta_restarts[i - 1];
    const bool pinned = static_cast<size_t>(5) * R <= h->pin_cap;      // small transfers through pinned memory: no blocking copies
    double* stage = pinned ? h->h_pin : st.data();
    if (pinned) std::memcpy(stage, st.data(), sizeof(double) * R);
    CK(cudaMemcpyAsync(h->d_fs, stage, sizeof(double) * R, cudaMemcpyHostToDevice, h->stream));
    FitSmallParams fp;
    fp.X = h->dX; fp.y = h->dy; fp.n = h->n; fp.d = h->d; fp.kind = h->kind; fp.R = R; fp.mg = h->mg;
    fp.jitter = h->jitter; fp.lo = theta_lo; fp.hi = theta_hi; fp.starts = h->d_fs; fp.xs = h->d_xs;
    fp.per_run = h->d_fs + R;
    fit_small_kernel<<<R, FS_THREADS, FS_SMEM, h->stream>>>(fp);
    LAUNCH_CHECK("fit_small_kernel");
    CK(cudaMemcpyAsync(pinned ? h->h_pin + R : out.data(), h->d_fs + R, sizeof(double) * 4 * R, cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
    if (pinned) std::memcpy(out.data(), h->h_pin + R, sizeof(double) * 4 * R);
    int best = 0;
    for (int i = 1; i < R; ++i) if (-out[4 * i + 2] < -out[4 * best + 2]) best = i;     // first minimum of -LML
    if (per_run) std::memcpy(per_run, out.data(), sizeof(double) * 4 * R);
    if (theta_star) *theta_star = out[4 * best + 1];
    if (lml_star) *lml_star = out[4 * best + 2];
    if (!h->fit_apply) return PLX_OK;
    return plx_gp_set_theta(h, out[4 * best + 1]);
  }
  std::vector<Lbfgsb1D> runs(R);
  std::vector<double> starts(R);
  for (int r = 0; r < R; ++r) {
    starts[r] = (r == 0) ? theta0 : theta_restarts[r - 1];
    runs[r].lo = theta_lo; runs[r].hi = theta_hi;
    runs[r].start(starts[r]);
  }
  std::vector<double> xs(R), fs(R), gs(R);
  std::vector<int> who(R);
  for (;;) {
    int m = 0;
    for (int r = 0; r < R; ++r) if (!runs[r].done) { who[m] = r; xs[m] = runs[r].x_eval; ++m; }
    if (m == 0) break;
    const int rc = eval_lml(h, xs.data(), m, fs.data(), gs.data());
    if (rc != PLX_OK) return rc;
    for (int i = 0; i < m; ++i) runs[who[i]].feed(-fs[i], -gs[i]);   // minimise -LML
  }
  int best = 0;
  for (int r = 1; r < R; ++r) if (runs[r].f < runs[best].f) best = r;     // first minimum, like numpy.argmin
  if (per_run)
    for (int r = 0; r < R; ++r) {
      per_run[4 * r + 0] = starts[r]; per_run[4 * r + 1] = runs[r].x;
      per_run[4 * r + 2] = -runs[r].f; per_run[4 * r + 3] = runs[r].nfev;
    }
  if (theta_star) *theta_star = runs[best].x;
  if (lml_star) *lml_star = -runs[best].f;
  if (!h->fit_apply) return PLX_OK;       // runs only (a rank's share of a sharded fit)
  return plx_gp_set_theta(h, runs[best].x);
}

// ---------------------------------------------------------------------------------------------
// sweep
// ---------------------------------------------------------------------------------------------
static void fill_params(plx_bo* h, SweepParams& p, const double* Xc_dev, long long N, long long idx_base, int acq,
                        double kappa, double eps, double y_max) {
  p.Xc = Xc_dev; p.N = N; p.idx_base = idx_base;
  p.d = h->d; p.n = h->n; p.npad = h->npad; p.n16 = (h->n + 15) / 16 * 16;
  p.X = h->dX; p.Xt = h->dXt; p.alpha = h->d_alpha_state; p.Z = h->dZ; p.ell = std::exp(h->theta);
  p.kind = h->kind; p.mg = h->mg; p.acq = acq; p.kappa = kappa; p.eps = eps; p.y_max = y_max;
  p.Ks = h->dKs; p.out_mean = p.out_std = p.out_acq = nullptr;
  p.planes = nullptr; p.plane_stride = 0;
  p.xrow = p.xnc = p.xmean = nullptr;         // set by launch_sweep_i8 once the rows are built
  p.blk_val = h->d_blk_val; p.blk_idx = h->d_blk_idx; p.clamped = h->d_clamped; p.dbg = h->sweep_dbg;
}

// 3-D map over digit planes: `cols` bytes per row, `rows` rows per plane, I8_S planes; box = 64 bytes x box_rows x all planes
static int make_map_i8(plx_bo* h, CUtensorMap* map, signed char* base, size_t cols, size_t rows, int box_rows,
                       int box_planes = I8_S) {
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(I8_S)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(cols) * rows};
  cuuint32_t box[3] = {I8_KC, static_cast<cuuint32_t>(box_rows), static_cast<cuuint32_t>(box_planes)};
  cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, I8_KC == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, PLX_ERR_CUDA, "cuTensorMapEncodeTiled (int8 planes) failed with code " + std::to_string(r));
  return PLX_OK;
}

// plane buffers, row scales and tensor maps of the int8-sliced sweeps for the current padded size
static int ensure_i8(plx_bo* h) {
  const long long rows = static_cast<long long>(h->num_sms) * 128 * 2;
  const size_t npad = h->npad;
  int r;
  if (h->i8_rows != rows || h->i8_npad != h->npad || !h->d_kplanes) {
    if ((r = ensure(h, h->d_kplanes, h->kplanes_cap, static_cast<size_t>(I8_S) * rows * npad)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_zplanes, h->zplanes_cap, static_cast<size_t>(I8_S) * npad * npad)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_zscale, h->zscale_cap, npad)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_i8mean, h->i8mean_cap, static_cast<size_t>(rows))) != PLX_OK) return r;
    if ((r = make_map_i8(h, &h->tmA8, h->d_kplanes, npad, rows, 128)) != PLX_OK) return r;
    if ((r = make_map_i8(h, &h->tmB8, h->d_zplanes, npad, npad, 64)) != PLX_OK) return r;
    if ((r = make_map_i8(h, &h->tmA8p, h->d_kplanes, npad, rows, 128, 1)) != PLX_OK) return r;
    const int max_groups = h->num_sms / I8C_R;
    if ((r = ensure(h, h->d_i8part, h->i8part_cap, static_cast<size_t>(max_groups) * 2 * I8C_R * 256)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_i8sync, h->i8sync_cap, static_cast<size_t>(max_groups) * 8)) != PLX_OK) return r;
    h->i8_rows = rows; h->i8_npad = h->npad; h->i8_state_valid = false;
  }
  // operands of the tensor-path builders (d <= 32): centre, centred rows, their norms
  if ((r = ensure(h, h->d_xmean, h->xmean_cap, 32)) != PLX_OK) return r;
  if ((r = ensure(h, h->d_xrow, h->xrow_cap, npad * XROW_LD)) != PLX_OK) return r;
  if ((r = ensure(h, h->d_xnc, h->xnc_cap, npad)) != PLX_OK) return r;
  return PLX_OK;
}

// Variant 2 (sweep_i8.cuh): chunks of candidates; per chunk the digit planes of K*, then the sliced contraction.
static int launch_sweep_i8(plx_bo* h, const SweepParams& p, cudaEvent_t e0, cudaEvent_t e1) {
  cudaStream_t st = h->stream;
  const long long rows = static_cast<long long>(h->num_sms) * 128 * 2;
  const size_t npad = h->npad;
  int r;
  if ((r = ensure_i8(h)) != PLX_OK) return r;
  if (!h->i8_state_valid) {
    // columns of the K* planes beyond the observations are never written by the slicing kernel: keep them zero
    CK(cudaMemsetAsync(h->d_kplanes, 0, static_cast<size_t>(I8_S) * rows * npad, st));
    zslice_kernel<<<h->npad, 128, 0, st>>>(h->dZ, h->npad, h->n, reinterpret_cast<int8_t*>(h->d_zplanes), h->d_zscale);
    LAUNCH_CHECK("zslice_kernel");
    if (h->d <= 32) {
      // operands of the tensor-path builders: centre, centred rows, their norms (same stream, before any sweep)
      obs_centre_kernel<<<32, 256, 0, st>>>(h->dX, h->n, h->d, std::exp(h->theta), h->d_xmean);
      LAUNCH_CHECK("obs_centre_kernel");
      obs_rows_kernel<<<(h->npad + 127) / 128, 128, 0, st>>>(h->dX, h->n, h->d, h->npad, std::exp(h->theta), h->d_xmean,
                                                            h->d_xrow, h->d_xnc);
      LAUNCH_CHECK("obs_rows_kernel");
    }
    h->i8_state_valid = true;
  }
  if (e0) CK(cudaEventRecord(e0, st));
  const int impl_i8 = h->sweep_impl == 7 ? 3 : h->sweep_impl;
  // squared distances on the FP64 tensor path (build_tile_mma): at most 32 dimensions; by default in the cooperative
  // form from 16 dimensions up, where it also moves the cross-over with the fused form from 1536 to 1024 observations
  // (n = 1024, d = 32: 7.97 against 8.46 ms per 400 000 candidates; n = 512, d = 16: the fused form still wins)
  const bool mma_dist = h->d <= 32 && (h->i8_distance == 1 || (h->i8_distance == 2 && h->d >= 16));
  if (impl_i8 == 5 || impl_i8 == 6 || (impl_i8 == 3 && (p.n >= 1536 || (mma_dist && p.n >= 1024)))) {
    // cooperative forms: groups of four CTAs share a candidate tile so that the planes in flight stay in L2.
    // 6: the four are two clusters of two that fetch the A stream once per pair (TMA multicast); measured equal to 5
    // (8.3 ms against 8.2 ms per 151 552 candidates: the operand stream is not what bounds the kernel), so the
    // default stays with the plain cooperative form
    const bool gen = p.kind == K_GEN;                  // its own instantiations; the paired form has none
    const bool paired = impl_i8 == 6 && !gen;
    const int max_groups = h->num_sms / I8C_R;
    const int groups = static_cast<int>(std::min<long long>((p.N + 127) / 128, max_groups));
    if ((r = ensure(h, h->d_i8part, h->i8part_cap, static_cast<size_t>(max_groups) * 2 * I8C_R * 256)) != PLX_OK) return r;
    if ((r = ensure(h, h->d_i8sync, h->i8sync_cap, static_cast<size_t>(max_groups) * 8)) != PLX_OK) return r;
    CK(cudaMemsetAsync(h->d_i8sync, 0, sizeof(unsigned) * max_groups * 8, st));
    SweepParams q = p;
    q.dbg = p.dbg & ~15; q.planes = h->d_kplanes; q.plane_stride = static_cast<size_t>(rows) * npad;
    if (mma_dist) { q.xrow = h->d_xrow; q.xnc = h->d_xnc; q.xmean = h->d_xmean; }
    I8Params ip;
    ip.N = p.N; ip.idx_base = p.idx_base; ip.n = p.n; ip.zscale = h->d_zscale; ip.mean = nullptr;
    ip.acq = p.acq; ip.kappa = p.kappa; ip.eps = p.eps; ip.y_max = p.y_max;
    ip.out_mean = p.out_mean; ip.out_std = p.out_std; ip.out_acq = p.out_acq;
    ip.blk_val = h->d_blk_val; ip.blk_idx = h->d_blk_idx; ip.a_from_smem = h->i8_a_from_smem; ip.zplanes_dbg = h->d_zplanes; ip.clamped = h->d_clamped;
    // cooperative launch: the runtime starts the grid only when all its CTAs can be resident at once, which the
    // spin-waits between the CTAs of a group rely on (two such sweeps from different handles then run one after
    // the other instead of starving each other)
    if (paired) {
      void* args[] = {&h->tmA8p, &h->tmB8, &q, &ip, &h->d_i8part, &h->d_i8sync};
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(groups * I8C_R); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = I8F_SMEM; cfg.stream = st;
      cudaLaunchAttribute attrs[2];
      attrs[0].id = cudaLaunchAttributeClusterDimension;
      attrs[0].val.clusterDim.x = 2; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
      attrs[1].id = cudaLaunchAttributeCooperative;
      attrs[1].val.cooperative = 1;
      cfg.attrs = attrs;
      cfg.numAttrs = h->coop_cluster_ok ? 2 : 1;
      cudaError_t le = cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(sweep_i8p_kernel), args);
      if (le != cudaSuccess && cfg.numAttrs == 2) {
        // a driver that refuses cooperative + cluster: the grid has one CTA per SM at most, so on an otherwise idle
        // GPU all of it is resident anyway
        (void)cudaGetLastError();
        h->coop_cluster_ok = 0;
        cfg.numAttrs = 1;
        le = cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(sweep_i8p_kernel), args);
      }
      CK(le);
      LAUNCH_CHECK("sweep_i8p_kernel");
      sweep_merge_kernel<<<1, 32, 0, st>>>(h->d_blk_val, h->d_blk_idx, groups * I8C_R, h->d_best_val, h->d_best_idx);
      LAUNCH_CHECK("sweep_merge_kernel");
      if (e1) CK(cudaEventRecord(e1, st));
      return PLX_OK;
    }
    {
      void* args[] = {&h->tmA8, &h->tmB8, &q, &ip, &h->d_i8part, &h->d_i8sync};
      CK(cudaLaunchCooperativeKernel(gen ? reinterpret_cast<const void*>(sweep_i8c_gen_kernel) : reinterpret_cast<const void*>(sweep_i8c_kernel),
                                     dim3(groups * I8C_R), dim3(512), args,
                                     I8F_SMEM, st));
    }
    LAUNCH_CHECK("sweep_i8c_kernel");
    sweep_merge_kernel<<<1, 32, 0, st>>>(h->d_blk_val, h->d_blk_idx, groups * I8C_R, h->d_best_val, h->d_best_idx);
    LAUNCH_CHECK("sweep_merge_kernel");
    if (e1) CK(cudaEventRecord(e1, st));
    return PLX_OK;
  }
  if (impl_i8 != 2) {
    // fused form: one persistent launch, the planes of the next tile are built while the current one is contracted
    const int grid = static_cast<int>(std::min<long long>((p.N + 127) / 128, h->num_sms));
    SweepParams q = p;
    q.dbg = p.dbg & ~15; q.planes = h->d_kplanes; q.plane_stride = static_cast<size_t>(rows) * npad;
    if (h->i8_distance == 1 && h->d <= 32) { q.xrow = h->d_xrow; q.xnc = h->d_xnc; q.xmean = h->d_xmean; }
    I8Params ip;
    ip.N = p.N; ip.idx_base = p.idx_base; ip.n = p.n; ip.zscale = h->d_zscale; ip.mean = nullptr;
    ip.acq = p.acq; ip.kappa = p.kappa; ip.eps = p.eps; ip.y_max = p.y_max;
    ip.out_mean = p.out_mean; ip.out_std = p.out_std; ip.out_acq = p.out_acq;
    ip.blk_val = h->d_blk_val; ip.blk_idx = h->d_blk_idx; ip.a_from_smem = h->i8_a_from_smem; ip.zplanes_dbg = h->d_zplanes; ip.clamped = h->d_clamped;
    if (p.kind == K_GEN) sweep_i8f_gen_kernel<<<grid, 512, I8F_SMEM, st>>>(h->tmA8, h->tmB8, q, ip, h->d_i8mean);
    else sweep_i8f_kernel<<<grid, 512, I8F_SMEM, st>>>(h->tmA8, h->tmB8, q, ip, h->d_i8mean);
    LAUNCH_CHECK("sweep_i8f_kernel");
    sweep_merge_kernel<<<1, 32, 0, st>>>(h->d_blk_val, h->d_blk_idx, grid, h->d_best_val, h->d_best_idx);
    LAUNCH_CHECK("sweep_merge_kernel");
    if (e1) CK(cudaEventRecord(e1, st));
    return PLX_OK;
  }
  for (long long base = 0; base < p.N; base += rows) {
    const long long cnt = std::min(rows, p.N - base);
    const int grid = static_cast<int>(std::min<long long>((cnt + 127) / 128, h->num_sms));
    SweepParams q = p;
    q.Xc = p.Xc + base * p.d; q.N = cnt; q.dbg = p.dbg & ~15;
    q.planes = h->d_kplanes; q.plane_stride = static_cast<size_t>(rows) * npad;
    kstar_slices_kernel<<<grid, 256, KSL_SMEM, st>>>(q, h->d_i8mean);
    LAUNCH_CHECK("kstar_slices_kernel");
    I8Params ip;
    ip.N = cnt; ip.idx_base = p.idx_base + base; ip.n = p.n; ip.zscale = h->d_zscale; ip.mean = h->d_i8mean;
    ip.acq = p.acq; ip.kappa = p.kappa; ip.eps = p.eps; ip.y_max = p.y_max;
    ip.out_mean = p.out_mean ? p.out_mean + base : nullptr;
    ip.out_std = p.out_std ? p.out_std + base : nullptr;
    ip.out_acq = p.out_acq ? p.out_acq + base : nullptr;
    ip.blk_val = h->d_blk_val; ip.blk_idx = h->d_blk_idx; ip.a_from_smem = h->i8_a_from_smem; ip.zplanes_dbg = h->d_zplanes; ip.clamped = h->d_clamped;
    sweep_i8_kernel<<<grid, 192, I8_SMEM, st>>>(h->tmA8, h->tmB8, ip);
    LAUNCH_CHECK("sweep_i8_kernel");
    sweep_merge_kernel<<<1, 32, 0, st>>>(h->d_blk_val, h->d_blk_idx, grid, h->d_best_val, h->d_best_idx);
    LAUNCH_CHECK("sweep_merge_kernel");
  }
  if (e1) CK(cudaEventRecord(e1, st));
  return PLX_OK;
}

static int launch_sweep(plx_bo* h, const SweepParams& p, cudaEvent_t e0, cudaEvent_t e1) {
  const long long ntiles = (p.N + 127) / 128;
  const int grid = static_cast<int>(std::min<long long>(ntiles, h->num_sms));
  if (grid <= 0) return PLX_OK;
  // 3 (default): the int8-sliced sweep from 256 observations up (measured break-even between 128 and 256; fused
  // form up to 1535 observations, cooperative form -- four CTAs per tile, planes L2-resident -- from 1536, or from
  // 1024 where the tensor-path builders apply: launch_sweep_i8),
  // FP64 DMMA below and for the general-nu Matern kernel, whose Bessel evaluation dominates anyway
  if ((h->sweep_impl == 3 || h->sweep_impl == 7) && p.n <= SS_NMAX) {
    // small observation sets: a thread per candidate against the state in shared memory (sweep_small.cuh)
    const int g = static_cast<int>(std::min<long long>((p.N + SS_THREADS - 1) / SS_THREADS, 4LL * h->num_sms));
    if (e0) CK(cudaEventRecord(e0, h->stream));
    sweep_small_kernel<<<g, SS_THREADS, sweep_small_smem(p.n, p.d), h->stream>>>(p);
    LAUNCH_CHECK("sweep_small_kernel");
    if (e1) CK(cudaEventRecord(e1, h->stream));
    sweep_merge_kernel<<<1, 32, 0, h->stream>>>(h->d_blk_val, h->d_blk_idx, g, h->d_best_val, h->d_best_idx);
    LAUNCH_CHECK("sweep_merge_kernel");
    return PLX_OK;
  }
  const int impl = h->sweep_impl == 7 ? 3 : h->sweep_impl;       // 7 = 3 with the small-set kernel forced on (it is on by default)
  const bool sliced = impl == 2 || impl >= 4 || (impl == 3 && p.n >= 256);
  if (sliced && (p.kind != K_GEN || p.mg.tab != nullptr)) return launch_sweep_i8(h, p, e0, e1);   // general nu: with its value table only
  if (e0) CK(cudaEventRecord(e0, h->stream));
  if (h->sweep_impl >= 1) {
    sweep_tma_kernel<<<grid, 384, SWEEP_SMEM, h->stream>>>(h->tmZ, h->tmK, p);
    LAUNCH_CHECK("sweep_tma_kernel");
  } else {
    sweep_simple_kernel<<<grid, 256, SWEEP_SIMPLE_SMEM, h->stream>>>(p);
    LAUNCH_CHECK("sweep_simple_kernel");
  }
  if (e1) CK(cudaEventRecord(e1, h->stream));
  sweep_merge_kernel<<<1, 32, 0, h->stream>>>(h->d_blk_val, h->d_blk_idx, grid, h->d_best_val, h->d_best_idx);
  LAUNCH_CHECK("sweep_merge_kernel");
  return PLX_OK;
}

__global__ void sweep_reset_kernel(double* rec) {
  rec[0] = -INFINITY;
  reinterpret_cast<long long*>(rec)[1] = 0x7fffffffffffffffLL;
  reinterpret_cast<unsigned long long*>(rec)[2] = 0ULL;
}
// rec[3 .. 3 + d) <- the candidate row the record's index names (all candidates of the sweep in one device array)
__global__ void sweep_gather_row_kernel(double* rec, const double* __restrict__ Xc, long long idx_base, long long N, int d) {
  const long long i = reinterpret_cast<const long long*>(rec)[1] - idx_base;
  if (i < 0 || i >= N) return;
  for (int q = threadIdx.x; q < d; q += blockDim.x) rec[3 + q] = Xc[i * d + q];
}

static int reset_best(plx_bo* h) {
  sweep_reset_kernel<<<1, 1, 0, h->stream>>>(h->d_rec);
  LAUNCH_CHECK("sweep_reset_kernel");
  return PLX_OK;
}

// the record (and `extra` doubles after its three scalars) to the host: one copy, one synchronisation
static int fetch_record(plx_bo* h, int extra, double* best_value, long long* best_index) {
  CK(cudaMemcpyAsync(h->h_pin, h->d_rec, sizeof(double) * (3 + extra), cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  *best_value = h->h_pin[0];
  *best_index = reinterpret_cast<const long long*>(h->h_pin)[1];
  h->last_clamped = reinterpret_cast<const long long*>(h->h_pin)[2];
  return PLX_OK;
}

static int need_events(plx_bo* h, size_t count) {
  while (h->ev_pool.size() < count) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    h->ev_pool.push_back(e);
  }
  return PLX_OK;
}

static int check_sweep_args(plx_bo* h, const void* Xc, long long N, int acq, double* bv, long long* bi) {
  if (!h) return PLX_ERR_INVALID;
  if (N < 0 || (N > 0 && !Xc) || acq < 0 || acq > 2 || !bv || !bi) return fail(h, PLX_ERR_INVALID, "acq_sweep: bad arguments");
  if (!h->have_state) return fail(h, PLX_ERR_STATE, "acq_sweep: no posterior state (call plx_gp_fit or plx_gp_set_theta)");
  return PLX_OK;
}

extern "C" int plx_acq_sweep_dev(plx_bo_handle h, const double* Xc_dev, long long N, long long idx_base, int acq,
                                 double kappa, double eps, double y_max, double* best_value,
                                 long long* best_index, double* best_x) {
  int r = check_sweep_args(h, Xc_dev, N, acq, best_value, best_index);
  if (r != PLX_OK) return r;
  CK(cudaSetDevice(h->device));
  const long long l0 = h->launches;
  if ((r = need_events(h, 2)) != PLX_OK) return r;
  if ((r = reset_best(h)) != PLX_OK) return r;
  SweepParams p;
  fill_params(h, p, Xc_dev, N, idx_base, acq, kappa, eps, y_max);
  if ((r = launch_sweep(h, p, h->ev_pool[0], h->ev_pool[1])) != PLX_OK) return r;
  const bool row = best_x != nullptr && N > 0 && static_cast<size_t>(3 + h->d) <= h->rec_cap;
  if (row) {
    sweep_gather_row_kernel<<<1, 128, 0, h->stream>>>(h->d_rec, Xc_dev, idx_base, N, h->d);
    LAUNCH_CHECK("sweep_gather_row_kernel");
  }
  if ((r = fetch_record(h, row ? h->d : 0, best_value, best_index)) != PLX_OK) return r;
  float ms = 0.f;
  if (N > 0) CK(cudaEventElapsedTime(&ms, h->ev_pool[0], h->ev_pool[1]));
  h->last_ms = ms; h->last_launches = h->launches - l0;
  if (N == 0) { *best_index = -1; *best_value = NAN; return PLX_OK; }
  if (row) std::memcpy(best_x, h->h_pin + 3, sizeof(double) * h->d);
  else if (best_x) CK(cudaMemcpy(best_x, Xc_dev + (*best_index - idx_base) * h->d, sizeof(double) * h->d, cudaMemcpyDeviceToHost));
  return PLX_OK;
}

extern "C" int plx_acq_sweep(plx_bo_handle h, const double* Xc, long long N, long long idx_base, int acq,
                             double kappa, double eps, double y_max, double* best_value, long long* best_index,
                             double* best_x, double* mean, double* sd, double* acq_values) {
  int r = check_sweep_args(h, Xc, N, acq, best_value, best_index);
  if (r != PLX_OK) return r;
  CK(cudaSetDevice(h->device));
  const long long l0 = h->launches;
  if (N == 0) { *best_index = -1; *best_value = NAN; h->last_ms = 0; h->last_launches = 0; h->last_clamped = 0; return PLX_OK; }
  const int d = h->d;
  // chunk: a whole number of waves of 128-candidate tiles, at most ~64 MiB of candidates
  const long long wave = 128LL * h->num_sms;
  long long k = (64LL << 20) / (8LL * d * wave);
  if (k < 1) k = 1;
  const long long chunk = std::min<long long>(wave * k, (N + 127) / 128 * 128);
  const long long nchunks = (N + chunk - 1) / chunk;
  if (h->cand_cap < static_cast<size_t>(chunk) * d) {
    for (int b = 0; b < 2; ++b) { if (h->dcand[b]) CK(cudaFree(h->dcand[b])); h->dcand[b] = nullptr; }
    h->cand_cap = 0;
    for (int b = 0; b < 2; ++b) CK(cudaMalloc(&h->dcand[b], sizeof(double) * chunk * d));
    h->cand_cap = static_cast<size_t>(chunk) * d;
  }
  const bool dump = mean || sd || acq_values;
  if (dump && h->out_cap < static_cast<size_t>(chunk)) {
    for (int b = 0; b < 3; ++b) { if (h->d_out[b]) CK(cudaFree(h->d_out[b])); h->d_out[b] = nullptr; }
    h->out_cap = 0;
    for (int b = 0; b < 3; ++b) CK(cudaMalloc(&h->d_out[b], sizeof(double) * chunk));
    h->out_cap = static_cast<size_t>(chunk);
  }
  if ((r = need_events(h, static_cast<size_t>(2 * nchunks))) != PLX_OK) return r;
  if ((r = reset_best(h)) != PLX_OK) return r;
  for (long long c = 0; c < nchunks; ++c) {
    const int b = static_cast<int>(c & 1);
    const long long off = c * chunk, cn = std::min(chunk, N - off);
    CK(cudaStreamWaitEvent(h->copy_stream, h->ev_done[b], 0));          // kernel that last read dcand[b]
    CK(cudaMemcpyAsync(h->dcand[b], Xc + off * d, sizeof(double) * cn * d, cudaMemcpyHostToDevice, h->copy_stream));
    CK(cudaEventRecord(h->ev_copy[b], h->copy_stream));
    CK(cudaStreamWaitEvent(h->stream, h->ev_copy[b], 0));
    SweepParams p;
    fill_params(h, p, h->dcand[b], cn, idx_base + off, acq, kappa, eps, y_max);
    if (dump) { p.out_mean = h->d_out[0]; p.out_std = h->d_out[1]; p.out_acq = h->d_out[2]; }
    if ((r = launch_sweep(h, p, h->ev_pool[2 * c], h->ev_pool[2 * c + 1])) != PLX_OK) return r;
    CK(cudaEventRecord(h->ev_done[b], h->stream));
    if (dump) {
      if (mean) CK(cudaMemcpyAsync(mean + off, h->d_out[0], sizeof(double) * cn, cudaMemcpyDeviceToHost, h->stream));
      if (sd) CK(cudaMemcpyAsync(sd + off, h->d_out[1], sizeof(double) * cn, cudaMemcpyDeviceToHost, h->stream));
      if (acq_values) CK(cudaMemcpyAsync(acq_values + off, h->d_out[2], sizeof(double) * cn, cudaMemcpyDeviceToHost, h->stream));
    }
  }
  if ((r = fetch_record(h, 0, best_value, best_index)) != PLX_OK) return r;
  double total = 0.0;
  for (long long c = 0; c < nchunks; ++c) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->ev_pool[2 * c], h->ev_pool[2 * c + 1]));
    total += ms;
  }
  h->last_ms = total; h->last_launches = h->launches - l0;
  if (best_x) std::memcpy(best_x, Xc + (*best_index - idx_base) * d, sizeof(double) * d);
  return PLX_OK;
}

extern "C" int plx_candidates_set(plx_bo_handle h, const double* Xc, long long N, int d) {
  if (!h || N < 0 || d <= 0 || (N > 0 && !Xc)) return h ? fail(h, PLX_ERR_INVALID, "candidates_set: bad arguments") : PLX_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  h->resident_n = 0;
  int r;
  if ((r = ensure(h, h->d_resident, h->resident_cap, static_cast<size_t>(std::max<long long>(N, 1)) * d)) != PLX_OK) return r;
  if (N > 0) CK(cudaMemcpyAsync(h->d_resident, Xc, sizeof(double) * N * d, cudaMemcpyHostToDevice, h->stream));
  CK(sync_stream(h));
  h->resident_n = N; h->resident_first = 0; h->resident_d = d;
  return PLX_OK;
}

extern "C" int plx_candidates_draw_mt19937_range(plx_bo_handle h, const unsigned* key, int pos, const double* low,
                                                 const double* range, long long N, int d, long long first,
                                                 long long count, unsigned* key_out, int* pos_out) {
  if (!h || !key || !low || !range || !key_out || !pos_out || N < 0 || d <= 0 || pos < 0 || pos > 624 || first < 0 ||
      count < 0 || first + count > N)
    return h ? fail(h, PLX_ERR_INVALID, "candidates_draw_mt19937: bad arguments") : PLX_ERR_INVALID;
  if (pos & 1) return fail(h, PLX_ERR_UNSUPPORTED, "candidates_draw_mt19937: odd stream position (draw on the host instead)");
  if (d > 2048) return fail(h, PLX_ERR_UNSUPPORTED, "candidates_draw_mt19937: d exceeds 2048");
  if (static_cast<double>(N) * d * 2.0 + 1248.0 >= std::ldexp(1.0, MT_JUMP_KMAX + 1))
    return fail(h, PLX_ERR_UNSUPPORTED, "candidates_draw_mt19937: N * d beyond the jump table (2^40 values)");
  CK(cudaSetDevice(h->device));
  h->resident_n = 0; h->resident_first = 0;
  int r;
  if ((r = ensure(h, h->d_resident, h->resident_cap, static_cast<size_t>(std::max<long long>(count, 1)) * d)) != PLX_OK) return r;
  // scratch layout (doubles): [0, 2d) low | range, then 1248 words of generator state (in | out), then the position
  if ((r = ensure(h, h->d_misc, h->misc_cap, static_cast<size_t>(2 * d) + 640)) != PLX_OK) return r;
  double* d_lr = h->d_misc;
  unsigned* d_key = reinterpret_cast<unsigned*>(h->d_misc + 2 * d);
  int* d_pos = reinterpret_cast<int*>(d_key + 1248);
  CK(cudaMemcpyAsync(d_key, key, sizeof(unsigned) * 624, cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(d_lr, low, sizeof(double) * d, cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(d_lr + d, range, sizeof(double) * d, cudaMemcpyHostToDevice, h->stream));
  // One CTA per chunk of rows, two CTAs per SM.  A chunk pays up to ~1 ms for its jump, so short draws (below ~4 M
  // stream words per extra CTA) stay on fewer CTAs; the first chunk of a draw that starts at row 0 needs no jump.
  const long long words = 2 * count * d;
  long long chunks = std::min<long long>(2LL * h->num_sms, std::max<long long>(1, words >> 22));
  chunks = std::max<long long>(1, std::min(chunks, count));
  MtDrawParams mp;
  mp.key = d_key; mp.pos = pos; mp.low = d_lr; mp.range = d_lr + d; mp.d = d;
  mp.first_row = first; mp.rows = count; mp.rows_per_cta = (count + chunks - 1) / chunks;
  mp.out = h->d_resident; mp.total_words = pos + 2 * N * d;
  mp.key_out = d_key + 624; mp.pos_out = d_pos;
  const size_t smem = sizeof(unsigned) * MT_SEQ + sizeof(double) * 2 * d;
  mt19937_draw_kernel<<<static_cast<int>(chunks) + 1, MT_THREADS, smem, h->stream>>>(mp);
  LAUNCH_CHECK("mt19937_draw_kernel");
  CK(cudaMemcpyAsync(key_out, d_key + 624, sizeof(unsigned) * 624, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(pos_out, d_pos, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  h->resident_n = count; h->resident_first = first; h->resident_d = d;
  return PLX_OK;
}

extern "C" int plx_candidates_draw_mt19937(plx_bo_handle h, const unsigned* key, int pos, const double* low,
                                           const double* range, long long N, int d, unsigned* key_out, int* pos_out) {
  return plx_candidates_draw_mt19937_range(h, key, pos, low, range, N, d, 0, N, key_out, pos_out);
}

extern "C" int plx_candidates_get(plx_bo_handle h, double* out, long long first, long long count) {
  if (!h || !out || first < 0 || count < 0) return h ? fail(h, PLX_ERR_INVALID, "candidates_get: bad arguments") : PLX_ERR_INVALID;
  if (first < h->resident_first || first + count > h->resident_first + h->resident_n)
    return fail(h, PLX_ERR_INVALID, "candidates_get: rows outside the resident block");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(out, h->d_resident + (first - h->resident_first) * h->resident_d, sizeof(double) * count * h->resident_d,
                     cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  return PLX_OK;
}

extern "C" int plx_acq_sweep_resident_range(plx_bo_handle h, long long first, long long count, int acq, double kappa,
                                            double eps, double y_max, double* best_value, long long* best_index,
                                            double* best_x) {
  if (!h) return PLX_ERR_INVALID;
  if (h->resident_d != h->d || (h->resident_n > 0 && !h->d_resident))
    return fail(h, PLX_ERR_STATE, "acq_sweep_resident_range: no resident candidates of this dimension");
  if (first < h->resident_first || count < 0 || first + count > h->resident_first + h->resident_n)
    return fail(h, PLX_ERR_INVALID, "acq_sweep_resident_range: rows outside the resident block");
  return plx_acq_sweep_dev(h, h->d_resident + (first - h->resident_first) * h->d, count, first, acq, kappa, eps, y_max, best_value,
                           best_index, best_x);
}

extern "C" int plx_acq_sweep_resident(plx_bo_handle h, long long idx_base, int acq, double kappa, double eps,
                                      double y_max, double* best_value, long long* best_index, double* best_x) {
  if (!h) return PLX_ERR_INVALID;
  if (h->resident_d != h->d || (h->resident_n > 0 && !h->d_resident))
    return fail(h, PLX_ERR_STATE, "acq_sweep_resident: no resident candidates of this dimension (call plx_candidates_set)");
  return plx_acq_sweep_dev(h, h->d_resident, h->resident_n, idx_base + h->resident_first, acq, kappa, eps, y_max, best_value, best_index, best_x);
}

extern "C" int plx_kernel_values(plx_bo_handle h, int kernel, double nu, const double* r2, int n, double* out) {
  if (!h || !r2 || !out || n <= 0) return h ? fail(h, PLX_ERR_INVALID, "kernel_values: bad arguments") : PLX_ERR_INVALID;
  int kind;
  if (kernel_kind(kernel, nu, &kind) != PLX_OK || kind == K_GEN) return fail(h, PLX_ERR_UNSUPPORTED, "kernel_values: closed-form kernels only");
  CK(cudaSetDevice(h->device));
  int r;
  if ((r = ensure(h, h->d_misc, h->misc_cap, static_cast<size_t>(2) * n)) != PLX_OK) return r;
  double *d_in = h->d_misc, *d_out = h->d_misc + n;
  CK(cudaMemcpyAsync(d_in, r2, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
  kern_value_probe_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(d_in, n, kind, d_out);
  LAUNCH_CHECK("kern_value_probe_kernel");
  CK(cudaMemcpyAsync(out, d_out, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  return PLX_OK;
}

extern "C" int plx_measure_i8_peak(plx_bo_handle h, double* tops) {
  if (!h || !tops) return h ? fail(h, PLX_ERR_INVALID, "measure_i8_peak: bad arguments") : PLX_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  int r;
  if ((r = need_events(h, 2)) != PLX_OK) return r;
  const int repeats = 20000;
  i8_peak_kernel<<<h->num_sms, 128, 0, h->stream>>>(1000);
  LAUNCH_CHECK("i8_peak_kernel");
  CK(cudaEventRecord(h->ev_pool[0], h->stream));
  i8_peak_kernel<<<h->num_sms, 128, 0, h->stream>>>(repeats);
  LAUNCH_CHECK("i8_peak_kernel");
  CK(cudaEventRecord(h->ev_pool[1], h->stream));
  CK(sync_stream(h));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, h->ev_pool[0], h->ev_pool[1]));
  *tops = 2.0 * 128 * 128 * 128 * static_cast<double>(repeats) * h->num_sms / (ms * 1e-3) / 1e12;
  return PLX_OK;
}

extern "C" int plx_last_sweep_clamped(plx_bo_handle h, long long* count) {
  if (!h || !count) return PLX_ERR_INVALID;
  *count = h->last_clamped;
  return PLX_OK;
}

extern "C" int plx_last_sweep_stats(plx_bo_handle h, double* kernel_ms, long long* launches) {
  if (!h) return PLX_ERR_INVALID;
  if (kernel_ms) *kernel_ms = h->last_ms;
  if (launches) *launches = h->last_launches;
  return PLX_OK;
}

extern "C" int plx_lbfgsb1d_minimize(plx_fg_callback fg, void* user, double x0, double lo, double hi, double* x,
                                     double* f, int* n_evals, int* n_iters, int* status) {
  if (!fg || !(lo < hi)) return PLX_ERR_INVALID;
  Lbfgsb1D m;
  m.lo = lo; m.hi = hi;
  m.start(x0);
  while (!m.done) {
    double fv = 0.0, gv = 0.0;
    fg(m.x_eval, &fv, &gv, user);
    m.feed(fv, gv);
  }
  if (x) *x = m.x;
  if (f) *f = m.f;
  if (n_evals) *n_evals = m.nfev;
  if (n_iters) *n_iters = m.iter;
  if (status) *status = m.status;
  return PLX_OK;
}

// ---------------------------------------------------------------------------------------------
// refinement (acquisition_function.py:95-111) and point-wise acquisition value + gradient
// ---------------------------------------------------------------------------------------------
static int ensure_refine(plx_bo* h, int S) {
  if (S <= h->rf_cap_s && h->npad == h->rf_cap_npad && h->d == h->rf_cap_d) return PLX_OK;
  void* old[] = {h->rf_kv, h->rf_kp, h->rf_v, h->rf_w, h->rf_wpart, h->rf_vec, h->rf_bounds, h->rf_val, h->rf_grad, h->rf_sc, h->rf_active};
  for (void* p : old) if (p) CK(cudaFree(p));
  h->rf_kv = h->rf_kp = h->rf_v = h->rf_w = h->rf_wpart = h->rf_vec = h->rf_bounds = h->rf_val = h->rf_grad = nullptr;
  h->rf_sc = nullptr; h->rf_active = nullptr; h->rf_cap_s = 0;
  const size_t sn = static_cast<size_t>(S) * h->npad;
  CK(cudaMalloc(&h->rf_kv, sizeof(double) * sn));
  CK(cudaMalloc(&h->rf_kp, sizeof(double) * sn));
  CK(cudaMalloc(&h->rf_v, sizeof(double) * sn));
  CK(cudaMalloc(&h->rf_w, sizeof(double) * sn));
  CK(cudaMalloc(&h->rf_wpart, sizeof(double) * sn * ((h->npad + RF_WROWS - 1) / RF_WROWS)));
  CK(cudaMalloc(&h->rf_vec, sizeof(double) * rf_vec_stride(h->d) * S));
  CK(cudaMalloc(&h->rf_bounds, sizeof(double) * 2 * h->d));
  CK(cudaMalloc(&h->rf_val, sizeof(double) * S));
  CK(cudaMalloc(&h->rf_grad, sizeof(double) * S * h->d));
  CK(cudaMalloc(&h->rf_sc, sizeof(RefineScalars) * S));
  CK(cudaMalloc(&h->rf_active, sizeof(int)));
  h->rf_cap_s = S; h->rf_cap_npad = h->npad; h->rf_cap_d = h->d;
  return PLX_OK;
}

static void fill_refine(plx_bo* h, RefineParams& p, int S, int acq, double kappa, double eps, double y_max, int max_iter) {
  p.S = S; p.d = h->d; p.n = h->n; p.npad = h->npad; p.kind = h->kind; p.mg = h->mg; p.acq = acq; p.max_iter = max_iter;
  p.ell = std::exp(h->theta); p.kappa = kappa; p.eps = eps; p.y_max = y_max;
  p.X = h->dX; p.alpha = h->d_alpha_state; p.Z = h->dZ; p.bounds = h->rf_bounds;
  p.kv = h->rf_kv; p.kp = h->rf_kp; p.v = h->rf_v; p.w = h->rf_w; p.wpart = nullptr;
  p.vec = h->rf_vec; p.vec_stride = rf_vec_stride(h->d); p.sc = h->rf_sc;
  p.val_out = nullptr; p.grad_out = nullptr; p.active = h->rf_active;
}

// one round: everything a seed needs to evaluate the acquisition and its gradient at its trial point
static int launch_refine_round(plx_bo* h, const RefineParams& p) {
  cudaStream_t st = h->stream;
  refine_kstar_kernel<<<dim3((p.npad + 127) / 128, p.S), 128, 0, st>>>(p);
  LAUNCH_CHECK("refine_kstar_kernel");
  refine_v_kernel<<<(p.n + 7) / 8, 256, 0, st>>>(p);
  LAUNCH_CHECK("refine_v_kernel");
  RefineParams q = p;
  q.wpart = h->rf_wpart;
  refine_w_kernel<<<dim3((p.n + RF_WROWS - 1) / RF_WROWS, (p.n + 127) / 128), 128, 0, st>>>(q);
  LAUNCH_CHECK("refine_w_kernel");
  refine_step_kernel<<<p.S, 512, sizeof(double) * 2 * p.d, st>>>(q);
  LAUNCH_CHECK("refine_step_kernel");
  return PLX_OK;
}

static int upload_points(plx_bo* h, const double* pts, int S, const double* bounds) {
  // each seed's block starts with its current point x
  const size_t stride = rf_vec_stride(h->d);
  CK(cudaMemcpy2DAsync(h->rf_vec, stride * 8, pts, static_cast<size_t>(h->d) * 8, static_cast<size_t>(h->d) * 8, S,
                       cudaMemcpyHostToDevice, h->stream));
  if (bounds) CK(cudaMemcpyAsync(h->rf_bounds, bounds, sizeof(double) * 2 * h->d, cudaMemcpyHostToDevice, h->stream));
  return PLX_OK;
}

extern "C" int plx_acq_value_grad(plx_bo_handle h, const double* Xp, int S, int acq, double kappa, double eps,
                                  double y_max, double* values, double* grads) {
  if (!h || !Xp || S <= 0 || acq < 0 || acq > 2 || !values || !grads) return h ? fail(h, PLX_ERR_INVALID, "acq_value_grad: bad arguments") : PLX_ERR_INVALID;
  if (!h->have_state) return fail(h, PLX_ERR_STATE, "acq_value_grad: no posterior state");
  if (h->d > RF_DMAX) return fail(h, PLX_ERR_UNSUPPORTED, "acq_value_grad: d exceeds 2048");
  CK(cudaSetDevice(h->device));
  int r;
  if ((r = ensure_refine(h, S)) != PLX_OK) return r;
  if ((r = upload_points(h, Xp, S, nullptr)) != PLX_OK) return r;
  RefineParams p;
  fill_refine(h, p, S, acq, kappa, eps, y_max, 0);
  p.val_out = h->rf_val; p.grad_out = h->rf_grad;
  if ((r = launch_refine_round(h, p)) != PLX_OK) return r;
  CK(cudaMemcpyAsync(values, h->rf_val, sizeof(double) * S, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(grads, h->rf_grad, sizeof(double) * S * h->d, cudaMemcpyDeviceToHost, h->stream));
  CK(sync_stream(h));
  return PLX_OK;
}

extern "C" int plx_acq_refine(plx_bo_handle h, const double* seeds, int S, const double* bounds, int acq,
                              double kappa, double eps, double y_max, int max_iter, double* x_out, double* acq_out,
                              int* success_out, int* nfev_out) {
  if (!h || S < 0 || (S > 0 && (!seeds || !x_out || !acq_out || !success_out)) || !bounds || acq < 0 || acq > 2)
    return h ? fail(h, PLX_ERR_INVALID, "acq_refine: bad arguments") : PLX_ERR_INVALID;
  if (!h->have_state) return fail(h, PLX_ERR_STATE, "acq_refine: no posterior state");
  if (h->d > RF_DMAX) return fail(h, PLX_ERR_UNSUPPORTED, "acq_refine: d exceeds 2048");
  h->last_refine_rounds = 0;
  if (S == 0) return PLX_OK;
  CK(cudaSetDevice(h->device));
  if (max_iter <= 0) max_iter = 15000;                 // scipy's maxiter
  const int d = h->d;
  for (int q = 0; q < d; ++q)
    if (!(bounds[2 * q] <= bounds[2 * q + 1])) return fail(h, PLX_ERR_INVALID, "acq_refine: lower bound above upper bound");
  int r;
  if ((r = ensure_refine(h, S)) != PLX_OK) return r;
  // scipy clips x0 into the box before the first evaluation
  std::vector<double> x0(static_cast<size_t>(S) * d);
  for (int s = 0; s < S; ++s)
    for (int q = 0; q < d; ++q)
      x0[static_cast<size_t>(s) * d + q] = std::min(std::max(seeds[static_cast<size_t>(s) * d + q], bounds[2 * q]), bounds[2 * q + 1]);
  if ((r = upload_points(h, x0.data(), S, bounds)) != PLX_OK) return r;
  CK(cudaMemsetAsync(h->rf_sc, 0, sizeof(RefineScalars) * S, h->stream));
  CK(cudaMemcpyAsync(h->rf_active, &S, sizeof(int), cudaMemcpyHostToDevice, h->stream));
  RefineParams p;
  fill_refine(h, p, S, acq, kappa, eps, y_max, max_iter);
  const long long max_rounds = 15000;                  // scipy's maxfun
  int active = S;
  long long rounds = 0;
  if (h->refine_persist && h->n <= RF_PERSIST_NMAX) {
    // small observation sets: every seed start to finish in one launch (refine.cuh)
    refine_persist_kernel<<<S, 128, sizeof(double) * rf_persist_smem_doubles(p.n, p.npad, p.d), h->stream>>>(
        p, static_cast<int>(max_rounds));
    LAUNCH_CHECK("refine_persist_kernel");
    active = 0;
  }
  while (active > 0 && rounds < max_rounds) {
    for (int k = 0; k < 4 && rounds < max_rounds; ++k, ++rounds)
      if ((r = launch_refine_round(h, p)) != PLX_OK) return r;
    int* pinned_active = reinterpret_cast<int*>(h->h_pin);
    CK(cudaMemcpyAsync(pinned_active, h->rf_active, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
    active = *pinned_active;
  }
  h->last_refine_rounds = rounds;
  std::vector<RefineScalars> sc(S);
  const size_t stride = rf_vec_stride(d);
  const size_t sc_bytes = (sizeof(RefineScalars) * S + 7) / 8 * 8, x_bytes = sizeof(double) * S * d;
  if (sc_bytes + x_bytes <= h->pin_cap * sizeof(double)) {
    // both results into pinned memory: the copies queue behind the kernels and one synchronisation collects them
    unsigned char* pin = reinterpret_cast<unsigned char*>(h->h_pin);
    CK(cudaMemcpyAsync(pin, h->rf_sc, sizeof(RefineScalars) * S, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpy2DAsync(pin + sc_bytes, static_cast<size_t>(d) * 8, h->rf_vec, stride * 8, static_cast<size_t>(d) * 8, S,
                         cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
    std::memcpy(sc.data(), pin, sizeof(RefineScalars) * S);
    std::memcpy(x_out, pin + sc_bytes, x_bytes);
  } else {
    CK(cudaMemcpyAsync(sc.data(), h->rf_sc, sizeof(RefineScalars) * S, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpy2DAsync(x_out, static_cast<size_t>(d) * 8, h->rf_vec, stride * 8, static_cast<size_t>(d) * 8, S,
                         cudaMemcpyDeviceToHost, h->stream));
    CK(sync_stream(h));
  }
  for (int s = 0; s < S; ++s) {
    acq_out[s] = -sc[s].f;
    success_out[s] = (sc[s].phase == 2 && sc[s].status == 0) ? 1 : 0;
    if (nfev_out) nfev_out[s] = sc[s].nfev;
  }
  return PLX_OK;
}

// ---------------------------------------------------------------------------------------------
// plx_bo_reserve: allocate every workspace a problem of the given shape will need, so that the first
// plx_gp_set_data / plx_gp_fit / plx_acq_sweep* / plx_acq_refine of a group does not pay for cudaMalloc (a
// device-wide synchronising call) inside its latency.  Workspaces are keyed on the padded observation count
// (n rounded up to 128): a later problem with the same padded size, at most `runs` optimiser runs, `seeds`
// refinement seeds and `candidates` resident rows allocates nothing.  The kernels themselves are loaded when
// the handle is created (every kernel of the library is touched there).
// ---------------------------------------------------------------------------------------------
static int ensure_refine(plx_bo* h, int S);

extern "C" int plx_bo_reserve(plx_bo_handle h, int n, int d, int runs, int seeds, long long candidates) {
  if (!h || n <= 0 || d <= 0 || runs < 0 || seeds < 0 || candidates < 0)
    return h ? fail(h, PLX_ERR_INVALID, "bo_reserve: bad arguments") : PLX_ERR_INVALID;
  if (h->have_data) return fail(h, PLX_ERR_STATE, "bo_reserve: call it before plx_gp_set_data");
  CK(cudaSetDevice(h->device));
  const int n0 = h->n, d0 = h->d, npad0 = h->npad, nb0 = h->nb;
  h->n = n; h->d = d; h->npad = (n + 127) / 128 * 128; h->nb = h->npad / 128;
  int r = PLX_OK;
  const size_t mat = static_cast<size_t>(h->npad) * h->npad;
  do {
    if ((r = ensure(h, h->dX, h->x_cap, static_cast<size_t>(h->npad) * d)) != PLX_OK) break;
    if ((r = ensure(h, h->dy, h->y_cap, static_cast<size_t>(h->npad))) != PLX_OK) break;
    if ((r = ensure_batch(h, std::max(runs, 1))) != PLX_OK) break;
    if ((r = ensure(h, h->d_xs, h->xs_cap, static_cast<size_t>(std::max(runs, 1)) * n * d)) != PLX_OK) break;
    if ((r = ensure(h, h->d_fs, h->fs_cap, static_cast<size_t>(5) * std::max(runs, 2))) != PLX_OK) break;
    if ((r = ensure(h, h->dZ, h->z_cap, mat)) != PLX_OK) break;
    if ((r = ensure(h, h->d_alpha_state, h->as_cap, static_cast<size_t>(h->npad))) != PLX_OK) break;
    if ((r = ensure(h, h->dKs, h->ks_cap, static_cast<size_t>(h->num_sms) * 128 * h->npad)) != PLX_OK) break;
    if ((r = ensure(h, h->dXt, h->xt_cap, static_cast<size_t>(d) * h->npad)) != PLX_OK) break;
    if ((r = ensure(h, h->d_mgtab, h->mgtab_cap, static_cast<size_t>(2) * MG_NINT * MG_ROW)) != PLX_OK) break;   // general-nu tables (11 KB)
    if (n >= 256 && (r = ensure_i8(h)) != PLX_OK) break;
    if (seeds > 0 && (r = ensure_refine(h, seeds)) != PLX_OK) break;
    if (candidates > 0) {
      if ((r = ensure(h, h->d_resident, h->resident_cap, static_cast<size_t>(candidates) * d)) != PLX_OK) break;
      if ((r = ensure(h, h->d_misc, h->misc_cap, static_cast<size_t>(2 * d) + 640)) != PLX_OK) break;
    }
    if ((r = need_events(h, 2)) != PLX_OK) break;
  } while (false);
  h->n = n0; h->d = d0; h->npad = npad0; h->nb = nb0;
  if (r == PLX_OK) CK(cudaDeviceSynchronize());
  return r;
}

// ---------------------------------------------------------------------------------------------
// plx_acq_sweep_noalloc: the warm-up sweep for callers that must not allocate device memory.
//
// Who needs it: a host application that owns the whole GPU memory budget itself -- a serving or
// training process that has handed every byte to its own pool, or one running under a fixed
// MPS / cgroup memory limit -- and therefore forbids cudaMalloc inside library calls, and code in
// a latency-critical section where an allocation (a device-wide synchronising call) is not
// acceptable.  Such a caller passes every input and output as device pointers it already owns;
// the only thing the sweep still needs is scratch (the K* slabs, a zero-padded copy of L^-1 and
// the per-CTA bests).  That scratch lives in the file-scope array below: about 8.9 MB that every
// GPU which loads this library holds until the process exits.
//
// Limits: n_obs <= 256 (PLX_NOALLOC_MAX_OBS); any d, any number of candidates.  Larger problems
// are refused with PLX_ERR_INVALID; use the handle API for them.
//
// One-time set-up happens inside the FIRST call on each device (symbol address look-up, the
// shared-memory attribute of the kernel, and the driver's lazy loading of the kernels), so make
// that first call before any stream capture.  Later calls issue three kernel launches and nothing
// else.  All calls share the one array and it keeps the previous call's contents; every region a
// call reads is rewritten by that call first, but two calls that overlap on the GPU (different
// streams, or different host threads) race on it -- serialise them.  The one-time set-up itself takes no
// lock either: two host threads making their FIRST call on a device at the same moment both run it and
// store the same address (harmless, but make the first call from one thread).  The call leaves the
// caller's current CUDA device as it found it.  A failure is described by plx_last_error(NULL).
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int NOALLOC_MAX_OBS = 256;
constexpr int NOALLOC_GRID = 32;                                   // CTAs (one 128 x npad slab each)
constexpr size_t NOALLOC_Z = static_cast<size_t>(NOALLOC_MAX_OBS) * NOALLOC_MAX_OBS;
constexpr size_t NOALLOC_KS = static_cast<size_t>(NOALLOC_GRID) * 128 * NOALLOC_MAX_OBS;
constexpr size_t NOALLOC_DOUBLES = NOALLOC_Z + NOALLOC_KS + 2 * NOALLOC_GRID;
constexpr int NOALLOC_MAX_DEVICES = 64;
}  // namespace

static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];

// In host code the name g_noalloc_scratch is only a handle for the runtime, not an address the
// GPU understands: the real device address is looked up once per device and cached here.
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// zero-padded copy of the caller's L^-1 (n x n, leading dimension ldz) into the padded layout the
// sweep reads, plus the start value of the running best
__global__ void noalloc_pack_kernel(const double* __restrict__ Linv, int n, int ldz, int npad, double* Zpad,
                                    double* best_val, long long* best_idx) {
  const int total = npad * npad;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int i = e / npad, j = e - i * npad;
    Zpad[e] = (i < n && j <= i) ? Linv[static_cast<size_t>(i) * ldz + j] : 0.0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) { *best_val = -INFINITY; *best_idx = 0x7fffffffffffffffLL; }
}

static int noalloc_fail(int code, const char* msg) { g_noalloc_error = msg; return code; }

static const char* last_handleless_error() {
  const char* e = g_noalloc_error;
  return e ? e : g_create_error.c_str();
}

extern "C" int plx_acq_sweep_noalloc(int device, void* stream, const double* Xc_dev, long long N,
                                     long long idx_base, const double* X_dev, const double* alpha_dev,
                                     const double* Linv_dev, int ldz, int n, int d, int kernel, double nu,
                                     double theta, int acq, double kappa, double eps, double y_max,
                                     double* best_value_dev, long long* best_index_dev) {
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return noalloc_fail(PLX_ERR_INVALID, "sweep_noalloc: device index out of range");
  if (n > NOALLOC_MAX_OBS)
    return noalloc_fail(PLX_ERR_INVALID, "sweep_noalloc: n_obs exceeds the static scratch (max 256 observations); "
                                         "use plx_acq_sweep_dev for larger problems");
  if (n <= 0 || d <= 0 || N < 0 || ldz < n || acq < 0 || acq > 2 || !X_dev || !alpha_dev || !Linv_dev ||
      !best_value_dev || !best_index_dev || (N > 0 && !Xc_dev))
    return noalloc_fail(PLX_ERR_INVALID, "sweep_noalloc: bad arguments");
  int kind;
  if (kernel_kind(kernel, nu, &kind) != PLX_OK) return noalloc_fail(PLX_ERR_UNSUPPORTED, "sweep_noalloc: unsupported kernel / nu");
  // the caller's current device is put back on every path out of here
  struct DeviceGuard {
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  } guard;
  if (cudaGetDevice(&guard.prev) != cudaSuccess) guard.prev = -1;
  if (cudaSetDevice(device) != cudaSuccess) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: cudaSetDevice failed");
  if (g_noalloc_ptr[device] == nullptr) {
    // once per device, with that device current
    double* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: cudaGetSymbolAddress failed");
    e = cudaFuncSetAttribute(sweep_simple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SWEEP_SIMPLE_SMEM);
    if (e != cudaSuccess) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: cudaFuncSetAttribute failed");
    g_noalloc_ptr[device] = ptr;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  double* base = g_noalloc_ptr[device];
  double* Zpad = base;
  double* Ks = base + NOALLOC_Z;
  double* blk_val = Ks + NOALLOC_KS;
  long long* blk_idx = reinterpret_cast<long long*>(blk_val + NOALLOC_GRID);
  const int npad = (n + 127) / 128 * 128;
  noalloc_pack_kernel<<<64, 256, 0, st>>>(Linv_dev, n, ldz, npad, Zpad, best_value_dev, best_index_dev);
  if (cudaGetLastError() != cudaSuccess) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: launch of the pack kernel failed");
  if (N == 0) return PLX_OK;                    // outputs hold (-inf, INT64_MAX): no candidate
  SweepParams p{};
  p.Xc = Xc_dev; p.N = N; p.idx_base = idx_base;
  p.d = d; p.n = n; p.npad = npad; p.n16 = (n + 15) / 16 * 16;
  p.X = X_dev; p.Xt = nullptr; p.alpha = alpha_dev; p.Z = Zpad; p.ell = std::exp(theta);
  p.kind = kind; p.mg = (kind == K_GEN) ? make_matern_gen(nu) : MaternGen{}; p.acq = acq; p.kappa = kappa; p.eps = eps; p.y_max = y_max;
  p.Ks = Ks; p.out_mean = p.out_std = p.out_acq = nullptr;
  p.blk_val = blk_val; p.blk_idx = blk_idx; p.clamped = nullptr; p.dbg = 0;
  const long long ntiles = (N + 127) / 128;
  const int grid = static_cast<int>(std::min<long long>(ntiles, NOALLOC_GRID));
  sweep_simple_kernel<<<grid, 256, SWEEP_SIMPLE_SMEM, st>>>(p);
  if (cudaGetLastError() != cudaSuccess) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: launch of the sweep kernel failed");
  sweep_merge_kernel<<<1, 32, 0, st>>>(blk_val, blk_idx, grid, best_value_dev, best_index_dev);
  if (cudaGetLastError() != cudaSuccess) return noalloc_fail(PLX_ERR_CUDA, "sweep_noalloc: launch of the merge kernel failed");
  return PLX_OK;
}

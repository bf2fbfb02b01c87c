// extern "C" entry points of include/venture_b200.h: handle management,
// host<->device staging, and dispatch into the per-model site kernels.
#include <string.h>
#include <cmath>
#include "vk_common.cuh"
#include "weights.cuh"
#include "models.cuh"

namespace vk {
static thread_local std::string g_last_error;
void set_last_error(const std::string &m) { g_last_error = m; }
}  // namespace vk

using namespace vk;

#define VK_TRY try {
#define VK_CATCH                                                              \
  }                                                                           \
  catch (const vk::Error &e) { vk::set_last_error(e.what()); return e.code; } \
  catch (const std::exception &e) { vk::set_last_error(e.what()); return VK_ECUDA; } \
  return VK_OK;

static void check_handle(const vk_particles *h) {
  if (!h || !h->model) throw Error(VK_EINVAL, "null or destroyed vk_particles handle");
}
// every entry point passes through here: whatever it does next may change the weights, so the
// "min/max of lw are already in red[]" hint dies; vk_make_consistent may set it again on its way out
static void use_device(const vk_particles *h) {
  const_cast<vk_particles *>(h)->lw_minmax_fresh = false;
  const_cast<vk_particles *>(h)->lw_x0_done = false;
  VK_CUDA(cudaSetDevice(h->device));
}

// Device staging for a host output of n doubles: the handle's grow-only buffer
// (no cudaMalloc on the hot path).  finish() copies out and drains the stream;
// calls WITHOUT a host output stay stream-ordered (no synchronisation).
struct HostOut {
  vk_particles *h; double *host; size_t n;
  HostOut(vk_particles *h_, double *host_, size_t n_) : h(h_), host(host_), n(n_) {
    if (host && h->out_d.n < n) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->out_d.alloc(n); }
  }
  double *ptr() { return host ? h->out_d.p : nullptr; }
  void finish() {
    if (!host) return;
    VK_CUDA(cudaMemcpyAsync(host, h->out_d.p, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));
    vk::p2p_check_error(h);      // a peer-memory exchange that timed out is fatal for this particle set
  }
};

static vk_particles *new_handle(int kind, int64_t P, uint64_t seed, int device) {
  VK_REQUIRE(P >= 1 && P <= ((int64_t)1 << 30), VK_EINVAL, "n_particles must be in [1, 2^30]");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    throw Error(VK_ECUDA, std::string("no CUDA device available (this library has no CPU fallback): ") +
                              cudaGetErrorString(e));
  VK_REQUIRE(device >= 0 && device < n_dev, VK_EINVAL, "device index out of range");
  VK_CUDA(cudaSetDevice(device));
  vk_particles *h = new vk_particles();
  try {
    h->kind = kind; h->device = device; h->P = P; h->seed = seed;
    h->prof.on = getenv("VK_PROFILE") != nullptr;
    VK_CUDA(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device));
    VK_CUDA(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    h->lw.alloc((size_t)P);
    h->scratch_d.alloc((size_t)P);
    h->cdf.alloc((size_t)P);
    h->anc.alloc((size_t)P);
    h->red.alloc((size_t)(8 + 3 * h->sm_count * 8 + 64));
    h->counters.alloc(8);
    VK_CUDA(cudaMemsetAsync(h->lw.p, 0, h->lw.bytes(), h->stream));
    VK_CUDA(cudaMemsetAsync(h->anc.p, 0, h->anc.bytes(), h->stream));
    VK_CUDA(cudaMemsetAsync(h->counters.p, 0, h->counters.bytes(), h->stream));
  } catch (...) { delete h; throw; }
  return h;
}
static void finish_create(vk_particles *h, vk_particles **out) {
  VK_CUDA(cudaStreamSynchronize(h->stream));
  *out = h;
}

extern "C" {

const char *vk_last_error(void) { return vk::g_last_error.c_str(); }
int vk_version(void) { return 100; }
int vk_device_count(int *n_out) {
  VK_TRY
  VK_REQUIRE(n_out, VK_EINVAL, "n_out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { n = 0; cudaGetLastError(); }
  *n_out = n;
  VK_CATCH
}

int vk_create_normal_normal(int64_t P, uint64_t seed, int device, double prior_mu, double prior_sigma,
                            double obs_sigma, int64_t max_obs, vk_particles **out) {
  VK_TRY
  VK_REQUIRE(out, VK_EINVAL, "out is NULL");
  VK_REQUIRE(prior_sigma > 0 && obs_sigma > 0, VK_EINVAL, "normal: sigma must be positive");
  vk_particles *h = new_handle(VK_MODEL_NORMAL_NORMAL, P, seed, device);
  try { h->model = make_normal_normal(h, prior_mu, prior_sigma, obs_sigma, max_obs); }
  catch (...) { cudaStreamDestroy(h->own_stream); delete h; throw; }
  finish_create(h, out);
  VK_CATCH
}

int vk_create_program(int64_t P, uint64_t seed, int device, const int32_t *code, int64_t n_code, const double *consts,
                      int64_t n_consts, const int32_t *meta, int64_t n_meta, const double *data, int64_t n_data,
                      const double *affected, vk_particles **out) {
  VK_TRY
  VK_REQUIRE(out && code && consts && meta && data && affected, VK_EINVAL, "vk_create_program: NULL argument");
  VK_REQUIRE(n_code >= 0 && n_consts >= 0 && n_meta >= 0 && n_data >= 0, VK_EINVAL, "vk_create_program: negative length");
  vk_particles *h = new_handle(VK_MODEL_PROGRAM, P, seed, device);
  try { h->model = make_program(h, code, n_code, consts, n_consts, meta, n_meta, data, n_data, affected); }
  catch (...) { cudaStreamDestroy(h->own_stream); delete h; throw; }
  finish_create(h, out);
  VK_CATCH
}

int vk_program_resimulate(vk_particles *h, int32_t first_site) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(h->kind == VK_MODEL_PROGRAM, VK_EINVAL, "vk_program_resimulate: not a compiled program");
  program_resimulate_from(h->model, first_site);
  VK_CATCH
}

int vk_create_blr(int64_t P, uint64_t seed, int device, int32_t D, int64_t n_obs_capacity,
                  double prior_mu, double prior_sigma, vk_particles **out) {
  VK_TRY
  VK_REQUIRE(out, VK_EINVAL, "out is NULL");
  VK_REQUIRE(prior_sigma > 0, VK_EINVAL, "normal: sigma must be positive");
  vk_particles *h = new_handle(VK_MODEL_BLR, P, seed, device);
  try { h->model = make_blr(h, D, n_obs_capacity, prior_mu, prior_sigma); }
  catch (...) { cudaStreamDestroy(h->own_stream); delete h; throw; }
  finish_create(h, out);
  VK_CATCH
}

int vk_create_hmm(int64_t P, uint64_t seed, int device, int32_t S, int32_t O, int32_t n_trans_rows,
                  const double *trans_host, const double *emit_host, int32_t init_state, int32_t T_max,
                  vk_particles **out) {
  VK_TRY
  VK_REQUIRE(out && trans_host && emit_host, VK_EINVAL, "NULL argument");
  vk_particles *h = new_handle(VK_MODEL_HMM, P, seed, device);
  try { h->model = make_hmm(h, S, O, n_trans_rows, trans_host, emit_host, init_state, T_max); }
  catch (...) { cudaStreamDestroy(h->own_stream); delete h; throw; }
  finish_create(h, out);
  VK_CATCH
}

int vk_create_lda(int64_t P, uint64_t seed, int device, int32_t K, int32_t V, int32_t n_docs,
                  int32_t max_tokens, double hyper_shape, double hyper_rate, vk_particles **out) {
  VK_TRY
  VK_REQUIRE(out, VK_EINVAL, "out is NULL");
  vk_particles *h = new_handle(VK_MODEL_LDA, P, seed, device);
  try { h->model = make_lda(h, K, V, n_docs, max_tokens, hyper_shape, hyper_rate); }
  catch (...) { cudaStreamDestroy(h->own_stream); delete h; throw; }
  finish_create(h, out);
  VK_CATCH
}

int vk_destroy(vk_particles *h) {
  VK_TRY
  if (!h) return VK_OK;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  h->prof.resolve();
  { char who[64]; snprintf(who, sizeof who, "rank %s dev %d", getenv("RANK") ? getenv("RANK") : "-", h->device); h->prof.report(who); }
  vk::p2p_release(h);
  delete h->model; h->model = nullptr;
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
  VK_CATCH
}

int vk_kind(const vk_particles *h) { return h ? h->kind : 0; }
int64_t vk_num_particles(const vk_particles *h) { return h ? h->P : 0; }
int vk_set_seed(vk_particles *h, uint64_t seed) { VK_TRY check_handle(h); h->seed = seed; h->step = 0; VK_CATCH }
int vk_get_seed(const vk_particles *h, uint64_t *seed_out) { VK_TRY check_handle(h); VK_REQUIRE(seed_out, VK_EINVAL, "NULL"); *seed_out = h->seed; VK_CATCH }
int vk_get_step(const vk_particles *h, uint32_t *step_out) { VK_TRY check_handle(h); VK_REQUIRE(step_out, VK_EINVAL, "NULL"); *step_out = h->step; VK_CATCH }
int vk_set_step(vk_particles *h, uint32_t step) { VK_TRY check_handle(h); h->step = step; VK_CATCH }
int vk_set_particle_base(vk_particles *h, int64_t base) { VK_TRY check_handle(h); VK_REQUIRE(base >= 0, VK_EINVAL, "negative base"); h->p_base = base; VK_CATCH }
int vk_set_stream(vk_particles *h, void *cuda_stream) {
  VK_TRY
  check_handle(h); use_device(h);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
  VK_CATCH
}
int vk_synchronize(vk_particles *h) {
  VK_TRY check_handle(h); use_device(h);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  h->prof.resolve();
  vk::p2p_check_error(h);
  VK_CATCH
}

int vk_reset_to_prior(vk_particles *h) {
  VK_TRY check_handle(h); use_device(h);
  h->model->reset_to_prior();
  VK_CATCH
}

int vk_nn_observe(vk_particles *h, const double *y_host, int64_t n) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(h->kind == VK_MODEL_NORMAL_NORMAL, VK_EUNSUPPORTED, "vk_nn_observe: not a normal-normal batch");
  VK_REQUIRE(n >= 0 && (y_host || n == 0), VK_EINVAL, "bad observation buffer");
  if (n) normal_normal_observe(h->model, y_host, n);
  VK_CATCH
}
int vk_blr_observe(vk_particles *h, const double *X_rows_host, const uint8_t *y_host, int64_t n) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(h->kind == VK_MODEL_BLR, VK_EUNSUPPORTED, "vk_blr_observe: not a BLR batch");
  VK_REQUIRE(n >= 0 && ((X_rows_host && y_host) || n == 0), VK_EINVAL, "bad observation buffers");
  if (n) blr_observe(h->model, X_rows_host, y_host, n);
  VK_CATCH
}
int vk_hmm_observe(vk_particles *h, int32_t symbol) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(h->kind == VK_MODEL_HMM, VK_EUNSUPPORTED, "vk_hmm_observe: not an HMM batch");
  hmm_observe(h->model, symbol);
  VK_CATCH
}
int vk_lda_observe(vk_particles *h, const int32_t *doc_host, const int32_t *word_host, int64_t n) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(h->kind == VK_MODEL_LDA, VK_EUNSUPPORTED, "vk_lda_observe: not an LDA batch");
  VK_REQUIRE(n >= 0 && ((doc_host && word_host) || n == 0), VK_EINVAL, "bad observation buffers");
  if (n) lda_observe(h->model, doc_host, word_host, n);
  VK_CATCH
}

int vk_forget_observations(vk_particles *h, double *weights_out_host) {
  VK_TRY check_handle(h); use_device(h);
  HostOut o(h, weights_out_host, (size_t)h->P);
  h->model->forget_observations(o.ptr());
  o.finish();
  VK_CATCH
}
int vk_make_consistent(vk_particles *h, double *increments_out_host) {
  VK_TRY check_handle(h); use_device(h);
  HostOut o(h, increments_out_host, (size_t)h->P);
  h->model->make_consistent(o.ptr());     // may set h->lw_minmax_fresh
  o.finish();
  VK_CATCH
}
int vk_likelihood_weight(vk_particles *h) {
  VK_TRY check_handle(h); use_device(h);
  h->model->likelihood_weight();
  VK_CATCH
}
int vk_get_log_weights(vk_particles *h, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  VK_CUDA(cudaMemcpyAsync(out_host, h->lw.p, h->lw.bytes(), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  h->prof.resolve();
  vk::p2p_check_error(h);
  VK_CATCH
}
int vk_set_log_weights(vk_particles *h, const double *in_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(in_host, VK_EINVAL, "in_host is NULL");
  VK_CUDA(cudaMemcpyAsync(h->lw.p, in_host, h->lw.bytes(), cudaMemcpyHostToDevice, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_increment_log_weights(vk_particles *h, const double *in_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(in_host, VK_EINVAL, "in_host is NULL");
  VK_CUDA(cudaMemcpyAsync(h->scratch_d.p, in_host, h->lw.bytes(), cudaMemcpyHostToDevice, h->stream));
  launch_axpy(h, h->lw.p, h->scratch_d.p, h->P);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_logsumexp(vk_particles *h, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  device_logsumexp(h->lw.p, h->P, h->red.p, h->counters.p, h->sm_count, h->stream);
  VK_CUDA(cudaMemcpyAsync(out_host, h->red.p + 4, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_normalized_probs(vk_particles *h, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  launch_normalized_probs(h, h->scratch_d.p);
  VK_CUDA(cudaMemcpyAsync(out_host, h->scratch_d.p, h->lw.bytes(), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_logsumexp_array(const double *in_host, int64_t n, int device, double *out_host) {
  VK_TRY
  VK_REQUIRE(out_host && n >= 0 && (in_host || n == 0), VK_EINVAL, "bad arguments");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0) throw Error(VK_ECUDA, "no CUDA device available (no CPU fallback)");
  VK_CUDA(cudaSetDevice(device));
  int sm = 148;
  VK_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device));
  DevBuf<double> a, red; DevBuf<unsigned int> ctr;
  a.alloc((size_t)(n > 0 ? n : 1)); red.alloc((size_t)(8 + 3 * sm * 8 + 64)); ctr.alloc(8);
  VK_CUDA(cudaMemset(ctr.p, 0, ctr.bytes()));
  if (n) VK_CUDA(cudaMemcpy(a.p, in_host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
  device_logsumexp(a.p, n, red.p, ctr.p, sm, 0);
  VK_CUDA(cudaMemcpy(out_host, red.p + 4, sizeof(double), cudaMemcpyDeviceToHost));
  VK_CATCH
}

int vk_resample(vk_particles *h, int64_t n_new, int mode, const double *uniforms_host,
                int32_t *ancestors_out_host) {
  // red[0..2] produced with the peer exchange describe ALL ranks' weights: not this local resample's input
  const bool minmax_fresh = h && h->lw_minmax_fresh && !h->lw_x0_done;
  VK_TRY check_handle(h); use_device(h);
  h->lw_minmax_fresh = minmax_fresh;   // resample() consumes it
  DevBuf<double> u;
  if (uniforms_host) {
    size_t n = mode == VK_RESAMPLE_SYSTEMATIC ? 1 : (size_t)n_new;
    u.alloc(n);
    VK_CUDA(cudaMemcpyAsync(u.p, uniforms_host, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  }
  resample(h, n_new, mode, uniforms_host ? u.p : nullptr);
  if (ancestors_out_host)
    VK_CUDA(cudaMemcpyAsync(ancestors_out_host, h->anc.p, (size_t)n_new * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  // trace_set.py:207: resample ends with incorporate()
  h->model->make_consistent(nullptr);
  if (uniforms_host || ancestors_out_host) VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}

int vk_mh(vk_particles *h, int32_t scope, int32_t block, int32_t transitions, double *mean_nodes_out_host,
          int64_t *accepted_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(transitions >= 0, VK_EINVAL, "transitions must be >= 0");
  HostOut o(h, mean_nodes_out_host, (size_t)h->P);
  if (accepted_out_host && h->acc_d.n < (size_t)h->P) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->acc_d.alloc((size_t)h->P); }
  long long *acc = accepted_out_host ? h->acc_d.p : nullptr;
  if (acc && transitions == 0) VK_CUDA(cudaMemsetAsync(acc, 0, (size_t)h->P * sizeof(long long), h->stream));
  h->model->mh(scope, block, transitions, o.ptr(), acc);
  if (acc)
    VK_CUDA(cudaMemcpyAsync(accepted_out_host, acc, (size_t)h->P * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  o.finish();
  if (acc && !mean_nodes_out_host) VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}

// ---- peer-memory (NVLink) resampling across the GPUs of one node: csrc/p2p_resample.cuh -------------
int vk_p2p_export(vk_particles *h, unsigned char *handles_out, int32_t max_handles, int32_t *n_handles_out) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(handles_out && n_handles_out, VK_EINVAL, "vk_p2p_export: null output");
  *n_handles_out = vk::p2p_export(h, handles_out, max_handles);
  VK_CATCH
}
int vk_p2p_open(vk_particles *h, int32_t world, int32_t rank, const unsigned char *all_handles, int32_t n_handles) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(all_handles, VK_EINVAL, "vk_p2p_open: null handles");
  vk::p2p_open(h, world, rank, all_handles, n_handles);
  VK_CATCH
}
int vk_p2p_resample(vk_particles *h, int mode) {
  const bool minmax_fresh = h && h->lw_minmax_fresh, x0_done = h && h->lw_x0_done;
  VK_TRY check_handle(h); use_device(h);
  h->lw_minmax_fresh = minmax_fresh; h->lw_x0_done = x0_done;   // p2p_resample() consumes them
  vk::p2p_resample(h, mode);
  // trace_set.py:207: resample ends with incorporate()
  h->model->make_consistent(nullptr);
  VK_CATCH
}
int vk_p2p_status(vk_particles *h, int32_t *timed_out, double *lse_out) {
  VK_TRY check_handle(h); use_device(h);
  int e = 0; double l = 0.0;
  vk::p2p_status(h, &e, &l);
  if (timed_out) *timed_out = e;
  if (lse_out) *lse_out = l;
  VK_CATCH
}

int vk_pgibbs(vk_particles *h, int32_t scope, int32_t block, int32_t particles, int32_t transitions,
              double *mean_nodes_out_host, int64_t *accepted_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(transitions >= 0, VK_EINVAL, "transitions must be >= 0");
  HostOut o(h, mean_nodes_out_host, (size_t)h->P);
  if (accepted_out_host && h->acc_d.n < (size_t)h->P) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->acc_d.alloc((size_t)h->P); }
  long long *acc = accepted_out_host ? h->acc_d.p : nullptr;
  if (acc && transitions == 0) VK_CUDA(cudaMemsetAsync(acc, 0, (size_t)h->P * sizeof(long long), h->stream));
  h->model->pgibbs(scope, block, particles, transitions, o.ptr(), acc);
  if (acc)
    VK_CUDA(cudaMemcpyAsync(accepted_out_host, acc, (size_t)h->P * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  o.finish();
  if (acc && !mean_nodes_out_host) VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}

static int slice_call(vk_particles *h, int32_t scope, int32_t block, double w, double m, int doubling, int32_t transitions,
                      double *mean_nodes_out_host, int64_t *moved_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(transitions >= 0, VK_EINVAL, "transitions must be >= 0");
  HostOut o(h, mean_nodes_out_host, (size_t)h->P);
  if (moved_out_host && h->acc_d.n < (size_t)h->P) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->acc_d.alloc((size_t)h->P); }
  long long *acc = moved_out_host ? h->acc_d.p : nullptr;
  if (acc && transitions == 0) VK_CUDA(cudaMemsetAsync(acc, 0, (size_t)h->P * sizeof(long long), h->stream));
  h->model->slice(scope, block, w, m, doubling, transitions, o.ptr(), acc);
  if (acc)
    VK_CUDA(cudaMemcpyAsync(moved_out_host, acc, (size_t)h->P * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  o.finish();
  if (acc && !mean_nodes_out_host) VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_slice(vk_particles *h, int32_t scope, int32_t block, double w, double m, int32_t transitions,
             double *mean_nodes_out_host, int64_t *moved_out_host) {
  return slice_call(h, scope, block, w, m, 0, transitions, mean_nodes_out_host, moved_out_host);
}
int vk_slice_doubling(vk_particles *h, int32_t scope, int32_t block, double w, int32_t p, int32_t transitions,
                      double *mean_nodes_out_host, int64_t *moved_out_host) {
  return slice_call(h, scope, block, w, (double)p, 1, transitions, mean_nodes_out_host, moved_out_host);
}

static int gibbs_call(vk_particles *h, int32_t scope, int32_t block, int32_t transitions, double *mean_nodes_out_host,
                      int64_t *changed_out_host, bool map) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(transitions >= 0, VK_EINVAL, "transitions must be >= 0");
  HostOut o(h, mean_nodes_out_host, (size_t)h->P);
  if (changed_out_host && h->acc_d.n < (size_t)h->P) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->acc_d.alloc((size_t)h->P); }
  long long *acc = changed_out_host ? h->acc_d.p : nullptr;
  if (acc && transitions == 0) VK_CUDA(cudaMemsetAsync(acc, 0, (size_t)h->P * sizeof(long long), h->stream));
  if (map) h->model->emap(scope, block, transitions, o.ptr(), acc);
  else h->model->gibbs(scope, block, transitions, o.ptr(), acc);
  if (acc)
    VK_CUDA(cudaMemcpyAsync(changed_out_host, acc, (size_t)h->P * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  o.finish();
  if (acc && !mean_nodes_out_host) VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}

int vk_gibbs(vk_particles *h, int32_t scope, int32_t block, int32_t transitions, double *mean_nodes_out_host,
             int64_t *changed_out_host) {
  return gibbs_call(h, scope, block, transitions, mean_nodes_out_host, changed_out_host, false);
}
int vk_emap(vk_particles *h, int32_t scope, int32_t block, int32_t transitions, double *mean_nodes_out_host,
            int64_t *changed_out_host) {
  return gibbs_call(h, scope, block, transitions, mean_nodes_out_host, changed_out_host, true);
}

int vk_subsampled_mh(vk_particles *h, int32_t scope, int32_t block, int64_t Nbatch, int32_t k0,
                     double epsilon, int32_t use_delta_kernels, double delta_sigma, int32_t update_values,
                     int32_t transitions, double *mean_nodes_out_host, double *stats_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(transitions >= 0 && Nbatch >= 1 && k0 >= 0, VK_EINVAL, "bad subsampled_mh arguments");
  // both outputs share the handle's one output buffer: statistics in [0, 3P), node counts in [3P, 4P)
  const size_t P = (size_t)h->P;
  const bool any = mean_nodes_out_host || stats_out_host;
  if (any && h->out_d.n < 4 * P) { VK_CUDA(cudaStreamSynchronize(h->stream)); h->out_d.alloc(4 * P); }
  double *st = stats_out_host ? h->out_d.p : nullptr, *nd = mean_nodes_out_host ? h->out_d.p + 3 * P : nullptr;
  h->model->subsampled_mh(scope, block, Nbatch, k0, epsilon, use_delta_kernels, delta_sigma, update_values,
                          transitions, nd, st);
  if (st) VK_CUDA(cudaMemcpyAsync(stats_out_host, st, 3 * P * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (nd) VK_CUDA(cudaMemcpyAsync(mean_nodes_out_host, nd, P * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (any) { VK_CUDA(cudaStreamSynchronize(h->stream)); vk::p2p_check_error(h); }
  VK_CATCH
}

int vk_log_likelihood_at(vk_particles *h, int32_t scope, int32_t block, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  HostOut o(h, out_host, (size_t)h->P);
  h->model->log_likelihood_at(scope, block, false, o.ptr());
  o.finish();
  VK_CATCH
}
int vk_log_joint_at(vk_particles *h, int32_t scope, int32_t block, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  HostOut o(h, out_host, (size_t)h->P);
  h->model->log_likelihood_at(scope, block, true, o.ptr());
  o.finish();
  VK_CATCH
}

int vk_num_sites(const vk_particles *h, int64_t *n_out) { VK_TRY check_handle(h); VK_REQUIRE(n_out, VK_EINVAL, "NULL"); *n_out = h->model->num_sites(); VK_CATCH }
int vk_num_blocks(const vk_particles *h, int32_t scope, int64_t *n_out) { VK_TRY check_handle(h); VK_REQUIRE(n_out, VK_EINVAL, "NULL"); *n_out = h->model->num_blocks(scope); VK_CATCH }
int vk_read_site(vk_particles *h, int64_t site, double *out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out_host, VK_EINVAL, "out_host is NULL");
  HostOut o(h, out_host, (size_t)h->P);
  h->model->read_site(site, o.ptr());
  o.finish();
  VK_CATCH
}
int vk_write_site(vk_particles *h, int64_t site, const double *in_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(in_host, VK_EINVAL, "in_host is NULL");
  VK_CUDA(cudaMemcpyAsync(h->scratch_d.p, in_host, (size_t)h->P * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  h->model->write_site(site, h->scratch_d.p);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_set_precision(vk_particles *h, int32_t precision) { VK_TRY check_handle(h); use_device(h); h->model->set_precision(precision); VK_CATCH }
int vk_get_precision(const vk_particles *h, int32_t *precision_out) {
  VK_TRY check_handle(h); VK_REQUIRE(precision_out, VK_EINVAL, "NULL"); *precision_out = h->model->get_precision(); VK_CATCH
}
int vk_refresh(vk_particles *h) { VK_TRY check_handle(h); use_device(h); h->model->refresh(); VK_CATCH }

int vk_device_buffer(vk_particles *h, int32_t which, void **dev_ptr_out, int64_t *n_bytes_out) {
  VK_TRY check_handle(h);
  VK_REQUIRE(dev_ptr_out && n_bytes_out, VK_EINVAL, "NULL");
  if (which == 0) { *dev_ptr_out = h->lw.p; *n_bytes_out = (int64_t)h->lw.bytes(); }
  else if (which == -1) { *dev_ptr_out = h->anc.p; *n_bytes_out = (int64_t)h->anc.bytes(); }
  else if (which == -2) { *dev_ptr_out = h->cdf.p; *n_bytes_out = (int64_t)h->cdf.bytes(); }
  else { int rc = h->model->device_buffer(which, dev_ptr_out, n_bytes_out); if (rc) throw Error(rc, "no such device buffer"); }
  VK_CATCH
}


// ---- multi-GPU plumbing ---------------------------------------------------------
int vk_shard_weight_max(vk_particles *h, double *max_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(max_out_host, VK_EINVAL, "NULL");
  shard_weight_max(h);
  VK_CUDA(cudaMemcpyAsync(max_out_host, h->red.p, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_shard_weight_stats(vk_particles *h, double *out3_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(out3_host, VK_EINVAL, "NULL");
  shard_weight_max(h);
  VK_CUDA(cudaMemcpyAsync(out3_host, h->red.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_shard_build_cdf(vk_particles *h, double global_max, int32_t frac_bits, uint64_t *mass_out_host,
                       double *sumexp_out_host) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(mass_out_host && sumexp_out_host, VK_EINVAL, "NULL");
  VK_REQUIRE(frac_bits >= 1 && frac_bits <= 52, VK_EINVAL, "frac_bits must be in [1,52]");
  shard_build_cdf(h, global_max, frac_bits);
  VK_CUDA(cudaMemcpyAsync(mass_out_host, h->cdf.p + (h->P - 1), sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaMemcpyAsync(sumexp_out_host, h->red.p + 3, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_shard_ancestors(vk_particles *h, uint64_t offset, uint64_t total, int64_t n_new_total, int64_t k_begin,
                       int64_t k_count, int mode, uint32_t step, const double *uniforms_host, int32_t *anc_out_dev) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(anc_out_dev || k_count == 0, VK_EINVAL, "NULL");
  VK_REQUIRE(mode == VK_RESAMPLE_MULTINOMIAL || mode == VK_RESAMPLE_SYSTEMATIC, VK_EINVAL, "bad mode");
  DevBuf<double> u;
  if (uniforms_host) {
    size_t n = mode == VK_RESAMPLE_SYSTEMATIC ? 1 : (size_t)n_new_total;
    u.alloc(n);
    VK_CUDA(cudaMemcpyAsync(u.p, uniforms_host, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  }
  launch_ancestor_search_shard(h, offset, total, n_new_total, k_begin, k_count, mode, step,
                               uniforms_host ? u.p : nullptr, anc_out_dev);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_state_row_bytes(vk_particles *h, int64_t *bytes_out) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(bytes_out, VK_EINVAL, "NULL");
  *bytes_out = (int64_t)packed_row_bytes(h);
  VK_CATCH
}
int vk_pack_state(vk_particles *h, const int32_t *particles_dev, int64_t n, void *packed_out_dev) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(n == 0 || (particles_dev && packed_out_dev), VK_EINVAL, "NULL");
  pack_state(h, particles_dev, n, packed_out_dev, false);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_unpack_state(vk_particles *h, const int32_t *slots_dev, int64_t n, const void *packed_in_dev) {
  VK_TRY check_handle(h); use_device(h);
  VK_REQUIRE(n == 0 || (slots_dev && packed_in_dev), VK_EINVAL, "NULL");
  pack_state(h, slots_dev, n, const_cast<void *>(packed_in_dev), true);
  h->model->after_resample();
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_fill_log_weights(vk_particles *h, double value) {
  VK_TRY check_handle(h); use_device(h);
  launch_fill_even_weights(h, h->P, value, 0.0);
  VK_CATCH
}

int vk_blr_set_observation_shard(vk_particles *h, int64_t global_offset, int64_t n_total) {
  VK_TRY check_handle(h);
  VK_REQUIRE(h->kind == VK_MODEL_BLR, VK_EUNSUPPORTED, "not a BLR batch");
  blr_set_observation_shard(h->model, global_offset, n_total);
  VK_CATCH
}
#define VK_BLR_ONLY VK_REQUIRE(h->kind == VK_MODEL_BLR, VK_EUNSUPPORTED, "subsampled_mh is defined for the BLR model only")
int vk_ssmh_begin(vk_particles *h) { VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY; blr_ss_begin(h->model); VK_CATCH }
int vk_ssmh_propose(vk_particles *h, int32_t block, int32_t t_index, int32_t delta_mode, double delta_arg) {
  VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY;
  blr_ss_propose(h->model, block, t_index, delta_mode, delta_arg);
  VK_CATCH
}
int vk_ssmh_round(vk_particles *h, int64_t q0, int64_t q1, void **sums_dev_out) {
  VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY;
  VK_REQUIRE(sums_dev_out && q0 >= 0 && q1 >= q0, VK_EINVAL, "bad round");
  *sums_dev_out = blr_ss_round(h->model, q0, q1);
  VK_CUDA(cudaStreamSynchronize(h->stream));
  VK_CATCH
}
int vk_ssmh_decide(vk_particles *h, const void *sums_dev, int64_t n_end, int32_t k0, double epsilon,
                   int32_t *n_active_out) {
  VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY;
  VK_REQUIRE(sums_dev && n_active_out, VK_EINVAL, "NULL");
  *n_active_out = blr_ss_decide(h->model, (const double *)sums_dev, n_end, k0, epsilon);
  VK_CATCH
}
int vk_ssmh_commit(vk_particles *h) { VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY; blr_ss_commit(h->model); VK_CUDA(cudaStreamSynchronize(h->stream)); VK_CATCH }
int vk_ssmh_stats(vk_particles *h, int32_t transitions, double *stats_out_host) {
  VK_TRY check_handle(h); use_device(h); VK_BLR_ONLY;
  VK_REQUIRE(stats_out_host, VK_EINVAL, "NULL");
  HostOut o(h, stats_out_host, (size_t)h->P * 3);
  blr_ss_stats(h->model, transitions, o.ptr());
  o.finish();
  VK_CATCH
}

}  // extern "C"

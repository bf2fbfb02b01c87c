// Config C1: the normal-normal model of benchmarks/Makefile:6-9
//   assume x = normal(mu0, s0);  observe normal(x, s) = y_k   (k < n_obs)
// compiled to one site program:
//   scaffold(x): DRG {x, one lookup per observation}, absorbing {obs_k}.
//   likelihood_weight: x' ~ N(mu0,s0);  lw = sum_k logN(y_k; x', s)
//                      (backend/lite/trace.py:522-530, continuous.py:313-323)
//   resimulation_mh  : alpha = sum_k logN(y_k;x',s) - sum_k logN(y_k;x,s)
//                      (infer/mh.py:59-113,187-208; proposal from the prior
//                       contributes 0 to both weights)
//   makeConsistent   : increment = sum_{new k} logN(y_k; x, s)
//                      (backend/lite/trace.py:422-448)
// Layout: x[P], lw[P] particle-minor; one thread per particle, fully
// coalesced 8-byte accesses; y_k (shared by all particles) is broadcast
// through L1.  HBM-bound: 2*8 B per particle for likelihood_weight.
#include "vk_common.cuh"
#include "sp_math.cuh"

namespace vk {

struct NNParams { double mu0, s0, s; };

__device__ __forceinline__ double nn_loglik(const double *__restrict__ y, int64_t k0, int64_t k1,
                                            double x, double s) {
  double acc = 0.0;
  for (int64_t k = k0; k < k1; ++k) acc += normal_logpdf(y[k], x, s);
  return acc;
}

__global__ void nn_likelihood_weight_kernel(double *__restrict__ x, double *__restrict__ lw,
                                            const double *__restrict__ y, int64_t n_obs,
                                            int64_t P, int64_t p_base, NNParams prm, uint64_t seed,
                                            uint32_t step) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    Stream r(seed, (uint64_t)(p_base + p), step);
    double xn = normal_sim(r, prm.mu0, prm.s0);
    x[p] = xn;
    lw[p] = nn_loglik(y, 0, n_obs, xn, prm.s);
  }
}

__global__ void nn_make_consistent_kernel(const double *__restrict__ x, double *__restrict__ lw,
                                          double *__restrict__ incr, const double *__restrict__ y,
                                          int64_t k0, int64_t k1, int64_t P, NNParams prm) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    double w = nn_loglik(y, k0, k1, x[p], prm.s);
    if (w != w) w = -INFINITY;  // lite/trace.py:438-448
    if (incr) incr[p] = w;
    lw[p] += w;
  }
}

// mh.py:187-208 +-inf policy
__device__ __forceinline__ double mh_log_alpha(double rho, double xi) {
  if (rho == -INFINITY && xi == -INFINITY) return INFINITY;
  if (rho == INFINITY && xi == INFINITY) return -INFINITY;
  return xi - rho;
}

__global__ void nn_mh_kernel(double *__restrict__ x, const double *__restrict__ y, int64_t n_obs,
                             int64_t P, int64_t p_base, NNParams prm, uint64_t seed,
                             uint32_t step0, int K, long long *__restrict__ acc_out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    double xc = x[p];
    double rho = nn_loglik(y, 0, n_obs, xc, prm.s);
    long long acc = 0;
    for (int t = 0; t < K; ++t) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      (void)r.uniform();  // draw 0: block choice (one block in the default scope)
      double logu = log(r.uniform());  // draw 1: accept uniform (mh.py:85)
      double xn = normal_sim(r, prm.mu0, prm.s0);  // draws 2,3: proposal
      double xi = nn_loglik(y, 0, n_obs, xn, prm.s);
      double alpha = mh_log_alpha(rho, xi);
      if (logu < alpha) { xc = xn; rho = xi; ++acc; }
    }
    x[p] = xc;
    if (acc_out) acc_out[p] = acc;
  }
}

__global__ void nn_loglik_at_kernel(const double *__restrict__ x, const double *__restrict__ y,
                                    int64_t n_obs, int64_t P, NNParams prm, int joint,
                                    double *__restrict__ out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    double xc = x[p];
    double v = nn_loglik(y, 0, n_obs, xc, prm.s);
    if (joint) v += normal_logpdf(xc, prm.mu0, prm.s0);
    out[p] = v;
  }
}

__global__ void fill_const_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}

struct NormalNormalModel : Model {
  NNParams prm;
  int64_t max_obs, n_obs = 0, n_prop = 0;
  DevBuf<double> y;
  double *x = nullptr, *x_alt = nullptr;
  DevBuf<double> xa, xb;

  NormalNormalModel(vk_particles *h_, double mu0, double s0, double s, int64_t max_obs_)
      : Model(h_), max_obs(max_obs_) {
    prm.mu0 = mu0; prm.s0 = s0; prm.s = s;
    y.alloc((size_t)(max_obs > 0 ? max_obs : 1));
    xa.alloc((size_t)h->P); xb.alloc((size_t)h->P);
    x = xa.p; x_alt = xb.p;
    VK_CUDA(cudaMemsetAsync(x, 0, (size_t)h->P * sizeof(double), h->stream));
  }
  // the prior simulation that `assume` performs (regen.py eval); weights untouched
  void reset_to_prior() override {
    int grid = grid_for(h->P, 256, h->sm_count, 4);
    uint32_t step = h->next_step();
    nn_likelihood_weight_kernel<<<grid, 256, 0, h->stream>>>(x, h->scratch_d.p, y.p, 0, h->P, h->p_base, prm, h->seed, step);
    VK_CUDA(cudaGetLastError());
  }
  void observe(const double *y_host, int64_t n) {
    VK_REQUIRE(n_obs + n <= max_obs, VK_EINVAL, "normal-normal: observation capacity exceeded");
    VK_CUDA(cudaMemcpyAsync(y.p + n_obs, y_host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));
    n_obs += n;
  }
  void make_consistent(double *incr_dev) override {
    int grid = grid_for(h->P, 256, h->sm_count, 4);
    nn_make_consistent_kernel<<<grid, 256, 0, h->stream>>>(x, h->lw.p, incr_dev, y.p, n_prop, n_obs, h->P, prm);
    VK_CUDA(cudaGetLastError());
    n_prop = n_obs;
  }
  void likelihood_weight() override {
    // unpropagated observations are constrained first (TraceSet flows always
    // incorporate before inference); their increment is then overwritten.
    n_prop = n_obs;
    int grid = grid_for(h->P, 256, h->sm_count, 4);
    uint32_t step = h->next_step();
    nn_likelihood_weight_kernel<<<grid, 256, 0, h->stream>>>(x, h->lw.p, y.p, n_obs, h->P, h->p_base, prm, h->seed, step);
    VK_CUDA(cudaGetLastError());
  }
  void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "normal-normal: only the default scope exists");
    VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || block == VK_BLOCK_ALL || block == 0,
               VK_EINVAL, "normal-normal: block must be one/each/all/0");
    VK_REQUIRE(n_prop == n_obs, VK_ESTATE, "mh: unpropagated observations; call vk_make_consistent first");
    int grid = grid_for(h->P, 256, h->sm_count, 4);
    if (K > 0) {
      uint32_t step0 = h->next_step((uint32_t)K);
      nn_mh_kernel<<<grid, 256, 0, h->stream>>>(x, y.p, n_obs, h->P, h->p_base, prm, h->seed, step0, K, acc_dev);
    }
    if (mean_nodes_dev)  // |DRG| = x + one lookup node per observation (scaffold.py:67-68)
      fill_const_kernel<<<grid, 256, 0, h->stream>>>(mean_nodes_dev, h->P, K > 0 ? (double)(1 + n_obs) : 0.0);
    VK_CUDA(cudaGetLastError());
  }
  void log_likelihood_at(int scope, int block, bool joint, double *out_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "normal-normal: only the default scope exists");
    int grid = grid_for(h->P, 256, h->sm_count, 4);
    nn_loglik_at_kernel<<<grid, 256, 0, h->stream>>>(x, y.p, n_prop, h->P, prm, joint ? 1 : 0, out_dev);
    VK_CUDA(cudaGetLastError());
  }
  void forget_observations(double *w_dev) override {
    if (w_dev) log_likelihood_at(0, VK_BLOCK_ALL, false, w_dev);
    n_obs = n_prop = 0;
  }
  int64_t num_sites() const override { return 1; }
  int64_t num_blocks(int) const override { return 1; }
  void read_site(int64_t site, double *out_dev) override {
    VK_REQUIRE(site == 0, VK_EINVAL, "normal-normal: site 0 (x) is the only site");
    VK_CUDA(cudaMemcpyAsync(out_dev, x, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  void write_site(int64_t site, const double *in_dev) override {
    VK_REQUIRE(site == 0, VK_EINVAL, "normal-normal: site 0 (x) is the only site");
    VK_CUDA(cudaMemcpyAsync(x, in_dev, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  std::vector<StateArray> state_arrays(bool) override {
    StateArray a;
    a.cur = (void **)&x; a.alt = (void **)&x_alt; a.row_bytes = 0; a.n_rows = 1; a.elem_bytes = 8;
    return {a};
  }
  void resized(int64_t n_new, const std::vector<void *> &nc, const std::vector<void *> &na) override {
    xa.adopt(nc[0], (size_t)n_new); xb.adopt(na[0], (size_t)n_new);
    x = xa.p; x_alt = xb.p;
  }
  int device_buffer(int which, void **p, int64_t *bytes) override {
    if (which == 1) { *p = x; *bytes = (int64_t)h->P * 8; return VK_OK; }
    return VK_EINVAL;
  }
};

Model *make_normal_normal(vk_particles *h, double mu0, double s0, double s, int64_t max_obs) {
  return new NormalNormalModel(h, mu0, s0, s, max_obs);
}
void normal_normal_observe(Model *m, const double *y_host, int64_t n) {
  static_cast<NormalNormalModel *>(m)->observe(y_host, n);
}

}  // namespace vk

// Batched SP plugin surface: PSP.simulate(args) / PSP.logDensity(value, args)
// (backend/lite/psp.py:32-318) for the builtins the north star names, one
// application per thread.  The same device functions (sp_math.cuh) are the
// ones fused into the site kernels; this entry point exists so each primitive
// can be parity-tested (vs the oracle and scipy) and KS-tested on its own,
// like test/conformance/sps/ and test/properties/test_sps.py do with MockArgs
// (backend/lite/sp_use.py:30-62).
#include "vk_common.cuh"
#include "sp_math.cuh"

namespace vk {

static const int SP_MAX_DIM = 64;  // vector SPs via this entry point (per-thread local arrays)

__global__ void sp_simulate_kernel(int sp, const double *__restrict__ params, int n_params, int64_t n,
                                   int dim, uint64_t seed, uint32_t step, double *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double *q = params + i * n_params;
    Stream r(seed, (uint64_t)i, step);
    switch (sp) {
      case VK_SP_NORMAL: out[i] = normal_sim(r, q[0], q[1]); break;
      case VK_SP_GAMMA: out[i] = gamma_sim(r, q[0], q[1]); break;
      case VK_SP_BETA: out[i] = beta_sim(r, q[0], q[1]); break;
      case VK_SP_FLIP: out[i] = flip_sim(r, q[0]) ? 1.0 : 0.0; break;
      case VK_SP_LOG_FLIP: out[i] = log_flip_sim(r, q[0]) ? 1.0 : 0.0; break;
      case VK_SP_LOG_ODDS_FLIP: out[i] = log_odds_flip_sim(r, q[0]) ? 1.0 : 0.0; break;
      case VK_SP_CATEGORICAL: out[i] = (double)categorical_sim(r, q, n_params, 1); break;
      case VK_SP_LOG_CATEGORICAL: {  // utils.py:170-180: exp(lg - max) then categorical
        double m = -INFINITY, tmp[SP_MAX_DIM];
        for (int k = 0; k < n_params; ++k) m = fmax(m, q[k]);
        for (int k = 0; k < n_params; ++k) tmp[k] = m > -INFINITY ? exp(q[k] - m) : 0.0;
        out[i] = (double)categorical_sim(r, tmp, n_params, 1);
        break;
      }
      case VK_SP_DIRICHLET: dirichlet_sim(r, q, dim, out + i * dim); break;
      case VK_SP_SYM_DIRICHLET: {
        double al[SP_MAX_DIM];
        for (int k = 0; k < dim; ++k) al[k] = q[0];
        dirichlet_sim(r, al, dim, out + i * dim);
        break;
      }
      case VK_SP_SYM_DIR_CAT: {  // params: alpha, counts[dim]  (dirichlet.py:164-166)
        double wts[SP_MAX_DIM];
        for (int k = 0; k < dim; ++k) wts[k] = q[0] + q[1 + k];
        out[i] = (double)categorical_sim(r, wts, dim, 1);
        break;
      }
      case VK_SP_BETA_BERNOULLI: {  // params: a, b, yes, no  (discrete.py:440-443)
        double w = (q[0] + q[2]) / (q[0] + q[2] + q[1] + q[3]);
        out[i] = r.uniform() < w ? 1.0 : 0.0;
        break;
      }
      case VK_SP_DIR_CAT: {  // params: alpha[dim], counts[dim]  (dirichlet.py:164-166)
        double wts[SP_MAX_DIM];
        for (int k = 0; k < dim; ++k) wts[k] = q[k] + q[dim + k];
        out[i] = (double)categorical_sim(r, wts, dim, 1);
        break;
      }
      case VK_SP_UC_DIR_CAT: out[i] = (double)categorical_sim(r, q, dim, 1); break;     // params: theta[dim]  (dirichlet.py:271-273)
      case VK_SP_SUFF_BERNOULLI: out[i] = r.uniform() < q[0] ? 1.0 : 0.0; break;        // params: weight  (discrete.py:531-532)
    }
  }
}

__global__ void sp_log_density_kernel(int sp, const double *__restrict__ values,
                                      const double *__restrict__ params, int n_params, int64_t n, int dim,
                                      double *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double *q = params + i * n_params;
    bool vec = (sp == VK_SP_DIRICHLET || sp == VK_SP_SYM_DIRICHLET);
    double v = vec ? 0.0 : values[i];
    double res = NAN;
    switch (sp) {
      case VK_SP_NORMAL: res = normal_logpdf(v, q[0], q[1]); break;
      case VK_SP_GAMMA: res = gamma_logpdf(v, q[0], q[1]); break;
      case VK_SP_BETA: res = beta_logpdf(v, q[0], q[1]); break;
      case VK_SP_FLIP: res = flip_logpmf(v != 0.0, q[0]); break;
      case VK_SP_LOG_FLIP: res = log_flip_logpmf(v != 0.0, q[0]); break;
      case VK_SP_LOG_ODDS_FLIP: res = log_odds_flip_logpmf(v != 0.0, q[0]); break;
      case VK_SP_CATEGORICAL: res = categorical_logpmf((int)v, q, n_params, 1); break;
      case VK_SP_LOG_CATEGORICAL: res = log_categorical_logpmf((int)v, q, n_params, 1); break;
      case VK_SP_DIRICHLET: res = dirichlet_logpdf(values + i * dim, q, dim); break;
      case VK_SP_SYM_DIRICHLET: {
        double al[SP_MAX_DIM];
        for (int k = 0; k < dim; ++k) al[k] = q[0];
        res = dirichlet_logpdf(values + i * dim, al, dim);
        break;
      }
      case VK_SP_SYM_DIR_CAT: {  // dirichlet.py:168-173
        double tot = 0.0;
        for (int k = 0; k < dim; ++k) tot += q[1 + k];
        res = dircat_logpmf(q[1 + (int)v], q[0], tot, q[0] * (double)dim);
        break;
      }
      case VK_SP_BETA_BERNOULLI: res = betabern_logpmf(v != 0.0, q[0], q[1], q[2], q[3]); break;
      case VK_SP_DIR_CAT: {  // dirichlet.py:168-173
        double tot = 0.0, A = 0.0;
        for (int k = 0; k < dim; ++k) { tot += q[dim + k]; A += q[k]; }
        res = dircat_logpmf(q[dim + (int)v], q[(int)v], tot, A);
        break;
      }
      case VK_SP_UC_DIR_CAT: res = log(q[(int)v]); break;                               // dirichlet.py:275-277
      case VK_SP_SUFF_BERNOULLI: res = v != 0.0 ? log(q[0]) : log1p(-q[0]); break;       // discrete.py:534-538
    }
    out[i] = res;
  }
}

// logDensityOfData(aux): see include/venture_b200.h vk_sp_log_density_of_data
__device__ __forceinline__ double sp_data_term(int sp, const double *q, int dim) {
  switch (sp) {
    case VK_SP_SYM_DIR_CAT: return dircat_log_density_of_data(q, 0, q + 1, dim);
    case VK_SP_DIR_CAT: return dircat_log_density_of_data(q, 1, q + dim, dim);
    case VK_SP_BETA_BERNOULLI: return betabern_log_density_of_data(q[0], q[1], q[2], q[3]);
    case VK_SP_SUFF_BERNOULLI: return suff_bernoulli_log_density_of_data(q[0], q[1], q[2]);
  }
  return NAN;
}
__global__ void sp_log_density_of_data_kernel(int sp, const double *__restrict__ params, int n_params, int64_t n, int dim,
                                              double *__restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = sp_data_term(sp, params + i * n_params, dim);
}

// the uncollapsed makers' AAA kernels: a Gibbs draw of the made SP's parameter from its conjugate posterior;
// weight = the marginal likelihood of the data (PosteriorAAALKernel.weight, lkernel.py:137-156)
__global__ void sp_aaa_simulate_kernel(int sp, const double *__restrict__ params, int n_params, int64_t n, int dim,
                                       uint64_t seed, uint32_t step, double *__restrict__ out, double *__restrict__ wout) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double *q = params + i * n_params;
    Stream r(seed, (uint64_t)i, step);
    double post[SP_MAX_DIM];
    switch (sp) {
      case VK_SP_UC_DIR_CAT:        // dirichlet.py:250-263: theta ~ dirichlet(alpha + counts)
        for (int k = 0; k < dim; ++k) post[k] = q[dim + k] + q[k];
        dirichlet_sim(r, post, dim, out + i * dim);
        if (wout) wout[i] = sp_data_term(VK_SP_DIR_CAT, q, dim);
        break;
      case VK_SP_SYM_DIRICHLET:     // dirichlet.py:395-407: the symmetric maker
        for (int k = 0; k < dim; ++k) post[k] = q[1 + k] + q[0];
        dirichlet_sim(r, post, dim, out + i * dim);
        if (wout) wout[i] = sp_data_term(VK_SP_SYM_DIR_CAT, q, dim);
        break;
      case VK_SP_SUFF_BERNOULLI:    // discrete.py:502-510: weight ~ beta(a + yes, b + no)
        out[i] = beta_sim(r, q[0] + q[2], q[1] + q[3]);
        if (wout) wout[i] = sp_data_term(VK_SP_BETA_BERNOULLI, q, dim);
        break;
    }
  }
}

static void sp_check(int sp, int n_params, int64_t n, int dim) {
  VK_REQUIRE(sp >= VK_SP_NORMAL && sp <= VK_SP_SUFF_BERNOULLI, VK_EINVAL, "unknown SP id");
  VK_REQUIRE(n >= 0 && n_params >= 1 && dim >= 1, VK_EINVAL, "bad SP batch shape");
  VK_REQUIRE(dim <= SP_MAX_DIM, VK_EINVAL, "vector SPs through vk_sp_* are limited to 64 dimensions");
  if (sp == VK_SP_LOG_CATEGORICAL) VK_REQUIRE(n_params <= SP_MAX_DIM, VK_EINVAL, "log_categorical: at most 64 outcomes here");
  if (sp == VK_SP_DIRICHLET) VK_REQUIRE(n_params == dim, VK_EINVAL, "dirichlet: n_params must equal dim");
  if (sp == VK_SP_SYM_DIR_CAT) VK_REQUIRE(n_params == dim + 1, VK_EINVAL, "sym_dir_cat: params are alpha, counts[dim]");
  if (sp == VK_SP_BETA_BERNOULLI) VK_REQUIRE(n_params == 4, VK_EINVAL, "beta_bernoulli: params are a, b, yes, no");
  if (sp == VK_SP_DIR_CAT) VK_REQUIRE(n_params == 2 * dim, VK_EINVAL, "dir_cat: params are alpha[dim], counts[dim]");
  if (sp == VK_SP_UC_DIR_CAT) VK_REQUIRE(n_params == dim, VK_EINVAL, "uc_dir_cat: params are theta[dim]");
  if (sp == VK_SP_SUFF_BERNOULLI) VK_REQUIRE(n_params >= 1, VK_EINVAL, "suff_bernoulli: params are weight[, yes, no]");
}

static int sp_device(int device, int *sm) {
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0) throw Error(VK_ECUDA, "no CUDA device available (no CPU fallback)");
  VK_REQUIRE(device >= 0 && device < n_dev, VK_EINVAL, "device index out of range");
  VK_CUDA(cudaSetDevice(device));
  VK_CUDA(cudaDeviceGetAttribute(sm, cudaDevAttrMultiProcessorCount, device));
  return 0;
}

}  // namespace vk

using namespace vk;

extern "C" {

int vk_sp_simulate(int32_t sp, const double *params_host, int32_t n_params, int64_t n, int32_t dim,
                   uint64_t seed, uint32_t step, int device, double *values_out_host) {
  try {
    sp_check(sp, n_params, n, dim);
    VK_REQUIRE(params_host && values_out_host, VK_EINVAL, "NULL buffer");
    if (n == 0) return VK_OK;
    int sm = 148;
    sp_device(device, &sm);
    bool vec = (sp == VK_SP_DIRICHLET || sp == VK_SP_SYM_DIRICHLET);
    size_t n_out = (size_t)n * (vec ? dim : 1);
    DevBuf<double> p, o;
    p.alloc((size_t)n * n_params); o.alloc(n_out);
    VK_CUDA(cudaMemcpy(p.p, params_host, p.bytes(), cudaMemcpyHostToDevice));
    sp_simulate_kernel<<<grid_for(n, 128, sm, 16), 128>>>(sp, p.p, n_params, n, dim, seed, step, o.p);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaMemcpy(values_out_host, o.p, o.bytes(), cudaMemcpyDeviceToHost));
  } catch (const vk::Error &e) { vk::set_last_error(e.what()); return e.code; }
  return VK_OK;
}

int vk_sp_log_density(int32_t sp, const double *values_host, const double *params_host, int32_t n_params,
                      int64_t n, int32_t dim, int device, double *out_host) {
  try {
    sp_check(sp, n_params, n, dim);
    VK_REQUIRE(params_host && values_host && out_host, VK_EINVAL, "NULL buffer");
    if (n == 0) return VK_OK;
    int sm = 148;
    sp_device(device, &sm);
    bool vec = (sp == VK_SP_DIRICHLET || sp == VK_SP_SYM_DIRICHLET);
    DevBuf<double> p, v, o;
    p.alloc((size_t)n * n_params); v.alloc((size_t)n * (vec ? dim : 1)); o.alloc((size_t)n);
    VK_CUDA(cudaMemcpy(p.p, params_host, p.bytes(), cudaMemcpyHostToDevice));
    VK_CUDA(cudaMemcpy(v.p, values_host, v.bytes(), cudaMemcpyHostToDevice));
    sp_log_density_kernel<<<grid_for(n, 128, sm, 16), 128>>>(sp, v.p, p.p, n_params, n, dim, o.p);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaMemcpy(out_host, o.p, o.bytes(), cudaMemcpyDeviceToHost));
  } catch (const vk::Error &e) { vk::set_last_error(e.what()); return e.code; }
  return VK_OK;
}

int vk_sp_log_density_of_data(int32_t sp, const double *params_host, int32_t n_params, int64_t n, int32_t dim,
                              int device, double *out_host) {
  try {
    VK_REQUIRE(sp == VK_SP_SYM_DIR_CAT || sp == VK_SP_DIR_CAT || sp == VK_SP_BETA_BERNOULLI || sp == VK_SP_SUFF_BERNOULLI,
               VK_EINVAL, "logDensityOfData: SP keeps no sufficient statistics");
    VK_REQUIRE(n >= 0 && dim >= 1 && dim <= SP_MAX_DIM && params_host && out_host, VK_EINVAL, "bad SP batch");
    VK_REQUIRE(n_params == (sp == VK_SP_SYM_DIR_CAT ? dim + 1 : sp == VK_SP_DIR_CAT ? 2 * dim : sp == VK_SP_BETA_BERNOULLI ? 4 : 3),
               VK_EINVAL, "logDensityOfData: wrong number of parameters per row");
    if (n == 0) return VK_OK;
    int sm = 148;
    sp_device(device, &sm);
    DevBuf<double> p, o;
    p.alloc((size_t)n * n_params); o.alloc((size_t)n);
    VK_CUDA(cudaMemcpy(p.p, params_host, p.bytes(), cudaMemcpyHostToDevice));
    sp_log_density_of_data_kernel<<<grid_for(n, 128, sm, 16), 128>>>(sp, p.p, n_params, n, dim, o.p);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaMemcpy(out_host, o.p, o.bytes(), cudaMemcpyDeviceToHost));
  } catch (const vk::Error &e) { vk::set_last_error(e.what()); return e.code; }
  return VK_OK;
}

int vk_sp_aaa_simulate(int32_t sp, const double *params_host, int32_t n_params, int64_t n, int32_t dim, uint64_t seed,
                       uint32_t step, int device, double *values_out_host, double *weights_out_host) {
  try {
    VK_REQUIRE(sp == VK_SP_UC_DIR_CAT || sp == VK_SP_SYM_DIRICHLET || sp == VK_SP_SUFF_BERNOULLI, VK_EINVAL,
               "aaa_simulate: not an uncollapsed maker");
    VK_REQUIRE(n >= 0 && dim >= 1 && dim <= SP_MAX_DIM && params_host && values_out_host, VK_EINVAL, "bad SP batch");
    VK_REQUIRE(n_params == (sp == VK_SP_UC_DIR_CAT ? 2 * dim : sp == VK_SP_SYM_DIRICHLET ? dim + 1 : 4), VK_EINVAL,
               "aaa_simulate: wrong number of parameters per row");
    if (n == 0) return VK_OK;
    int sm = 148;
    sp_device(device, &sm);
    const size_t n_out = (size_t)n * (sp == VK_SP_SUFF_BERNOULLI ? 1 : dim);
    DevBuf<double> p, o, w;
    p.alloc((size_t)n * n_params); o.alloc(n_out); w.alloc((size_t)n);
    VK_CUDA(cudaMemcpy(p.p, params_host, p.bytes(), cudaMemcpyHostToDevice));
    sp_aaa_simulate_kernel<<<grid_for(n, 128, sm, 16), 128>>>(sp, p.p, n_params, n, dim, seed, step, o.p, w.p);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaMemcpy(values_out_host, o.p, o.bytes(), cudaMemcpyDeviceToHost));
    if (weights_out_host) VK_CUDA(cudaMemcpy(weights_out_host, w.p, w.bytes(), cudaMemcpyDeviceToHost));
  } catch (const vk::Error &e) { vk::set_last_error(e.what()); return e.code; }
  return VK_OK;
}

}  // extern "C"

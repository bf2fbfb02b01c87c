// Shared host/device plumbing of the C-ABI library: error convention,
// the particle-batch handle, device buffers, warp/block reductions.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/venture_b200.h"
#include "philox.cuh"

namespace vk {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const std::string &m);

#define VK_CUDA(expr)                                                        \
  do {                                                                       \
    cudaError_t _e = (expr);                                                 \
    if (_e != cudaSuccess) {                                                 \
      char _b[512];                                                          \
      snprintf(_b, sizeof _b, "CUDA error %s at %s:%d: %s",                  \
               cudaGetErrorName(_e), __FILE__, __LINE__,                     \
               cudaGetErrorString(_e));                                      \
      throw vk::Error(VK_ECUDA, _b);                                         \
    }                                                                        \
  } while (0)

#define VK_REQUIRE(cond, code, msg)                                          \
  do {                                                                       \
    if (!(cond)) throw vk::Error((code), (msg));                             \
  } while (0)

// RAII device array
template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  ~DevBuf() { release(); }
  void alloc(size_t count) {
    release();
    n = count;
    if (count) VK_CUDA(cudaMalloc((void **)&p, count * sizeof(T)));
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  void swap(DevBuf &o) { std::swap(p, o.p); std::swap(n, o.n); }
  // take ownership of a cudaMalloc'd array of `count` elements (frees what was held)
  void adopt(void *ptr, size_t count) { release(); p = (T *)ptr; n = count; }
  size_t bytes() const { return n * sizeof(T); }
};

// one state array that must follow its particle through resampling
struct StateArray {
  void **cur;        // address of the live device pointer
  void **alt;        // address of the back buffer pointer (same size)
  size_t row_bytes;  // bytes per particle when particle-major (rows)
  size_t n_rows;     // SoA: number of particle-minor rows of elem_bytes each
  size_t elem_bytes; // SoA element size (particle-minor layout); 0 => particle-major rows
  size_t live_bytes = 0;  // particle-major rows: leading bytes that hold data so far (0 => the whole row); a gather may skip the rest
};

struct Model;

// Per-kernel timing of a stream (VK_PROFILE=1): events between launches, resolved after a synchronisation.
// The interval that ENDS at a mark carries the mark's name; "(gap)" marks end host-side gaps.
struct StreamProfile {
  bool on = false;
  std::vector<std::pair<const char *, cudaEvent_t>> marks;
  std::vector<cudaEvent_t> pool;
  std::map<std::string, std::pair<double, long long>> totals;
  void mark(cudaStream_t st, const char *name) {
    if (!on) return;
    cudaEvent_t e;
    if (pool.empty()) cudaEventCreate(&e); else { e = pool.back(); pool.pop_back(); }
    cudaEventRecord(e, st);
    marks.push_back(std::make_pair(name, e));
  }
  void resolve() {                       // call after the stream drained
    if (!on) return;
    for (size_t i = 1; i < marks.size(); ++i) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second) == cudaSuccess) {
        auto &t = totals[marks[i].first];
        t.first += ms; t.second += 1;
      }
    }
    if (!marks.empty()) {
      for (size_t i = 0; i + 1 < marks.size(); ++i) pool.push_back(marks[i].second);
      marks.erase(marks.begin(), marks.end() - 1);
    }
  }
  void report(const char *who) {
    if (!on || totals.empty()) return;
    fprintf(stderr, "[vk profile %s] interval: count, mean us\n", who);
    for (auto &kv : totals)
      fprintf(stderr, "[vk profile %s]   %-28s %8lld %10.2f\n", who, kv.first.c_str(), kv.second.second,
              1e3 * kv.second.first / (double)kv.second.second);
  }
};

}  // namespace vk

// The opaque handle of include/venture_b200.h.
struct vk_particles {
  int kind = 0;
  int device = 0;
  int sm_count = 148;
  int64_t P = 0;
  int64_t p_base = 0;   // global index of local particle 0 (multi-GPU sharding)
  uint64_t seed = 0;
  uint32_t step = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  vk::DevBuf<double> lw;        // log weights, P
  vk::DevBuf<double> scratch_d; // P doubles (increments, probs)
  vk::DevBuf<double> out_d;     // grow-only staging for host outputs
  vk::DevBuf<long long> acc_d;  // accept counters staging (vk_mh)
  vk::DevBuf<uint64_t> cdf;     // fixed-point inclusive weight prefix sums, P
  vk::DevBuf<int32_t> anc;      // ancestors, P
  vk::DevBuf<double> red;       // small reduction scratch
  vk::DevBuf<uint64_t> red_u;   // scan block sums
  vk::DevBuf<double> tile_exp;  // per-tile sums of exp(lw - max)
  vk::DevBuf<unsigned int> counters;
  vk::Model *model = nullptr;
  // red[0..2] = (max, min, NaN flag) of lw as it is NOW: set by a model whose make_consistent reduced the
  // weights it wrote, cleared at the entry of every API call; consumed by vk_resample / vk_p2p_resample
  bool lw_minmax_fresh = false;
  // ... and they describe the weights of ALL ranks: the producer's last block already ran exchange 0 of the
  // peer-memory resample (p2p_common.cuh); same lifetime as lw_minmax_fresh
  bool lw_x0_done = false;
  void *p2p = nullptr;          // vk::P2P (p2p_resample.cuh): peer-memory resampling state
  vk::StreamProfile prof;       // VK_PROFILE=1
  uint32_t next_step(uint32_t n = 1) { uint32_t s = step; step += n; return s; }
};

namespace vk {

// Per-model-kind behaviour behind the C ABI.
struct Model {
  vk_particles *h;
  explicit Model(vk_particles *h_) : h(h_) {}
  virtual ~Model() {}
  virtual void make_consistent(double *incr_dev) = 0;   // adds to h->lw, writes increments
  virtual void likelihood_weight() = 0;                 // overwrites h->lw
  virtual void reset_to_prior() = 0;                    // resimulate latents, weights untouched
  virtual void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) = 0;
  // enumerative Gibbs over discrete sites (infer/egibbs.py); changed_dev counts transitions that moved
  virtual void gibbs(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) {
    throw Error(VK_EUNSUPPORTED, "gibbs: this model has no enumerable sites compiled");
  }
  // `(emap scope block K)`: as gibbs, the candidate of largest weight instead of a draw (infer/egibbs.py:117-121)
  virtual void emap(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) {
    throw Error(VK_EUNSUPPORTED, "emap: compiled for the boolean sites of programs of the model compiler only");
  }
  // particle Gibbs over an ordered scope (infer/pgibbs.py): conditional SMC with `particles` inner particles
  virtual void pgibbs(int scope, int block, int particles, int K, double *mean_nodes_dev, long long *acc_dev) {
    throw Error(VK_EUNSUPPORTED, "pgibbs: compiled for the HMM model's ordered state sequence only");
  }
  // slice sampling with stepping out on a continuous one-dimensional site (infer/slice_sample.py); moved_dev counts
  // transitions that changed the value
  virtual void slice(int scope, int block, double w, double m, int doubling, int K, double *mean_nodes_dev, long long *moved_dev) {
    throw Error(VK_EUNSUPPORTED, "slice: compiled for programs of the model compiler only (vk_create_program)");
  }
  virtual void log_likelihood_at(int scope, int block, bool joint, double *out_dev) {
    throw Error(VK_EUNSUPPORTED, "log_likelihood_at/log_joint_at not defined for this model");
  }
  virtual void subsampled_mh(int scope, int block, int64_t Nbatch, int k0, double eps,
                             int use_delta, double delta_sigma, int update_values, int K,
                             double *mean_nodes_dev, double *stats_dev) {
    throw Error(VK_EUNSUPPORTED, "subsampled_mh is compiled for the logistic-regression model and for programs of the model compiler");
  }
  virtual void forget_observations(double *w_dev) { throw Error(VK_EUNSUPPORTED, "forget not defined for this model"); }
  virtual int64_t num_sites() const = 0;
  virtual int64_t num_blocks(int scope) const = 0;
  virtual void read_site(int64_t site, double *out_dev) = 0;
  virtual void write_site(int64_t site, const double *in_dev) = 0;
  virtual void refresh() {}
  // VK_FP64 (parity, default) or VK_FP32 (throughput): models without an fp32 kernel refuse VK_FP32
  virtual void set_precision(int prec) {
    if (prec != VK_FP64) throw Error(VK_EUNSUPPORTED, "this model kind has no fp32 mode in this version");
  }
  virtual int get_precision() const { return VK_FP64; }
  // arrays gathered by ancestor index on resample; after the gather the model
  // swaps cur/alt itself via StateArray.  alloc_alt = false: do not allocate back buffers that do not exist
  // yet (the caller installs its own: resample with n_new != P).
  virtual std::vector<StateArray> state_arrays(bool alloc_alt = true) = 0;
  // the same description with NO side effect at all (validation before anything is committed)
  virtual std::vector<StateArray> state_arrays_peek() { return state_arrays(false); }
  // resample changed the particle count: the live arrays are new_cur[i] and the back buffers new_alt[i]
  // (cudaMalloc'd for n_new particles, in state_arrays() order; *cur / *alt already point at new_cur[i] / stale
  // memory).  The model frees its old buffers, adopts the new ones and drops every cache sized by the old count.
  virtual void resized(int64_t n_new, const std::vector<void *> &new_cur, const std::vector<void *> &new_alt) = 0;
  // state_arrays() for a gather that can patch one byte into every row it copies: a model whose newest
  // column is still pending in a particle-minor array (HMM: hist[.][T-1] lives in cur[]) reports
  // (row array index, source SoA array index, byte offset in the row) instead of materialising the column
  // first; patch[0] < 0 = nothing to patch.  Only for callers that copy EVERY row (the peer-memory gather).
  virtual std::vector<StateArray> state_arrays_patched(int patch[3]) { patch[0] = -1; return state_arrays(); }
  // a model may gather its whole state by ancestor in one fused pass of its own; returns 0 (not done: the
  // generic per-array gather over state_arrays() runs), 1 (state gathered) or 3 (state gathered AND the
  // evened-out weights lw[k] = red[4] - log_n_new written, so the caller skips its fill pass).
  // lw_out may be null (no fill wanted).
  virtual int gather(const int32_t *anc_dev, int64_t n_new, double *lw_out, const double *red, double log_n_new) {
    (void)anc_dev; (void)n_new; (void)lw_out; (void)red; (void)log_n_new; return 0;
  }
  virtual void after_resample() {}
  virtual int device_buffer(int which, void **p, int64_t *bytes) { return VK_EINVAL; }
};

// ---- device helpers -------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum, result valid in every thread. `sh` holds >= 33 doubles.
__device__ __forceinline__ double block_sum(double v, double *sh) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect sh from a previous use
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double t = (lane < nw) ? sh[lane] : 0.0;
  t = warp_sum(t);
  return t;
}

inline int grid_for(int64_t n, int threads, int sm_count, int ctas_per_sm) {
  int64_t need = (n + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

}  // namespace vk

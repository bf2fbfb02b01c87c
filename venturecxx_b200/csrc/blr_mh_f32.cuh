// fp32 THROUGHPUT mode of the logistic-regression single-site MH loop ("the build has fp64
// (parity) and fp32 (throughput) modes"; north_star: log densities within 1e-5 relative in fp32 mode).
//
// Same site program, same Philox streams and the same fp64 proposals (site, w', log u) as blr_mh_kernel
// (infer/mh.py:84-89, 187-208; regen.py:120-127; detach.py:72-83).  What changes is the arithmetic of the
// N-observation sum  xi = sum_i log_logistic(eta_i + delta x_ij):
//   * the logit row is staged in shared memory as float (4N bytes: 4 CTAs per SM instead of 2),
//   * Xs is read from a float copy (half the L2 traffic per transition),
//   * log_logistic(v) = min(v,0) - ln2 * lg2(1 + ex2(-|v| log2 e)) on the MUFU pipe; the four terms of a
//     float4 share ONE lg2: sum_k lg2(1 + t_k) = lg2(prod_k (1 + t_k)), the product is in [1,16]
//     (1.25 MUFU + 7 fp32 instructions per observation instead of ~20 fp64 instructions and two table
//     lookups; the first version spent 2 MUFU per observation and ran at 72 % of the MUFU pipe),
//   * each thread keeps 5 fp32 partial sums of 10 terms each (at C2); the thread's total is widened to
//     fp64 before the block-wide sum, so the rounding of the accumulation stays ~1e-8 relative to xi.
// w itself, the proposals and the accept test stay fp64.  rho is recomputed from the staged row when a
// particle is loaded, so xi and rho are the same function of nearly the same arguments and the systematic
// part of the MUFU error cancels in xi - rho.  The float row is widened back to the fp64 logit cache when
// the launch ends; the model marks the cache "low precision" and refreshes it from w (fp64) before any
// fp64 consumer (log_likelihood_at, resample, make_consistent) reads it.
#pragma once

namespace vk {

static const int BLR_F32_THREADS = 256;
static const int BLR_F32_TILE = 4 * BLR_F32_THREADS;   // one float4 per thread per iteration

__device__ __forceinline__ float f32_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f32_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// min(v,0) into m; returns 1 + 2^(-|v| log2 e) in [1,2]
__device__ __forceinline__ float f32_one_plus_exp_neg_abs(float v, float &m) {
  m += fminf(v, 0.0f);
  return 1.0f + f32_ex2(-1.4426950408889634f * fabsf(v));
}

__host__ __device__ inline size_t blr_mh_f32_smem_base(int D) {
  return (size_t)(((D + 1) & ~1) + 34 + 64 + 2 * BLR_MH_CHUNK) * sizeof(double) + (size_t)BLR_MH_CHUNK * sizeof(int);
}

// One tile: four observations of this thread.  min(v,0) into m0..m3, lg2 of the product of the four
// 1 + exp(-|v|) into l.
__device__ __forceinline__ void blr_f32_tile(const float4 e, float &m0, float &m1, float &m2, float &m3, float &l) {
  const float q0 = f32_one_plus_exp_neg_abs(e.x, m0) * f32_one_plus_exp_neg_abs(e.y, m1);
  const float q1 = f32_one_plus_exp_neg_abs(e.z, m2) * f32_one_plus_exp_neg_abs(e.w, m3);
  l += f32_lg2(q0 * q1);
}
__device__ __forceinline__ float4 blr_f32_axpy(float df, const float4 x, float4 e) {
  e.x = fmaf(df, x.x, e.x); e.y = fmaf(df, x.y, e.y); e.z = fmaf(df, x.z, e.z); e.w = fmaf(df, x.w, e.w);
  return e;
}

// sum_i log_logistic(row_i + df * x_i) over the padded row (WITH_X = false: df = 0, no x).  All threads
// return the same fp64 value.  `xn` holds this thread's first x tile on entry and the first tile of the
// NEXT transition's column (`xnext`, already offset by tid) on exit; the last iteration is peeled so the
// loop's prefetch address is a plain pointer increment.
template <bool WITH_X>
__device__ __forceinline__ double blr_f32_row_sum(const float4 *__restrict__ row, const float4 *__restrict__ xcol,
                                                  const float4 *__restrict__ xnext, float4 &xn, float df, int iters,
                                                  double *slots, int parity) {
  float m0 = 0.f, m1 = 0.f, m2 = 0.f, m3 = 0.f, l = 0.f;
  const float4 *xp = xcol + BLR_F32_THREADS;      // row, xcol, xnext are already offset by tid
  for (int it = 0; it < iters - 1; ++it, row += BLR_F32_THREADS, xp += BLR_F32_THREADS) {
    float4 e = *row;
    if (WITH_X) { const float4 x = xn; xn = __ldg(xp); e = blr_f32_axpy(df, x, e); }
    blr_f32_tile(e, m0, m1, m2, m3, l);
  }
  {
    float4 e = *row;
    if (WITH_X) { const float4 x = xn; xn = __ldg(xnext); e = blr_f32_axpy(df, x, e); }
    blr_f32_tile(e, m0, m1, m2, m3, l);
  }
  const float part = fmaf(-0.6931471805599453f, l, (m0 + m1) + (m2 + m3));
  return block_sum_1sync<BLR_F32_THREADS / 32>((double)part, slots, parity);
}

template <int MINB>
__global__ void __launch_bounds__(BLR_F32_THREADS, MINB)
blr_mh_f32_kernel(double *__restrict__ w, double *__restrict__ eta, double *__restrict__ ll,
                  const float *__restrict__ Xs32, BLRParams prm, int64_t P, int64_t p_base,
                  uint64_t seed, uint32_t step0, int n_prop_steps, int block_mode,
                  long long *__restrict__ acc_out) {
  constexpr int THREADS = BLR_F32_THREADS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int Dp = (prm.D + 1) & ~1;
  double *w_sh = reinterpret_cast<double *>(smem_raw);   // D
  double *red = w_sh + Dp;                               // 34
  double *slots = red + 34;                              // 2 x 32 per-warp partial sums
  double *prop_w = slots + 64;                           // BLR_MH_CHUNK proposed values
  double *prop_logu = prop_w + BLR_MH_CHUNK;             // BLR_MH_CHUNK log-uniforms
  int *prop_site = (int *)(prop_logu + BLR_MH_CHUNK);    // BLR_MH_CHUNK sites
  float *eta_sh = (float *)(prop_site + BLR_MH_CHUNK);   // Npad floats, 16-byte aligned
  const int N = (int)prm.N;
  const int Npad = (N + BLR_F32_TILE - 1) / BLR_F32_TILE * BLR_F32_TILE;
  const int iters = Npad / BLR_F32_TILE;
  const int tid = threadIdx.x;
  float4 *row4 = reinterpret_cast<float4 *>(eta_sh);
  const float4 *X4 = reinterpret_cast<const float4 *>(Xs32);
  const int ld4 = (int)(prm.Xld / 4);
  int parity = 0;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    __syncthreads();
    double *eta_g = eta + p * prm.Ncap;
    if (tid < prm.D) w_sh[tid] = w[(int64_t)tid * P + p];
    {
      // Ncap is even and rows are 16-byte aligned: read double2, store float2
      const double2 *src = reinterpret_cast<const double2 *>(eta_g);
      float2 *dst = reinterpret_cast<float2 *>(eta_sh);
      for (int i = tid; i < Npad / 2; i += THREADS) {
        float2 v = make_float2((float)BLR_PAD_ETA, (float)BLR_PAD_ETA);
        if (2 * i < N) { double2 d = src[i]; v.x = (float)d.x; if (2 * i + 1 < N) v.y = (float)d.y; }
        dst[i] = v;
      }
    }
    __syncthreads();
    float4 xn = make_float4(0.f, 0.f, 0.f, 0.f);
    double rho = blr_f32_row_sum<false>(row4 + tid, nullptr, nullptr, xn, 0.0f, iters, slots, parity);
    parity ^= 1;
    long long acc = 0;
    bool dirty = false;
    for (int c0 = 0; c0 < n_prop_steps; c0 += BLR_MH_CHUNK) {
      int cn = n_prop_steps - c0 < BLR_MH_CHUNK ? n_prop_steps - c0 : BLR_MH_CHUNK;
      __syncthreads();
      for (int t = tid; t < cn; t += THREADS) {
        Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)(c0 + t));
        double u_sel = r.uniform();
        double u_acc = r.uniform();
        int j;
        if (block_mode == VK_BLOCK_ONE) { j = (int)(u_sel * (double)prm.D); if (j >= prm.D) j = prm.D - 1; }
        else if (block_mode == VK_BLOCK_EACH) j = (c0 + t) % prm.D;
        else j = block_mode;
        prop_site[t] = j;
        prop_logu[t] = log(u_acc);
        prop_w[t] = normal_sim(r, prm.mu0, prm.s0);
      }
      __syncthreads();
      xn = __ldg(X4 + tid + prop_site[0] * ld4);   // column offsets in float4 units: D * Xld / 4 < 2^31
      for (int t = 0; t < cn; ++t) {
        const int j = prop_site[t];
        const int off = j * ld4;
        const double w_new = prop_w[t];
        const float df = (float)(w_new - w_sh[j]);
        const float4 *xcol4 = X4 + off;
        const float4 *xnext = X4 + tid + prop_site[t + 1 < cn ? t + 1 : t] * ld4;
        double xi = blr_f32_row_sum<true>(row4 + tid, xcol4 + tid, xnext, xn, df, iters, slots, parity);
        parity ^= 1;
        double alpha = blr_log_alpha(rho, xi);
        if (prop_logu[t] < alpha) {   // uniform across the CTA
          for (int k = tid; k < Npad / 4; k += THREADS) {
            float4 e = row4[k];
            const float4 x = __ldg(xcol4 + k);
            e.x = fmaf(df, x.x, e.x); e.y = fmaf(df, x.y, e.y); e.z = fmaf(df, x.z, e.z); e.w = fmaf(df, x.w, e.w);
            row4[k] = e;
          }
          if (tid == 0) w_sh[j] = w_new;
          rho = xi; ++acc; dirty = true;
          __syncthreads();
        }
      }
    }
    if (dirty) {
      double2 *dst = reinterpret_cast<double2 *>(eta_g);
      const float2 *src = reinterpret_cast<const float2 *>(eta_sh);
      for (int i = tid; i < (N + 1) / 2; i += THREADS) { float2 v = src[i]; dst[i] = make_double2((double)v.x, (double)v.y); }
      if (tid < prm.D) w[(int64_t)tid * P + p] = w_sh[tid];
    }
    if (tid == 0) ll[p] = rho;   // fp32-accurate; the model refreshes it before an fp64 consumer reads it
    if (acc_out && tid == 0) acc_out[p] = acc;
  }
}

__global__ void blr_to_float_kernel(const double *__restrict__ in, float *__restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = (float)in[i];
}

}  // namespace vk

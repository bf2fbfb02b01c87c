// Generic particle kernels for COMPILED programs (venturecxx_b200/compiler.py): one register
// machine interprets the site programs the compiler derived from the reference's scaffold rules
// (backend/lite/scaffold.py:149-316) for ANY fixed-structure program over the scalar SPs of SURVEY 8(a) A12 --
// no .cu file per model.
//
//   likelihood_weight   trace.py:522-530   PRIOR section (every site from its prior, in evaluation order), then
//                                          lw := sum of the observation groups' log densities
//   makeConsistent      trace.py:422-448   lw += log density of the observations (a leaf observation's rho is 0)
//   resimulation_mh     mh.py:59-113, 167-208; detach.py:39-83; regen.py:58-161
//                                          per transition and particle: block choice, rho = log densities absorbing
//                                          at the site's border under the old value, SIM, xi under the new value,
//                                          accept iff log u < xi - rho (with MHOperator's +-inf policy)
//   log_likelihood_at / log_joint_at       dispatch.py:331-361
//
// Mapping.  LANES = 1: one thread per particle (few observations).  LANES = 32: one warp per particle, lanes take
// the rows of an observation group in turn and the warp adds their sums (many observations per particle);
// every lane runs the row-independent code redundantly with the same Philox stream, so no broadcast is needed.
// Registers live in local memory (the register FILE of the machine is indexed dynamically); code, constants and
// group tables are read through the read-only path and are the same addresses across a warp (one transaction).
// RNG protocol of a transition = the hand-written kernels': Stream(seed, particle, step0 + t): draw 0 block
// choice, draw 1 accept uniform, then the proposal's draws -- a compiled normal-normal model reproduces
// nn_mh_kernel's chains bit for bit (tests/test_gpu_program.py).
//
// The same interpreter restated on the CPU is the oracle (oracle/oracle.c orc_vm_*).
#include "vk_common.cuh"
#include "sp_math.cuh"
#include "subsample_common.cuh"

namespace vk {

enum VMOp {      // dense: the interpreter's switch compiles to one jump table (mirrors compiler.py OP)
  VM_CONST = 0, VM_SITE = 1, VM_DATA = 2, VM_ADD = 3, VM_SUB = 4, VM_MUL = 5, VM_DIV = 6, VM_NEG = 7, VM_EXP = 8,
  VM_LOG = 9, VM_SQRT = 10, VM_POW = 11, VM_LOGISTIC = 12, VM_ABS = 13, VM_MIN = 14, VM_MAX = 15, VM_LOG1P = 16,
  VM_EXPM1 = 17, VM_LOG_LOGISTIC = 18, VM_LOGIT = 19, VM_STORE = 20, VM_COND = 21, VM_SELECT = 22, VM_MOV = 23,
  VM_ARGN = 24, VM_LOOKUP = 25,
  VM_SIM_NORMAL = 26, VM_SIM_GAMMA = 27, VM_SIM_BETA = 28, VM_SIM_UNIFORM = 29, VM_SIM_EXPON = 30, VM_SIM_FLIP = 31,
  VM_SIM_LOG_FLIP = 32, VM_SIM_LOG_ODDS_FLIP = 33, VM_SIM_CATEGORICAL = 34,
  VM_LD_NORMAL = 35, VM_LD_GAMMA = 36, VM_LD_BETA = 37, VM_LD_UNIFORM = 38, VM_LD_EXPON = 39, VM_LD_FLIP = 40,
  VM_LD_LOG_FLIP = 41, VM_LD_LOG_ODDS_FLIP = 42, VM_LD_CATEGORICAL = 43
};
static const int VM_WORDS = 4;      // op, dst, a, b
static const int VM_MAX_REGS = 128;
static const int VM_LEAN_REGS = 32;   // programs with at most this many registers and none of the extended instructions
static const int VM_MAX_ARGN = 32;  // outcomes of a categorical

// device-resident program (all pointers into one allocation each)
struct VMProg {
  const int32_t *code;       // VM_WORDS words per instruction
  const double *consts;
  const double *data;        // group tables, row-major
  const int32_t *dir;        // per section: first instruction, count
  const int32_t *site_tab;   // per site: SIM, CHILD, OWN section, number of groups, offset into site_groups
  const int32_t *site_groups;
  const int32_t *group_tab;  // per group: prologue section, row section, rows, columns, data offset
  const double *affected;    // per site: nodes of Lite's scaffold (scaffold.py:67-68)
  const int32_t *ss_tab;     // per site: kind (0 no border, 1 border, 2 refused), local sections, offset into ss_roots
  const int32_t *ss_roots;   // pairs: (group, row) | (-1, child site), in the order of Lite's childrenAt(globalBorder)
  const double *ss_nodes;    // per site: nodes of the scaffold cut at the global border
  const int32_t *scope_tab;  // per scope: number of blocks, offset into scope_sites (scope 0 = default: every site)
  const int32_t *scope_sites;   // the blocks' sites (one site per block)
  const int32_t *cand_tab;   // per site: enumeration candidates (count, offset into cand_consts); count 0 = not enumerable
  const int32_t *cand_consts;   // indices into consts
  int n_sites, n_groups, sec_prior, sec_joint;
};

// one straight-line section.  row: the current row of a group table (nullptr outside row programs).
// ov_site / ov_val: read site ov_site as ov_val instead of from memory (the OLD value of a proposal in flight).
template <bool FULL>
__device__ __forceinline__ void vm_exec(const VMProg &pg, int section, double *R, double *sites, int64_t P, int64_t p,
                                        const double *row, Stream *rng, double &acc, bool store,
                                        int ov_site = -1, double ov_val = 0.0) {
  const int first = __ldg(&pg.dir[2 * section]), n = __ldg(&pg.dir[2 * section + 1]);
  const int32_t *ip = pg.code + (size_t)first * VM_WORDS;
  bool flag = false;                      // (biplex p a b) = COND p; SELECT dst a b
  int argn = 0;                           // ARGN n: the length of the register block the next variadic instruction reads
  for (int i = 0; i < n; ++i, ip += VM_WORDS) {
    const int op = __ldg(ip), d = __ldg(ip + 1), a = __ldg(ip + 2), b = __ldg(ip + 3);   // (one int4 load measured 20 % slower)
    switch (op) {
      case VM_CONST: R[d] = __ldg(&pg.consts[a]); break;
      case VM_SITE: R[d] = a == ov_site ? ov_val : sites[(int64_t)a * P + p]; break;
      case VM_DATA: R[d] = __ldg(&row[a]); break;
      case VM_ADD: R[d] = R[a] + R[b]; break;
      case VM_SUB: R[d] = R[a] - R[b]; break;
      case VM_MUL: R[d] = R[a] * R[b]; break;
      case VM_DIV: R[d] = R[a] / R[b]; break;
      case VM_NEG: R[d] = -R[a]; break;
      case VM_EXP: R[d] = exp(R[a]); break;
      case VM_LOG: R[d] = vlog(R[a]); break;
      case VM_SQRT: R[d] = sqrt(R[a]); break;
      case VM_POW: R[d] = pow(R[a], R[b]); break;
      case VM_LOGISTIC: R[d] = logistic(R[a]); break;
      case VM_ABS: R[d] = fabs(R[a]); break;
      case VM_MIN: R[d] = fmin(R[a], R[b]); break;
      case VM_MAX: R[d] = fmax(R[a], R[b]); break;
      case VM_LOG1P: R[d] = vlog1p(R[a]); break;
      case VM_EXPM1: R[d] = expm1(R[a]); break;
      case VM_LOG_LOGISTIC: R[d] = log_logistic(R[a]); break;
      case VM_LOGIT: R[d] = logit(R[a]); break;
      case VM_MOV: if (FULL) { R[d] = R[a]; } break;
      case VM_ARGN: if (FULL) { argn = a; } break;
      case VM_LOOKUP: if (FULL) {                     // (lookup <array> i): block R[a .. a + argn), index R[b]
        const int j = (int)R[b];
        R[d] = (j >= 0 && j < argn && (double)j == R[b]) ? R[a + j] : NAN;
        break;
      }
      case VM_SIM_CATEGORICAL: if (FULL) {            // simulateCategorical (utils.py:101-104): normalise, one draw; block = weights, objects
        // (categorical_sim of sp_math.cuh on the normalised weights, without materialising them)
        double den = 0.0;
        for (int j = 0; j < argn; ++j) den += R[a + j];
        const double uni = 1.0 / (double)argn;
        double den2 = 0.0;
        for (int j = 0; j < argn; ++j) den2 += den > 0.0 ? R[a + j] / den : uni;
        const double u = rng->uniform();
        int idx;
        if (!(den2 > 0.0)) { idx = (int)(u * (double)argn); if (idx >= argn) idx = argn - 1; }
        else {
          const double t = u * den2;
          double run = 0.0;
          int last = 0;
          idx = -1;
          for (int j = 0; j < argn; ++j) {
            const double pj = den > 0.0 ? R[a + j] / den : uni;
            if (pj > 0.0) last = j;
            run += pj;
            if (t < run) { idx = j; break; }
          }
          if (idx < 0) idx = last;
        }
        R[d] = R[a + argn + idx];
        break;
      }
      case VM_LD_CATEGORICAL: if (FULL) {             // logDensityCategorical (utils.py:106-118): the mass of the objects equal to the value
        double den = 0.0, pm = 0.0;
        for (int j = 0; j < argn; ++j) den += R[a + j];
        for (int j = 0; j < argn; ++j)
          if (R[a + argn + j] == R[d]) pm += den > 0.0 ? R[a + j] / den : 1.0 / (double)argn;
        acc += pm == 0.0 ? -INFINITY : log(pm);
        break;
      }
      case VM_COND: if (FULL) { flag = R[a] != 0.0; } break;
      case VM_SELECT: if (FULL) { R[d] = flag ? R[a] : R[b]; } break;
      case VM_SIM_NORMAL: R[d] = normal_sim(*rng, R[a], R[b]); break;
      case VM_SIM_GAMMA: R[d] = gamma_sim(*rng, R[a], R[b]); break;
      case VM_SIM_BETA: R[d] = beta_sim(*rng, R[a], R[b]); break;
      case VM_SIM_UNIFORM: R[d] = uniform_sim(*rng, R[a], R[b]); break;
      case VM_SIM_EXPON: R[d] = expon_sim(*rng, R[a]); break;
      case VM_SIM_FLIP: R[d] = flip_sim(*rng, R[a]) ? 1.0 : 0.0; break;
      case VM_SIM_LOG_FLIP: R[d] = log_flip_sim(*rng, R[a]) ? 1.0 : 0.0; break;
      case VM_SIM_LOG_ODDS_FLIP: R[d] = log_odds_flip_sim(*rng, R[a]) ? 1.0 : 0.0; break;
      case VM_LD_NORMAL: acc += normal_logpdf(R[d], R[a], R[b]); break;
      case VM_LD_GAMMA: acc += gamma_logpdf(R[d], R[a], R[b]); break;
      case VM_LD_BETA: acc += beta_logpdf(R[d], R[a], R[b]); break;
      case VM_LD_UNIFORM: acc += uniform_logpdf(R[d], R[a], R[b]); break;
      case VM_LD_EXPON: acc += expon_logpdf(R[d], R[a]); break;
      case VM_LD_FLIP: acc += flip_logpmf(R[d] != 0.0, R[a]); break;
      case VM_LD_LOG_FLIP: acc += log_flip_logpmf(R[d] != 0.0, R[a]); break;
      case VM_LD_LOG_ODDS_FLIP: acc += log_odds_flip_logpmf(R[d] != 0.0, R[a]); break;
      case VM_STORE: if (store) sites[(int64_t)a * P + p] = R[d]; break;
      default: break;
    }
  }
}

// log density of group g's observations (prologue once, then the rows); LANES lanes share the rows
template <int LANES, bool FULL>
__device__ __forceinline__ double vm_group(const VMProg &pg, int g, double *R, double *sites, int64_t P, int64_t p, int lane) {
  const int32_t *gt = pg.group_tab + 5 * g;
  const int sec_pro = __ldg(gt), sec_row = __ldg(gt + 1), n_rows = __ldg(gt + 2), n_cols = __ldg(gt + 3), off = __ldg(gt + 4);
  double dummy = 0.0;
  vm_exec<FULL>(pg, sec_pro, R, sites, P, p, nullptr, nullptr, dummy, false);
  double part = 0.0;
  for (int r = lane; r < n_rows; r += LANES)
    vm_exec<FULL>(pg, sec_row, R, sites, P, p, pg.data + off + (size_t)r * n_cols, nullptr, part, false);
  if (LANES == 32) part = warp_sum(part);
  return part;
}

// what absorbs at site k's border: child sites (CHILD section) + the observation groups below k
template <int LANES, bool FULL>
__device__ __forceinline__ double vm_children(const VMProg &pg, int k, double *R, double *sites, int64_t P, int64_t p, int lane) {
  const int32_t *st = pg.site_tab + 5 * k;
  double acc = 0.0;
  vm_exec<FULL>(pg, __ldg(st + 1), R, sites, P, p, nullptr, nullptr, acc, false);
  const int ng = __ldg(st + 3), go = __ldg(st + 4);
  for (int j = 0; j < ng; ++j) acc += vm_group<LANES, FULL>(pg, __ldg(&pg.site_groups[go + j]), R, sites, P, p, lane);
  return acc;
}

template <int LANES, bool FULL>
__device__ __forceinline__ double vm_all_groups(const VMProg &pg, double *R, double *sites, int64_t P, int64_t p, int lane) {
  double acc = 0.0;
  for (int g = 0; g < pg.n_groups; ++g) acc += vm_group<LANES, FULL>(pg, g, R, sites, P, p, lane);
  return acc;
}

__device__ __forceinline__ double vm_log_alpha(double rho, double xi) {      // mh.py:187-208
  if (rho == -INFINITY && xi == -INFINITY) return INFINITY;
  if (rho == INFINITY && xi == INFINITY) return -INFINITY;
  return xi - rho;
}

enum { VM_MODE_PRIOR = 0, VM_MODE_LW = 1, VM_MODE_CONSISTENT = 2, VM_MODE_LOGLIK = 3, VM_MODE_RESIM = 4 };

// mode PRIOR: sites from the prior.  LW: the same, lw := log density of the observations.  CONSISTENT: lw += that,
// increments out.  LOGLIK: out := xi of the scaffold of `block` (VK_BLOCK_ALL: every observation), + the
// principal nodes' own densities when joint.
template <int LANES, bool FULL>
__global__ void __launch_bounds__(256, 4)
prog_kernel(VMProg pg, double *sites, double *lw, double *out, int64_t P, int64_t p_base, uint64_t seed, uint32_t step,
            int mode, int block, int joint) {
  double R[FULL ? VM_MAX_REGS : VM_LEAN_REGS];
  const int lane = LANES == 32 ? (threadIdx.x & 31) : 0;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES, n_units = (int64_t)gridDim.x * blockDim.x / LANES;
  for (int64_t p = unit; p < P; p += n_units) {
    double dummy = 0.0;
    if (mode == VM_MODE_PRIOR || mode == VM_MODE_LW) {
      Stream r(seed, (uint64_t)(p_base + p), step);
      vm_exec<FULL>(pg, pg.sec_prior, R, sites, P, p, nullptr, &r, dummy, lane == 0);
      if (LANES == 32) __syncwarp();
    }
    if (mode == VM_MODE_RESIM) {              // sites block.. from their priors GIVEN the current values of the earlier ones
      Stream r(seed, (uint64_t)(p_base + p), step);
      for (int k = block; k < pg.n_sites; ++k) {
        vm_exec<FULL>(pg, __ldg(&pg.site_tab[5 * k]), R, sites, P, p, nullptr, &r, dummy, lane == 0);
        if (LANES == 32) __syncwarp();
      }
      continue;
    }
    if (mode == VM_MODE_PRIOR) continue;
    double v;
    if (mode == VM_MODE_LOGLIK && block >= 0) {
      v = vm_children<LANES, FULL>(pg, block, R, sites, P, p, lane);
      if (joint) vm_exec<FULL>(pg, __ldg(&pg.site_tab[5 * block + 2]), R, sites, P, p, nullptr, nullptr, v, false);
    } else {
      v = vm_all_groups<LANES, FULL>(pg, R, sites, P, p, lane);
      if (mode == VM_MODE_LOGLIK && joint) vm_exec<FULL>(pg, pg.sec_joint, R, sites, P, p, nullptr, nullptr, v, false);
    }
    if (lane != 0) continue;
    if (mode == VM_MODE_LW) lw[p] = v;
    else if (mode == VM_MODE_CONSISTENT) {
      if (v != v) v = -INFINITY;              // lite/trace.py:438-448
      lw[p] += v;
      if (out) out[p] = v;
    } else out[p] = v;
  }
}

template <int LANES, bool FULL>
__global__ void __launch_bounds__(256, 4)
prog_mh_kernel(VMProg pg, double *sites, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps, int scope,
               int block_mode, double *nodes_out, long long *acc_out) {
  double R[FULL ? VM_MAX_REGS : VM_LEAN_REGS];
  const int lane = LANES == 32 ? (threadIdx.x & 31) : 0;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES, n_units = (int64_t)gridDim.x * blockDim.x / LANES;
  const int S = __ldg(&pg.scope_tab[2 * scope]);                       // blocks of the scope (one site each)
  const int32_t *sc_sites = pg.scope_sites + __ldg(&pg.scope_tab[2 * scope + 1]);
  for (int64_t p = unit; p < P; p += n_units) {
    long long acc = 0;
    double nodes = 0.0;
    for (int t = 0; t < n_steps; ++t) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      const double u_sel = r.uniform();                   // draw 0: block choice
      const double logu = log(r.uniform());               // draw 1: accept uniform (mh.py:85)
      int k;
      if (block_mode == VK_BLOCK_ONE) { k = (int)(u_sel * (double)S); if (k >= S) k = S - 1; k = __ldg(&sc_sites[k]); }
      else if (block_mode == VK_BLOCK_EACH) k = __ldg(&sc_sites[t % S]);
      else k = block_mode;
      const double rho = vm_children<LANES, FULL>(pg, k, R, sites, P, p, lane);
      const double old = sites[(int64_t)k * P + p];
      if (LANES == 32) __syncwarp();                      // every lane has read the old value
      double dummy = 0.0;
      vm_exec<FULL>(pg, __ldg(&pg.site_tab[5 * k]), R, sites, P, p, nullptr, &r, dummy, lane == 0);
      if (LANES == 32) __syncwarp();                      // the new value is in memory for every lane
      const double xi = vm_children<LANES, FULL>(pg, k, R, sites, P, p, lane);
      if (logu < vm_log_alpha(rho, xi)) ++acc;
      else { if (LANES == 32) __syncwarp(); if (lane == 0) sites[(int64_t)k * P + p] = old; if (LANES == 32) __syncwarp(); }
      nodes += __ldg(&pg.affected[k]);
    }
    if (lane == 0) {
      if (acc_out) acc_out[p] = acc;
      // `one`: the mean over the transitions of the chosen sites' counts (dispatch.py:109-120); other block modes
      // have a particle-independent value, filled by the host
      if (nodes_out && block_mode == VK_BLOCK_ONE) nodes_out[p] = n_steps > 0 ? nodes / (double)n_steps : 0.0;
    }
  }
}

// `(slice scope block w m transitions)`: StepOutSliceOperator under mixMH (infer/slice_sample.py:39-139, dispatch.py:190-194).
// f(x) = the regen weight of the scaffold with the site forced to x (DeterministicLKernel): the site's own log density
// + everything absorbing at its border (slice_sample.py:26-37).  Draw layout of a transition: draw 0 block choice, then
// in call order: the slice height (py_rng.uniform(0,1), :72), the interval position U and the step split V (:98-103),
// one uniform per shrink step (:49-58), the accept uniform of mixMH (mh.py:85; logAlpha is 0, :85-90).
template <int LANES, bool FULL>
__global__ void __launch_bounds__(256, 4)
prog_slice_kernel(VMProg pg, double *sites, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps, int scope,
                  int block_mode, double w, double m, int doubling, double *nodes_out, long long *moved_out) {
  double R[FULL ? VM_MAX_REGS : VM_LEAN_REGS];
  const int lane = LANES == 32 ? (threadIdx.x & 31) : 0;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES, n_units = (int64_t)gridDim.x * blockDim.x / LANES;
  const int S = __ldg(&pg.scope_tab[2 * scope]);                       // blocks of the scope (one site each)
  const int32_t *sc_sites = pg.scope_sites + __ldg(&pg.scope_tab[2 * scope + 1]);
  for (int64_t p = unit; p < P; p += n_units) {
    long long moved = 0;
    double nodes = 0.0;
    for (int t = 0; t < n_steps; ++t) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      const double u_sel = r.uniform();
      int k;
      if (block_mode == VK_BLOCK_ONE) { k = (int)(u_sel * (double)S); if (k >= S) k = S - 1; k = __ldg(&sc_sites[k]); }
      else if (block_mode == VK_BLOCK_EACH) k = __ldg(&sc_sites[t % S]);
      else k = block_mode;
      double *cell = sites + (int64_t)k * P + p;
      const int sec_own = __ldg(&pg.site_tab[5 * k + 2]);
      auto f = [&](double x) -> double {
        if (LANES == 32) __syncwarp();
        if (lane == 0) *cell = x;
        if (LANES == 32) __syncwarp();
        double v = vm_children<LANES, FULL>(pg, k, R, sites, P, p, lane);
        vm_exec<FULL>(pg, sec_own, R, sites, P, p, nullptr, nullptr, v, false);
        return v;
      };
      const double x0 = *cell;
      const double f0 = f(x0);
      const double logy = f0 + log(r.uniform());
      const double U = r.uniform();
      double L = x0 - w * U, Rr = L + w;
      if (doubling) {                                      // DoublingSliceOperator.findInterval (:155-176), m = doublings p
        double fl = f(L), fr = f(Rr);
        for (double K = m; K > 0.0; K -= 1.0) {
          if (logy >= fl && logy >= fr) break;
          if (fl != fl || fr != fr) break;
          const double dist = Rr - L;
          if (r.uniform() < 0.5) { L -= dist; fl = f(L); }
          else { Rr += dist; fr = f(Rr); }
        }
      } else {
        const double V = r.uniform();
        double J = floor(m * V), K = (m - 1.0) - J;
        for (int it = 1; J > 0.0 && it < 10000; ++it) {
          const double fl = f(L);
          if (logy >= fl || fl != fl) break;
          L -= w; J -= 1.0;
        }
        for (int it = 1; K > 0.0 && it < 10000; ++it) {
          const double fr = f(Rr);
          if (logy >= fr || fr != fr) break;
          Rr += w; K -= 1.0;
        }
      }
      // DoublingSliceOperator.legalProposal (:178-196): would the doubling started at x1 have produced this interval
      auto legal = [&](double xc, double Ll, double Rl) -> bool {
        bool D = false;
        double fr = f(Rl), fl = f(Ll);
        while ((Rl - Ll) > 1.1 * w) {
          const double M = (Ll + Rl) / 2.0;
          if ((x0 < M && xc >= M) || (x0 >= M && xc < M)) D = true;
          if (xc < M) { Rl = M; fr = f(Rl); } else { Ll = M; fl = f(Ll); }
          if (D && logy >= fl && logy >= fr) return false;
        }
        return true;
      };
      double x1 = x0;
      for (int it = 1; it < 10000; ++it) {
        const double Us = r.uniform();
        const double xc = L + Us * (Rr - L);
        const double fx = f(xc);
        if (logy <= fx && (!doubling || legal(xc, L, Rr))) { x1 = xc; break; }
        if (xc < x0) L = xc; else Rr = xc;
      }
      const double logu = log(r.uniform());
      if (!(logu < 0.0)) x1 = x0;                          // alpha = 0 (slice_sample.py:85-90): accepted unless u == 1
      if (LANES == 32) __syncwarp();
      if (lane == 0) *cell = x1;
      if (LANES == 32) __syncwarp();
      if (x1 != x0) ++moved;
      nodes += __ldg(&pg.affected[k]);
    }
    if (lane == 0) {
      if (moved_out) moved_out[p] = moved;
      if (nodes_out && block_mode == VK_BLOCK_ONE) nodes_out[p] = n_steps > 0 ? nodes / (double)n_steps : 0.0;
    }
  }
}

// `(subsampled_mh scope block Nbatch k0 epsilon ...)` for a compiled program: subsampledMixMH + sequentialTest
// (infer/subsampled_mh.py:76-171).  The compiler found every site's global border and local sections
// (findGlobalBorder :326-368; compiler.py _subsample_site): a section is one row of an observation group or one
// latent child, and dllh_i = its log density under the proposed value - under the old one (evalOneLocalSection
// :377-409 refreshes the section to the old value first, so nothing cached is read: Lite followed by makeConsistent).
// Two register files: Ro holds the groups' row-independent nodes under the old value, Rn under the new one; a row
// program runs once on each.  Draws: 0 block choice, 1 accept uniform, then the proposal; the sections are visited in
// the keyed Feistel order of (seed, step) -- the same order for every particle, as in the logistic-regression kernels.
// LANES = 32: lanes take the positions of a minibatch in turn and the warp adds (sum, sum of squares).
template <int LANES, bool FULL>
__global__ void __launch_bounds__(256, 4)
prog_ssmh_kernel(VMProg pg, double *sites, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps, int scope,
                 int block_mode, double Nbatch, int k0, double eps, double *nodes_out, double *stats_out) {
  double Ro[FULL ? VM_MAX_REGS : VM_LEAN_REGS], Rn[FULL ? VM_MAX_REGS : VM_LEAN_REGS];
  const int lane = LANES == 32 ? (threadIdx.x & 31) : 0;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES, n_units = (int64_t)gridDim.x * blockDim.x / LANES;
  const int S = __ldg(&pg.scope_tab[2 * scope]);                       // blocks of the scope (one site each)
  const int32_t *sc_sites = pg.scope_sites + __ldg(&pg.scope_tab[2 * scope + 1]);
  for (int64_t p = unit; p < P; p += n_units) {
    double accepted = 0.0, visited = 0.0, tlast = 0.0, nodes = 0.0;
    for (int t = 0; t < n_steps; ++t) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      const double u_sel = r.uniform();
      const double logu = log(r.uniform());
      int k;
      if (block_mode == VK_BLOCK_ONE) { k = (int)(u_sel * (double)S); if (k >= S) k = S - 1; k = __ldg(&sc_sites[k]); }
      else if (block_mode == VK_BLOCK_EACH) k = __ldg(&sc_sites[t % S]);
      else k = block_mode;
      const int kind = __ldg(&pg.ss_tab[3 * k]), Nloc = __ldg(&pg.ss_tab[3 * k + 1]), off = __ldg(&pg.ss_tab[3 * k + 2]);
      const int32_t *st = pg.site_tab + 5 * k;
      nodes += __ldg(&pg.ss_nodes[k]);
      double *cell = sites + (int64_t)k * P + p;
      const double old = *cell;
      double dummy = 0.0;
      bool accept;
      if (kind != 1) {                                     // no global border: regular MH, alpha > log u (:91-95)
        const double rho = vm_children<LANES, FULL>(pg, k, Ro, sites, P, p, lane);
        if (LANES == 32) __syncwarp();
        vm_exec<FULL>(pg, __ldg(st), Ro, sites, P, p, nullptr, &r, dummy, lane == 0);
        if (LANES == 32) __syncwarp();
        const double xi = vm_children<LANES, FULL>(pg, k, Ro, sites, P, p, lane);
        accept = (xi - rho) > logu;
      } else {
        const int ng = __ldg(st + 3), go = __ldg(st + 4);
        for (int j = 0; j < ng; ++j)
          vm_exec<FULL>(pg, __ldg(&pg.group_tab[5 * __ldg(&pg.site_groups[go + j])]), Ro, sites, P, p, nullptr, nullptr, dummy, false);
        if (LANES == 32) __syncwarp();
        vm_exec<FULL>(pg, __ldg(st), Rn, sites, P, p, nullptr, &r, dummy, lane == 0);      // the proposal: SIM_k
        if (LANES == 32) __syncwarp();
        for (int j = 0; j < ng; ++j)
          vm_exec<FULL>(pg, __ldg(&pg.group_tab[5 * __ldg(&pg.site_groups[go + j])]), Rn, sites, P, p, nullptr, nullptr, dummy, false);
        const double N = (double)Nloc;
        const double mu0 = (logu - 0.0) / N;               // alpha of the global section: nothing absorbs in it
        const FeistelKey key = make_feistel_key(seed, step0 + (uint32_t)t, (uint64_t)Nloc);
        double n = 0.0, kk = 0.0, cum = 0.0, cum2 = 0.0, ts_t = 0.0;
        accept = false;
        while (n < N) {
          const double n_end = fmin(n + Nbatch, N);
          double s1 = 0.0, s2 = 0.0;
          for (int64_t q = (int64_t)n + lane; q < (int64_t)n_end; q += LANES) {
            const int64_t i = (int64_t)feistel_perm((uint64_t)q, key);
            const int g = __ldg(&pg.ss_roots[2 * (off + i)]), rr = __ldg(&pg.ss_roots[2 * (off + i) + 1]);
            double a = 0.0, b = 0.0;
            if (g >= 0) {
              const int32_t *gt = pg.group_tab + 5 * g;
              const double *row = pg.data + __ldg(gt + 4) + (size_t)rr * __ldg(gt + 3);
              vm_exec<FULL>(pg, __ldg(gt + 1), Rn, sites, P, p, row, nullptr, a, false);
              vm_exec<FULL>(pg, __ldg(gt + 1), Ro, sites, P, p, row, nullptr, b, false);
            } else {                                       // a latent child: its own density given its parents
              const int sec = __ldg(&pg.site_tab[5 * rr + 2]);
              vm_exec<FULL>(pg, sec, Rn, sites, P, p, nullptr, nullptr, a, false);
              vm_exec<FULL>(pg, sec, Ro, sites, P, p, nullptr, nullptr, b, false, k, old);
            }
            const double d = a - b;
            s1 += d; s2 += d * d;
          }
          if (LANES == 32) { s1 = warp_sum(s1); s2 = warp_sum(s2); }
          cum += s1; cum2 += s2; kk += 1.0; n = n_end;
          const double mx = cum / n, mx2 = cum2 / n;
          if (kk < (double)k0 && n < N) continue;          // no test during the first k0 minibatches (:146-148)
          if (n == N) { accept = mx >= mu0; break; }
          const double sx = sqrt((1.0 - (n - 1.0) / (N - 1.0)) * (mx2 - mx * mx) / (n - 1.0));
          double q;
          if (sx == 0.0) q = mx >= mu0 ? 1.0 : 0.0;
          else { ts_t = (mx - mu0) / sx; q = dev_student_t_cdf(ts_t, n - 1.0); }
          if (q < eps) { accept = false; break; }
          if (q > 1.0 - eps) { accept = true; break; }
        }
        visited += n;
        tlast = ts_t != 0.0 ? ts_t : (accept ? INFINITY : -INFINITY);      // :170
      }
      if (accept) accepted += 1.0;
      else { if (LANES == 32) __syncwarp(); if (lane == 0) *cell = old; if (LANES == 32) __syncwarp(); }
    }
    if (lane == 0) {
      if (stats_out) {
        stats_out[3 * p] = accepted;
        stats_out[3 * p + 1] = n_steps > 0 ? visited / (double)n_steps : 0.0;
        stats_out[3 * p + 2] = tlast;
      }
      if (nodes_out) nodes_out[p] = n_steps > 0 ? nodes / (double)n_steps : 0.0;
    }
  }
}

// `(gibbs scope block transitions)` on a BOOLEAN site: EnumerativeGibbsOperator under mixMH (infer/egibbs.py:42-107).  The
// candidates are [true, false] (discrete.py:78-86, 121-125, 156-164; a candidate of probability 0 is left out there and gets
// weight 0 here: the same choice); each is weighted by the regen weight with the value forced (DeterministicLKernel: the
// site's own log density + what absorbs at its border), the new value is drawn by sampleLogCategorical (utils.py:170-180:
// exp(xi - max), normalised, one uniform = draw 2) and always accepted (logAlpha = 0; draw 1 is the accept uniform).
template <int LANES, bool FULL>
__global__ void __launch_bounds__(256, 4)
prog_gibbs_kernel(VMProg pg, double *sites, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps, int scope,
                  int block_mode, int map, double *nodes_out, long long *changed_out) {
  double R[FULL ? VM_MAX_REGS : VM_LEAN_REGS];
  const int lane = LANES == 32 ? (threadIdx.x & 31) : 0;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES, n_units = (int64_t)gridDim.x * blockDim.x / LANES;
  const int S = __ldg(&pg.scope_tab[2 * scope]);
  const int32_t *sc_sites = pg.scope_sites + __ldg(&pg.scope_tab[2 * scope + 1]);
  for (int64_t p = unit; p < P; p += n_units) {
    long long changed = 0;
    double nodes = 0.0;
    for (int t = 0; t < n_steps; ++t) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      const double u_sel = r.uniform();
      const double logu = log(r.uniform());
      int k;
      if (block_mode == VK_BLOCK_ONE) { k = (int)(u_sel * (double)S); if (k >= S) k = S - 1; k = __ldg(&sc_sites[k]); }
      else if (block_mode == VK_BLOCK_EACH) k = __ldg(&sc_sites[t % S]);
      else k = block_mode;
      double *cell = sites + (int64_t)k * P + p;
      const int sec_own = __ldg(&pg.site_tab[5 * k + 2]);
      const double old = *cell;
      const int nc = __ldg(&pg.cand_tab[2 * k]);
      const int32_t *cc = pg.cand_consts + __ldg(&pg.cand_tab[2 * k + 1]);
      double xi[VM_MAX_ARGN], ps[VM_MAX_ARGN];
      double mx = -INFINITY;
      for (int c = 0; c < nc; ++c) {
        if (LANES == 32) __syncwarp();
        if (lane == 0) *cell = __ldg(&pg.consts[__ldg(&cc[c])]);
        if (LANES == 32) __syncwarp();
        double v = vm_children<LANES, FULL>(pg, k, R, sites, P, p, lane);
        vm_exec<FULL>(pg, sec_own, R, sites, P, p, nullptr, nullptr, v, false);
        xi[c] = v;
        mx = fmax(mx, v);
      }
      double den = 0.0;
      for (int c = 0; c < nc; ++c) { ps[c] = mx > -INFINITY ? exp(xi[c] - mx) : 0.0; den += ps[c]; }
      for (int c = 0; c < nc; ++c) ps[c] = den > 0.0 ? ps[c] / den : 1.0 / (double)nc;
      int idx = 0;
      if (map) {                          // emap (EnumerativeMAPOperator, egibbs.py:117-121): the first candidate of maximal weight, no draw
        for (int c = 0; c < nc; ++c) if (xi[c] == mx) { idx = c; break; }
      } else idx = categorical_sim(r, ps, nc, 1);
      double nv = __ldg(&pg.consts[__ldg(&cc[idx])]);
      if (!(logu < 0.0)) nv = old;
      if (LANES == 32) __syncwarp();
      if (lane == 0) *cell = nv;
      if (LANES == 32) __syncwarp();
      if (nv != old) ++changed;
      nodes += __ldg(&pg.affected[k]);
    }
    if (lane == 0) {
      if (changed_out) changed_out[p] = changed;
      if (nodes_out && block_mode == VK_BLOCK_ONE) nodes_out[p] = n_steps > 0 ? nodes / (double)n_steps : 0.0;
    }
  }
}

__global__ void prog_fill_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = v;
}

struct ProgramModel : Model {
  std::vector<double> affected_host;   // per site: nodes of the regular scaffold, then of the scaffold cut at the global border
  std::vector<int> ss_kind, site_bool;
  std::vector<int32_t> scope_tab_host, scope_sites_host, cand_tab_host, cand_consts_host;
  int n_scopes = 1;
  bool full = true;        // false: the lean kernels (<= VM_LEAN_REGS registers, core instructions) interpret this program
  DevBuf<int32_t> code, tables;       // tables: dir | site_tab | site_groups | group_tab
  DevBuf<double> consts, data, affected;
  DevBuf<double> sa, sb;
  double *sites = nullptr, *sites_alt = nullptr;
  VMProg pg;
  int S = 0, G = 0, n_regs = 0, n_sections = 0;
  int64_t max_rows = 0;
  bool propagated = false;

  ProgramModel(vk_particles *h_, const int32_t *code_h, int64_t n_code, const double *consts_h, int64_t n_consts,
               const int32_t *meta, int64_t n_meta, const double *data_h, int64_t n_data, const double *affected_h)
      : Model(h_) {
    VK_REQUIRE(n_meta >= 6, VK_EINVAL, "program: truncated meta block");
    S = meta[0]; G = meta[1]; n_regs = meta[2]; n_sections = meta[3];
    VK_REQUIRE(S >= 1 && G >= 0 && n_sections >= 2, VK_EINVAL, "program: bad header");
    VK_REQUIRE(n_regs <= VM_MAX_REGS, VK_EUNSUPPORTED, "program: too many registers for the generic kernel");
    VK_REQUIRE(n_code % VM_WORDS == 0, VK_EINVAL, "program: code length is not a whole number of instructions");
    // meta: header(6) | per site (sim, child, own, n_groups) | site group lists | per group (pro, row, rows, cols, off) | directory
    std::vector<int32_t> site_tab((size_t)S * 5), site_groups, group_tab((size_t)(G > 0 ? G : 1) * 5, 0);
    int64_t at = 6;
    VK_REQUIRE(n_meta >= at + 4 * (int64_t)S, VK_EINVAL, "program: truncated site table");
    int total_groups = 0;
    for (int k = 0; k < S; ++k) {
      for (int j = 0; j < 4; ++j) site_tab[(size_t)k * 5 + j] = meta[at + 4 * k + j];
      site_tab[(size_t)k * 5 + 4] = total_groups;
      VK_REQUIRE(meta[at + 4 * k + 3] >= 0, VK_EINVAL, "program: negative group count");
      total_groups += meta[at + 4 * k + 3];
    }
    at += 4 * (int64_t)S;
    VK_REQUIRE(n_meta >= at + total_groups + 5 * (int64_t)G + 2 * (int64_t)n_sections, VK_EINVAL, "program: truncated tables");
    for (int j = 0; j < total_groups; ++j) {
      VK_REQUIRE(meta[at + j] >= 0 && meta[at + j] < G, VK_EINVAL, "program: group index out of range");
      site_groups.push_back(meta[at + j]);
    }
    at += total_groups;
    for (int g = 0; g < G; ++g) {
      for (int j = 0; j < 5; ++j) group_tab[(size_t)g * 5 + j] = meta[at + 5 * g + j];
      const int64_t rows = meta[at + 5 * g + 2], cols = meta[at + 5 * g + 3], off = meta[at + 5 * g + 4];
      VK_REQUIRE(rows >= 0 && cols >= 1 && off >= 0 && off + rows * cols <= n_data, VK_EINVAL, "program: group table out of range");
      if (rows > max_rows) max_rows = rows;
    }
    at += 5 * (int64_t)G;
    std::vector<int32_t> dir(meta + at, meta + at + 2 * (int64_t)n_sections);
    at += 2 * (int64_t)n_sections;
    // subsampled MH block: per site (kind, local sections, offset), then the sections as pairs
    VK_REQUIRE(n_meta >= at + 3 * (int64_t)S, VK_EINVAL, "program: truncated subsampled-MH table");
    std::vector<int32_t> ss_tab(meta + at, meta + at + 3 * (int64_t)S);
    at += 3 * (int64_t)S;
    VK_REQUIRE(n_meta >= at + 1, VK_EINVAL, "program: truncated subsampled-MH table");
    const int64_t n_roots = meta[at];
    at += 1;
    VK_REQUIRE(n_roots >= 0 && n_meta >= at + 2 * n_roots + 1, VK_EINVAL, "program: truncated local-section list");
    std::vector<int32_t> ss_roots(meta + at, meta + at + 2 * n_roots);
    ss_roots.push_back(0);
    at += 2 * n_roots;
    // scopes: count; per scope (blocks, offset); the blocks' sites
    n_scopes = meta[at];
    at += 1;
    VK_REQUIRE(n_scopes >= 1 && n_scopes <= 64 && n_meta >= at + 2 * (int64_t)n_scopes, VK_EINVAL, "program: bad scope table");
    scope_tab_host.assign(meta + at, meta + at + 2 * (int64_t)n_scopes);
    at += 2 * (int64_t)n_scopes;
    int64_t n_scope_sites = 0;
    for (int sc = 0; sc < n_scopes; ++sc) n_scope_sites += scope_tab_host[(size_t)2 * sc] > 0 ? scope_tab_host[(size_t)2 * sc] : 0;
    VK_REQUIRE(n_meta >= at + n_scope_sites + 2 * (int64_t)S, VK_EINVAL, "program: truncated scope / candidate tables");
    scope_sites_host.assign(meta + at, meta + at + n_scope_sites);
    at += n_scope_sites;
    // enumeration candidates: per site (count, offset), then indices into the constants
    cand_tab_host.assign(meta + at, meta + at + 2 * (int64_t)S);
    at += 2 * (int64_t)S;
    cand_consts_host.assign(meta + at, meta + n_meta);
    cand_consts_host.push_back(0);
    for (int k = 0; k < S; ++k) {
      const int nc = cand_tab_host[(size_t)2 * k], off = cand_tab_host[(size_t)2 * k + 1];
      VK_REQUIRE(nc >= 0 && nc <= VM_MAX_ARGN && off >= 0 && (size_t)off + (size_t)nc < cand_consts_host.size() + (nc == 0), VK_EINVAL,
                 "program: bad candidate entry");
      for (int j = 0; j < nc; ++j)
        VK_REQUIRE(cand_consts_host[(size_t)off + j] >= 0 && cand_consts_host[(size_t)off + j] < n_consts, VK_EINVAL, "program: candidate constant out of range");
    }
    for (int sc = 0; sc < n_scopes; ++sc) {
      const int nb = scope_tab_host[(size_t)2 * sc], off = scope_tab_host[(size_t)2 * sc + 1];
      VK_REQUIRE(nb >= (sc == 0 ? S : 1) && off >= 0 && (size_t)off + (size_t)nb <= scope_sites_host.size(), VK_EINVAL, "program: bad scope entry");
      for (int i = off; i < off + nb; ++i)
        VK_REQUIRE(scope_sites_host[(size_t)i] >= 0 && scope_sites_host[(size_t)i] < S, VK_EINVAL, "program: scope site out of range");
    }
    for (int k = 0; k < S; ++k) VK_REQUIRE(scope_sites_host[(size_t)scope_tab_host[1] + k] == k, VK_EINVAL, "program: the default scope lists every site in order");
    for (int k = 0; k < S; ++k) {
      const int kind = ss_tab[(size_t)3 * k], n = ss_tab[(size_t)3 * k + 1], off = ss_tab[(size_t)3 * k + 2];
      VK_REQUIRE(kind >= 0 && kind <= 2 && n >= 0 && off >= 0 && (int64_t)off + n <= n_roots, VK_EINVAL, "program: bad subsampled-MH entry");
      VK_REQUIRE(kind != 1 || n > 1, VK_EINVAL, "program: a global border has more than one local section");
      for (int i = off; i < off + n; ++i) {
        const int g = ss_roots[(size_t)2 * i], r = ss_roots[(size_t)2 * i + 1];
        if (g >= 0) VK_REQUIRE(g < G && r >= 0 && r < group_tab[(size_t)g * 5 + 2], VK_EINVAL, "program: local section out of range");
        else VK_REQUIRE(r >= 0 && r < S, VK_EINVAL, "program: local section out of range");
      }
      ss_kind.push_back(kind);
    }
    const int64_t n_instr = n_code / VM_WORDS;
    for (int s = 0; s < n_sections; ++s)
      VK_REQUIRE(dir[2 * s] >= 0 && dir[2 * s + 1] >= 0 && (int64_t)dir[2 * s] + dir[2 * s + 1] <= n_instr, VK_EINVAL,
                 "program: section out of range");
    int argn_seen = 0;
    full = n_regs > VM_LEAN_REGS;
    for (int64_t i = 0; i < n_instr; ++i) {          // every operand index is checked once, here
      const int32_t *w = code_h + i * VM_WORDS;
      const int op = w[0];
      if (op == VM_MOV || op == VM_ARGN || op == VM_COND || op == VM_SELECT || op == VM_LOOKUP || op == VM_SIM_CATEGORICAL ||
          op == VM_LD_CATEGORICAL) full = true;
      auto reg = [&](int r) { VK_REQUIRE(r >= 0 && r < n_regs, VK_EINVAL, "program: register out of range"); };
      reg(w[1]);
      if (op == VM_ARGN) { VK_REQUIRE(w[2] >= 1 && w[2] <= VM_MAX_ARGN, VK_EINVAL, "program: block length out of range"); argn_seen = w[2]; continue; }
      if (op == VM_LOOKUP || op == VM_SIM_CATEGORICAL || op == VM_LD_CATEGORICAL) {     // a register block, announced by ARGN
        const int len = op == VM_LOOKUP ? argn_seen : 2 * argn_seen;
        VK_REQUIRE(argn_seen >= 1 && w[2] >= 0 && w[2] + len <= n_regs, VK_EINVAL, "program: register block out of range");
        if (op == VM_LOOKUP) reg(w[3]);
        argn_seen = 0;
        continue;
      }
      if (op == VM_CONST) VK_REQUIRE(w[2] >= 0 && w[2] < n_consts, VK_EINVAL, "program: constant out of range");
      else if (op == VM_SITE || op == VM_STORE) VK_REQUIRE(w[2] >= 0 && w[2] < S, VK_EINVAL, "program: site out of range");
      else if (op == VM_DATA) VK_REQUIRE(w[2] >= 0, VK_EINVAL, "program: column out of range");
      else { reg(w[2]); reg(w[3]); }
    }
    for (int g = 0; g < G; ++g) {                    // DATA columns are checked against the table of the group that reads them
      const int pro = group_tab[(size_t)g * 5], row = group_tab[(size_t)g * 5 + 1], cols = group_tab[(size_t)g * 5 + 3];
      VK_REQUIRE(pro >= 0 && pro < n_sections && row >= 0 && row < n_sections, VK_EINVAL, "program: bad section id");
      for (int i = dir[2 * pro]; i < dir[2 * pro] + dir[2 * pro + 1]; ++i)
        VK_REQUIRE(code_h[(int64_t)i * VM_WORDS] != VM_DATA, VK_EINVAL, "program: a group prologue reads no table column");
      for (int i = dir[2 * row]; i < dir[2 * row] + dir[2 * row + 1]; ++i)
        if (code_h[(int64_t)i * VM_WORDS] == VM_DATA)
          VK_REQUIRE(code_h[(int64_t)i * VM_WORDS + 2] < cols, VK_EINVAL, "program: column out of range");
    }
    for (int k = 0; k < S; ++k)
      for (int j = 0; j < 3; ++j)
        VK_REQUIRE(site_tab[(size_t)k * 5 + j] >= 0 && site_tab[(size_t)k * 5 + j] < n_sections, VK_EINVAL, "program: bad section id");
    site_bool.assign((size_t)S, 0);           // a site is boolean when its SIM section simulates it with one of the flip SPs
    for (int k = 0; k < S; ++k) {
      const int sec = site_tab[(size_t)k * 5];
      for (int i = dir[2 * sec]; i < dir[2 * sec] + dir[2 * sec + 1]; ++i) {
        const int32_t *w = code_h + (int64_t)i * VM_WORDS;
        if (w[0] == VM_STORE && w[2] == k && i > dir[2 * sec]) {
          const int prev = code_h[(int64_t)(i - 1) * VM_WORDS];
          site_bool[(size_t)k] = prev == VM_SIM_FLIP || prev == VM_SIM_LOG_FLIP || prev == VM_SIM_LOG_ODDS_FLIP;
        }
      }
    }
    VK_REQUIRE(meta[4] >= 0 && meta[4] < n_sections && meta[5] >= 0 && meta[5] < n_sections, VK_EINVAL, "program: bad section id");
    // upload
    std::vector<int32_t> all;
    const size_t o_dir = 0;
    all.insert(all.end(), dir.begin(), dir.end());
    const size_t o_site = all.size();
    all.insert(all.end(), site_tab.begin(), site_tab.end());
    const size_t o_sg = all.size();
    all.insert(all.end(), site_groups.begin(), site_groups.end());
    all.push_back(0);
    const size_t o_gt = all.size();
    all.insert(all.end(), group_tab.begin(), group_tab.end());
    const size_t o_ss = all.size();
    all.insert(all.end(), ss_tab.begin(), ss_tab.end());
    const size_t o_roots = all.size();
    all.insert(all.end(), ss_roots.begin(), ss_roots.end());
    const size_t o_sct = all.size();
    all.insert(all.end(), scope_tab_host.begin(), scope_tab_host.end());
    const size_t o_scs = all.size();
    all.insert(all.end(), scope_sites_host.begin(), scope_sites_host.end());
    const size_t o_ct = all.size();
    all.insert(all.end(), cand_tab_host.begin(), cand_tab_host.end());
    const size_t o_cc = all.size();
    all.insert(all.end(), cand_consts_host.begin(), cand_consts_host.end());
    code.alloc((size_t)(n_code > 0 ? n_code : 1)); tables.alloc(all.size());
    consts.alloc((size_t)(n_consts > 0 ? n_consts : 1)); data.alloc((size_t)(n_data > 0 ? n_data : 1)); affected.alloc((size_t)2 * S);
    VK_CUDA(cudaMemcpyAsync(code.p, code_h, (size_t)n_code * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(tables.p, all.data(), all.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    if (n_consts > 0) VK_CUDA(cudaMemcpyAsync(consts.p, consts_h, (size_t)n_consts * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (n_data > 0) VK_CUDA(cudaMemcpyAsync(data.p, data_h, (size_t)n_data * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(affected.p, affected_h, (size_t)2 * S * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));      // the host vectors die with this constructor
    sa.alloc((size_t)S * h->P); sb.alloc((size_t)S * h->P);
    sites = sa.p; sites_alt = sb.p;
    VK_CUDA(cudaMemsetAsync(sites, 0, sa.bytes(), h->stream));
    pg.code = code.p; pg.consts = consts.p; pg.data = data.p;
    pg.dir = tables.p + o_dir; pg.site_tab = tables.p + o_site; pg.site_groups = tables.p + o_sg; pg.group_tab = tables.p + o_gt;
    pg.affected = affected.p; pg.ss_nodes = affected.p + S;
    pg.ss_tab = tables.p + o_ss; pg.ss_roots = tables.p + o_roots;
    pg.scope_tab = tables.p + o_sct; pg.scope_sites = tables.p + o_scs;
    pg.cand_tab = tables.p + o_ct; pg.cand_consts = tables.p + o_cc;
    affected_host.assign(affected_h, affected_h + 2 * S);
    pg.n_sites = S; pg.n_groups = G; pg.sec_prior = meta[4]; pg.sec_joint = meta[5];
  }
  bool warp_mode() const { return max_rows >= 64; }
  int grid(bool warp) const { return grid_for(warp ? h->P * 32 : h->P, 256, h->sm_count, 8); }
  void launch(int mode, double *out, uint32_t step, int block = VK_BLOCK_ALL, int joint = 0) {
    const bool wm = warp_mode();
    if (!full) {       // small register file (it lives in local memory: stays in L1), core instruction set only
      if (wm) prog_kernel<32, false><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->lw.p, out, h->P, h->p_base, h->seed, step, mode, block, joint);
      else prog_kernel<1, false><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->lw.p, out, h->P, h->p_base, h->seed, step, mode, block, joint);
    } else {
      if (wm) prog_kernel<32, true><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->lw.p, out, h->P, h->p_base, h->seed, step, mode, block, joint);
      else prog_kernel<1, true><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->lw.p, out, h->P, h->p_base, h->seed, step, mode, block, joint);
    }
    VK_CUDA(cudaGetLastError());
  }
  void reset_to_prior() override { launch(VM_MODE_PRIOR, nullptr, h->next_step()); }
  void resimulate_from(int first) {
    VK_REQUIRE(first >= 0 && first <= S, VK_EINVAL, "program: first site out of range");
    if (first < S) launch(VM_MODE_RESIM, nullptr, h->next_step(), first, 0);
  }
  void likelihood_weight() override { propagated = true; launch(VM_MODE_LW, nullptr, h->next_step()); }
  void make_consistent(double *incr_dev) override {
    if (propagated) {
      if (incr_dev) VK_CUDA(cudaMemsetAsync(incr_dev, 0, (size_t)h->P * sizeof(double), h->stream));
      return;
    }
    launch(VM_MODE_CONSISTENT, incr_dev, 0);
    propagated = true;
  }
  // blocks of a scope (one site each); scope 0 = default (every site), then the scopes made by `tag`
  int scope_blocks(int scope) const {
    VK_REQUIRE(scope >= 0 && scope < n_scopes, VK_EINVAL, "program: scope out of range (scope 0 = default, then the tagged scopes in order of appearance)");
    return scope_tab_host[(size_t)2 * scope];
  }
  int scope_site(int scope, int b) const { return scope_sites_host[(size_t)scope_tab_host[(size_t)2 * scope + 1] + (size_t)b]; }
  // block: one / each stay as they are (the kernels pick inside the scope); all only in the default scope; an explicit
  // block index of the scope becomes its SITE index
  int block_of(int scope, int block, bool allow_all) const {
    const int nb = scope_blocks(scope);
    if (block == VK_BLOCK_ONE || block == VK_BLOCK_EACH) return block;
    if (block == VK_BLOCK_ALL) {
      VK_REQUIRE(allow_all && scope == 0, VK_EUNSUPPORTED, "program: block 'all' is compiled for log_likelihood_at / log_joint_at in the default scope only");
      return block;
    }
    VK_REQUIRE(block >= 0 && block < nb, VK_EINVAL, "program: block must be one / each / a block index of the scope");
    return scope_site(scope, block);
  }
  // transloop's value for a particle-independent block choice (dispatch.py:109-128)
  void fill_nodes(double *mean_nodes_dev, int scope, int block, int K, const double *counts) {
    if (!mean_nodes_dev || block == VK_BLOCK_ONE) return;
    double v = 0.0;
    if (K > 0 && block == VK_BLOCK_EACH) {         // scaffolder_loop: the blocks' counts summed, FLOOR-divided by #blocks (dispatch.py:122-128)
      const int nb = scope_blocks(scope);
      double sum = 0.0;
      for (int b = 0; b < nb; ++b) sum += counts[(size_t)scope_site(scope, b)];
      v = floor(sum / (double)nb);
    } else if (K > 0) v = counts[(size_t)block];
    prog_fill_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(mean_nodes_dev, h->P, v);
  }
  void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) override {
    block = block_of(scope, block, false);
    VK_REQUIRE(propagated || G == 0, VK_ESTATE, "mh: unpropagated observations; call vk_make_consistent first");
    const int n_steps = block == VK_BLOCK_EACH ? K * scope_blocks(scope) : K;
    const uint32_t step0 = h->next_step((uint32_t)n_steps);
    if (!full) {       // small register file (it lives in local memory: stays in L1), core instruction set only
      if (warp_mode()) prog_mh_kernel<32, false><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, mean_nodes_dev, acc_dev);
      else prog_mh_kernel<1, false><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, mean_nodes_dev, acc_dev);
    } else {
      if (warp_mode()) prog_mh_kernel<32, true><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, mean_nodes_dev, acc_dev);
      else prog_mh_kernel<1, true><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, mean_nodes_dev, acc_dev);
    }
    fill_nodes(mean_nodes_dev, scope, block, K, affected_host.data());
    VK_CUDA(cudaGetLastError());
  }
  // `each` is not a block of subsampled_mh's indexer (BlockScaffoldIndexer: one or a block, dispatch.py:315-322)
  void subsampled_mh(int scope, int block, int64_t Nbatch, int k0, double eps, int use_delta, double delta_sigma,
                     int update_values, int K, double *mean_nodes_dev, double *stats_dev) override {
    (void)use_delta; (void)delta_sigma; (void)update_values;   // none of the compiled SPs has a delta kernel (psp.py:281; continuous.py:352-353)
    block = block_of(scope, block, false);
    VK_REQUIRE(block != VK_BLOCK_EACH, VK_EINVAL, "subsampled_mh: block must be one or a block of the scope");
    VK_REQUIRE(eps >= 0.0 && eps < 0.5, VK_EINVAL, "subsampled_mh: epsilon must be in [0, 0.5)");
    VK_REQUIRE(propagated || G == 0, VK_ESTATE, "subsampled_mh: unpropagated observations; call vk_make_consistent first");
    for (int b = 0; b < scope_blocks(scope); ++b) {
      const int k = scope_site(scope, b);
      if (block == VK_BLOCK_ONE || block == k)
        VK_REQUIRE(ss_kind[(size_t)k] != 2, VK_EUNSUPPORTED,
                   "subsampled_mh: Invalid global border or overlapping local sections at a site of the block "
                   "(the reference raises SubsampledScaffoldError / warns NotApplicable)");
    }
    const uint32_t step0 = h->next_step((uint32_t)K);
    if (!full) {       // small register file (it lives in local memory: stays in L1), core instruction set only
      if (warp_mode()) prog_ssmh_kernel<32, false><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, K, scope, block, (double)Nbatch, k0, eps, mean_nodes_dev, stats_dev);
      else prog_ssmh_kernel<1, false><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, K, scope, block, (double)Nbatch, k0, eps, mean_nodes_dev, stats_dev);
    } else {
      if (warp_mode()) prog_ssmh_kernel<32, true><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, K, scope, block, (double)Nbatch, k0, eps, mean_nodes_dev, stats_dev);
      else prog_ssmh_kernel<1, true><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, K, scope, block, (double)Nbatch, k0, eps, mean_nodes_dev, stats_dev);
    }
    VK_CUDA(cudaGetLastError());
  }
  void gibbs(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) override {
    enumerate(scope, block, K, 0, mean_nodes_dev, changed_dev);
  }
  void emap(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) override {
    enumerate(scope, block, K, 1, mean_nodes_dev, changed_dev);
  }
  void enumerate(int scope, int block, int K, int map, double *mean_nodes_dev, long long *changed_dev) {
    block = block_of(scope, block, false);
    VK_REQUIRE(propagated || G == 0, VK_ESTATE, "gibbs: unpropagated observations; call vk_make_consistent first");
    for (int b = 0; b < scope_blocks(scope); ++b) {
      const int k = scope_site(scope, b);
      if (block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || block == k)
        VK_REQUIRE(cand_tab_host[(size_t)2 * k] > 0, VK_EUNSUPPORTED,
                   "gibbs: a site of the block is not boolean or categorical -- only those are enumerated for compiled programs "
                   "(the reference's PSP.enumerateValues raises for continuous SPs)");
    }
    const int n_steps = block == VK_BLOCK_EACH ? K * scope_blocks(scope) : K;
    const uint32_t step0 = h->next_step((uint32_t)n_steps);
    if (!full) {       // small register file (it lives in local memory: stays in L1), core instruction set only
      if (warp_mode()) prog_gibbs_kernel<32, false><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, map, mean_nodes_dev, changed_dev);
      else prog_gibbs_kernel<1, false><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, map, mean_nodes_dev, changed_dev);
    } else {
      if (warp_mode()) prog_gibbs_kernel<32, true><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, map, mean_nodes_dev, changed_dev);
      else prog_gibbs_kernel<1, true><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, map, mean_nodes_dev, changed_dev);
    }
    fill_nodes(mean_nodes_dev, scope, block, K, affected_host.data());
    VK_CUDA(cudaGetLastError());
  }
  void slice(int scope, int block, double w, double m, int doubling, int K, double *mean_nodes_dev, long long *moved_dev) override {
    block = block_of(scope, block, false);
    VK_REQUIRE(w > 0.0 && m >= (doubling ? 0.0 : 1.0) && m == floor(m), VK_EINVAL,
               "slice: w must be positive, m a whole number (at least 1 steps out, at least 0 doublings)");
    VK_REQUIRE(propagated || G == 0, VK_ESTATE, "slice: unpropagated observations; call vk_make_consistent first");
    const int n_steps = block == VK_BLOCK_EACH ? K * scope_blocks(scope) : K;
    const uint32_t step0 = h->next_step((uint32_t)n_steps);
    if (!full) {       // small register file (it lives in local memory: stays in L1), core instruction set only
      if (warp_mode()) prog_slice_kernel<32, false><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, w, m, doubling, mean_nodes_dev, moved_dev);
      else prog_slice_kernel<1, false><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, w, m, doubling, mean_nodes_dev, moved_dev);
    } else {
      if (warp_mode()) prog_slice_kernel<32, true><<<grid(true), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, w, m, doubling, mean_nodes_dev, moved_dev);
      else prog_slice_kernel<1, true><<<grid(false), 256, 0, h->stream>>>(pg, sites, h->P, h->p_base, h->seed, step0, n_steps, scope, block, w, m, doubling, mean_nodes_dev, moved_dev);
    }
    fill_nodes(mean_nodes_dev, scope, block, K, affected_host.data());
    VK_CUDA(cudaGetLastError());
  }
  void log_likelihood_at(int scope, int block, bool joint, double *out_dev) override {
    block = block_of(scope, block, true);
    VK_REQUIRE(block == VK_BLOCK_ALL || block >= 0, VK_EINVAL, "log_likelihood_at: block must be all or a block of the scope");
    launch(VM_MODE_LOGLIK, out_dev, 0, block, joint ? 1 : 0);
  }
  int64_t num_sites() const override { return S; }
  int64_t num_blocks(int scope) const override { return scope_blocks(scope); }
  void read_site(int64_t site, double *out_dev) override {
    VK_REQUIRE(site >= 0 && site < S, VK_EINVAL, "program: site out of range");
    VK_CUDA(cudaMemcpyAsync(out_dev, sites + site * h->P, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  void write_site(int64_t site, const double *in_dev) override {
    VK_REQUIRE(site >= 0 && site < S, VK_EINVAL, "program: site out of range");
    VK_CUDA(cudaMemcpyAsync(sites + site * h->P, in_dev, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  std::vector<StateArray> state_arrays(bool) override {
    StateArray a{(void **)&sites, (void **)&sites_alt, 0, (size_t)S, 8};
    return {a};
  }
  void resized(int64_t n_new, const std::vector<void *> &nc, const std::vector<void *> &na) override {
    sa.adopt(nc[0], (size_t)n_new * S); sb.adopt(na[0], (size_t)n_new * S);
    sites = sa.p; sites_alt = sb.p;
  }
  int device_buffer(int which, void **p, int64_t *bytes) override {
    if (which == 1) { *p = sites; *bytes = (int64_t)sa.bytes(); return VK_OK; }
    return VK_EINVAL;
  }
};

Model *make_program(vk_particles *h, const int32_t *code, int64_t n_code, const double *consts, int64_t n_consts,
                    const int32_t *meta, int64_t n_meta, const double *data, int64_t n_data, const double *affected) {
  return new ProgramModel(h, code, n_code, consts, n_consts, meta, n_meta, data, n_data, affected);
}

void program_resimulate_from(Model *m, int first_site) { static_cast<ProgramModel *>(m)->resimulate_from(first_site); }

}  // namespace vk

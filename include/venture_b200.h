/*
 * venture_b200.h -- C ABI of the B200-native particle-parallel inference path
 * for Venture (probcomp/Venturecxx).
 *
 * One `vk_particles` handle is one device-resident batch of P weighted traces
 * ("particles") of ONE fixed-structure model, laid out structure-of-arrays in
 * HBM.  It replaces, for that batch, what the reference reaches through
 *   TraceSet.traces.map('<method>')        python/lib/engine/trace_set.py:148-164
 * i.e. the per-trace backend methods exported by Puma's Boost.Python module
 *   backend/new_cxx/src/pytrace.cxx:330-366
 * and by the Lite trace
 *   backend/lite/trace.py:422-448 (makeConsistent), :509-520 (primitive_infer,
 *   log_likelihood_at, log_joint_at), :522-530 (likelihood_weight)
 * plus the TraceSet-level weight bookkeeping and resampling
 *   python/lib/engine/trace_set.py:202-235, 287-294
 *   backend/lite/utils.py:70-91, 152-180.
 *
 * Conventions
 *   - every entry point returns 0 on success and a nonzero VK_E* code on
 *     failure; vk_last_error() then holds a message (thread-local).  The
 *     Python shim raises it as VentureException("evaluation", msg), the way
 *     pytrace.cxx:236-242,331-332 translates Puma's C++ exceptions.
 *   - plain pointers and sizes only.  Pointers named *_host are host memory
 *     (pageable or pinned).  A call that is given a host INPUT or OUTPUT
 *     buffer issues its copies on the handle's stream and returns after the
 *     stream has drained, so results are ready and inputs may be reused.
 *     Calls without host buffers (vk_likelihood_weight, vk_resample / vk_mh
 *     with NULL outputs, vk_reset_to_prior, ...) are STREAM-ORDERED: they
 *     return once enqueued; vk_synchronize() or any later host-output call
 *     waits for them, and an asynchronous CUDA error is reported there.
 *   - the handle owns all device memory.  One caller at a time per handle
 *     (the reference engine is single-threaded per TraceSet).
 *   - arithmetic is IEEE double (the reference is double everywhere).
 *   - there is NO CPU fallback: without a CUDA device every compute call
 *     fails with VK_ECUDA.
 */
#ifndef VENTURE_B200_H
#define VENTURE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vk_particles vk_particles;

enum vk_status {
  VK_OK = 0,
  VK_EINVAL = 1,      /* bad argument  -> VentureException("invalid_argument") */
  VK_ECUDA = 2,       /* CUDA runtime / launch failure                        */
  VK_EUNSUPPORTED = 3,/* operation not defined for this model kind            */
  VK_ESTATE = 4       /* call sequence error (e.g. infer before observe)      */
};

enum vk_model_kind {
  VK_MODEL_NORMAL_NORMAL = 1, /* x ~ normal(mu0,s0); observe normal(x,s) = y_k          */
  VK_MODEL_BLR = 2,           /* w_j ~ normal(mu0,s0); observe log_odds_flip(w.x_i)=y_i */
  VK_MODEL_HMM = 3,           /* examples/hmm.vnt:31-60                                 */
  VK_MODEL_LDA = 4,           /* examples/lda.vnt:18-41                                 */
  VK_MODEL_PROGRAM = 5        /* any fixed-structure program compiled by compiler.py    */
};

/* block argument of an infer command (backend/lite/infer/dispatch.py:86-107) */
#define VK_BLOCK_ONE  (-1)  /* a uniformly chosen block per transition */
#define VK_BLOCK_ALL  (-2)  /* all blocks of the scope as one proposal  */
#define VK_BLOCK_EACH (-3)  /* every block in order, one proposal each  */
#define VK_BLOCK_ORDERED (-4) /* all blocks of an ordered scope, in order (pgibbs) */

/* arithmetic of the N-observation sums inside the MH hot loop ("fp64 (parity) and fp32
 * (throughput) modes").  State (latent values, logit caches, weights), proposals and accept tests are
 * fp64 in both. */
enum vk_precision {
  VK_FP64 = 0, /* default: every log density within 1e-12 relative of Lite's       */
  VK_FP32 = 1  /* log densities within 1e-5 relative (tests/test_gpu_fp32_mode.py) */
};

enum vk_resample_mode {
  VK_RESAMPLE_MULTINOMIAL = 0, /* trace_set.py:209-219 (i.i.d. categorical draws) */
  VK_RESAMPLE_SYSTEMATIC = 1   /* one uniform, stratified positions (k+u)/P       */
};

/* ---- library ---------------------------------------------------------- */
const char *vk_last_error(void);
int vk_version(void);
int vk_device_count(int *n_out);

/* ---- construction (replaces Backend.trace_constructor()(seed) for P traces:
 *      python/lib/shortcuts/__init__.py:90-149, trace_set.py:38-48) --------- */
int vk_create_normal_normal(int64_t n_particles, uint64_t seed, int device,
                            double prior_mu, double prior_sigma,
                            double obs_sigma, int64_t max_obs,
                            vk_particles **out);
/* A COMPILED fixed-structure program (venturecxx_b200/compiler.py): what
 * constructScaffold (backend/lite/scaffold.py:149-316) + detach / regen (detach.py:31-120,
 * regen.py:49-161) reduce to for every random choice of a program over the scalar SPs, as
 * straight-line code of a register machine interpreted by ONE pair of kernels
 * (csrc/k_program.cu).  Replaces Trace.eval of the assume / observe directives
 * (lite/trace.py:397-420) for all particles.
 *   code      4 int32 per instruction: op, dst, a, b (op codes: compiler.py OP)
 *   consts    literals of the assumes
 *   meta      n_sites, n_groups, n_registers, n_sections, PRIOR section, JOINT section;
 *             per site: SIM, CHILD, OWN section, number of observation groups below it;
 *             the sites' group lists; per group: prologue section, row section, rows,
 *             columns (literals + observed value), offset into data; per section: first
 *             instruction, count; per site, for subsampled MH (infer/subsampled_mh.py:326-368):
 *             kind (0 no global border, 1 border found, 2 refused), number of local sections,
 *             offset into the section list; the number of local sections in all, then the local
 *             sections as pairs (group, row) | (-1, child site), in the order of the border node's
 *             children; the number of scopes, per scope (blocks, offset into the site list), the
 *             blocks' sites -- scope 0 is the default scope (every site, in order), the others are
 *             what `(tag scope block expr)` made (lite/scope.py:25-52; one random choice per block);
 *             entry points taking (scope, block) read block as an index into the scope's list;
 *             per site (enumeration candidates, offset), then the candidates as indices into consts
 *             ([true, false] for the flip family, the objects of a categorical, none otherwise)
 *   data      the groups' tables, row-major
 *   affected  2 * n_sites doubles.  Per site: Scaffold.numAffectedNodes (scaffold.py:67-68), what
 *             resimulation_mh reports (dispatch.py:109-120); then per site the same count for the
 *             scaffold cut at the global border, what subsampled_mh reports
 * Every index is validated here; the kernels do not re-check. */
int vk_create_program(int64_t n_particles, uint64_t seed, int device, const int32_t *code,
                      int64_t n_code, const double *consts, int64_t n_consts,
                      const int32_t *meta, int64_t n_meta, const double *data, int64_t n_data,
                      const double *affected, vk_particles **out);
/* Sites first_site.. of a compiled program from their priors given the CURRENT values of the
 * earlier sites: what evaluating further `assume` directives does to an existing trace
 * (lite/trace.py:397-401 eval -> regen.py:181-230 evalFamily), for all particles. */
int vk_program_resimulate(vk_particles *h, int32_t first_site);
/* X rows are D doubles; capacity n_obs_capacity observations. */
int vk_create_blr(int64_t n_particles, uint64_t seed, int device, int32_t D,
                  int64_t n_obs_capacity, double prior_mu, double prior_sigma,
                  vk_particles **out);
/* trans: n_trans_rows x S row-major, emit: S x O row-major (unnormalised
 * categorical weights, normalised per call like lite/utils.py:101-118);
 * state(1) = init_state (1-based, hmm.vnt:55-60); T_max = history length. */
int vk_create_hmm(int64_t n_particles, uint64_t seed, int device, int32_t S,
                  int32_t O, int32_t n_trans_rows, const double *trans_host,
                  const double *emit_host, int32_t init_state, int32_t T_max,
                  vk_particles **out);
/* tokens are observed later with vk_lda_observe; doc d holds at most
 * max_tokens_per_doc tokens.  Hyperpriors alpha ~ gamma(hyper_shape, hyper_rate). */
int vk_create_lda(int64_t n_particles, uint64_t seed, int device, int32_t K,
                  int32_t V, int32_t n_docs, int32_t max_tokens,
                  double hyper_shape, double hyper_rate, vk_particles **out);
int vk_destroy(vk_particles *h);

int vk_kind(const vk_particles *h);
int64_t vk_num_particles(const vk_particles *h);
/* has_own_prng / set_seed / get_seed  (lite/trace.py:89-92, trace_set.py) */
int vk_set_seed(vk_particles *h, uint64_t seed);
int vk_get_seed(const vk_particles *h, uint64_t *seed_out);
/* RNG step counter: every infer-level operation consumes one or more steps
 * of the counter-based Philox stream (DESIGN.md "RNG contract"). */
int vk_get_step(const vk_particles *h, uint32_t *step_out);
int vk_set_step(vk_particles *h, uint32_t step);
/* global index of this batch's particle 0 when the particle set is sharded
 * over several GPUs (streams are addressed by GLOBAL particle index, so
 * results do not depend on the GPU count). */
int vk_set_particle_base(vk_particles *h, int64_t base);
/* TraceSet.reset_to_prior (trace_set.py:199-200): resimulate every latent
 * from its prior; weights untouched.  Also the initial `assume` simulation:
 * call once after vk_create_* (and after vk_set_particle_base). */
int vk_reset_to_prior(vk_particles *h);
/* run on a caller-provided CUDA stream (cudaStream_t passed as void*),
 * e.g. torch.cuda.current_stream().cuda_stream; NULL = the handle's own. */
int vk_set_stream(vk_particles *h, void *cuda_stream);
int vk_synchronize(vk_particles *h);

/* ---- observe (engine.observe -> trace.observe; values are shared by all
 *      particles; they stay "unpropagated" until vk_make_consistent,
 *      lite/trace.py:415-448) -------------------------------------------- */
int vk_nn_observe(vk_particles *h, const double *y_host, int64_t n);
int vk_blr_observe(vk_particles *h, const double *X_rows_host,
                   const uint8_t *y_host, int64_t n);
int vk_hmm_observe(vk_particles *h, int32_t symbol /* 1-based */);
int vk_lda_observe(vk_particles *h, const int32_t *doc_host,
                   const int32_t *word_host, int64_t n);

/* engine.forget of every observe directive (trace_set.py:157-161 ->
 * lite/trace.py:483-493 unobserve): drops all observations; returns per
 * particle the weight unconstrain() reports (sum of their logDensity);
 * log weights are NOT changed, as in the reference. out may be NULL. */
int vk_forget_observations(vk_particles *h, double *weights_out_host);

/* ---- weights (trace_set.py:287-294; engine/inference.py:218-234) ------- */
/* TraceSet.incorporate: lw[p] += makeConsistent()[p]; increments optionally
 * copied out (host, P doubles; may be NULL). NaN increment -> -inf. */
int vk_make_consistent(vk_particles *h, double *increments_out_host);
/* TraceSet.likelihood_weight: resimulate all latents from the prior,
 * lw[p] := sum_obs logDensity (overwrite). */
int vk_likelihood_weight(vk_particles *h);
int vk_get_log_weights(vk_particles *h, double *out_host);
int vk_set_log_weights(vk_particles *h, const double *in_host);
int vk_increment_log_weights(vk_particles *h, const double *in_host);
/* lite/utils.py:70-91 over the particle log-weights. */
int vk_logsumexp(vk_particles *h, double *out_host);
/* lite/utils.py:159-168 (all -inf -> zeros). */
int vk_normalized_probs(vk_particles *h, double *out_host);
/* stand-alone logsumexp of a host array (n may be 0 -> -inf). */
int vk_logsumexp_array(const double *in_host, int64_t n, int device,
                       double *out_host);
/* vk_logsumexp_array for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  No counterpart in the reference.
 * Allocates nothing and calls no allocating routine: its scratch is a 131 328-byte static __device__ array inside
 * the library (csrc/k_static.cu; every GPU that loads the library holds it until exit).
 *   in_dev      n doubles in device memory of the CURRENT device (cudaSetDevice is the caller's)
 *   n           0 <= n <= 16 777 216 (2^24: one scratch record per 4096 elements, 4096 records); larger n answers
 *               VK_EINVAL with the limit in vk_last_error(); n = 0 writes -inf
 *   out_dev     one double in device memory: lite/utils.py:70-91 with its NaN / +-inf cases
 *   cuda_stream cudaStream_t to enqueue on (NULL = the default stream); stream-ordered, returns without
 *               synchronising
 * Every call shares the one scratch array: calls that overlap on the GPU (different streams, threads or processes
 * on one context) race and give wrong results -- serialise them.  The first call on a device looks up the array's
 * device address (cudaGetSymbolAddress); make that call outside stream capture. */
int vk_logsumexp_static(const double *in_dev, int64_t n, double *out_dev,
                        void *cuda_stream);

/* ---- resample (trace_set.py:202-235, utils.py:93-104,152-157,170-180) ---
 * Draw n_new ancestors proportional to exp(lw); copy ancestor state; all new
 * weights = logsumexp(lw) - log(n_new).  n_new may differ from the current particle count (the
 * reference's programs start with `(resample 1000)` on one particle, benchmarks/Makefile:6-9): the
 * handle then holds n_new particles afterwards (vk_num_particles), every per-particle output of later
 * calls has n_new entries, and global particle indices (RNG streams) are p_base + [0, n_new).  A
 * particle set sharded over peer GPUs (vk_p2p_open) cannot change its size.
 * uniforms_host: NULL = draw from the handle's Philox stream; otherwise 1
 * (systematic) or n_new (multinomial) uniforms in [0,1) to use instead.
 * ancestors_out_host: NULL or n_new int32. */
int vk_resample(vk_particles *h, int64_t n_new, int mode,
                const double *uniforms_host, int32_t *ancestors_out_host);

/* ---- inference (infer/dispatch.py:130-136, mh.py:59-113,167-208) --------
 * K single-site resimulation-MH transitions per particle.  scope 0 =
 * "default".  block = VK_BLOCK_ONE / VK_BLOCK_EACH / a block index.
 * mean_nodes_out_host: NULL or P doubles = transloop's return value
 * (mean #affected nodes per transition, dispatch.py:109-120).
 * accepted_out_host: NULL or P int64 accept counters (profiling analogue of
 * mh.py:98-111). */
int vk_mh(vk_particles *h, int32_t scope, int32_t block, int32_t transitions,
          double *mean_nodes_out_host, int64_t *accepted_out_host);

/* ---- sharded resample over NVLink peer memory ------------------------------------
 * python/lib/engine/trace_set.py:202-235 for a particle set block-partitioned over the GPUs of one node,
 * one process per GPU: rank g owns global particles [g*P, (g+1)*P) (vk_set_particle_base).  No NCCL
 * collective and no host synchronisation on the path: three scalars per rank travel through mailboxes in
 * peer memory (the two exchanges ride in the tails of the kernels that produce their payload, so a sharded
 * resample is the same four launches as a single-GPU one), ancestor indices are stored straight into the
 * owning rank's array, ancestor rows are pulled with peer loads as soon as their index has arrived
 * (csrc/p2p_common.cuh, csrc/p2p_resample.cuh have the protocol).  Result bit-identical to vk_resample
 * on a single GPU holding all world*P particles.
 * Failure: a peer that never arrives at an exchange trips a wall-clock bound (VK_P2P_TIMEOUT_S, default 20);
 * the step is then abandoned on this rank (weights poisoned with NaN, nothing consumed from the stale
 * mailbox) and the NEXT host-synchronising call on the handle (vk_synchronize, any call with a host output)
 * fails with VK_ECUDA -- the particle set is invalid from then on.
 *   vk_p2p_export  allocates the mailbox and writes this handle's CUDA IPC handles (64 bytes each:
 *                  mailbox, ancestor array, both buffers of every state array) to handles_out.
 *   vk_p2p_open    all_handles = the exports of ranks 0..world-1 concatenated (exchange them with any
 *                  host mechanism, e.g. torch.distributed.all_gather_object); opens the peers' memory.
 *   vk_p2p_resample  the resample itself (mode as vk_resample); stream-ordered, returns immediately.
 *                  Every rank must call it the same number of times.
 *   vk_p2p_status  synchronises and reports without failing; *timed_out = 1 if a peer never arrived at an
 *                  exchange; *lse_out = global logsumexp of the weights before the last resample.
 * The same peer tables carry the observation-sharded subsampled MH (E3): open them on handles whose
 * particles are REPLICATED (particle base 0 on every rank, vk_blr_set_observation_shard) and
 * vk_subsampled_mh exchanges its per-particle partial sums through peer mailboxes inside its kernel. */
int vk_p2p_export(vk_particles *h, unsigned char *handles_out, int32_t max_handles, int32_t *n_handles_out);
int vk_p2p_open(vk_particles *h, int32_t world, int32_t rank, const unsigned char *all_handles, int32_t n_handles);
int vk_p2p_resample(vk_particles *h, int mode);
int vk_p2p_status(vk_particles *h, int32_t *timed_out, double *lse_out);

/* Particle Gibbs, `(pgibbs scope ordered particles transitions)`: infer/dispatch.py:197-212 ->
 * infer/pgibbs.py:49-126 (PGibbsOperator under mixMH; registered at inference_sps.py:269-300): conditional
 * SMC over the ordered blocks of the scope inside every trace, `particles` new inner particles next to the
 * retained path, one MH accept of the chosen path.  Compiled for the HMM model (the state sequence,
 * hmm.vnt's scope `seq` = scope 0 here; 1..16 inner particles, at most 32 time steps); other model kinds
 * answer VK_EUNSUPPORTED.  accepted_out_host: NULL or P int64 = transitions that replaced the path. */
int vk_pgibbs(vk_particles *h, int32_t scope, int32_t block, int32_t particles, int32_t transitions,
              double *mean_nodes_out_host, int64_t *accepted_out_host);

/* Slice sampling with stepping out, `(slice scope block w m transitions)`: infer/dispatch.py:190-194 ->
 * infer/slice_sample.py:39-139 (StepOutSliceOperator under mixMH; registered at inference_sps.py:500-516).
 * The site must be continuous and one-dimensional.  Compiled for programs of the model compiler
 * (vk_create_program); the hand-described model kinds answer VK_EUNSUPPORTED.
 * mean_nodes_out_host as vk_mh; moved_out_host: NULL or P int64 = transitions that changed the value. */
int vk_slice(vk_particles *h, int32_t scope, int32_t block, double w, double m, int32_t transitions,
             double *mean_nodes_out_host, int64_t *moved_out_host);
/* Slice sampling with doubling, `(slice_doubling scope block w p transitions)`: infer/dispatch.py:195-199 ->
 * DoublingSliceOperator, infer/slice_sample.py:148-196 (findInterval: at most p doublings of the interval, one uniform each;
 * legalProposal: Neal's acceptance test for a point of a doubled interval).  Otherwise as vk_slice. */
int vk_slice_doubling(vk_particles *h, int32_t scope, int32_t block, double w, int32_t p, int32_t transitions,
                      double *mean_nodes_out_host, int64_t *moved_out_host);

/* Enumerative Gibbs, `(gibbs scope block transitions)`: infer/dispatch.py:175-179 ->
 * infer/egibbs.py:42-107 (EnumerativeGibbsOperator under mixMH).  For a discrete site every candidate
 * value is weighted by its regen weight, the new value is drawn proportionally and always accepted.
 * Compiled for the HMM's state sites (scope 0) and the LDA z sites (scope 1 = "z", block = token index,
 * or scope 0 with block = 2 + token index; the default scope's gamma sites cannot be enumerated and
 * give VK_EUNSUPPORTED, as PSP.enumerateValues raises in the reference) and for the BOOLEAN sites (flip, log_flip,
 * log_odds_flip) and CATEGORICAL sites of compiled programs, in any scope.  Other models and sites: VK_EUNSUPPORTED.
 * mean_nodes_out_host as vk_mh; changed_out_host: NULL or P int64 = transitions that changed the value. */
int vk_gibbs(vk_particles *h, int32_t scope, int32_t block, int32_t transitions,
             double *mean_nodes_out_host, int64_t *changed_out_host);
/* `(emap scope block transitions)`: infer/dispatch.py:180-184 -> EnumerativeMAPOperator, infer/egibbs.py:117-121 -- as
 * vk_gibbs, but the candidate of largest regen weight (the first of equal ones) is taken instead of a draw.  Compiled
 * for the boolean sites of compiled programs; other models answer VK_EUNSUPPORTED. */
int vk_emap(vk_particles *h, int32_t scope, int32_t block, int32_t transitions,
            double *mean_nodes_out_host, int64_t *changed_out_host);

/* `(subsampled_mh scope block Nbatch k0 epsilon useDeltaKernels deltaKernelArgs updateValues transitions)`:
 * infer/dispatch.py:149-152, 315-322 -> infer/subsampled_mh.py:76-171 (subsampledMixMH, sequentialTest).
 * mean_nodes_out_host: NULL or P doubles = transloop's return value, the nodes of the scaffold cut at the global
 * border (logistic regression: 3).  stats_out_host: NULL or 3*P doubles per call = (accepted transitions, mean local
 * sections visited per transition, t statistic of the last sequential test, +-inf when it computed none, :170).
 * Logistic regression: one cooperative, persistent kernel runs every transition of the call (proposal, minibatch
 * rounds, sequential tests, accept) on the device (csrc/blr_subsampled_persistent.cuh; D <= 32, P <= 4096 -- larger
 * D runs the host-driven round loop).  With observations sharded over peer GPUs
 * (vk_blr_set_observation_shard + vk_p2p_open on replicated particles) the same call exchanges the
 * per-particle partial sums through peer memory and every rank takes identical decisions.
 * Compiled programs (vk_create_program): the compiler finds each site's global border and local sections
 * (findGlobalBorder :326-368) -- rows of observation groups and latent children; a site without a border takes a
 * regular MH step (:91-95); a site whose structure the reference rejects (SubsampledScaffoldError, "Invalid global
 * border") or mis-handles (overlapping local sections) answers VK_EUNSUPPORTED.  Sections are evaluated from the
 * current values (no per-section cache): the reference followed by its makeConsistent (:411-441).  block: one or a
 * site index.  The other model kinds answer VK_EUNSUPPORTED. */
int vk_subsampled_mh(vk_particles *h, int32_t scope, int32_t block,
                     int64_t Nbatch, int32_t k0, double epsilon,
                     int32_t use_delta_kernels, double delta_sigma,
                     int32_t update_values, int32_t transitions,
                     double *mean_nodes_out_host, double *stats_out_host);

/* infer/dispatch.py:331-361 -- non-mutating. out_host: P doubles. */
int vk_log_likelihood_at(vk_particles *h, int32_t scope, int32_t block,
                         double *out_host);
int vk_log_joint_at(vk_particles *h, int32_t scope, int32_t block,
                    double *out_host);

/* ---- state access (extractValue / parity-test injection) ----------------
 * site ids are model specific (DESIGN.md "site tables"); values are doubles
 * (integers are converted), P per site. */
int vk_num_sites(const vk_particles *h, int64_t *n_out);
int vk_read_site(vk_particles *h, int64_t site, double *out_host);
int vk_write_site(vk_particles *h, int64_t site, const double *in_host);
/* after vk_write_site: recompute derived caches (logits, counts). */
int vk_refresh(vk_particles *h);
/* Precision of the hot loop.  VK_FP32 is implemented by the logistic-regression model's
 * resimulation_mh with block one / each / <site> (rows that fit shared memory as float); every other
 * entry point, and every other model kind, computes in fp64 -- other kinds answer VK_EUNSUPPORTED to
 * VK_FP32.  After an fp32 call the logit cache holds float-rounded values; the library recomputes it from
 * the fp64 latents before vk_log_likelihood_at / vk_log_joint_at / vk_resample / vk_pack_state read it
 * (the reference has no counterpart: Lite and Puma are double throughout). */
int vk_set_precision(vk_particles *h, int32_t precision);
int vk_get_precision(const vk_particles *h, int32_t *precision_out);
/* numBlocksInScope("default")  (lite/trace.py) */
int vk_num_blocks(const vk_particles *h, int32_t scope, int64_t *n_out);

/* ---- raw device buffers, for NCCL plumbing in the host layer ------------
 * which: 0 = log weights (P doubles), 1.. = model state arrays.
 * Returns device pointer, element size, rows (per-particle stride semantics
 * in DESIGN.md).  The pointers stay valid until the next resample. */
int vk_device_buffer(vk_particles *h, int32_t which, void **dev_ptr_out,
                     int64_t *n_bytes_out);

/* ---- multi-GPU plumbing ------------------------------------
 * One process per GPU owns one vk_particles shard; the host layer
 * (venturecxx_b200/distributed.py) moves the few scalars and the packed
 * ancestor rows with NCCL (torch.distributed).  Pointers named *_dev are
 * device pointers supplied by the caller (e.g. torch tensors' data_ptr).
 *
 * Sharded resample (E2): local max -> allreduce(max) -> local fixed-point CDF
 * against the GLOBAL max -> allgather(mass) -> ancestor search for the children
 * whose threshold falls in this shard's mass interval -> pack rows ->
 * all_to_all -> unpack into slots. */
int vk_shard_weight_max(vk_particles *h, double *max_out_host);
/* out3_host = (max ignoring NaN, min ignoring NaN, 1.0 if any weight is NaN): what utils.logsumexp
 * (lite/utils.py:77-91) needs from every shard for its +-inf / NaN cases. */
int vk_shard_weight_stats(vk_particles *h, double *out3_host);
int vk_shard_build_cdf(vk_particles *h, double global_max, int32_t frac_bits,
                       uint64_t *mass_out_host, double *sumexp_out_host);
/* anc_out_dev[t] (t < k_count) = local ancestor of child k_begin+t of
 * n_new_total, or -1 if its threshold lies outside [offset, offset+mass). */
int vk_shard_ancestors(vk_particles *h, uint64_t offset, uint64_t total,
                       int64_t n_new_total, int64_t k_begin, int64_t k_count,
                       int mode, uint32_t step, const double *uniforms_host,
                       int32_t *anc_out_dev);
int vk_state_row_bytes(vk_particles *h, int64_t *bytes_out);
int vk_pack_state(vk_particles *h, const int32_t *particles_dev, int64_t n,
                  void *packed_out_dev);
int vk_unpack_state(vk_particles *h, const int32_t *slots_dev, int64_t n,
                    const void *packed_in_dev);
int vk_fill_log_weights(vk_particles *h, double value);

/* Sharded subsampled MH (E3): observations are sharded, particles replicated.
 * This shard holds global observation indices [global_offset, global_offset+N)
 * of n_total.  The round-level primitives let the host all-reduce the
 * (sum dllh, sum dllh^2) pairs between vk_ssmh_round and vk_ssmh_decide;
 * vk_subsampled_mh is the same loop on one GPU. */
int vk_blr_set_observation_shard(vk_particles *h, int64_t global_offset,
                                 int64_t n_total);
int vk_ssmh_begin(vk_particles *h);
int vk_ssmh_propose(vk_particles *h, int32_t block, int32_t t_index,
                    int32_t delta_mode, double delta_arg);
int vk_ssmh_round(vk_particles *h, int64_t q0, int64_t q1, void **sums_dev_out);
int vk_ssmh_decide(vk_particles *h, const void *sums_dev, int64_t n_end,
                   int32_t k0, double epsilon, int32_t *n_active_out);
int vk_ssmh_commit(vk_particles *h);
int vk_ssmh_stats(vk_particles *h, int32_t transitions, double *stats_out_host);

/* ---- SP plugin surface, batched (backend/lite/psp.py:32-318:
 *      simulate(args) / logDensity(value,args)) ---------------------------
 * params_host: n x n_params row-major.  For vector-valued SPs (dirichlet)
 * values are n x dim.  Draws use Philox(seed, particle=i, step). */
enum vk_sp_id {
  VK_SP_NORMAL = 1, VK_SP_GAMMA = 2, VK_SP_BETA = 3, VK_SP_FLIP = 4,
  VK_SP_LOG_FLIP = 5, VK_SP_LOG_ODDS_FLIP = 6, VK_SP_CATEGORICAL = 7,
  VK_SP_LOG_CATEGORICAL = 8, VK_SP_DIRICHLET = 9, VK_SP_SYM_DIRICHLET = 10,
  VK_SP_SYM_DIR_CAT = 11, VK_SP_BETA_BERNOULLI = 12,
  /* made SPs of the remaining Dirichlet / Bernoulli makers:
   * DIR_CAT        make_dir_cat's CDirCatOutputPSP, asymmetric alpha: params alpha[dim], counts[dim]
   *                (dirichlet.py:116-206)
   * UC_DIR_CAT     make_uc_dir_cat / make_uc_sym_dir_cat's UDirCatOutputPSP: params theta[dim]
   *                (dirichlet.py:265-293, 355-415)
   * SUFF_BERNOULLI make_uc_beta_bernoulli's SuffBernoulliOutputPSP: params weight[, yes, no]
   *                (discrete.py:513-545)
   * The MAKERS' own simulate / logDensity are VK_SP_DIRICHLET / VK_SP_SYM_DIRICHLET / VK_SP_BETA
   * (theta ~ dirichlet(alpha), logDensityDirichlet(theta, alpha); weight ~ beta(a, b)). */
  VK_SP_DIR_CAT = 13, VK_SP_UC_DIR_CAT = 14, VK_SP_SUFF_BERNOULLI = 15
};
int vk_sp_simulate(int32_t sp, const double *params_host, int32_t n_params,
                   int64_t n, int32_t dim, uint64_t seed, uint32_t step,
                   int device, double *values_out_host);
int vk_sp_log_density(int32_t sp, const double *values_host,
                      const double *params_host, int32_t n_params, int64_t n,
                      int32_t dim, int device, double *out_host);

/* logDensityOfData(aux) of a made SP from its sufficient statistics (psp.py:266-270;
 * dirichlet.py:192-201 CDirCatOutputPSP, discrete.py:453-459 CBetaBernoulliOutputPSP, :537-541
 * SuffBernoulliOutputPSP): what DeterministicMakerAAALKernel.weight (lkernel.py:158-171) and
 * PosteriorAAALKernel.weight = marginalLogDensityOfData (lkernel.py:137-156) evaluate.
 * params per row: SYM_DIR_CAT alpha, counts[dim]; DIR_CAT alpha[dim], counts[dim];
 * BETA_BERNOULLI a, b, yes, no; SUFF_BERNOULLI weight, yes, no. */
int vk_sp_log_density_of_data(int32_t sp, const double *params_host, int32_t n_params,
                              int64_t n, int32_t dim, int device, double *out_host);
/* The uncollapsed makers' AAA LKernels (UDirCatAAALKernel dirichlet.py:250-263,
 * USymDirCatAAALKernel :395-407, UBetaBernoulliAAALKernel discrete.py:502-510): row i draws the
 * made SP's parameter from its conjugate posterior given the statistics -- theta ~
 * dirichlet(alpha + counts) (UC_DIR_CAT: params alpha[dim], counts[dim]; SYM_DIRICHLET id for
 * the symmetric maker: params alpha, counts[dim]) or weight ~ beta(a + yes, b + no)
 * (SUFF_BERNOULLI: params a, b, yes, no) -- from Philox(seed, i, step), and reports the kernel's
 * weight, the marginal likelihood of the data (nullable). */
int vk_sp_aaa_simulate(int32_t sp, const double *params_host, int32_t n_params, int64_t n,
                       int32_t dim, uint64_t seed, uint32_t step, int device,
                       double *values_out_host, double *weights_out_host);

#ifdef __cplusplus
}
#endif
#endif /* VENTURE_B200_H */

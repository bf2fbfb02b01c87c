// Factories and model-specific entry points behind the C ABI.
#pragma once
#include "vk_common.cuh"

namespace vk {
Model *make_normal_normal(vk_particles *h, double mu0, double s0, double s, int64_t max_obs);
void normal_normal_observe(Model *m, const double *y_host, int64_t n);

Model *make_blr(vk_particles *h, int D, int64_t Ncap, double mu0, double s0);
void blr_observe(Model *m, const double *X_rows_host, const uint8_t *y_host, int64_t n);

void blr_set_observation_shard(Model *m, int64_t lo, int64_t total);
void blr_ss_begin(Model *m);
void blr_ss_propose(Model *m, int block, int t_index, int delta_mode, double eps);
double *blr_ss_round(Model *m, int64_t q0, int64_t q1);
int blr_ss_decide(Model *m, const double *sums_dev, int64_t n_end, int k0, double eps);
void blr_ss_commit(Model *m);
void blr_ss_stats(Model *m, int K, double *out_dev);

Model *make_hmm(vk_particles *h, int S, int O, int n_trans_rows, const double *trans_host,
                const double *emit_host, int init_state, int T_max);
void hmm_observe(Model *m, int symbol);

Model *make_lda(vk_particles *h, int K, int V, int n_docs, int max_tokens, double hyper_shape,
                double hyper_rate);
void lda_observe(Model *m, const int32_t *doc_host, const int32_t *word_host, int64_t n);

Model *make_program(vk_particles *h, const int32_t *code, int64_t n_code, const double *consts, int64_t n_consts,
                    const int32_t *meta, int64_t n_meta, const double *data, int64_t n_data, const double *affected);

void program_resimulate_from(Model *m, int first_site);

void launch_axpy(vk_particles *h, double *y, const double *x, int64_t n);  // y += x
}  // namespace vk

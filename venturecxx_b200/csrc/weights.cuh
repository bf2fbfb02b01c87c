// Host-side entry points of k_weights.cu (see that file for semantics).
#pragma once
#include "vk_common.cuh"
#include "p2p_common.cuh"

namespace vk {
void device_logsumexp(const double *a, int64_t n, double *red, unsigned int *counter,
                      int sm_count, cudaStream_t st);
int frac_bits_for(int64_t P_total);
void build_fixed_cdf(vk_particles *h, const double *red_or_null, double m, int frac_bits);
int gather_state(vk_particles *h, const int32_t *anc_dev, int64_t n_new, double *lw_out, double log_n_new,
                 std::vector<StateArray> *explicit_arrs = nullptr);
void resample(vk_particles *h, int64_t n_new, int mode, const double *uniforms_dev);
void launch_normalized_probs(vk_particles *h, double *out_dev);
void launch_ancestor_search_single(vk_particles *h, int64_t n_new, int mode, uint32_t step,
                                   const double *uniforms_dev, int grid);
void launch_ancestor_search_shard(vk_particles *h, uint64_t offset, uint64_t total, int64_t n_new,
                                  int64_t k_begin, int64_t k_count, int mode, uint32_t step,
                                  const double *uniforms_dev, int32_t *anc_out_dev);
void launch_fill_even_weights(vk_particles *h, int64_t n, double lse, double log_n_new);
size_t packed_row_bytes(vk_particles *h);
void pack_state(vk_particles *h, const int32_t *idx_dev, int64_t n, void *packed_dev, bool unpack);
void shard_weight_max(vk_particles *h);
void shard_build_cdf(vk_particles *h, double global_max, int frac_bits);
int p2p_export(vk_particles *h, unsigned char *handles_out, int max_handles);
void p2p_open(vk_particles *h, int world, int rank, const unsigned char *all_handles, int n_handles);
bool p2p_is_open(vk_particles *h);
bool p2p_begin_exchange(vk_particles *h, int phase, P2PXchg *x);
void p2p_check_error(vk_particles *h);
bool p2p_ss_view(vk_particles *h, P2PSSView *v);
void p2p_resample(vk_particles *h, int mode);
void p2p_status(vk_particles *h, int *err_out, double *lse_out);
void p2p_release(vk_particles *h);
}  // namespace vk

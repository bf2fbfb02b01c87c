// Entry points not yet implemented in this build step.
#include "vk_common.cuh"
#include "models.cuh"
namespace vk {
Model *make_hmm(vk_particles *, int, int, int, const double *, const double *, int, int) { throw Error(VK_EUNSUPPORTED, "hmm: not built yet"); }
void hmm_observe(Model *, int) { throw Error(VK_EUNSUPPORTED, "hmm: not built yet"); }
Model *make_lda(vk_particles *, int, int, int, int, double, double) { throw Error(VK_EUNSUPPORTED, "lda: not built yet"); }
void lda_observe(Model *, const int32_t *, const int32_t *, int64_t) { throw Error(VK_EUNSUPPORTED, "lda: not built yet"); }
}
extern "C" {
int vk_shard_weight_max(vk_particles *, double *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
int vk_shard_weight_mass(vk_particles *, double, int32_t, uint64_t *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
int vk_sp_simulate(int32_t, const double *, int32_t, int64_t, int32_t, uint64_t, uint32_t, int, double *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
int vk_sp_log_density(int32_t, const double *, const double *, int32_t, int64_t, int32_t, int, double *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
}

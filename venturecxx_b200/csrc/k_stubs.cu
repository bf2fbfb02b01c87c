// Entry points not yet implemented in this build step.
#include "vk_common.cuh"
extern "C" {
int vk_shard_weight_max(vk_particles *, double *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
int vk_shard_weight_mass(vk_particles *, double, int32_t, uint64_t *) { vk::set_last_error("not built yet"); return VK_EUNSUPPORTED; }
}

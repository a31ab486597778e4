// placeholder while the first slice goes to the GPU (replaced by the LBVH implementation)
#include "scene_impl.h"
namespace gb {
cudaError_t collision_init(Scene&) { return cudaSuccess; }
void collision_destroy(Scene&) {}
cudaError_t collision_dcd(Scene&, bool) { return cudaSuccess; }
cudaError_t collision_ccd(Scene&) { return cudaSuccess; }
cudaError_t collision_prepare(Scene&) { return cudaSuccess; }
cudaError_t collision_update_infos(Scene&, int) { return cudaSuccess; }
int64_t collision_fetch_contacts(Scene&, gaia_contact*, int64_t) { return 0; }
int64_t collision_leaf_order(Scene&, int, int32_t*, int64_t) { return 0; }
bool collision_contact_view(Scene&, ContactView&, const uint8_t**) { return false; }
}

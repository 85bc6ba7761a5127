// vcm.cu — placeholder until the wavefront VCM/UPG lands.
#include "render_common.cuh"
namespace b2pt {
int render_vcm(b2pt_scene_s*, const RenderConfig&, cudaStream_t, B2ptStats*) { set_error("VCM/UPG: not built yet"); return B2PT_ERR_INVALID; }
}

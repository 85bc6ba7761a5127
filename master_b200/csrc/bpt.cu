// bpt.cu — placeholder until the wavefront BPT lands (next milestone).
#include "render_common.cuh"
namespace b2pt {
int render_bpt(b2pt_scene_s*, const RenderConfig&, cudaStream_t, B2ptStats*) { set_error("BPT: not built yet"); return B2PT_ERR_INVALID; }
}

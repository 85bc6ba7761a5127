// grid.cu — placeholder until the sorted hash grid lands.
#include "common.cuh"
extern "C" {
int b2pt_grid_build(const float*, size_t, float, b2pt_grid_t*) { b2pt::set_error("grid: not built yet"); return B2PT_ERR_INVALID; }
void b2pt_grid_destroy(b2pt_grid_t) {}
int b2pt_grid_query(b2pt_grid_t, const float*, size_t, uint32_t, uint32_t*, uint32_t*) { b2pt::set_error("grid: not built yet"); return B2PT_ERR_INVALID; }
}

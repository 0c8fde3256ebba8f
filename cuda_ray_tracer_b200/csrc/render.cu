// render.cu — placeholder until the wavefront renderer lands (next commit).
#include "common.cuh"
int rt_render_free_state(rt_scene*) { return RT_OK; }
extern "C" int rt_render(rt_scene*, int, int, int, unsigned, const rt_tile_set*, uint8_t*, int32_t*, uint8_t*, rt_counters*, void*) {
  rt_set_error("rt_render: not implemented yet"); return RT_ERR_INVALID;
}
extern "C" int rt_render_host(rt_scene*, int, int, int, unsigned, const rt_tile_set*, uint8_t*, int32_t*, uint8_t*, rt_counters*) {
  rt_set_error("rt_render_host: not implemented yet"); return RT_ERR_INVALID;
}
extern "C" int rt_trace(rt_scene*, int, const float*, const float*, unsigned, int32_t*, float*, rt_counters*, void*) {
  rt_set_error("rt_trace: not implemented yet"); return RT_ERR_INVALID;
}

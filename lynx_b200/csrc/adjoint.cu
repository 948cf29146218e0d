// placeholder, replaced below
#include "common.cuh"
extern "C" size_t lynx_bwd_workspace_bytes(const lynx_program_t*, int64_t) { return 0; }
extern "C" int lynx_track_particles_bwd(const lynx_program_t*, const float*, int64_t, const float*, int64_t, int64_t, int64_t, const float*, int64_t, const double*, const double*, const float*, float*, float*, int64_t, void*, lynx_stream_t) { return LYNX_E_BADARG; }
extern "C" int lynx_parambeam_bwd(const lynx_program_t*, const float*, int64_t, const float*, const float*, int64_t, const float*, const float*, const float*, const float*, float*, float*, float*, lynx_stream_t) { return LYNX_E_BADARG; }

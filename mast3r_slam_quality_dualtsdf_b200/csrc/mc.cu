// mc.cu -- marching-cubes extraction (placeholder until the kernel lands).
#include "common.cuh"
extern "C" int b200tsdf_extract_mesh(b200tsdf_t *h, int which, float *verts_dev, int64_t tri_cap,
                                     int64_t *n_tris_host, void *stream) {
  b200_set_error("b200tsdf_extract_mesh: not built yet");
  return B200TSDF_E_INVALID;
}

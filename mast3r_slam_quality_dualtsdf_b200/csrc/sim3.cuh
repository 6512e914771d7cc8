// sim3.cuh -- the one Sim3 operation the TSDF path needs on the device.
#pragma once

// Sim3 act = s * R(q) x + t with pose = [t(3), q(xyzw), s]; float32 operation
// order of oracle/tsdf_oracle.c:364-376 (restating the reference's own
// actSO3/actSim3, backend/src/gn_kernels.cu:195-221).
__device__ __forceinline__ void sim3_act(const float *__restrict__ P, float x0, float x1, float x2,
                                         float &y0, float &y1, float &y2) {
  const float q0 = P[3], q1 = P[4], q2 = P[5], q3 = P[6], s = P[7];
  float uv0 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q1, x2), __fmul_rn(q2, x1)));
  float uv1 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q2, x0), __fmul_rn(q0, x2)));
  float uv2 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q0, x1), __fmul_rn(q1, x0)));
  float a0 = __fadd_rn(__fadd_rn(x0, __fmul_rn(q3, uv0)),
                       __fsub_rn(__fmul_rn(q1, uv2), __fmul_rn(q2, uv1)));
  float a1 = __fadd_rn(__fadd_rn(x1, __fmul_rn(q3, uv1)),
                       __fsub_rn(__fmul_rn(q2, uv0), __fmul_rn(q0, uv2)));
  float a2 = __fadd_rn(__fadd_rn(x2, __fmul_rn(q3, uv2)),
                       __fsub_rn(__fmul_rn(q0, uv1), __fmul_rn(q1, uv0)));
  y0 = __fadd_rn(__fmul_rn(a0, s), P[0]);
  y1 = __fadd_rn(__fmul_rn(a1, s), P[1]);
  y2 = __fadd_rn(__fmul_rn(a2, s), P[2]);
}

// Task-head transforms (elementwise, per event row).
//
// The `_forward` of the reconstruction heads in the reference,
// models/task/reconstruction.py:11-112, applied to the output of the task's affine
// layer (models/task/task.py:293).  eps = float32 machine epsilon
// (utilities/maths.py:6-8).
#include <float.h>
#include <math.h>

#include "common.cuh"

namespace gn {

__global__ void task_head_kernel(const float* __restrict__ y, int64_t M, int p, int kind,
                                 float* __restrict__ out) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= M) return;
  const float eps = FLT_EPSILON;
  const float* in = y + r * p;
  switch (kind) {
    case GN_HEAD_ENERGY: {
      // softplus(x, beta=0.05, threshold=20) + eps   (torch.nn.functional.softplus)
      for (int c = 0; c < p; ++c) {
        const float v = in[c], bv = 0.05f * v;
        const float sp = bv > 20.f ? v : log1pf(expf(bv)) / 0.05f;
        out[r * p + c] = sp + eps;
      }
      break;
    }
    case GN_HEAD_DIRECTION_KAPPA: {
      const float a = in[0], b = in[1], c = in[2];
      const float kappa = sqrtf(a * a + b * b + c * c) + eps;
      out[r * 4 + 0] = a / kappa;
      out[r * 4 + 1] = b / kappa;
      out[r * 4 + 2] = c / kappa;
      out[r * 4 + 3] = kappa;
      break;
    }
    case GN_HEAD_ZENITH: {
      out[r] = (1.f / (1.f + expf(-in[0]))) * 3.14159265358979323846f;
      break;
    }
    case GN_HEAD_ZENITH_KAPPA: {
      out[r * 2 + 0] = (1.f / (1.f + expf(-in[0]))) * 3.14159265358979323846f;
      out[r * 2 + 1] = fabsf(in[1]) + eps;
      break;
    }
    case GN_HEAD_AZIMUTH_KAPPA: {
      const float a = in[0], b = in[1];
      float ang = atan2f(b, a);
      if (ang < 0.f) ang += 6.28318530717958647692f;
      out[r * 2 + 0] = ang;
      out[r * 2 + 1] = sqrtf(a * a + b * b) + eps;
      break;
    }
    default: {
      for (int c = 0; c < p; ++c) out[r * p + c] = in[c];
    }
  }
}

}  // namespace gn

using namespace gn;

extern "C" int gn_task_head(const float* y, int64_t M, int p, int kind, float* out, void* stream) {
  GN_CHECK_ARG(M >= 0 && p >= 1, "gn_task_head: bad sizes");
  GN_CHECK_ARG(kind >= GN_HEAD_IDENTITY && kind <= GN_HEAD_AZIMUTH_KAPPA,
               "gn_task_head: unknown head %d", kind);
  const int need = kind == GN_HEAD_DIRECTION_KAPPA ? 3
                 : (kind == GN_HEAD_ZENITH_KAPPA || kind == GN_HEAD_AZIMUTH_KAPPA) ? 2
                 : 1;
  GN_CHECK_ARG(p >= need, "gn_task_head: head %d needs %d inputs, got %d", kind, need, p);
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(y && out, "gn_task_head: null pointer");
  task_head_kernel<<<(unsigned)ceil_div(M, 128), 128, 0, as_stream(stream)>>>(y, M, p, kind, out);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

// Host launchers (one kernel family per translation unit, see launch_impl.cuh).
#include "launch_impl.cuh"

namespace b200kv {

int launch_adam(const MtaCall& c, const OptScalars& s, const float* gscale) {
  if (c.staged) {
    const int r = launch_adam_staged(c, s, gscale);
    if (r >= 0) return r;
  }
  AdamOp op{s.beta1, s.beta2, s.eps, s.rescale, s.clip};
  return launch_update(c, op, gscale, nullptr);
}

int launch_adagrad(const MtaCall& c, const OptScalars& s, const float* gscale) {
  AdagradOp op{s.eps, s.rescale, s.clip};
  return launch_update(c, op, gscale, nullptr);
}

int launch_adamw(const MtaCall& c, const OptScalars& s, const float* rescale_dev) {
  AdamWOp op{s.beta1, s.beta2, s.eps, s.clip, rescale_dev};
  return launch_update(c, op, nullptr, nullptr);
}

}  // namespace b200kv

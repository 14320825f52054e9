// Host launchers (one kernel family per translation unit, see launch_impl.cuh).
#include "launch_impl.cuh"

namespace b200kv {

int launch_sgd(const MtaCall& c, bool has_mom, const OptScalars& s, const float* gscale) {
  if (c.staged) {
    const int r = launch_sgd_staged(c, has_mom, s, gscale);
    if (r >= 0) return r;
  }
  if (has_mom) {
    SgdOp<true> op{s.momentum, s.rescale, s.clip};
    return launch_update(c, op, gscale, nullptr);
  }
  SgdOp<false> op{0.f, s.rescale, s.clip};
  return launch_update(c, op, gscale, nullptr);
}

}  // namespace b200kv

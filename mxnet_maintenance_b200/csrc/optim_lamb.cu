// Host launchers (one kernel family per translation unit, see launch_impl.cuh).
#include "launch_impl.cuh"

namespace b200kv {

int launch_lamb1(const MtaCall& c, const OptScalars& s, const float* gscale, float* partials, bool store_u) {
  if (store_u) {
    Lamb1Op<true> op{s.beta1, s.beta2, s.eps, s.rescale, s.clip, s.bias_correction};
    return launch_update(c, op, gscale, partials);
  }
  Lamb1Op<false> op{s.beta1, s.beta2, s.eps, s.rescale, s.clip, s.bias_correction};
  return launch_update(c, op, gscale, partials);
}

int launch_lamb2u(const MtaCall& c, const float* ratio) {
  Lamb2UOp op{ratio};
  return launch_update(c, op, nullptr, nullptr);
}

int launch_lamb2(const MtaCall& c, const OptScalars& s, const float* ratio) {
  Lamb2Op op{s.eps, s.bias_correction, ratio};
  return launch_update(c, op, nullptr, nullptr);
}

int launch_sgd_plain(const MtaCall& c, const OptScalars& s) {
  SgdPlainOp op{s.rescale, s.clip};
  return launch_update(c, op, nullptr, nullptr);
}

int launch_lamb_phase1(const MtaCall& c, const OptScalars& s) {
  KV_CHECK(c.dtype == kF32, "lamb_update_phase1 is float32");
  LambPhase1Op op{s.beta1, s.beta2, s.eps, s.rescale, s.clip, s.bias_correction};
  return launch_ng<LambPhase1Op, float, float>(c, op, nullptr, nullptr);
}

int launch_lamb_phase2(const MtaCall& c, const float* r1, const float* r2, float lower, float upper) {
  KV_CHECK(c.dtype == kF32, "lamb_update_phase2 is float32");
  LambPhase2Op op{r1, r2, lower, upper};
  return launch_ng<LambPhase2Op, float, float>(c, op, nullptr, nullptr);
}

int launch_sumsq(const MtaCall& c, float* partials) {
  SumSqOp op;
  if (c.dtype == kF32) return launch_ng<SumSqOp, float, float>(c, op, nullptr, partials);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<SumSqOp, __half, __half>(c, op, nullptr, partials);
}

int launch_nonfinite(const MtaCall& c, float* partials) {
  NonFiniteOp op;
  if (c.dtype == kF32) return launch_ng<NonFiniteOp, float, float>(c, op, nullptr, partials);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<NonFiniteOp, __half, __half>(c, op, nullptr, partials);
}

}  // namespace b200kv

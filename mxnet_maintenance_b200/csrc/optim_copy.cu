// Host launchers (one kernel family per translation unit, see launch_impl.cuh).
#include "launch_impl.cuh"

namespace b200kv {

int launch_copy(const MtaCall& c, bool to_w, bool to_out, bool w_is_f32) {
  if (c.staged && to_w && !to_out && !w_is_f32) {
    const int r = launch_copy_staged(c);
    if (r >= 0) return r;
  }
#define GO(TW, TO)                                                                          \
  do {                                                                                      \
    CopyOp<TW, TO> op;                                                                      \
    if (c.dtype == kF32) return launch_ng<CopyOp<TW, TO>, float, float>(c, op, nullptr, nullptr); \
    KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);                             \
    if (w_is_f32) return launch_ng<CopyOp<TW, TO>, __half, float>(c, op, nullptr, nullptr);  \
    return launch_ng<CopyOp<TW, TO>, __half, __half>(c, op, nullptr, nullptr);               \
  } while (0)
  if (to_w && to_out) GO(true, true);
  if (to_w) GO(true, false);
  KV_CHECK(to_out, "launch_copy: nothing to write");
  GO(false, true);
#undef GO
}

int launch_scale(const MtaCall& c, const float* gscale) {
  CopyOp<false, true> op;
  if (c.dtype == kF32) return launch_ng<CopyOp<false, true>, float, float>(c, op, gscale, nullptr);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<CopyOp<false, true>, __half, __half>(c, op, gscale, nullptr);
}

int launch_master_to_out(const MtaCall& c) {
  MasterToOutOp op;
  KV_CHECK(c.dtype == kF16, "master_to_out is for fp16 outputs");
  return launch_ng<MasterToOutOp, __half, float>(c, op, nullptr, nullptr);
}

int launch_zero(const MtaCall& c) {
  ZeroOp op;
  if (c.dtype == kF32) return launch_ng<ZeroOp, float, float>(c, op, nullptr, nullptr);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<ZeroOp, __half, __half>(c, op, nullptr, nullptr);
}

}  // namespace b200kv

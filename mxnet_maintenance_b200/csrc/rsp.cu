// Row-sparse keys (placeholder until the sort / segmented-reduce kernels land).
#include "store.h"

namespace b200kv {

struct RspKey {};

void Store::init_rsp(int, DLManagedTensor*) { KV_CHECK(false, "row-sparse keys: not built yet"); }
void Store::push_rsp(int, int, DLManagedTensor* const*, DLManagedTensor* const*) {
  KV_CHECK(false, "row-sparse keys: not built yet");
}
int64_t Store::rsp_merged_size(int) { KV_CHECK(false, "row-sparse keys: not built yet"); return 0; }
void Store::rsp_merged_copy(int, DLManagedTensor*, DLManagedTensor*) {
  KV_CHECK(false, "row-sparse keys: not built yet");
}
int64_t Store::row_sparse_pull(int, DLManagedTensor*, DLManagedTensor*, DLManagedTensor*) {
  KV_CHECK(false, "row-sparse keys: not built yet");
  return 0;
}
void Store::set_hyper_rsp(int, float, float, int) {}

}  // namespace b200kv

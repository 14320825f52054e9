// Multi-tensor-apply tables shared by host and device code.
//
// One kernel launch covers every parameter array of a model: the launch gets a device-resident
// table of per-tensor base pointers / sizes / scalars and a list of fixed-size chunks; a block
// works on one chunk at a time.  This replaces the reference's by-value MultiSGDKernelParam /
// MultiLAMBKernelParam structs (src/operator/optimizer_op-inl.h:207-223, contrib/multi_lamb-inl.h:191-213)
// that cap a launch at 60 / 45 tensors and make thread i touch element i of *every* tensor.
#pragma once
#include <stdint.h>

namespace b200kv {

// elements per chunk: 256 threads x 4 elements x 4 unrolled vectors
constexpr int kChunk = 4096;
constexpr int kThreads = 256;
constexpr int kMaxCopies = 8;   // local gradient / output copies per key handled in one launch

struct Chunk {          // 16 bytes
  int32_t tensor;       // row of the tensor table
  int32_t len;          // elements in this chunk (<= kChunk)
  int64_t off;          // element offset inside the tensor
};

// The table is one blob of 8-byte words laid out as consecutive sections of `nt` words each:
//   grad[0..ngrad) | out[0..nout) | w | s0 | s1 | x | (lr,wd) | (aux0,aux1)
// Pointers are per-tensor element-0 addresses; a section may hold nulls when unused.
// Cross-GPU flag blocks of the peer-memory path (see optim.cuh): flags[k] is rank k's block
// [ready[8] | done[8] | error | cta counter] as mapped into this process.
struct P2PFlags {
  uint32_t* flags[kMaxCopies];
  int world, rank;
  uint32_t epoch;
};

struct MtaArgs {
  const uint64_t* blob;
  const Chunk* chunks;
  int nchunks;
  int nt;
  int ngrad;
  int nout;
  // optional in-kernel cross-GPU barriers: bit 0 = wait until every rank's gradient bucket is
  // ready before the first load; bit 1 = after the last store signal "my shard has landed" and
  // wait for every rank's signal
  const P2PFlags* sync;
  uint32_t epoch;
  int sync_mode;   // kSync* bits
};
// sync_mode bits: the two barriers, and NVLS multicast addressing of the gradient / out sections
// (multimem.ld_reduce / multimem.st; vector path only — multicast buffers are 16-byte aligned
// bucket slots whose sizes the store pads to whole vectors)
enum SyncBits { kSyncWaitReady = 1, kSyncSignalDone = 2, kSyncMcOut = 4, kSyncMcGrad = 8 };

inline size_t mta_blob_words(int nt, int ngrad, int nout) {
  return static_cast<size_t>(nt) * static_cast<size_t>(ngrad + nout + 6);
}

// kSecX: optional fp32 scratch array per tensor (LAMB keeps its update direction there between passes)
enum Section { kSecW = 0, kSecS0 = 1, kSecS1 = 2, kSecX = 3, kSecLrWd = 4, kSecAux = 5 };

}  // namespace b200kv

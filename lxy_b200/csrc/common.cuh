// common.cuh — shared declarations of liblxyb200 (context, error plumbing, device formats).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/lxy_b200.h"

namespace lxy {

// ---- error plumbing ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
#define LXY_CUDA(expr)                                                                     \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      ::lxy::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__,   \
                       __LINE__);                                                          \
      return LXY_ERR_CUDA;                                                                 \
    }                                                                                      \
  } while (0)
#define LXY_TRY(expr)          \
  do {                         \
    int rc_ = (expr);          \
    if (rc_ != LXY_OK) return rc_; \
  } while (0)

// ---- device formats ---------------------------------------------------------------------------
// Contigs are grouped into blocks of kBlk consecutive ids; a link (a < b) lives in tile
// (a / kBlk, b / kBlk) and is stored as the two block-local ids, one byte each.
constexpr int kBlk = 256;
constexpr int kBlkShift = 8;
// The decay table T of scaff.go:252-302 times 100 (exact integers), indexed by d + kBias where
// d = pos a - pos b in [-20, 20]; everything else clamps to kMiss.
constexpr int kBias = 20;
constexpr int kMiss = 41;
constexpr int kLutLen = 48;  // u8 entries, padded

enum WeightMode : int { W_UNIT = 0, W_INT32 = 1, W_F64 = 2 };

// A growable device buffer (never shrinks; freed with the context).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);
  void release();
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Tile-sorted link store of the scaffolding scorers.
struct ScaffLinks {
  bool ready = false;
  bool gprofile = false;
  int32_t n = 0;            // ids
  int64_t L = 0;            // records of the store (after summing duplicates, when that was done)
  int64_t L_in = 0;         // records handed to set_links
  int path = LXY_PATH_STREAM;  // what a LXY_PATH_AUTO context chose for this store (STREAM or PAIRS)
  int32_t nblk = 0;         // ceil(n / kBlk)
  int64_t ntiles = 0;       // nblk * (nblk + 1) / 2
  int wmode = W_UNIT;
  bool wide_pos = false;    // positions need 32 bits (n > kMaxN16)
  DevBuf rec;               // R: uint16 (a_loc << 8 | b_loc) sorted by (tile, a_loc, b_loc)
                            // G: uint2 {a_loc << 8 | b_loc | ..., xa, xb} see scaffg
  DevBuf w;                 // int32 or double per record (absent for W_UNIT)
  DevBuf tile_start;        // int64[ntiles + 1]
  // G-profile only
  DevBuf len;               // uint32[n]
  DevBuf xa, xb;            // uint32 in-contig offsets per record, same order as rec
  uint32_t unit = 1;
  uint64_t total_len = 0;
  // pair-table path (cfg.path == LXY_PATH_PAIRS): dense symmetric n x n table, int32 or double
  DevBuf dense;
  bool dense_ready = false;
  uint64_t version = 0;     // bumped by every build: cached launch graphs of older stores are dropped
};

// A scoring call captured as a CUDA graph: a GA loop (or a host that scores single orderings) calls
// with the same buffers over and over, and for small problems the 15 or so runtime calls of one
// scoring pass cost more host time than the kernels take on the device.
struct ScoreGraph {
  const void* perm = nullptr;
  void* out = nullptr;
  void* x100 = nullptr;
  int32_t P = 0;
  uint64_t version = 0;
  int seen = 0;                    // identical calls so far (captured on the second)
  bool failed = false;             // capture was refused once: stay on the direct path
  cudaGraphExec_t exec = nullptr;
  lxy_timing timing{};
  uint64_t stamp = 0;              // last use (eviction)
  const void* scratch[6] = {};     // the context's scratch buffers the graph writes to: a later, larger
                                   // call may have reallocated them
};

struct PhaseLinks {
  bool ready = false;
  int32_t n = 0;
  int64_t L = 0;            // records kept (valid step only)
  int64_t L_in = 0;
  DevBuf rec;               // uint2 (lo, hi) sorted by lo, see phase_kernels.cu
  DevBuf w;                 // int32 w * m
  int64_t sum_w = 0;        // sum of w * m over kept records
  bool narrow = true;       // max |w * m| * chunk < 2^31: 32-bit chunk sums
};

}  // namespace lxy

struct lxy_ctx {
  lxy_cfg cfg{};
  std::mutex mu;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;   // side stream (head-of-ordering look-ups)
  cudaEvent_t ev_join = nullptr;
  cudaEvent_t ev_fork = nullptr;
  cudaEvent_t ev[8]{};
  cudaEvent_t ev_slice[8]{};        // upload slices of a packed phasing population (api.cu)
  bool capturing = false;           // the scoring pass is being captured into a graph
  lxy::ScoreGraph graphs[4];
  uint64_t graph_clock = 0;
  int sm_count = 0;
  int max_smem_optin = 0;
  lxy_timing timing{};
  int timing_pending = 0;     // 1: device-only call, 2: call with host copies (events not read yet)
  lxy::ScaffLinks scaff;
  lxy::PhaseLinks phase;
  // per-call scratch
  lxy::DevBuf d_perm, d_ori, d_out, d_x100, d_posA, d_posB, d_partial, d_corr, d_flag, d_genes,
      d_gbits, d_tmp0, d_tmp1, d_tmp2, d_tmp3, d_start;
  int* h_deferred = nullptr;  // pinned: validation flags of calls not yet reported (LXY_FLAG_DEFER_ERRORS)
  int n_deferred = 0;
  void* h_pin = nullptr;      // pinned staging for populations / scores
  size_t h_pin_cap = 0;
  // multi-GPU (comm.cu): NCCL communicator of this context's device, collectives run on `stream`
  void* comm = nullptr;       // ncclComm_t
  int32_t world = 1, rank = 0;
  lxy::DevBuf d_shard, d_gather;  // send / receive buffers of the score collectives
  uint64_t agree_version = 0; // link-store version the ranks last agreed an accumulation mode for
  int agree_f64 = 0;          // some rank's shard carries non-integral weights
};

namespace lxy {
// Guards the process-global "kernel attributes set for device d" flags next to the launch helpers
// (two contexts of one device may be driven from two threads; the context mutex does not cover them).
std::mutex& config_mutex();
int pin_reserve(lxy_ctx* c, size_t bytes);
// Validation flags of the call just queued: read back now (a host sync) or, with LXY_FLAG_DEFER_ERRORS,
// copied asynchronously and reported by report_deferred_flags() (lxy_ctx_sync).
int read_flags(lxy_ctx* c, const char* what, const char* where);
int report_deferred_flags(lxy_ctx* c);
void drop_score_graphs(lxy_ctx* c);
// timing event k of the context on its main stream; inside a capture the record must be an
// "external" one to become a node whose time stamp can be read after the graph ran
inline cudaError_t record_timing_event(lxy_ctx* c, int k) {
  return cudaEventRecordWithFlags(c->ev[k], c->stream, c->capturing ? cudaEventRecordExternal : cudaEventRecordDefault);
}

// link_store.cu
struct LinkInput {
  int32_t n = 0;
  int64_t L = 0;
  const uint32_t *a = nullptr, *b = nullptr, *xa = nullptr, *xb = nullptr;
  const double* w = nullptr;
  const uint32_t* len = nullptr;  // host pointer; non-null selects the G-profile
  uint32_t unit = 1;
  bool on_device = false;         // a, b, w, xa, xb are device pointers
};
int build_scaff_links(lxy_ctx* c, const LinkInput& in);
int build_phase_links(lxy_ctx* c, int32_t n, int64_t L, const uint32_t* a, const uint32_t* b,
                      const double* w);
int links_get(lxy_ctx* c, int32_t id1, int32_t id2, double* out);

// scaff_kernels.cu
int scaff_score(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P,
                double* out_dev, int64_t* x100_dev);
int scaff_positions(lxy_ctx* c, const int32_t* perm_dev, int32_t P, int32_t n, int32_t* pos_dev);
int scaff_link_distances(lxy_ctx* c, int32_t n, int64_t L, const uint32_t* a_dev,
                         const uint32_t* b_dev, const int32_t* perm_dev, int64_t* dist_dev);

// scaffg_kernels.cu
int scaffg_score(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P,
                 double* out_dev, int64_t* x100_dev);
int scaffg_layout64(lxy_ctx* c, const int32_t* perm_dev, int32_t P, uint64_t* start_dev,
                    int32_t* pos_dev);
int scaffg_link_distances(lxy_ctx* c, int64_t L, const uint32_t* a, const uint32_t* xa,
                          const uint32_t* b, const uint32_t* xb, const uint64_t* start_dev,
                          const int32_t* pos_dev, const uint8_t* ori_dev, uint64_t* dist_dev);

// comm.cu
int sharded_score_dev(lxy_ctx* c, int32_t mode, const int32_t* perm_dev, int32_t P, double* out_dev,
                      int64_t* x100_dev, bool check_flags);
int scaff_dispatch(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev, int64_t* x100_dev);
int stage_h2d(lxy_ctx* c, void* dst_dev, const void* src, size_t bytes, size_t pin_off);

// pair_kernels.cu
int build_pair_table(lxy_ctx* c);
int pairs_score(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev, int64_t* x100_dev);

// phase_kernels.cu
int phase_score(lxy_ctx* c, const uint8_t* genes_dev, int32_t P, int64_t* out_i64_dev,
                double* out_f64_dev);
int phase_score_any(lxy_ctx* c, const uint8_t* genes_dev, const uint32_t* packed_dev, int32_t P,
                    int64_t* out_i64_dev, double* out_f64_dev);
}  // namespace lxy

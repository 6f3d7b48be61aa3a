// api.cu — the extern "C" boundary of liblxyb200.so (declared in include/lxy_b200.h).
#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>
#include <numeric>
#include <vector>

#include "common.cuh"

namespace lxy {

static thread_local std::string g_err;

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
}

std::mutex& config_mutex() {
  static std::mutex mu;
  return mu;
}

void drop_score_graphs(lxy_ctx* c) {
  for (auto& g : c->graphs) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    g = ScoreGraph();
  }
}

int pin_reserve(lxy_ctx* c, size_t bytes) {
  if (bytes <= c->h_pin_cap) return LXY_OK;
  if (c->h_pin) cudaFreeHost(c->h_pin);
  c->h_pin = nullptr;
  c->h_pin_cap = 0;
  size_t want = bytes + (bytes >> 2) + 4096;
  LXY_CUDA(cudaMallocHost(&c->h_pin, want));
  c->h_pin_cap = want;
  return LXY_OK;
}

// Host -> device copy of one call's inputs.  Caller memory that is already page-locked
// (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) is handed to the copy engine as it
// is; pageable memory (Go slices through cgo) goes through the context's pinned staging buffer at
// offset pin_off, in 2 MB chunks copied by up to four host threads, each chunk's DMA queued as soon as
// it is staged so the copy engine overlaps the remaining memcpys.  The caller's memory is not retained
// past the return of the entry point in either case (the entry points synchronise before returning).
int stage_h2d(lxy_ctx* c, void* dst_dev, const void* src, size_t bytes, size_t pin_off) {  // declared in common.cuh
  if (bytes == 0) return LXY_OK;
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, src) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
    LXY_CUDA(cudaMemcpyAsync(dst_dev, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return LXY_OK;
  }
  (void)cudaGetLastError();  // unregistered host memory reports an error on older drivers
  char* pin = (char*)c->h_pin + pin_off;
  constexpr size_t kChunk = (size_t)2 << 20;
  const size_t nchunks = (bytes + kChunk - 1) / kChunk;
  unsigned hw = std::thread::hardware_concurrency();
  const size_t nthreads = std::min<size_t>(std::min<size_t>(4, hw ? hw : 1), nchunks);
  std::atomic<int> err(0);
  auto work = [&](size_t t) {
    if (t != 0 && cudaSetDevice(c->cfg.device) != cudaSuccess) { err = 1; return; }
    for (size_t i = t; i < nchunks; i += nthreads) {
      const size_t off = i * kChunk, len = std::min(kChunk, bytes - off);
      std::memcpy(pin + off, (const char*)src + off, len);
      if (cudaMemcpyAsync((char*)dst_dev + off, pin + off, len, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) err = 1;
    }
  };
  std::vector<std::thread> pool;
  for (size_t t = 1; t < nthreads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  if (err) { set_error("host -> device staging copy failed: %s", cudaGetErrorString(cudaGetLastError())); return LXY_ERR_CUDA; }
  return LXY_OK;
}

struct Guard {  // context mutex + device selection for one entry point
  std::lock_guard<std::mutex> lk;
  bool ok;
  explicit Guard(lxy_ctx* c) : lk(c->mu), ok(cudaSetDevice(c->cfg.device) == cudaSuccess) {
    if (!ok) set_error("cudaSetDevice(%d) failed", c->cfg.device);
  }
};

static float ev_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) { cudaGetLastError(); return 0.f; }
  return ms;
}

}  // namespace lxy

namespace lxy {
static int flags_to_status(int flag, const char* what, const char* where) {
  if (flag & 1) { set_error("%s: gene value outside [0, n)%s", what, where); return LXY_ERR_BAD_GENE; }
  if (flag & 2) { set_error("%s: a population row is not a permutation%s", what, where); return LXY_ERR_BAD_GENE; }
  if (flag & 4) { set_error("%s: placement offsets overflow the device format%s", what, where); return LXY_ERR_RANGE; }
  return LXY_OK;
}
constexpr int kDeferredSlots = 64;
int report_deferred_flags(lxy_ctx* c) {   // the stream has been synchronised
  int all = 0;
  for (int i = 0; i < c->n_deferred; ++i) all |= c->h_deferred[i];
  const int n = c->n_deferred;
  c->n_deferred = 0;
  if (!all) return LXY_OK;
  char where[96];
  snprintf(where, sizeof where, " (deferred: in one of the %d scoring calls since the last sync)", n);
  return flags_to_status(all, "scoring", where);
}
int read_flags(lxy_ctx* c, const char* what, const char* where) {
  if (c->cfg.flags & LXY_FLAG_DEFER_ERRORS) {
    if (!c->h_deferred) LXY_CUDA(cudaMallocHost((void**)&c->h_deferred, sizeof(int) * kDeferredSlots));
    if (c->n_deferred == kDeferredSlots) {   // every slot holds an unreported call: report them first
      LXY_CUDA(cudaStreamSynchronize(c->stream));
      LXY_TRY(report_deferred_flags(c));
    }
    LXY_CUDA(cudaMemcpyAsync(&c->h_deferred[c->n_deferred++], c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    return LXY_OK;
  }
  int flag = 0;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return flags_to_status(flag, what, where);
}
}  // namespace lxy

namespace lxy {
__global__ void k_widen_u16(const uint16_t* __restrict__ in, int64_t count, int32_t* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < count; i += stride) out[i] = (int32_t)in[i];
}
}  // namespace lxy

// The scorer of the context's R-profile store: the dense pair table when the store has one
// (LXY_PATH_PAIRS, or LXY_PATH_AUTO's choice), else the record stream.
namespace lxy {
int scaff_dispatch(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev, int64_t* x100_dev) {
  if (c->scaff.ready && c->scaff.path == LXY_PATH_PAIRS) return pairs_score(c, perm_dev, P, out_dev, x100_dev);
  return scaff_score(c, perm_dev, nullptr, P, out_dev, x100_dev);
}
}  // namespace lxy

using namespace lxy;

#define CHECK_CTX(c)                                                    \
  if (!(c)) { set_error("null context"); return LXY_ERR_ARG; }          \
  Guard guard_(c);                                                      \
  if (!guard_.ok) return LXY_ERR_CUDA;

static void collect_timing(lxy_ctx* c, bool host_copies) {
  c->timing.ms_layout = ev_ms(c->ev[0], c->ev[1]);
  c->timing.ms_score = ev_ms(c->ev[1], c->ev[2]);
  c->timing.ms_finish = ev_ms(c->ev[2], c->ev[3]);
  c->timing.ms_total = ev_ms(c->ev[0], c->ev[3]);
  c->timing.ms_h2d = host_copies ? ev_ms(c->ev[4], c->ev[0]) : 0.f;
  c->timing.ms_d2h = host_copies ? ev_ms(c->ev[3], c->ev[5]) : 0.f;
}

extern "C" {

int lxy_abi_version(void) { return LXY_ABI_VERSION; }
const char* lxy_last_error(void) { return g_err.c_str(); }

int lxy_device_count(int32_t* count) {
  if (!count) { set_error("null count"); return LXY_ERR_ARG; }
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    *count = 0;
    set_error("cudaGetDeviceCount failed: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return LXY_ERR_CUDA;
  }
  *count = n;
  return LXY_OK;
}

int lxy_ctx_create(const lxy_cfg* cfg, lxy_ctx** out) {
  if (!out) { set_error("null out"); return LXY_ERR_ARG; }
  *out = nullptr;
  lxy_cfg cf{};
  if (cfg) cf = *cfg;
  if (cf.acc_mode < LXY_ACC_AUTO || cf.acc_mode > LXY_ACC_INT) { set_error("bad acc_mode"); return LXY_ERR_ARG; }
  if (cf.path != LXY_PATH_AUTO && cf.path != LXY_PATH_STREAM && cf.path != LXY_PATH_PAIRS) { set_error("bad path"); return LXY_ERR_ARG; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) {
    set_error("no CUDA device (%s); liblxyb200 has no CPU fallback",
              e != cudaSuccess ? cudaGetErrorString(e) : "count 0");
    cudaGetLastError();
    return LXY_ERR_CUDA;
  }
  if (cf.device < 0 || cf.device >= ndev) { set_error("device %d out of range [0,%d)", cf.device, ndev); return LXY_ERR_ARG; }
  LXY_CUDA(cudaSetDevice(cf.device));
  cudaDeviceProp prop;
  LXY_CUDA(cudaGetDeviceProperties(&prop, cf.device));
  if (prop.major < 10) {
    set_error("device %d is sm_%d%d; liblxyb200 is built for sm_100a only", cf.device, prop.major, prop.minor);
    return LXY_ERR_CUDA;
  }
  lxy_ctx* c = new (std::nothrow) lxy_ctx();
  if (!c) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  c->cfg = cf;
  c->sm_count = prop.multiProcessorCount;
  c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    set_error("cudaStreamCreate failed");
    delete c;
    return LXY_ERR_CUDA;
  }
  if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess) {
    set_error("cudaStreamCreate failed");
    lxy_ctx_destroy(c);
    return LXY_ERR_CUDA;
  }
  for (auto& ev : c->ev) {
    if (cudaEventCreate(&ev) != cudaSuccess) { set_error("cudaEventCreate failed"); lxy_ctx_destroy(c); return LXY_ERR_CUDA; }
  }
  *out = c;
  return LXY_OK;
}

void lxy_ctx_destroy(lxy_ctx* c) {
  if (!c) return;
  if (c->comm) lxy_comm_destroy(c);
  cudaSetDevice(c->cfg.device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->stream2) cudaStreamSynchronize(c->stream2);
  DevBuf* bufs[] = {&c->scaff.rec, &c->scaff.w, &c->scaff.tile_start, &c->scaff.len, &c->scaff.xa, &c->scaff.xb, &c->scaff.dense,
                    &c->phase.rec, &c->phase.w, &c->d_perm, &c->d_ori, &c->d_out, &c->d_x100,
                    &c->d_posA, &c->d_posB, &c->d_partial, &c->d_corr, &c->d_flag, &c->d_genes,
                    &c->d_gbits, &c->d_tmp0, &c->d_tmp1, &c->d_tmp2, &c->d_tmp3, &c->d_start, &c->d_shard, &c->d_gather};
  for (DevBuf* b : bufs) b->release();
  if (c->h_deferred) cudaFreeHost(c->h_deferred);
  if (c->h_pin) cudaFreeHost(c->h_pin);
  for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
  lxy::drop_score_graphs(c);
  for (auto& e : c->ev_slice) if (e) cudaEventDestroy(e);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->stream2) cudaStreamDestroy(c->stream2);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int lxy_ctx_get_timing(lxy_ctx* c, lxy_timing* out) {
  if (!c || !out) { set_error("null argument"); return LXY_ERR_ARG; }
  CHECK_CTX(c);
  if (c->timing_pending) {
    LXY_CUDA(cudaEventSynchronize(c->ev[c->timing_pending == 2 ? 5 : 3]));
    collect_timing(c, c->timing_pending == 2);
    c->timing_pending = 0;
  }
  *out = c->timing;
  return LXY_OK;
}

int lxy_ctx_stream(lxy_ctx* c, void** s) {
  if (!c || !s) { set_error("null argument"); return LXY_ERR_ARG; }
  *s = (void*)c->stream;
  return LXY_OK;
}

int lxy_ctx_sync(lxy_ctx* c) {
  CHECK_CTX(c);
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return report_deferred_flags(c);   // LXY_FLAG_DEFER_ERRORS: the validation results of the calls since the last sync
}

int lxy_ctx_path(lxy_ctx* c, int32_t* path, int64_t* stream_records) {
  CHECK_CTX(c);
  if (!c->scaff.ready) { set_error("ctx_path: links not set"); return LXY_ERR_STATE; }
  if (path) *path = c->scaff.path;
  if (stream_records) *stream_records = c->scaff.L;
  return LXY_OK;
}

// ---- scaffolding --------------------------------------------------------------------------------
int lxy_ctx_set_links(lxy_ctx* c, int32_t n_ids, int64_t n_links, const uint32_t* a,
                      const uint32_t* b, const double* w) {
  CHECK_CTX(c);
  LinkInput in;
  in.n = n_ids; in.L = n_links; in.a = a; in.b = b; in.w = w;
  return build_scaff_links(c, in);
}

int lxy_ctx_set_links_dev(lxy_ctx* c, int32_t n_ids, int64_t n_links, const uint32_t* a_dev,
                          const uint32_t* b_dev, const double* w_dev) {
  CHECK_CTX(c);
  LinkInput in;
  in.n = n_ids; in.L = n_links; in.a = a_dev; in.b = b_dev; in.w = w_dev; in.on_device = true;
  return build_scaff_links(c, in);
}


// Host-buffer entry points copy their results back and wait for them: their flags are read now.
static int finish_flags(lxy_ctx* c, const char* what) {
  int flag = 0;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return lxy::flags_to_status(flag, what, "");
}
// Device-pointer entry points: now, or deferred to lxy_ctx_sync (LXY_FLAG_DEFER_ERRORS).
static int finish_flags_dev(lxy_ctx* c, const char* what) { return lxy::read_flags(c, what, ""); }


int lxy_scaff_score_batch_dev(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                              int64_t* x100_dev) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!perm_dev || !out_dev))) { set_error("score_batch_dev: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  LXY_TRY(scaff_dispatch(c, perm_dev, P, out_dev, x100_dev));
  c->timing_pending = 1;
  // With LXY_FLAG_NO_VALIDATE nothing has to come back to the host: the call is asynchronous on
  // the context's stream (lxy_ctx_sync / lxy_ctx_get_timing wait for it).
  if (c->cfg.flags & LXY_FLAG_NO_VALIDATE) return LXY_OK;
  return finish_flags_dev(c, "scaff score");
}

int lxy_scaff_score_batch(lxy_ctx* c, const int32_t* perm, int32_t P, double* out, int64_t* x100) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!perm || !out))) { set_error("score_batch: bad argument"); return LXY_ERR_ARG; }
  if (!c->scaff.ready) { set_error("score_batch: links not set"); return LXY_ERR_STATE; }
  if (P == 0) return LXY_OK;
  const size_t n = (size_t)c->scaff.n;
  const size_t pbytes = sizeof(int32_t) * n * (size_t)P;
  const size_t obytes = 16 * (size_t)P;
  LXY_TRY(pin_reserve(c, pbytes + obytes));
  LXY_TRY(c->d_perm.reserve(pbytes));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  // pinned staging: the copy below is then a true async DMA (and the caller's memory is not
  // retained — cgo rule)
  LXY_CUDA(cudaEventRecord(c->ev[4], c->stream));
  LXY_TRY(stage_h2d(c, c->d_perm.p, perm, pbytes, 0));
  LXY_TRY(scaff_dispatch(c, c->d_perm.as<int32_t>(), P, c->d_out.as<double>(), c->d_x100.as<int64_t>()));
  char* hout = (char*)c->h_pin + pbytes;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaEventRecord(c->ev[5], c->stream));
  c->timing_pending = 2;
  LXY_TRY(finish_flags(c, "scaff score"));
  std::memcpy(out, hout, 8 * (size_t)P);
  if (x100) std::memcpy(x100, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

// 16-bit genes (n_ids <= 65,536): half the bytes over PCIe; widened to the kernels' int32 on the device.
int lxy_scaff_score_batch_u16(lxy_ctx* c, const uint16_t* perm, int32_t P, double* out, int64_t* x100) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!perm || !out))) { set_error("score_batch_u16: bad argument"); return LXY_ERR_ARG; }
  if (!c->scaff.ready) { set_error("score_batch_u16: links not set"); return LXY_ERR_STATE; }
  if (c->scaff.n > 65536) { set_error("score_batch_u16: %d ids do not fit 16-bit genes", c->scaff.n); return LXY_ERR_RANGE; }
  if (P == 0) return LXY_OK;
  const size_t cnt = (size_t)c->scaff.n * (size_t)P;
  const size_t pbytes = sizeof(uint16_t) * cnt;
  const size_t poff = (pbytes + 15) & ~(size_t)15;
  LXY_TRY(pin_reserve(c, poff + 16 * (size_t)P));
  LXY_TRY(c->d_tmp3.reserve(pbytes));
  LXY_TRY(c->d_perm.reserve(sizeof(int32_t) * cnt));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  LXY_CUDA(cudaEventRecord(c->ev[4], c->stream));
  LXY_TRY(stage_h2d(c, c->d_tmp3.p, perm, pbytes, 0));
  lxy::k_widen_u16<<<(unsigned)std::min<size_t>((cnt + 255) / 256, (size_t)c->sm_count * 16), 256, 0, c->stream>>>(c->d_tmp3.as<uint16_t>(), (int64_t)cnt,
                                                                                     c->d_perm.as<int32_t>());
  LXY_CUDA(cudaGetLastError());
  LXY_TRY(scaff_dispatch(c, c->d_perm.as<int32_t>(), P, c->d_out.as<double>(), c->d_x100.as<int64_t>()));
  char* hout = (char*)c->h_pin + poff;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaEventRecord(c->ev[5], c->stream));
  c->timing_pending = 2;
  LXY_TRY(finish_flags(c, "scaff score"));
  std::memcpy(out, hout, 8 * (size_t)P);
  if (x100) std::memcpy(x100, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

int lxy_scaff_positions(lxy_ctx* c, const int32_t* perm, int32_t P, int32_t* pos) {
  CHECK_CTX(c);
  if (!c->scaff.ready) { set_error("positions: links not set"); return LXY_ERR_STATE; }
  if (P < 0 || (P > 0 && (!perm || !pos))) { set_error("positions: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  const size_t bytes = sizeof(int32_t) * (size_t)c->scaff.n * (size_t)P;
  LXY_TRY(c->d_perm.reserve(bytes));
  LXY_TRY(c->d_tmp0.reserve(bytes));
  LXY_CUDA(cudaMemcpyAsync(c->d_perm.p, perm, bytes, cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaMemsetAsync(c->d_tmp0.p, 0xFF, bytes, c->stream));
  LXY_TRY(scaff_positions(c, c->d_perm.as<int32_t>(), P, c->scaff.n, c->d_tmp0.as<int32_t>()));
  LXY_CUDA(cudaMemcpyAsync(pos, c->d_tmp0.p, bytes, cudaMemcpyDeviceToHost, c->stream));
  return finish_flags(c, "positions");
}

int lxy_scaff_link_distances(lxy_ctx* c, int32_t n, int64_t L, const uint32_t* a, const uint32_t* b,
                             const int32_t* perm, int64_t* dist) {
  CHECK_CTX(c);
  if (n <= 0 || L < 0 || !perm || (L > 0 && (!a || !b || !dist))) { set_error("link_distances: bad argument"); return LXY_ERR_ARG; }
  LXY_TRY(c->d_perm.reserve(sizeof(int32_t) * (size_t)n));
  LXY_TRY(c->d_tmp0.reserve(4 * (size_t)L + 16));
  LXY_TRY(c->d_tmp1.reserve(4 * (size_t)L + 16));
  LXY_TRY(c->d_tmp2.reserve(8 * (size_t)L + 16));
  LXY_CUDA(cudaMemcpyAsync(c->d_perm.p, perm, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  if (L > 0) {
    LXY_CUDA(cudaMemcpyAsync(c->d_tmp0.p, a, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
    LXY_CUDA(cudaMemcpyAsync(c->d_tmp1.p, b, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
  }
  LXY_TRY(scaff_link_distances(c, n, L, c->d_tmp0.as<uint32_t>(), c->d_tmp1.as<uint32_t>(),
                               c->d_perm.as<int32_t>(), c->d_tmp2.as<int64_t>()));
  if (L > 0) LXY_CUDA(cudaMemcpyAsync(dist, c->d_tmp2.p, 8 * (size_t)L, cudaMemcpyDeviceToHost, c->stream));
  return finish_flags(c, "link_distances");
}

int lxy_links_get(lxy_ctx* c, int32_t id1, int32_t id2, double* out) {
  CHECK_CTX(c);
  if (!out) { set_error("links_get: null out"); return LXY_ERR_ARG; }
  return links_get(c, id1, id2, out);
}

// ---- G-profile ----------------------------------------------------------------------------------
int lxy_ctx_set_links_g(lxy_ctx* c, int32_t n_ids, int64_t n_links, const uint32_t* a,
                        const uint32_t* xa, const uint32_t* b, const uint32_t* xb, const double* w,
                        const uint32_t* len, uint32_t unit) {
  CHECK_CTX(c);
  if (!len) { set_error("set_links_g: null len"); return LXY_ERR_ARG; }
  LinkInput in;
  in.n = n_ids; in.L = n_links; in.a = a; in.b = b; in.w = w; in.xa = xa; in.xb = xb;
  in.len = len; in.unit = unit;
  return build_scaff_links(c, in);
}

int lxy_ctx_set_links_g_dev(lxy_ctx* c, int32_t n_ids, int64_t n_links, const uint32_t* a_dev,
                            const uint32_t* xa_dev, const uint32_t* b_dev, const uint32_t* xb_dev,
                            const double* w_dev, const uint32_t* len, uint32_t unit) {
  CHECK_CTX(c);
  if (!len) { set_error("set_links_g_dev: null len"); return LXY_ERR_ARG; }
  LinkInput in;
  in.n = n_ids; in.L = n_links; in.a = a_dev; in.b = b_dev; in.w = w_dev; in.xa = xa_dev; in.xb = xb_dev;
  in.len = len; in.unit = unit; in.on_device = true;
  return build_scaff_links(c, in);
}

int lxy_scaffg_score_batch(lxy_ctx* c, const int32_t* perm, const uint8_t* ori, int32_t P,
                           double* out, int64_t* x100) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!perm || !out))) { set_error("scaffg_score_batch: bad argument"); return LXY_ERR_ARG; }
  if (!c->scaff.ready || !c->scaff.gprofile) { set_error("scaffg_score_batch: G-profile links not set"); return LXY_ERR_STATE; }
  if (P == 0) return LXY_OK;
  const size_t n = (size_t)c->scaff.n;
  const size_t pbytes = sizeof(int32_t) * n * (size_t)P, obytes = ori ? n * (size_t)P : 0;
  const size_t ooff = (pbytes + obytes + 15) & ~(size_t)15;
  LXY_TRY(pin_reserve(c, ooff + 16 * (size_t)P));
  LXY_TRY(c->d_perm.reserve(pbytes));
  LXY_TRY(c->d_ori.reserve(obytes + 16));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  LXY_CUDA(cudaEventRecord(c->ev[4], c->stream));
  LXY_TRY(stage_h2d(c, c->d_perm.p, perm, pbytes, 0));
  if (ori) LXY_TRY(stage_h2d(c, c->d_ori.p, ori, obytes, pbytes));
  LXY_TRY(scaffg_score(c, c->d_perm.as<int32_t>(), ori ? c->d_ori.as<uint8_t>() : nullptr, P,
                       c->d_out.as<double>(), c->d_x100.as<int64_t>()));
  char* hout = (char*)c->h_pin + ooff;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaEventRecord(c->ev[5], c->stream));
  c->timing_pending = 2;
  LXY_TRY(finish_flags(c, "scaffg score"));
  std::memcpy(out, hout, 8 * (size_t)P);
  if (x100) std::memcpy(x100, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

int lxy_scaffg_score_batch_dev(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P,
                               double* out_dev, int64_t* x100_dev) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!perm_dev || !out_dev))) { set_error("scaffg_score_batch_dev: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  LXY_TRY(scaffg_score(c, perm_dev, ori_dev, P, out_dev, x100_dev));
  c->timing_pending = 1;
  if (c->cfg.flags & LXY_FLAG_NO_VALIDATE) return LXY_OK;
  return finish_flags_dev(c, "scaffg score");
}

int lxy_scaffg_layout(lxy_ctx* c, const int32_t* perm, int32_t P, uint64_t* start_by_pos,
                      int32_t* pos_by_contig) {
  CHECK_CTX(c);
  if (!c->scaff.ready || !c->scaff.gprofile) { set_error("scaffg_layout: G-profile links not set"); return LXY_ERR_STATE; }
  if (P < 0 || (P > 0 && (!perm || !start_by_pos || !pos_by_contig))) { set_error("scaffg_layout: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  const size_t cnt = (size_t)c->scaff.n * (size_t)P;
  LXY_TRY(c->d_perm.reserve(4 * cnt));
  LXY_TRY(c->d_tmp0.reserve(8 * cnt));
  LXY_TRY(c->d_tmp1.reserve(4 * cnt));
  LXY_CUDA(cudaMemcpyAsync(c->d_perm.p, perm, 4 * cnt, cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaMemsetAsync(c->d_tmp1.p, 0xFF, 4 * cnt, c->stream));
  LXY_TRY(scaffg_layout64(c, c->d_perm.as<int32_t>(), P, c->d_tmp0.as<uint64_t>(), c->d_tmp1.as<int32_t>()));
  LXY_CUDA(cudaMemcpyAsync(start_by_pos, c->d_tmp0.p, 8 * cnt, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(pos_by_contig, c->d_tmp1.p, 4 * cnt, cudaMemcpyDeviceToHost, c->stream));
  return finish_flags(c, "scaffg_layout");
}

int lxy_scaffg_link_distances(lxy_ctx* c, int64_t L, const uint32_t* a, const uint32_t* xa,
                              const uint32_t* b, const uint32_t* xb, const int32_t* perm,
                              const uint8_t* ori, uint64_t* dist) {
  CHECK_CTX(c);
  if (!c->scaff.ready || !c->scaff.gprofile) { set_error("scaffg_link_distances: G-profile links not set"); return LXY_ERR_STATE; }
  if (L < 0 || !perm || (L > 0 && (!a || !xa || !b || !xb || !dist))) { set_error("scaffg_link_distances: bad argument"); return LXY_ERR_ARG; }
  const size_t n = (size_t)c->scaff.n;
  LXY_TRY(c->d_perm.reserve(4 * n));
  LXY_TRY(c->d_ori.reserve(n + 16));
  LXY_TRY(c->d_posA.reserve(8 * n));
  LXY_TRY(c->d_posB.reserve(4 * n));
  LXY_TRY(c->d_tmp0.reserve(16 * (size_t)L + 64));
  LXY_TRY(c->d_tmp2.reserve(8 * (size_t)L + 64));
  LXY_CUDA(cudaMemcpyAsync(c->d_perm.p, perm, 4 * n, cudaMemcpyHostToDevice, c->stream));
  if (ori) LXY_CUDA(cudaMemcpyAsync(c->d_ori.p, ori, n, cudaMemcpyHostToDevice, c->stream));
  LXY_TRY(scaffg_layout64(c, c->d_perm.as<int32_t>(), 1, c->d_posA.as<uint64_t>(), c->d_posB.as<int32_t>()));
  uint32_t* da = c->d_tmp0.as<uint32_t>();
  if (L > 0) {
    LXY_CUDA(cudaMemcpyAsync(da, a, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
    LXY_CUDA(cudaMemcpyAsync(da + L, xa, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
    LXY_CUDA(cudaMemcpyAsync(da + 2 * L, b, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
    LXY_CUDA(cudaMemcpyAsync(da + 3 * L, xb, 4 * (size_t)L, cudaMemcpyHostToDevice, c->stream));
  }
  LXY_TRY(scaffg_link_distances(c, L, da, da + L, da + 2 * L, da + 3 * L, c->d_posA.as<uint64_t>(),
                                c->d_posB.as<int32_t>(), ori ? c->d_ori.as<uint8_t>() : nullptr,
                                c->d_tmp2.as<uint64_t>()));
  if (L > 0) LXY_CUDA(cudaMemcpyAsync(dist, c->d_tmp2.p, 8 * (size_t)L, cudaMemcpyDeviceToHost, c->stream));
  return finish_flags(c, "scaffg_link_distances");
}

// ---- phasing ------------------------------------------------------------------------------------
int lxy_ctx_set_phase_links(lxy_ctx* c, int32_t n_ids, int64_t n_links, const uint32_t* a,
                            const uint32_t* b, const double* w) {
  CHECK_CTX(c);
  return build_phase_links(c, n_ids, n_links, a, b, w);
}

int lxy_phase_score_batch_dev(lxy_ctx* c, const uint8_t* genes_dev, int32_t P, int64_t* out_i64_dev,
                              double* out_f64_dev) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!genes_dev || (!out_i64_dev && !out_f64_dev)))) { set_error("phase_score_batch_dev: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  LXY_TRY(phase_score(c, genes_dev, P, out_i64_dev, out_f64_dev));
  c->timing_pending = 1;
  if (c->cfg.flags & LXY_FLAG_NO_VALIDATE) return LXY_OK;
  return finish_flags_dev(c, "phase score");
}

int lxy_phase_score_batch(lxy_ctx* c, const uint8_t* genes, int32_t P, int64_t* out_i64,
                          double* out_f64) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!genes || (!out_i64 && !out_f64)))) { set_error("phase_score_batch: bad argument"); return LXY_ERR_ARG; }
  if (!c->phase.ready) { set_error("phase_score_batch: links not set"); return LXY_ERR_STATE; }
  if (P == 0) return LXY_OK;
  const size_t gbytes = (size_t)c->phase.n * (size_t)P;
  LXY_TRY(pin_reserve(c, gbytes + 16 * (size_t)P));
  LXY_TRY(c->d_genes.reserve(gbytes));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  LXY_CUDA(cudaEventRecord(c->ev[4], c->stream));
  LXY_TRY(stage_h2d(c, c->d_genes.p, genes, gbytes, 0));
  LXY_TRY(phase_score(c, c->d_genes.as<uint8_t>(), P, c->d_x100.as<int64_t>(), c->d_out.as<double>()));
  char* hout = (char*)c->h_pin + gbytes;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaEventRecord(c->ev[5], c->stream));
  c->timing_pending = 2;
  LXY_TRY(finish_flags(c, "phase score"));
  if (out_f64) std::memcpy(out_f64, hout, 8 * (size_t)P);
  if (out_i64) std::memcpy(out_i64, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

// Row-packed populations: 1/8 of the bytes of the byte-per-gene entry point over PCIe.
int lxy_phase_score_batch_bits_dev(lxy_ctx* c, const uint32_t* words_dev, int32_t P, int64_t* out_i64_dev,
                                   double* out_f64_dev) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!words_dev || (!out_i64_dev && !out_f64_dev)))) { set_error("phase_score_batch_bits_dev: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  LXY_TRY(phase_score_any(c, nullptr, words_dev, P, out_i64_dev, out_f64_dev));
  c->timing_pending = 1;
  return LXY_OK;   // a packed bit is 0 or 1 by construction: nothing to validate
}

int lxy_phase_score_batch_bits(lxy_ctx* c, const uint32_t* words, int32_t P, int64_t* out_i64,
                               double* out_f64) {
  CHECK_CTX(c);
  if (P < 0 || (P > 0 && (!words || (!out_i64 && !out_f64)))) { set_error("phase_score_batch_bits: bad argument"); return LXY_ERR_ARG; }
  if (!c->phase.ready) { set_error("phase_score_batch_bits: links not set"); return LXY_ERR_STATE; }
  if (P == 0) return LXY_OK;
  const size_t W = ((size_t)c->phase.n + 31) / 32;
  const size_t gbytes = 4 * W * (size_t)P;
  LXY_TRY(pin_reserve(c, gbytes + 16 * (size_t)P));
  LXY_TRY(c->d_genes.reserve(gbytes));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  LXY_CUDA(cudaEventRecord(c->ev[4], c->stream));
  // Page-locked caller memory and a population of several candidate blocks: the rows go up in slices on
  // the side stream while the slices already there are scored (candidates are independent; a slice is a
  // multiple of the 128 candidates one warp scores), so the copy engine and the SMs overlap: 38 MB of
  // genes at 300,000 blocks x 1,024 genomes cost 1.5 ms over PCIe against 1.0 ms of scoring.
  cudaPointerAttributes attr;
  const bool pinned = cudaPointerGetAttributes(&attr, words) == cudaSuccess && attr.type == cudaMemoryTypeHost;
  if (!pinned) (void)cudaGetLastError();
  if (pinned && P >= 512) {
    const int32_t nsl = P / 256 < 8 ? P / 256 : 8;
    const int32_t rows = ((P + nsl - 1) / nsl + 127) / 128 * 128;
    for (auto& e : c->ev_slice)
      if (!e) LXY_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    LXY_CUDA(cudaEventRecord(c->ev_fork, c->stream));          // the device buffer may still be read by the previous call
    LXY_CUDA(cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
    int32_t k = 0;
    for (int32_t p0 = 0; p0 < P; p0 += rows, ++k) {
      const int32_t cnt = P - p0 < rows ? P - p0 : rows;
      LXY_CUDA(cudaMemcpyAsync(c->d_genes.as<uint32_t>() + (size_t)p0 * W, words + (size_t)p0 * W, 4 * W * (size_t)cnt,
                               cudaMemcpyHostToDevice, c->stream2));
      LXY_CUDA(cudaEventRecord(c->ev_slice[k], c->stream2));
    }
    k = 0;
    for (int32_t p0 = 0; p0 < P; p0 += rows, ++k) {
      const int32_t cnt = P - p0 < rows ? P - p0 : rows;
      LXY_CUDA(cudaStreamWaitEvent(c->stream, c->ev_slice[k], 0));
      LXY_TRY(phase_score_any(c, nullptr, c->d_genes.as<uint32_t>() + (size_t)p0 * W, cnt, c->d_x100.as<int64_t>() + p0,
                              c->d_out.as<double>() + p0));
    }
  } else {
    LXY_TRY(stage_h2d(c, c->d_genes.p, words, gbytes, 0));
    LXY_TRY(phase_score_any(c, nullptr, c->d_genes.as<uint32_t>(), P, c->d_x100.as<int64_t>(), c->d_out.as<double>()));
  }
  char* hout = (char*)c->h_pin + gbytes;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaEventRecord(c->ev[5], c->stream));
  c->timing_pending = 2;
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  if (out_f64) std::memcpy(out_f64, hout, 8 * (size_t)P);
  if (out_i64) std::memcpy(out_i64, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

// ---- rank ---------------------------------------------------------------------------------------
int lxy_rank(const double* s, int32_t P, int32_t* order) {
  if (P < 0 || (P > 0 && (!s || !order))) { set_error("rank: bad argument"); return LXY_ERR_ARG; }
  std::vector<int32_t> idx((size_t)P);
  std::iota(idx.begin(), idx.end(), 0);
  std::stable_sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) { return s[x] < s[y]; });
  if (P) std::memcpy(order, idx.data(), sizeof(int32_t) * (size_t)P);
  return LXY_OK;
}

}  // extern "C"

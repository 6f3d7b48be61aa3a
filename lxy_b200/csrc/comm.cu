// comm.cu — multi-GPU scoring behind the C ABI: one lxy_ctx per device, an NCCL communicator per
// context, the collectives issued on the context's own stream.
//
// north_star: the population is sharded across GPUs for large populations (links replicated,
// ncclAllGather of the per-rank score slices), the links are sharded for huge link sets (every rank
// scores the whole population against its share of the records, ncclAllReduce of the exact int64
// partial sums — the sum over records of DESIGN.md §2 splits over any partition of the records), and
// only the per-generation score vector moves between GPUs.  The reference's host is one Go process
// (/root/reference/src/lxy/scaff/scaff.go:88-94, go-galib's GAParallel with 7 goroutines), so the
// communicator can be created for all devices of one process (lxy_comm_init_all) as well as per rank
// (lxy_comm_unique_id + lxy_comm_init_rank: torchrun / MPI style).
//
// NCCL is bound at run time (dlopen of libnccl.so.2; the copy torch already mapped is reused when the
// caller is a torch process): a single-GPU user of liblxyb200 never needs the library.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace lxy {
namespace {

struct Nccl {
  void* h = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

Nccl* nccl() {
  static Nccl lib;
  static std::once_flag once;
  std::call_once(once, [] {
    // One NCCL per process: the copy that is already mapped (a torch process), else the one the host
    // names in LXY_NCCL_LIB (the Python binding points it at torch's bundled copy, so that a later
    // `import torch` finds its own libnccl.so.2 under that soname and not an older system one), else
    // the system library.
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);
    if (!h && getenv("LXY_NCCL_LIB")) h = dlopen(getenv("LXY_NCCL_LIB"), RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    Nccl l;
    l.h = h;
#define LXY_SYM(name) l.name = (decltype(l.name))dlsym(h, "nccl" #name)
    LXY_SYM(GetUniqueId); LXY_SYM(CommInitRank); LXY_SYM(CommDestroy); LXY_SYM(AllGather);
    LXY_SYM(AllReduce); LXY_SYM(GroupStart); LXY_SYM(GroupEnd); LXY_SYM(GetErrorString);
#undef LXY_SYM
    if (l.GetUniqueId && l.CommInitRank && l.CommDestroy && l.AllGather && l.AllReduce && l.GroupStart &&
        l.GroupEnd && l.GetErrorString)
      lib = l;
  });
  return lib.h ? &lib : nullptr;
}

#define LXY_NCCL(expr)                                                                        \
  do {                                                                                        \
    ncclResult_t r_ = (expr);                                                                 \
    if (r_ != ncclSuccess) {                                                                  \
      ::lxy::set_error("%s failed: %s", #expr, nccl()->GetErrorString(r_));                   \
      return LXY_ERR_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

int need_nccl() {
  if (nccl()) return LXY_OK;
  set_error("libnccl.so.2 could not be loaded (%s)", dlerror() ? dlerror() : "symbols missing");
  return LXY_ERR_STATE;
}

struct Slices {  // first[r] of the balanced contiguous split, r in [0, world]
  int32_t first[65];
  int32_t world;
  int32_t qmax;
};

Slices make_slices(int32_t total, int32_t world) {
  Slices s{};
  s.world = world;
  const int32_t q = total / world, rem = total % world;
  for (int r = 0; r <= world; ++r) s.first[r] = r * q + std::min(r, rem);
  s.qmax = q + (rem ? 1 : 0);
  return s;
}

// gathered[r] = [qmax doubles | qmax int64 | flag word] of rank r  ->  out[P], x100[P], flags OR-ed
__global__ void k_unpack_slices(Slices sl, const uint8_t* __restrict__ gathered, size_t stride,
                                int32_t P, double* __restrict__ out, int64_t* __restrict__ x100,
                                int* __restrict__ flag) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    int f = 0;
    for (int r = 0; r < sl.world; ++r) f |= *(const int*)(gathered + (size_t)r * stride + 16 * (size_t)sl.qmax);
    *flag = f;
  }
  if (p >= P) return;
  int r = 0;
  while (p >= sl.first[r + 1]) ++r;
  const uint8_t* base = gathered + (size_t)r * stride;
  const int i = p - sl.first[r];
  out[p] = ((const double*)base)[i];
  if (x100) x100[p] = ((const int64_t*)(base + 8 * (size_t)sl.qmax))[i];
}

// The flag word travels with the scores so that every rank returns the same status.
__global__ void k_pack_flag(const int* __restrict__ flag, int* __restrict__ dst) { *dst = *flag; }

// link-sharded, integer mode: buf[0..P) partial x100 of this rank, buf[P] the flag bits spread over
// three 20-bit counters (the all-reduce adds them: any rank's bit survives as a non-zero counter)
__global__ void k_links_pack(int32_t P, const int* __restrict__ flag, int64_t* __restrict__ buf) {
  const int f = *flag;
  buf[P] = (int64_t)(f & 1) | ((int64_t)((f >> 1) & 1) << 20) | ((int64_t)((f >> 2) & 1) << 40);
}
__global__ void k_links_finish_int(int32_t P, const int64_t* __restrict__ buf, double* __restrict__ out,
                                   int64_t* __restrict__ x100, int* __restrict__ flag) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    const int64_t v = buf[P];
    *flag = ((v & 0xFFFFF) ? 1 : 0) | (((v >> 20) & 0xFFFFF) ? 2 : 0) | (((v >> 40) & 0xFFFFF) ? 4 : 0);
  }
  if (p >= P) return;
  const int64_t x = buf[p];
  out[p] = -((double)x / 100.0);
  if (x100) x100[p] = x;
}
// link-sharded, fp64 mode: gathered[r] = [P doubles (-partial/100) | flag]; fixed rank order
__global__ void k_links_finish_f64(int32_t P, int32_t world, const double* __restrict__ gathered,
                                   double* __restrict__ out, int64_t* __restrict__ x100, int* __restrict__ flag) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    int f = 0;
    for (int r = 0; r < world; ++r) f |= *(const int*)(gathered + (size_t)r * (P + 1) + P);
    *flag = f;
  }
  if (p >= P) return;
  double s = 0.0;
  for (int r = 0; r < world; ++r) s += gathered[(size_t)r * (P + 1) + p];
  out[p] = s;
  if (x100) x100[p] = INT64_MIN;
}

int check_flags(lxy_ctx* c, const char* what) { return read_flags(c, what, " (on some rank)"); }
int check_flags_now(lxy_ctx* c, const char* what) {   // host-buffer entry points: results and errors are waited for
  int flag = 0;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  if (flag & 1) { set_error("%s: gene value outside [0, n) (on some rank)", what); return LXY_ERR_BAD_GENE; }
  if (flag & 2) { set_error("%s: a population row is not a permutation (on some rank)", what); return LXY_ERR_BAD_GENE; }
  if (flag & 4) { set_error("%s: placement offsets overflow the device format", what); return LXY_ERR_RANGE; }
  return LXY_OK;
}

// Do all shards accumulate in integers?  Agreed once per link store (a host-synchronous all-reduce).
int agree_mode(lxy_ctx* c) {
  if (c->agree_version == c->scaff.version) return LXY_OK;
  LXY_TRY(c->d_gather.reserve(64));
  int32_t mine = (c->scaff.wmode == W_F64) ? 1 : 0, any = 0;
  LXY_CUDA(cudaMemcpyAsync(c->d_gather.p, &mine, 4, cudaMemcpyHostToDevice, c->stream));
  if (c->world > 1 || c->comm)
    LXY_NCCL(nccl()->AllReduce(c->d_gather.p, c->d_gather.p, 1, ncclInt32, ncclMax, (ncclComm_t)c->comm, c->stream));
  LXY_CUDA(cudaMemcpyAsync(&any, c->d_gather.p, 4, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  c->agree_f64 = any;
  c->agree_version = c->scaff.version;
  return LXY_OK;
}

struct Guard {
  std::lock_guard<std::mutex> lk;
  bool ok;
  explicit Guard(lxy_ctx* c) : lk(c->mu), ok(cudaSetDevice(c->cfg.device) == cudaSuccess) {
    if (!ok) set_error("cudaSetDevice(%d) failed", c->cfg.device);
  }
};

}  // namespace

// The sharded scoring pass on the context's stream.  perm_dev: the whole population, resident.
// check_flags: read the combined error flags back (a host sync) and turn them into a status.
int sharded_score_dev(lxy_ctx* c, int32_t mode, const int32_t* perm_dev, int32_t P, double* out_dev,
                      int64_t* x100_dev, bool want_check) {
  ScaffLinks& S = c->scaff;
  if (!S.ready || S.gprofile) { set_error("sharded score: R-profile links not set"); return LXY_ERR_STATE; }
  if (mode != LXY_SHARD_POPULATION && mode != LXY_SHARD_LINKS) { set_error("sharded score: bad mode"); return LXY_ERR_ARG; }
  if (P <= 0) return LXY_OK;
  cudaStream_t st = c->stream;
  const int32_t world = c->world, rank = c->rank;
  if (world > 1 && !c->comm) { set_error("sharded score: no communicator"); return LXY_ERR_STATE; }
  const size_t n = (size_t)S.n;
  LXY_TRY(c->d_flag.reserve(64));
  if (mode == LXY_SHARD_POPULATION) {
    const Slices sl = make_slices(P, world);
    const int32_t first = sl.first[rank], cnt = sl.first[rank + 1] - first;
    const size_t stride = 16 * (size_t)sl.qmax + 16;   // doubles | int64 | flag word (+ pad)
    LXY_TRY(c->d_shard.reserve(stride));
    LXY_TRY(c->d_gather.reserve(stride * (size_t)world));
    uint8_t* send = c->d_shard.as<uint8_t>();
    if (cnt > 0) {
      LXY_TRY(scaff_dispatch(c, perm_dev + (size_t)first * n, cnt, (double*)send, (int64_t*)(send + 8 * (size_t)sl.qmax)));
    } else {
      LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
    }
    k_pack_flag<<<1, 1, 0, st>>>(c->d_flag.as<int>(), (int*)(send + 16 * (size_t)sl.qmax));
    LXY_CUDA(cudaGetLastError());
    const uint8_t* gathered = send;
    if (c->comm) {
      LXY_NCCL(nccl()->AllGather(send, c->d_gather.p, stride, ncclInt8, (ncclComm_t)c->comm, st));
      gathered = c->d_gather.as<uint8_t>();
    }
    k_unpack_slices<<<(P + 255) / 256, 256, 0, st>>>(sl, gathered, stride, P, out_dev, x100_dev, c->d_flag.as<int>());
    LXY_CUDA(cudaGetLastError());
  } else {
    LXY_TRY(agree_mode(c));
    if (!c->agree_f64) {
      LXY_TRY(c->d_shard.reserve(8 * (size_t)(P + 2)));
      LXY_TRY(c->d_gather.reserve(8 * (size_t)(P + 2)));
      int64_t* buf = c->d_shard.as<int64_t>();
      LXY_TRY(scaff_dispatch(c, perm_dev, P, c->d_gather.as<double>(), buf));   // fitness of the shard alone: unused
      k_links_pack<<<1, 1, 0, st>>>(P, c->d_flag.as<int>(), buf);
      LXY_CUDA(cudaGetLastError());
      if (c->comm) LXY_NCCL(nccl()->AllReduce(buf, buf, (size_t)P + 1, ncclInt64, ncclSum, (ncclComm_t)c->comm, st));
      k_links_finish_int<<<(P + 255) / 256, 256, 0, st>>>(P, buf, out_dev, x100_dev, c->d_flag.as<int>());
      LXY_CUDA(cudaGetLastError());
    } else {
      // some shard has real weights: every rank accumulates in fp64 (its integer partial, when it has
      // one, is exact in fp64 up to 2^53) and the partial fitness values are added in rank order
      const size_t row = 8 * (size_t)(P + 1);
      LXY_TRY(c->d_shard.reserve(row + 8 * (size_t)P));
      LXY_TRY(c->d_gather.reserve(row * (size_t)world));
      double* send = c->d_shard.as<double>();
      LXY_TRY(scaff_dispatch(c, perm_dev, P, send, (int64_t*)(send + P + 1)));
      k_pack_flag<<<1, 1, 0, st>>>(c->d_flag.as<int>(), (int*)(send + P));
      LXY_CUDA(cudaGetLastError());
      const double* gathered = send;
      if (c->comm) {
        LXY_NCCL(nccl()->AllGather(send, c->d_gather.p, row, ncclInt8, (ncclComm_t)c->comm, st));
        gathered = c->d_gather.as<double>();
      }
      k_links_finish_f64<<<(P + 255) / 256, 256, 0, st>>>(P, c->comm ? world : 1, gathered, out_dev, x100_dev, c->d_flag.as<int>());
      LXY_CUDA(cudaGetLastError());
    }
  }
  c->timing.launches += 2;
  c->timing_pending = 1;
  if (!want_check || (c->cfg.flags & LXY_FLAG_NO_VALIDATE)) return LXY_OK;
  return check_flags(c, "sharded scaff score");
}

}  // namespace lxy

using namespace lxy;

extern "C" {

int lxy_comm_unique_id(void* id128) {
  if (!id128) { set_error("comm_unique_id: null id"); return LXY_ERR_ARG; }
  LXY_TRY(need_nccl());
  static_assert(sizeof(ncclUniqueId) == LXY_COMM_ID_BYTES, "NCCL unique id size");
  ncclUniqueId id;
  LXY_NCCL(nccl()->GetUniqueId(&id));
  std::memcpy(id128, &id, sizeof(id));
  return LXY_OK;
}

int lxy_comm_init_rank(lxy_ctx* c, int32_t world, int32_t rank, const void* id128) {
  if (!c || !id128) { set_error("comm_init_rank: null argument"); return LXY_ERR_ARG; }
  if (world < 1 || world > 64 || rank < 0 || rank >= world) { set_error("comm_init_rank: need 0 <= rank < world <= 64"); return LXY_ERR_ARG; }
  LXY_TRY(need_nccl());
  Guard g(c);
  if (!g.ok) return LXY_ERR_CUDA;
  if (c->comm) { set_error("comm_init_rank: the context already has a communicator"); return LXY_ERR_STATE; }
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof(id));
  ncclComm_t comm = nullptr;
  LXY_NCCL(nccl()->CommInitRank(&comm, world, id, rank));
  c->comm = comm;
  c->world = world;
  c->rank = rank;
  c->agree_version = 0;
  return LXY_OK;
}

int lxy_comm_init_all(lxy_ctx* const* ctxs, int32_t world) {
  if (!ctxs || world < 1 || world > 64) { set_error("comm_init_all: need 1 <= world <= 64 contexts"); return LXY_ERR_ARG; }
  for (int i = 0; i < world; ++i) {
    if (!ctxs[i]) { set_error("comm_init_all: null context %d", i); return LXY_ERR_ARG; }
    if (ctxs[i]->comm) { set_error("comm_init_all: context %d already has a communicator", i); return LXY_ERR_STATE; }
    for (int j = 0; j < i; ++j)
      if (ctxs[j]->cfg.device == ctxs[i]->cfg.device) { set_error("comm_init_all: contexts %d and %d share device %d", j, i, ctxs[i]->cfg.device); return LXY_ERR_ARG; }
  }
  LXY_TRY(need_nccl());
  ncclUniqueId id;
  LXY_NCCL(nccl()->GetUniqueId(&id));
  std::vector<ncclComm_t> comms((size_t)world, nullptr);
  LXY_NCCL(nccl()->GroupStart());
  for (int i = 0; i < world; ++i) {
    if (cudaSetDevice(ctxs[i]->cfg.device) != cudaSuccess) { nccl()->GroupEnd(); set_error("cudaSetDevice failed"); return LXY_ERR_CUDA; }
    ncclResult_t r = nccl()->CommInitRank(&comms[(size_t)i], world, id, i);
    if (r != ncclSuccess) { nccl()->GroupEnd(); set_error("ncclCommInitRank failed: %s", nccl()->GetErrorString(r)); return LXY_ERR_CUDA; }
  }
  LXY_NCCL(nccl()->GroupEnd());
  for (int i = 0; i < world; ++i) {
    std::lock_guard<std::mutex> lk(ctxs[i]->mu);
    ctxs[i]->comm = comms[(size_t)i];
    ctxs[i]->world = world;
    ctxs[i]->rank = i;
    ctxs[i]->agree_version = 0;
  }
  return LXY_OK;
}

int lxy_comm_destroy(lxy_ctx* c) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  Guard g(c);
  if (!g.ok) return LXY_ERR_CUDA;
  if (c->comm) {
    cudaStreamSynchronize(c->stream);
    if (nccl()) nccl()->CommDestroy((ncclComm_t)c->comm);
    c->comm = nullptr;
  }
  c->world = 1;
  c->rank = 0;
  return LXY_OK;
}

int lxy_comm_info(lxy_ctx* c, int32_t* world, int32_t* rank) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(c->mu);
  if (world) *world = c->world;
  if (rank) *rank = c->rank;
  return LXY_OK;
}

int lxy_comm_slice(int32_t total, int32_t world, int32_t rank, int32_t* first, int32_t* count) {
  if (total < 0 || world < 1 || world > 64 || rank < 0 || rank >= world) { set_error("comm_slice: bad argument"); return LXY_ERR_ARG; }
  const Slices s = make_slices(total, world);
  if (first) *first = s.first[rank];
  if (count) *count = s.first[rank + 1] - s.first[rank];
  return LXY_OK;
}

int lxy_comm_link_owners(int32_t n_ids, int64_t n_links, const uint32_t* a, const uint32_t* b, int32_t world,
                         int32_t* owner) {
  if (n_ids <= 0 || n_links < 0 || world < 1 || world > 64 || (n_links > 0 && (!a || !b || !owner))) {
    set_error("comm_link_owners: bad argument");
    return LXY_ERR_ARG;
  }
  const int64_t nblk = ((int64_t)n_ids + kBlk - 1) / kBlk, ntiles = nblk * (nblk + 1) / 2;
  for (int64_t i = 0; i < n_links; ++i) {
    if (a[i] >= (uint32_t)n_ids || b[i] >= (uint32_t)n_ids) { set_error("comm_link_owners: a link id is outside [0, %d)", n_ids); return LXY_ERR_ARG; }
    const int64_t lo = std::min(a[i], b[i]) >> kBlkShift, hi = std::max(a[i], b[i]) >> kBlkShift;
    const int64_t t = lo * nblk - lo * (lo - 1) / 2 + (hi - lo);       // tile (lo, hi) of the upper triangle, row-major
    owner[i] = (int32_t)(((t + 1) * world - 1) / ntiles);               // tiles [ntiles r / world, ntiles (r + 1) / world)
  }
  return LXY_OK;
}

int lxy_scaff_score_sharded_dev(lxy_ctx* c, int32_t mode, const int32_t* perm_dev, int32_t P,
                                double* out_dev, int64_t* x100_dev) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  Guard g(c);
  if (!g.ok) return LXY_ERR_CUDA;
  if (P < 0 || (P > 0 && (!perm_dev || !out_dev))) { set_error("score_sharded_dev: bad argument"); return LXY_ERR_ARG; }
  return sharded_score_dev(c, mode, perm_dev, P, out_dev, x100_dev, true);
}

int lxy_scaff_score_sharded(lxy_ctx* c, int32_t mode, const int32_t* perm, int32_t P, double* out,
                            int64_t* x100) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  Guard g(c);
  if (!g.ok) return LXY_ERR_CUDA;
  if (P < 0 || (P > 0 && (!perm || !out))) { set_error("score_sharded: bad argument"); return LXY_ERR_ARG; }
  if (!c->scaff.ready) { set_error("score_sharded: links not set"); return LXY_ERR_STATE; }
  if (mode != LXY_SHARD_POPULATION && mode != LXY_SHARD_LINKS) { set_error("score_sharded: bad mode"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  if (c->world > 1 && !c->comm) { set_error("score_sharded: no communicator"); return LXY_ERR_STATE; }
  cudaStream_t st = c->stream;
  const size_t n = (size_t)c->scaff.n, rowb = 4 * n;
  const int32_t world = c->world, rank = c->rank;
  // upload split: uniform q rows per rank, so that the gathered rows are contiguous
  const int32_t q = (P + world - 1) / world;
  const int32_t up_first = std::min(P, rank * q), up_cnt = std::min(P, (rank + 1) * q) - up_first;
  const Slices sl = make_slices(P, world);
  const size_t pbytes = rowb * (size_t)q * (size_t)world;
  LXY_TRY(c->d_perm.reserve(pbytes));
  LXY_TRY(c->d_out.reserve(8 * (size_t)P));
  LXY_TRY(c->d_x100.reserve(8 * (size_t)P));
  LXY_CUDA(cudaEventRecord(c->ev[4], st));
  if (mode == LXY_SHARD_LINKS) {
    // 1/world of the rows over PCIe, the rest over NVLink
    LXY_TRY(pin_reserve(c, rowb * (size_t)q + 16 * (size_t)P));
    uint8_t* mine = c->d_perm.as<uint8_t>() + (size_t)rank * q * rowb;
    if (up_cnt > 0) LXY_TRY(stage_h2d(c, mine, perm + (size_t)up_first * n, rowb * (size_t)up_cnt, 0));
    if (c->comm && world > 1)
      LXY_NCCL(nccl()->AllGather(mine, c->d_perm.p, rowb * (size_t)q, ncclInt8, (ncclComm_t)c->comm, st));
  } else {
    // only the rows this rank scores
    const int32_t first = sl.first[rank], cnt = sl.first[rank + 1] - first;
    LXY_TRY(pin_reserve(c, rowb * (size_t)std::max(cnt, 1) + 16 * (size_t)P));
    if (cnt > 0) LXY_TRY(stage_h2d(c, c->d_perm.as<uint8_t>() + (size_t)first * rowb, perm + (size_t)first * n, rowb * (size_t)cnt, 0));
  }
  LXY_TRY(sharded_score_dev(c, mode, c->d_perm.as<int32_t>(), P, c->d_out.as<double>(), c->d_x100.as<int64_t>(), false));
  const size_t used = (mode == LXY_SHARD_LINKS) ? rowb * (size_t)q
                                               : rowb * (size_t)std::max(sl.first[rank + 1] - sl.first[rank], 1);
  char* hout = (char*)c->h_pin + used;
  LXY_CUDA(cudaMemcpyAsync(hout, c->d_out.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaMemcpyAsync(hout + 8 * (size_t)P, c->d_x100.p, 8 * (size_t)P, cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaEventRecord(c->ev[5], st));
  c->timing_pending = 2;
  LXY_TRY(check_flags_now(c, "sharded scaff score"));   // synchronises; the flags are checked even with NO_VALIDATE (they are then 0)
  std::memcpy(out, hout, 8 * (size_t)P);
  if (x100) std::memcpy(x100, hout + 8 * (size_t)P, 8 * (size_t)P);
  return LXY_OK;
}

}  // extern "C"

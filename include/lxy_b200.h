/* lxy_b200.h — C ABI of liblxyb200.so: LXY's GA fitness evaluation on one B200 (sm_100a).
 *
 * This is the drop-in boundary.  The reference has no FFI; the hooks it exposes for this path are
 * the two package-private Go functions handed to the genomes as `sfunc`
 *     func score(g *GAOrderedIntGenome) float64         /root/reference/src/lxy/scaff/scaff.go:246-310
 *     func score(g *GAFixedBitstringGenome) float64     /root/reference/src/lxy/phase/phase.go:171-207
 * reached through Genome.Score()  (scaff/genome.go:153-159, phase/genome.go:87-93), both reading
 * the link table sequtil.Links through Links.Get (src/sequtil/links.go:190-221).  A cgo shim binds
 * the entry points below (see INTEGRATION.md for the Go side).
 *
 * Conventions: every function returns 0 (LXY_OK) or a negative lxy_status; nothing throws or
 * aborts across the boundary; the message of the last failure on the calling thread is
 * lxy_last_error().  All pointers are HOST pointers unless the name ends in _dev.  The library
 * copies what it needs before returning and never retains caller memory (cgo pointer rule).
 * A context belongs to one CUDA device; entry points take the context mutex, so concurrent
 * callers (go-galib runs Score() from 7 goroutines, scaff.go:88) are serialised, not corrupted.
 * There is NO CPU fallback: without a CUDA device lxy_ctx_create fails with LXY_ERR_CUDA.
 */
#ifndef LXY_B200_H_
#define LXY_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LXY_ABI_VERSION 2

typedef enum lxy_status {
  LXY_OK = 0,
  LXY_ERR_ARG = -1,      /* null pointer, negative size, id out of range                     */
  LXY_ERR_CUDA = -2,     /* a CUDA runtime call failed (message holds cudaGetErrorString)     */
  LXY_ERR_STATE = -3,    /* e.g. scoring before links were set                                */
  LXY_ERR_BAD_GENE = -4, /* a row of the population is not a permutation of 0..n-1 / not 0,1  */
  LXY_ERR_RANGE = -5,    /* value does not fit the device format (n too large, overflow)      */
  LXY_ERR_NOMEM = -6
} lxy_status;

/* accumulation mode of the scaffolding scorers */
#define LXY_ACC_AUTO 0 /* exact integers when every weight is integral (|w| < 2^31), else fp64 */
#define LXY_ACC_F64 1  /* always fp64 accumulation                                            */
#define LXY_ACC_INT 2  /* require integral weights (error otherwise)                          */

#define LXY_FLAG_NO_VALIDATE 1u /* skip the on-device permutation / bit check                  */
#define LXY_FLAG_NO_STREAM1 2u  /* score batches of <= 8 orderings with the batch kernel too    \
                                   (default: the lane = record streaming kernel, HBM-bound)     */

#define LXY_FLAG_NO_GRAPH 8u    /* never replay a repeated scoring call from a captured CUDA graph */
#define LXY_FLAG_DEFER_ERRORS 16u /* device-pointer scoring calls (lxy_*_dev) do not wait for their validation  \
                                     result: the on-device checks still run in every call, their flags are        \
                                     copied to the host asynchronously and reported by the next lxy_ctx_sync()    \
                                     (LXY_ERR_BAD_GENE / LXY_ERR_RANGE for "one of the calls since the last       \
                                     sync").  Calls then queue back to back on the context's stream, as CUDA's    \
                                     own asynchronous errors do; host-buffer entry points return results and    \
                                     therefore errors synchronously as before. */

/* which kernels score the R-profile (lxy_scaff_score_batch*); every path returns the same integers */
#define LXY_PATH_AUTO 0   /* default (a zero-initialised lxy_cfg): chosen at lxy_ctx_set_links* from what is  \
                             known then - the dense pair table when it fits (4 n^2 bytes, 8 n^2 with real      \
                             weights, at most 16 GiB and a quarter of the free device memory) and the link   \
                             set is dense enough for it to win (>= 256 records per contig); else the record \
                             stream, with the records of a pair summed into one int32-weighted record        \
                             (Links.Add, links.go:149-177) when that at least halves the stream.            \
                             lxy_ctx_path reports the choice.                                               */
#define LXY_PATH_STREAM 1 /* stream every record as given for every candidate (north_star design)             */
#define LXY_PATH_PAIRS 2  /* aggregate the records into a dense pair table at set_links and score a         \
                             candidate with the reference's own 6n table gathers (cost independent of the    \
                             number of records)                                                             */

typedef struct lxy_cfg {
  int32_t device;   /* CUDA ordinal                                                          */
  int32_t acc_mode; /* LXY_ACC_*                                                             */
  uint32_t flags;   /* LXY_FLAG_*                                                            */
  int32_t path;     /* LXY_PATH_*                                                            */
} lxy_cfg;

typedef struct lxy_ctx lxy_ctx;

/* Device time of the most recent scoring call, measured with CUDA events on the context's own
 * stream (bench.py reads these; they are the roofline numerators' denominators). */
typedef struct lxy_timing {
  float ms_total;    /* first kernel start -> last kernel end (excludes host<->device copies)  */
  float ms_layout;   /* ordering -> placement tables (inverse permutation / prefix scan)       */
  float ms_score;    /* the link-streaming fitness kernel (the dominant kernel)                */
  float ms_finish;   /* head-of-ordering correction + fixed-order reduction                    */
  float ms_h2d;      /* host->device copy of the population (0 for *_dev entry points)         */
  float ms_d2h;      /* device->host copy of the scores                                        */
  int32_t launches;  /* kernels launched by the call                                           */
  int32_t grid;      /* CTAs of the dominant kernel                                            */
  int64_t link_bytes; /* bytes of link records the dominant kernel read per pass              */
  int64_t table_bytes; /* bytes of placement tables it staged into shared memory              */
} lxy_timing;

int lxy_abi_version(void);
const char* lxy_last_error(void);
int lxy_device_count(int32_t* count);

int lxy_ctx_create(const lxy_cfg* cfg, lxy_ctx** out);
void lxy_ctx_destroy(lxy_ctx* ctx);
int lxy_ctx_get_timing(lxy_ctx* ctx, lxy_timing* out);
int lxy_ctx_stream(lxy_ctx* ctx, void** cuda_stream); /* cudaStream_t the kernels run on       */
int lxy_ctx_sync(lxy_ctx* ctx);
/* The path the scaffolding scorer of this context takes (LXY_PATH_STREAM or LXY_PATH_PAIRS; for a
 * LXY_PATH_AUTO context the choice made by the last lxy_ctx_set_links*), and the number of records the
 * streaming kernels read per candidate (smaller than n_links when duplicates were summed). */
int lxy_ctx_path(lxy_ctx* ctx, int32_t* path, int64_t* stream_records);

/* ---- scaffolding, R-profile = the reference scorer (scaff.go:246-310) -------------------------
 * Link records (a, b, w): ids in [0, n_ids), w nullable (= 1 per record, what
 * ScaffoldLinksFromSam adds per read pair, src/sequtil/alignment.go:656-660).  Records of the same
 * pair accumulate (Links.Add, links.go:149-177); (a,b) == (b,a) (links.go:130-132,162-164).
 * fitness(p) = - sum_records w * T[|pos a - pos b|] * (1 + [min(pos a, pos b) > 0]),
 * T = {1:1, 2:.5, 3:.33, 5:.2, 11:.1, 20:.05}, pos = inverse of the ordering. */
int lxy_ctx_set_links(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a,
                      const uint32_t* b, const double* w);
/* Same, with a_dev / b_dev / w_dev already resident on the context's device (e.g. produced by a
 * device-side SAM -> links step); the arrays are read, not retained. */
int lxy_ctx_set_links_dev(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a_dev,
                          const uint32_t* b_dev, const double* w_dev);
/* population perm[P][n_ids] row-major (Gene of GAOrderedIntGenome, genome.go:20-26, narrowed to
 * int32); out[P] fitness (lower is better, scaff.go:308); x100 (nullable) receives the exact
 * integer sum_records 100*T*w*m when the context runs in integer mode, INT64_MIN otherwise. */
int lxy_scaff_score_batch(lxy_ctx* ctx, const int32_t* perm, int32_t P, double* out,
                          int64_t* x100);
int lxy_scaff_score_batch_dev(lxy_ctx* ctx, const int32_t* perm_dev, int32_t P, double* out_dev,
                              int64_t* x100_dev);
/* The same scores from 16-bit gene values (n_ids <= 65,536, else LXY_ERR_RANGE): half the bytes over PCIe
 * for a caller whose orderings live in host memory - scaff/genome.go keeps them as []int, so the shim
 * narrows while it copies the rows it copies anyway (go/scaff/gpu.go).  On the dense pair-table path
 * (LXY_PATH_AUTO at BASELINE config 3) the copy is most of a call. */
int lxy_scaff_score_batch_u16(lxy_ctx* ctx, const uint16_t* perm, int32_t P, double* out, int64_t* x100);
/* Integer checks: pos[P][n_ids] (placement offset of every contig = its ordering index in the
 * R-profile) and, for ONE candidate, |pos a - pos b| of caller-supplied records in their order. */
int lxy_scaff_positions(lxy_ctx* ctx, const int32_t* perm, int32_t P, int32_t* pos);
int lxy_scaff_link_distances(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a,
                             const uint32_t* b, const int32_t* perm, int64_t* dist);
/* Aggregated pair weight as Links.Get returns it (links.go:190-221): 0 for an absent pair,
 * LXY_ERR_ARG (and *out = -1) for an unknown id. */
int lxy_links_get(lxy_ctx* ctx, int32_t id1, int32_t id2, double* out);

/* ---- scaffolding, G-profile (north_star generalisation; no reference counterpart) -------------
 * contig lengths len[n_ids], records (a, xa, b, xb, w) with in-contig offsets, per-candidate
 * orientation bytes ori[P][n_ids] indexed by ORDERING POSITION (nullable = all forward).
 * start = exclusive prefix sum of len over the ordering; coord = start + (ori ? len-1-x : x);
 * bucket = |coord a - coord b| / unit; fitness as above with T[bucket].  len==1, x==0, unit==1
 * reproduces the R-profile bit for bit. */
int lxy_ctx_set_links_g(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a,
                        const uint32_t* xa, const uint32_t* b, const uint32_t* xb,
                        const double* w, const uint32_t* len, uint32_t unit);
/* Same with the record arrays already on the device (len stays a host array of n_ids entries). */
int lxy_ctx_set_links_g_dev(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a_dev,
                            const uint32_t* xa_dev, const uint32_t* b_dev, const uint32_t* xb_dev,
                            const double* w_dev, const uint32_t* len, uint32_t unit);
int lxy_scaffg_score_batch(lxy_ctx* ctx, const int32_t* perm, const uint8_t* ori, int32_t P,
                           double* out, int64_t* x100);
int lxy_scaffg_score_batch_dev(lxy_ctx* ctx, const int32_t* perm_dev, const uint8_t* ori_dev,
                               int32_t P, double* out_dev, int64_t* x100_dev);
int lxy_scaffg_layout(lxy_ctx* ctx, const int32_t* perm, int32_t P, uint64_t* start_by_pos,
                      int32_t* pos_by_contig);
int lxy_scaffg_link_distances(lxy_ctx* ctx, int64_t n_links, const uint32_t* a,
                              const uint32_t* xa, const uint32_t* b, const uint32_t* xb,
                              const int32_t* perm, const uint8_t* ori, uint64_t* dist);

/* ---- phasing (phase.go:171-207) ---------------------------------------------------------------
 * ids are gene indices (phase.go:185); records (a, b, w) with signed integral w (read-pair counts,
 * links.go:401-412).  genes[P][n_ids] bytes, 0 = false.
 * fitness(p) = - sum_{records, |a-b| in steps} w * (g[a]==g[b] ? +1 : -1) * (1 + [min(a,b) > 0]),
 * steps = {1..11, 15, 20, 25, 30}.  out_i64 is exact; out_f64 = (double)out_i64. */
int lxy_ctx_set_phase_links(lxy_ctx* ctx, int32_t n_ids, int64_t n_links, const uint32_t* a,
                            const uint32_t* b, const double* w);
int lxy_phase_score_batch(lxy_ctx* ctx, const uint8_t* genes, int32_t P, int64_t* out_i64,
                          double* out_f64);
int lxy_phase_score_batch_dev(lxy_ctx* ctx, const uint8_t* genes_dev, int32_t P,
                              int64_t* out_i64_dev, double* out_f64_dev);
/* The same scores from a ROW-PACKED population: words[P][ceil(n_ids / 32)], bit (i & 31) of word
 * [p][i >> 5] = gene i of candidate p (bits past n_ids are ignored).  One eighth of the bytes over PCIe:
 * the Go shim packs the []bool rows it copies anyway. */
int lxy_phase_score_batch_bits(lxy_ctx* ctx, const uint32_t* words, int32_t P, int64_t* out_i64,
                               double* out_f64);
int lxy_phase_score_batch_bits_dev(lxy_ctx* ctx, const uint32_t* words_dev, int32_t P,
                                   int64_t* out_i64_dev, double* out_f64_dev);
/* Allocation-free variant of lxy_phase_score_batch_bits_dev.  It has NO counterpart in the reference: it
 * was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same scores, same
 * asynchronous behaviour on the context's stream; the call itself performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.  Its scratch (the per-candidate accumulator and the bit-sliced
 * gene words) lives in one static device array of LXY_PHASE_NOALLOC_SCRATCH_BYTES that every GPU which
 * loads the library holds until the process exits, so the size is bounded:
 *     n_ids * Ppad / 8 + 8 * Ppad <= LXY_PHASE_NOALLOC_SCRATCH_BYTES,
 * Ppad = P rounded up to 32 (P <= 32), 64 (P <= 64) or a multiple of 128; beyond that it returns
 * LXY_ERR_RANGE and leaves the outputs untouched (use lxy_phase_score_batch_bits_dev).  The links must
 * have been set before (lxy_ctx_set_phase_links allocates; this call does not).  All calls on one device
 * share the array: it keeps the last call's contents (the call clears what must start at zero), and two
 * calls that overlap on the GPU - two contexts of one device, i.e. two streams - race on it. */
#define LXY_PHASE_NOALLOC_SCRATCH_BYTES (8u << 20)
int lxy_phase_score_batch_bits_noalloc_dev(lxy_ctx* ctx, const uint32_t* words_dev, int32_t P,
                                           int64_t* out_i64_dev, double* out_f64_dev);

/* ---- multi-GPU: one context per device, NCCL on the context's own stream ----------------------------
 * north_star: "the population is sharded across GPUs for large populations, and the links are sharded
 * with an NCCL allreduce of partial fitness for huge link sets; only the per-generation score vector
 * moves between GPUs".  The reference's host is ONE Go process (scaff.go:88-94): it creates one context
 * per device and either calls lxy_comm_init_all, or - one process per GPU (torchrun, MPI) - rank 0 calls
 * lxy_comm_unique_id, ships the 128 bytes to the other ranks by any means, and every rank calls
 * lxy_comm_init_rank.  NCCL (libnccl.so.2) is loaded on first use; single-GPU callers never need it.
 * All ranks must make the same sharded calls in the same order; with several contexts in one process
 * each context needs its own host thread (a goroutine locked to an OS thread), because a call returns
 * only when its collective has completed.  Errors found on any rank (bad gene, range) are returned on
 * every rank.  world <= 64. */
#define LXY_COMM_ID_BYTES 128
int lxy_comm_unique_id(void* id128);
int lxy_comm_init_rank(lxy_ctx* ctx, int32_t world, int32_t rank, const void* id128);
int lxy_comm_init_all(lxy_ctx* const* ctxs, int32_t world); /* rank i = ctxs[i], one process      */
int lxy_comm_destroy(lxy_ctx* ctx);
int lxy_comm_info(lxy_ctx* ctx, int32_t* world, int32_t* rank); /* (1, 0) without a communicator  */
/* Rows [first, first + count) of `total` that rank `rank` of `world` scores (balanced contiguous split). */
int lxy_comm_slice(int32_t total, int32_t world, int32_t rank, int32_t* first, int32_t* count);

/* How to split the records for LXY_SHARD_LINKS (any split gives the same scores; this one is the fast
 * one): owner[i] = the rank that should hold record (a[i], b[i]) when every rank takes a contiguous
 * range of the tiles of the sorted store (256 x 256 contig blocks of the upper triangle, row-major),
 * tiles [T r / world, T (r + 1) / world).  A rank then owns whole tiles, so the scoring kernel stages
 * each tile's placement tables once per pass instead of once per rank; with a random 1/world of every
 * tile on every rank that staging is amortised over world times fewer records (measured at 8 GPUs on
 * BASELINE configs[3]: 38.1 ms instead of 33.7 ms per step). */
int lxy_comm_link_owners(int32_t n_ids, int64_t n_links, const uint32_t* a, const uint32_t* b,
                         int32_t world, int32_t* owner);

#define LXY_SHARD_POPULATION 0 /* links replicated on every rank; rank r scores its slice of the        \
                                  population, ncclAllGather of the score slices                        */
#define LXY_SHARD_LINKS 1      /* every rank holds a share of the records (its own lxy_ctx_set_links*)   \
                                  and scores the whole population against it; exact int64 partial sums   \
                                  are combined with ncclAllReduce (bit-identical to the unsharded result), \
                                  fp64 partials are all-gathered and added in rank order                */
/* perm[P][n_ids] is the WHOLE population on every rank (host pointer; _dev: resident on this rank's
 * device); out[P] / x100[P] receive the whole result on every rank.  The host entry point moves each
 * population row over PCIe once per box, not once per GPU: in LXY_SHARD_LINKS mode a rank uploads 1/world
 * of the rows and the rest arrives by ncclAllGather over NVLink; in LXY_SHARD_POPULATION mode a rank
 * uploads only the rows it scores. */
int lxy_scaff_score_sharded(lxy_ctx* ctx, int32_t mode, const int32_t* perm, int32_t P, double* out,
                            int64_t* x100);
int lxy_scaff_score_sharded_dev(lxy_ctx* ctx, int32_t mode, const int32_t* perm_dev, int32_t P,
                                double* out_dev, int64_t* x100_dev);

/* ---- rank / elite (go-galib sorts by Score(); lower is better, ties keep population order) --- */
int lxy_rank(const double* scores, int32_t P, int32_t* order);

/* ---- host-side link table: sequtil.Links as the fitness path needs it --------------------------
 * LoadLinks (links.go:427-491): lines starting with '#' skipped, fields split at every single space,
 * ids by first appearance, a malformed value reads as 0, Set semantics (the last row of a pair wins,
 * (a,b) == (b,a)), LXY_ERR_STATE when nothing was loaded; LXY_ERR_ARG where the Go code would panic
 * (empty line, fewer than three fields).  lxy_links_id(.., create != 0) is Links.ID (links.go:100-110);
 * lxy_links_name is Decode for one id (links.go:276-286, NULL for an unknown id); lxy_links_value is
 * Links.Get (links.go:190-221: absent pair 0; unknown id -> -1 and LXY_ERR_ARG); lxy_links_set / _add
 * are Set / Add (links.go:113-177); lxy_links_write is Write (links.go:242-264). */
typedef struct lxy_links lxy_links;
int lxy_links_new(lxy_links** out);
int lxy_links_load(const char* path, lxy_links** out);
void lxy_links_free(lxy_links* l);
int lxy_links_write(const lxy_links* l, const char* path);
int32_t lxy_links_size(const lxy_links* l);
int64_t lxy_links_npairs(const lxy_links* l);
int32_t lxy_links_id(lxy_links* l, const char* name, int32_t create);
const char* lxy_links_name(const lxy_links* l, int32_t id);
int lxy_links_set(lxy_links* l, int32_t id1, int32_t id2, double v);
int lxy_links_add(lxy_links* l, int32_t id1, int32_t id2, double v);
int lxy_links_value(const lxy_links* l, int32_t id1, int32_t id2, double* out);
int lxy_links_records(const lxy_links* l, uint32_t* a, uint32_t* b, double* w); /* npairs entries, a <= b */
/* kind 0: the table becomes the scaffolding link store (lxy_ctx_set_links); 1: the phasing store. */
int lxy_ctx_set_links_from(lxy_ctx* ctx, const lxy_links* l, int32_t kind);

/* ---- LoadLinks on the device (lxy_b200/csrc/links_parse.cu) -----------------------------------------
 * The same table as lxy_links_load (names -> ids by first appearance; one record per pair, in order
 * of the pair's first row, holding the value of its LAST row; the same error returns), built by
 * kernels: the text crosses PCIe once, lines, fields, name hashing, decimal -> double conversion,
 * id assignment and the last-row-wins reduction all run on the GPU.  Values outside the exactly
 * rounded fast path (inf / nan / hex / 20+ digits / malformed) are resolved by the host loader's own
 * conversion, so the two loaders agree on every bit.  lxy_links_load_gpu returns the table as a
 * host object; lxy_ctx_load_links also makes it the context's link store (kind 0: scaffolding, the
 * record arrays never leave the device; kind 1: phasing) - `out` and `stats` are nullable. */
typedef struct lxy_load_stats {
  int64_t bytes;       /* file size                                                          */
  int64_t lines;       /* lines incl. comments                                               */
  int64_t records;     /* data lines                                                         */
  int64_t pairs;       /* distinct pairs = records of the table                              */
  int32_t names;       /* distinct names = Size()                                            */
  int32_t slow_values; /* value tokens resolved on the host                                  */
  int32_t attempts;    /* 1 unless a 64-bit name-hash collision was detected and re-seeded   */
  int32_t reserved;
  float ms_read;       /* file -> host memory                                                */
  float ms_h2d;        /* text host -> device                                                */
  float ms_device;     /* all kernels, from text resident to records resident                */
  float ms_total;      /* wall clock of the call's parsing part                              */
} lxy_load_stats;
int lxy_links_load_gpu(const char* path, int32_t device, lxy_links** out, lxy_load_stats* stats);
int lxy_ctx_load_links(lxy_ctx* ctx, const char* path, int32_t kind, lxy_links** out,
                       lxy_load_stats* stats);

/* ---- ScaffoldLinksFromSam (src/sequtil/alignment.go:624-676): SAM text -> contig link table --------
 * The step before LoadLinks in a scaffolding run (`lxy scaff links`, scaff/cmd.go:225).  Semantics of
 * the reference loop, quirks included: lines starting with '@' are skipped; every other line is split
 * at tabs; consecutive lines with the same field 0 (read id) form a group; when the NEXT read id
 * arrives, a group of exactly two hits adds 1 to the pair (ID(field 2 of hit 1), ID(field 2 of hit
 * 2)) - names receive their ids at that moment, in that order, same-contig pairs included; the group
 * open at end of file is never added (the loop has no final flush).  LXY_ERR_ARG where the Go code
 * would panic (empty line: line[0] at :648; fewer than three fields: arr[2] at :668).  An empty
 * result is not an error (the reference writes the empty table).
 *   lxy_links_from_sam           the host loop (lxy_b200/csrc/links_io.cu)
 *   lxy_links_from_sam_gpu       the same table built by kernels (links_parse.cu): tab split, group
 *                                detection, name ids and pair counts on the device
 *   lxy_scaffold_links_from_sam  ScaffoldLinksFromSam(samPath, outPath): device table + Write
 *   lxy_ctx_load_sam             SAM -> the context's scaffolding link store, the records never leave
 *                                the device; `out` (nullable) receives the table (names for Decode). */
int lxy_links_from_sam(const char* sam_path, lxy_links** out);
int lxy_links_from_sam_gpu(const char* sam_path, int32_t device, lxy_links** out, lxy_load_stats* stats);
int lxy_scaffold_links_from_sam(const char* sam_path, const char* out_path, int32_t device,
                                lxy_load_stats* stats);
int lxy_ctx_load_sam(lxy_ctx* ctx, const char* sam_path, lxy_links** out, lxy_load_stats* stats);

/* ---- TabulateVariantBlockLinks (src/sequtil/links.go:372-420): read pairs -> phasing link table -------
 * The accumulation step of BlockLinksFromSam (src/sequtil/alignment.go:461-555, `lxy phase links`), i.e.
 * what produces every phasing `.links` file.  For n_pairs read pairs of ONE chromosome: mate 1 of pair p
 * holds the entries blk1[4 * off1[p] .. 4 * off1[p+1]), mate 2 blk2[4 * off2[p] .. 4 * off2[p+1]); an
 * entry is the map entry GetVariantBlocks builds (alignment.go:265-323): {block number, #ref, #alt,
 * #other} (block numbers distinct within a mate).  For every (k1 of mate 1, k2 of mate 2) with k1 > k2
 * and #other == 0 on both sides, Links.Add(ID("chrom_k1"), ID("chrom_k2"), +1) when both blocks were read
 * ref-major or both alt-major (strictly), else -1.  *ct = number of such additions, *balance = their sum
 * (the function's return values).  device >= 0: the tabulation runs on that GPU (count / scan / emit /
 * radix sort / reduce by key, varblock_links.cu); device < 0: the host loop.  Both leave the same table:
 * new names get their ids in ascending (k1, k2) order of the pairs they first occur in, id1 before id2
 * (the reference's order is Go map iteration order, random from run to run - compare tables by name). */
int lxy_links_tabulate_varblocks(lxy_links* l, const char* chrom, int64_t n_pairs, const int64_t* off1,
                                 const int32_t* blk1, const int64_t* off2, const int32_t* blk2,
                                 int32_t device, int64_t* ct, int64_t* balance);

/* ---- device-resident GA generation loop --------------------------------------------------------
 * Replaces the loop `gao.Optimize(1); best := gao.Best()` of Scaffold / Phase (scaff.go:105-126,
 * phase.go:74-95) and the genome operators it drives (scaff/genome.go:47-91,93-97,125-138,161-209;
 * phase/genome.go:37-72,95-131): selection, PMX / two-point crossover, the shift / switch / invert
 * mutators, scoring of the new individuals, sort and truncation all run on the device; per
 * generation only the child count (4 bytes) crosses PCIe.  Every random decision is a pure
 * function of (seed, generation, slot, draw) — include/lxy_ga_plan.h — so runs are reproducible and
 * comparable with the CPU oracle.  The context must have its links set (lxy_ctx_set_links for
 * LXY_GA_SCAFF, lxy_ctx_set_phase_links for LXY_GA_PHASE) and outlive the engine. */
#define LXY_GA_SCAFF 0 /* GAOrderedIntGenome: permutations of 0..n-1, scored by scaff.score */
#define LXY_GA_PHASE 1 /* GAFixedBitstringGenome: n bits, scored by phase.score           */

typedef struct lxy_ga_params {
  int32_t kind;      /* LXY_GA_*                                                               */
  int32_t popsize;   /* gao.Init(40, genome) scaff.go:94; Init(10, ...) phase.go:63            */
  double p_breed;    /* GAParameter.PBreed  (.2 scaff.go:85, .7 phase.go:53)                   */
  double p_mutate;   /* GAParameter.PMutate (.6 scaff.go:84, .7 phase.go:52)                   */
  double tour_p;     /* NewGATournamentSelector(0.7, 5) scaff.go:82, phase.go:50               */
  int32_t tour_k;    /* contestants, <= 16                                                     */
  uint32_t mutators; /* LXY_MUT_* mask (lxy_ga_plan.h); scaff.go:72-78 registers all three,    \
                        phase.go:42-46 invert + switch                                        */
  uint64_t seed;
} lxy_ga_params;

typedef struct lxy_ga_stats {
  int64_t generations;     /* generations run so far                                           */
  int64_t scored;          /* individuals scored so far (the reference's `scores` counter)     */
  int32_t last_children;   /* children appended by the most recent generation                  */
  float ms_last;           /* device time of the most recent generation                        */
  float ms_operators;      /*   plan + children (selection, crossover, mutation)               */
  float ms_score;          /*   scoring of the children                                        */
  float ms_select;         /*   sort + truncation + gather                                     */
} lxy_ga_stats;

typedef struct lxy_ga lxy_ga;

int lxy_ga_create(lxy_ctx* ctx, const lxy_ga_params* prm, lxy_ga** out);
void lxy_ga_destroy(lxy_ga* ga);
/* GARandomInitializer (scaff.go:81): every individual is a copy of `first` (nullable: 0..n-1 /
 * all false) followed by Randomize() (n random transpositions, scaff/genome.go:130-138; n random
 * bits, phase/genome.go:61-72) unless randomize == 0.  first: int32[n] (scaff) / uint8[n] (phase).
 * The population is scored and sorted before the call returns. */
int lxy_ga_init(lxy_ga* ga, const void* first, int32_t randomize);
/* Population from the caller (popsize rows, host memory); scored and sorted. */
int lxy_ga_set_population(lxy_ga* ga, const void* genes);
int lxy_ga_optimize(lxy_ga* ga, int32_t generations);
/* Shard the scoring of lxy_ga_init / lxy_ga_set_population / lxy_ga_optimize over the context's
 * communicator (LXY_GA_SCAFF only): mode LXY_SHARD_POPULATION / LXY_SHARD_LINKS, or -1 for local scoring
 * (the default).  Every rank runs the same deterministic plan and makes the same children; only the
 * score vector crosses NVLink. */
int lxy_ga_set_sharding(lxy_ga* ga, int32_t mode);
/* One generation in two halves, for callers that score the children themselves (e.g. sharded
 * over several GPUs): begin makes the children on the device, the caller fills the score arrays
 * (fitness, lower is better; x100 optional, unused by the engine), end sorts and truncates. */
int lxy_ga_step_begin(lxy_ga* ga, int32_t* n_children, void** children_dev, double** scores_dev,
                      int64_t** x100_dev);
int lxy_ga_step_end(lxy_ga* ga);
int lxy_ga_score_children(lxy_ga* ga, int32_t first, int32_t count); /* the local scorer on [first, first+count) */
/* Best() (scaff.go:109): gene (int32[n] / uint8[n], nullable) and fitness of the first individual. */
int lxy_ga_best(lxy_ga* ga, void* gene, double* score);
/* The sorted population: genes[popsize][n] (nullable), scores[popsize] (nullable). */
int lxy_ga_population(lxy_ga* ga, void* genes, double* scores);
/* Test hooks: the children of the generation begun last (host copies), their plan entries. */
int lxy_ga_children(lxy_ga* ga, void* genes, void* plan /* lxy_ga_child[n_children], nullable */);
int lxy_ga_get_stats(lxy_ga* ga, lxy_ga_stats* out);


/* ---- solution-quality metrics for whole populations ---------------------------------------------
 * lxy_eval_scaffolding = EvalScaffolding (scaff.go:329-384) for P orderings against one key:
 * key_rank[c] = index of contig c in the key order (the reference's keyOrder map), out[P][2] =
 * (fraction of index triples i<j<k in key order — or in reversed key order, whichever is larger —,
 * the same over adjacent triples).  O(n^2) per ordering instead of the reference's O(n^3), same
 * numbers.  n <= 65536.
 * lxy_eval_phasing = EvalPhasing (phase.go:211-263): out[P][4] = (score, firstScore, secondScore,
 * thirdScore), including the reference's neighbour-counter quirk (phase.go:225-235). */
int lxy_eval_scaffolding(lxy_ctx* ctx, const int32_t* perm, const int32_t* key_rank, int32_t n,
                         int32_t P, double* out);
int lxy_eval_phasing(lxy_ctx* ctx, const uint8_t* genes, const uint8_t* key, int32_t n, int32_t P,
                     double* out);
/* lxy_eval_phasing_dev = EvalPhasingDev (phase.go:266-325), the metric `lxy phase infer` logs
 * (phase/cmd.go:152): index[i] is the integer after '_' in the name of block i (Atoi of
 * Split(name, "_")[1], :279-280; 0 when it does not parse), the n blocks are those present in both
 * the phasing and the key (the others are skipped, :289-299).  out[P][3] = matches / comparisons over
 * all ordered pairs, over the pairs whose indices differ by <= 2, and by <= 9 (i == j included).
 * O(n log n + P n) instead of the reference's O(P n^2), same numbers.  n <= 2^26. */
int lxy_eval_phasing_dev(lxy_ctx* ctx, const uint8_t* genes, const uint8_t* key, const int32_t* index,
                         int32_t n, int32_t P, double* out);

/* ---- the two workflows, file to file -----------------------------------------------------------
 * lxy_scaffold_file = Scaffold (scaff.go:68-141) as `lxy scaff infer` calls it (scaff/cmd.go:135-202,
 * without the quality metrics and plots): LoadLinks, the GA with the reference's parameters when
 * prm is NULL (population 40, PBreed .2, PMutate .6, tournament(0.7, 5), shift + switch + invert),
 * Decode, WriteScaffolding (one contig name per line).  lxy_phase_file = Phase (phase.go:34-111):
 * population 10, PBreed .7, PMutate .7, invert + switch; output `name 1` / `name 0`, one block per line, as writePhasing emits and
 * readPhasing parses (phase.go:119-167).
 * best_score (nullable) receives Best().Score(). */
int lxy_scaffold_file(const char* links_path, const char* out_path, int32_t iterations,
                      const lxy_ga_params* prm, int32_t device, double* best_score);
int lxy_phase_file(const char* links_path, const char* out_path, int32_t iterations,
                   const lxy_ga_params* prm, int32_t device, double* best_score);

#ifdef __cplusplus
}
#endif
#endif /* LXY_B200_H_ */

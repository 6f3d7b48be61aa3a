// links_io.cu — the host-side link table of liblxyb200: sequtil.Links as far as the fitness path
// needs it (name <-> id maps, the upper-triangular pair table, the text loader / writer), so that
// `LoadLinks -> score / GA -> Decode` runs inside the library.  Pure host C++ (no kernels).
//
// Follows, file:line relative to /root/reference/:
//   Links            src/sequtil/links.go:14-28   (idKey / idKeyRev / maxid / data, smaller id on top)
//   addKey / ID      :51-62, :100-110             (ids by first appearance, dense from 0)
//   Set / Add / Get  :113-145, :149-177, :190-221 (unknown id -> error and -1; absent pair -> 0)
//   Size / Decode    :267-269, :276-286
//   Write            :242-264                     ("# name:id ..." header, then "name name %f")
//   LoadLinks        :427-491                     ('#' lines skipped; strings.Split(line, " "); the
//                                                  ParseFloat error is ignored (-> 0); Set: the last row
//                                                  of a pair wins and (a,b) == (b,a); error when empty)
// Where the Go code would panic (empty line: line[0] at :450; fewer than three fields: arr[2] at
// :473) the loader returns LXY_ERR_ARG.
//   ScaffoldLinksFromSam  src/sequtil/alignment.go:624-676  (lxy_links_from_sam: the table the
//                                                  function builds before links.Write; see there)
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/lxy_ga_plan.h"
#include "common.cuh"
#include "links_table.h"

using namespace lxy;

namespace {

int32_t add_key(lxy_links* l, const std::string& key) {  // links.go:51-62
  const int32_t id = (int32_t)l->name_of.size();
  l->id_of.emplace(key, id);
  l->name_of.push_back(key);
  return id;
}

int32_t id_create(lxy_links* l, const std::string& key) {  // links.go:100-110 / :458-470
  auto it = l->id_of.find(key);
  return it != l->id_of.end() ? it->second : add_key(l, key);
}

int put(lxy_links* l, int32_t id1, int32_t id2, double v, bool accumulate) {
  const int32_t n = (int32_t)l->name_of.size();
  if (id1 < 0 || id2 < 0 || id1 >= n || id2 >= n) {  // links.go:116-127 / :152-160
    set_error("links: an id is not present in the Links object");
    return LXY_ERR_ARG;
  }
  const uint32_t lo = (uint32_t)(id1 < id2 ? id1 : id2), hi = (uint32_t)(id1 < id2 ? id2 : id1);
  const uint64_t key = ((uint64_t)lo << 32) | hi;
  l->ensure_slots();
  auto it = l->slot.find(key);
  if (it == l->slot.end()) {
    l->slot.emplace(key, l->a.size());
    l->a.push_back(lo);
    l->b.push_back(hi);
    l->w.push_back(v);
  } else if (accumulate) {
    l->w[it->second] += v;
  } else {
    l->w[it->second] = v;
  }
  return LXY_OK;
}

}  // namespace

namespace lxy {
// strconv.ParseFloat(tok, 64) with the error ignored (links.go:473): on a syntax error the Go
// function returns 0; on a range error it returns +-Inf, which is what strtod gives as well.
double parse_float_or_zero(const char* s, size_t len) {
  if (len == 0) return 0.0;
  std::string tok(s, len);
  for (char c : tok)  // Go rejects leading/trailing blanks and hex floats without 'p'; keep it simple:
    if (c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f') return 0.0;
  char* end = nullptr;
  errno = 0;
  const double v = std::strtod(tok.c_str(), &end);
  if (end != tok.c_str() + tok.size()) return 0.0;
  return v;
}

}  // namespace lxy

extern "C" {

int lxy_links_new(lxy_links** out) {
  if (!out) { set_error("links_new: null out"); return LXY_ERR_ARG; }
  *out = new (std::nothrow) lxy_links();
  if (!*out) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  return LXY_OK;
}

void lxy_links_free(lxy_links* l) { delete l; }

int lxy_links_load(const char* path, lxy_links** out) {
  if (!path || !out) { set_error("links_load: null argument"); return LXY_ERR_ARG; }
  *out = nullptr;
  FILE* f = std::fopen(path, "rb");
  if (!f) { set_error("links_load: cannot open %s: %s", path, std::strerror(errno)); return LXY_ERR_ARG; }
  std::string text;
  {
    char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, got);
  }
  std::fclose(f);
  lxy_links* l = new (std::nothrow) lxy_links();
  if (!l) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  size_t pos = 0, lineno = 0;
  while (pos < text.size()) {  // bufio.Scanner lines: split at '\n', one trailing '\r' dropped
    size_t eol = text.find('\n', pos);
    if (eol == std::string::npos) eol = text.size();
    size_t len = eol - pos;
    if (len > 0 && text[pos + len - 1] == '\r') --len;
    const char* line = text.data() + pos;
    pos = eol + 1;
    ++lineno;
    if (len == 0) {  // line[0] panics in the reference (links.go:450)
      set_error("links_load: %s:%zu: empty line (the reference panics here)", path, lineno);
      delete l;
      return LXY_ERR_ARG;
    }
    if (line[0] == '#') continue;
    // strings.Split(line, " "): every single space separates, empty fields are kept
    const char* f0 = line;
    const char* e = line + len;
    const char* s1 = (const char*)std::memchr(f0, ' ', (size_t)(e - f0));
    const char* s2 = s1 ? (const char*)std::memchr(s1 + 1, ' ', (size_t)(e - s1 - 1)) : nullptr;
    if (!s1 || !s2) {  // arr[2] panics in the reference (links.go:473)
      set_error("links_load: %s:%zu: fewer than three fields (the reference panics here)", path, lineno);
      delete l;
      return LXY_ERR_ARG;
    }
    const char* f2 = s2 + 1;
    const char* s3 = (const char*)std::memchr(f2, ' ', (size_t)(e - f2));
    const int32_t ida = id_create(l, std::string(f0, (size_t)(s1 - f0)));
    const int32_t idb = id_create(l, std::string(s1 + 1, (size_t)(s2 - s1 - 1)));
    put(l, ida, idb, parse_float_or_zero(f2, (size_t)((s3 ? s3 : e) - f2)), false);  // Set (links.go:477)
  }
  if (l->name_of.empty()) {  // links.go:485-487
    set_error("links_load: Loaded empty links object...");
    delete l;
    return LXY_ERR_STATE;
  }
  *out = l;
  return LXY_OK;
}

// ScaffoldLinksFromSam (alignment.go:624-676) up to links.Write: '@' lines skipped (:648; an empty
// line panics on line[0]), strings.Split(line, "\t") (:650), consecutive lines sharing arr[0] form a
// group, and when the NEXT read id arrives a group of exactly two hits does
// Add(ID(arr[2] of hit 1), ID(arr[2] of hit 2), 1) (:652-660) — names get their ids there, in that
// order.  The group open at end of file is never added (:672-674 has no flush).  arr[2] on a line
// with fewer than three fields panics (:668) -> LXY_ERR_ARG.  An empty result is not an error (the
// Go function writes the empty table).
int lxy_links_from_sam(const char* path, lxy_links** out) {
  if (!path || !out) { set_error("links_from_sam: null argument"); return LXY_ERR_ARG; }
  *out = nullptr;
  FILE* f = std::fopen(path, "rb");
  if (!f) { set_error("links_from_sam: cannot open %s: %s", path, std::strerror(errno)); return LXY_ERR_ARG; }
  std::string text;
  {
    char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, got);
  }
  std::fclose(f);
  lxy_links* l = new (std::nothrow) lxy_links();
  if (!l) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  size_t pos = 0, lineno = 0;
  const char* cur = "";   // currentID
  size_t cur_len = 0;
  const char* hit[2] = {nullptr, nullptr};  // chrHits[0..1]; further hits only count
  size_t hit_len[2] = {0, 0};
  size_t nhits = 0;
  while (pos < text.size()) {
    size_t eol = text.find('\n', pos);
    if (eol == std::string::npos) eol = text.size();
    size_t len = eol - pos;
    if (len > 0 && text[pos + len - 1] == '\r') --len;
    const char* line = text.data() + pos;
    pos = eol + 1;
    ++lineno;
    if (len == 0) {
      set_error("links_from_sam: %s:%zu: empty line (the reference panics here)", path, lineno);
      delete l;
      return LXY_ERR_ARG;
    }
    if (line[0] == '@') continue;
    const char* e = line + len;
    const char* t1 = (const char*)std::memchr(line, '\t', len);
    const size_t qlen = t1 ? (size_t)(t1 - line) : len;
    if (qlen != cur_len || std::memcmp(line, cur, qlen) != 0) {
      if (nhits == 2) {
        const int32_t ida = id_create(l, std::string(hit[0], hit_len[0]));
        const int32_t idb = id_create(l, std::string(hit[1], hit_len[1]));
        put(l, ida, idb, 1.0, true);
      }
      cur = line;
      cur_len = qlen;
      nhits = 0;
    }
    const char* t2 = t1 ? (const char*)std::memchr(t1 + 1, '\t', (size_t)(e - t1 - 1)) : nullptr;
    if (!t2) {
      set_error("links_from_sam: %s:%zu: fewer than three fields (the reference panics here)", path, lineno);
      delete l;
      return LXY_ERR_ARG;
    }
    const char* t3 = (const char*)std::memchr(t2 + 1, '\t', (size_t)(e - t2 - 1));
    if (nhits < 2) {
      hit[nhits] = t2 + 1;
      hit_len[nhits] = (size_t)((t3 ? t3 : e) - (t2 + 1));
    }
    ++nhits;
  }
  *out = l;
  return LXY_OK;
}

int lxy_links_write(const lxy_links* l, const char* path) {  // links.go:242-264
  if (!l || !path) { set_error("links_write: null argument"); return LXY_ERR_ARG; }
  FILE* f = std::fopen(path, "wb");
  if (!f) { set_error("links_write: cannot open %s: %s", path, std::strerror(errno)); return LXY_ERR_ARG; }
  std::fputc('#', f);
  for (size_t i = 0; i < l->name_of.size(); ++i) std::fprintf(f, " %s:%zu", l->name_of[i].c_str(), i);
  std::fputc('\n', f);
  for (size_t k = 0; k < l->a.size(); ++k)
    std::fprintf(f, "%s %s %f\n", l->name_of[l->a[k]].c_str(), l->name_of[l->b[k]].c_str(), l->w[k]);
  std::fclose(f);
  return LXY_OK;
}

int32_t lxy_links_size(const lxy_links* l) { return l ? (int32_t)l->name_of.size() : 0; }  // Size()
int64_t lxy_links_npairs(const lxy_links* l) { return l ? (int64_t)l->a.size() : 0; }

int32_t lxy_links_id(lxy_links* l, const char* name, int32_t create) {  // ID() creates (links.go:100-110)
  if (!l || !name) return -1;
  auto it = l->id_of.find(name);
  if (it != l->id_of.end()) return it->second;
  return create ? add_key(l, name) : -1;
}

const char* lxy_links_name(const lxy_links* l, int32_t id) {  // Decode, one element (links.go:276-286)
  if (!l || id < 0 || id >= (int32_t)l->name_of.size()) {
    set_error("Error decoding links, an input id was not recognized.");
    return nullptr;
  }
  return l->name_of[(size_t)id].c_str();
}

int lxy_links_set(lxy_links* l, int32_t id1, int32_t id2, double v) {
  if (!l) { set_error("links_set: null"); return LXY_ERR_ARG; }
  return put(l, id1, id2, v, false);
}
int lxy_links_add(lxy_links* l, int32_t id1, int32_t id2, double v) {
  if (!l) { set_error("links_add: null"); return LXY_ERR_ARG; }
  return put(l, id1, id2, v, true);
}
int lxy_links_value(const lxy_links* l, int32_t id1, int32_t id2, double* out) {  // Get (links.go:190-221)
  if (!l || !out) { set_error("links_value: null"); return LXY_ERR_ARG; }
  const int32_t n = (int32_t)l->name_of.size();
  if (id1 < 0 || id2 < 0 || id1 >= n || id2 >= n) {
    *out = -1.0;
    set_error("links: an id is not present in the Links object");
    return LXY_ERR_ARG;
  }
  const uint32_t lo = (uint32_t)(id1 < id2 ? id1 : id2), hi = (uint32_t)(id1 < id2 ? id2 : id1);
  l->ensure_slots();
  auto it = l->slot.find(((uint64_t)lo << 32) | hi);
  *out = it == l->slot.end() ? 0.0 : l->w[it->second];
  return LXY_OK;
}

int lxy_links_records(const lxy_links* l, uint32_t* a, uint32_t* b, double* w) {
  if (!l) { set_error("links_records: null"); return LXY_ERR_ARG; }
  const size_t m = l->a.size();
  if (m && (!a || !b || !w)) { set_error("links_records: null output"); return LXY_ERR_ARG; }
  if (m) {
    std::memcpy(a, l->a.data(), 4 * m);
    std::memcpy(b, l->b.data(), 4 * m);
    std::memcpy(w, l->w.data(), 8 * m);
  }
  return LXY_OK;
}

// The table becomes the context's resident link store: kind 0 = scaffolding (lxy_ctx_set_links),
// kind 1 = phasing (lxy_ctx_set_phase_links; ids are gene indices, phase.go:185).
int lxy_ctx_set_links_from(lxy_ctx* c, const lxy_links* l, int32_t kind) {
  if (!c || !l) { set_error("set_links_from: null argument"); return LXY_ERR_ARG; }
  const int32_t n = (int32_t)l->name_of.size();
  const int64_t m = (int64_t)l->a.size();
  if (kind == 1) return lxy_ctx_set_phase_links(c, n, m, l->a.data(), l->b.data(), l->w.data());
  return lxy_ctx_set_links(c, n, m, l->a.data(), l->b.data(), l->w.data());
}

}  // extern "C"

// ---- the two workflows, file to file --------------------------------------------------------------
// Scaffold (scaff.go:68-141) as `lxy scaff infer` drives it (scaff/cmd.go:135-202 minus the quality
// metrics and plots): LoadLinks -> GA (the reference's parameters unless prm is given) -> Decode ->
// WriteScaffolding (one contig name per line, scaff.go:24-39).  Phase (phase.go:34-111):
// LoadLinks -> GA over bitstrings -> DecodePhasing -> writePhasing (`name 1` / `name 0`, space separated, phase.go:151-167;
// readPhasing, phase.go:119-149, accepts exactly that).
namespace {

int run_ga_file(const char* links_path, const char* out_path, int32_t iterations, const lxy_ga_params* prm_in,
                int32_t device, int32_t kind, double* best_score) {
  if (!links_path || !out_path || iterations < 0) { set_error("infer: bad argument"); return LXY_ERR_ARG; }
  lxy_cfg cfg{};
  cfg.device = device;
  lxy_ctx* ctx = nullptr;
  LXY_TRY(lxy_ctx_create(&cfg, &ctx));
  lxy_links* l = nullptr;  // LoadLinks on the device; the records go straight into the link store
  int rc = lxy_ctx_load_links(ctx, links_path, kind, &l, nullptr);
  if (rc != LXY_OK) { lxy_ctx_destroy(ctx); return rc; }
  lxy_ga* ga = nullptr;
  const int32_t n = lxy_links_size(l);
  std::vector<int32_t> gene32;
  std::vector<uint8_t> gene8;
  if (rc == LXY_OK) {
    lxy_ga_params prm{};
    if (prm_in) {
      prm = *prm_in;
    } else if (kind == LXY_GA_SCAFF) {  // scaff.go:72-94
      prm.popsize = 40; prm.p_breed = 0.2; prm.p_mutate = 0.6; prm.tour_p = 0.7; prm.tour_k = 5;
      prm.mutators = LXY_MUT_SHIFT | LXY_MUT_SWITCH | LXY_MUT_INVERT; prm.seed = 1;
    } else {                            // phase.go:42-63
      prm.popsize = 10; prm.p_breed = 0.7; prm.p_mutate = 0.7; prm.tour_p = 0.7; prm.tour_k = 5;
      prm.mutators = LXY_MUT_INVERT | LXY_MUT_SWITCH; prm.seed = 1;
    }
    prm.kind = kind;
    rc = lxy_ga_create(ctx, &prm, &ga);
  }
  if (rc == LXY_OK) rc = lxy_ga_init(ga, nullptr, 1);  // ids 0..n-1 are links.IntIDs() (scaff.go:89) / all false (phase.go:59)
  if (rc == LXY_OK) rc = lxy_ga_optimize(ga, iterations);
  if (rc == LXY_OK) {
    double s = 0;
    if (kind == LXY_GA_SCAFF) { gene32.resize((size_t)n); rc = lxy_ga_best(ga, gene32.data(), &s); }
    else { gene8.resize((size_t)n); rc = lxy_ga_best(ga, gene8.data(), &s); }
    if (best_score) *best_score = s;
  }
  if (rc == LXY_OK) {
    FILE* f = std::fopen(out_path, "wb");
    if (!f) {
      set_error("infer: cannot open %s: %s", out_path, std::strerror(errno));
      rc = LXY_ERR_ARG;
    } else {
      for (int32_t i = 0; i < n; ++i) {
        if (kind == LXY_GA_SCAFF) std::fprintf(f, "%s\n", lxy_links_name(l, gene32[(size_t)i]));
        else std::fprintf(f, "%s %d\n", lxy_links_name(l, i), gene8[(size_t)i] ? 1 : 0);
      }
      std::fclose(f);
    }
  }
  if (ga) lxy_ga_destroy(ga);
  if (ctx) lxy_ctx_destroy(ctx);
  lxy_links_free(l);
  return rc;
}

}  // namespace

extern "C" {
int lxy_scaffold_file(const char* links_path, const char* out_path, int32_t iterations,
                      const lxy_ga_params* prm, int32_t device, double* best_score) {
  return run_ga_file(links_path, out_path, iterations, prm, device, LXY_GA_SCAFF, best_score);
}
int lxy_phase_file(const char* links_path, const char* out_path, int32_t iterations,
                   const lxy_ga_params* prm, int32_t device, double* best_score) {
  return run_ga_file(links_path, out_path, iterations, prm, device, LXY_GA_PHASE, best_score);
}
}

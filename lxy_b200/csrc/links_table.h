// links_table.h — the host-side link table object behind `lxy_links*` (shared by links_io.cu, the host
// loader, and links_parse.cu, the device loader).  sequtil.Links, /root/reference/src/sequtil/links.go:14-28.
#pragma once
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

struct lxy_links {
  std::unordered_map<std::string, int32_t> id_of;  // idKey
  std::vector<std::string> name_of;                // idKeyRev (dense: id == index)
  std::vector<uint32_t> a, b;                      // a <= b, in order of first appearance of the pair
  std::vector<double> w;
  // (lo << 32 | hi) -> index into a/b/w.  The device loader fills a/b/w in bulk and leaves the map
  // to be built by the first Set / Add / Get that needs it (slot_ready == false).
  mutable std::unordered_map<uint64_t, size_t> slot;
  mutable bool slot_ready = true;
  void ensure_slots() const {
    if (slot_ready) return;
    slot.clear();
    slot.reserve(a.size());
    for (size_t k = 0; k < a.size(); ++k) slot.emplace(((uint64_t)a[k] << 32) | b[k], k);
    slot_ready = true;
  }
};

namespace lxy {
// strconv.ParseFloat(tok, 64) with the error ignored (links.go:473), host version (links_io.cu)
double parse_float_or_zero(const char* s, size_t len);
}

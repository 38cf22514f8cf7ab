// Host-side state-slot manager (f4) behind the C ABI: free list over max_num_seqs slots
// and the (request id -> {sequence id -> slot}) map of vLLM 0.10.0's ConstantSizeCache, which the
// reference's HymetaCacheManager / GLACacheManager inherit (vllm_hymeta/models/hymeta_cache.py:24-63,
// model_for_7B/gla_cache.py:17-32; read back at modeling_hymeta.py:1206-1208).  Pure host code (no
// CUDA calls): one call resolves a whole batch and reports, as flat arrays, the slots of the run,
// the sibling forks to copy and the freshly admitted slots to zero, so the caller issues ONE
// hymeta_state_copy_slot per fork and ONE hymeta_state_zero_slots per batch
// (modeling_hymeta.py:1202-1227 does this per request per layer in Python).
// Slot order is the reference's: LIFO pop from the back of the free list, release appends.
#include <string.h>

#include <string>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "common.cuh"

struct hymeta_slot_mgr {
  int max_slots;
  std::vector<int32_t> free_list;
  // insertion-ordered sequences per request (the Python dict's order decides the fork source)
  std::unordered_map<std::string, std::vector<std::pair<int64_t, int32_t>>> map;
};

namespace {
// ids arrive packed: n NUL-terminated strings back to back (one marshalling step on the caller's side)
inline const char* next_id(const char* p) { return p + strlen(p) + 1; }
void release_ids(hymeta_slot_mgr* m, const char* ids, int n) {
  for (int i = 0; i < n; ++i, ids = next_id(ids)) {
    auto it = m->map.find(ids);
    if (it == m->map.end()) continue;
    for (auto& kv : it->second) m->free_list.push_back(kv.second);
    m->map.erase(it);
  }
}
}  // namespace

extern "C" hymeta_slot_mgr* hymeta_slot_mgr_create(int max_slots) {
  if (max_slots < 0) return nullptr;
  auto* m = new hymeta_slot_mgr;
  m->max_slots = max_slots;
  m->free_list.resize(max_slots);
  for (int i = 0; i < max_slots; ++i) m->free_list[i] = i;
  return m;
}

extern "C" void hymeta_slot_mgr_destroy(hymeta_slot_mgr* m) { delete m; }

extern "C" int hymeta_slot_mgr_release(hymeta_slot_mgr* m, const char* finished,
                                       int n_finished) {
  if (!m || n_finished < 0 || (n_finished > 0 && !finished)) return HYMETA_ERR_INVALID_ARG;
  release_ids(m, finished, n_finished);
  return HYMETA_OK;
}

extern "C" int hymeta_slot_mgr_prepare(hymeta_slot_mgr* m, const char* req_ids,
                                       const int64_t* seq_ids, int n, const char* finished,
                                       int n_finished, int32_t* slots_out, int32_t* copy_from,
                                       int32_t* copy_to, int* n_copies, int32_t* new_slots,
                                       int* n_new) {
  if (!m || n < 0 || n_finished < 0 || (n > 0 && (!req_ids || !seq_ids || !slots_out)) ||
      (n_finished > 0 && !finished) || !n_copies || !n_new ||
      (n > 0 && (!copy_from || !copy_to || !new_slots)))
    return HYMETA_ERR_INVALID_ARG;
  release_ids(m, finished, n_finished);
  std::unordered_set<std::string> done;
  const char* f = finished;
  for (int i = 0; i < n_finished; ++i, f = next_id(f)) done.insert(f);
  int nc = 0, nn = 0;
  const char* rp = req_ids;
  for (int i = 0; i < n; ++i, rp = next_id(rp)) {
    const std::string r(rp);
    if (!done.empty() && done.count(r)) {  // finished before it ever ran
      slots_out[i] = -1;
      continue;
    }
    auto it = m->map.find(r);
    if (it == m->map.end()) {
      if (m->free_list.empty()) { *n_copies = nc; *n_new = nn; return HYMETA_ERR_NO_SLOT; }
      const int32_t s = m->free_list.back();
      m->free_list.pop_back();
      m->map[r].emplace_back(seq_ids[i], s);
      new_slots[nn++] = s;
      slots_out[i] = s;
      continue;
    }
    int32_t found = -1;
    for (auto& kv : it->second)
      if (kv.first == seq_ids[i]) { found = kv.second; break; }
    if (found < 0) {  // sibling of a running request (n>1 sampling / beam): fork the first one
      if (m->free_list.empty()) { *n_copies = nc; *n_new = nn; return HYMETA_ERR_NO_SLOT; }
      found = m->free_list.back();
      m->free_list.pop_back();
      copy_from[nc] = it->second.front().second;
      copy_to[nc++] = found;
      it->second.emplace_back(seq_ids[i], found);
    }
    slots_out[i] = found;
  }
  *n_copies = nc;
  *n_new = nn;
  return HYMETA_OK;
}

extern "C" int hymeta_slot_mgr_free_count(const hymeta_slot_mgr* m) {
  return m ? (int)m->free_list.size() : HYMETA_ERR_INVALID_ARG;
}

extern "C" int hymeta_slot_mgr_free_list(const hymeta_slot_mgr* m, int32_t* out, int cap) {
  if (!m || cap < 0 || (cap > 0 && !out)) return HYMETA_ERR_INVALID_ARG;
  const int n = (int)m->free_list.size() < cap ? (int)m->free_list.size() : cap;
  for (int i = 0; i < n; ++i) out[i] = m->free_list[i];
  return (int)m->free_list.size();
}

extern "C" int hymeta_slot_mgr_slot_of(const hymeta_slot_mgr* m, const char* req_id,
                                       int64_t seq_id) {
  if (!m || !req_id) return HYMETA_ERR_INVALID_ARG;
  auto it = m->map.find(req_id);
  if (it == m->map.end()) return -1;
  for (auto& kv : it->second)
    if (kv.first == seq_id) return kv.second;
  return -1;
}

extern "C" int hymeta_slot_mgr_live_count(const hymeta_slot_mgr* m) {
  if (!m) return HYMETA_ERR_INVALID_ARG;
  int n = 0;
  for (auto& e : m->map) n += (int)e.second.size();
  return n;
}

// Enumerate live (request, sequence, slot) triples: request ids NUL-separated into ids_buf (one per
// triple).  Returns the number of triples; *ids_bytes = bytes needed.  Nothing is written beyond the
// given capacities, so a first call with zero capacities sizes the buffers.
extern "C" int hymeta_slot_mgr_live(const hymeta_slot_mgr* m, char* ids_buf, size_t ids_cap,
                                    size_t* ids_bytes, int64_t* seq_ids, int32_t* slots, int cap) {
  if (!m || cap < 0 || !ids_bytes) return HYMETA_ERR_INVALID_ARG;
  size_t off = 0;
  int n = 0;
  for (auto& e : m->map)
    for (auto& kv : e.second) {
      const size_t len = e.first.size() + 1;
      if (ids_buf && off + len <= ids_cap) memcpy(ids_buf + off, e.first.c_str(), len);
      off += len;
      if (n < cap && seq_ids && slots) { seq_ids[n] = kv.first; slots[n] = kv.second; }
      ++n;
    }
  *ids_bytes = off;
  return n;
}

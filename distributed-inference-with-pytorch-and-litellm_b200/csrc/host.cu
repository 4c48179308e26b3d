// Host-only entry points: library info, the bucket planner (a-1) and chunk planning.
#include <algorithm>
#include <utility>
#include <vector>
#include "launch.cuh"

using namespace ddpb200;

static_assert(sizeof(ddpb200_seg) == 48, "ddpb200_seg is part of the ABI");

extern "C" DDPB200_API int ddpb200_abi_version(void) { return DDPB200_ABI_VERSION; }

extern "C" DDPB200_API const char* ddpb200_error_string(int code) {
  if (code == 0) return "ok";
  if (code == DDPB200_E_INVALID) return "ddp_b200: invalid argument";
  if (code == DDPB200_E_UNSUPPORTED) return "ddp_b200: dtype combination not supported";
  if (code == DDPB200_E_TOO_LARGE) return "ddp_b200: size exceeds the stated limit";
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "ddp_b200: unknown error";
}

extern "C" DDPB200_API int ddpb200_sm_count(void) {
  int err = 0;
  const int sms = sm_count_cached(&err);
  return err ? -err : sms;
}

// Greedy size-capped bucketing. A bucket under construction per (dtype, device) key; keys are
// kept in first-seen order so that the leftover buckets come out in a defined order (see below).
extern "C" DDPB200_API int ddpb200_bucket_assign(const int64_t* numel, const int32_t* elem_size,
                                     const int64_t* key, const uint8_t* expect_sparse,
                                     const int64_t* tensor_indices, int64_t n,
                                     const uint64_t* size_limits, int64_t n_limits,
                                     int64_t* out_indices, int64_t* out_bucket_start,
                                     uint64_t* out_bucket_limit, int64_t* out_n_buckets) {
  if (n <= 0 || n_limits <= 0 || !numel || !elem_size || !size_limits || !out_indices ||
      !out_bucket_start || !out_bucket_limit || !out_n_buckets)
    return DDPB200_E_INVALID;

  struct Open { int64_t key; std::vector<int64_t> idx; uint64_t bytes; uint64_t limit; int64_t next_limit; };
  std::vector<Open> open;                                        // one per key, first-seen order
  std::vector<std::pair<std::vector<int64_t>, uint64_t>> done;  // (indices, limit)

  for (int64_t i = 0; i < n; ++i) {
    const int64_t ti = tensor_indices ? tensor_indices[i] : i;
    if (ti < 0) return DDPB200_E_INVALID;
    if (expect_sparse && expect_sparse[ti]) {                    // own bucket, no size limit
      done.emplace_back(std::vector<int64_t>{ti}, 0);
      continue;
    }
    const int64_t k = key ? key[i] : 0;
    Open* o = nullptr;
    for (auto& cand : open) if (cand.key == k) { o = &cand; break; }
    if (!o) { open.push_back(Open{k, {}, 0, 0, 0}); o = &open.back(); }
    o->idx.push_back(ti);
    o->bytes += (uint64_t)numel[i] * (uint64_t)elem_size[i];
    o->limit = size_limits[o->next_limit];
    if (o->bytes >= o->limit) {
      done.emplace_back(std::move(o->idx), o->limit);
      o->idx.clear();
      o->bytes = 0;
      if (o->next_limit + 1 < n_limits) ++o->next_limit;
    }
  }
  // torch walks a std::unordered_map here; with libstdc++ that visits keys newest-first, and
  // the order is observable when tensor_indices is given (no final sort), so it is reproduced.
  for (auto it = open.rbegin(); it != open.rend(); ++it)
    if (!it->idx.empty()) done.emplace_back(std::move(it->idx), it->limit);

  if (!tensor_indices) {
    std::stable_sort(done.begin(), done.end(), [](const auto& a, const auto& b) {
      return *std::min_element(a.first.begin(), a.first.end()) <
             *std::min_element(b.first.begin(), b.first.end());
    });
  }

  int64_t pos = 0;
  for (size_t b = 0; b < done.size(); ++b) {
    out_bucket_start[b] = pos;
    out_bucket_limit[b] = done[b].second;
    for (int64_t ti : done[b].first) out_indices[pos++] = ti;
  }
  out_bucket_start[done.size()] = pos;
  *out_n_buckets = (int64_t)done.size();
  return 0;
}

extern "C" DDPB200_API int ddpb200_bucket_layout(const int64_t* numel, const int64_t* indices,
                                     const int64_t* bucket_start, int64_t n_buckets,
                                     int64_t* out_offset, int64_t* out_bucket_numel) {
  if (n_buckets < 0 || !numel || !indices || !bucket_start || !out_offset || !out_bucket_numel)
    return DDPB200_E_INVALID;
  for (int64_t b = 0; b < n_buckets; ++b) {
    int64_t off = 0;
    for (int64_t i = bucket_start[b]; i < bucket_start[b + 1]; ++i) {
      out_offset[i] = off;
      off += numel[indices[i]];
    }
    out_bucket_numel[b] = off;
  }
  return 0;
}

extern "C" DDPB200_API int ddpb200_plan_chunks(ddpb200_seg* host_table, int64_t n, int64_t chunk_elems,
                                   int64_t* out_total_chunks) {
  if (n < 0 || (n > 0 && !host_table) || chunk_elems <= 0 || chunk_elems % kTileElems != 0 ||
      !out_total_chunks)
    return DDPB200_E_INVALID;
  int64_t c = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (host_table[i].numel < 0) return DDPB200_E_INVALID;
    host_table[i].chunk_begin = c;
    c += (host_table[i].numel + chunk_elems - 1) / chunk_elems;
  }
  *out_total_chunks = c;
  return 0;
}

// Aim for >= 8 chunks per CTA of the persistent grid so the cyclic schedule ends level, within
// [2048, 65536] elements (a power of two times the tile).
extern "C" DDPB200_API int64_t ddpb200_suggest_chunk_elems(int64_t total_numel, int32_t sm_count) {
  if (sm_count <= 0) sm_count = 148;
  const int64_t ctas = (int64_t)sm_count * kCtasPerSm;
  int64_t chunk = 65536;
  while (chunk > kTileElems && total_numel / chunk < ctas * 8) chunk >>= 1;
  return chunk;
}

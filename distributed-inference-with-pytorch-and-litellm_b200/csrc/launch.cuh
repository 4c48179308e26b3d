// Host-side launch helpers shared by the entry points: persistent-grid sizing and the
// dispatch between a device-resident segment table and a by-value (launch-parameter) one.
#pragma once
#include "common.cuh"

namespace ddpb200 {

// SM count of the current device, cached per device ordinal.
inline int sm_count_cached(int* err) {
  static int cache[64] = {0};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { *err = (int)e; return 0; }
  if (dev >= 0 && dev < 64 && cache[dev] > 0) return cache[dev];
  int sms = 0;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) { *err = (int)e; return 0; }
  if (dev >= 0 && dev < 64) cache[dev] = sms;
  return sms;
}

// One CTA (or several, up to kCtasPerSm) per SM; CTAs walk the chunk list cyclically so that
// neighbouring SMs stream neighbouring DRAM pages at the same time.
inline int persistent_grid(int64_t total_chunks, int* err, int ctas_per_sm = kCtasPerSm) {
  const int sms = sm_count_cached(err);
  if (sms <= 0) return 0;
  const int64_t cap = (int64_t)sms * ctas_per_sm;
  return (int)(total_chunks < cap ? total_chunks : cap);
}

inline bool valid_chunking(int64_t n, int64_t total_chunks, int64_t chunk_elems) {
  return n >= 0 && n <= INT32_MAX && total_chunks >= 0 && chunk_elems > 0 &&
         (chunk_elems % kTileElems) == 0;
}

// Calls f(tab, chunks_sub, seg_offset, chunk_offset) once for a device table, or once per
// slice of at most kMaxValSegs segments for a host table. Slices keep absolute chunk_begin
// values; kernels subtract the slice's first chunk_begin. seg_offset / chunk_offset locate the
// slice's per-segment outputs and per-chunk scratch inside the caller's arrays.
template <class F>
inline int for_each_table_slice(const ddpb200_seg* table, int table_on_device, int64_t n,
                                int64_t total_chunks, F&& f) {
  if (table_on_device) {
    DevTable t{table, (int)n};
    return f(t, total_chunks, (int64_t)0, (int64_t)0);
  }
  const int64_t base0 = n > 0 ? table[0].chunk_begin : 0;
  int64_t i = 0;
  while (i < n) {
    const int64_t m = (n - i) < kMaxValSegs ? (n - i) : kMaxValSegs;
    const int64_t cb = table[i].chunk_begin;
    const int64_t ce = (i + m < n) ? table[i + m].chunk_begin : base0 + total_chunks;
    int rc;
    if (m <= 32) {
      ValTable<32> t; t.n = (int)m;
      for (int64_t k = 0; k < m; ++k) t.seg[k] = table[i + k];
      rc = f(t, ce - cb, i, cb - base0);
    } else if (m <= 128) {
      ValTable<128> t; t.n = (int)m;
      for (int64_t k = 0; k < m; ++k) t.seg[k] = table[i + k];
      rc = f(t, ce - cb, i, cb - base0);
    } else {
      ValTable<kMaxValSegs> t; t.n = (int)m;
      for (int64_t k = 0; k < m; ++k) t.seg[k] = table[i + k];
      rc = f(t, ce - cb, i, cb - base0);
    }
    if (rc != 0) return rc;
    i += m;
  }
  return 0;
}

}  // namespace ddpb200

// Kernel (a) pack / unpack (+scale, +cast) and kernel (b) multi-tensor L2 norm.
//
// One persistent launch covers every segment of a bucket (or of the whole model). A CTA
// takes chunk c, c+grid, c+2*grid, ...; a chunk is a contiguous run of one segment, so every
// warp request is a fully coalesced run of 1 KB (256-bit accesses) or 512 B. Roofline: HBM.
//   pack/unpack : s_src + s_dst bytes per element
//   norm        : s bytes per element (read only)
// The sum of squares, when asked for, is reduced warp-shuffle -> shared-memory tree -> one
// float per chunk -> "last CTA" fold per segment (common.cuh), i.e. single pass, one launch.
#include <mutex>
#include "launch.cuh"

namespace ddpb200 {

// One chunk through G-element groups; returns the number of elements done (the ragged end is
// left to the caller's scalar loop). NORM accumulates the squares of the values as stored.
template <int SDT, int DDT, bool STORE, bool NORM, int VECB>
__device__ __forceinline__ int64_t stream_chunk_vec(const char* src, char* dst, int64_t len, float scale,
                                                    float (&acc)[4]) {
  constexpr bool any_f32 = (SDT == DDPB200_F32 || (STORE && DDT == DDPB200_F32));
  constexpr int G = GroupSize<VECB, any_f32>::G;
  const int64_t groups_end = len / G * G;
#pragma unroll 4
  for (int64_t i = (int64_t)threadIdx.x * G; i < groups_end; i += kThreads * G) {
    float x[G];
    GroupIO<SDT, G>::load(src, i, x);
#pragma unroll
    for (int k = 0; k < G; ++k) {
      float y = __fmul_rn(x[k], scale);
      // four independent accumulators: a single one is a 4-cycle dependent chain per element
      if constexpr (NORM) { y = round_through<STORE ? DDT : SDT>(y); acc[k & 3] = __fmaf_rn(y, y, acc[k & 3]); }
      x[k] = y;
    }
    if constexpr (STORE) GroupIO<DDT, G>::store(dst, i, x);
  }
  return groups_end;
}

// SRC / DST: which ptr[] slot is read / written. STORE=false -> norm only.
template <class Tab, int SDT, int DDT, int SRC, int DST, bool STORE, bool NORM, bool DEVSCALE>
__global__ void __launch_bounds__(kThreads, kStreamCtasPerSm)
stream_kernel(const __grid_constant__ Tab tab, int64_t total_chunks, int64_t chunk_elems,
              float scale, const float* __restrict__ scale_ptr, float* __restrict__ sumsq_out,
              float* __restrict__ chunk_partials, uint32_t* ticket) {
  constexpr bool any_f32 = (SDT == DDPB200_F32 || (STORE && DDT == DDPB200_F32));
  constexpr int G32 = GroupSize<32, any_f32>::G, G16 = GroupSize<16, any_f32>::G;
  __shared__ float red[kThreads / 32];

  if constexpr (DEVSCALE) scale = __ldg(scale_ptr);
  const int64_t chunk_base = tab.begin(0);

  SegCursor cur;
  for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const int s = find_seg_cached(tab, c + chunk_base, chunk_base + total_chunks, cur);
    const int64_t start = (c + chunk_base - cur.b) * chunk_elems;
    const int64_t rest = tab.numel(s) - start;
    const int64_t len = rest < chunk_elems ? rest : chunk_elems;
    const char* src = static_cast<const char*>(tab.ptr(s, SRC)) + start * ElemOf<SDT>::size;
    char* dst = nullptr;
    if constexpr (STORE) dst = static_cast<char*>(tab.ptr(s, DST)) + start * ElemOf<DDT>::size;

    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    bool a32 = group_aligned<SDT, G32>(src), a16 = group_aligned<SDT, G16>(src);
    if constexpr (STORE) { a32 = a32 && group_aligned<DDT, G32>(dst); a16 = a16 && group_aligned<DDT, G16>(dst); }
    int64_t done = 0;
    if (a32) done = stream_chunk_vec<SDT, DDT, STORE, NORM, 32>(src, dst, len, scale, acc);
    else if (a16) done = stream_chunk_vec<SDT, DDT, STORE, NORM, 16>(src, dst, len, scale, acc);
    // ragged end, or a segment whose base is not 16-byte aligned (odd element offset inside a bucket)
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
      float y = __fmul_rn(load1<SDT>(src, i), scale);
      if constexpr (NORM) { y = round_through<STORE ? DDT : SDT>(y); acc[0] = __fmaf_rn(y, y, acc[0]); }
      if constexpr (STORE) store1<DDT>(dst, i, y);
    }

    if constexpr (NORM) {
      const float total = block_sum(__fadd_rn(__fadd_rn(acc[0], acc[1]), __fadd_rn(acc[2], acc[3])), red);
      if (threadIdx.x == 0) chunk_partials[c] = total;
    }
  }

  if constexpr (NORM) finish_sumsq(tab, chunk_base, chunk_partials, sumsq_out, total_chunks, ticket);
}

template <int SDT, int DDT, int SRC, int DST, bool STORE, bool NORM, bool DEVSCALE>
static int launch_stream(const ddpb200_seg* table, int table_loc, int64_t n, int64_t total_chunks,
                         int64_t chunk_elems, float scale, const float* scale_ptr, float* sumsq_out,
                         float* chunk_partials, uint32_t* ticket, cudaStream_t st) {
  if (!valid_chunking(n, total_chunks, chunk_elems) || (n > 0 && !table)) return DDPB200_E_INVALID;
  if (NORM && n > 0 && (!sumsq_out || !ticket || (total_chunks > 0 && !chunk_partials)))
    return DDPB200_E_INVALID;
  if (n == 0) return 0;
  return for_each_table_slice(table, table_loc == DDPB200_TABLE_DEVICE, n, total_chunks,
    [&](auto& tab, int64_t chunks, int64_t seg_off, int64_t chunk_off) -> int {
      using Tab = std::remove_reference_t<decltype(tab)>;
      int err = 0;
      int grid = persistent_grid(chunks, &err, kStreamCtasPerSm);
      if (err) return err;
      if (grid == 0) {
        if (!NORM) return 0;
        grid = 1;  // all segments empty: still write zeros to sumsq_out
      }
      stream_kernel<Tab, SDT, DDT, SRC, DST, STORE, NORM, DEVSCALE><<<grid, kThreads, 0, st>>>(
          tab, chunks, chunk_elems, scale, scale_ptr, NORM ? sumsq_out + seg_off : nullptr,
          NORM ? chunk_partials + chunk_off : nullptr, ticket);
      return (int)cudaGetLastError();
    });
}

}  // namespace ddpb200

using namespace ddpb200;

extern "C" DDPB200_API int ddpb200_pack_scale_cast(const ddpb200_seg* table, int32_t table_loc, int64_t n,
                                       int64_t total_chunks, int64_t chunk_elems,
                                       int32_t src_dtype, int32_t dst_dtype, float scale,
                                       float* sumsq_out, float* chunk_partials, uint32_t* ticket,
                                       void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define DDPB200_PACK_CASE(S, D)                                                                  \
  if (src_dtype == S && dst_dtype == D) {                                                        \
    return sumsq_out                                                                             \
      ? launch_stream<S, D, 0, 1, true, true, false>(table, table_loc, n, total_chunks,          \
            chunk_elems, scale, nullptr, sumsq_out, chunk_partials, ticket, st)                  \
      : launch_stream<S, D, 0, 1, true, false, false>(table, table_loc, n, total_chunks,         \
            chunk_elems, scale, nullptr, nullptr, nullptr, nullptr, st);                         \
  }
  DDPB200_PACK_CASE(DDPB200_F32, DDPB200_F32)
  DDPB200_PACK_CASE(DDPB200_BF16, DDPB200_BF16)
  DDPB200_PACK_CASE(DDPB200_F32, DDPB200_BF16)
  DDPB200_PACK_CASE(DDPB200_BF16, DDPB200_F32)
#undef DDPB200_PACK_CASE
  return DDPB200_E_UNSUPPORTED;
}

extern "C" DDPB200_API int ddpb200_l2norm_sumsq(const ddpb200_seg* table, int32_t table_loc, int64_t n,
                                    int64_t total_chunks, int64_t chunk_elems, int32_t dtype,
                                    float* sumsq_out, float* chunk_partials, uint32_t* ticket,
                                    void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DDPB200_F32)
    return launch_stream<DDPB200_F32, DDPB200_F32, 1, 1, false, true, false>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, nullptr, sumsq_out, chunk_partials, ticket, st);
  if (dtype == DDPB200_BF16)
    return launch_stream<DDPB200_BF16, DDPB200_BF16, 1, 1, false, true, false>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, nullptr, sumsq_out, chunk_partials, ticket, st);
  return DDPB200_E_UNSUPPORTED;
}

extern "C" DDPB200_API int ddpb200_scale_by_device_scalar(const ddpb200_seg* table, int32_t table_loc,
                                              int64_t n, int64_t total_chunks,
                                              int64_t chunk_elems, int32_t dtype,
                                              const float* coef, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!coef) return DDPB200_E_INVALID;
  if (dtype == DDPB200_F32)
    return launch_stream<DDPB200_F32, DDPB200_F32, 1, 1, true, false, true>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, coef, nullptr, nullptr, nullptr, st);
  if (dtype == DDPB200_BF16)
    return launch_stream<DDPB200_BF16, DDPB200_BF16, 1, 1, true, false, true>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, coef, nullptr, nullptr, nullptr, st);
  return DDPB200_E_UNSUPPORTED;
}

// ---- total norm + clip coefficient ----------------------------------------------------
namespace ddpb200 {

__global__ void __launch_bounds__(kThreads, 1)
norm_finalize_kernel(const float* __restrict__ sumsq, int64_t n, int mode, float max_norm,
                     float* __restrict__ out) {
  __shared__ double red[kThreads / 32];
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += kThreads) {
    float ss = sumsq[i];
    if (mode == DDPB200_NORM_TORCH_BF16) {
      // torch._foreach_norm returns each tensor's norm in the grad dtype; vector_norm then
      // squares those rounded values in fp32 (clip_grad.py:96,106-108)
      const float nt = round_through_bf16(__fsqrt_rn(ss));
      ss = __fmul_rn(nt, nt);
    }
    acc += (double)ss;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < kThreads / 32; ++w) t += red[w];
    float total = __fsqrt_rn((float)t);
    float coef;
    if (mode == DDPB200_NORM_TORCH_BF16) {
      total = round_through_bf16(total);
      const float d = round_through_bf16(__fadd_rn(total, 1e-6f));   // total_norm + 1e-6
      coef = round_through_bf16(__fdiv_rn(max_norm, d));              // max_norm / (...)
    } else {
      coef = __fdiv_rn(max_norm, __fadd_rn(total, 1e-6f));
    }
    coef = (coef != coef) ? coef : fminf(coef, 1.0f);                 // clamp(max=1.0) keeps NaN
    out[0] = total;
    out[1] = coef;
  }
}

}  // namespace ddpb200

extern "C" DDPB200_API int ddpb200_l2norm_finalize(const float* sumsq, int64_t n, int32_t mode,
                                       float max_norm, float* out, void* stream) {
  if (n < 0 || !out || (n > 0 && !sumsq)) return DDPB200_E_INVALID;
  if (mode != DDPB200_NORM_EXACT_F32 && mode != DDPB200_NORM_TORCH_BF16) return DDPB200_E_INVALID;
  norm_finalize_kernel<<<1, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(sumsq, n, mode, max_norm, out);
  return (int)cudaGetLastError();
}

// ---- L2 norm for callers that cannot bring scratch memory ------------------------------------------
// Same kernel as ddpb200_l2norm_sumsq; the per-chunk partial sums and the ticket counter live in a
// fixed array that is part of this library's device image instead of in caller memory.

namespace ddpb200 {

constexpr int64_t kStaticTicketBytes = 256;   // the ticket sits alone in the first 256 bytes
constexpr int64_t kStaticScratchBytes = kStaticTicketBytes + (int64_t)DDPB200_STATIC_SCRATCH_MAX_CHUNKS * sizeof(float);

// One copy per GPU that loads the library, resident until the process exits (1 MiB + 256 B).
static __device__ __align__(256) unsigned char g_norm_scratch[kStaticScratchBytes];

// In host code `g_norm_scratch` is only a handle for the runtime, not a device address: the
// address has to be asked for, per device, with that device current.
static std::mutex g_scratch_mutex;
static void* g_scratch_addr[64] = {nullptr};

static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return DDPB200_E_INVALID;
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_norm_scratch);      // first call on this device: set-up
    if (e != cudaSuccess) return (int)e;
    if (!p) return DDPB200_E_INVALID;
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace ddpb200

extern "C" DDPB200_API int ddpb200_l2norm_sumsq_static(const ddpb200_seg* table, int32_t table_loc,
                                                       int64_t n, int64_t total_chunks,
                                                       int64_t chunk_elems, int32_t dtype,
                                                       float* sumsq_out, void* stream) {
  if (total_chunks > DDPB200_STATIC_SCRATCH_MAX_CHUNKS) return DDPB200_E_TOO_LARGE;
  // every argument is checked before anything is enqueued: a refused call leaves the shared array alone
  if (!valid_chunking(n, total_chunks, chunk_elems) || (n > 0 && (!table || !sumsq_out))) return DDPB200_E_INVALID;
  if (dtype != DDPB200_F32 && dtype != DDPB200_BF16) return DDPB200_E_UNSUPPORTED;
  if (n == 0) return 0;
  void* base = nullptr;
  const int rc = static_scratch_for_current_device(&base);
  if (rc != 0) return rc;
  uint32_t* ticket = static_cast<uint32_t*>(base);
  float* partials = reinterpret_cast<float*>(static_cast<unsigned char*>(base) + kStaticTicketBytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps whatever the previous call left in it. The partials are fully rewritten before
  // they are read; the ticket is the one thing that must start at zero, so it is cleared here, on
  // the caller's stream, rather than trusted.
  cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(uint32_t), st);
  if (e != cudaSuccess) return (int)e;
  if (dtype == DDPB200_F32)
    return launch_stream<DDPB200_F32, DDPB200_F32, 1, 1, false, true, false>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, nullptr, sumsq_out, partials, ticket, st);
  if (dtype == DDPB200_BF16)
    return launch_stream<DDPB200_BF16, DDPB200_BF16, 1, 1, false, true, false>(
        table, table_loc, n, total_chunks, chunk_elems, 1.0f, nullptr, sumsq_out, partials, ticket, st);
  return DDPB200_E_UNSUPPORTED;
}

// Shared device helpers for the sm_100a multi-tensor streaming kernels.
//
// All three kernels on the hot path (pack/unpack, L2 norm, clip+AdamW) are HBM-bound
// streaming passes with no data reuse, so there is no shared-memory tiling and no tensor
// core work here: the job is to keep enough 16-byte requests in flight per SM, fully
// coalesced, and to spend as few issue slots per byte as possible.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/ddp_b200.h"

namespace ddpb200 {

constexpr int kThreads   = 256;   // threads per CTA
constexpr int kTileElems = 2048;  // elements one CTA moves per inner iteration (8 per thread)
constexpr int kCtasPerSm = 4;     // persistent grid = SMs * kCtasPerSm
constexpr int kMaxValSegs = 448;  // segments that fit a by-value (launch-parameter) table

// ---- segment tables ---------------------------------------------------------------
// Device-resident table: built once, reused every step (static pointers).
struct DevTable {
  const ddpb200_seg* seg;
  int n;
  __device__ __forceinline__ int64_t begin(int i) const { return __ldg(&seg[i].chunk_begin); }
  __device__ __forceinline__ int64_t numel(int i) const { return __ldg(&seg[i].numel); }
  __device__ __forceinline__ void* ptr(int i, int k) const {
    return (void*)__ldg((const unsigned long long*)&seg[i].ptr[k]);
  }
};

// By-value table: travels in the launch parameters (constant bank), so a host-side table
// whose pointers change every step (fresh autograd grads) needs no H2D copy and has no
// lifetime to manage. CUDA >= 12.1 allows 32 KB of parameters on sm_70+.
template <int K>
struct ValTable {
  ddpb200_seg seg[K];
  int n;
  __device__ __forceinline__ int64_t begin(int i) const { return seg[i].chunk_begin; }
  __device__ __forceinline__ int64_t numel(int i) const { return seg[i].numel; }
  __device__ __forceinline__ void* ptr(int i, int k) const { return seg[i].ptr[k]; }
};

// Last segment whose chunk_begin <= c (upper_bound - 1); empty segments are skipped because
// they share chunk_begin with their successor.
template <class Tab>
__device__ __forceinline__ int find_seg(const Tab& tab, int64_t c) {
  int lo = 0, hi = tab.n;  // answer in [lo, hi)
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (tab.begin(mid) <= c) lo = mid; else hi = mid;
  }
  return lo;
}

// ---- element types -----------------------------------------------------------------
template <int DT> struct ElemOf;
template <> struct ElemOf<DDPB200_F32>  { using type = float;          static constexpr int size = 4; };
template <> struct ElemOf<DDPB200_BF16> { using type = __nv_bfloat16;  static constexpr int size = 2; };

__device__ __forceinline__ float bf16_bits_to_f32(uint32_t hi16) { return __uint_as_float(hi16 << 16); }

// round-to-nearest-even fp32 -> bf16 -> fp32 (what a store+reload through a bf16 tensor does)
__device__ __forceinline__ float round_through_bf16(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}
template <int DT> __device__ __forceinline__ float round_through(float x) {
  if constexpr (DT == DDPB200_BF16) return round_through_bf16(x); else return x;
}

// ---- vector groups -------------------------------------------------------------------
// A "group" is G contiguous elements owned by one thread; G is chosen per kernel so that the
// widest stream moves 16 bytes per request (G=4 if any fp32 stream is present, else 8).
// Streaming cache policy: every byte is touched once per launch, so loads and stores carry
// the evict-first (.cs) hint and do not displace each other's lines in L2.
template <int DT, int G> struct GroupIO;

template <> struct GroupIO<DDPB200_F32, 4> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[4]) {
    float4 t = __ldcs(reinterpret_cast<const float4*>(static_cast<const float*>(p) + i));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[4]) {
    __stcs(reinterpret_cast<float4*>(static_cast<float*>(p) + i), make_float4(v[0], v[1], v[2], v[3]));
  }
};

template <> struct GroupIO<DDPB200_BF16, 8> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[8]) {
    uint4 t = __ldcs(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p) + i));
    v[0] = bf16_bits_to_f32(t.x & 0xffffu); v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = bf16_bits_to_f32(t.y & 0xffffu); v[3] = __uint_as_float(t.y & 0xffff0000u);
    v[4] = bf16_bits_to_f32(t.z & 0xffffu); v[5] = __uint_as_float(t.z & 0xffff0000u);
    v[6] = bf16_bits_to_f32(t.w & 0xffffu); v[7] = __uint_as_float(t.w & 0xffff0000u);
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[8]) {
    uint4 t;
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 c = __floats2bfloat162_rn(v[4], v[5]);
    __nv_bfloat162 d = __floats2bfloat162_rn(v[6], v[7]);
    t.x = *reinterpret_cast<uint32_t*>(&a); t.y = *reinterpret_cast<uint32_t*>(&b);
    t.z = *reinterpret_cast<uint32_t*>(&c); t.w = *reinterpret_cast<uint32_t*>(&d);
    __stcs(reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p) + i), t);
  }
};

template <> struct GroupIO<DDPB200_BF16, 4> {  // bf16 stream beside an fp32 one: 8-byte requests
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[4]) {
    uint2 t = __ldcs(reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(p) + i));
    v[0] = bf16_bits_to_f32(t.x & 0xffffu); v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = bf16_bits_to_f32(t.y & 0xffffu); v[3] = __uint_as_float(t.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[4]) {
    uint2 t;
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
    t.x = *reinterpret_cast<uint32_t*>(&a); t.y = *reinterpret_cast<uint32_t*>(&b);
    __stcs(reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p) + i), t);
  }
};

// scalar element access (misaligned segments and tails)
template <int DT> __device__ __forceinline__ float load1(const void* p, int64_t i) {
  if constexpr (DT == DDPB200_F32) return static_cast<const float*>(p)[i];
  else return __bfloat162float(static_cast<const __nv_bfloat16*>(p)[i]);
}
template <int DT> __device__ __forceinline__ void store1(void* p, int64_t i, float x) {
  if constexpr (DT == DDPB200_F32) static_cast<float*>(p)[i] = x;
  else static_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(x);
}

template <int DT, int G> __device__ __forceinline__ bool group_aligned(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) % (G * ElemOf<DT>::size)) == 0;
}

// ---- reductions ------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}
__device__ __forceinline__ double warp_sum(double x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// Sum over the CTA; result valid in thread 0. Fixed tree: lanes by xor-shuffle, then warps
// in index order, so the value depends only on the inputs (bitwise reproducible).
__device__ __forceinline__ float block_sum(float x, float* smem /* kThreads/32 floats */) {
  x = warp_sum(x);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // smem may still be read from a previous call
  if (lane == 0) smem[warp] = x;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) r += smem[w];
  }
  return r;
}

// "Last CTA done" hand-off: every CTA publishes its chunk partials, takes a ticket, and the
// CTA that draws the last ticket folds the partials per segment in a fixed order. One launch,
// no float atomics, no second kernel. The ticket counter is left at zero for the next launch.
template <class Tab>
__device__ __forceinline__ void finish_sumsq(const Tab& tab, int64_t chunk_base,
                                             const float* chunk_partials, float* sumsq_out,
                                             int64_t total_chunks, uint32_t* ticket) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int s = warp; s < tab.n; s += kThreads / 32) {
    const int64_t b = tab.begin(s) - chunk_base;
    const int64_t e = (s + 1 < tab.n) ? tab.begin(s + 1) - chunk_base : total_chunks;
    double acc = 0.0;
    for (int64_t c = b + lane; c < e; c += 32) acc += (double)__ldcg(chunk_partials + c);
    acc = warp_sum(acc);
    if (lane == 0) sumsq_out[s] = (float)acc;
  }
  if (threadIdx.x == 0) *ticket = 0u;
}

}  // namespace ddpb200

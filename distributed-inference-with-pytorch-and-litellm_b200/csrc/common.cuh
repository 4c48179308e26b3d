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
constexpr int kTileElems = 2048;  // granularity of chunk sizes (elements)
constexpr int kCtasPerSm = 3;     // persistent grid = SMs * kCtasPerSm (measured: 2-4 is the plateau, 8 loses 15 %)
constexpr int kStreamCtasPerSm = 4;  // pack / norm: one or two streams, few registers: 4 CTAs/SM keep more loads in flight
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

// A CTA's chunks come c, c+grid, c+2*grid, ...: with chunks much shorter than the large tensors, the next
// chunk is usually in the segment of the last one (or the following segment), so the binary search is
// only taken when this cheap test fails. cur.s < 0 = nothing cached yet.
struct SegCursor { int s = -1; int64_t b = 0, e = 0; };   // segment s covers absolute chunks [b, e)
template <class Tab>
__device__ __forceinline__ int find_seg_cached(const Tab& tab, int64_t c_abs, int64_t end_abs, SegCursor& cur) {
  if (cur.s < 0 || c_abs < cur.b || c_abs >= cur.e) {
    cur.s = find_seg(tab, c_abs);
    cur.b = tab.begin(cur.s);
    cur.e = (cur.s + 1 < tab.n) ? tab.begin(cur.s + 1) : end_abs;
  }
  return cur.s;
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
// A "group" is G contiguous elements owned by one thread. sm_100a has 256-bit global loads and
// stores (LDG.E.256 / STG.E.256), which is also the only width that takes an L2 eviction hint:
// every byte here is touched once per launch, so loads and stores are marked no-allocate in L1
// and evict-first in L2 and do not displace each other's lines. A kernel picks G from the widest
// stream (32 bytes of it per group); narrower streams of the same group use 128-bit accesses.
// Pointers in the segment tables are generic; they are converted to the global window once.
template <int W> struct Raw { uint32_t r[W]; };   // W 32-bit words

__device__ __forceinline__ Raw<8> ldg256(const void* p) {
  Raw<8> x;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]), "=r"(x.r[4]), "=r"(x.r[5]),
                 "=r"(x.r[6]), "=r"(x.r[7]) : "l"(__cvta_generic_to_global(p)));
  return x;
}
__device__ __forceinline__ void stg256(void* p, const Raw<8>& x) {
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(__cvta_generic_to_global(p)), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]),
                  "r"(x.r[4]), "r"(x.r[5]), "r"(x.r[6]), "r"(x.r[7]) : "memory");
}
__device__ __forceinline__ Raw<4> ldg128(const void* p) {
  Raw<4> x;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "l"(__cvta_generic_to_global(p)));
  return x;
}
__device__ __forceinline__ void stg128(void* p, const Raw<4>& x) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(__cvta_generic_to_global(p)), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]) : "memory");
}
__device__ __forceinline__ Raw<2> ldg64(const void* p) {
  Raw<2> x;
  asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(x.r[0]), "=r"(x.r[1]) : "l"(__cvta_generic_to_global(p)));
  return x;
}
__device__ __forceinline__ void stg64(void* p, const Raw<2>& x) {
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};"
               :: "l"(__cvta_generic_to_global(p)), "r"(x.r[0]), "r"(x.r[1]) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {   // a -> low half
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

// GroupIO<DT, G>: load G elements at element index i of a stream into fp32 registers / store back.
template <int DT, int G> struct GroupIO;

template <> struct GroupIO<DDPB200_F32, 8> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[8]) {
    Raw<8> t = ldg256(static_cast<const float*>(p) + i);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(t.r[k]);
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[8]) {
    Raw<8> t;
#pragma unroll
    for (int k = 0; k < 8; ++k) t.r[k] = __float_as_uint(v[k]);
    stg256(static_cast<float*>(p) + i, t);
  }
};
template <> struct GroupIO<DDPB200_F32, 4> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[4]) {
    Raw<4> t = ldg128(static_cast<const float*>(p) + i);
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __uint_as_float(t.r[k]);
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[4]) {
    Raw<4> t;
#pragma unroll
    for (int k = 0; k < 4; ++k) t.r[k] = __float_as_uint(v[k]);
    stg128(static_cast<float*>(p) + i, t);
  }
};
template <> struct GroupIO<DDPB200_BF16, 16> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[16]) {
    Raw<8> t = ldg256(static_cast<const __nv_bfloat16*>(p) + i);
#pragma unroll
    for (int k = 0; k < 8; ++k) { v[2 * k] = bf16_bits_to_f32(t.r[k] & 0xffffu); v[2 * k + 1] = __uint_as_float(t.r[k] & 0xffff0000u); }
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[16]) {
    Raw<8> t;
#pragma unroll
    for (int k = 0; k < 8; ++k) t.r[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
    stg256(static_cast<__nv_bfloat16*>(p) + i, t);
  }
};
template <> struct GroupIO<DDPB200_BF16, 8> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[8]) {
    Raw<4> t = ldg128(static_cast<const __nv_bfloat16*>(p) + i);
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[2 * k] = bf16_bits_to_f32(t.r[k] & 0xffffu); v[2 * k + 1] = __uint_as_float(t.r[k] & 0xffff0000u); }
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[8]) {
    Raw<4> t;
#pragma unroll
    for (int k = 0; k < 4; ++k) t.r[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
    stg128(static_cast<__nv_bfloat16*>(p) + i, t);
  }
};
template <> struct GroupIO<DDPB200_BF16, 4> {
  static __device__ __forceinline__ void load(const void* p, int64_t i, float (&v)[4]) {
    Raw<2> t = ldg64(static_cast<const __nv_bfloat16*>(p) + i);
#pragma unroll
    for (int k = 0; k < 2; ++k) { v[2 * k] = bf16_bits_to_f32(t.r[k] & 0xffffu); v[2 * k + 1] = __uint_as_float(t.r[k] & 0xffff0000u); }
  }
  static __device__ __forceinline__ void store(void* p, int64_t i, const float (&v)[4]) {
    Raw<2> t;
#pragma unroll
    for (int k = 0; k < 2; ++k) t.r[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
    stg64(static_cast<__nv_bfloat16*>(p) + i, t);
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

// Elements per group when the widest stream moves VECB bytes per access (VECB = 32 or 16).
template <int VECB, bool ANY_F32> struct GroupSize { static constexpr int G = VECB / (ANY_F32 ? 4 : 2); };

// x / c with IEEE round-to-nearest, for a divisor that is constant over the launch: reciprocal,
// residual, correction (Markstein). rc must be the correctly rounded 1/c. Bit-identical to
// __fdiv_rn(x, c) for finite normal quotients (checked exhaustively over all 2^23 significands for
// 492 bias-correction constants in tests/test_oracle_golden.py::test_division_by_constant); the
// select keeps x = +inf -> +inf. 3 issue slots instead of ~10.
__device__ __forceinline__ float div_by_const(float x, float c, float rc) {
  const float q0 = __fmul_rn(x, rc);
  const float r = __fmaf_rn(-q0, c, x);
  const float q = __fmaf_rn(r, rc, q0);
  return (fabsf(q0) == __int_as_float(0x7f800000)) ? q0 : q;
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

// Kernel (c), bulk-copy variant: the fused clip+AdamW pass with its four input streams staged through
// shared memory by the TMA engine (cp.async.bulk global -> shared, completion on an mbarrier) in a ring of
// NSTAGE tiles, instead of LDG into registers. Every copy reads this GPU's own memory, unicast.
//
// Why it exists: the register version (adamw.cu) is latency-bound — 64 registers x 256 threads x 4 CTAs fill
// the register file, every warp alternates "four loads in flight" with "compute + store", and the bytes in
// flight per SM therefore depend on how fast the SM gets through the compute phase: 4 CTAs/SM win under the
// power cap (1.7 GHz), 3 CTAs/SM win at 1.965 GHz (profiles/r02_adamw_kbench.txt). Here one producer lane
// per CTA keeps NSTAGE tiles (NSTAGE x 16 KB) in flight whatever the eight consumer warps are doing; loads
// cost no LSU issue slots and no registers; stores still go straight from registers (STG.128).
//
// All-bf16 or all-fp32, fast modes only (16 bytes of every stream per thread and tile; the arithmetic is
// adamw_group_packed of adamw_math.cuh, bit-identical to adamw.cu); a chunk whose four pointers are not 16-byte aligned, and the ragged last elements of a segment, are
// read from global memory by the consumers as in adamw.cu. Selected at run time (see launch_adamw_bulk).
#include <stdlib.h>
#include "launch.cuh"
#include "adamw_math.cuh"

namespace ddpb200 {

constexpr int kBulkConsumers = kThreads;              // 8 consumer warps, 16 bytes of each stream per thread and tile
constexpr int kBulkStreamBytes = kBulkConsumers * 16; // one stream's tile: 4 KB = 2048 bf16 / 1024 fp32 elements
constexpr int kBulkThreads = kThreads + 32;           // + 1 producer warp (one lane works)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.shared::cta.b64 st, [%0];\n}" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}" :: "r"(bar), "r"(parity) : "memory");
}
// 1-D bulk copy global -> this CTA's shared memory (no multicast), completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(dst), "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ Raw<4> lds128(uint32_t addr) {
  Raw<4> x;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "r"(addr));
  return x;
}
template <int DT> __device__ __forceinline__ void unpack16(const Raw<4>& t, float (&v)[16 / ElemOf<DT>::size]) {
  if constexpr (DT == DDPB200_F32) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __uint_as_float(t.r[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[2 * k] = bf16_bits_to_f32(t.r[k] & 0xffffu); v[2 * k + 1] = __uint_as_float(t.r[k] & 0xffff0000u); }
  }
}

// staged: p, m and v are 16-byte aligned and go through shared memory; g_staged: so does the gradient. A gradient
// that alone is misaligned (a bucket view behind an odd-sized tensor of the same bucket) is read from global memory
// element by element by the consumers while the other three streams stay staged.
struct BulkChunk { char *P, *G, *M, *V; int64_t len; bool staged, g_staged; };

template <int ES, class Tab>
__device__ __forceinline__ BulkChunk bulk_chunk(const Tab& tab, int64_t c, int64_t chunk_base, int64_t total_chunks,
                                                int64_t chunk_elems, SegCursor& cur) {
  const int s = find_seg_cached(tab, c + chunk_base, chunk_base + total_chunks, cur);
  const int64_t start = (c + chunk_base - cur.b) * chunk_elems;
  const int64_t rest = tab.numel(s) - start;
  BulkChunk k;
  k.len = rest < chunk_elems ? rest : chunk_elems;
  k.P = static_cast<char*>(tab.ptr(s, 0)) + start * ES;
  k.G = static_cast<char*>(tab.ptr(s, 1)) + start * ES;
  k.M = static_cast<char*>(tab.ptr(s, 2)) + start * ES;
  k.V = static_cast<char*>(tab.ptr(s, 3)) + start * ES;
  k.staged = ((reinterpret_cast<uintptr_t>(k.P) | reinterpret_cast<uintptr_t>(k.M) | reinterpret_cast<uintptr_t>(k.V)) & 15) == 0;
  k.g_staged = k.staged && (reinterpret_cast<uintptr_t>(k.G) & 15) == 0;
  return k;
}

template <class Tab, int DT, int MODE, int NSTAGE, int CT>
__global__ void __launch_bounds__(kBulkThreads, CT)
adamw_bulk_kernel(const __grid_constant__ Tab tab, int64_t total_chunks, int64_t chunk_elems, AdamScalars a,
                  const float* __restrict__ lr_ptr, const float* __restrict__ clip_ptr, uint32_t flags) {
  constexpr int ES = ElemOf<DT>::size;
  constexpr int G = 16 / ES;                                        // elements per thread and tile
  constexpr int kBulkTile = kBulkConsumers * G;                     // elements per tile and stream
  constexpr uint32_t kStreamBytes = kBulkStreamBytes;
  constexpr uint32_t kStageBytes = 4 * kStreamBytes;
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t data = smem_u32(smem);
  const uint32_t full = data + NSTAGE * kStageBytes, empty = full + NSTAGE * 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(full + 8 * s, 1); mbar_init(empty + 8 * s, kBulkConsumers / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int64_t chunk_base = tab.begin(0);
  SegCursor cur;
  int stage = 0;
  uint32_t phase = 0;

  if (warp == kBulkConsumers / 32) {
    // ---- producer: one lane walks the same tile sequence as the consumers, NSTAGE tiles ahead -------------
    if (lane != 0) return;
    for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
      const BulkChunk k = bulk_chunk<ES>(tab, c, chunk_base, total_chunks, chunk_elems, cur);
      for (int64_t t0 = 0; t0 < k.len; t0 += kBulkTile) {
        const int64_t cnt = (k.len - t0) < kBulkTile ? (k.len - t0) : kBulkTile;
        const uint32_t bytes = k.staged ? (uint32_t)(cnt / G * G) * ES : 0u;
        mbar_wait(empty + 8 * stage, phase ^ 1);                    // the consumers are done with this slot
        const uint32_t bar = full + 8 * stage, dst = data + stage * kStageBytes;
        if (bytes) {
          mbar_arrive_expect_tx(bar, (k.g_staged ? 4 : 3) * bytes);
          if (k.g_staged) bulk_g2s(dst, k.G + t0 * ES, bytes, bar);
          bulk_g2s(dst + kStreamBytes, k.P + t0 * ES, bytes, bar);
          bulk_g2s(dst + 2 * kStreamBytes, k.M + t0 * ES, bytes, bar);
          bulk_g2s(dst + 3 * kStreamBytes, k.V + t0 * ES, bytes, bar);
        } else {
          mbar_arrive(bar);                                         // nothing staged: the consumers read global memory
        }
        if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
      }
    }
    return;
  }

  // ---- consumers ---------------------------------------------------------------------------------------
  const AdamRuntime r = make_adam_runtime(a, lr_ptr, clip_ptr, nullptr, flags);
  const AdamPacked pk = make_adam_packed(a, r);
  for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const BulkChunk k = bulk_chunk<ES>(tab, c, chunk_base, total_chunks, chunk_elems, cur);
    for (int64_t t0 = 0; t0 < k.len; t0 += kBulkTile) {
      const int64_t cnt = (k.len - t0) < kBulkTile ? (k.len - t0) : kBulkTile;
      const int64_t vec = k.staged ? cnt / G * G : 0;
      mbar_wait(full + 8 * stage, phase);                           // the tile has landed
      const int64_t i = (int64_t)threadIdx.x * G;
      float p[G], g[G], m[G], v[G];
      const bool mine = i < vec;
      if (mine) {
        const uint32_t src = data + stage * kStageBytes + threadIdx.x * 16;
        if (k.g_staged) {
          unpack16<DT>(lds128(src), g);
        } else {
#pragma unroll
          for (int e = 0; e < G; ++e) g[e] = load1<DT>(k.G, t0 + i + e);
        }
        unpack16<DT>(lds128(src + kStreamBytes), p);
        unpack16<DT>(lds128(src + 2 * kStreamBytes), m);
        unpack16<DT>(lds128(src + 3 * kStreamBytes), v);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + 8 * stage);                // values are in registers: hand the slot back
      if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
      if (mine) {
        adamw_group_packed<DT, MODE, G>(p, g, m, v, a, r, pk);
        GroupIO<DT, G>::store(k.P, t0 + i, p);
        GroupIO<DT, G>::store(k.M, t0 + i, m);
        GroupIO<DT, G>::store(k.V, t0 + i, v);
      }
      // the ragged end of a segment, or a chunk that could not be staged: element by element from global memory
      for (int64_t j = t0 + vec + threadIdx.x; j < t0 + cnt; j += kBulkConsumers) {
        float pe = load1<DT>(k.P, j), ge = load1<DT>(k.G, j), me = load1<DT>(k.M, j), ve = load1<DT>(k.V, j);
        adamw_elem<DT, MODE>(pe, ge, me, ve, a, r);
        store1<DT>(k.P, j, pe);
        store1<DT>(k.M, j, me);
        store1<DT>(k.V, j, ve);
      }
    }
  }
}

// 0 = not applicable (the caller falls through to the register kernel), else launched (or failed): *rc holds the code
template <int DT, int MODE>
int launch_adamw_bulk(const ddpb200_seg* table, int table_loc, int64_t n, int64_t total_chunks, int64_t chunk_elems,
                      const AdamScalars& a, const float* lr_ptr, const float* clip, uint32_t flags, cudaStream_t st,
                      int stages, int ctas, int* rc) {
  if (table_loc != DDPB200_TABLE_DEVICE) return 0;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) { *rc = err; return 1; }
  DevTable tab{table, (int)n};
  const int64_t cap = (int64_t)sms * ctas;
  const int grid = (int)(total_chunks < cap ? total_chunks : cap);
  if (grid == 0) { *rc = 0; return 1; }
#define DDPB200_BULK_LAUNCH(NS, CTV)                                                                       \
  {                                                                                                       \
    auto kern = adamw_bulk_kernel<DevTable, DT, MODE, NS, CTV>;                                           \
    const int smem = NS * 4 * kBulkStreamBytes + 2 * NS * 8;                                              \
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);        \
    if (e != cudaSuccess) { *rc = (int)e; return 1; }                                                     \
    kern<<<grid, kBulkThreads, smem, st>>>(tab, total_chunks, chunk_elems, a, lr_ptr, clip, flags);        \
    *rc = (int)cudaGetLastError();                                                                        \
    return 1;                                                                                             \
  }
#define DDPB200_BULK_STAGES(CTV)                                                                            \
  { if (stages == 1) DDPB200_BULK_LAUNCH(1, CTV) if (stages == 2) DDPB200_BULK_LAUNCH(2, CTV)                \
    if (stages == 3) DDPB200_BULK_LAUNCH(3, CTV) if (stages == 4) DDPB200_BULK_LAUNCH(4, CTV)                \
    if (stages == 5 && CTV < 3) DDPB200_BULK_LAUNCH(5, CTV) if (stages == 6 && CTV < 3) DDPB200_BULK_LAUNCH(6, CTV) }
  if (ctas == 1) DDPB200_BULK_STAGES(1)
  if (ctas == 2) DDPB200_BULK_STAGES(2)
  if (ctas == 3) DDPB200_BULK_STAGES(3)
#undef DDPB200_BULK_STAGES
#undef DDPB200_BULK_LAUNCH
  return 0;
}

#define DDPB200_BULK_INST(DT, MODE)                                                                                  \
  template int launch_adamw_bulk<DT, MODE>(const ddpb200_seg*, int, int64_t, int64_t, int64_t, const AdamScalars&,   \
                                           const float*, const float*, uint32_t, cudaStream_t, int, int, int*);
DDPB200_BULK_INST(DDPB200_BF16, kModeFastClip)
DDPB200_BULK_INST(DDPB200_BF16, kModeFastPlain)
DDPB200_BULK_INST(DDPB200_F32, kModeFastClip)
DDPB200_BULK_INST(DDPB200_F32, kModeFastPlain)
#undef DDPB200_BULK_INST

}  // namespace ddpb200

// Bucket allreduce fused with the gradient-norm pass, over NVLink peer memory / NVSwitch multicast.
//
// Replaces, for the bucket arena of one dtype, the pair  ncclAllReduce(SUM)  (c10d Reducer,
// TORCH/include/torch/csrc/distributed/c10d/reducer.hpp:524)  +  torch._foreach_norm
// (TORCH/nn/utils/clip_grad.py:96)  by ONE kernel per rank that moves every gradient byte once:
//
//   chunk c belongs to rank c % world. The owner reads the chunk from EVERY rank's arena
//     - peer mode:      one 256-bit load per rank over NVLink, summed in fp32 in rank order 0..n-1
//     - multicast mode: multimem.ld_reduce (the NVSwitch adds the n copies, fp32 accumulation)
//   rounds once to the bucket dtype, accumulates the squares of what it stores, and writes the
//   result into every rank's arena (n peer stores, or one multimem.st that the switch replicates).
//   The chunk's sum of squares goes to every rank's partials array, so that after a barrier each
//   rank folds the same numbers in the same order (ddpb200_fold_chunk_partials): norms are
//   bitwise identical on all ranks, and so is every averaged gradient (one owner computes it).
//
// The 1/world_size scale of the pack pass can be folded in (scale != 1): the caller then skips the
// in-place scale of gradients that already live in the arena (gradient_as_bucket_view). Peer mode and the
// scalar ends apply it per rank BEFORE the sum, round(fp32(g) * scale) to the bucket dtype, i.e. exactly
// what the pack kernel would have stored — same bits as pack + allreduce for any world size. Multicast
// mode can only scale the switch's sum afterwards; for a power-of-two scale that is the same number
// (scaling by 2^-k commutes with every rounding) except where a scaled bf16/fp32 value would be
// subnormal, so the host passes scale != 1 in multicast mode only for power-of-two world sizes.
//
// NVLink bytes per GPU and direction for S bytes of gradients: peer mode 2(n-1)/n * S (as a ring),
// multicast mode (1 + 1/n) * S. Multicast wins from 4 ranks up; at 2 ranks peer mode moves less.
// The arena must be symmetric memory (same offset on every rank); the caller provides the barriers
// (all ranks packed before this kernel; all ranks finished it before anyone reads the arena).
#include <stdlib.h>
#include "launch.cuh"
#include "peer.cuh"

namespace ddpb200 {

template <int DT> __device__ __forceinline__ void unpack128(const Raw<4>& t, float (&v)[16 / ElemOf<DT>::size]) {
  if constexpr (DT == DDPB200_F32) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __uint_as_float(t.r[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[2 * k] = bf16_bits_to_f32(t.r[k] & 0xffffu); v[2 * k + 1] = __uint_as_float(t.r[k] & 0xffff0000u); }
  }
}

// MC: 1 = multicast (multimem) for the vector body, 0 = peer loads/stores. W: compile-time bound on
// the world size (2, 4 or 8), so that a 2-rank job does not carry registers for 8 peers.
// SCATTER: the owner keeps the reduced chunk to itself (local store only): the reduce-scatter half, for
// the sharded update (sharded_update.cu) which sends updated PARAMETERS back instead of gradients.
// A chunk's owner is (absolute chunk id) % world, so ownership does not depend on how a step slices
// the table into launches.
template <int DT, int MC, int W, bool SCATTER, int UNROLL = 4>
__global__ void __launch_bounds__(kThreads, (MC || W <= 2) ? 3 : 2)
fused_allreduce_kernel(DevTable tab, int64_t total_chunks, int64_t chunk_elems, const __grid_constant__ PeerMap pm,
                       float scale) {
  constexpr int ES = ElemOf<DT>::size;
  constexpr int G = 32 / ES;            // 256-bit groups in peer mode
  constexpr int H = 16 / ES;            // 128-bit groups in multicast mode
  __shared__ float red[kThreads / 32];
  const int64_t chunk_base = tab.begin(0);
  const int world = pm.world;
  char* const local = pm.arena[pm.rank];
  const bool scaled = scale != 1.0f;
  constexpr int kAlign = MC ? 16 : 32;    // bytes: the vector body needs the arena offset aligned to its access width

  // element by element through the peer mappings: the head that brings a misaligned segment onto the
  // vector body's alignment (the offset is the same on every rank, so all copies align together), and
  // the ragged end
  auto scalar_range = [&](int64_t off, int64_t b, int64_t e, float& acc0) {
    for (int64_t i = b + threadIdx.x; i < e; i += kThreads) {
      float sum = 0.f;
      for (int p = 0; p < world; ++p) {
        float x = load1<DT>(pm.arena[p] + off, i);
        if (scaled) x = round_through<DT>(__fmul_rn(x, scale));
        sum = __fadd_rn(sum, x);
      }
      sum = round_through<DT>(sum);
      acc0 = __fmaf_rn(sum, sum, acc0);
      if constexpr (SCATTER) store1<DT>(local + off, i, sum);
      else for (int p = 0; p < world; ++p) store1<DT>(pm.arena[p] + off, i, sum);
    }
  };

  const int64_t c0 = ((pm.rank - chunk_base % world) % world + world) % world;   // first relative chunk this rank owns
  for (int64_t c = c0 + (int64_t)world * blockIdx.x; c < total_chunks; c += (int64_t)world * gridDim.x) {
    const int s = find_seg(tab, c + chunk_base);
    const int64_t start = (c + chunk_base - tab.begin(s)) * chunk_elems;
    const int64_t rest = tab.numel(s) - start;
    const int64_t len = rest < chunk_elems ? rest : chunk_elems;
    const int64_t off = (static_cast<char*>(tab.ptr(s, 1)) - local) + start * ES;     // byte offset in the arena
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int64_t mis = off % kAlign;
    const int64_t head = mis ? ((kAlign - mis) / ES < len ? (kAlign - mis) / ES : len) : 0;
    if (head) scalar_range(off, 0, head, acc[0]);
    int64_t done = head;

    if constexpr (MC != 0) {
      const int64_t groups_end = head + (len - head) / H * H;
#pragma unroll UNROLL
      for (int64_t i = head + (int64_t)threadIdx.x * H; i < groups_end; i += kThreads * H) {
        Raw<4> t;
        if constexpr (DT == DDPB200_F32) t = mc_ld_reduce_f32x4(pm.mc + off + i * ES);
        else t = mc_ld_reduce_bf16x8(pm.mc + off + i * ES);
        float v[H];
        unpack128<DT>(t, v);
        if (scaled) {                                       // power-of-two scale of the switch's sum (see the header)
#pragma unroll
          for (int k = 0; k < H; ++k) v[k] = __fmul_rn(v[k], scale);
          if constexpr (DT == DDPB200_F32) {
#pragma unroll
            for (int k = 0; k < 4; ++k) t.r[k] = __float_as_uint(v[k]);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) t.r[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
            unpack128<DT>(t, v);
          }
        }
#pragma unroll
        for (int k = 0; k < H; ++k) acc[k & 3] = __fmaf_rn(v[k], v[k], acc[k & 3]);
        if constexpr (SCATTER) stg128(local + off + i * ES, t);
        else mc_st_128(pm.mc + off + i * ES, t);
      }
      done = groups_end;
    } else {
      const int64_t groups_end = head + (len - head) / G * G;
#pragma unroll 1
      for (int64_t i = head + (int64_t)threadIdx.x * G; i < groups_end; i += kThreads * G) {
        Raw<8> raw[W];
#pragma unroll
        for (int p = 0; p < W; ++p)                                        // all loads in flight first
          if (p < world) raw[p] = ldg256(pm.arena[p] + off + i * ES);
        float sum[G];
#pragma unroll
        for (int k = 0; k < G; ++k) sum[k] = 0.f;
#pragma unroll
        for (int p = 0; p < W; ++p)
          if (p < world) {                                                 // rank order: same bits everywhere
            if constexpr (DT == DDPB200_F32) {
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                float x = __uint_as_float(raw[p].r[k]);
                if (scaled) x = __fmul_rn(x, scale);
                sum[k] = __fadd_rn(sum[k], x);
              }
            } else {
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                uint32_t w = raw[p].r[k];
                if (scaled)                                    // what the pack kernel would have stored for this rank
                  w = pack_bf16x2(__fmul_rn(bf16_bits_to_f32(w & 0xffffu), scale), __fmul_rn(__uint_as_float(w & 0xffff0000u), scale));
                sum[2 * k] = __fadd_rn(sum[2 * k], bf16_bits_to_f32(w & 0xffffu));
                sum[2 * k + 1] = __fadd_rn(sum[2 * k + 1], __uint_as_float(w & 0xffff0000u));
              }
            }
          }
        Raw<8> out;
        if constexpr (DT == DDPB200_F32) {
#pragma unroll
          for (int k = 0; k < 8; ++k) { acc[k & 3] = __fmaf_rn(sum[k], sum[k], acc[k & 3]); out.r[k] = __float_as_uint(sum[k]); }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            out.r[k] = pack_bf16x2(sum[2 * k], sum[2 * k + 1]);
            const float lo = bf16_bits_to_f32(out.r[k] & 0xffffu), hi = __uint_as_float(out.r[k] & 0xffff0000u);
            acc[(2 * k) & 3] = __fmaf_rn(lo, lo, acc[(2 * k) & 3]);
            acc[(2 * k + 1) & 3] = __fmaf_rn(hi, hi, acc[(2 * k + 1) & 3]);
          }
        }
        if constexpr (SCATTER) {
          stg256(local + off + i * ES, out);
        } else {
#pragma unroll
          for (int p = 0; p < W; ++p)
            if (p < world) stg256(pm.arena[p] + off + i * ES, out);
        }
      }
      done = groups_end;
    }
    if (done < len) scalar_range(off, done, len, acc[0]);     // ragged end

    const float total = block_sum(__fadd_rn(__fadd_rn(acc[0], acc[1]), __fadd_rn(acc[2], acc[3])), red);
    if (threadIdx.x == 0)
      for (int p = 0; p < world; ++p) pm.partials[p][c] = total;
  }
  __threadfence_system();               // peer stores are out before the kernel is reported complete
}

// Per-segment fold of the chunk partials, fixed order (the tail of finish_sumsq as its own launch:
// here the partials come from other GPUs, so a cross-GPU barrier sits between the two kernels).
__global__ void __launch_bounds__(kThreads, 1)
fold_partials_kernel(DevTable tab, int64_t total_chunks, const float* __restrict__ partials,
                     float* __restrict__ sumsq_out) {
  const int64_t chunk_base = tab.begin(0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (kThreads / 32) * gridDim.x;
  for (int s = blockIdx.x * (kThreads / 32) + warp; s < tab.n; s += nwarps) {
    const int64_t b = tab.begin(s) - chunk_base;
    const int64_t e = (s + 1 < tab.n) ? tab.begin(s + 1) - chunk_base : total_chunks;
    double acc = 0.0;
    for (int64_t c = b + lane; c < e; c += 32) acc += (double)__ldcg(partials + c);
    acc = warp_sum(acc);
    if (lane == 0) sumsq_out[s] = (float)acc;
  }
}

}  // namespace ddpb200

namespace ddpb200 {

template <bool SCATTER>
static int launch_fused_allreduce(const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks,
                                  int64_t chunk_elems, int32_t dtype, const ddpb200_peer_map* peers,
                                  int32_t use_multicast, float scale, void* stream) {
  if (!valid_chunking(n, total_chunks, chunk_elems) || !dev_table || !peers || !(scale > 0.f)) return DDPB200_E_INVALID;
  PeerMap pm;
  if (!make_peer_map(peers, true, &pm)) return DDPB200_E_INVALID;
  if (use_multicast && !pm.mc) return DDPB200_E_INVALID;
  if (n == 0 || total_chunks == 0) return 0;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) return err;
  const int64_t mine = (total_chunks + pm.world - 1) / pm.world;                // chunks this rank owns (upper bound)
  // tuning knobs, read at every launch (a getenv costs ~100 ns against a multi-millisecond kernel) so that
  // tools/arbench.py can sweep them inside one process
  const int env_ctas = [] { const char* e = getenv("DDPB200_AR_CTAS"); return e ? atoi(e) : 0; }();
  // measured at 2 ranks (profiles/r01_arbench.txt): peer mode is NVLink-bound from 1 CTA/SM (3.71 ms vs 3.79 at 2)
  const int ctas = env_ctas > 0 ? env_ctas : (use_multicast ? 2 : 1);
  const int64_t cap = (int64_t)sms * ctas;
  const int grid = (int)(mine < cap ? (mine > 0 ? mine : 1) : cap);
  DevTable tab{dev_table, (int)n};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int env_unroll = [] { const char* e = getenv("DDPB200_AR_UNROLL"); return e ? atoi(e) : 0; }();
#define DDPB200_AR_LAUNCH(DT, MCV, WV)                                                                \
  fused_allreduce_kernel<DT, MCV, WV, SCATTER><<<grid, kThreads, 0, st>>>(tab, total_chunks, chunk_elems, pm, scale)
#define DDPB200_AR_CASE(DT)                                                                          \
  if (dtype == DT) {                                                                                 \
    if (use_multicast && env_unroll == 8)                                                            \
      fused_allreduce_kernel<DT, 1, 2, SCATTER, 8><<<grid, kThreads, 0, st>>>(tab, total_chunks, chunk_elems, pm, scale); \
    else if (use_multicast && env_unroll == 2)                                                       \
      fused_allreduce_kernel<DT, 1, 2, SCATTER, 2><<<grid, kThreads, 0, st>>>(tab, total_chunks, chunk_elems, pm, scale); \
    else if (use_multicast) DDPB200_AR_LAUNCH(DT, 1, 2);                                             \
    else if (pm.world <= 2) DDPB200_AR_LAUNCH(DT, 0, 2);                                             \
    else if (pm.world <= 4) DDPB200_AR_LAUNCH(DT, 0, 4);                                             \
    else DDPB200_AR_LAUNCH(DT, 0, 8);                                                                \
    return (int)cudaGetLastError();                                                                  \
  }
  DDPB200_AR_CASE(DDPB200_BF16)
  DDPB200_AR_CASE(DDPB200_F32)
#undef DDPB200_AR_CASE
#undef DDPB200_AR_LAUNCH
  return DDPB200_E_UNSUPPORTED;
}

}  // namespace ddpb200

using namespace ddpb200;

extern "C" DDPB200_API int ddpb200_allreduce_sumsq_chunks(const ddpb200_seg* dev_table, int64_t n,
                                                          int64_t total_chunks, int64_t chunk_elems,
                                                          int32_t dtype, const ddpb200_peer_map* peers,
                                                          int32_t use_multicast, float scale, void* stream) {
  return launch_fused_allreduce<false>(dev_table, n, total_chunks, chunk_elems, dtype, peers, use_multicast, scale, stream);
}

extern "C" DDPB200_API int ddpb200_reduce_scatter_sumsq_chunks(const ddpb200_seg* dev_table, int64_t n,
                                                               int64_t total_chunks, int64_t chunk_elems,
                                                               int32_t dtype, const ddpb200_peer_map* peers,
                                                               int32_t use_multicast, float scale, void* stream) {
  return launch_fused_allreduce<true>(dev_table, n, total_chunks, chunk_elems, dtype, peers, use_multicast, scale, stream);
}

extern "C" DDPB200_API int ddpb200_fold_chunk_partials(const ddpb200_seg* dev_table, int64_t n,
                                                       int64_t total_chunks, const float* chunk_partials,
                                                       float* sumsq_out, void* stream) {
  if (n < 0 || n > INT32_MAX || total_chunks < 0 || (n > 0 && (!dev_table || !sumsq_out)) ||
      (total_chunks > 0 && !chunk_partials))
    return DDPB200_E_INVALID;
  if (n == 0) return 0;
  DevTable tab{dev_table, (int)n};
  const int grid = (int)((n + 7) / 8 < 64 ? (n + 7) / 8 : 64);
  fold_partials_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(tab, total_chunks, chunk_partials, sumsq_out);
  return (int)cudaGetLastError();
}

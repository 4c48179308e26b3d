// Sharded update (ZeRO-1 style; SURVEY.md §8e "natural extension", §8f-4): the second half of
//   reduce-scatter  ->  global norm  ->  AdamW on the owned shard  ->  all-gather of parameters
// as ONE kernel per rank over NVLink peer memory. The first half is fused_allreduce.cu in SCATTER mode
// (the owner of chunk c = c % world keeps the reduced gradients to itself and every rank gets the
// chunk's sum of squares). Here the owner of a chunk reads its reduced gradients, its parameters and
// its slice of exp_avg / exp_avg_sq from LOCAL memory, applies exactly the arithmetic of adamw.cu, and
// stores the new parameters into EVERY rank's parameter arena (peer stores, or one multimem.st that
// the NVSwitch replicates). Per GPU the optimizer touches 14/n bytes per parameter of HBM instead of
// 14, holds 1/n of the optimizer state, and NVLink carries the same bytes an allreduce would.
//
// Replaces, together with SCATTER mode: bucket allreduce (reducer.hpp:524) + clip_grad_norm_
// (clip_grad.py:186-233) + AdamW.step (adam.py:215-272) of the replicated path; parameters after a
// step are bit-identical to that path (same element arithmetic on the same rounded gradients).
// Parameter and gradient arenas must be symmetric memory with the SAME layout (same byte offsets).
#include <stdlib.h>
#include "launch.cuh"
#include "adamw_math.cuh"
#include "peer.cuh"

namespace ddpb200 {

template <int DT, int MC, int W, int MODE>
__global__ void __launch_bounds__(kThreads, 2)
sharded_adamw_kernel(DevTable tab, int64_t total_chunks, int64_t chunk_elems, const char* grad_local,
                     const __grid_constant__ PeerMap pp, char* m_shard, char* v_shard, AdamScalars a,
                     const float* __restrict__ lr_ptr, const float* __restrict__ clip_ptr, uint32_t flags) {
  constexpr int ES = ElemOf<DT>::size;
  // multicast stores are 128-bit: one group per lane keeps a warp's store a single contiguous 512 B
  // (two 16-byte halves of a 32-byte group per lane measured 3x slower: half-filled NVLink packets)
  constexpr int G = (MC ? 16 : 32) / ES;
  const AdamRuntime r = make_adam_runtime(a, lr_ptr, clip_ptr, nullptr, flags);
  const AdamPacked pk = make_adam_packed(a, r);
  const int64_t chunk_base = tab.begin(0);
  const int world = pp.world;
  char* const param_local = pp.arena[pp.rank];
  const int64_t c0 = ((pp.rank - chunk_base % world) % world + world) % world;

  for (int64_t c = c0 + (int64_t)world * blockIdx.x; c < total_chunks; c += (int64_t)world * gridDim.x) {
    const int s = find_seg(tab, c + chunk_base);
    const int64_t start = (c + chunk_base - tab.begin(s)) * chunk_elems;
    const int64_t rest = tab.numel(s) - start;
    const int64_t len = rest < chunk_elems ? rest : chunk_elems;
    const int64_t off = (static_cast<char*>(tab.ptr(s, 1)) - grad_local) + start * ES;   // same offset in both arenas
    const int64_t slot = ((c + chunk_base) / world) * chunk_elems * ES;                   // this chunk's place in the shard
    const char* Gp = grad_local + off;
    char* P = param_local + off;
    char* M = m_shard + slot;
    char* V = v_shard + slot;

    int64_t done = 0;
    if ((off % 32) == 0) {
      const int64_t groups_end = len / G * G;
#pragma unroll 2
      for (int64_t i = (int64_t)threadIdx.x * G; i < groups_end; i += kThreads * G) {
        float p[G], g[G], m[G], v[G];
        GroupIO<DT, G>::load(Gp, i, g);
        GroupIO<DT, G>::load(P, i, p);
        GroupIO<DT, G>::load(M, i, m);
        GroupIO<DT, G>::load(V, i, v);
        if constexpr (MODE != kModeGeneral) {
          adamw_group_packed<DT, MODE, G>(p, g, m, v, a, r, pk);
        } else {
#pragma unroll
          for (int k = 0; k < G; ++k) adamw_elem<DT, MODE>(p[k], g[k], m[k], v[k], a, r);
        }
        GroupIO<DT, G>::store(M, i, m);
        GroupIO<DT, G>::store(V, i, v);
        if constexpr (MC) {                                   // the switch replicates the store to every rank
          Raw<4> out;
          if constexpr (DT == DDPB200_F32) {
#pragma unroll
            for (int k = 0; k < 4; ++k) out.r[k] = __float_as_uint(p[k]);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) out.r[k] = pack_bf16x2(p[2 * k], p[2 * k + 1]);
          }
          mc_st_128(pp.mc + off + i * ES, out);
        } else {
          Raw<8> out;
          if constexpr (DT == DDPB200_F32) {
#pragma unroll
            for (int k = 0; k < 8; ++k) out.r[k] = __float_as_uint(p[k]);
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) out.r[k] = pack_bf16x2(p[2 * k], p[2 * k + 1]);
          }
#pragma unroll
          for (int q = 0; q < W; ++q)
            if (q < world) stg256(pp.arena[q] + off + i * ES, out);
        }
      }
      done = groups_end;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
      float p = load1<DT>(P, i), g = load1<DT>(Gp, i), m = load1<DT>(M, i), v = load1<DT>(V, i);
      adamw_elem<DT, MODE>(p, g, m, v, a, r);
      store1<DT>(M, i, m);
      store1<DT>(V, i, v);
      for (int q = 0; q < world; ++q) store1<DT>(pp.arena[q] + off, i, p);
    }
  }
  __threadfence_system();
}

}  // namespace ddpb200

using namespace ddpb200;

extern "C" DDPB200_API int ddpb200_sharded_adamw_allgather(
    const ddpb200_seg* dev_table, int64_t n, int64_t total_chunks, int64_t chunk_elems, int32_t dtype,
    const void* grad_arena_local, const ddpb200_peer_map* param_peers, void* exp_avg_shard,
    void* exp_avg_sq_shard, double lr, const float* lr_ptr, double beta1, double beta2, double eps,
    double weight_decay, int64_t step, const float* clip_coef, uint32_t flags, int32_t use_multicast,
    void* stream) {
  if (!valid_chunking(n, total_chunks, chunk_elems) || !dev_table || !grad_arena_local || !exp_avg_shard ||
      !exp_avg_sq_shard || step < 1)
    return DDPB200_E_INVALID;
  PeerMap pp;
  if (!make_peer_map(param_peers, false, &pp)) return DDPB200_E_INVALID;
  if (use_multicast && !pp.mc) return DDPB200_E_INVALID;
  if (n == 0 || total_chunks == 0) return 0;
  float bc[2];
  ddpb200_adamw_bias_corrections(beta1, beta2, step, bc);
  AdamScalars a = make_adam_scalars(lr, beta1, beta2, eps, weight_decay, bc);
  int mode = kModeGeneral;
  const bool plain = a.weight_decay != 0.f && !(flags & (DDPB200_ADAMW_MAXIMIZE | DDPB200_ADAMW_STORE_GRAD));
  if (plain && clip_coef && (flags & DDPB200_ADAMW_ROUND_CLIPPED_GRAD)) mode = kModeFastClip;
  else if (plain && !clip_coef) mode = kModeFastPlain;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) return err;
  const int64_t mine = (total_chunks + pp.world - 1) / pp.world;
  const int64_t cap = (int64_t)sms * 2;
  const int grid = (int)(mine < cap ? (mine > 0 ? mine : 1) : cap);
  DevTable tab{dev_table, (int)n};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const char* gl = static_cast<const char*>(grad_arena_local);
  char* ms = static_cast<char*>(exp_avg_shard);
  char* vs = static_cast<char*>(exp_avg_sq_shard);
#define DDPB200_SH_LAUNCH(DT, MCV, WV, MD)                                                              \
  sharded_adamw_kernel<DT, MCV, WV, MD><<<grid, kThreads, 0, st>>>(tab, total_chunks, chunk_elems, gl, pp, ms, vs, \
                                                                 a, lr_ptr, clip_coef, flags)
#define DDPB200_SH_MODE(DT, MCV, WV)                                                                    \
  { if (mode == kModeFastClip) DDPB200_SH_LAUNCH(DT, MCV, WV, kModeFastClip);                           \
    else if (mode == kModeFastPlain) DDPB200_SH_LAUNCH(DT, MCV, WV, kModeFastPlain);                    \
    else DDPB200_SH_LAUNCH(DT, MCV, WV, kModeGeneral); }
#define DDPB200_SH_CASE(DT)                                                                             \
  if (dtype == DT) {                                                                                    \
    if (use_multicast) DDPB200_SH_MODE(DT, 1, 2)                                                        \
    else if (pp.world <= 2) DDPB200_SH_MODE(DT, 0, 2)                                                   \
    else if (pp.world <= 4) DDPB200_SH_MODE(DT, 0, 4)                                                   \
    else DDPB200_SH_MODE(DT, 0, 8)                                                                      \
    return (int)cudaGetLastError();                                                                     \
  }
  DDPB200_SH_CASE(DDPB200_BF16)
  DDPB200_SH_CASE(DDPB200_F32)
#undef DDPB200_SH_CASE
#undef DDPB200_SH_MODE
#undef DDPB200_SH_LAUNCH
  return DDPB200_E_UNSUPPORTED;
}

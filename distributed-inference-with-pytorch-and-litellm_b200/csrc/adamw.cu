// Kernel (c): fused unscale + clip + AdamW, multi-tensor, one persistent launch for the whole
// parameter list. Streams p, g, m, v once through registers: 28 B/param fp32, 14 B/param bf16,
// 22 B/param for bf16 p/g with fp32 state. Roofline: HBM. No shared memory, no tensor cores.
//
// Arithmetic follows ADAM_MODE::ADAMW of TORCH/include/ATen/native/cuda/fused_adam_utils.cuh:41-93
// in fp32 ("opmath") with every rounding step spelled out as an intrinsic, so that the compiler
// cannot contract or reassociate anything and the CPU oracle (oracle/ddp_oracle.c) can follow it
// bit for bit. The bias corrections 1 - beta^step are computed once on the host, in double from
// the fp32-rounded betas, and rounded to fp32 (fused_adam_utils.cuh:131-150 takes the same
// power on the device in fp32).
//
// What the first ncu capture (profiles/r01_*) showed and what was done about it:
//   * 57 issue slots per element, 67 % issue-slot utilisation at 5.5 TB/s: the kernel was close to
//     issue-bound. The option flags (clip / grad_scale / maximize / store-grad / weight-decay) are
//     therefore compile-time in the common cases (FAST modes), and sqrt(v)/sqrt(bc2) — a division
//     by a per-launch constant — uses the exact 3-instruction reciprocal+residual form
//     (div_by_const, common.cuh) instead of the ~10-instruction IEEE division sequence.
//   * 256-bit global accesses with L2 evict-first (sm_100a) halve the LSU instruction count.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include "launch.cuh"
#include "adamw_math.cuh"

namespace ddpb200 {
// Default of the TMA-staged variant (adamw_bulk.cu) for all-bf16 fast modes on a device table: 2 stages x 2 CTAs/SM
// = 64 KB in flight per SM. Llama-3.2-1B set, burst (10 launches) / 300 launches back to back, same box
// (profiles/r02_adamw_kbench.txt): 2.68 / 2.72 ms against 2.82 / 2.78 for the register kernel; more in flight
// loses (3 x 2: 2.87 / 2.85; 4 x 1 wins the burst, 2.63, and loses under the power cap, 2.89). 0 stages = off.
constexpr int kBulkDefaultStages = 2, kBulkDefaultCtas = 2;
}

namespace ddpb200 {

// One chunk through G-element groups (vector accesses); returns how many elements were done, the
// ragged end is left to the caller's scalar loop.
// GELEM: the gradient stream alone is not aligned to the group (a bucket view behind an odd-sized tensor of the
// same bucket, against parameters and state from the allocator): it is read, and in the general mode written,
// element by element — L1 merges the eight accesses of a group — while p, m and v keep their vector accesses.
template <int PDT, int GDT, int SDT, int MODE, int VECB, bool PACKED, bool GELEM = false>
__device__ __forceinline__ int64_t adamw_chunk_vec(char* P, char* Gp, char* M, char* V, int64_t len,
                                                   const AdamScalars& a, const AdamRuntime& r,
                                                   const AdamPacked& pk, bool store_g) {
  constexpr bool any_f32 = (PDT == DDPB200_F32 || GDT == DDPB200_F32 || SDT == DDPB200_F32);
  constexpr int G = GroupSize<VECB, any_f32>::G;
  const int64_t groups_end = len / G * G;
#pragma unroll 1
  for (int64_t i = (int64_t)threadIdx.x * G; i < groups_end; i += kThreads * G) {
    float p[G], g[G], m[G], v[G];
    if constexpr (GELEM) {
#pragma unroll
      for (int k = 0; k < G; ++k) g[k] = load1<GDT>(Gp, i + k);
    } else {
      GroupIO<GDT, G>::load(Gp, i, g);
    }
    GroupIO<PDT, G>::load(P, i, p);
    GroupIO<SDT, G>::load(M, i, m);
    GroupIO<SDT, G>::load(V, i, v);
    if constexpr (PACKED && MODE != kModeGeneral) {
      adamw_group_packed<GDT, MODE, G>(p, g, m, v, a, r, pk);
    } else {
#pragma unroll
      for (int k = 0; k < G; ++k) adamw_elem<GDT, MODE>(p[k], g[k], m[k], v[k], a, r);
    }
    GroupIO<PDT, G>::store(P, i, p);
    GroupIO<SDT, G>::store(M, i, m);
    GroupIO<SDT, G>::store(V, i, v);
    if (MODE == kModeGeneral && store_g) {
      if constexpr (GELEM) {
#pragma unroll
        for (int k = 0; k < G; ++k) store1<GDT>(Gp, i + k, g[k]);
      } else {
        GroupIO<GDT, G>::store(Gp, i, g);
      }
    }
  }
  return groups_end;
}

template <int PDT, int GDT, int SDT, int G>
__device__ __forceinline__ bool all_aligned(const void* P, const void* Gp, const void* M, const void* V) {
  return group_aligned<PDT, G>(P) && group_aligned<GDT, G>(Gp) && group_aligned<SDT, G>(M) &&
         group_aligned<SDT, G>(V);
}

// MAXVEC: widest access the kernel may use (32 = 256-bit, 16 = 128-bit: fewer live registers)
template <class Tab, int PDT, int GDT, int SDT, int MODE, int CT, int MAXVEC = 32, bool PACKED = true>
__global__ void __launch_bounds__(kThreads, CT)
adamw_kernel(const __grid_constant__ Tab tab, int64_t total_chunks, int64_t chunk_elems,
             AdamScalars a, const float* __restrict__ lr_ptr, const float* __restrict__ clip_ptr,
             const float* __restrict__ grad_scale_ptr, const float* __restrict__ found_inf_ptr,
             const float* __restrict__ step_ptr, uint32_t flags) {
  constexpr bool any_f32 = (PDT == DDPB200_F32 || GDT == DDPB200_F32 || SDT == DDPB200_F32);
  if (found_inf_ptr && __ldg(found_inf_ptr) == 1.0f) return;      // GradScaler skip (cuh:137-139)
  if (step_ptr) device_bias_corrections(a, __ldg(step_ptr));      // step count kept on the device (AMP: no host sync)
  const AdamRuntime r = make_adam_runtime(a, lr_ptr, clip_ptr, grad_scale_ptr, flags);
  const bool store_g = (flags & DDPB200_ADAMW_STORE_GRAD) != 0;
  const AdamPacked pk = make_adam_packed(a, r);
  const int64_t chunk_base = tab.begin(0);

  SegCursor cur;
  for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const int s = find_seg_cached(tab, c + chunk_base, chunk_base + total_chunks, cur);
    const int64_t start = (c + chunk_base - cur.b) * chunk_elems;
    const int64_t rest = tab.numel(s) - start;
    const int64_t len = rest < chunk_elems ? rest : chunk_elems;
    char* P = static_cast<char*>(tab.ptr(s, 0)) + start * ElemOf<PDT>::size;
    char* Gp = static_cast<char*>(tab.ptr(s, 1)) + start * ElemOf<GDT>::size;
    char* M = static_cast<char*>(tab.ptr(s, 2)) + start * ElemOf<SDT>::size;
    char* V = static_cast<char*>(tab.ptr(s, 3)) + start * ElemOf<SDT>::size;

    int64_t done = 0;
    if (MAXVEC >= 32 && all_aligned<PDT, GDT, SDT, GroupSize<32, any_f32>::G>(P, Gp, M, V))
      done = adamw_chunk_vec<PDT, GDT, SDT, MODE, 32, PACKED>(P, Gp, M, V, len, a, r, pk, store_g);
    else if (all_aligned<PDT, GDT, SDT, GroupSize<16, any_f32>::G>(P, Gp, M, V))
      done = adamw_chunk_vec<PDT, GDT, SDT, MODE, 16, PACKED>(P, Gp, M, V, len, a, r, pk, store_g);
    else if (all_aligned<PDT, GDT, SDT, GroupSize<16, any_f32>::G>(P, P, M, V))          // only the gradient is off
      done = adamw_chunk_vec<PDT, GDT, SDT, MODE, 16, PACKED, true>(P, Gp, M, V, len, a, r, pk, store_g);
    // ragged end, or a segment whose base is not even 16-byte aligned (odd offset inside a bucket)
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
      float p = load1<PDT>(P, i), g = load1<GDT>(Gp, i), m = load1<SDT>(M, i), v = load1<SDT>(V, i);
      adamw_elem<GDT, MODE>(p, g, m, v, a, r);
      store1<PDT>(P, i, p);
      store1<SDT>(M, i, m);
      store1<SDT>(V, i, v);
      if (MODE == kModeGeneral && store_g) store1<GDT>(Gp, i, g);
    }
  }
}

template <int PDT, int GDT, int SDT, int MODE>
static int launch_adamw(const ddpb200_seg* table, int table_loc, int64_t n, int64_t total_chunks,
                        int64_t chunk_elems, const AdamScalars& a, const float* lr_ptr,
                        const float* clip, const float* grad_scale, const float* found_inf,
                        const float* step_ptr, uint32_t flags, cudaStream_t st) {
  return for_each_table_slice(table, table_loc == DDPB200_TABLE_DEVICE, n, total_chunks,
    [&](auto& tab, int64_t chunks, int64_t, int64_t) -> int {
      using Tab = std::remove_reference_t<decltype(tab)>;
      int err = 0;
      const int sms = sm_count_cached(&err);
      if (err) return err;
      // CTAs per SM and access width, measured on B200 at Llama-1B size (tools/kbench.py,
      // profiles/r01_kbench_ctas.txt), burst (10 launches) / sustained (600 launches, power cap active):
      //   all-bf16 fast modes: 256-bit x 3 CTAs/SM 6.28 / 5.94 TB/s; 128-bit x 4 CTAs/SM 6.22 / 6.20 TB/s
      //                        (128-bit x 3: 5.54, x 5: 5.54 with spills; 256-bit x 4: 5.05 with spills)
      //   -> 128-bit x 4: the narrower groups need fewer live registers, the fourth CTA hides latency
      //      when the power cap pulls the SM clock down, which is the state a training run lives in.
      //   anything with an fp32 stream, and the general mode: 256-bit x 2 CTAs/SM (6.6 TB/s fp32).
      constexpr bool any_f32 = (PDT == DDPB200_F32 || GDT == DDPB200_F32 || SDT == DDPB200_F32);
      constexpr bool narrow = !any_f32 && MODE != kModeGeneral;
      constexpr int CT = narrow ? 4 : 2;
      constexpr int MAXVEC = narrow ? 16 : 32;
      // tuning: DDPB200_ADAMW_CTAS = CTAs per SM actually launched (<= CT; fractional: 3.5 = half the SMs hold 4
      // CTAs, half 3), read at every launch so that tools/kbench.py can sweep it in one process
      const double env_ct = [] { const char* e = getenv("DDPB200_ADAMW_CTAS"); return e ? atof(e) : 0.0; }();
      const int64_t cap = (env_ct > 0.0 && env_ct <= CT) ? (int64_t)(sms * env_ct) : (int64_t)sms * CT;
      const int grid = (int)(chunks < cap ? chunks : cap);
      if (grid == 0) return 0;
      // A/B switch for tools/kbench.py only: DDPB200_ADAMW_WIDE=1 selects 256-bit x 3 CTAs/SM
      static const bool wide = [] { const char* e = getenv("DDPB200_ADAMW_WIDE"); return e && atoi(e) == 1; }();
      if constexpr (narrow) {
        if (wide) {
          const int64_t wcap = (int64_t)sms * 3;
          adamw_kernel<Tab, PDT, GDT, SDT, MODE, 3, 32><<<(int)(chunks < wcap ? chunks : wcap), kThreads, 0, st>>>(
              tab, chunks, chunk_elems, a, lr_ptr, clip, grad_scale, found_inf, step_ptr, flags);
          return (int)cudaGetLastError();
        }
      }
      // A/B switch for tools/kbench.py only: DDPB200_ADAMW_PACKED=0 selects the scalar element arithmetic
      // (IEEE intrinsics per element) that round 1 shipped, for the all-bf16 fast modes
      static const int packed_env = [] { const char* e = getenv("DDPB200_ADAMW_PACKED"); return e ? atoi(e) : -1; }();
      // fp32 streams: the packed path's 119 registers cost more than its issue slots save (5.46 ms against 5.29 for
      // the scalar arithmetic on the Llama-1B set, profiles/r02_kbench_all.txt), so it is the default only for all-bf16
      const bool scalar_math = packed_env == 0 || (packed_env < 0 && !narrow);
      if constexpr (MODE != kModeGeneral) {
        if (scalar_math) {
          adamw_kernel<Tab, PDT, GDT, SDT, MODE, CT, MAXVEC, false><<<grid, kThreads, 0, st>>>(
              tab, chunks, chunk_elems, a, lr_ptr, clip, grad_scale, found_inf, step_ptr, flags);
          return (int)cudaGetLastError();
        }
      }
      adamw_kernel<Tab, PDT, GDT, SDT, MODE, CT, MAXVEC><<<grid, kThreads, 0, st>>>(
          tab, chunks, chunk_elems, a, lr_ptr, clip, grad_scale, found_inf, step_ptr, flags);
      return (int)cudaGetLastError();
    });
}

}  // namespace ddpb200

namespace ddpb200 {
// adamw_bulk.cu
template <int DT, int MODE>
int launch_adamw_bulk(const ddpb200_seg* table, int table_loc, int64_t n, int64_t total_chunks, int64_t chunk_elems,
                      const AdamScalars& a, const float* lr_ptr, const float* clip, uint32_t flags, cudaStream_t st,
                      int stages, int ctas, int* rc);
}  // namespace ddpb200

using namespace ddpb200;

extern "C" DDPB200_API int ddpb200_adamw_bias_corrections(double beta1, double beta2, int64_t step, float* out) {
  if (!out || step < 1) return DDPB200_E_INVALID;
  // betas rounded to fp32 first: the kernel's recurrences use fp32 betas, and the corrections
  // must cancel against them (fused_adam_utils.cuh:131-150 does the same cast before pow)
  const double bc1 = 1.0 - pow((double)(float)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)(float)beta2, (double)step);
  out[0] = (float)bc1;
  out[1] = (float)sqrt(bc2);
  return 0;
}

extern "C" DDPB200_API int ddpb200_fused_clip_adamw(const ddpb200_seg* table, int32_t table_loc, int64_t n,
                                        int64_t total_chunks, int64_t chunk_elems,
                                        int32_t param_dtype, int32_t grad_dtype,
                                        int32_t state_dtype, double lr, const float* lr_ptr,
                                        double beta1, double beta2, double eps,
                                        double weight_decay, int64_t step, const float* step_ptr,
                                        const float* clip_coef, const float* grad_scale, const float* found_inf,
                                        uint32_t flags, void* stream) {
  if (!valid_chunking(n, total_chunks, chunk_elems) || (n > 0 && !table) || (step < 1 && !step_ptr))
    return DDPB200_E_INVALID;
  if (step < 1) step = 1;          // step_ptr given: the device value is used, this only seeds the host-side scalars
  if (n == 0) return 0;
  float bc[2];
  ddpb200_adamw_bias_corrections(beta1, beta2, step, bc);
  AdamScalars a = make_adam_scalars(lr, beta1, beta2, eps, weight_decay, bc);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // common cases get compile-time options; anything else takes the general kernel
  int mode = kModeGeneral;
  const bool plain = !grad_scale && a.weight_decay != 0.f &&
                     !(flags & (DDPB200_ADAMW_MAXIMIZE | DDPB200_ADAMW_STORE_GRAD));
  if (plain && clip_coef && (flags & DDPB200_ADAMW_ROUND_CLIPPED_GRAD)) mode = kModeFastClip;
  else if (plain && !clip_coef) mode = kModeFastPlain;
  // DDPB200_ADAMW_BULK="stages,ctas" takes the all-bf16 / all-fp32 fast modes on a device table through the TMA-staged
  // variant (adamw_bulk.cu); "0" switches it off; read at every launch so that tools/kbench.py can sweep it
  if (mode != kModeGeneral && param_dtype == grad_dtype && grad_dtype == state_dtype && !found_inf && !step_ptr) {
    // all-fp32 stays on the register kernel by default: 5.19 / 5.25 ms (burst / sustained) against 5.32 / 5.34 staged
    int stages = param_dtype == DDPB200_BF16 ? kBulkDefaultStages : 0, ctas = kBulkDefaultCtas;
    if (const char* e = getenv("DDPB200_ADAMW_BULK")) {
      if (sscanf(e, "%d,%d", &stages, &ctas) != 2) stages = 0;
    }
    if (stages > 0) {
      int rc = 0, took = 0;
#define DDPB200_BULK_CALL(DT, MD) launch_adamw_bulk<DT, MD>(table, table_loc, n, total_chunks, chunk_elems, a, lr_ptr, clip_coef, flags, st, stages, ctas, &rc)
      if (param_dtype == DDPB200_BF16)
        took = mode == kModeFastClip ? DDPB200_BULK_CALL(DDPB200_BF16, kModeFastClip) : DDPB200_BULK_CALL(DDPB200_BF16, kModeFastPlain);
      else
        took = mode == kModeFastClip ? DDPB200_BULK_CALL(DDPB200_F32, kModeFastClip) : DDPB200_BULK_CALL(DDPB200_F32, kModeFastPlain);
#undef DDPB200_BULK_CALL
      if (took) return rc;
    }
  }
#define DDPB200_ADAMW_CASE(P, G, S)                                                                   \
  if (param_dtype == P && grad_dtype == G && state_dtype == S) {                                      \
    if (mode == kModeFastClip)                                                                        \
      return launch_adamw<P, G, S, kModeFastClip>(table, table_loc, n, total_chunks, chunk_elems, a,  \
                                                  lr_ptr, clip_coef, grad_scale, found_inf, step_ptr, flags, st); \
    if (mode == kModeFastPlain)                                                                       \
      return launch_adamw<P, G, S, kModeFastPlain>(table, table_loc, n, total_chunks, chunk_elems, a, \
                                                   lr_ptr, clip_coef, grad_scale, found_inf, step_ptr, flags, st); \
    return launch_adamw<P, G, S, kModeGeneral>(table, table_loc, n, total_chunks, chunk_elems, a,     \
                                               lr_ptr, clip_coef, grad_scale, found_inf, step_ptr, flags, st);  \
  }
  DDPB200_ADAMW_CASE(DDPB200_F32, DDPB200_F32, DDPB200_F32)
  DDPB200_ADAMW_CASE(DDPB200_BF16, DDPB200_BF16, DDPB200_BF16)
  DDPB200_ADAMW_CASE(DDPB200_BF16, DDPB200_BF16, DDPB200_F32)
  DDPB200_ADAMW_CASE(DDPB200_F32, DDPB200_BF16, DDPB200_F32)
#undef DDPB200_ADAMW_CASE
  return DDPB200_E_UNSUPPORTED;
}

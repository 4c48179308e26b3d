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
#include <math.h>
#include "launch.cuh"

namespace ddpb200 {

struct AdamScalars {
  float lr, beta1, beta2, eps, lr_wd_valid;  // lr_wd_valid: 1 if weight decay is applied
  float weight_decay, bc1, bc2_sqrt;
};

template <int PDT, int GDT, int SDT> struct AdamGroup {
  static constexpr bool any_f32 = (PDT == DDPB200_F32 || GDT == DDPB200_F32 || SDT == DDPB200_F32);
  static constexpr int G = any_f32 ? 4 : 8;
  static constexpr int U = 8 / G;
};

// one element; p, m, v updated in place, g returns the value to store back (if asked)
template <int GDT>
__device__ __forceinline__ void adamw_elem(float& p, float& g, float& m, float& v,
                                           const AdamScalars& a, float lr_wd, float step_size,
                                           float inv_scale_valid, float grad_scale,
                                           float clip_valid, float clip_coef, uint32_t flags) {
  if (inv_scale_valid != 0.f) g = __fdiv_rn(g, grad_scale);
  if (clip_valid != 0.f) {
    g = __fmul_rn(g, clip_coef);
    if (flags & DDPB200_ADAMW_ROUND_CLIPPED_GRAD) g = round_through<GDT>(g);
  }
  float gm = (flags & DDPB200_ADAMW_MAXIMIZE) ? -g : g;
  if (a.lr_wd_valid != 0.f) p = __fmaf_rn(-lr_wd, p, p);            // p -= lr*wd*p
  m = __fmaf_rn(a.beta1, m, __fmaf_rn(-a.beta1, gm, gm));            // lerp(m, g, 1-b1)
  const float gg = __fmul_rn(gm, gm);
  v = __fmaf_rn(a.beta2, v, __fmaf_rn(-a.beta2, gg, gg));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), a.bc2_sqrt), a.eps);
  p = __fsub_rn(p, __fdiv_rn(__fmul_rn(step_size, m), denom));
}

template <class Tab, int PDT, int GDT, int SDT>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
adamw_kernel(const __grid_constant__ Tab tab, int64_t total_chunks, int64_t chunk_elems,
             AdamScalars a, const float* __restrict__ lr_ptr, const float* __restrict__ clip_ptr,
             const float* __restrict__ grad_scale_ptr, const float* __restrict__ found_inf_ptr,
             uint32_t flags) {
  constexpr int G = AdamGroup<PDT, GDT, SDT>::G;
  constexpr int U = AdamGroup<PDT, GDT, SDT>::U;

  if (found_inf_ptr && __ldg(found_inf_ptr) == 1.0f) return;      // GradScaler skip (cuh:137-139)
  if (lr_ptr) a.lr = __ldg(lr_ptr);
  const float lr_wd = __fmul_rn(a.lr, a.weight_decay);
  const float step_size = __fdiv_rn(a.lr, a.bc1);
  const float gs_valid = grad_scale_ptr ? 1.f : 0.f;
  const float gs = grad_scale_ptr ? __ldg(grad_scale_ptr) : 1.f;
  const float cl_valid = clip_ptr ? 1.f : 0.f;
  const float cl = clip_ptr ? __ldg(clip_ptr) : 1.f;
  const bool store_g = (flags & DDPB200_ADAMW_STORE_GRAD) != 0;
  const int64_t chunk_base = tab.begin(0);

  for (int64_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const int s = find_seg(tab, c + chunk_base);
    const int64_t start = (c + chunk_base - tab.begin(s)) * chunk_elems;
    const int64_t rest = tab.numel(s) - start;
    const int64_t len = rest < chunk_elems ? rest : chunk_elems;
    char* P = static_cast<char*>(tab.ptr(s, 0)) + start * ElemOf<PDT>::size;
    char* Gp = static_cast<char*>(tab.ptr(s, 1)) + start * ElemOf<GDT>::size;
    char* M = static_cast<char*>(tab.ptr(s, 2)) + start * ElemOf<SDT>::size;
    char* V = static_cast<char*>(tab.ptr(s, 3)) + start * ElemOf<SDT>::size;

    const bool vec = group_aligned<PDT, G>(P) && group_aligned<GDT, G>(Gp) &&
                     group_aligned<SDT, G>(M) && group_aligned<SDT, G>(V);
    int64_t scalar_from = 0;
    if (vec) {
      const int64_t full = len / kTileElems * kTileElems;
      for (int64_t t0 = 0; t0 < full; t0 += kTileElems) {
        float p[U][G], g[U][G], m[U][G], v[U][G];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = t0 + u * (kThreads * G) + threadIdx.x * G;
          GroupIO<GDT, G>::load(Gp, i, g[u]);
          GroupIO<PDT, G>::load(P, i, p[u]);
          GroupIO<SDT, G>::load(M, i, m[u]);
          GroupIO<SDT, G>::load(V, i, v[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = t0 + u * (kThreads * G) + threadIdx.x * G;
#pragma unroll
          for (int k = 0; k < G; ++k)
            adamw_elem<GDT>(p[u][k], g[u][k], m[u][k], v[u][k], a, lr_wd, step_size, gs_valid, gs,
                            cl_valid, cl, flags);
          GroupIO<PDT, G>::store(P, i, p[u]);
          GroupIO<SDT, G>::store(M, i, m[u]);
          GroupIO<SDT, G>::store(V, i, v[u]);
          if (store_g) GroupIO<GDT, G>::store(Gp, i, g[u]);
        }
      }
      const int64_t groups_end = len / G * G;
      for (int64_t i = full + (int64_t)threadIdx.x * G; i < groups_end; i += kThreads * G) {
        float p[G], g[G], m[G], v[G];
        GroupIO<GDT, G>::load(Gp, i, g);
        GroupIO<PDT, G>::load(P, i, p);
        GroupIO<SDT, G>::load(M, i, m);
        GroupIO<SDT, G>::load(V, i, v);
#pragma unroll
        for (int k = 0; k < G; ++k)
          adamw_elem<GDT>(p[k], g[k], m[k], v[k], a, lr_wd, step_size, gs_valid, gs, cl_valid, cl, flags);
        GroupIO<PDT, G>::store(P, i, p);
        GroupIO<SDT, G>::store(M, i, m);
        GroupIO<SDT, G>::store(V, i, v);
        if (store_g) GroupIO<GDT, G>::store(Gp, i, g);
      }
      scalar_from = groups_end;
    }
    for (int64_t i = scalar_from + threadIdx.x; i < len; i += kThreads) {
      float p = load1<PDT>(P, i), g = load1<GDT>(Gp, i), m = load1<SDT>(M, i), v = load1<SDT>(V, i);
      adamw_elem<GDT>(p, g, m, v, a, lr_wd, step_size, gs_valid, gs, cl_valid, cl, flags);
      store1<PDT>(P, i, p);
      store1<SDT>(M, i, m);
      store1<SDT>(V, i, v);
      if (store_g) store1<GDT>(Gp, i, g);
    }
  }
}

template <int PDT, int GDT, int SDT>
static int launch_adamw(const ddpb200_seg* table, int table_loc, int64_t n, int64_t total_chunks,
                        int64_t chunk_elems, const AdamScalars& a, const float* lr_ptr,
                        const float* clip, const float* grad_scale, const float* found_inf,
                        uint32_t flags, cudaStream_t st) {
  return for_each_table_slice(table, table_loc == DDPB200_TABLE_DEVICE, n, total_chunks,
    [&](auto& tab, int64_t chunks, int64_t, int64_t) -> int {
      using Tab = std::remove_reference_t<decltype(tab)>;
      int err = 0;
      const int grid = persistent_grid(chunks, &err);
      if (err) return err;
      if (grid == 0) return 0;
      adamw_kernel<Tab, PDT, GDT, SDT><<<grid, kThreads, 0, st>>>(
          tab, chunks, chunk_elems, a, lr_ptr, clip, grad_scale, found_inf, flags);
      return (int)cudaGetLastError();
    });
}

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
                                        double weight_decay, int64_t step, const float* clip_coef,
                                        const float* grad_scale, const float* found_inf,
                                        uint32_t flags, void* stream) {
  if (!valid_chunking(n, total_chunks, chunk_elems) || (n > 0 && !table) || step < 1)
    return DDPB200_E_INVALID;
  if (n == 0) return 0;
  float bc[2];
  ddpb200_adamw_bias_corrections(beta1, beta2, step, bc);
  AdamScalars a;
  a.lr = (float)lr; a.beta1 = (float)beta1; a.beta2 = (float)beta2; a.eps = (float)eps;
  a.weight_decay = (float)weight_decay; a.lr_wd_valid = (weight_decay != 0.0) ? 1.f : 0.f;
  a.bc1 = bc[0]; a.bc2_sqrt = bc[1];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define DDPB200_ADAMW_CASE(P, G, S)                                                       \
  if (param_dtype == P && grad_dtype == G && state_dtype == S)                            \
    return launch_adamw<P, G, S>(table, table_loc, n, total_chunks, chunk_elems, a, lr_ptr, \
                                 clip_coef, grad_scale, found_inf, flags, st);
  DDPB200_ADAMW_CASE(DDPB200_F32, DDPB200_F32, DDPB200_F32)
  DDPB200_ADAMW_CASE(DDPB200_BF16, DDPB200_BF16, DDPB200_BF16)
  DDPB200_ADAMW_CASE(DDPB200_BF16, DDPB200_BF16, DDPB200_F32)
  DDPB200_ADAMW_CASE(DDPB200_F32, DDPB200_BF16, DDPB200_F32)
#undef DDPB200_ADAMW_CASE
  return DDPB200_E_UNSUPPORTED;
}

/*
 * ddp_oracle.c — CPU restatement of the DDP gradient-sync + clip + AdamW arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in the product path may import, link or call this file;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg do,
 * and only as the checker or the timed CPU baseline.
 *
 * The upstream repo holds no implementation of this path: the arithmetic lives in
 * its un-vendored dependency torch ("torch>=1.8.0", REF/requirements.txt:2; installed and
 * authoritative here: torch 2.11.0). Each function cites the torch file:line it restates
 * (TORCH/ = torch 2.11.0 site-packages). The oracle is pinned against golden vectors produced by
 * running that torch code in the build container (tests/golden/make_golden.py).
 *
 * Plain scalar C, fp32 "opmath" with explicit fmaf()/sqrtf() and -ffp-contract=off so every
 * rounding step is the one written down.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_F32 0
#define ORACLE_BF16 1

/* ---- bf16 <-> fp32, round to nearest even (c10/util/BFloat16.h round_to_nearest_even) ---- */
static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if (f != f) return 0x7fc0;
  const uint32_t bias = 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)((u + bias) >> 16);
}
static inline float ld(const void* p, int64_t i, int dt) {
  return dt == ORACLE_F32 ? ((const float*)p)[i] : bf16_to_f32(((const uint16_t*)p)[i]);
}
static inline void st(void* p, int64_t i, int dt, float x) {
  if (dt == ORACLE_F32) ((float*)p)[i] = x; else ((uint16_t*)p)[i] = f32_to_bf16(x);
}
static inline float rnd(float x, int dt) { return dt == ORACLE_F32 ? x : bf16_to_f32(f32_to_bf16(x)); }

float oracle_round_bf16(float x) { return rnd(x, ORACLE_BF16); }

/* ---- a-1: bucket assignment --------------------------------------------------------------
 * Restates compute_bucket_assignment_by_size (decl TORCH/include/torch/csrc/distributed/c10d/
 * reducer.hpp:586-598; BucketAccumulator reducer.hpp:39-43; Python call site
 * TORCH/nn/parallel/distributed.py:1237-1244). Tensors are visited in the given order; each
 * (dtype,device) key owns one open bucket and a cursor into size_limits; a bucket closes when
 * bytes >= limit; the cursor then advances unless already at the last limit; sparse tensors are
 * closed alone with limit 0; leftovers are appended; without tensor_indices the result is sorted
 * by each bucket's minimum index.
 * bucket_of[i] receives the bucket number of the i-th VISITED tensor; order_in_bucket is implied
 * by visit order. Returns the number of buckets, or -1 on bad input. */
int64_t oracle_bucket_assign(const int64_t* numel, const int32_t* elem_size, const int64_t* key,
                             const uint8_t* expect_sparse, const int64_t* tensor_indices,
                             int64_t n, const uint64_t* limits, int64_t n_limits,
                             int64_t* bucket_of, uint64_t* bucket_limit) {
  if (n <= 0 || n_limits <= 0) return -1;
  /* per-key state, keys discovered in visit order */
  int64_t* keys = (int64_t*)malloc(sizeof(int64_t) * n);
  int64_t* cursor = (int64_t*)calloc(n, sizeof(int64_t));
  uint64_t* bytes = (uint64_t*)calloc(n, sizeof(uint64_t));
  int64_t* open_id = (int64_t*)malloc(sizeof(int64_t) * n); /* provisional id of the open bucket, -1 none */
  int64_t* close_rank = (int64_t*)malloc(sizeof(int64_t) * n); /* provisional id -> order of closing */
  uint64_t* prov_limit = (uint64_t*)malloc(sizeof(uint64_t) * n);
  int64_t* prov_of = (int64_t*)malloc(sizeof(int64_t) * n);
  int64_t nkeys = 0, nprov = 0, nclosed = 0;

  for (int64_t i = 0; i < n; ++i) {
    const int64_t ti = tensor_indices ? tensor_indices[i] : i;
    if (expect_sparse && expect_sparse[ti]) {
      prov_of[i] = nprov; prov_limit[nprov] = 0; close_rank[nprov] = nclosed++; ++nprov;
      continue;
    }
    const int64_t k = key ? key[i] : 0;
    int64_t ks = -1;
    for (int64_t j = 0; j < nkeys; ++j) if (keys[j] == k) { ks = j; break; }
    if (ks < 0) { ks = nkeys++; keys[ks] = k; cursor[ks] = 0; bytes[ks] = 0; open_id[ks] = -1; }
    if (open_id[ks] < 0) { open_id[ks] = nprov++; }
    prov_of[i] = open_id[ks];
    bytes[ks] += (uint64_t)numel[i] * (uint64_t)elem_size[i];
    prov_limit[open_id[ks]] = limits[cursor[ks]];
    if (bytes[ks] >= limits[cursor[ks]]) {
      close_rank[open_id[ks]] = nclosed++;
      open_id[ks] = -1; bytes[ks] = 0;
      if (cursor[ks] + 1 < n_limits) ++cursor[ks];
    }
  }
  /* leftovers: torch iterates a std::unordered_map, which libstdc++ visits newest key first
   * (checked against torch in tests/golden/make_golden.py's mixed-dtype cases) */
  for (int64_t j = nkeys - 1; j >= 0; --j)
    if (open_id[j] >= 0) close_rank[open_id[j]] = nclosed++;

  /* final numbering: closing order, or by minimum tensor index when no tensor_indices */
  int64_t* final_of_prov = (int64_t*)malloc(sizeof(int64_t) * (nprov ? nprov : 1));
  if (tensor_indices) {
    for (int64_t b = 0; b < nprov; ++b) final_of_prov[b] = close_rank[b];
  } else {
    /* min index of a provisional bucket = first visited tensor in it (indices ascend) */
    int64_t* minidx = (int64_t*)malloc(sizeof(int64_t) * (nprov ? nprov : 1));
    for (int64_t b = 0; b < nprov; ++b) minidx[b] = INT64_MAX;
    for (int64_t i = 0; i < n; ++i) if (i < minidx[prov_of[i]]) minidx[prov_of[i]] = i;
    for (int64_t b = 0; b < nprov; ++b) {
      int64_t r = 0;
      for (int64_t c = 0; c < nprov; ++c)
        if (minidx[c] < minidx[b] || (minidx[c] == minidx[b] && close_rank[c] < close_rank[b])) ++r;
      final_of_prov[b] = r;
    }
    free(minidx);
  }
  for (int64_t i = 0; i < n; ++i) bucket_of[i] = final_of_prov[prov_of[i]];
  for (int64_t b = 0; b < nprov; ++b) bucket_limit[final_of_prov[b]] = prov_limit[b];

  free(keys); free(cursor); free(bytes); free(open_id); free(close_rank); free(prov_limit);
  free(prov_of); free(final_of_prov);
  return nprov;
}

/* ---- a-2 / a-4: pack and unpack --------------------------------------------------------------
 * Reducer::mark_variable_ready_dense writes bucket_view = grad * (1/world_size)
 * (TORCH/include/torch/csrc/distributed/c10d/reducer.hpp:276, div_factor_ reducer.hpp:497-500):
 * exactly round_dst(fp32(g) * float(1.0/ws)) (probe). copy_bucket_to_grad
 * (reducer.hpp:328-334) is the same with scale 1. */
void oracle_scale_cast(const void* src, int src_dt, void* dst, int dst_dt, int64_t n, float scale) {
  for (int64_t i = 0; i < n; ++i) st(dst, i, dst_dt, ld(src, i, src_dt) * scale);
}

/* ---- a-5: grad norm and clip coefficient -------------------------------------------------------
 * torch._foreach_norm(grads, 2) (TORCH/nn/utils/clip_grad.py:96): per-tensor sqrt(sum g^2) with
 * a wide accumulator. Returned in double; the caller applies dtype rounding. */
double oracle_sumsq(const void* g, int dt, int64_t n) {
  double acc = 0.0;
  for (int64_t i = 0; i < n; ++i) { const double x = (double)ld(g, i, dt); acc += x * x; }
  return acc;
}

/* total = vector_norm(stack(norms)) (clip_grad.py:106-108); coef = max_norm / (total + 1e-6)
 * (:165); clamp(max=1) (:169). torch_bf16 != 0 rounds every intermediate tensor to bf16, as
 * happens when the grads (hence the norms) are bf16. out[0]=total, out[1]=coef. */
void oracle_total_norm_coef(const double* sumsq, int64_t n, int torch_bf16, float max_norm, float* out) {
  double acc = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    if (torch_bf16) {
      const float nt = rnd(sqrtf((float)sumsq[i]), ORACLE_BF16);
      acc += (double)(nt * nt);
    } else {
      acc += sumsq[i];
    }
  }
  float total = sqrtf((float)acc);
  float coef;
  if (torch_bf16) {
    total = rnd(total, ORACLE_BF16);
    const float d = rnd(total + 1e-6f, ORACLE_BF16);
    coef = rnd(max_norm / d, ORACLE_BF16);
  } else {
    coef = max_norm / (total + 1e-6f);
  }
  if (coef == coef && coef > 1.0f) coef = 1.0f;
  out[0] = total;
  out[1] = coef;
}

/* g *= coef in the grad dtype (torch._foreach_mul_, clip_grad.py:174) */
void oracle_scale_inplace(void* g, int dt, int64_t n, float coef) {
  for (int64_t i = 0; i < n; ++i) st(g, i, dt, ld(g, i, dt) * coef);
}

/* ---- a-6: AdamW ---------------------------------------------------------------------------------
 * Bias corrections for the fused arithmetic: 1 - beta^step with beta ALREADY ROUNDED TO fp32,
 * because the recurrences below use fp32 betas (TORCH/include/ATen/native/cuda/
 * fused_adam_utils.cuh:131-150 casts beta to opmath before pow). Using the double beta here
 * would leave v_hat = (1-float(b2)) g^2 / (1-b2) off by 1.3e-5 at step 1. The power itself is
 * taken in double and rounded once. out[0]=bc1, out[1]=sqrt(bc2). */
void oracle_bias_corrections(double beta1, double beta2, int64_t step, float* out) {
  const double b1 = (double)(float)beta1, b2 = (double)(float)beta2;
  out[0] = (float)(1.0 - pow(b1, (double)step));
  out[1] = (float)sqrt(1.0 - pow(b2, (double)step));
}

/* Fused element math, ADAM_MODE::ADAMW, amsgrad=false
 * (TORCH/include/ATen/native/cuda/fused_adam_utils.cuh:41-93), preceded by the GradScaler unscale
 * (.cuh:52-55) and the clip multiply of clip_grad.py:174 (rounded to the grad dtype when
 * round_clipped != 0, which is what an in-place _foreach_mul_ on the grads leaves behind).
 * flags: bit0 round clipped grad, bit1 store grad back, bit2 maximize. */
void oracle_fused_clip_adamw(void* p, int p_dt, void* g, int g_dt, void* m, void* v, int s_dt,
                             int64_t n, double lr_d, double beta1_d, double beta2_d, double eps_d,
                             double wd_d, int64_t step, int has_clip, float clip_coef,
                             int has_grad_scale, float grad_scale, uint32_t flags) {
  float bc[2];
  oracle_bias_corrections(beta1_d, beta2_d, step, bc);
  const float lr = (float)lr_d, beta1 = (float)beta1_d, beta2 = (float)beta2_d;
  const float eps = (float)eps_d, wd = (float)wd_d;
  const float lr_wd = lr * wd;
  const float step_size = lr / bc[0];
  for (int64_t i = 0; i < n; ++i) {
    float pi = ld(p, i, p_dt), gi = ld(g, i, g_dt), mi = ld(m, i, s_dt), vi = ld(v, i, s_dt);
    if (has_grad_scale) gi = gi / grad_scale;
    if (has_clip) { gi = gi * clip_coef; if (flags & 1u) gi = rnd(gi, g_dt); }
    const float gm = (flags & 4u) ? -gi : gi;
    if (wd_d != 0.0) pi = fmaf(-lr_wd, pi, pi);
    mi = fmaf(beta1, mi, fmaf(-beta1, gm, gm));
    const float gg = gm * gm;
    vi = fmaf(beta2, vi, fmaf(-beta2, gg, gg));
    const float denom = sqrtf(vi) / bc[1] + eps;
    pi = pi - (step_size * mi) / denom;
    st(p, i, p_dt, pi); st(m, i, s_dt, mi); st(v, i, s_dt, vi);
    if (flags & 2u) st(g, i, g_dt, gi);
  }
}

/* torch's default (non-fused) element math, _single_tensor_adam with decoupled weight decay
 * (TORCH/optim/adam.py:419 p.mul_(1-lr*wd); :457 lerp_; :476 mul_/addcmul_; :531-547 sqrt, div,
 * add, addcdiv_). Every tensor op rounds its result to the tensor dtype; scalars are Python
 * doubles cast to opmath fp32 inside each op. The placement of the one fused multiply-add
 * (addcmul) and the (value*t1)/t2 order of addcdiv are those of torch's CPU kernels, found by
 * bit-comparison (tests/golden/make_golden.py); torch's vectorised CPU sqrt is not correctly
 * rounded in ~0.6 % of lanes, so fp32 agreement is to 1 ulp there, not bitwise. Kept to measure the spread between torch's own two
 * paths (fused vs default), which bounds the tolerance the parity tests state. */
void oracle_single_tensor_adamw(void* p, int p_dt, const void* g, int g_dt, void* m, void* v,
                                int s_dt, int64_t n, double lr, double beta1, double beta2,
                                double eps, double wd, int64_t step) {
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  const float decay = (float)(1.0 - lr * wd);
  const float w = (float)(1.0 - beta1);
  const float b2 = (float)beta2, omb2 = (float)(1.0 - beta2);
  const float bc2_sqrt = (float)sqrt(bc2);
  const float neg_step_size = (float)(-(lr / bc1));
  const float epsf = (float)eps;
  for (int64_t i = 0; i < n; ++i) {
    float pi = ld(p, i, p_dt), gi = ld(g, i, g_dt), mi = ld(m, i, s_dt), vi = ld(v, i, s_dt);
    pi = rnd(pi * decay, p_dt);                              /* param.mul_(1 - lr*wd)          */
    mi = rnd(w < 0.5f ? fmaf(w, gi - mi, mi)                /* exp_avg.lerp_(grad, 1-beta1)   */
                      : gi - (gi - mi) * (1.0f - w), s_dt);
    vi = rnd(vi * b2, s_dt);                                 /* exp_avg_sq.mul_(beta2)         */
    vi = rnd(fmaf(omb2 * gi, gi, vi), s_dt);                 /* .addcmul_(grad, grad, 1-beta2) */
    float denom = rnd(sqrtf(vi), s_dt);                      /* exp_avg_sq.sqrt()              */
    denom = rnd(denom / bc2_sqrt, s_dt);                     /* / bias_correction2_sqrt        */
    denom = rnd(denom + epsf, s_dt);                         /* .add_(eps)                     */
    pi = rnd(pi + (neg_step_size * mi) / denom, p_dt);       /* param.addcdiv_(m, denom, -ss)  */
    st(p, i, p_dt, pi); st(m, i, s_dt, mi); st(v, i, s_dt, vi);
  }
}

/* ---- check of the kernel's division by a per-launch constant -----------------------------------
 * csrc/common.cuh div_by_const computes x / c as q0 = x*rc; r = fma(-q0, c, x); q = fma(r, rc, q0)
 * with rc = RN(1/c). This counts, over ALL 2^23 significands of x in [1, 2), how often that differs
 * from the IEEE quotient x / c (scaling x by a power of two scales every step exactly). The oracle's
 * own AdamW above uses the plain '/' operator; this function exists so a CPU test can show that the
 * two are the same function for the constants the kernel meets. */
int64_t oracle_div_by_const_mismatches(float c) {
  const float rc = 1.0f / c;
  int64_t bad = 0;
  for (uint32_t mant = 0; mant < (1u << 23); ++mant) {
    const uint32_t u = 0x3f800000u | mant;
    float x;
    memcpy(&x, &u, 4);
    const float q0 = x * rc;
    const float r = fmaf(-q0, c, x);
    const float q = fmaf(r, rc, q0);
    if (q != x / c) ++bad;
  }
  return bad;
}

// Standalone tuning harness (not part of the product): times variants of the bf16 fused AdamW
// streaming loop on flat arrays of Llama-3.2-1B size, to decide vector width, prefetch depth and
// occupancy for csrc/adamw.cu from measurements rather than guesses.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o adamw_variants adamw_variants.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

struct Sc { float lr_wd, beta1, beta2, eps, bc2s, step_size, clip; };

__device__ __forceinline__ float lo(uint32_t x) { return __uint_as_float(x << 16); }
__device__ __forceinline__ float hi(uint32_t x) { return __uint_as_float(x & 0xffff0000u); }
__device__ __forceinline__ uint32_t pk(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float rbf(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

template <bool MARK>
__device__ __forceinline__ void elem(float& p, float g, float& m, float& v, const Sc& a) {
  g = rbf(__fmul_rn(g, a.clip));
  p = __fmaf_rn(-a.lr_wd, p, p);
  m = __fmaf_rn(a.beta1, m, __fmaf_rn(-a.beta1, g, g));
  const float gg = __fmul_rn(g, g);
  v = __fmaf_rn(a.beta2, v, __fmaf_rn(-a.beta2, gg, gg));
  float s = __fsqrt_rn(v);
  float q;
  if (MARK) {  // correctly rounded x / c via reciprocal + residual (Markstein); rc passed in eps slot trick
    const float rc = __frcp_rn(a.bc2s);
    float q0 = __fmul_rn(s, rc);
    float r = __fmaf_rn(-q0, a.bc2s, s);
    q = __fmaf_rn(r, rc, q0);
  } else {
    q = __fdiv_rn(s, a.bc2s);
  }
  const float denom = __fadd_rn(q, a.eps);
  p = __fsub_rn(p, __fdiv_rn(__fmul_rn(a.step_size, m), denom));
}

template <int W> struct Raw { uint32_t r[W]; };

template <int W, int HINT>
__device__ __forceinline__ Raw<W> ldv(const void* p) {
  Raw<W> x;
  if constexpr (W == 4) {
    if (HINT == 0)
      asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "l"(p));
    else
      asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "l"(p));
  } else {
    if (HINT == 0)
      asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]), "=r"(x.r[4]), "=r"(x.r[5]), "=r"(x.r[6]), "=r"(x.r[7]) : "l"(p));
    else
      asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]), "=r"(x.r[4]), "=r"(x.r[5]), "=r"(x.r[6]), "=r"(x.r[7]) : "l"(p));
  }
  return x;
}
template <int W, int HINT>
__device__ __forceinline__ void stv(void* p, const Raw<W>& x) {
  if constexpr (W == 4) {
    if (HINT == 0)
      asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]) : "memory");
    else
      asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]) : "memory");
  } else {
    if (HINT == 0)
      asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]), "r"(x.r[4]), "r"(x.r[5]), "r"(x.r[6]), "r"(x.r[7]) : "memory");
    else
      asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]), "r"(x.r[4]), "r"(x.r[5]), "r"(x.r[6]), "r"(x.r[7]) : "memory");
  }
}

// W = 32-bit words per access (4 = 128-bit, 8 = 256-bit). MODE: 0 = full math, 1 = copy only
// (p,m,v = their inputs xor g: same traffic, trivial math), PF = register prefetch of next tile.
template <int W, int HINT, int MODE, bool PF, bool MARK, int CTAS>
__global__ void __launch_bounds__(256, CTAS)
k_adamw(uint16_t* __restrict__ P, const uint16_t* __restrict__ G, uint16_t* __restrict__ M, uint16_t* __restrict__ V,
        int64_t n, int64_t chunk, Sc a) {
  constexpr int E = 2 * W;                     // elements per access
  constexpr int64_t TILE = 256 * E;
  const int64_t nchunks = (n + chunk - 1) / chunk;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int64_t start = c * chunk;
    const int64_t len = (n - start) < chunk ? (n - start) : chunk;
    const int64_t ntiles = len / TILE;          // harness: sizes are multiples of TILE
    const int64_t base = start + (int64_t)threadIdx.x * E;
    Raw<W> rp, rg, rm, rv;
    if (PF && ntiles > 0) {
      rg = ldv<W, HINT>(G + base); rp = ldv<W, HINT>(P + base); rm = ldv<W, HINT>(M + base); rv = ldv<W, HINT>(V + base);
    }
    for (int64_t t = 0; t < ntiles; ++t) {
      const int64_t i = base + t * TILE;
      Raw<W> cp, cg, cm, cv;
      if (PF) {
        cp = rp; cg = rg; cm = rm; cv = rv;
        if (t + 1 < ntiles) {
          rg = ldv<W, HINT>(G + i + TILE); rp = ldv<W, HINT>(P + i + TILE);
          rm = ldv<W, HINT>(M + i + TILE); rv = ldv<W, HINT>(V + i + TILE);
        }
      } else {
        cg = ldv<W, HINT>(G + i); cp = ldv<W, HINT>(P + i); cm = ldv<W, HINT>(M + i); cv = ldv<W, HINT>(V + i);
      }
#pragma unroll
      for (int w = 0; w < W; ++w) {
        if (MODE == 1) {
          cp.r[w] ^= cg.r[w]; cm.r[w] ^= cg.r[w]; cv.r[w] ^= cg.r[w];
        } else {
          float p0 = lo(cp.r[w]), p1 = hi(cp.r[w]), m0 = lo(cm.r[w]), m1 = hi(cm.r[w]);
          float v0 = lo(cv.r[w]), v1 = hi(cv.r[w]);
          elem<MARK>(p0, lo(cg.r[w]), m0, v0, a);
          elem<MARK>(p1, hi(cg.r[w]), m1, v1, a);
          cp.r[w] = pk(p0, p1); cm.r[w] = pk(m0, m1); cv.r[w] = pk(v0, v1);
        }
      }
      stv<W, HINT>(P + i, cp); stv<W, HINT>(M + i, cm); stv<W, HINT>(V + i, cv);
    }
  }
}

template <int W, int HINT, int MODE, bool PF, bool MARK, int CTAS>
float run(const char* name, uint16_t* P, uint16_t* G, uint16_t* M, uint16_t* V, int64_t n, int64_t chunk, int sms) {
  Sc a{5e-7f, 0.9f, 0.999f, 1e-8f, 0.0316f, 5e-4f, 0.5f};
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = sms * CTAS;
  for (int i = 0; i < 3; ++i) k_adamw<W, HINT, MODE, PF, MARK, CTAS><<<grid, 256>>>(P, G, M, V, n, chunk, a);
  CK(cudaDeviceSynchronize());
  const int reps = 10;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < reps; ++i) k_adamw<W, HINT, MODE, PF, MARK, CTAS><<<grid, 256>>>(P, G, M, V, n, chunk, a);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= reps;
  printf("%-46s W=%d hint=%d pf=%d ctas=%d chunk=%6lld : %.3f ms  %.0f GB/s\n", name, W * 32, HINT, (int)PF, CTAS,
         (long long)chunk, ms, 14.0 * n / ms / 1e6);
  return ms;
}

__global__ void fill(uint16_t* p, int64_t n, uint32_t seed, float scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = (uint32_t)i * 2654435761u + seed; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    float f = ((h & 0xffff) / 65536.0f - 0.5f) * scale;
    __nv_bfloat16 b = __float2bfloat16_rn(f);
    p[i] = *reinterpret_cast<uint16_t*>(&b);
  }
}

int main() {
  const int64_t n = 1235814400LL / 8192 * 8192;
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  uint16_t *P, *G, *M, *V;
  CK(cudaMalloc(&P, n * 2)); CK(cudaMalloc(&G, n * 2)); CK(cudaMalloc(&M, n * 2)); CK(cudaMalloc(&V, n * 2));
  fill<<<sms * 8, 256>>>(P, n, 1, 0.04f); fill<<<sms * 8, 256>>>(G, n, 2, 0.02f);
  CK(cudaMemset(M, 0, n * 2)); CK(cudaMemset(V, 0, n * 2));
  CK(cudaDeviceSynchronize());
  printf("n=%lld sms=%d\n", (long long)n, sms);
  const int64_t C = 65536;
  // memory-system ceiling for this 4-read / 3-write mix (no math)
  run<4, 0, 1, false, false, 4>("copy-mix 128b cs", P, G, M, V, n, C, sms);
  run<4, 1, 1, false, false, 4>("copy-mix 128b na", P, G, M, V, n, C, sms);
  run<8, 0, 1, false, false, 4>("copy-mix 256b na", P, G, M, V, n, C, sms);
  run<8, 1, 1, false, false, 4>("copy-mix 256b na+efl2", P, G, M, V, n, C, sms);
  run<8, 1, 1, true, false, 4>("copy-mix 256b na+efl2 prefetch", P, G, M, V, n, C, sms);
  run<8, 1, 1, false, false, 8>("copy-mix 256b na+efl2 8cta", P, G, M, V, n, C, sms);
  run<8, 1, 1, false, false, 2>("copy-mix 256b na+efl2 2cta", P, G, M, V, n, C, sms);
  // full math
  run<4, 0, 0, false, false, 4>("adamw 128b cs (current)", P, G, M, V, n, C, sms);
  run<4, 0, 0, true, false, 4>("adamw 128b cs prefetch", P, G, M, V, n, C, sms);
  run<4, 0, 0, true, false, 3>("adamw 128b cs prefetch 3cta", P, G, M, V, n, C, sms);
  run<8, 1, 0, false, false, 4>("adamw 256b efl2", P, G, M, V, n, C, sms);
  run<8, 1, 0, false, false, 3>("adamw 256b efl2 3cta", P, G, M, V, n, C, sms);
  run<8, 1, 0, true, false, 3>("adamw 256b efl2 prefetch 3cta", P, G, M, V, n, C, sms);
  run<8, 1, 0, true, false, 2>("adamw 256b efl2 prefetch 2cta", P, G, M, V, n, C, sms);
  run<8, 1, 0, true, true, 3>("adamw 256b efl2 prefetch 3cta markstein", P, G, M, V, n, C, sms);
  run<8, 1, 0, false, true, 4>("adamw 256b efl2 4cta markstein", P, G, M, V, n, C, sms);
  run<4, 0, 0, false, true, 4>("adamw 128b cs markstein", P, G, M, V, n, C, sms);
  run<4, 0, 0, true, true, 4>("adamw 128b cs prefetch markstein", P, G, M, V, n, C, sms);
  run<8, 1, 0, true, true, 3>("adamw 256b pf 3cta mark chunk16k", P, G, M, V, n, 16384, sms);
  run<8, 1, 0, true, true, 3>("adamw 256b pf 3cta mark chunk256k", P, G, M, V, n, 262144, sms);
  return 0;
}

// Device-side self-checks of the hand-written arithmetic in adamw_math.cuh / common.cuh against the
// IEEE intrinsics they replace. Not on the hot path: tests/test_gpu_parity.py calls these through the C
// ABI to show, on the GPU that runs the kernels, that inside the packed path's box of magnitudes
//   * the written-out square root equals __fsqrt_rn        (every float in the box, exhaustively),
//   * the whole denominator sqrt(v)/c + eps equals the scalar form for a given (c, eps),
//   * the written-out quotient p - a/b equals __fsub_rn(p, __fdiv_rn(a, b))   (pseudo-random sweep),
//   * div_by_const(x, c, 1/c) equals __fdiv_rn(x, c) for every significand of x, for a list of c.
// Each check counts mismatches into a device counter and records the first offending input.
#include "launch.cuh"
#include "adamw_math.cuh"

namespace ddpb200 {

__device__ __forceinline__ void report(unsigned long long* out, unsigned long long a, unsigned long long b) {
  if (atomicAdd(out, 1ull) == 0ull) { out[1] = a; out[2] = b; }
}

// v = every bit pattern in [first, first + count): -neg_sqrt_scaled(v) * 2^-24 vs __fsqrt_rn(v), and, with
// c > 0, the packed denominator vs the scalar one.
__global__ void __launch_bounds__(256)
check_sqrt_kernel(uint32_t first, uint64_t count, float c, float eps, unsigned long long* out) {
  const float rc = 1.0f / c;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t u0 = first + (uint32_t)i;
    const uint32_t u1 = u0 ^ 1u;                                        // the pair's other lane: a neighbour
    const float2 v = make_float2(__uint_as_float(u0), __uint_as_float(u1));
    const float2 nr = neg_sqrt_scaled(v);
    const float want0 = __fsqrt_rn(v.x), want1 = __fsqrt_rn(v.y);
    const float got0 = __fmul_rn(-nr.x, 0x1p-24f), got1 = __fmul_rn(-nr.y, 0x1p-24f);
    if (__float_as_uint(got0) != __float_as_uint(want0)) report(out, u0, __float_as_uint(got0));
    if (__float_as_uint(got1) != __float_as_uint(want1)) report(out, u1, __float_as_uint(got1));
    if (c > 0.f) {
      const float2 q0 = __fmul2_rn(nr, splat(rc));
      const float2 rem = __ffma2_rn(q0, splat(-c), nr);
      const float2 nd = __ffma2_rn(rem, splat(rc), q0);
      const float2 nb = __ffma2_rn(nd, splat(0x1p-24f), splat(-eps));
      const float d0 = __fadd_rn(div_by_const(want0, c, rc), eps);
      if (__float_as_uint(-nb.x) != __float_as_uint(d0)) report(out, u0, __float_as_uint(-nb.x));
    }
  }
}

__device__ __forceinline__ uint32_t mix(uint64_t x) {                    // splitmix64 -> 32 bits
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return (uint32_t)((x ^ (x >> 31)) >> 16);
}

// a float with a random significand and sign and an exponent field drawn uniformly from [elo, ehi]
__device__ __forceinline__ float rnd_float(uint64_t k, uint32_t elo, uint32_t ehi, bool sign) {
  const uint32_t r = mix(k), e = elo + mix(k ^ 0x55aa55aa12345678ull) % (ehi - elo + 1);
  return __uint_as_float((sign ? (r & 0x80000000u) : 0u) | (e << 23) | (r & 0x007fffffu));
}

// count pseudo-random (p, a, b) triples inside the box: |a| in {0} U [2^-149, 2^20), b in [2^-40, 2^21),
// |p| in [2^-100, 2^40); every 61st a is zero and every 7th a is denormal.
__global__ void __launch_bounds__(256)
check_div_kernel(uint64_t seed, uint64_t count, unsigned long long* out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t k = seed + 3 * i;
    float a0 = rnd_float(k, 1, 127 + 19, true), a1 = rnd_float(k ^ 0xabcdefull, 0, 127 + 19, true);
    if (i % 7 == 0) a0 = __uint_as_float(__float_as_uint(a0) & 0x807fffffu);     // denormal (or zero)
    if (i % 61 == 0) a1 = 0.f;
    const float b0 = rnd_float(k + 1, 127 - 40, 127 + 20, false), b1 = rnd_float(k + 1 ^ 0x1234567ull, 127 - 40, 127 + 20, false);
    const float p0 = rnd_float(k + 2, 127 - 100, 127 + 39, true), p1 = rnd_float(k + 2 ^ 0x7654321ull, 127 - 100, 127 + 39, true);
    const float2 got = sub_quotient(make_float2(p0, p1), make_float2(a0, a1), make_float2(-b0, -b1));
    const float w0 = __fsub_rn(p0, __fdiv_rn(a0, b0)), w1 = __fsub_rn(p1, __fdiv_rn(a1, b1));
    if (__float_as_uint(got.x) != __float_as_uint(w0)) report(out, ((unsigned long long)__float_as_uint(a0) << 32) | __float_as_uint(b0), __float_as_uint(p0));
    if (__float_as_uint(got.y) != __float_as_uint(w1)) report(out, ((unsigned long long)__float_as_uint(a1) << 32) | __float_as_uint(b1), __float_as_uint(p1));
  }
}

// for every c in cs[0..nc) and every significand of x in [1, 2): div_by_const(x, c, RN(1/c)) vs __fdiv_rn(x, c)
__global__ void __launch_bounds__(256)
check_div_by_const_kernel(const float* __restrict__ cs, int nc, unsigned long long* out) {
  for (int j = blockIdx.y; j < nc; j += gridDim.y) {
    const float c = cs[j], rc = __frcp_rn(c);
    for (uint32_t mant = blockIdx.x * blockDim.x + threadIdx.x; mant < (1u << 23); mant += gridDim.x * blockDim.x) {
      const float x = __uint_as_float(0x3f800000u | mant);
      const float got = div_by_const(x, c, rc), want = __fdiv_rn(x, c);
      if (__float_as_uint(got) != __float_as_uint(want)) report(out, ((unsigned long long)__float_as_uint(c) << 32) | __float_as_uint(x), __float_as_uint(got));
    }
  }
}

}  // namespace ddpb200

using namespace ddpb200;

extern "C" DDPB200_API int ddpb200_selfcheck_sqrt(uint32_t first_bits, uint64_t count, float c, float eps,
                                                  unsigned long long* out3, void* stream) {
  if (!out3) return DDPB200_E_INVALID;
  if (count == 0) return 0;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) return err;
  check_sqrt_kernel<<<sms * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(first_bits, count, c, eps, out3);
  return (int)cudaGetLastError();
}

extern "C" DDPB200_API int ddpb200_selfcheck_div(uint64_t seed, uint64_t count, unsigned long long* out3, void* stream) {
  if (!out3) return DDPB200_E_INVALID;
  if (count == 0) return 0;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) return err;
  check_div_kernel<<<sms * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(seed, count, out3);
  return (int)cudaGetLastError();
}

extern "C" DDPB200_API int ddpb200_selfcheck_div_by_const(const float* constants, int32_t n, unsigned long long* out3,
                                                          void* stream) {
  if (!out3 || n < 0 || (n > 0 && !constants)) return DDPB200_E_INVALID;
  if (n == 0) return 0;
  int err = 0;
  const int sms = sm_count_cached(&err);
  if (err) return err;
  dim3 grid(sms, n < 64 ? n : 64);
  check_div_by_const_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(constants, n, out3);
  return (int)cudaGetLastError();
}

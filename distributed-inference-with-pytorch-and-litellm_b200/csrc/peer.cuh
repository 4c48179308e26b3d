// Peer-memory plumbing shared by the fused allreduce (fused_allreduce.cu) and the sharded update
// (sharded_update.cu): the device-side peer map and the NVSwitch multicast (multimem) accessors.
#pragma once
#include "common.cuh"

namespace ddpb200 {

struct PeerMap {                        // device-side copy of ddpb200_peer_map
  char* arena[DDPB200_MAX_PEERS];
  float* partials[DDPB200_MAX_PEERS];
  char* mc;
  int rank, world;
};

__device__ __forceinline__ Raw<4> mc_ld_reduce_bf16x8(const void* p) {
  Raw<4> x;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "l"(__cvta_generic_to_global(p)) : "memory");
  return x;
}
__device__ __forceinline__ Raw<4> mc_ld_reduce_f32x4(const void* p) {
  Raw<4> x;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(x.r[0]), "=r"(x.r[1]), "=r"(x.r[2]), "=r"(x.r[3]) : "l"(__cvta_generic_to_global(p)) : "memory");
  return x;
}
__device__ __forceinline__ void mc_st_128(void* p, const Raw<4>& x) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(__cvta_generic_to_global(p)), "r"(x.r[0]), "r"(x.r[1]), "r"(x.r[2]), "r"(x.r[3]) : "memory");
}

// fills the device-side peer map from the ABI struct; false if it is malformed
inline bool make_peer_map(const ddpb200_peer_map* peers, bool need_partials, PeerMap* pm) {
  if (!peers || peers->world < 1 || peers->world > DDPB200_MAX_PEERS || peers->rank < 0 ||
      peers->rank >= peers->world)
    return false;
  for (int p = 0; p < DDPB200_MAX_PEERS; ++p) {
    pm->arena[p] = p < peers->world ? static_cast<char*>(peers->arena[p]) : nullptr;
    pm->partials[p] = p < peers->world ? peers->partials[p] : nullptr;
    if (p < peers->world && (!pm->arena[p] || (need_partials && !pm->partials[p]))) return false;
  }
  pm->mc = static_cast<char*>(peers->arena_multicast);
  pm->rank = peers->rank;
  pm->world = peers->world;
  return true;
}


}  // namespace ddpb200

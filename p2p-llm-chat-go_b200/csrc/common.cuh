// common.cuh — device-side helpers shared by the sm_100a kernels (PTX wrappers, number formats).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"

#define DEVINL __device__ __forceinline__

namespace b200 {


// ---- shared-memory / mbarrier / bulk-copy (TMA engine) wrappers -------------------------------------
DEVINL uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

DEVINL void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
DEVINL void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
DEVINL void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
DEVINL void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
DEVINL bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
DEVINL void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
DEVINL uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
DEVINL uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 1-D bulk async copy global -> shared through the TMA unit (SASS: UBLKCP); completes on an mbarrier.
DEVINL void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
// asks L2 to fetch [p, p+bytes) from HBM (bytes % 16 == 0); a hint, nothing waits on it
DEVINL void l2_prefetch(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// 16-byte read-only load that asks L2 to keep the line (evict_last): small tensors read by every CTA of every step
// (norm weights) must survive the gigabytes of evict_first weight traffic streaming past them
DEVINL float4 ldg_keep(const float4* p, uint64_t policy) {
  float4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p), "l"(policy));
  return v;
}
DEVINL void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Programmatic dependent launch: everything before pdl_wait() overlaps the previous kernel's tail.
DEVINL void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
DEVINL void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

DEVINL unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---- number formats ---------------------------------------------------------------------------------
DEVINL float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
DEVINL float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
DEVINL uint16_t f32_to_bf16_bits(float f) {  // RNE, same integer recipe as the CPU oracle
  uint32_t u = __float_as_uint(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
DEVINL float bf16_round(float f) { return __uint_as_float((uint32_t)f32_to_bf16_bits(f) << 16); }
// silu(g) * u of the tensor-core (bf16) path, the oracle's formula operation for operation (expf, IEEE divide).  The
// approximate forms (__expf, __fdividef: ~1e-6 relative) save 0.1 ms of a 9.5 ms prefill, but at the BASELINE
// configuration the Q4_K_M greedy continuation then left the oracle's at token 52 on a near-tie (the tolerance test
// still passed); with these it is identical for all 128 tokens, so they stay.
DEVINL float silu_mul(float g, float u) { return (g / (1.0f + expf(-g))) * u; }
DEVINL float fp16_bits_to_f32(uint16_t h) { return __half2float(__ushort_as_half(h)); }

DEVINL float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
DEVINL float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// order-preserving float -> uint key; (key<<32 | ~idx) makes argmax a single 64-bit atomicMax with
// lowest-index tie-break.
DEVINL unsigned long long argmax_key(float v, uint32_t idx) {
  uint32_t u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned long long)(0xffffffffu - idx);
}
DEVINL uint32_t argmax_key_index(unsigned long long key) { return 0xffffffffu - (uint32_t)(key & 0xffffffffull); }

DEVINL int dp4a_ss(int a, int b, int c) {
  int d;
  asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
DEVINL int dp4a_su(int a, uint32_t b, int c) {  // signed a-bytes x unsigned b-bytes
  int d;
  asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

}  // namespace b200

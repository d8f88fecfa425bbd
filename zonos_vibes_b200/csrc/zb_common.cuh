// Shared device/host helpers for libzonos_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/zonos_b200.h"

#define ZB_WARP 32
#define ZB_NUM_SMS 148          // B200: 2 dies x 74 SMs
#define ZB_PAGE 128             // KV page size in tokens
#define ZB_DH 128               // head dim supported by the attention kernels

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------------
// host side error plumbing
// ---------------------------------------------------------------------------------------------
struct ZbErr {
  std::string msg;
  int fail(int code, const char* fmt, ...) __attribute__((format(printf, 3, 4)));
};

#define ZB_CUDA(err_obj, expr)                                                                      \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess)                                                                          \
      return (err_obj).fail(ZB_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,              \
                            cudaGetErrorString(_e));                                                \
  } while (0)

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
// Checked build (-DZB_BOUNDS, scripts/build_checked.sh -> libzonos_b200_checked.so, loaded with ZB_LIB=...): index and capacity
// assertions at the address computations of the kernels; a violation prints where and traps (the launch fails, the test fails).
// compute-sanitizer is closed on the GPU pool this was developed on; this is the substitute.  The production build has none of it.
#ifdef ZB_BOUNDS
#define ZB_ASSERT(c)                                                                                                              \
  do {                                                                                                                            \
    if (!(c)) {                                                                                                                   \
      printf("ZB_ASSERT failed: %s  (%s:%d, block %d,%d,%d thread %d)\n", #c, __FILE__, __LINE__, (int)blockIdx.x, (int)blockIdx.y, \
             (int)blockIdx.z, (int)threadIdx.x);                                                                                  \
      __trap();                                                                                                                   \
    }                                                                                                                             \
  } while (0)
#else
#define ZB_ASSERT(c) do { } while (0)
#endif
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// two packed bf16 (little end = element 0) -> fp32; bf16->fp32 is a 16-bit shift.
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// streaming 128-bit load: read-only path, do not allocate in L1 (weights are read exactly once per step).
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// dot of 8 bf16 weights (one uint4) with 8 bf16 activations (one uint4), fp32 accumulate.
__device__ __forceinline__ float dot8(const uint4& w, const uint4& x, float acc) {
  acc = fmaf(bf16lo(w.x), bf16lo(x.x), acc);
  acc = fmaf(bf16hi(w.x), bf16hi(x.x), acc);
  acc = fmaf(bf16lo(w.y), bf16lo(x.y), acc);
  acc = fmaf(bf16hi(w.y), bf16hi(x.y), acc);
  acc = fmaf(bf16lo(w.z), bf16lo(x.z), acc);
  acc = fmaf(bf16hi(w.z), bf16hi(x.z), acc);
  acc = fmaf(bf16lo(w.w), bf16lo(x.w), acc);
  acc = fmaf(bf16hi(w.w), bf16hi(x.w), acc);
  return acc;
}

// Programmatic dependent launch (PDL).  Both are no-ops when the kernel was launched without the
// programmatic-stream-serialization attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait_prior_grid() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// L2 prefetch of a byte range (fire and forget; size multiple of 16, address 16-aligned).
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// Barrier policies for code shared between stand-alone kernels (whole CTA) and the persistent decode kernel, where
// the 256 consumer threads synchronise on named barrier 1 while the producer warp runs its own loop.
struct SyncCta {
  static __device__ __forceinline__ void sync() { __syncthreads(); }
};
struct SyncConsumers {
  static __device__ __forceinline__ void sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
};

// ---------------------------------------------------------------------------------------------
// device-resident generation state (one per context); every step kernel reads it, k_update writes it.
// ---------------------------------------------------------------------------------------------
struct ZbDevState {
  int pos;             // KV write index == reference seqlen_offset (same for all rows)
  int col;             // delayed-codes column that is the NEXT input (== last written column)
  int done;            // 1 once every utterance has remaining_steps <= 0: step kernels return at once
  int step;            // decode-loop iteration counter (Philox offset)
  int L;               // length of the delayed buffer (T + n_q)
  int B;               // utterances
  int apply_eos_bias;  // model.py:266-267,280 (off for the prefill sample, model.py:255-256)
  int use_history;     // repetition penalty history available (off for the prefill sample)
  int eos_logic;       // run the EOS wind-down of model.py:283-301 in k_update
  int compat_scatter;  // masked_scatter_ sequential routing
  float cfg_scale;
  unsigned long long seed;
  zb_sampling sp;
};

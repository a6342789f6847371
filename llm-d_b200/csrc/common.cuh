// Shared device helpers for libb200ep (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200ep.h"

#ifndef __CUDA_ARCH__
#define B200EP_HOST 1
#endif

namespace b200ep {

constexpr int kMaxRanks = B200EP_MAX_RANKS;
constexpr int kMaxLocalExperts = B200EP_MAX_LOCAL_EXPERTS;
constexpr int kMaxTopk = B200EP_MAX_TOPK;
constexpr int kNumSlots = B200EP_NUM_SLOTS;
constexpr int kGroup = 128;  // FP8 quantisation group (DEEPEP_QUANT_BLOCK_SIZE, deepep_ll.py:28)

// ---- symmetric control block (first kCtrlBytes of every rank's region; peers write it)
constexpr int kCombChunks = 8;     // the overlapped combine signals completion per chunk of source tokens
// Chunk c of a source rank with T tokens = tokens [comb_chunk_begin(T, c), comb_chunk_begin(T, c + 1)): about a
// sixth each, then smaller, so that what is left to reduce after the LAST flag is 1/32 of the tokens (the
// reduction of the early chunks runs under the transfer; the last one is pure tail).
__host__ __device__ inline int comb_chunk_begin(int T, int c) {
  const int frac32[kCombChunks + 1] = {0, 5, 10, 15, 20, 24, 28, 31, 32};
  if (c >= kCombChunks) return T;
  return (int)(((long long)T * frac32[c] / 32) / 32 * 32);  // multiple of 32 tokens
}
enum FlagCat { kFlagHtNotify = 0, kFlagHtData, kFlagHtComb, kFlagLlCount, kFlagLlData, kFlagLlComb, kFlagHtCombChunk0,
               kNumFlagCat = kFlagHtCombChunk0 + kCombChunks };
constexpr size_t kFlagStride = 128;  // one flag per 128-byte line
constexpr size_t kCtrlFlagsOff = 0;
constexpr size_t kCtrlFlagsBytes = (size_t)kNumFlagCat * kNumSlots * kMaxRanks * kFlagStride;
// counts travel as self-validating 8-byte words (call tag << 32 | count): no flag, no system fence
constexpr size_t kCtrlHtTokOff = 32768;      // u64 [slot][src]: tokens of the source rank in this call
constexpr size_t kCtrlHtCntRankOff = 65536;  // u64 [slot][src][kMaxRanks]
constexpr size_t kCtrlHtCntExpOff = 81920;   // u64 [slot][src][kMaxLocalExperts]
constexpr size_t kCtrlLlCntOff = kCtrlHtCntExpOff + (size_t)kNumSlots * kMaxRanks * kMaxLocalExperts * 8;
constexpr size_t kCtrlBytes = 1 << 20;
static_assert(kCtrlFlagsBytes <= kCtrlHtTokOff, "flags overflow");
static_assert(kCtrlHtCntRankOff + (size_t)kNumSlots * kMaxRanks * kMaxRanks * 8 <= kCtrlHtCntExpOff, "rank counts overflow");
// low-latency counts: u64 [slot][src][kMaxLocalExperts] words (call tag << 32 | count), each word valid by itself
static_assert(kCtrlLlCntOff % 8 == 0 && kCtrlLlCntOff + (size_t)kNumSlots * kMaxRanks * kMaxLocalExperts * 8 <= kCtrlBytes, "ctrl overflow");

__host__ __device__ inline size_t flag_off(int cat, int slot, int src) {
  return kCtrlFlagsOff + (((size_t)cat * kNumSlots + slot) * kMaxRanks + src) * kFlagStride;
}

// ---- per-slot metadata in PRIVATE device memory (never written by peers)
struct HtMeta {
  int32_t rank_prefix[kMaxRanks + 1];  // rows [p[s], p[s+1]) of my receive region came from rank s
  int32_t send_base[kMaxRanks];        // first row I own in rank d's receive region
  int32_t num_recv;
  int32_t pad[5];
  int32_t send_count[kMaxRanks];       // tokens I send to rank d
  int32_t src_tokens[kMaxRanks];       // tokens rank s holds in this call (its chunk boundaries follow from it)
  int32_t recv_per_expert[kMaxLocalExperts];
};

struct Peers {
  char* base[kMaxRanks];
};

// Region layout of one high-throughput slot (identical on every rank).
struct HtPlan {
  size_t off_x, off_scales, off_idx, off_w, off_src, off_stage;
  size_t off_side;      // [cap_rows][side_bytes]: what travels with a row besides its payload (see ht.cu)
  int side_bytes;       // K*12 + 4 + scale_bytes rounded up to 16
  size_t off_qstage;    // local only: quantised rows [t_cap][qrow_bytes] staged before the ordered push
  size_t off_sendlist;  // local only: int32 [world][t_cap], token sent to rank d at position p
  int qrow_bytes;       // row_bytes + scale_bytes rounded up to 16
  int t_cap;       // tokens per rank the slot can take
  int cap_rows;    // world * t_cap receive rows
  int row_bytes;   // payload bytes per token (H for fp8, 2H for bf16)
  int scale_bytes; // H/128*4 or 0
  int H, K, x_kind;
};

// Region layout of one low-latency slot.
struct LlPlan {
  size_t off_x, off_scales, off_src, off_stage;
  int cap;  // world * max_tok slots per local expert
  int g4, cap4;  // packed-UE8M0 scales: int32 [e_local][g4 = ceil(H/512)][cap4 = cap rounded up to 4]
  int e_local;
  int row_bytes, scale_bytes;
  int H, K, max_tok, use_fp8;
};

// ------------------------------------------------------------------ PTX wrappers
#ifdef __CUDACC__

__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ unsigned atom_add_acq_rel_gpu(unsigned* p, unsigned v) {
  unsigned r;
  asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], %2;" : "=r"(r) : "l"(p), "r"(v) : "memory");
  return r;
}
__device__ __forceinline__ uint64_t ld_acquire_gpu_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int4 ld_nc_v4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int4 ld_cg_v4(const void* p) {  // L2 only: data written by peers / other kernels
  int4 r;
  asm volatile("ld.global.cg.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_na_v4(void* p, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-stream-serialization
// attribute may start while its predecessor in the stream is still running; it must not touch anything
// the predecessor writes before pdl_wait() (which returns when the predecessor has completed and its
// writes are visible).  The predecessor calls pdl_launch_dependents() to let the dependent grid start
// early.  Both are no-ops in a normally launched kernel.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// TMA bulk copy shared -> global (local or peer), tracked by the thread's bulk async-group.
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all of this thread's groups have finished READING shared memory (buffer reusable)
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// all but the most recent group have finished reading shared memory (double-buffered stages)
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
// all of this thread's groups are complete (writes performed)
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// at most N of this thread's most recent groups still pending (reading / at all)
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
// TMA bulk copy global -> shared, completion (bytes) signalled on an mbarrier of this CTA
__device__ __forceinline__ void bulk_g2s(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(sdst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// order this thread's generic-proxy shared-memory writes before later async-proxy (TMA) reads
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Spin until *flag >= want (flag written by another GPU).  Traps instead of hanging forever.
__device__ __forceinline__ void wait_flag_sys(const uint64_t* flag, uint64_t want, int what, int me, int peer) {
  long long t0 = clock64();
  while (ld_acquire_sys_u64(flag) < want) {
    if (clock64() - t0 > 60000000000LL) {  // ~30 s at 2 GHz
      printf("b200ep: rank %d timed out waiting for flag kind %d from rank %d (want %llu, have %llu)\n", me, what,
             peer, (unsigned long long)want, (unsigned long long)ld_acquire_sys_u64(flag));
      __trap();
    }
    __nanosleep(20);
  }
}

// A word another GPU writes as (tag << 32 | value) in ONE 8-byte store needs no flag and no fence: spin until the
// tag is the expected one, the value comes with it.  Traps instead of hanging forever.
__device__ __forceinline__ int wait_tagged_word(const uint64_t* word, unsigned tag, int what, int me, int peer) {
  uint64_t v = ld_relaxed_sys_u64(word);
  if ((unsigned)(v >> 32) == tag) return (int)(uint32_t)v;
  const long long t0 = clock64();
  for (;;) {
    v = ld_relaxed_sys_u64(word);
    if ((unsigned)(v >> 32) == tag) return (int)(uint32_t)v;
    if (clock64() - t0 > 60000000000LL) {
      printf("b200ep: rank %d timed out waiting for a count word (kind %d) from rank %d (want tag %u, have %llu)\n", me, what,
             peer, tag, (unsigned long long)v);
      __trap();
    }
  }
}

// Call epochs live in PRIVATE device memory (LocalScratch::epoch) and are advanced by the kernels
// themselves, never passed as launch arguments: a captured CUDA graph replays correctly.
//
// "All CTAs of this kernel have finished their stores" -> signal every peer with the next epoch.
// Call with all threads of the CTA after the CTA's last store (and after bulk_wait_all()).
// counter: private device word, 0 before the kernel, reset to 0 by the last CTA.
__device__ __forceinline__ void cta_done_then_signal(unsigned* counter, unsigned long long* epoch_ptr,
                                                     const Peers& peers, int world, int cat, int slot, int me) {
  // A system-scope fence costs ~5 us on B200 (measured: an empty kernel whose CTAs each fence.sys),
  // so exactly ONE is issued per kernel, by the last CTA.  Every other CTA releases its stores at
  // gpu scope through the counter; the PTX memory model's causality order is transitive across
  // scopes: stores -> bar.sync -> atom.acq_rel.gpu (CTA i) -> atom.acq_rel.gpu (last CTA)
  // -> fence.acq_rel.sys -> st.relaxed.sys flag -> ld.acquire.sys on the peer.
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atom_add_acq_rel_gpu(counter, 1u);
    if (prev == gridDim.x - 1) {
      *counter = 0;
      const uint64_t epoch = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
      fence_acq_rel_sys();
      for (int d = 0; d < world; ++d)
        st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(peers.base[d] + flag_off(cat, slot, me)), epoch);
      *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) = epoch;
    }
  }
}

// One CTA publishes to all peers: stores -> bar.sync -> (thread 0) one fence.acq_rel.sys -> W relaxed flags.
__device__ __forceinline__ void single_cta_signal(unsigned long long* epoch_ptr, const Peers& peers, int world, int cat,
                                                  int slot, int me) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint64_t epoch = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
    fence_acq_rel_sys();
    for (int d = 0; d < world; ++d)
      st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(peers.base[d] + flag_off(cat, slot, me)), epoch);
    *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) = epoch;
  }
}

// All CTAs wait for every source rank's flag (flags live in MY region).  The epoch to wait for is
// the one this rank's own send of the same call already stored (stream order guarantees that).
__device__ __forceinline__ void cta_wait_all_sources(const char* my_base, int world, int cat, int slot, int me,
                                                     const unsigned long long* epoch_ptr) {
  if ((int)threadIdx.x < world) {
    const uint64_t epoch = *reinterpret_cast<const volatile unsigned long long*>(epoch_ptr);
    wait_flag_sys(reinterpret_cast<const uint64_t*>(my_base + flag_off(cat, slot, threadIdx.x)), epoch, cat, me,
                  threadIdx.x);
  }
  __syncthreads();
}

// ---- FP8 e4m3 quantisation of 8 bf16 values held in an int4 (arithmetic "div", see oracle/fp8.py)
__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ float absmax8(const float (&f)[8]) {
  float m = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) m = fmaxf(m, fabsf(f[i]));
  return m;
}
__device__ __forceinline__ float group_scale(float amax, bool ue8m0) {
  // RN(max(amax, 1e-10) / 448) by Markstein's correction with the constant r = RN(1/448)
  // (448 = 1.75 * 2^8: its significand is not all ones, so the corrected quotient is exact).
  const float a = fmaxf(amax, 1e-10f);
  const float r448 = 1.0f / 448.0f;  // folded at compile time, correctly rounded
  const float q0 = __fmul_rn(a, r448);
  float s = __fmaf_rn(__fmaf_rn(-q0, 448.0f, a), r448, q0);
  if (ue8m0) {  // 2^ceil(log2 max(s, 1e-10)), exact on the bit pattern (vLLM clamps the scale again here)
    uint32_t u = __float_as_uint(fmaxf(s, 1e-10f));
    u = (u & 0x7f800000u) + ((u & 0x007fffffu) ? 0x00800000u : 0u);
    s = __uint_as_float(u);
  }
  return s;
}
// x / s rounded to nearest, without an IEEE division per element: r = RN(1/s) once per group,
// q0 = x*r, e = fma(-q0, s, x) (exact residual), q = fma(e, r, q0)  (Markstein).  Equal to
// __fdiv_rn(x, s) for every input that can reach a non-zero e4m3 code (tests/test_div_trick.py;
// byte parity with vLLM's CUDA op is checked on the GPU in tests/test_gpu_layout_quant.py).
__device__ __forceinline__ float div_by_scale(float x, float s, float r) {
  const float q0 = __fmul_rn(x, r);
  const float e = __fmaf_rn(-q0, s, x);
  return copysignf(__fmaf_rn(e, r, q0), x);  // keep the sign of a zero quotient, as the division does
}
__device__ __forceinline__ uint2 quant8(const float (&f)[8], float s) {
  const float r = __frcp_rn(s);
  uint32_t o[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float2 a = make_float2(div_by_scale(f[4 * i], s, r), div_by_scale(f[4 * i + 1], s, r));
    float2 b = make_float2(div_by_scale(f[4 * i + 2], s, r), div_by_scale(f[4 * i + 3], s, r));
    uint32_t lo = __nv_cvt_float2_to_fp8x2(a, __NV_SATFINITE, __NV_E4M3);
    uint32_t hi = __nv_cvt_float2_to_fp8x2(b, __NV_SATFINITE, __NV_E4M3);
    o[i] = lo | (hi << 16);
  }
  return make_uint2(o[0], o[1]);
}
// ---- 32 bytes (16 bf16) per lane: one LDG.E.256 per lane, 8 lanes per 128-element group.  Halves the
// per-group overhead (reciprocal, scale, shuffles) per element compared with 16-byte lanes.
struct alignas(32) V8 {
  uint32_t w[8];
};
__device__ __forceinline__ V8 ld_nc_v8(const void* p) {
  V8 r;
  asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]),
                 "=r"(r.w[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ float group_amax8(float m) {
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
  return m;
}
// |x| maximum over the 16 bf16 values of `v`.  Done on packed bf16 pairs (HMNMX2 with |.| operands:
// a bf16 maximum is exact, NaNs are skipped as fmaxf does), widened to fp32 once at the end:
// ~18 instructions instead of ~48 for unpack + fabs + fmax per element.
__device__ __forceinline__ float absmax16(const V8& v) {
  __nv_bfloat162 m = __habs2(*reinterpret_cast<const __nv_bfloat162*>(&v.w[0]));
#pragma unroll
  for (int i = 1; i < 8; ++i) m = __hmax2(m, __habs2(*reinterpret_cast<const __nv_bfloat162*>(&v.w[i])));
  const uint32_t u = *reinterpret_cast<const uint32_t*>(&m);
  return fmaxf(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
// Quantise the 16 bf16 values of `v` with the scale of their group -> 16 e4m3 bytes.
// x / s by Markstein's correction (see div_by_scale) on PAIRS of values with the packed fp32
// instructions of sm_100 (FMUL2 / FFMA2, round to nearest per lane, i.e. the same results as the
// scalar forms): q0 = x*r, e = fma(q0, -s, x), q = fma(e, r, q0).  The sign of a zero quotient
// (x = -0, or an underflowing negative x) is restored on the e4m3 bytes instead of with one
// copysign per value: byte |= sign bit of the source bf16.
__device__ __forceinline__ uint4 quant16(const V8& v, float s, float r) {  // r = RN(1/s)
  const float2 r2 = make_float2(r, r), ns2 = make_float2(-s, -s);
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t wa = v.w[2 * i], wb = v.w[2 * i + 1];
    const float2 a = make_float2(__uint_as_float(wa << 16), __uint_as_float(wa & 0xffff0000u));
    const float2 b = make_float2(__uint_as_float(wb << 16), __uint_as_float(wb & 0xffff0000u));
    const float2 qa0 = __fmul2_rn(a, r2), qb0 = __fmul2_rn(b, r2);
    const float2 qa = __ffma2_rn(__ffma2_rn(qa0, ns2, a), r2, qa0);
    const float2 qb = __ffma2_rn(__ffma2_rn(qb0, ns2, b), r2, qb0);
    const uint32_t lo = __nv_cvt_float2_to_fp8x2(qa, __NV_SATFINITE, __NV_E4M3);
    const uint32_t hi = __nv_cvt_float2_to_fp8x2(qb, __NV_SATFINITE, __NV_E4M3);
    // sign bytes: bf16 sign bits live in bytes 1 and 3 of each source word
    o[i] = (lo | (hi << 16)) | (__byte_perm(wa, wb, 0x7531) & 0x80808080u);
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ uint4 quant16(const V8& v, float s) { return quant16(v, s, __frcp_rn(s)); }
// packed |x| maximum of two vectors (32 values) in one chain, widened once
__device__ __forceinline__ float absmax32(const V8& a, const V8& b) {
  __nv_bfloat162 m = __hmax2(__habs2(*reinterpret_cast<const __nv_bfloat162*>(&a.w[0])),
                             __habs2(*reinterpret_cast<const __nv_bfloat162*>(&b.w[0])));
#pragma unroll
  for (int i = 1; i < 8; ++i) {
    m = __hmax2(m, __habs2(*reinterpret_cast<const __nv_bfloat162*>(&a.w[i])));
    m = __hmax2(m, __habs2(*reinterpret_cast<const __nv_bfloat162*>(&b.w[i])));
  }
  const uint32_t u = *reinterpret_cast<const uint32_t*>(&m);
  return fmaxf(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
// amax over the 4 lanes that hold one 128-element group (32 values per lane)
__device__ __forceinline__ float group_amax4(float m) {
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
  return m;
}
// amax over the 16 lanes that hold one 128-element group
__device__ __forceinline__ float group_amax16(float m) {
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
  return m;
}

#endif  // __CUDACC__

// ------------------------------------------------------------------ host side
void set_error(const char* fmt, ...);
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) only when a kernel needs more than it already has:
// the call costs several microseconds and the data kernels are launched every step.
cudaError_t ensure_dynamic_smem(const void* kernel, int device, int bytes);
#define B200EP_CUDA(call)                                                                        \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      b200ep::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return B200EP_ERR_CUDA;                                                                    \
    }                                                                                            \
  } while (0)

}  // namespace b200ep

// Private scratch layout (device memory owned by the context, not peer visible)
namespace b200ep {
constexpr int kMaxLayoutBlocks = 1024;
constexpr int kMaxExperts = kMaxRanks * kMaxLocalExperts;
struct LocalScratch {
  unsigned done[128];                                        // CTA-completion counters / work queues, one set per kernel family
  // routing-layout kernel: all of it lives on the device and is rewound by the kernel's last block,
  // so the launch takes no per-call host counter (replays under a CUDA graph)
  unsigned long long layout_ticket;                          // next logical block id of the running call
  unsigned layout_done;                                      // blocks of the running call that have finished
  unsigned layout_pad;
  unsigned long long layout_epoch;                           // calls completed so far (tags layout_pub)
  unsigned long long layout_pub[kMaxLayoutBlocks][kMaxRanks]; // (epoch << 32 | count) per block and rank
  int layout_hist[kMaxExperts];                              // per-expert totals of the running call, zero between calls
  unsigned long long trace[128];                             // B200EP_TRACE=1: %globaltimer stamps of the last kernels (b200ep_debug_trace)
  HtMeta ht[kNumSlots];
  unsigned long long epoch[kNumFlagCat][kNumSlots];          // call epochs, advanced on the device
};
// private scratch of the expert-grouped permute's plan kernel (permute.cu): block ticket, finished blocks, call tag and the
// per-(block, local expert) totals the blocks publish to each other, tagged by the call (never cleared)
constexpr size_t kPermPubWords = (size_t)1 << 19;
struct PermScratch {
  unsigned long long ticket;
  unsigned done, pad0;
  unsigned long long epoch;
  unsigned long long pad1[5];
  unsigned long long pub[kPermPubWords];
};
// pinned host mirror of the count exchange result
// Every word is (low 32 bits of the call epoch << 32 | value): one 8-byte store is atomic over PCIe, so
// the host can tell a fresh word from a stale one without the device issuing a system fence first.
constexpr int kHostWordsPerSlot = 1 + (kMaxRanks + 1) + kMaxLocalExperts;  // num_recv, rank_prefix[], per_expert[]
struct HostMirror {
  volatile unsigned long long word[kNumSlots][kHostWordsPerSlot];
  int32_t num_recv_cache[kNumSlots];  // host-side copy made by b200ep_ht_notify_wait (sizes later grids)
};
__host__ __device__ inline unsigned long long host_word(unsigned long long epoch, int32_t v) {
  return ((epoch & 0xffffffffull) << 32) | (unsigned long long)(uint32_t)v;
}
}  // namespace b200ep

struct b200ep_ctx {
  int rank, world, device, num_sms, sm_limit;
  char* sym;
  size_t sym_bytes, slot_bytes;
  b200ep::Peers peers;
  bool ipc_open[b200ep::kMaxRanks];
  bool connected;
  b200ep::LocalScratch* local;   // device
  b200ep::PermScratch* perm;     // device
  char* ll_local;                // device scratch for low-latency tables
  size_t ll_local_bytes;
  b200ep::HostMirror* host;      // pinned, mapped
  b200ep::HostMirror* host_dev;  // device alias of `host`
  uint64_t epoch[b200ep::kNumFlagCat][b200ep::kNumSlots];
  b200ep::HtPlan ht_plan[b200ep::kNumSlots];
  b200ep::LlPlan ll_plan[b200ep::kNumSlots];
  int ht_e_local[b200ep::kNumSlots];  // E/world of the last notify on the slot
  int ht_tokens[b200ep::kNumSlots];   // T of the last dispatch on the slot
};

namespace b200ep {
inline char* slot_base(const b200ep_ctx* c, const char* base, int slot) {
  return const_cast<char*>(base) + kCtrlBytes + (size_t)slot * c->slot_bytes;
}
int make_ht_plan(const b200ep_ctx* c, int H, int K, int x_kind, HtPlan* p);
int make_ll_plan(const b200ep_ctx* c, int H, int K, int E, int max_tok, int use_fp8, LlPlan* p);
}  // namespace b200ep

// High-throughput (prefill) dispatch and combine (north-star kernels 2 and 3, HT mode).
//
// Replaces deep_ep.Buffer.dispatch / .combine as vLLM calls them
// (vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ht.py:139-162,364-376)
// and folds in the per-token-group FP8 quantisation vLLM runs before the dispatch
// (deepep_ht.py:282-289 -> quantization/utils/fp8_utils.py:510-591).  CPU
// restatement: oracle/dispatch.py (ht_dispatch), oracle/combine.py (ht_combine),
// oracle/fp8.py (formula "div").
//
// Data path (no ring buffers, no receiver-side copy):
//   notify   every rank writes its per-destination and per-expert counts into
//            every peer's control block and raises a flag; after the wait each
//            rank knows the whole W x W count matrix, hence the first row it owns
//            in every destination's receive region (canonical order: source rank
//            ascending, source token ascending).  One rank: a single local kernel.
//   dispatch persistent CTAs of four 128-thread teams, one row per team at a time:
//            32-byte loads (LDG.E.256), all of a pass in flight -> fp32 amax per
//            128 group (8-lane shuffles) -> scale by a Markstein-corrected
//            reciprocal (== IEEE division) -> e4m3 into a double-buffered shared
//            memory stage -> ONE TMA bulk copy (cp.async.bulk) per destination rank
//            straight into the peer's receive region at the row's final position;
//            top-k ids are rewritten to destination-local ids on the way.  recv_x is
//            then read in place by the experts.
//   combine  expert ranks push result rows (bf16) with the same staging + bulk copy
//            into the source rank's stage [token][dst rank], destinations
//            interleaved; the source rank sums the <= W rows of a token in fp32 in
//            rank order and writes bf16 once.  With all phases in one call the push
//            and the reduction overlap in one kernel (per-chunk flags).
// Cross-GPU ordering: payload stores -> cp.async.bulk.wait_group -> CTA barrier ->
// atom.acq_rel.gpu on a CTA counter -> (last CTA) ONE fence.acq_rel.sys ->
// st.relaxed.sys flag in every peer; consumers spin with ld.acquire.sys.
#include <stdlib.h>

#include "common.cuh"

namespace b200ep {

// ------------------------------------------------------------------ notify
__global__ void __launch_bounds__(256)
    ht_notify_send_kernel(Peers peers, int W, int me, int slot, int e_local, const int32_t* __restrict__ tok_per_rank,
                          const int32_t* __restrict__ tok_per_expert, unsigned long long* epoch_ptr) {
  for (int d = 0; d < W; ++d) {
    char* ctrl = peers.base[d];
    int32_t* cr = reinterpret_cast<int32_t*>(ctrl + kCtrlHtCntRankOff) + ((size_t)slot * kMaxRanks + me) * kMaxRanks;
    int32_t* ce = reinterpret_cast<int32_t*>(ctrl + kCtrlHtCntExpOff) + ((size_t)slot * kMaxRanks + me) * kMaxLocalExperts;
    if ((int)threadIdx.x < W) cr[threadIdx.x] = tok_per_rank[threadIdx.x];
    for (int i = threadIdx.x; i < e_local; i += blockDim.x) ce[i] = tok_per_expert[d * e_local + i];
  }
  single_cta_signal(epoch_ptr, peers, W, kFlagHtNotify, slot, me);
}

// Derive this rank's receive offsets / totals from the count matrix and post them to the host.
// counts come from the control block (filled by every rank's notify_send) or, for a single rank,
// straight from the layout arrays (no exchange, no flags, no fence).
__device__ __forceinline__ void notify_finish(const int32_t* cr, int cr_stride, const int32_t* ce, int ce_stride, int W,
                                              int me, int slot, int e_local, HtMeta* meta, HostMirror* host,
                                              unsigned long long epoch) {
  const int tid = threadIdx.x;
  volatile unsigned long long* hw = host->word[slot];
  if (tid == 0) {
    int p = 0;
    for (int s = 0; s < W; ++s) {
      meta->rank_prefix[s] = p;
      hw[1 + s] = host_word(epoch, p);
      p += __ldcg(&cr[s * cr_stride + me]);
    }
    meta->rank_prefix[W] = p;
    hw[1 + W] = host_word(epoch, p);
    meta->num_recv = p;
    hw[0] = host_word(epoch, p);
  }
  if (tid >= 32 && tid < 32 + W) {
    const int d = tid - 32;
    int b = 0;
    for (int s = 0; s < me; ++s) b += __ldcg(&cr[s * cr_stride + d]);
    meta->send_base[d] = b;
    meta->send_count[d] = __ldcg(&cr[me * cr_stride + d]);
  }
  for (int i = tid; i < e_local; i += blockDim.x) {
    int n = 0;
    for (int s = 0; s < W; ++s) n += __ldcg(&ce[(size_t)s * ce_stride + i]);
    meta->recv_per_expert[i] = n;
    hw[1 + kMaxRanks + 1 + i] = host_word(epoch, n);
  }
}

__global__ void __launch_bounds__(256)
    ht_notify_recv_kernel(const char* my_base, int W, int me, int slot, int e_local, HtMeta* meta, HostMirror* host,
                          const unsigned long long* epoch_ptr) {
  cta_wait_all_sources(my_base, W, kFlagHtNotify, slot, me, epoch_ptr);
  const int32_t* cr = reinterpret_cast<const int32_t*>(my_base + kCtrlHtCntRankOff) + (size_t)slot * kMaxRanks * kMaxRanks;
  const int32_t* ce =
      reinterpret_cast<const int32_t*>(my_base + kCtrlHtCntExpOff) + (size_t)slot * kMaxRanks * kMaxLocalExperts;
  notify_finish(cr, kMaxRanks, ce, kMaxLocalExperts, W, me, slot, e_local, meta, host,
                *reinterpret_cast<const volatile unsigned long long*>(epoch_ptr));
}

// world == 1: the whole count "exchange" in one small kernel.
__global__ void __launch_bounds__(256)
    ht_notify_local_kernel(const int32_t* __restrict__ tok_per_rank, const int32_t* __restrict__ tok_per_expert, int slot,
                           int e_local, HtMeta* meta, HostMirror* host, unsigned long long* epoch_ptr) {
  const unsigned long long epoch = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
  notify_finish(tok_per_rank, 1, tok_per_expert, e_local, 1, 0, slot, e_local, meta, host, epoch);
  __syncthreads();
  if (threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) = epoch;
}

// ------------------------------------------------------------------ dispatch
struct HtDispatchArgs {
  Peers peers;
  HtPlan plan;
  const void* x;
  const float* x_scales;
  const void* ids;
  const float* weights;
  const uint8_t* in_rank;
  const int32_t* send_off;
  const HtMeta* meta;
  unsigned* done;
  size_t slot_off;  // kCtrlBytes + slot * slot_bytes
  unsigned long long* epoch_ptr;
  int T, W, me, slot, e_local, ue8m0, ids_i64, nstage;
};

constexpr int kTeamThreads = 128;  // a team of 4 warps moves one row at a time
constexpr int kTeams = 4;          // teams per CTA (512 threads); one sys-scope fence per CTA at the end
constexpr int kVpt = 8;            // 16-byte vectors in flight per thread (8 x 128 x 16 B = 16 KB per pass)

__device__ __forceinline__ void team_bar(int team) {
  asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(kTeamThreads) : "memory");
}

// Load one pass of a row into registers (all loads in flight), nothing else.
__device__ __forceinline__ void load_pass(const int4* src, int nvec, int v0, int ttid, int4 (&raw)[kVpt]) {
#pragma unroll
  for (int u = 0; u < kVpt; ++u) {
    const int v = v0 + u * kTeamThreads + ttid;
    raw[u] = v < nvec ? ld_nc_v4(src + v) : make_int4(0, 0, 0, 0);
  }
}

// Registers -> shared-memory stage; KIND B200EP_X_BF16_QUANT quantises on the way
// (16 consecutive lanes x 8 bf16 = one 128-element group).
template <int KIND>
__device__ __forceinline__ void store_pass(const int4 (&raw)[kVpt], int nvec, int v0, int ttid, char* s_row,
                                           float* s_scale, bool ue8m0) {
#pragma unroll
  for (int u = 0; u < kVpt; ++u) {
    const int v = v0 + u * kTeamThreads + ttid;
    if (v0 + u * kTeamThreads >= nvec) break;  // uniform over the team
    if (KIND == B200EP_X_BF16_QUANT) {
      float f[8];
      unpack8(raw[u], f);
      const float s = group_scale(group_amax16(absmax8(f)), ue8m0);
      if (v < nvec) {
        *reinterpret_cast<uint2*>(s_row + (size_t)v * 8) = quant8(f, s);
        if ((ttid & 15) == 0) s_scale[v >> 4] = s;
      }
    } else {
      if (v < nvec) reinterpret_cast<int4*>(s_row)[v] = raw[u];
    }
  }
}

// Quantising passes with 32-byte lanes (KIND == B200EP_X_BF16_QUANT): 4 x 128 x 32 B = 16 KB per pass.
constexpr int kVpt32 = 2;
__device__ __forceinline__ void load_pass32(const V8* src, int nv32, int v0, int ttid, V8 (&raw)[kVpt32]) {
#pragma unroll
  for (int u = 0; u < kVpt32; ++u) {
    const int v = v0 + u * kTeamThreads + ttid;
    if (v < nv32) raw[u] = ld_nc_v8(src + v);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) raw[u].w[i] = 0u;
    }
  }
}
__device__ __forceinline__ void quant_pass32(const V8 (&raw)[kVpt32], int nv32, int v0, int ttid, char* s_row,
                                             float* s_scale, bool ue8m0) {
#pragma unroll
  for (int u = 0; u < kVpt32; ++u) {
    const int v = v0 + u * kTeamThreads + ttid;
    if (v0 + u * kTeamThreads >= nv32) break;  // uniform over the team
    const float s = group_scale(group_amax8(absmax16(raw[u])), ue8m0);
    if (v < nv32) {
      reinterpret_cast<uint4*>(s_row)[v] = quant16(raw[u], s);
      if ((ttid & 7) == 0) s_scale[v >> 3] = s;
    }
  }
}

// Persistent CTAs of four 128-thread teams.  Per row a team: issues every load of the row (up to
// 16 KB in flight), waits until its previous TMA bulk copies have read the stage, quantises into
// the stage, and warp 0 fans the row out with ONE cp.async.bulk per destination rank while warp 1
// writes the per-destination metadata.  The next row's loads overlap the bulk copies of this one.
// kH > 0: hidden size known at compile time (the DeepSeek shape), which folds the per-vector bounds
// checks of the quantising passes; kH == 0: any hidden size.
template <int KIND, int kH = 0>
__global__ void __launch_bounds__(kTeams* kTeamThreads, KIND == B200EP_X_BF16_QUANT ? 3 : 2) ht_dispatch_send_kernel(const __grid_constant__ HtDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  const int tid = threadIdx.x, team = tid / kTeamThreads, ttid = tid % kTeamThreads;
  const int lane = tid & 31, twarp = ttid >> 5;
  const int W = a.W, K = a.plan.K, H = kH > 0 ? kH : a.plan.H;
  const int row_bytes = (kH > 0 && KIND == B200EP_X_BF16_QUANT) ? kH : a.plan.row_bytes;
  const int scale_bytes = (kH > 0 && KIND == B200EP_X_BF16_QUANT) ? kH / kGroup * 4 : a.plan.scale_bytes;
  const int stage_bytes = (row_bytes + scale_bytes + 127) / 128 * 128;
  char* s_team = s_mem + (size_t)team * a.nstage * stage_bytes;
  const int in_row_bytes = KIND == B200EP_X_BF16_QUANT ? 2 * H : row_bytes;
  const int nvec = in_row_bytes / 16;
  int n = 0;  // rows this team has staged so far: stage = n % nstage, one bulk group per row and lane

  for (int t = blockIdx.x * kTeams + team; t < a.T; t += gridDim.x * kTeams) {
    // destination set of this token: lane d holds rank d's membership and row (every warp computes it)
    int my_in = 0, my_row = -1;
    if (lane < W) {
      my_in = a.in_rank[(size_t)t * W + lane];
      if (my_in) my_row = a.meta->send_base[lane] + a.send_off[(size_t)t * W + lane];
    }
    const unsigned dmask = __ballot_sync(0xffffffffu, my_in);
    if (dmask == 0) continue;  // uniform over the team

    char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * stage_bytes;
    float* s_scale = reinterpret_cast<float*>(s_row + row_bytes);
    ++n;
    if (KIND == B200EP_X_BF16_QUANT) {
      const V8* src = reinterpret_cast<const V8*>(reinterpret_cast<const char*>(a.x) + (size_t)t * in_row_bytes);
      const int nv32 = nvec / 2;
      V8 raw[kVpt32];
      load_pass32(src, nv32, 0, ttid, raw);
      if (twarp == 0) {  // the bulk copies that last used this stage have finished reading it
        if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
      }
      team_bar(team);
      quant_pass32(raw, nv32, 0, ttid, s_row, s_scale, a.ue8m0 != 0);
      for (int v0 = kTeamThreads * kVpt32; v0 < nv32; v0 += kTeamThreads * kVpt32) {  // rows wider than 16 KB
        load_pass32(src, nv32, v0, ttid, raw);
        quant_pass32(raw, nv32, v0, ttid, s_row, s_scale, a.ue8m0 != 0);
      }
    } else {
      const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const char*>(a.x) + (size_t)t * in_row_bytes);
      int4 raw[kVpt];
      load_pass(src, nvec, 0, ttid, raw);
      if (twarp == 0) {
        if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
      }
      team_bar(team);
      store_pass<KIND>(raw, nvec, 0, ttid, s_row, s_scale, false);
      for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
        load_pass(src, nvec, v0, ttid, raw);
        store_pass<KIND>(raw, nvec, v0, ttid, s_row, s_scale, false);
      }
    }
    if (KIND == B200EP_X_FP8) {
      const int ng = H / kGroup;
      for (int g = ttid; g < ng; g += kTeamThreads) s_scale[g] = a.x_scales[(size_t)t * ng + g];
    }
    fence_proxy_async_smem();
    team_bar(team);
    if (twarp == 0) {
      if (my_in) {
        char* dst = a.peers.base[lane] + a.slot_off;
        bulk_s2g(dst + a.plan.off_x + (size_t)my_row * row_bytes, s_row, row_bytes);
        if (scale_bytes && (scale_bytes & 15) == 0)
          bulk_s2g(dst + a.plan.off_scales + (size_t)my_row * scale_bytes, s_scale, scale_bytes);
        *reinterpret_cast<int32_t*>(dst + a.plan.off_src + (size_t)my_row * 4) = t;
      }
      bulk_commit();  // every lane, every row (possibly empty): group count == rows staged
      if (scale_bytes && (scale_bytes & 15)) {  // scale vector not a 16-byte multiple: plain stores
        const int ng = scale_bytes / 4;
        for (int d = 0; d < W; ++d) {
          if (!((dmask >> d) & 1u)) continue;
          const int row = __shfl_sync(0xffffffffu, my_row, d);
          float* dsts = reinterpret_cast<float*>(a.peers.base[d] + a.slot_off + a.plan.off_scales + (size_t)row * scale_bytes);
          for (int g = lane; g < ng; g += 32) dsts[g] = s_scale[g];
        }
        __syncwarp();
      }
    } else if (twarp == 1) {
      // top-k ids -> destination-local ids; weights zeroed where the expert lives elsewhere
      long long id = -1;
      float w = 0.f;
      if (lane < K) {
        id = a.ids_i64 ? reinterpret_cast<const long long*>(a.ids)[(size_t)t * K + lane]
                       : (long long)reinterpret_cast<const int*>(a.ids)[(size_t)t * K + lane];
        w = a.weights ? a.weights[(size_t)t * K + lane] : 0.f;
      }
      const int owner = id >= 0 ? (int)(id / a.e_local) : -1;
      for (int d = 0; d < W; ++d) {
        if (!((dmask >> d) & 1u)) continue;
        const int row = __shfl_sync(0xffffffffu, my_row, d);
        if (lane < K) {
          char* dst = a.peers.base[d] + a.slot_off;
          const bool mine = owner == d;
          reinterpret_cast<long long*>(dst + a.plan.off_idx)[(size_t)row * K + lane] = mine ? id - (long long)d * a.e_local : -1;
          reinterpret_cast<float*>(dst + a.plan.off_w)[(size_t)row * K + lane] = mine ? w : 0.f;
        }
      }
    }
  }
  if ((ttid >> 5) == 0) bulk_wait_all();
  if (W > 1) cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtData, a.slot, a.me);  // a single rank has nobody to tell
}

// ---- warp-granular quantising dispatch (KIND == B200EP_X_BF16_QUANT, rows up to 8192 elements).
// ncu on the team kernel at EP=1: 21 % of the stall samples were team barriers and 24 % first uses of
// loaded data, issue slots 56 % busy.  Here every warp owns whole rows and a private shared-memory
// stage, so nothing but __syncwarp separates load -> quantise -> TMA fan-out, and 24 independent
// warps per SM (one 768-thread CTA, <= 85 registers) keep 4 x 32 B x 32 lanes each in flight.
constexpr int kWarpCtaWarps = 24;
__global__ void __launch_bounds__(kWarpCtaWarps * 32, 1) ht_dispatch_send_warp_kernel(const __grid_constant__ HtDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int W = a.W, K = a.plan.K, H = a.plan.H;
  const int row_bytes = a.plan.row_bytes, scale_bytes = a.plan.scale_bytes;
  const int stage_bytes = (row_bytes + scale_bytes + 127) / 128 * 128;
  char* s_row = s_mem + (size_t)warp * stage_bytes;
  float* s_scale = reinterpret_cast<float*>(s_row + row_bytes);
  const int nv32 = H / 16;
  const bool ue8m0 = a.ue8m0 != 0;

  for (int t = blockIdx.x * kWarpCtaWarps + warp; t < a.T; t += gridDim.x * kWarpCtaWarps) {
    int my_in = 0, my_row = -1;
    if (lane < W) {
      my_in = a.in_rank[(size_t)t * W + lane];
      if (my_in) my_row = a.meta->send_base[lane] + a.send_off[(size_t)t * W + lane];
    }
    const unsigned dmask = __ballot_sync(0xffffffffu, my_in);
    if (dmask == 0) continue;
    const V8* src = reinterpret_cast<const V8*>(reinterpret_cast<const __nv_bfloat16*>(a.x) + (size_t)t * H);
    bool stage_free = false;
    for (int v0 = 0; v0 < nv32; v0 += 32 * kVpt32) {
      V8 raw[kVpt32];
#pragma unroll
      for (int u = 0; u < kVpt32; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nv32) raw[u] = ld_nc_v8(src + v);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) raw[u].w[i] = 0u;
        }
      }
      if (!stage_free) {  // loads are in flight; now make sure the previous row's copies have read the stage
        bulk_wait_read0();
        __syncwarp();
        stage_free = true;
      }
#pragma unroll
      for (int u = 0; u < kVpt32; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v0 + u * 32 >= nv32) break;  // uniform over the warp
        const float s = group_scale(group_amax8(absmax16(raw[u])), ue8m0);
        if (v < nv32) {
          reinterpret_cast<uint4*>(s_row)[v] = quant16(raw[u], s);
          if ((lane & 7) == 0) s_scale[v >> 3] = s;
        }
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (my_in) {
      char* dst = a.peers.base[lane] + a.slot_off;
      bulk_s2g(dst + a.plan.off_x + (size_t)my_row * row_bytes, s_row, row_bytes);
      if ((scale_bytes & 15) == 0) bulk_s2g(dst + a.plan.off_scales + (size_t)my_row * scale_bytes, s_scale, scale_bytes);
      *reinterpret_cast<int32_t*>(dst + a.plan.off_src + (size_t)my_row * 4) = t;
    }
    bulk_commit();
    // metadata: top-k ids -> destination-local ids; weights zeroed where the expert lives elsewhere
    long long id = -1;
    float w = 0.f;
    if (lane < K) {
      id = a.ids_i64 ? reinterpret_cast<const long long*>(a.ids)[(size_t)t * K + lane]
                     : (long long)reinterpret_cast<const int*>(a.ids)[(size_t)t * K + lane];
      w = a.weights ? a.weights[(size_t)t * K + lane] : 0.f;
    }
    const int owner = id >= 0 ? (int)(id / a.e_local) : -1;
    for (int d = 0; d < W; ++d) {
      if (!((dmask >> d) & 1u)) continue;
      const int row = __shfl_sync(0xffffffffu, my_row, d);
      char* dst = a.peers.base[d] + a.slot_off;
      if (lane < K) {
        const bool mine = owner == d;
        reinterpret_cast<long long*>(dst + a.plan.off_idx)[(size_t)row * K + lane] = mine ? id - (long long)d * a.e_local : -1;
        reinterpret_cast<float*>(dst + a.plan.off_w)[(size_t)row * K + lane] = mine ? w : 0.f;
      }
      if (scale_bytes & 15) {  // scale vector not a 16-byte multiple: plain stores
        float* dsts = reinterpret_cast<float*>(dst + a.plan.off_scales + (size_t)row * scale_bytes);
        for (int g = lane; g < scale_bytes / 4; g += 32) dsts[g] = s_scale[g];
      }
    }
  }
  bulk_wait_all();
  if (W > 1) cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtData, a.slot, a.me);
}

// ---- staged dispatch (W >= 3): quantise once, then push destination by destination.
// With the direct kernel every rank fans each token out to a random subset of ranks; at W = 8
// the fabric then behaves like an input-queued switch with head-of-line blocking (measured 54 %
// of the peer-copy peak against 86 % for the permutation-ordered combine push).  The staged form
// keeps the single quantisation pass but orders the NVLink traffic: rank r sends to r+1, r+2, ...
// in interleaved rotation, so at any moment the ranks push along a permutation.

// Phase 1: quantise (or copy) every token that leaves this rank into the local q-stage
// [t][row | scales] and record send_list[d][p] = t.  Purely local: no flags, no system fence.
template <int KIND>
__global__ void __launch_bounds__(kTeams* kTeamThreads, 2) ht_dispatch_stage_kernel(const __grid_constant__ HtDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  const int tid = threadIdx.x, team = tid / kTeamThreads, ttid = tid % kTeamThreads;
  const int lane = tid & 31, twarp = ttid >> 5;
  const int W = a.W, H = a.plan.H;
  const int row_bytes = a.plan.row_bytes, scale_bytes = a.plan.scale_bytes, qrow = a.plan.qrow_bytes;
  const int stage_bytes = (qrow + 127) / 128 * 128;
  char* s_team = s_mem + (size_t)team * a.nstage * stage_bytes;
  const int in_row_bytes = KIND == B200EP_X_BF16_QUANT ? 2 * H : row_bytes;
  const int nvec = in_row_bytes / 16;
  char* my_slot = a.peers.base[a.me] + a.slot_off;
  int32_t* send_list = reinterpret_cast<int32_t*>(my_slot + a.plan.off_sendlist);
  int n = 0;

  for (int t = blockIdx.x * kTeams + team; t < a.T; t += gridDim.x * kTeams) {
    int my_in = 0;
    if (lane < W) my_in = a.in_rank[(size_t)t * W + lane];
    const unsigned dmask = __ballot_sync(0xffffffffu, my_in);
    if (dmask == 0) continue;  // uniform over the team
    if (twarp == 1 && my_in) send_list[(size_t)lane * a.plan.t_cap + a.send_off[(size_t)t * W + lane]] = t;

    const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const char*>(a.x) + (size_t)t * in_row_bytes);
    int4 raw[kVpt];
    load_pass(src, nvec, 0, ttid, raw);
    char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * stage_bytes;
    float* s_scale = reinterpret_cast<float*>(s_row + row_bytes);
    ++n;
    if (ttid == 0) {
      if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
    }
    team_bar(team);
    store_pass<KIND>(raw, nvec, 0, ttid, s_row, s_scale, a.ue8m0 != 0);
    for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
      load_pass(src, nvec, v0, ttid, raw);
      store_pass<KIND>(raw, nvec, v0, ttid, s_row, s_scale, a.ue8m0 != 0);
    }
    if (KIND == B200EP_X_FP8) {
      const int ng = H / kGroup;
      for (int g = ttid; g < ng; g += kTeamThreads) s_scale[g] = a.x_scales[(size_t)t * ng + g];
    }
    if (scale_bytes && ttid < (qrow - row_bytes - scale_bytes) / 4) s_scale[scale_bytes / 4 + ttid] = 0.f;  // pad
    fence_proxy_async_smem();
    team_bar(team);
    if (ttid == 0) {
      bulk_s2g(my_slot + a.plan.off_qstage + (size_t)t * qrow, s_row, qrow);
      bulk_commit();
    }
  }
  if (ttid == 0) bulk_wait_all();
}

// Phase 2: push.  Items interleave the W destinations starting at rank me+1 (self last).
__global__ void __launch_bounds__(kTeams* kTeamThreads, 2) ht_dispatch_push_kernel(const __grid_constant__ HtDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  const int tid = threadIdx.x, team = tid / kTeamThreads, ttid = tid % kTeamThreads;
  const int lane = tid & 31, twarp = ttid >> 5;
  const int W = a.W, K = a.plan.K;
  const int row_bytes = a.plan.row_bytes, scale_bytes = a.plan.scale_bytes, qrow = a.plan.qrow_bytes;
  const int stage_bytes = (qrow + 127) / 128 * 128;
  char* s_team = s_mem + (size_t)team * a.nstage * stage_bytes;
  const int nvec = qrow / 16;
  const char* my_slot = a.peers.base[a.me] + a.slot_off;
  const int32_t* send_list = reinterpret_cast<const int32_t*>(my_slot + a.plan.off_sendlist);
  const char* qstage = my_slot + a.plan.off_qstage;
  int n = 0;
  int maxlen = 0;
  for (int d = 0; d < W; ++d) maxlen = max(maxlen, a.meta->send_count[d]);
  const long long items = (long long)maxlen * W;

  for (long long i = (long long)blockIdx.x * kTeams + team; i < items; i += (long long)gridDim.x * kTeams) {
    const int d = (a.me + 1 + (int)(i % W)) % W;
    const int v = (int)(i / W);
    if (v >= a.meta->send_count[d]) continue;  // uniform over the team
    const int t = send_list[(size_t)d * a.plan.t_cap + v];
    const int row = a.meta->send_base[d] + v;
    const int4* src = reinterpret_cast<const int4*>(qstage + (size_t)t * qrow);
    int4 raw[kVpt];
#pragma unroll
    for (int u = 0; u < kVpt; ++u) {
      const int vv = u * kTeamThreads + ttid;
      if (vv < nvec) raw[u] = ld_cg_v4(src + vv);  // written by the previous kernel: read through L2
    }
    char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * stage_bytes;
    ++n;
    if (ttid == 0) {
      if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
    }
    team_bar(team);
#pragma unroll
    for (int u = 0; u < kVpt; ++u) {
      const int vv = u * kTeamThreads + ttid;
      if (vv < nvec) reinterpret_cast<int4*>(s_row)[vv] = raw[u];
    }
    for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
      for (int vv = v0 + ttid; vv < min(nvec, v0 + kTeamThreads * kVpt); vv += kTeamThreads)
        reinterpret_cast<int4*>(s_row)[vv] = ld_cg_v4(src + vv);
    }
    fence_proxy_async_smem();
    team_bar(team);
    char* dst = a.peers.base[d] + a.slot_off;
    if (ttid == 0) {
      bulk_s2g(dst + a.plan.off_x + (size_t)row * row_bytes, s_row, row_bytes);
      if (scale_bytes && (scale_bytes & 15) == 0)
        bulk_s2g(dst + a.plan.off_scales + (size_t)row * scale_bytes, s_row + row_bytes, scale_bytes);
      bulk_commit();
      *reinterpret_cast<int32_t*>(dst + a.plan.off_src + (size_t)row * 4) = t;
    } else if (twarp == 1) {
      if (lane < K) {
        const long long id = a.ids_i64 ? reinterpret_cast<const long long*>(a.ids)[(size_t)t * K + lane]
                                       : (long long)reinterpret_cast<const int*>(a.ids)[(size_t)t * K + lane];
        const float w = a.weights ? a.weights[(size_t)t * K + lane] : 0.f;
        const bool mine = id >= 0 && (int)(id / a.e_local) == d;
        reinterpret_cast<long long*>(dst + a.plan.off_idx)[(size_t)row * K + lane] = mine ? id - (long long)d * a.e_local : -1;
        reinterpret_cast<float*>(dst + a.plan.off_w)[(size_t)row * K + lane] = mine ? w : 0.f;
      }
    } else if (twarp == 2 && scale_bytes && (scale_bytes & 15)) {
      float* dsts = reinterpret_cast<float*>(dst + a.plan.off_scales + (size_t)row * scale_bytes);
      const float* ss = reinterpret_cast<const float*>(s_row + row_bytes);
      for (int g = lane; g < scale_bytes / 4; g += 32) dsts[g] = ss[g];
    }
  }
  if (ttid == 0) bulk_wait_all();
  cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtData, a.slot, a.me);
}

__global__ void wait_sources_kernel(const char* my_base, int W, int cat, int slot, int me,
                                    const unsigned long long* epoch_ptr) {
  cta_wait_all_sources(my_base, W, cat, slot, me, epoch_ptr);
}

// ------------------------------------------------------------------ combine
struct HtCombineArgs {
  Peers peers;
  HtPlan plan;
  const void* y;    // bf16 [M, H]
  void* out;        // bf16 [T, H]
  const uint8_t* in_rank;
  const int32_t* send_off;
  const HtMeta* meta;
  unsigned* done;
  size_t slot_off;
  unsigned long long* epoch_ptr;
  unsigned* chunk_done;  // [kCombChunks] private counters of the overlapped kernel
  unsigned* red_next;    // [kCombChunks] next reduce item of each chunk, [kCombChunks] = warps finished, [+1] = role tickets
  unsigned long long epoch_value;  // overlapped kernel only: the epoch this call publishes (host mirror; HT is never graph-captured)
  int T, W, me, slot, nstage;
};

// Expert rank: result row m goes back to (source rank, source token) = stage[t][me] on that rank.
// Rows whose source is this rank stay where they are; the reduce reads dev_y directly.
// Same team structure as the dispatch: loads of the next row overlap the bulk copy of this one.
__global__ void __launch_bounds__(kTeams* kTeamThreads, 2) ht_combine_send_kernel(const __grid_constant__ HtCombineArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  const int tid = threadIdx.x, team = tid / kTeamThreads, ttid = tid % kTeamThreads;
  const int W = a.W;
  const int row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  char* s_team = s_mem + (size_t)team * a.nstage * row_bytes;
  int n = 0;
  const int32_t* src_idx = reinterpret_cast<const int32_t*>(a.peers.base[a.me] + a.slot_off + a.plan.off_src);
  // Work order: consecutive items go to DIFFERENT source ranks, starting at rank me+1, so that at
  // any moment the W ranks push along a permutation instead of all converging on rank 0 (incast).
  int maxlen = 0;
  for (int k = 1; k < W; ++k) {
    const int s = (a.me + k) % W;
    maxlen = max(maxlen, a.meta->rank_prefix[s + 1] - a.meta->rank_prefix[s]);
  }
  const long long items = (long long)maxlen * (W - 1);

  for (long long i = (long long)blockIdx.x * kTeams + team; i < items; i += (long long)gridDim.x * kTeams) {
    const int s = (a.me + 1 + (int)(i % (W - 1))) % W;
    const int v = (int)(i / (W - 1));
    const int lo = a.meta->rank_prefix[s];
    if (v >= a.meta->rank_prefix[s + 1] - lo) continue;  // uniform over the team
    const int m = lo + v;
    const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const char*>(a.y) + (size_t)m * row_bytes);
    int4 raw[kVpt];
    load_pass(src, nvec, 0, ttid, raw);
    char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * row_bytes;
    ++n;
    if (ttid == 0) {
      if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
    }
    team_bar(team);
    store_pass<B200EP_X_BF16>(raw, nvec, 0, ttid, s_row, nullptr, false);
    for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
      load_pass(src, nvec, v0, ttid, raw);
      store_pass<B200EP_X_BF16>(raw, nvec, v0, ttid, s_row, nullptr, false);
    }
    fence_proxy_async_smem();
    team_bar(team);
    if (ttid == 0) {
      const int t = __ldcg(src_idx + m);
      bulk_s2g(a.peers.base[s] + a.slot_off + a.plan.off_stage + ((size_t)t * W + a.me) * row_bytes, s_row, row_bytes);
      bulk_commit();
    }
  }
  if (ttid == 0) bulk_wait_all();
  cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtComb, a.slot, a.me);
}

__device__ __forceinline__ void acc8(float (&acc)[8], const int4& v) {
  float f[8];
  unpack8(v, f);
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
}
__device__ __forceinline__ int4 pack8_bf16(const float (&acc)[8]) {
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 p = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
    o[i] = *reinterpret_cast<uint32_t*>(&p);
  }
  return make_int4((int)o[0], (int)o[1], (int)o[2], (int)o[3]);
}

// Source rank: out[t] = bf16( sum_{d in ranks(t), ascending} fp32(row_d) ).
// One warp per (token, half row); the row pointers are resolved once per item, then the warp walks
// the half row in batches of 4 vectors per lane with the loads of up to four source rows in flight
// together (16 x 16 B per lane).  Accumulation order stays d = 0..W-1.
constexpr int kRedVec = 4;
// kMaxSrc = ranks the kernel is built for: the 1- and 2-rank builds hold at most 2 x 4 vectors of source
// data, fit 64 registers and run 4 CTAs per SM (ncu on the generic build at EP=1: 128 registers,
// 16 warps per SM, issue slots 35 % busy).
template <int kMaxSrc>
__global__ void __launch_bounds__(256, kMaxSrc == 1 ? 4 : (kMaxSrc == 2 ? 3 : 2)) ht_combine_recv_kernel(const __grid_constant__ HtCombineArgs a) {
  constexpr int kGrp = kMaxSrc < 4 ? kMaxSrc : 4;
  if (a.W > 1) cta_wait_all_sources(a.peers.base[a.me], a.W, kFlagHtComb, a.slot, a.me, a.epoch_ptr);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int W = a.W;
  const int row_bytes = 2 * a.plan.H;
  const int nvec = row_bytes / 16;
  const int nsplit = nvec >= 256 ? 2 : 1;
  const int per = (nvec / nsplit + 31) / 32 * 32;  // vectors per item, multiple of 32
  const char* stage = a.peers.base[a.me] + a.slot_off + a.plan.off_stage;
  const int my_first = a.meta->rank_prefix[a.me];
  const long long items = (long long)a.T * nsplit;

  for (long long it = (long long)blockIdx.x * nwarp + warp; it < items; it += (long long)gridDim.x * nwarp) {
    const int t = (int)(it / nsplit), vbeg = (int)(it % nsplit) * per;
    const int vend = min(nvec, vbeg + per);
    const char* rowp = nullptr;
    if (lane < W && a.in_rank[(size_t)t * W + lane]) {
      rowp = lane == a.me ? reinterpret_cast<const char*>(a.y) + (size_t)(my_first + a.send_off[(size_t)t * W + lane]) * row_bytes
                          : stage + ((size_t)t * W + lane) * row_bytes;
    }
    const unsigned dmask = __ballot_sync(0xffffffffu, rowp != nullptr);
    const char* rows[kMaxSrc];
#pragma unroll
    for (int d = 0; d < kMaxSrc; ++d)
      rows[d] = reinterpret_cast<const char*>(__shfl_sync(0xffffffffu, (unsigned long long)rowp, d));
    int4* dst = reinterpret_cast<int4*>(reinterpret_cast<char*>(a.out) + (size_t)t * row_bytes);
    for (int v0 = vbeg; v0 < vend; v0 += 32 * kRedVec) {
      float acc[kRedVec][8];
#pragma unroll
      for (int j = 0; j < kRedVec; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
#pragma unroll
      for (int d0 = 0; d0 < kMaxSrc; d0 += kGrp) {
        if ((dmask >> d0) == 0) break;
        int4 raw[kGrp][kRedVec];
#pragma unroll
        for (int dd = 0; dd < kGrp; ++dd) {
          if ((dmask >> (d0 + dd)) & 1u) {
#pragma unroll
            for (int j = 0; j < kRedVec; ++j) {
              const int v = v0 + j * 32 + lane;
              if (v < vend) raw[dd][j] = ld_cg_v4(rows[d0 + dd] + (size_t)v * 16);
            }
          }
        }
#pragma unroll
        for (int dd = 0; dd < kGrp; ++dd) {
          if ((dmask >> (d0 + dd)) & 1u) {
#pragma unroll
            for (int j = 0; j < kRedVec; ++j) {
              const int v = v0 + j * 32 + lane;
              if (v < vend) acc8(acc[j], raw[dd][j]);
            }
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kRedVec; ++j) {
        const int v = v0 + j * 32 + lane;
        if (v < vend) st_na_v4(dst + v, pack8_bf16(acc[j]));
      }
    }
  }
}

// ---- overlapped combine (production path, phases == ALL, W > 1).
// One launch, one CTA per SM, CTAs alternate between two roles:
//   even CTAs  push result rows back (as ht_combine_send_kernel) but chunk by chunk of 512 source
//              tokens; after draining a chunk every push CTA arrives on a per-chunk counter and the
//              last one raises that chunk's flags on all peers (one system fence per chunk, nobody waits);
//   odd CTAs   reduce (as ht_combine_recv_kernel) chunk c as soon as every peer's chunk-c flag is up.
// The fp32 reduction of the early chunks therefore runs under the NVLink transfer of the later
// ones instead of after it.  Push CTAs never wait, so the kernel cannot deadlock; the separate
// SEND / RECV kernels remain for phase-by-phase callers (single-GPU emulation, recv hooks).
__global__ void __launch_bounds__(kTeams* kTeamThreads, 1) ht_combine_fused_kernel(const __grid_constant__ HtCombineArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_lo[kMaxRanks][kCombChunks + 1];
  const int tid = threadIdx.x;
  const int W = a.W;
  const int row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  // CTAs alternate push / reduce-only; push CTAs join the reduction once their rows are out and
  // reduce items are handed out dynamically.  (Measured at EP=8: 1:1 static 483 us, 3:1 502 us --
  // with fewer reducers the reduction falls behind the transfer and becomes a tail.)
  // CTAs alternate push / reduce-only by block parity.  Placement matters: putting all push CTAs on
  // the low block indices (SMs 0..73, one die) cost 793 -> 1014 us per step at EP=8, handing roles out
  // by start-order tickets 1168 us.  To stay safe when the GPU is shared with other kernels (DBO) and
  // not every CTA can be resident, a reduce-only CTA first checks that all push CTAs of this rank are
  // running; if they are not within ~10 us it EXITS (reduce items come from a queue, so the push CTAs
  // finish the reduction themselves) instead of holding a slot a push CTA may be waiting for.
  const int nred = gridDim.x >> 1, nsend = gridDim.x - nred;
  const int role = blockIdx.x & 1;
  const int ridx = blockIdx.x >> 1;
  unsigned* started = &a.red_next[kCombChunks + 1];
  auto finish = [&]() {  // every warp of the grid, exactly once: the last one rewinds the counters
    if ((tid & 31) == 0) {
      const unsigned prev = atomicAdd(&a.red_next[kCombChunks], 1u);
      if (prev == gridDim.x * (blockDim.x >> 5) - 1)
        for (int c = 0; c <= kCombChunks + 1; ++c) a.red_next[c] = 0;
    }
  };
  if (role == 0) {
    if (tid == 0) atomicAdd(started, 1u);
  } else {
    __shared__ int s_go;
    if (tid == 0) {
      const long long t0 = clock64();
      while (*reinterpret_cast<volatile unsigned*>(started) < (unsigned)nsend && clock64() - t0 < 20000) {
      }
      s_go = *reinterpret_cast<volatile unsigned*>(started) >= (unsigned)nsend;
    }
    __syncthreads();
    if (!s_go) {
      finish();
      return;
    }
  }
  const int32_t* src_idx = reinterpret_cast<const int32_t*>(a.peers.base[a.me] + a.slot_off + a.plan.off_src);

  if (role == 0) {
    // ------------------------------------------------ push role
    const int team = tid / kTeamThreads, ttid = tid % kTeamThreads;
    char* s_team = s_mem + (size_t)team * a.nstage * row_bytes;
    int n = 0;
    // chunk bounds of every remote source's row range: rows are ascending in source token
    if (tid < (W - 1) * (kCombChunks + 1)) {
      const int k = tid / (kCombChunks + 1), c = tid % (kCombChunks + 1);
      const int s = (a.me + 1 + k) % W;
      int lo = a.meta->rank_prefix[s], hi = a.meta->rank_prefix[s + 1];
      if (c == kCombChunks) {
        lo = hi;
      } else {
        const int want = c * kCombChunkTokens;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (__ldcg(src_idx + mid) < want) lo = mid + 1; else hi = mid;
        }
      }
      s_lo[k][c] = lo;
    }
    __syncthreads();
    // Completion of chunk c is published one chunk late (after chunk c+1 has been issued), so the
    // wait for the last rows of a chunk to land overlaps the next chunk's transfers.
    auto arrive = [&](int c) {
      __syncthreads();
      if (tid == 0) {
        const unsigned prev = atom_add_acq_rel_gpu(&a.chunk_done[c], 1u);
        if (prev == (unsigned)nsend - 1) {
          a.chunk_done[c] = 0;
          fence_acq_rel_sys();
          for (int d = 0; d < W; ++d)
            if (d != a.me)
              st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(a.peers.base[d] + flag_off(kFlagHtCombChunk0 + c, a.slot, a.me)), a.epoch_value);
          if (c == kCombChunks - 1) *reinterpret_cast<volatile unsigned long long*>(a.epoch_ptr) = a.epoch_value;
        }
      }
    };
    for (int c = 0; c < kCombChunks; ++c) {
      int issued = 0;  // bulk groups this team's issuing thread committed in chunk c
      int maxlen = 0;
      for (int k = 0; k < W - 1; ++k) maxlen = max(maxlen, s_lo[k][c + 1] - s_lo[k][c]);
      const long long items = (long long)maxlen * (W - 1);
      for (long long i = (long long)ridx * kTeams + team; i < items; i += (long long)nsend * kTeams) {
        const int k = (int)(i % (W - 1));
        const int m = s_lo[k][c] + (int)(i / (W - 1));
        if (m >= s_lo[k][c + 1]) continue;  // uniform over the team
        const int s = (a.me + 1 + k) % W;
        const int4* src = reinterpret_cast<const int4*>(reinterpret_cast<const char*>(a.y) + (size_t)m * row_bytes);
        int4 raw[kVpt];
        load_pass(src, nvec, 0, ttid, raw);
        char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * row_bytes;
        ++n;
        if (ttid == 0) {
          if (a.nstage == 2) bulk_wait_read1(); else bulk_wait_read0();
        }
        team_bar(team);
        store_pass<B200EP_X_BF16>(raw, nvec, 0, ttid, s_row, nullptr, false);
        for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
          load_pass(src, nvec, v0, ttid, raw);
          store_pass<B200EP_X_BF16>(raw, nvec, v0, ttid, s_row, nullptr, false);
        }
        fence_proxy_async_smem();
        team_bar(team);
        if (ttid == 0) {
          const int t = __ldcg(src_idx + m);
          bulk_s2g(a.peers.base[s] + a.slot_off + a.plan.off_stage + ((size_t)t * W + a.me) * row_bytes, s_row, row_bytes);
          bulk_commit();
          ++issued;
        }
      }
      if (c > 0) {
        // everything older than this chunk's (>= 4) most recent groups has landed -> chunk c-1 is complete here
        if (ttid == 0) {
          if (issued >= 4) asm volatile("cp.async.bulk.wait_group 4;" ::: "memory");
          else bulk_wait_all();
        }
        arrive(c - 1);
      }
    }
    if (ttid == 0) bulk_wait_all();
    arrive(kCombChunks - 1);
  }
  {
    // ------------------------------------------------ reduce (every CTA; push CTAs arrive late)
    (void)nred;
    const int lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    const int nsplit = nvec >= 256 ? 2 : 1;
    const int per = (nvec / nsplit + 31) / 32 * 32;
    const char* stage = a.peers.base[a.me] + a.slot_off + a.plan.off_stage;
    const int my_first = a.meta->rank_prefix[a.me];
    for (int c = 0; c < kCombChunks; ++c) {
      const int t0 = c * kCombChunkTokens;
      const int t1 = c == kCombChunks - 1 ? a.T : min(a.T, t0 + kCombChunkTokens);
      if (tid < W && tid != a.me)
        wait_flag_sys(reinterpret_cast<const uint64_t*>(a.peers.base[a.me] + flag_off(kFlagHtCombChunk0 + c, a.slot, tid)),
                      a.epoch_value, kFlagHtCombChunk0 + c, a.me, tid);
      __syncthreads();
      if (t1 <= t0) continue;
      const int items = (t1 - t0) * nsplit;
      for (;;) {
        int it = 0;
        if (lane == 0) it = (int)atomicAdd(&a.red_next[c], 1u);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= items) break;
        const int t = t0 + it / nsplit, vbeg = (it % nsplit) * per;
        const int vend = min(nvec, vbeg + per);
        const char* rowp = nullptr;
        if (lane < W && a.in_rank[(size_t)t * W + lane]) {
          rowp = lane == a.me ? reinterpret_cast<const char*>(a.y) + (size_t)(my_first + a.send_off[(size_t)t * W + lane]) * row_bytes
                              : stage + ((size_t)t * W + lane) * row_bytes;
        }
        const unsigned dmask = __ballot_sync(0xffffffffu, rowp != nullptr);
        const char* rows[kMaxRanks];
#pragma unroll
        for (int d = 0; d < kMaxRanks; ++d)
          rows[d] = reinterpret_cast<const char*>(__shfl_sync(0xffffffffu, (unsigned long long)rowp, d));
        int4* dst = reinterpret_cast<int4*>(reinterpret_cast<char*>(a.out) + (size_t)t * row_bytes);
        for (int v0 = vbeg; v0 < vend; v0 += 32 * kRedVec) {
          float acc[kRedVec][8];
#pragma unroll
          for (int j = 0; j < kRedVec; ++j)
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
#pragma unroll
          for (int d0 = 0; d0 < kMaxRanks; d0 += 4) {
            if ((dmask >> d0) == 0) break;
            int4 raw[4][kRedVec];
#pragma unroll
            for (int dd = 0; dd < 4; ++dd) {
              if ((dmask >> (d0 + dd)) & 1u) {
#pragma unroll
                for (int j = 0; j < kRedVec; ++j) {
                  const int v = v0 + j * 32 + lane;
                  if (v < vend) raw[dd][j] = ld_cg_v4(rows[d0 + dd] + (size_t)v * 16);
                }
              }
            }
#pragma unroll
            for (int dd = 0; dd < 4; ++dd) {
              if ((dmask >> (d0 + dd)) & 1u) {
#pragma unroll
                for (int j = 0; j < kRedVec; ++j) {
                  const int v = v0 + j * 32 + lane;
                  if (v < vend) acc8(acc[j], raw[dd][j]);
                }
              }
            }
          }
#pragma unroll
          for (int j = 0; j < kRedVec; ++j) {
            const int v = v0 + j * 32 + lane;
            if (v < vend) st_na_v4(dst + v, pack8_bf16(acc[j]));
          }
        }
      }
    }
    finish();  // the last warp of the whole grid to run out of work rewinds the counters for the next call
  }
}

static int grid_for(const b200ep_ctx* c, int items, int per_cta, int ctas_per_sm) {
  int sms = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
  int g = (items + per_cta - 1) / per_cta;
  int cap = sms * ctas_per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return g;
}

}  // namespace b200ep

using namespace b200ep;

static int check_slot(b200ep_ctx* c, int slot, const char* who) {
  if (!c || slot < 0 || slot >= kNumSlots) {
    set_error("%s: bad context or slot", who);
    return B200EP_ERR_ARG;
  }
  if (!c->connected) {
    set_error("%s: peers not connected (call b200ep_import_handles / b200ep_set_peer_ptrs first)", who);
    return B200EP_ERR_STATE;
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_notify(b200ep_ctx* c, int slot, const int32_t* dev_tokens_per_rank,
                                const int32_t* dev_tokens_per_expert, int E, int phases, void* stream) {
  int rc = check_slot(c, slot, "ht_notify");
  if (rc) return rc;
  if (E <= 0 || E % c->world || E / c->world > kMaxLocalExperts) {
    set_error("ht_notify: bad E=%d for world=%d", E, c->world);
    return B200EP_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  const int el = E / c->world;
  if (c->world == 1) {  // nothing to exchange: one kernel, issued with the first phase the caller asks for
    if (phases & B200EP_PHASE_SEND) {
      if (!dev_tokens_per_rank || !dev_tokens_per_expert) {
        set_error("ht_notify: null counts");
        return B200EP_ERR_ARG;
      }
      ++c->epoch[kFlagHtNotify][slot];
      c->ht_e_local[slot] = el;
      ht_notify_local_kernel<<<1, 256, 0, st>>>(dev_tokens_per_rank, dev_tokens_per_expert, slot, el, &c->local->ht[slot],
                                               c->host_dev, &c->local->epoch[kFlagHtNotify][slot]);
      B200EP_CUDA(cudaGetLastError());
    }
    return B200EP_OK;
  }
  if (phases & B200EP_PHASE_SEND) {
    if (!dev_tokens_per_rank || !dev_tokens_per_expert) {
      set_error("ht_notify: null counts");
      return B200EP_ERR_ARG;
    }
    ++c->epoch[kFlagHtNotify][slot];  // host mirror of the device epoch, for b200ep_ht_notify_wait
    c->ht_e_local[slot] = el;
    ht_notify_send_kernel<<<1, 256, 0, st>>>(c->peers, c->world, c->rank, slot, el, dev_tokens_per_rank,
                                             dev_tokens_per_expert, &c->local->epoch[kFlagHtNotify][slot]);
    B200EP_CUDA(cudaGetLastError());
  }
  if (phases & B200EP_PHASE_RECV) {
    ht_notify_recv_kernel<<<1, 256, 0, st>>>(c->sym, c->world, c->rank, slot, el, &c->local->ht[slot], c->host_dev,
                                             &c->local->epoch[kFlagHtNotify][slot]);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_dispatch(b200ep_ctx* c, int slot, const void* dev_x, const float* dev_x_scales, int x_kind,
                                  int ue8m0, const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                                  const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H, int K,
                                  int E, int phases, void* stream) {
  int rc = check_slot(c, slot, "ht_dispatch");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  if (phases & B200EP_PHASE_SEND) {
    HtPlan plan;
    rc = make_ht_plan(c, H, K, x_kind, &plan);
    if (rc) return rc;
    if (T < 0 || T > plan.t_cap) {
      set_error("ht_dispatch: T=%d exceeds slot capacity %d tokens/rank (H=%d K=%d kind=%d, region %zu bytes)", T,
                plan.t_cap, H, K, x_kind, c->sym_bytes);
      return B200EP_ERR_CAPACITY;
    }
    if (E <= 0 || E % c->world) {
      set_error("ht_dispatch: bad E=%d", E);
      return B200EP_ERR_ARG;
    }
    if (plan.row_bytes % 16) {
      set_error("ht_dispatch: row of %d bytes is not a multiple of 16", plan.row_bytes);
      return B200EP_ERR_ARG;
    }
    if (T > 0 && (!dev_x || !dev_topk_ids || !dev_is_token_in_rank || !dev_send_off ||
                  (x_kind == B200EP_X_FP8 && !dev_x_scales))) {
      set_error("ht_dispatch: null pointer");
      return B200EP_ERR_ARG;
    }
    c->ht_plan[slot] = plan;
    c->ht_tokens[slot] = T;
    HtDispatchArgs a;
    a.peers = c->peers;
    a.plan = plan;
    a.x = dev_x;
    a.x_scales = dev_x_scales;
    a.ids = dev_topk_ids;
    a.weights = dev_topk_weights;
    a.in_rank = dev_is_token_in_rank;
    a.send_off = dev_send_off;
    a.meta = &c->local->ht[slot];
    a.done = &c->local->done[0 + slot];
    a.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
    a.epoch_ptr = &c->local->epoch[kFlagHtData][slot];
    a.T = T;
    a.W = c->world;
    a.me = c->rank;
    a.slot = slot;
    a.e_local = E / c->world;
    a.ue8m0 = ue8m0;
    a.ids_i64 = ids_i64;
    const size_t stage_bytes = (size_t)(plan.qrow_bytes + 127) / 128 * 128;
    a.nstage = 2 * kTeams * stage_bytes <= 100 * 1024 ? 2 : 1;  // two CTAs per SM must still fit
    const size_t smem = (size_t)kTeams * a.nstage * stage_bytes;
    const int grid = grid_for(c, T, kTeams, x_kind == B200EP_X_BF16_QUANT ? 3 : 2);
    auto launch = [&](auto kern, int g) -> int {
      B200EP_CUDA(ensure_dynamic_smem((const void*)kern, c->device, (int)smem));
      kern<<<g, kTeams * kTeamThreads, smem, st>>>(a);
      B200EP_CUDA(cudaGetLastError());
      return B200EP_OK;
    };
    // Optional (B200EP_HT_STAGED=1): quantise once into the local q-stage, then push in permutation
    // order.  Measured at EP=8 it is SLOWER than the direct fan-out (326 vs 250 us per dispatch), so
    // the direct kernel stays the default; the staged form is kept for experiments.
    static const int staged_env = getenv("B200EP_HT_STAGED") ? atoi(getenv("B200EP_HT_STAGED")) : 0;
    const bool staged = staged_env != 0 && c->world > 1;
    // Experiment, off by default: measured 31.0 us at EP=1 against 26.5 us for the team kernel (fewer
    // instructions, 9.8 M vs 11.9 M, but each warp serialises load -> quantise over 3.5 passes per row).
    static const int warp_env = getenv("B200EP_HT_WARP_KERNEL") ? atoi(getenv("B200EP_HT_WARP_KERNEL")) : 0;
    if (!staged && x_kind == B200EP_X_BF16_QUANT && warp_env && kWarpCtaWarps * stage_bytes <= 220 * 1024) {
      const size_t wsmem = (size_t)kWarpCtaWarps * stage_bytes;
      B200EP_CUDA(ensure_dynamic_smem((const void*)ht_dispatch_send_warp_kernel, c->device, (int)wsmem));
      ht_dispatch_send_warp_kernel<<<grid_for(c, T, kWarpCtaWarps, 1), kWarpCtaWarps * 32, wsmem, st>>>(a);
      B200EP_CUDA(cudaGetLastError());
    } else if (!staged) {
      if (x_kind == B200EP_X_BF16_QUANT && H == 7168) rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16_QUANT, 7168>, grid);
      else if (x_kind == B200EP_X_BF16_QUANT) rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16_QUANT>, grid);
      else if (x_kind == B200EP_X_FP8) rc = launch(ht_dispatch_send_kernel<B200EP_X_FP8>, grid);
      else rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16>, grid);
    } else {
      if (x_kind == B200EP_X_BF16_QUANT) rc = launch(ht_dispatch_stage_kernel<B200EP_X_BF16_QUANT>, grid);
      else if (x_kind == B200EP_X_FP8) rc = launch(ht_dispatch_stage_kernel<B200EP_X_FP8>, grid);
      else rc = launch(ht_dispatch_stage_kernel<B200EP_X_BF16>, grid);
      if (rc) return rc;
      rc = launch(ht_dispatch_push_kernel, grid_for(c, T * 6, kTeams, 2));
    }
    if (rc) return rc;
  }
  if ((phases & B200EP_PHASE_RECV) && c->world > 1) {  // a single rank has nobody to wait for
    wait_sources_kernel<<<1, 32, 0, st>>>(c->sym, c->world, kFlagHtData, slot, c->rank, &c->local->epoch[kFlagHtData][slot]);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_views(b200ep_ctx* c, int slot, b200ep_ht_view* out) {
  if (!c || !out || slot < 0 || slot >= kNumSlots) {
    set_error("ht_views: bad argument");
    return B200EP_ERR_ARG;
  }
  const HtPlan& p = c->ht_plan[slot];
  if (p.H == 0) {
    set_error("ht_views: no dispatch has run on slot %d", slot);
    return B200EP_ERR_STATE;
  }
  char* b = slot_base(c, c->sym, slot);
  out->recv_x = b + p.off_x;
  out->recv_x_scales = p.scale_bytes ? reinterpret_cast<float*>(b + p.off_scales) : nullptr;
  out->recv_topk_idx = reinterpret_cast<int64_t*>(b + p.off_idx);
  out->recv_topk_weights = reinterpret_cast<float*>(b + p.off_w);
  out->recv_src_idx = reinterpret_cast<int32_t*>(b + p.off_src);
  out->rank_prefix = c->local->ht[slot].rank_prefix;
  out->recv_per_expert = c->local->ht[slot].recv_per_expert;
  out->capacity_rows = p.cap_rows;
  return B200EP_OK;
}

extern "C" int b200ep_ht_combine(b200ep_ctx* c, int slot, const void* dev_y, void* dev_out,
                                 const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H,
                                 int phases, void* stream) {
  int rc = check_slot(c, slot, "ht_combine");
  if (rc) return rc;
  const HtPlan& plan = c->ht_plan[slot];
  if (plan.H == 0 || plan.H != H || T != c->ht_tokens[slot]) {
    set_error("ht_combine: no matching dispatch on slot %d (H=%d/%d, T=%d/%d)", slot, H, plan.H, T, c->ht_tokens[slot]);
    return B200EP_ERR_STATE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  HtCombineArgs a;
  a.peers = c->peers;
  a.plan = plan;
  a.y = dev_y;
  a.out = dev_out;
  a.in_rank = dev_is_token_in_rank;
  a.send_off = dev_send_off;
  a.meta = &c->local->ht[slot];
  a.done = &c->local->done[2 + slot];
  a.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
  a.T = T;
  a.W = c->world;
  a.me = c->rank;
  a.slot = slot;
  a.epoch_ptr = &c->local->epoch[kFlagHtComb][slot];
  const int row_bytes = 2 * H;
  a.chunk_done = &c->local->done[16 + slot * kCombChunks];
  a.red_next = &c->local->done[32 + slot * 16];
  a.epoch_value = 0;
  static const int overlap_env = getenv("B200EP_HT_OVERLAP") ? atoi(getenv("B200EP_HT_OVERLAP")) : 1;
  if (phases == B200EP_PHASE_ALL && c->world > 1 && overlap_env) {
    // production path: push and reduce overlapped in one launch (see ht_combine_fused_kernel)
    if (T > 0 && (!dev_out || !dev_is_token_in_rank || !dev_send_off)) {
      set_error("ht_combine: null pointer");
      return B200EP_ERR_ARG;
    }
    a.epoch_value = ++c->epoch[kFlagHtComb][slot];
    a.nstage = 2 * (size_t)kTeams * row_bytes <= 200 * 1024 ? 2 : 1;
    const size_t smem = (size_t)kTeams * a.nstage * row_bytes;
    int grid = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
    if (grid < 4) grid = 4;
    B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_fused_kernel, c->device, (int)smem));
    ht_combine_fused_kernel<<<grid, kTeams * kTeamThreads, smem, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
    return B200EP_OK;
  }
  if ((phases & B200EP_PHASE_SEND) && c->world > 1) {  // a single rank pushes nothing: the reduce reads dev_y directly
    ++c->epoch[kFlagHtComb][slot];  // host mirror of the device epoch (used by the overlapped path)
    a.nstage = 2 * (size_t)kTeams * row_bytes <= 100 * 1024 ? 2 : 1;
    int ctas_per_sm = 2;
    if (a.nstage == 1 && 2 * (size_t)kTeams * row_bytes <= 200 * 1024) {  // wide rows: one CTA per SM, still two stages
      a.nstage = 2;
      ctas_per_sm = 1;
    }
    // M lives on the device; the host mirror of the last notify gives the grid size
    const int M = c->host->num_recv_cache[slot];
    const size_t smem = (size_t)kTeams * a.nstage * row_bytes;
    B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_send_kernel, c->device, (int)smem));
    ht_combine_send_kernel<<<grid_for(c, M, kTeams, ctas_per_sm), kTeams * kTeamThreads, smem, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
  }
  if (phases & B200EP_PHASE_RECV) {
    if (T > 0 && (!dev_out || !dev_is_token_in_rank || !dev_send_off)) {
      set_error("ht_combine: null pointer");
      return B200EP_ERR_ARG;
    }
    const int nsplit = row_bytes / 16 >= 256 ? 2 : 1;
    if (c->world == 1) ht_combine_recv_kernel<1><<<grid_for(c, T * nsplit, 8, 4), 256, 0, st>>>(a);
    else if (c->world == 2) ht_combine_recv_kernel<2><<<grid_for(c, T * nsplit, 8, 3), 256, 0, st>>>(a);
    else ht_combine_recv_kernel<kMaxRanks><<<grid_for(c, T * nsplit, 8, 2), 256, 0, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}


// One host call for the whole dispatch: layout -> count exchange -> host wait -> payload (+ wait).
// Same kernels as the separate entry points; exists because at EP=1 the step is bound by the
// caller's per-call overhead (ctypes crossings), not by the GPU.
extern "C" int b200ep_ht_dispatch_all(b200ep_ctx* c, int slot, const void* dev_x, const float* dev_x_scales, int x_kind,
                                      int ue8m0, const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                                      int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                                      uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, int T, int H, int K, int E,
                                      int* num_recv_tokens, int32_t* per_expert, int32_t* rank_prefix, int timeout_ms,
                                      void* stream) {
  int rc = b200ep_layout(c, dev_topk_ids, ids_i64, T, K, E, dev_tokens_per_rank, dev_tokens_per_expert,
                         dev_is_token_in_rank, dev_send_off, stream);
  if (rc) return rc;
  rc = b200ep_ht_notify(c, slot, dev_tokens_per_rank, dev_tokens_per_expert, E, B200EP_PHASE_ALL, stream);
  if (rc) return rc;
  if (timeout_ms >= 0) {
    rc = b200ep_ht_notify_wait(c, slot, num_recv_tokens, per_expert, rank_prefix, timeout_ms);
    if (rc) return rc;
  } else {
    // host-sync-free mode: the caller sized its views for a worst case; counts stay on the device
    // (b200ep_ht_views: rank_prefix[world] = rows received, recv_per_expert).  Later grids assume the worst.
    HtPlan plan;
    rc = make_ht_plan(c, H, K, x_kind, &plan);
    if (rc) return rc;
    c->host->num_recv_cache[slot] = plan.cap_rows;
  }
  return b200ep_ht_dispatch(c, slot, dev_x, dev_x_scales, x_kind, ue8m0, dev_topk_ids, ids_i64, dev_topk_weights,
                            dev_is_token_in_rank, dev_send_off, T, H, K, E, B200EP_PHASE_ALL, stream);
}

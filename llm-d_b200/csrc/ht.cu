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
//            every peer's control block and raises a flag (notify.cuh).  On the
//            production path this is the tail of the routing-layout kernel and the
//            prologue of the dispatch kernel: no launch of its own.
//   dispatch W > 1: persistent CTAs of four 128-thread teams, one row per team at a
//            time: 32-byte loads (LDG.E.256), all of a pass in flight -> fp32 amax per
//            128 group (8-lane shuffles) -> scale by a Markstein-corrected reciprocal
//            (== IEEE division, packed FFMA2) -> e4m3 into a double-buffered shared
//            memory stage -> TMA bulk copies (cp.async.bulk) per destination rank
//            straight into the peer's receive region at the row's final position
//            (payload, scales, and the per-destination top-k ids / weights rows,
//            which are staged in shared memory too).
//            W == 1: nothing crosses a link, so rows go from registers straight to
//            the receive region (no staging, no barriers; HBM-bound).
//   combine  expert ranks return result rows (bf16).  The rows a rank received from
//            source s are contiguous in its receive order, and they land on s in
//            the same order in stage[dst rank][send position]: the push is a
//            software DMA -- one thread per CTA drives TMA loads (global -> shared)
//            and TMA stores (shared -> peer) of multi-row pieces through a 4-stage
//            ring, no data touches registers.  The source rank sums the <= W rows
//            of a token in fp32 in rank order and writes bf16 once.  With all
//            phases in one call every CTA does both: thread 0 drives the DMA, the
//            other warps reduce chunks of 512 tokens as their per-chunk flags come up.
// Cross-GPU ordering: payload stores -> cp.async.bulk.wait_group -> atom.acq_rel.gpu
// on a counter -> (last arrival) ONE fence.acq_rel.sys -> st.relaxed.sys flag in
// every peer; consumers spin with ld.acquire.sys.
#include "ht_types.cuh"
#include "notify.cuh"

namespace b200ep {

// ------------------------------------------------------------------ notify (phase-by-phase callers)
__global__ void __launch_bounds__(256)
    ht_notify_send_kernel(Peers peers, int W, int me, int slot, int e_local, int T, const int32_t* __restrict__ tok_per_rank,
                          const int32_t* __restrict__ tok_per_expert, unsigned long long* epoch_ptr) {
  notify_publish(peers, W, me, slot, e_local, T, tok_per_rank, tok_per_expert, epoch_ptr);
}

__global__ void __launch_bounds__(256)
    ht_notify_recv_kernel(const char* my_base, int W, int me, int slot, int e_local, HtMeta* meta, HostMirror* host,
                          const unsigned long long* epoch_ptr) {
  notify_finish(const_cast<char*>(my_base), W, me, slot, e_local, meta, host,
                *reinterpret_cast<const volatile unsigned long long*>(epoch_ptr));
}

// world == 1: the whole count "exchange" in one small kernel.
__global__ void __launch_bounds__(256)
    ht_notify_local_kernel(const int32_t* __restrict__ tok_per_rank, const int32_t* __restrict__ tok_per_expert, int slot,
                           int e_local, int T, HtMeta* meta, HostMirror* host, unsigned long long* epoch_ptr) {
  const unsigned long long epoch = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
  notify_finish_local(tok_per_rank, tok_per_expert, slot, e_local, T, meta, host, epoch);
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
  HtMeta* meta;
  HostMirror* host;
  unsigned* done;
  size_t slot_off;  // kCtrlBytes + slot * slot_bytes
  unsigned long long* epoch_ptr;                  // data flags
  const unsigned long long* notify_epoch_ptr;     // count flags (prologue)
  int T, W, me, slot, e_local, ue8m0, ids_i64, nstage;
  int prologue;   // 1: wait for the peers' counts and derive the offsets here (no notify-recv launch)
  int rotate;     // 1: lane l of the issuing warp serves destination (me + 1 + l) % W instead of l
  unsigned long long* trace;  // B200EP_TRACE=1: %globaltimer stamps go here (measurement aid), else null
};

// What travels with a row besides its payload, as ONE record of plan.side_bytes (16-byte multiple), identical
// for every destination: [ top-k ids K x i64 (global) | top-k weights K x f32 | source token i32 | scales ].
// The sender stages it next to the row and issues exactly two TMA bulk copies per (row, destination); the
// receiver's wait kernel unpacks the records into the contiguous recv_topk_idx (destination-local ids, -1
// elsewhere) / recv_topk_weights (0 elsewhere) / recv_src_idx / recv_x_scales arrays vLLM reads.  (Per
// destination three more small bulk copies or plain stores cost ~6 % of the push: fabric probe, 625 vs 667 GB/s.)
__host__ __device__ inline int side_w_off(int K) { return K * 8; }
__host__ __device__ inline int side_src_off(int K) { return K * 12; }
__host__ __device__ inline int side_scale_off(int K) { return K * 12 + 4; }

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
__device__ __forceinline__ void store_pass(const int4 (&raw)[kVpt], int nvec, int v0, int ttid, char* s_row) {
#pragma unroll
  for (int u = 0; u < kVpt; ++u) {
    const int v = v0 + u * kTeamThreads + ttid;
    if (v < nvec) reinterpret_cast<int4*>(s_row)[v] = raw[u];
  }
}

// Quantising passes with 32-byte lanes (KIND == B200EP_X_BF16_QUANT): 2 x 128 x 32 B = 8 KB per pass.
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

__device__ __forceinline__ long long load_topk_id(const void* ids, int ids_i64, size_t i) {
  return ids_i64 ? reinterpret_cast<const long long*>(ids)[i] : (long long)reinterpret_cast<const int*>(ids)[i];
}

// Persistent CTAs of four 128-thread teams.  Per row a team: issues every load of the row (up to
// 16 KB in flight), waits until its previous TMA bulk copies have read the stage, quantises into
// the stage (warp 1 also builds the per-destination top-k rows there), and warp 0 fans the row out:
// lane d issues the bulk copies for destination rank d.  The next row's loads overlap the copies.
// kH > 0: hidden size known at compile time (the DeepSeek shape), which folds the per-vector bounds
// checks of the quantising passes; kH == 0: any hidden size.
template <int KIND, int kH = 0>
__global__ void __launch_bounds__(kTeams* kTeamThreads, KIND == B200EP_X_BF16_QUANT ? 3 : 2)
    ht_dispatch_send_kernel(const __grid_constant__ HtDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_send_base[kMaxRanks];
  const int tid = threadIdx.x, team = tid / kTeamThreads, ttid = tid % kTeamThreads;
  const int lane = tid & 31, twarp = ttid >> 5;
  const int W = a.W, K = a.plan.K, H = kH > 0 ? kH : a.plan.H;
  const int row_bytes = (kH > 0 && KIND == B200EP_X_BF16_QUANT) ? kH : a.plan.row_bytes;
  const int scale_bytes = (kH > 0 && KIND == B200EP_X_BF16_QUANT) ? kH / kGroup * 4 : a.plan.scale_bytes;
  const int side_bytes = (K * 12 + 4 + scale_bytes + 15) / 16 * 16;
  const int stage_bytes = (row_bytes + side_bytes + 127) / 128 * 128;
  char* s_team = s_mem + (size_t)team * a.nstage * stage_bytes;
  const int in_row_bytes = KIND == B200EP_X_BF16_QUANT ? 2 * H : row_bytes;
  const int nvec = in_row_bytes / 16;
  const char* my_base = a.peers.base[a.me];
  int n = 0;  // rows this team has staged so far: stage = n % nstage, one bulk group per row and lane

  const bool trace = a.trace != nullptr && blockIdx.x == 0 && tid == 0;
  auto stamp = [&](int i) { if (trace) a.trace[i] = globaltimer_ns(); };
  stamp(20);
  pdl_wait();  // launched early behind the layout kernel (b200ep_ht_dispatch_all): its outputs are needed from here on
  stamp(21);
  if (a.prologue) {
    // the peers' counts of THIS call (their layout kernels published them; ours did too, which is
    // what advanced *notify_epoch_ptr): wait, then every CTA derives its own send offsets
    const unsigned long long nep = *reinterpret_cast<const volatile unsigned long long*>(a.notify_epoch_ptr);
    if (tid < W) {  // lane d: rows the lower ranks send to rank d come before mine (waits word by word, no flag)
      int b = 0;
      for (int s = 0; s < a.me; ++s) b += notify_rank_count(const_cast<char*>(my_base), a.slot, s, tid, (unsigned)nep, a.me);
      s_send_base[tid] = b;
    }
    if (blockIdx.x == 0)  // receive-side bookkeeping and the host mirror (the host may be waiting for it)
      notify_finish(const_cast<char*>(my_base), W, a.me, a.slot, a.e_local, a.meta, a.host, nep);
  } else if (tid < W) {
    s_send_base[tid] = a.meta->send_base[tid];
  }
  __syncthreads();
  stamp(22);

  for (int t = blockIdx.x * kTeams + team; t < a.T; t += gridDim.x * kTeams) {
    // destination set of this token: lane d holds rank d's membership and row (every warp computes it)
    // lane l serves destination (me + 1 + l) % W when a.rotate is set: the lanes' copies enter the TMA queue in lane
    // order, so without the rotation every rank would address rank 0 first, then rank 1, ...
    int my_in = 0, my_row = -1;
    const int my_dst = a.rotate ? (a.me + 1 + lane) % W : lane;
    if (lane < W) {
      my_in = a.in_rank[(size_t)t * W + my_dst];
      if (my_in) my_row = s_send_base[my_dst] + a.send_off[(size_t)t * W + my_dst];
    }
    const unsigned dmask = __ballot_sync(0xffffffffu, my_in);
    if (dmask == 0) continue;  // uniform over the team

    // warp 1: this token's top-k ids and weights (loads in flight with the row)
    long long id = -1;
    float w = 0.f;
    if (twarp == 1 && lane < K) {
      id = load_topk_id(a.ids, a.ids_i64, (size_t)t * K + lane);
      w = a.weights ? a.weights[(size_t)t * K + lane] : 0.f;
    }

    char* s_row = s_team + (size_t)(a.nstage == 2 ? (n & 1) : 0) * stage_bytes;
    char* s_side = s_row + row_bytes;
    float* s_scale = reinterpret_cast<float*>(s_side + side_scale_off(K));
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
      for (int v0 = kTeamThreads * kVpt32; v0 < nv32; v0 += kTeamThreads * kVpt32) {  // rows wider than 8 KB
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
      store_pass(raw, nvec, 0, ttid, s_row);
      for (int v0 = kTeamThreads * kVpt; v0 < nvec; v0 += kTeamThreads * kVpt) {
        load_pass(src, nvec, v0, ttid, raw);
        store_pass(raw, nvec, v0, ttid, s_row);
      }
    }
    if (KIND == B200EP_X_FP8) {
      const int ng = H / kGroup;
      for (int g = ttid; g < ng; g += kTeamThreads) s_scale[g] = a.x_scales[(size_t)t * ng + g];
    }
    if (twarp == 1) {  // the rest of the side record: ids and weights as they are, and the source token
      if (lane < K) {
        reinterpret_cast<long long*>(s_side)[lane] = id;
        reinterpret_cast<float*>(s_side + side_w_off(K))[lane] = w;
      }
      if (lane == 0) *reinterpret_cast<int32_t*>(s_side + side_src_off(K)) = t;
    }
    fence_proxy_async_smem();
    team_bar(team);
    if (twarp == 0) {
      if (my_in) {
        char* dst = a.peers.base[my_dst] + a.slot_off;
        bulk_s2g(dst + a.plan.off_x + (size_t)my_row * row_bytes, s_row, row_bytes);
        bulk_s2g(dst + a.plan.off_side + (size_t)my_row * side_bytes, s_side, side_bytes);
      }
      bulk_commit();  // every lane, every row (possibly empty): group count == rows staged
    }
  }
  stamp(23);
  if ((ttid >> 5) == 0) bulk_wait_all();
  stamp(24);
  if (W > 1) cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtData, a.slot, a.me);  // a single rank has nobody to tell
  stamp(25);
}

// ---- world == 1, quantising kind: rows never leave the GPU, so there is nothing to stage.
// A warp takes (row, segment of up to 2 x 1024 values).  A lane owns 32 CONSECUTIVE values (two 32-byte
// loads, 64 bytes), so a 128-value group is 4 lanes: two shuffles for the amax, and the scale, its
// reciprocal and the address arithmetic are paid once per 32 values (with 16 values per lane that
// per-group overhead was ~40 of ~140 instructions per vector).  e4m3 bytes leave as one 32-byte store
// per lane (1 KB contiguous per warp), the scale from the group's first lane.  No shared memory, no
// barriers; one item per warp and as many CTAs as items, so the block scheduler balances the SMs
// (persistent warps with a static split or an atomic queue and register double-buffering were measured
// slower: 24.7 / 27.0 us against 22.7 us -- fewer resident warps cost more than the prefetch gains).
struct V16 {
  V8 a, b;
};
__device__ __forceinline__ void st_v8(void* p, const uint4& lo, const uint4& hi) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w),
               "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w)
               : "memory");
}
template <int kH, bool kUe8m0, int kLocalU = 1, int kMinCtas = 6>
__global__ void __launch_bounds__(256, kMinCtas) ht_dispatch_local_kernel(const __grid_constant__ HtDispatchArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int H = kH > 0 ? kH : a.plan.H, K = a.plan.K;
  const int np = H / 32, ng = H / kGroup;  // 64-byte pairs per row, groups per row (4 pairs each)
  const int nseg = (np + 32 * kLocalU - 1) / (32 * kLocalU);
  const long long it = (long long)blockIdx.x * 8 + warp;
  if (it >= (long long)a.T * nseg) return;
  const int t = (int)(it / nseg), seg = (int)(it % nseg);
  const V8* src = reinterpret_cast<const V8*>(reinterpret_cast<const __nv_bfloat16*>(a.x) + (size_t)t * H);
  V16 raw[kLocalU];
#pragma unroll
  for (int u = 0; u < kLocalU; ++u) {
    const int p = (seg * kLocalU + u) * 32 + lane;
    if (p < np) {
      raw[u].a = ld_nc_v8(src + 2 * p);
      raw[u].b = ld_nc_v8(src + 2 * p + 1);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) raw[u].a.w[i] = raw[u].b.w[i] = 0u;
    }
  }
  // for the row's first segment, the token's top-k entries: in flight together with the row
  long long id = -1;
  float w = 0.f;
  if (seg == 0 && lane < K) {
    id = load_topk_id(a.ids, a.ids_i64, (size_t)t * K + lane);
    w = a.weights ? a.weights[(size_t)t * K + lane] : 0.f;
  }
  // Everything above reads the caller's inputs only.  When this kernel was launched early behind the layout
  // kernel (programmatic dependent launch, b200ep_ht_dispatch_all) its loads are already in flight while the
  // layout runs; the layout's output -- this token's row, -1 if it has no valid expert -- is read after the wait.
  pdl_wait();
  const int row = *reinterpret_cast<const volatile int32_t*>(a.send_off + t);
  if (row < 0) return;  // uniform over the warp
  char* base = a.peers.base[a.me] + a.slot_off;
  char* out_x = base + a.plan.off_x + (size_t)row * H;
  float* out_s = reinterpret_cast<float*>(base + a.plan.off_scales) + (size_t)row * ng;
#pragma unroll
  for (int u = 0; u < kLocalU; ++u) {
    const int p = (seg * kLocalU + u) * 32 + lane;
    if ((seg * kLocalU + u) * 32 >= np) break;  // uniform over the warp
    const float s = group_scale(group_amax4(absmax32(raw[u].a, raw[u].b)), kUe8m0);
    const float r = __frcp_rn(s);
    if (p < np) {
      st_v8(out_x + (size_t)p * 32, quant16(raw[u].a, s, r), quant16(raw[u].b, s, r));
      if ((lane & 3) == 0) out_s[p >> 2] = s;
    }
  }
  if (seg == 0) {
    if (lane < K) {
      const bool mine = id >= 0 && id < a.e_local;
      reinterpret_cast<long long*>(base + a.plan.off_idx)[(size_t)row * K + lane] = mine ? id : -1;
      reinterpret_cast<float*>(base + a.plan.off_w)[(size_t)row * K + lane] = mine ? w : 0.f;
    }
    if (lane == 0) reinterpret_cast<int32_t*>(base + a.plan.off_src)[row] = t;
  }
}

// RECV phase of the dispatch: wait until every source rank's rows have landed, then unpack the side
// records of the rows received into the contiguous arrays the caller reads (the number of rows is
// read on the device, so the host-sync-free form needs nothing from the host).
struct HtRecvArgs {
  HtPlan plan;
  char* my_base;  // this rank's region
  size_t slot_off;
  const HtMeta* meta;
  const unsigned long long* epoch_ptr;
  int W, me, slot, e_local;
};
__global__ void __launch_bounds__(256) ht_dispatch_recv_kernel(const __grid_constant__ HtRecvArgs a) {
  if (a.W > 1) cta_wait_all_sources(a.my_base, a.W, kFlagHtData, a.slot, a.me, a.epoch_ptr);
  const int K = a.plan.K, ng = a.plan.scale_bytes / 4, side_bytes = a.plan.side_bytes;
  char* base = a.my_base + a.slot_off;
  const char* side0 = base + a.plan.off_side;
  const long long M = *reinterpret_cast<const volatile int32_t*>(&a.meta->num_recv);
  // flat index spaces, one element per thread and step: every load is independent of the others
  const long long me_g = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  long long* out_idx = reinterpret_cast<long long*>(base + a.plan.off_idx);
  float* out_w = reinterpret_cast<float*>(base + a.plan.off_w);
  for (long long i = me_g; i < M * K; i += nth) {
    const long long row = i / K;
    const int k = (int)(i % K);
    const char* side = side0 + (size_t)row * side_bytes;
    const long long id = __ldcg(reinterpret_cast<const long long*>(side) + k);
    const float w = __ldcg(reinterpret_cast<const float*>(side + side_w_off(K)) + k);
    const bool mine = id >= 0 && (int)(id / a.e_local) == a.me;
    out_idx[i] = mine ? id - (long long)a.me * a.e_local : -1;
    out_w[i] = mine ? w : 0.f;
  }
  float* out_s = reinterpret_cast<float*>(base + a.plan.off_scales);
  for (long long i = me_g; i < M * ng; i += nth) {
    const long long row = i / ng;
    out_s[i] = __ldcg(reinterpret_cast<const float*>(side0 + (size_t)row * side_bytes + side_scale_off(K)) + (int)(i % ng));
  }
  int32_t* out_src = reinterpret_cast<int32_t*>(base + a.plan.off_src);
  for (long long i = me_g; i < M; i += nth)
    out_src[i] = __ldcg(reinterpret_cast<const int32_t*>(side0 + (size_t)i * side_bytes + side_src_off(K)));
}

// ------------------------------------------------------------------ combine (argument block: ht_types.cuh)

// ---- software DMA: one thread moves pieces global -> shared -> peer global with TMA bulk copies.
// kDmaStages stages; loads run kDmaAhead pieces ahead of stores; a stage is loaded again only after the
// store that last used it has finished READING it (cp.async.bulk.wait_group.read): when load g is issued
// at least g - kDmaAhead stores are committed, so "at most kDmaStages - kDmaAhead - 1 pending" covers store g - kDmaStages.
constexpr int kDmaStages = 3;
constexpr int kDmaAhead = 1;
// After every store the driver waits until at most kDmaPending of its groups are still incomplete: with 2-row
// pieces that is > 100 KB per SM on the wire (the fabric needs ~25 KB per SM), so it never starves the link, and
// it makes "piece number stores - kDmaPending and everything older has LANDED" known at every step -- which is
// what the per-chunk flags of the overlapped combine need.  (Waiting only at chunk boundaries, down to two
// pending groups, drained the pipeline eight times per call: 470 us against 410 us for the push alone at EP=8.)
// The queue must stay SHALLOW: 600-680 GB/s x ~3 us of latency is ~2 MB for the whole GPU, i.e. one 14 KB row per
// SM.  With 4 pending groups of 2-row pieces per driver, 25 MB (of 58 MB at EP=2) sat in the fabric's queues: chunk 0
// was issued by 16 us but had landed only at 50 us, every flag came late and the reduction could not hide.
constexpr int kDmaPending = 2;
struct DmaPipe {
  char* stage;
  uint64_t* bar;
  unsigned stage_bytes;
  int loads, stores;
  char* dst[kDmaStages];
  unsigned bytes[kDmaStages];
};
__device__ __forceinline__ void dma_load(DmaPipe& p, const void* src, char* dst, unsigned bytes) {
  const int g = p.loads++, st = g % kDmaStages;
  if (g >= kDmaStages) bulk_wait_read<kDmaStages - kDmaAhead - 1>();
  p.dst[st] = dst;
  p.bytes[st] = bytes;
  mbar_expect_tx(&p.bar[st], bytes);
  bulk_g2s(p.stage + (size_t)st * p.stage_bytes, src, bytes, &p.bar[st]);
}
__device__ __forceinline__ void dma_store(DmaPipe& p) {
  const int g = p.stores++, st = g % kDmaStages;
  mbar_wait(&p.bar[st], (unsigned)(g / kDmaStages) & 1u);
  bulk_s2g(p.dst[st], p.stage + (size_t)st * p.stage_bytes, p.bytes[st]);
  bulk_commit();
  bulk_wait<kDmaPending>();
}
// pieces known to have landed (all of this thread's stores older than the kDmaPending most recent ones)
__device__ __forceinline__ int dma_landed(const DmaPipe& p) { return p.stores - kDmaPending; }

// Row ranges of chunk c (comb_chunk_begin of the SOURCE's token count) inside every remote source's block
// of this rank's receive order: s_lo[k][c] for source (me+1+k) % W.
__device__ __forceinline__ void combine_chunk_bounds(const HtCombineArgs& a, const int32_t* src_idx,
                                                     int (*s_lo)[kCombChunks + 1]) {
  const int W = a.W, tid = threadIdx.x;
  if (tid < (W - 1) * (kCombChunks + 1)) {
    const int k = tid / (kCombChunks + 1), c = tid % (kCombChunks + 1);
    const int s = (a.me + 1 + k) % W;
    int lo = a.meta->rank_prefix[s], hi = a.meta->rank_prefix[s + 1];
    if (c == kCombChunks) {
      lo = hi;
    } else {
      const int want = comb_chunk_begin(a.meta->src_tokens[s], c);
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldcg(src_idx + mid) < want) lo = mid + 1; else hi = mid;
      }
    }
    s_lo[k][c] = lo;
  }
}

// The driver thread's part of chunk c: pieces interleave the remote sources starting at rank me+1, so that
// at any moment the W ranks push along a permutation (all ranks converging on rank 0 first was measured 2.2x
// slower).  Pieces come from a queue (one atomic per piece): the SMs do not get the same share of the fabric,
// and with a static split the slowest driver decided when every chunk -- and the whole push -- was complete
// (trace at EP=2: CTA 0 done at 73 us, the kernel's last piece at ~105 us).  Returns the number of pieces stored.
template <typename AfterStore>
__device__ __forceinline__ int combine_push_chunk(const HtCombineArgs& a, DmaPipe& p, int (*s_lo)[kCombChunks + 1], int c,
                                                  AfterStore after_store) {
  const int W = a.W, R = a.rows_per_piece, row_bytes = 2 * a.plan.H;
  int maxp = 0;
  for (int k = 0; k < W - 1; ++k) maxp = max(maxp, (s_lo[k][c + 1] - s_lo[k][c] + R - 1) / R);
  const long long total = (long long)maxp * (W - 1);
  int loaded = 0, stored = 0;
  bool more = true;
  for (;;) {
    if (more && loaded - stored < kDmaAhead) {
      // next valid piece from the queue
      bool found = false;
      for (;;) {
        const long long i = (long long)atomicAdd(&a.piece_next[c], 1u);
        if (i >= total) break;
        const int k = (int)(i % (W - 1)), j = (int)(i / (W - 1));
        const int lo = s_lo[k][c] + j * R, hi = min(s_lo[k][c + 1], lo + R);
        if (lo >= hi) continue;
        const int s = (a.me + 1 + k) % W;
        const char* src = reinterpret_cast<const char*>(a.y) + (size_t)lo * row_bytes;
        char* dst = a.peers.base[s] + a.slot_off + stage_row_off(a.plan, a.me, lo - a.meta->rank_prefix[s], row_bytes);
        dma_load(p, src, dst, (unsigned)(hi - lo) * (unsigned)row_bytes);
        ++loaded;
        found = true;
        break;
      }
      if (found) continue;
      more = false;
    }
    if (stored < loaded) {
      dma_store(p);
      ++stored;
      after_store();
    } else if (!more) {
      break;
    }
  }
  return stored;
}

// Phase SEND on its own (phase-by-phase callers): push everything, then signal once.
__global__ void __launch_bounds__(128, 1) ht_combine_send_kernel(const __grid_constant__ HtCombineArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_lo[kMaxRanks][kCombChunks + 1];
  __shared__ __align__(8) uint64_t s_bar[kDmaStages];
  const int32_t* src_idx = reinterpret_cast<const int32_t*>(a.peers.base[a.me] + a.slot_off + a.plan.off_src);
  combine_chunk_bounds(a, src_idx, s_lo);
  if (threadIdx.x == 0) {
    for (int i = 0; i < kDmaStages; ++i) mbar_init(&s_bar[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    DmaPipe p;
    p.stage = s_mem;
    p.bar = s_bar;
    p.stage_bytes = a.stage_bytes;
    p.loads = p.stores = 0;
    for (int c = 0; c < kCombChunks; ++c) combine_push_chunk(a, p, s_lo, c, []() {});
    bulk_wait_all();
    // the last driver to leave the queues rewinds them for the next call
    if (atom_add_acq_rel_gpu(&a.piece_next[kCombChunks], 1u) == gridDim.x - 1)
      for (int c = 0; c <= kCombChunks; ++c) a.piece_next[c] = 0;
  }
  cta_done_then_signal(a.done, a.epoch_ptr, a.peers, a.W, kFlagHtComb, a.slot, a.me);
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

// out[t][vbeg..vend) = bf16( sum_{d in ranks(t), ascending} fp32(row_d) ): one warp; the row pointers are
// resolved once, then the warp walks the range in batches of kRedVec vectors per lane with the loads
// of up to four source rows in flight together.  kMaxSrc = ranks the code is built for.
constexpr int kRedVec = 4;
// where the rows of token t are: lane d resolves rank d's row (its own rows straight from y, the others from the
// stage), everybody gets all pointers; dmask = ranks that hold a row of this token
template <int kMaxSrc>
__device__ __forceinline__ unsigned reduce_resolve(const HtCombineArgs& a, const char* stage, int my_first, int t, int lane,
                                                   const char* (&rows)[kMaxSrc]) {
  const int W = a.W, row_bytes = 2 * a.plan.H;
  const char* rowp = nullptr;
  if (lane < W && a.in_rank[(size_t)t * W + lane]) {
    const int pos = a.send_off[(size_t)t * W + lane];
    rowp = (lane == a.me && !a.self_staged) ? reinterpret_cast<const char*>(a.y) + (size_t)(my_first + pos) * row_bytes
                                            : stage + ((size_t)lane * a.plan.t_cap + pos) * row_bytes;
  }
  const unsigned dmask = __ballot_sync(0xffffffffu, rowp != nullptr);
#pragma unroll
  for (int d = 0; d < kMaxSrc; ++d)
    rows[d] = reinterpret_cast<const char*>(__shfl_sync(0xffffffffu, (unsigned long long)rowp, d));
  return dmask;
}
template <int kMaxSrc>
__device__ __forceinline__ void reduce_accumulate(const HtCombineArgs& a, const char* const (&rows)[kMaxSrc], unsigned dmask,
                                                  int t, int vbeg, int vend, int lane) {
  constexpr int kGrp = kMaxSrc < 4 ? kMaxSrc : 4;
  const int row_bytes = 2 * a.plan.H;
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
template <int kMaxSrc>
__device__ __forceinline__ void reduce_item(const HtCombineArgs& a, const char* stage, int my_first, int t, int vbeg,
                                            int vend, int lane) {
  const char* rows[kMaxSrc];
  const unsigned dmask = reduce_resolve<kMaxSrc>(a, stage, my_first, t, lane, rows);
  reduce_accumulate<kMaxSrc>(a, rows, dmask, t, vbeg, vend, lane);
}

// Phase RECV on its own: wait for every peer's push, then reduce (static partition of the items).
template <int kMaxSrc>
__global__ void __launch_bounds__(256, kMaxSrc == 2 ? 3 : 2) ht_combine_recv_kernel(const __grid_constant__ HtCombineArgs a) {
  if (a.W > 1) cta_wait_all_sources(a.peers.base[a.me], a.W, kFlagHtComb, a.slot, a.me, a.epoch_ptr);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int nvec = 2 * a.plan.H / 16;
  const int nsplit = nvec >= 256 ? 2 : 1;
  const int per = (nvec / nsplit + 31) / 32 * 32;  // vectors per item, multiple of 32
  const char* stage = a.peers.base[a.me] + a.slot_off + a.plan.off_stage;
  const int my_first = a.meta->rank_prefix[a.me];
  const long long items = (long long)a.T * nsplit;
  for (long long it = (long long)blockIdx.x * nwarp + warp; it < items; it += (long long)gridDim.x * nwarp) {
    const int t = (int)(it / nsplit), vbeg = (int)(it % nsplit) * per;
    reduce_item<kMaxSrc>(a, stage, my_first, t, vbeg, min(nvec, vbeg + per), lane);
  }
}

// world == 1: the combine is a gather of this rank's own result rows, out[t] = bf16(0 + y[row(t)]) (the
// fp32 accumulate with a single term), zeros for tokens that were sent nowhere.  Same item shape as
// the local dispatch: (row, segment of 4 x 1 KB), 32-byte loads and stores, four in flight per lane.
constexpr int kCombU = 4;  // 32-byte vectors per lane and item
template <bool kPersistent>
__global__ void __launch_bounds__(256, kPersistent ? 3 : 4) ht_combine_local_kernel(const __grid_constant__ HtCombineArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int H = a.plan.H, nv32 = H / 16;
  const int nseg = (nv32 + 32 * kCombU - 1) / (32 * kCombU);
  const long long items = (long long)a.T * nseg, stride = kPersistent ? (long long)gridDim.x * 8 : items;
  long long it = (long long)blockIdx.x * 8 + warp;
  if (it >= items) return;
  auto load = [&](long long i, V8 (&raw)[kCombU]) {
    const int row = i < items ? __ldg(a.send_off + (int)(i / nseg)) : -1;
    const int seg = (int)(i % nseg);
    const V8* src = reinterpret_cast<const V8*>(reinterpret_cast<const __nv_bfloat16*>(a.y) + (size_t)(row < 0 ? 0 : row) * H);
#pragma unroll
    for (int u = 0; u < kCombU; ++u) {
      const int v = (seg * kCombU + u) * 32 + lane;
      if (row >= 0 && v < nv32) raw[u] = ld_nc_v8(src + v);
      else {
#pragma unroll
        for (int i2 = 0; i2 < 8; ++i2) raw[u].w[i2] = 0u;
      }
    }
  };
  V8 raw[kCombU], nxt[kCombU];
  load(it, raw);
  for (; it < items; it += stride) {
    const int t = (int)(it / nseg), seg = (int)(it % nseg);
    if (kPersistent) load(it + stride, nxt);  // next item's loads in flight while this one is converted and stored
    V8* dst = reinterpret_cast<V8*>(reinterpret_cast<__nv_bfloat16*>(a.out) + (size_t)t * H);
#pragma unroll
    for (int u = 0; u < kCombU; ++u) {
      const int v = (seg * kCombU + u) * 32 + lane;
      if (v >= nv32) continue;
      V8 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float lo = __fadd_rn(0.f, __uint_as_float(raw[u].w[i] << 16));
        const float hi = __fadd_rn(0.f, __uint_as_float(raw[u].w[i] & 0xffff0000u));
        __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
        o.w[i] = *reinterpret_cast<uint32_t*>(&p);
      }
      asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst + v), "r"(o.w[0]),
                   "r"(o.w[1]), "r"(o.w[2]), "r"(o.w[3]), "r"(o.w[4]), "r"(o.w[5]), "r"(o.w[6]), "r"(o.w[7])
                   : "memory");
    }
#pragma unroll
    for (int u = 0; u < kCombU; ++u) raw[u] = nxt[u];
  }
}

// ---- overlapped combine (production path, phases == ALL, W > 1): one launch, one CTA per SM.
//   thread 0 of every CTA drives the DMA push chunk by chunk of 512 source tokens; after chunk c+1 is
//            issued it waits until chunk c has landed (all but the most recent groups complete),
//            arrives on chunk c's counter, and the last CTA to arrive raises that chunk's flags on all
//            peers (one system fence per chunk; the pushing never waits for anybody);
//   the other warps (and warp 0 once its pushes are out) reduce chunk c as soon as every peer's
//            chunk-c flag is up; items come from a per-chunk queue.
// The fp32 reduction of the early chunks runs under the NVLink transfer of the later ones, and every
// SM keeps pushing for the whole transfer.  Nothing here waits for one of this rank's OWN CTAs, so
// the kernel completes with any number of resident CTAs (the GPU may be shared under DBO).
// The call epoch is read from device memory by every warp at entry and advanced by the last warp of
// the grid to finish, together with the rewind of the queues.
__global__ void __launch_bounds__(512, 1) ht_combine_fused_kernel(const __grid_constant__ HtCombineArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_lo[kMaxRanks][kCombChunks + 1];
  __shared__ __align__(8) uint64_t s_bar[kDmaStages];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W = a.W;
  const unsigned long long want = *reinterpret_cast<const volatile unsigned long long*>(a.epoch_ptr) + 1;
  const int32_t* src_idx = reinterpret_cast<const int32_t*>(a.peers.base[a.me] + a.slot_off + a.plan.off_src);
  const bool trace = a.trace != nullptr && blockIdx.x == 0;
  auto stamp = [&](int i) { if (trace) a.trace[i] = globaltimer_ns(); };
  if (tid == 0) stamp(30);
  combine_chunk_bounds(a, src_idx, s_lo);
  if (tid == 0) {
    for (int i = 0; i < kDmaStages; ++i) mbar_init(&s_bar[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) stamp(31);

  if (tid == 0) {
    // ------------------------------------------------ push (DMA driver)
    DmaPipe p;
    p.stage = s_mem;
    p.bar = s_bar;
    p.stage_bytes = a.stage_bytes;
    p.loads = p.stores = 0;
    auto arrive = [&](int c) {
      const unsigned prev = atom_add_acq_rel_gpu(&a.chunk_done[c], 1u);
      if (prev == gridDim.x - 1) {
        a.chunk_done[c] = 0;
        if (a.trace) a.trace[70 + c] = globaltimer_ns();
        fence_acq_rel_sys();
        for (int d = 0; d < W; ++d)
          if (d != a.me)
            st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(a.peers.base[d] + flag_off(kFlagHtCombChunk0 + c, a.slot, a.me)), want);
        if (a.trace) a.trace[80 + c] = globaltimer_ns();
      }
    };
    int end_seq[kCombChunks];  // stores issued by the end of chunk c: chunk c has landed once dma_landed() >= end_seq[c]
    int next_arrive = 0;
    for (int c = 0; c < kCombChunks; ++c) {
      combine_push_chunk(a, p, s_lo, c, [&]() {
        while (next_arrive < c && end_seq[next_arrive] <= dma_landed(p)) arrive(next_arrive++);
      });
      end_seq[c] = p.stores;
      stamp(40 + c);
    }
    bulk_wait_all();
    while (next_arrive < kCombChunks) arrive(next_arrive++);
    stamp(32);
  }
  __syncwarp();

  // ------------------------------------------------ reduce (warps 1..15 of every CTA; warp 0 only drives the DMA)
  // An item = (token, 2 KB piece of its row): one batch of loads per source row, so an item costs about one
  // memory round trip.  Items of a chunk are split statically over the reducer warps of the grid, and a warp
  // resolves the row pointers of its first item BEFORE it waits for the chunk's flags: what is left on the
  // critical path after the last flag is one round trip and the stores.  (With half rows handed out from a
  // queue an item took ~5 us -- atomic, index loads, 3.5 batches -- and the last chunk ~10 us.)
  if (warp > 0) {
    const int nvec = 2 * a.plan.H / 16;
    const int per = 32 * kRedVec;
    const int nsplit = (nvec + per - 1) / per;
    const char* stage = a.peers.base[a.me] + a.slot_off + a.plan.off_stage;
    const int my_first = a.meta->rank_prefix[a.me];
    const int nred = gridDim.x * ((blockDim.x >> 5) - 1), w_g = blockIdx.x * ((blockDim.x >> 5) - 1) + warp - 1;
    for (int c = 0; c < kCombChunks; ++c) {
      const int t0 = comb_chunk_begin(a.T, c), t1 = comb_chunk_begin(a.T, c + 1);
      const int items = (t1 - t0) * nsplit;
      int it = w_g;
      const char* rows[kMaxRanks];
      unsigned dmask = 0;
      if (it < items) dmask = reduce_resolve<kMaxRanks>(a, stage, my_first, t0 + it / nsplit, lane, rows);
      if (lane < W && lane != a.me)
        wait_flag_sys(reinterpret_cast<const uint64_t*>(a.peers.base[a.me] + flag_off(kFlagHtCombChunk0 + c, a.slot, lane)), want,
                      kFlagHtCombChunk0 + c, a.me, lane);
      __syncwarp();
      if (warp == 1 && lane == 0) stamp(50 + c);
      while (it < items) {
        const int vbeg = (it % nsplit) * per;
        reduce_accumulate<kMaxRanks>(a, rows, dmask, t0 + it / nsplit, vbeg, min(nvec, vbeg + per), lane);
        it += nred;
        if (it < items) dmask = reduce_resolve<kMaxRanks>(a, stage, my_first, t0 + it / nsplit, lane, rows);
      }
      if (warp == 1 && lane == 0) stamp(60 + c);
    }
  }
  // the last warp of the whole grid to finish rewinds the queues and advances the epoch (every warp has read the
  // old epoch by then)
  if (lane == 0) {
    const unsigned prev = atom_add_acq_rel_gpu(&a.red_next[kCombChunks], 1u);
    if (prev == gridDim.x * (blockDim.x >> 5) - 1) {
      a.red_next[kCombChunks] = 0;
      for (int c = 0; c <= kCombChunks; ++c) a.piece_next[c] = 0;
      *reinterpret_cast<volatile unsigned long long*>(a.epoch_ptr) = want;
    }
  }
}

static int grid_for(const b200ep_ctx* c, long long items, int per_cta, int ctas_per_sm) {
  int sms = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
  long long g = (items + per_cta - 1) / per_cta;
  long long cap = (long long)sms * ctas_per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace b200ep

using namespace b200ep;

namespace b200ep {
int launch_layout(b200ep_ctx* c, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                  int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert, uint8_t* dev_is_token_in_rank,
                  int32_t* dev_send_off, const NotifyTail& tail, cudaStream_t st);
}

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
                                const int32_t* dev_tokens_per_expert, int T, int E, int phases, void* stream) {
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
      ht_notify_local_kernel<<<1, 256, 0, st>>>(dev_tokens_per_rank, dev_tokens_per_expert, slot, el, T, &c->local->ht[slot],
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
    ht_notify_send_kernel<<<1, 256, 0, st>>>(c->peers, c->world, c->rank, slot, el, T, dev_tokens_per_rank,
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

// Launch, optionally as a programmatic dependent of the kernel queued just before it on the stream.
template <typename Kern, typename Args>
static cudaError_t launch_kernel(Kern kern, int grid, int block, size_t smem, cudaStream_t st, bool pdl, const Args& a) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, a);
}

// SEND / RECV launches of the dispatch.  prologue: the kernel itself waits for the peers' counts
// (production path, b200ep_ht_dispatch_all) instead of relying on an earlier notify-recv launch.
static int dispatch_launch(b200ep_ctx* c, int slot, const void* dev_x, const float* dev_x_scales, int x_kind, int ue8m0,
                           const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                           const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H, int K, int E,
                           int phases, int prologue, cudaStream_t st) {
  int rc;
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
    if (T > 0 && ((uintptr_t)dev_x & 31)) {
      set_error("ht_dispatch: x must be 32-byte aligned");
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
    a.host = c->host_dev;
    a.done = &c->local->done[0 + slot];
    a.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
    a.epoch_ptr = &c->local->epoch[kFlagHtData][slot];
    a.notify_epoch_ptr = &c->local->epoch[kFlagHtNotify][slot];
    a.T = T;
    a.W = c->world;
    a.me = c->rank;
    a.slot = slot;
    a.e_local = E / c->world;
    a.ue8m0 = ue8m0;
    a.ids_i64 = ids_i64;
    a.prologue = (prologue && c->world > 1) ? 1 : 0;
    // measured at EP=8: dispatch push 248 -> 221 us, whole step 763 -> 734 us (no difference at EP=2)
    static const int rotate_env = getenv("B200EP_HT_ROTATE") ? atoi(getenv("B200EP_HT_ROTATE")) : 1;
    a.rotate = rotate_env;
    static const int trace_env = getenv("B200EP_TRACE") ? atoi(getenv("B200EP_TRACE")) : 0;
    a.trace = trace_env ? c->local->trace : nullptr;
    // called from b200ep_ht_dispatch_all the layout kernel is queued right in front: start early behind it
    static const int pdl_env = getenv("B200EP_PDL") ? atoi(getenv("B200EP_PDL")) : 1;
    const bool pdl = prologue && pdl_env;
    if (c->world == 1 && x_kind == B200EP_X_BF16_QUANT) {
      // single rank: registers -> receive region, no staging
      a.nstage = 0;
      // one 64-byte pair per lane and item, 6 CTAs per SM (40 registers): measured 18.4 us at 4096 x 7168 against
      // 20.4 us with two pairs per lane at 3 CTAs per SM -- resident warps matter more than bytes in flight per warp
      const int U = 1;
      const int nseg = (H / 32 + 32 * U - 1) / (32 * U);
      const long long items = (long long)T * nseg;
      const int grid = (int)((items + 7) / 8 > 0 ? (items + 7) / 8 : 1);
      if (H == 7168) {
        if (ue8m0) B200EP_CUDA(launch_kernel(ht_dispatch_local_kernel<7168, true>, grid, 256, 0, st, pdl, a));
        else B200EP_CUDA(launch_kernel(ht_dispatch_local_kernel<7168, false>, grid, 256, 0, st, pdl, a));
      } else {
        if (ue8m0) B200EP_CUDA(launch_kernel(ht_dispatch_local_kernel<0, true>, grid, 256, 0, st, pdl, a));
        else B200EP_CUDA(launch_kernel(ht_dispatch_local_kernel<0, false>, grid, 256, 0, st, pdl, a));
      }
      B200EP_CUDA(cudaGetLastError());
    } else {
      const size_t stage_bytes = (size_t)(plan.row_bytes + plan.side_bytes + 127) / 128 * 128;
      static const int ctas_env = getenv("B200EP_HT_DISPATCH_CTAS") ? atoi(getenv("B200EP_HT_DISPATCH_CTAS")) : 0;
      const int max_ctas = x_kind == B200EP_X_BF16_QUANT ? 3 : 2;
      const int ctas_per_sm = (ctas_env > 0 && ctas_env < max_ctas) ? ctas_env : max_ctas;
      a.nstage = 2 * kTeams * stage_bytes * max_ctas <= 216 * 1024 ? 2 : 1;  // every resident CTA must still fit
      static const int nstage_env = getenv("B200EP_HT_DISPATCH_STAGES") ? atoi(getenv("B200EP_HT_DISPATCH_STAGES")) : 0;
      if (nstage_env == 1) a.nstage = 1;
      const size_t smem = (size_t)kTeams * a.nstage * stage_bytes;
      if (smem > 220 * 1024) {
        set_error("ht_dispatch: a row of %d bytes does not fit the shared-memory stage", plan.row_bytes);
        return B200EP_ERR_ARG;
      }
      const int grid = grid_for(c, T, kTeams, ctas_per_sm);
      auto launch = [&](auto kern) -> int {
        B200EP_CUDA(ensure_dynamic_smem((const void*)kern, c->device, (int)smem));
        B200EP_CUDA(launch_kernel(kern, grid, kTeams * kTeamThreads, smem, st, pdl, a));
        return B200EP_OK;
      };
      if (x_kind == B200EP_X_BF16_QUANT && H == 7168) rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16_QUANT, 7168>);
      else if (x_kind == B200EP_X_BF16_QUANT) rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16_QUANT>);
      else if (x_kind == B200EP_X_FP8) rc = launch(ht_dispatch_send_kernel<B200EP_X_FP8>);
      else rc = launch(ht_dispatch_send_kernel<B200EP_X_BF16>);
      if (rc) return rc;
    }
  }
  if (phases & B200EP_PHASE_RECV) {
    const HtPlan& plan = c->ht_plan[slot];
    if (plan.H == 0) {
      set_error("ht_dispatch: RECV phase without a SEND phase on slot %d", slot);
      return B200EP_ERR_STATE;
    }
    // the single-rank quantising kernel writes the final arrays itself: nothing to wait for, nothing to unpack
    if (!(c->world == 1 && plan.x_kind == B200EP_X_BF16_QUANT)) {
      HtRecvArgs r;
      r.plan = plan;
      r.my_base = c->sym;
      r.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
      r.meta = &c->local->ht[slot];
      r.epoch_ptr = &c->local->epoch[kFlagHtData][slot];
      r.W = c->world;
      r.me = c->rank;
      r.slot = slot;
      r.e_local = E / c->world;
      // rows are counted on the device (they do not depend on what THIS rank sent): two CTAs per SM stride over them
      ht_dispatch_recv_kernel<<<grid_for(c, 1LL << 40, 8, 2), 256, 0, st>>>(r);
      B200EP_CUDA(cudaGetLastError());
    }
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_dispatch(b200ep_ctx* c, int slot, const void* dev_x, const float* dev_x_scales, int x_kind,
                                  int ue8m0, const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                                  const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H, int K,
                                  int E, int phases, void* stream) {
  int rc = check_slot(c, slot, "ht_dispatch");
  if (rc) return rc;
  B200EP_CUDA(cudaSetDevice(c->device));
  return dispatch_launch(c, slot, dev_x, dev_x_scales, x_kind, ue8m0, dev_topk_ids, ids_i64, dev_topk_weights,
                         dev_is_token_in_rank, dev_send_off, T, H, K, E, phases, 0, (cudaStream_t)stream);
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

namespace b200ep {
int ht_combine_args(b200ep_ctx* c, int slot, const void* dev_y, void* dev_out, const uint8_t* dev_is_token_in_rank,
                    const int32_t* dev_send_off, int T, int H, HtCombineArgs* out) {
  const HtPlan& plan = c->ht_plan[slot];
  if (plan.H == 0 || plan.H != H || T != c->ht_tokens[slot]) {
    set_error("ht_combine: no matching dispatch on slot %d (H=%d/%d, T=%d/%d)", slot, H, plan.H, T, c->ht_tokens[slot]);
    return B200EP_ERR_STATE;
  }
  HtCombineArgs& a = *out;
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
  static const int trace_env = getenv("B200EP_TRACE") ? atoi(getenv("B200EP_TRACE")) : 0;
  a.trace = trace_env ? c->local->trace : nullptr;
  a.chunk_done = &c->local->done[16 + slot * kCombChunks];
  a.red_next = &c->local->done[32 + slot * 16];
  a.piece_next = &c->local->done[72 + slot * 16];
  a.self_staged = 0;
  const int row_bytes = 2 * H;
  // DMA pieces: whole rows, up to 45 KB at EP >= 4 (three DeepSeek rows) and 15 KB below.  Measured at EP=8 (whole step,
  // same build): 1 row 763 us, 2 rows 740 us, 3 rows 709 us together with the rotated dispatch -- the combine ends ~25 us
  // earlier, the push kernel alone does not change (410 us): fewer, larger pieces leave the reducing warps more of the
  // L2 / HBM bandwidth and cost fewer queue atomics and mbarrier round trips per byte.  At EP=2 larger pieces were
  // slightly slower (219.8 vs 218.0 us), hence the split.
  const int piece_target = (c->world >= 4 ? 45 : 15) * 1024;
  a.rows_per_piece = piece_target / row_bytes;
  if (a.rows_per_piece < 1) a.rows_per_piece = 1;
  static const int piece_env = getenv("B200EP_HT_PIECE_ROWS") ? atoi(getenv("B200EP_HT_PIECE_ROWS")) : 0;
  if (piece_env > 0) a.rows_per_piece = piece_env;
  while (a.rows_per_piece > 1 && (size_t)a.rows_per_piece * row_bytes * kDmaStages > 200 * 1024) --a.rows_per_piece;
  a.stage_bytes = (unsigned)a.rows_per_piece * (unsigned)row_bytes;
  return B200EP_OK;
}

// RECV phase at W > 1: wait for every peer's push, then the fp32 reduction in rank order.
int ht_combine_launch_recv(b200ep_ctx* c, const HtCombineArgs& a, cudaStream_t st) {
  const int row_bytes = 2 * a.plan.H;
  const int nsplit = row_bytes / 16 >= 256 ? 2 : 1;
  if (c->world == 2) ht_combine_recv_kernel<2><<<grid_for(c, (long long)a.T * nsplit, 8, 3), 256, 0, st>>>(a);
  else ht_combine_recv_kernel<kMaxRanks><<<grid_for(c, (long long)a.T * nsplit, 8, 2), 256, 0, st>>>(a);
  B200EP_CUDA(cudaGetLastError());
  return B200EP_OK;
}
}  // namespace b200ep

extern "C" int b200ep_ht_combine(b200ep_ctx* c, int slot, const void* dev_y, void* dev_out,
                                 const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H,
                                 int phases, void* stream) {
  int rc = check_slot(c, slot, "ht_combine");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  HtCombineArgs a;
  rc = ht_combine_args(c, slot, dev_y, dev_out, dev_is_token_in_rank, dev_send_off, T, H, &a);
  if (rc) return rc;
  const int row_bytes = 2 * H;
  const size_t dma_smem = (size_t)kDmaStages * a.stage_bytes;
  const bool any = phases & (B200EP_PHASE_SEND | B200EP_PHASE_RECV);
  if (any && T > 0 && (!dev_out || !dev_is_token_in_rank || !dev_send_off) && (phases & B200EP_PHASE_RECV)) {
    set_error("ht_combine: null pointer");
    return B200EP_ERR_ARG;
  }
  if (c->world == 1) {  // a single rank pushes nothing: the gather reads dev_y directly
    if (phases & B200EP_PHASE_RECV) {
      if (T > 0 && (((uintptr_t)dev_y | (uintptr_t)dev_out) & 31)) {
        set_error("ht_combine: y and out must be 32-byte aligned");
        return B200EP_ERR_ARG;
      }
      const int nseg = (H / 16 + 32 * kCombU - 1) / (32 * kCombU);
      const long long items = (long long)T * nseg;
      ht_combine_local_kernel<false><<<(int)((items + 7) / 8 > 0 ? (items + 7) / 8 : 1), 256, 0, st>>>(a);
      B200EP_CUDA(cudaGetLastError());
    }
    return B200EP_OK;
  }
  if ((phases & B200EP_PHASE_SEND) && (!dev_y || ((uintptr_t)dev_y & 15))) {
    // (dev_y may only be null when this rank received nothing; a null pointer is never dereferenced then)
    if (dev_y || c->ht_tokens[slot] < 0) {
      set_error("ht_combine: y must be 16-byte aligned");
      return B200EP_ERR_ARG;
    }
  }
  if (dma_smem > 200 * 1024) {
    set_error("ht_combine: a row of %d bytes does not fit the DMA stages", row_bytes);
    return B200EP_ERR_ARG;
  }
  int grid = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
  static const int overlap_env = getenv("B200EP_HT_OVERLAP") ? atoi(getenv("B200EP_HT_OVERLAP")) : 1;
  if (phases == B200EP_PHASE_ALL && overlap_env) {
    // production path: push and reduce overlapped in one launch (see ht_combine_fused_kernel)
    B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_fused_kernel, c->device, (int)dma_smem));
    ht_combine_fused_kernel<<<grid, 512, dma_smem, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
    return B200EP_OK;
  }
  if (phases & B200EP_PHASE_SEND) {
    B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_send_kernel, c->device, (int)dma_smem));
    ht_combine_send_kernel<<<grid, 128, dma_smem, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
  }
  if (phases & B200EP_PHASE_RECV) return ht_combine_launch_recv(c, a, st);
  return B200EP_OK;
}

// One host call for the whole dispatch (deep_ep.Buffer.get_dispatch_layout + Buffer.dispatch,
// deepep_ht.py:121-162): the layout kernel (whose last block publishes the counts), the dispatch kernel
// (which waits for the peers' counts itself) and the wait for the payload are all queued BEFORE the
// host waits for the counts, so the host round trip runs under the dispatch kernel instead of in front of it.
extern "C" int b200ep_ht_dispatch_all(b200ep_ctx* c, int slot, const void* dev_x, const float* dev_x_scales, int x_kind,
                                      int ue8m0, const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                                      int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                                      uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, int T, int H, int K, int E,
                                      int* num_recv_tokens, int32_t* per_expert, int32_t* rank_prefix, int timeout_ms,
                                      void* stream) {
  int rc = check_slot(c, slot, "ht_dispatch_all");
  if (rc) return rc;
  if (E <= 0 || E % c->world || E / c->world > kMaxLocalExperts) {
    set_error("ht_dispatch_all: bad E=%d for world=%d", E, c->world);
    return B200EP_ERR_ARG;
  }
  {  // validate the shape before anything is queued: a failure after the counts are out would leave the peers waiting
    HtPlan plan;
    rc = make_ht_plan(c, H, K, x_kind, &plan);
    if (rc) return rc;
    if (T < 0 || T > plan.t_cap) {
      set_error("ht_dispatch: T=%d exceeds slot capacity %d tokens/rank (H=%d K=%d kind=%d, region %zu bytes)", T,
                plan.t_cap, H, K, x_kind, c->sym_bytes);
      return B200EP_ERR_CAPACITY;
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  NotifyTail tail;
  tail.peers = c->peers;
  tail.meta = &c->local->ht[slot];
  tail.host = c->host_dev;
  tail.epoch_ptr = &c->local->epoch[kFlagHtNotify][slot];
  tail.mode = c->world == 1 ? kTailLocal : kTailPublish;
  tail.me = c->rank;
  tail.slot = slot;
  ++c->epoch[kFlagHtNotify][slot];  // host mirror of the device epoch, for b200ep_ht_notify_wait
  c->ht_e_local[slot] = E / c->world;
  rc = launch_layout(c, dev_topk_ids, ids_i64, T, K, E, dev_tokens_per_rank, dev_tokens_per_expert, dev_is_token_in_rank,
                     dev_send_off, tail, st);
  if (rc) return rc;
  rc = dispatch_launch(c, slot, dev_x, dev_x_scales, x_kind, ue8m0, dev_topk_ids, ids_i64, dev_topk_weights,
                       dev_is_token_in_rank, dev_send_off, T, H, K, E, B200EP_PHASE_ALL, 1, st);
  if (rc) return rc;
  if (timeout_ms >= 0) return b200ep_ht_notify_wait(c, slot, num_recv_tokens, per_expert, rank_prefix, timeout_ms);
  // host-sync-free mode: the caller sized its views for a worst case; counts stay on the device
  // (b200ep_ht_views: rank_prefix[world] = rows received, recv_per_expert)
  return B200EP_OK;
}

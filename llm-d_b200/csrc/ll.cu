// Low-latency (decode) dispatch and combine (north-star kernels 2 and 3, LL mode).
//
// Replaces deep_ep.Buffer.low_latency_dispatch / low_latency_combine as vLLM calls
// them (vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ll.py:292-314,
// 399-408).  Output layout = BatchedExperts, the one
// prepare_finalize/batched.py:82-152 builds: packed_recv_x[e_local, slot, :] with
// the tokens of an expert packed at slots [0, count).  CPU restatement:
// oracle/dispatch.py (ll_dispatch), oracle/combine.py (ll_combine).
//
// No host synchronisation, fixed shapes, no allocation, counters keyed by a call
// epoch instead of being reset: the launches replay unchanged under a CUDA graph.
//
// Order is canonical (source rank ascending, source token ascending) and needs no
// atomics: every rank first publishes its per-expert counts to all peers (1 KB each),
// so a sender can compute the final slot of each of its (token, expert) rows,
//     slot = sum_{s < me} count[s][e] + (position of the token among my tokens for e),
// and writes the row ONCE, straight to that slot on the owner (no receive-side pack).
//   COUNT  one CTA: ballot/popc positions per expert, counts -> peers, flag.
//   SEND   one CTA per token: quantise into shared memory while the counts are in
//          flight, wait for the count flags, then K TMA bulk copies (cp.async.bulk)
//          smem -> peer; last CTA raises the data flags.
//   RECV   wait for the data flags (DeepEP's "recv hook").
// Combine: owners push occupied result rows back into stage[token][k] on the source
// (same staging + bulk copy); the source does sum_k w_k * y_k in fp32 (multiply, then
// add, k ascending) and writes bf16 once.
#include <stdlib.h>

#include "common.cuh"

namespace b200ep {

constexpr int kLlMaxExperts = kMaxLocalExperts;  // global experts in LL mode (count row of the control block)
constexpr int kLlMaxTokens = 1024;

struct LlScratch {  // private, per slot
  int32_t cnt_mine[kLlMaxExperts];
  int32_t pos[kLlMaxTokens * kMaxTopk];
  int32_t send_slot[kLlMaxTokens * kMaxTopk];
  int32_t recv_count[kMaxLocalExperts];
  int32_t layout_range[kMaxLocalExperts * kMaxRanks * 2];
};
static_assert(sizeof(LlScratch) * kNumSlots <= (4u << 20), "ll scratch");

// Per-expert counts travel as self-validating words (call tag << 32 | count), one 8-byte store per (expert, peer):
// the receiver polls the words it needs, so the count exchange costs no system fence and no flag round trip.
__device__ __forceinline__ uint64_t* ll_count_word(char* base, int slot, int src, int e) {
  return reinterpret_cast<uint64_t*>(base + kCtrlLlCntOff) + ((size_t)slot * kMaxRanks + src) * kLlMaxExperts + e;
}
__device__ __forceinline__ uint64_t ll_count_pack(unsigned long long tag, int n) {
  return ((tag & 0xffffffffull) << 32) | (uint32_t)n;
}

// Positions and counts from per-expert token bit masks: every (token, k) entry ORs its token's bit
// into mask[e] (shared-memory atomics -- the resulting mask does not depend on arrival order), then
// pos(t, e) = number of set bits below t and count[e] = popcount(mask[e]).  O(T*K) work, no scan loops.
template <typename IdT>
__global__ void __launch_bounds__(1024)
    ll_count_kernel(Peers peers, int W, int me, int slot, const IdT* __restrict__ ids, int T, int K, int E,
                    LlScratch* sc, unsigned long long* epoch_ptr) {
  extern __shared__ unsigned s_mask[];  // [E][nw]
  const int tid = threadIdx.x;
  const int nw = (T + 31) / 32;
  for (int i = tid; i < E * nw; i += blockDim.x) s_mask[i] = 0u;
  __syncthreads();
  for (int i = tid; i < T * K; i += blockDim.x) {
    const long long v = (long long)ids[i];
    if (v >= 0 && v < E) atomicOr(&s_mask[(int)v * nw + (i / K) / 32], 1u << ((i / K) & 31));
  }
  __syncthreads();
  for (int i = tid; i < T * K; i += blockDim.x) {
    const long long v = (long long)ids[i];
    if (v >= 0 && v < E) {
      const int t = i / K;
      const unsigned* m = &s_mask[(int)v * nw];
      int p = __popc(m[t >> 5] & ((1u << (t & 31)) - 1u));
      for (int w = 0; w < (t >> 5); ++w) p += __popc(m[w]);
      sc->pos[i] = p;
    }
  }
  const unsigned long long tag = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
  for (int e = tid; e < E; e += blockDim.x) {
    int n = 0;
    for (int w = 0; w < nw; ++w) n += __popc(s_mask[e * nw + w]);
    sc->cnt_mine[e] = n;
    for (int d = 0; d < W; ++d) st_relaxed_sys_u64(ll_count_word(peers.base[d], slot, me, e), ll_count_pack(tag, n));
  }
  __syncthreads();
  if (tid == 0) *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) = tag;
}

struct LlDispatchArgs {
  Peers peers;
  LlPlan plan;
  const __nv_bfloat16* x;
  const void* ids;
  LlScratch* sc;
  unsigned* done;
  size_t slot_off;
  unsigned long long *epoch_cnt_ptr, *epoch_data_ptr;
  int T, W, me, slot, E, ue8m0, ids_i64, fused, wait_data;
  unsigned long long* trace;  // B200EP_TRACE=1: %globaltimer stamps go here (measurement aid), else null
};

// fused == 1: the kernel also does the COUNT phase (production path, one launch for count + send):
//   * the extra last CTA histograms this rank's ids, writes the counts to every peer and raises the
//     count flags (its system fence overlaps the other CTAs' quantisation);
//   * every token CTA computes the positions of its own token (how many earlier tokens of this rank
//     chose the same expert) by itself, so it depends on nothing local.
// wait_data == 1: the last CTA to finish also waits for every source's data flag, i.e. the kernel
// completes when this rank's receive region is complete (no separate RECV launch).
template <bool kFp8>
__global__ void __launch_bounds__(256) ll_dispatch_send_kernel(const __grid_constant__ LlDispatchArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_pos[kMaxTopk];
  __shared__ long long s_myid[kMaxTopk];
  __shared__ int s_slot[kMaxTopk];  // packed-UE8M0 mode: flat index (e_local * g4) * cap4 + slot of entry k, -1 if none
  __shared__ int s_dst[kMaxTopk];
  const int tid = threadIdx.x, lane = tid & 31;
  const int W = a.W, K = a.plan.K, H = a.plan.H, e_local = a.plan.e_local, cap = a.plan.cap;
  const int row_bytes = a.plan.row_bytes, scale_bytes = a.plan.scale_bytes;
  char* s_row = s_mem;
  float* s_scale = reinterpret_cast<float*>(s_mem + row_bytes);
  const bool packed = kFp8 && a.ue8m0 == 2;
  const char* my_base = a.peers.base[a.me];
  char* my_base_w = a.peers.base[a.me];
  auto load_id = [&](int t, int k) -> long long {
    return a.ids_i64 ? reinterpret_cast<const long long*>(a.ids)[(size_t)t * K + k]
                     : (long long)reinterpret_cast<const int*>(a.ids)[(size_t)t * K + k];
  };
  // fused: block 0 is the count CTA (dispatched first, so the token CTAs that wait for its flags can
  // never keep it off the GPU), blocks 1.. are token CTAs
  const int ntok_ctas = a.fused ? (int)gridDim.x - 1 : (int)gridDim.x;
  const int tok_cta = a.fused ? (int)blockIdx.x - 1 : (int)blockIdx.x;
  const bool trace = a.trace != nullptr && tid == 0;
  auto stamp = [&](int i) { if (trace) a.trace[i] = globaltimer_ns(); };
  bool st_q = false, st_i = false;
  if (blockIdx.x == 0) stamp(0);
  if (tok_cta == 0) stamp(2);

  if (a.fused && blockIdx.x == 0) {
    // ---------------- count role (one CTA): per-expert histogram -> every peer, count flags
    int* s_cnt = reinterpret_cast<int*>(s_mem);
    for (int e = tid; e < a.E; e += blockDim.x) s_cnt[e] = 0;
    __syncthreads();
    for (int i = tid; i < a.T * K; i += blockDim.x) {
      const long long v = load_id(i / K, i % K);
      if (v >= 0 && v < a.E) atomicAdd(&s_cnt[(int)v], 1);
    }
    __syncthreads();
    const unsigned long long tag = *reinterpret_cast<volatile unsigned long long*>(a.epoch_cnt_ptr) + 1;
    for (int i = tid; i < a.E * W; i += blockDim.x) {  // every (expert, peer) word: one 8-byte store each, no fence, no flag
      const int e = i / W, d = i % W;
      st_relaxed_sys_u64(ll_count_word(a.peers.base[d], a.slot, a.me, e), ll_count_pack(tag, s_cnt[e]));
    }
    __syncthreads();
    if (tid == 0) *reinterpret_cast<volatile unsigned long long*>(a.epoch_cnt_ptr) = tag;
    stamp(1);
  } else {
    // count flags of this call carry (data epoch before this kernel) + 1: both advance once per dispatch,
    // and the data epoch cannot move before every CTA of this kernel is past its waits
    const unsigned want_cnt = (unsigned)(*reinterpret_cast<const volatile unsigned long long*>(a.epoch_data_ptr) + 1);
    bool counts_ready = false;
    for (int t = tok_cta; t < a.T || !counts_ready; t += ntok_ctas) {
      const bool have_token = t < a.T;
      if (have_token) {
        if (tid < K) bulk_wait_read0();  // issuing threads: previous copies have read the stage
        __syncthreads();
        if (a.fused) {
          if (tid < K) {
            s_pos[tid] = 0;
            s_myid[tid] = load_id(t, tid);
          }
          __syncthreads();
          // how many earlier tokens of this rank chose each of my experts (== the COUNT kernel's pos)
          for (int j0 = 0; j0 < t; j0 += blockDim.x) {
            const int j = j0 + tid;
            unsigned hit = 0;  // bit k: token j also routes to my k-th expert
            if (j < t) {
              for (int kk = 0; kk < K; ++kk) {
                const long long v = load_id(j, kk);
                if (v < 0) continue;
                for (int k = 0; k < K; ++k)
                  if (v == s_myid[k]) hit |= 1u << k;
              }
            }
            for (int k = 0; k < K; ++k) {
              const unsigned ball = __ballot_sync(0xffffffffu, (hit >> k) & 1u);
              if (lane == 0 && ball) atomicAdd(&s_pos[k], __popc(ball));
            }
          }
        }
        const int4* src = reinterpret_cast<const int4*>(a.x + (size_t)t * H);
        const int nvec = H / 8;
        if (kFp8) {
          for (int v0 = 0; v0 < nvec; v0 += 256) {
            const int v = v0 + tid;
            int4 raw = v < nvec ? ld_nc_v4(src + v) : make_int4(0, 0, 0, 0);
            float f[8];
            unpack8(raw, f);
            const float s = group_scale(group_amax16(absmax8(f)), a.ue8m0 != 0);
            if (v < nvec) {
              *reinterpret_cast<uint2*>(s_row + (size_t)v * 8) = quant8(f, s);
              if ((lane & 15) == 0) s_scale[v >> 4] = s;
            }
          }
        } else {
          for (int v = tid; v < nvec; v += 256) reinterpret_cast<int4*>(s_row)[v] = ld_nc_v4(src + v);
        }
        fence_proxy_async_smem();
      }
      if (tok_cta == 0 && !st_q) { stamp(3); st_q = true; }
      if (!counts_ready) {
        counts_ready = true;
        if (tok_cta == 0) {  // receive-side bookkeeping of this rank: counts and per-source ranges (waits word by word)
          for (int el = tid; el < e_local; el += blockDim.x) {
            int p = 0;
            for (int s = 0; s < W; ++s) {
              const int n = wait_tagged_word(ll_count_word(my_base_w, a.slot, s, a.me * e_local + el), want_cnt, kFlagLlCount, a.me, s);
              a.sc->layout_range[(el * W + s) * 2] = p;
              a.sc->layout_range[(el * W + s) * 2 + 1] = n;
              p += n;
            }
            a.sc->recv_count[el] = p;
          }
        }
      }
      __syncthreads();
      if (have_token && tid < K) {
        const long long id = load_id(t, tid);
        int slot_idx = -1;
        if (id >= 0 && id < a.E) {
          const int e = (int)id, d = e / e_local, el = e % e_local;
          int base = 0;  // rows of lower ranks come first: wait for exactly the count words this entry needs
          for (int s = 0; s < a.me; ++s) base += wait_tagged_word(ll_count_word(my_base_w, a.slot, s, e), want_cnt, kFlagLlCount, a.me, s);
          slot_idx = base + (a.fused ? s_pos[tid] : a.sc->pos[t * K + tid]);
          if (slot_idx < cap) {
            char* dst = a.peers.base[d] + a.slot_off;
            const size_t r = (size_t)el * cap + slot_idx;
            bulk_s2g(dst + a.plan.off_x + r * row_bytes, s_row, row_bytes);
            if (kFp8 && !packed && (scale_bytes & 15) == 0) bulk_s2g(dst + a.plan.off_scales + r * scale_bytes, s_scale, scale_bytes);
            bulk_commit();
            if (kFp8 && !packed && (scale_bytes & 15)) {
              float* ds = reinterpret_cast<float*>(dst + a.plan.off_scales + r * scale_bytes);
              for (int g = 0; g < scale_bytes / 4; ++g) ds[g] = s_scale[g];
            }
            *reinterpret_cast<int32_t*>(dst + a.plan.off_src + r * 4) = t * K + tid;
          } else {
            printf("b200ep: rank %d ll_dispatch overflow: expert %d slot %d >= capacity %d\n", a.me, e, slot_idx, cap);
            __trap();
          }
        }
        a.sc->send_slot[t * K + tid] = slot_idx;
        if (packed) {
          s_slot[tid] = slot_idx >= 0 ? ((int)(id % e_local) * a.plan.g4) * a.plan.cap4 + slot_idx : -1;
          s_dst[tid] = slot_idx >= 0 ? (int)(id / e_local) : 0;
        }
      }
      if (tok_cta == 0 && !st_i) { stamp(4); st_i = true; }
      if (kFp8 && packed && have_token) {
        // Packed UE8M0 scales (what DeepGEMM reads on sm_100, vLLM fp8_utils.py:643-714): the scales are
        // powers of two, so only the biased exponent byte travels, four groups per int32, written straight
        // into the owner's [e_local][H/512][slots] array (slot index fastest: "MN-major").
        __syncthreads();
        const int g4 = a.plan.g4, ng = scale_bytes / 4;
        for (int i = tid; i < K * g4; i += blockDim.x) {
          const int k = i / g4, j = i % g4;
          if (s_slot[k] < 0) continue;
          uint32_t wd = 0;
#pragma unroll
          for (int b = 0; b < 4; ++b)
            if (4 * j + b < ng) wd |= ((__float_as_uint(s_scale[4 * j + b]) >> 23) & 0xffu) << (8 * b);
          reinterpret_cast<uint32_t*>(a.peers.base[s_dst[k]] + a.slot_off + a.plan.off_scales)[(size_t)s_slot[k] + (size_t)j * a.plan.cap4] = wd;
        }
      }
    }
    if (tid < K) bulk_wait_all();
    if (tok_cta == 0) stamp(5);
  }
  // every CTA (count role included): completion counter, last one raises the data flags
  __syncthreads();
  if (tid == 0) {
    const unsigned prev = atom_add_acq_rel_gpu(a.done, 1u);
    if (prev == gridDim.x - 1) {
      *a.done = 0;
      const uint64_t epoch = *reinterpret_cast<volatile unsigned long long*>(a.epoch_data_ptr) + 1;
      stamp(6);
      fence_acq_rel_sys();
      for (int d = 0; d < W; ++d)
        st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(a.peers.base[d] + flag_off(kFlagLlData, a.slot, a.me)), epoch);
      *reinterpret_cast<volatile unsigned long long*>(a.epoch_data_ptr) = epoch;
      stamp(7);
      if (a.wait_data)
        for (int s = 0; s < W; ++s)
          wait_flag_sys(reinterpret_cast<const uint64_t*>(my_base + flag_off(kFlagLlData, a.slot, s)), epoch, kFlagLlData,
                        a.me, s);
      stamp(8);
    }
  }
}

struct LlCombineArgs {
  Peers peers;
  LlPlan plan;
  const void* y;  // bf16 [e_local, cap, H]
  void* out;      // bf16 [T, H]
  const void* ids;
  const float* weights;
  const LlScratch* sc;
  unsigned* done;
  unsigned* work;  // [3] push queue head, completed blocks, warps finished
  size_t slot_off;
  unsigned long long* epoch_ptr;
  int T, W, me, slot, E, ids_i64, fused;
  unsigned long long* trace;  // B200EP_TRACE=1: %globaltimer stamps go here (measurement aid), else null
};

template <typename Wait>
__device__ __forceinline__ void ll_reduce_items(const LlCombineArgs& a, Wait wait);

// a.fused == 1 (all phases in one call): after its pushes every CTA waits for the peers' flags and
// takes part in the reduction, so the whole combine is one launch.
__global__ void __launch_bounds__(512) ll_combine_send_kernel(const __grid_constant__ LlCombineArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  // the epoch this call publishes; the stored epoch only advances after every warp of the grid has read it
  const unsigned long long want = *reinterpret_cast<const volatile unsigned long long*>(a.epoch_ptr) + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const bool trace = a.trace != nullptr && threadIdx.x == 0;
  auto stamp = [&](int i) { if (trace) a.trace[i] = globaltimer_ns(); };
  if (blockIdx.x == 0) stamp(10);
  const int W = a.W, cap = a.plan.cap, e_local = a.plan.e_local;
  const int row_bytes = 2 * a.plan.H;
  char* s_row = s_mem + (size_t)warp * row_bytes;
  const int32_t* src_info = reinterpret_cast<const int32_t*>(a.peers.base[a.me] + a.slot_off + a.plan.off_src);
  // Occupied rows only: exclusive prefix of recv_count over the local experts (one warp, shared memory).
  __shared__ int s_pref[kMaxLocalExperts + 1];
  __shared__ __align__(8) uint64_t s_bar[16];  // one mbarrier per warp (TMA loads of its stage)
  unsigned uses = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) mbar_init(&s_bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    const int per = (e_local + 31) / 32;
    int sum = 0;
    for (int i = lane * per; i < min(e_local, (lane + 1) * per); ++i) sum += a.sc->recv_count[i];
    int incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum;
    for (int i = lane * per; i < min(e_local, (lane + 1) * per); ++i) {
      s_pref[i] = run;
      run += a.sc->recv_count[i];
    }
    if (lane == 31) s_pref[e_local] = incl;
  }
  __syncthreads();
  const int total = s_pref[e_local];
  // Rows are handed out from a queue (a.work[0]), eight per CTA per grab (one atomic per CTA, not per
  // warp), so whichever CTAs are resident do all the pushing; completion is counted in rows
  // (a.work[1]), not in CTAs.
  constexpr int kGrab = 8;
  __shared__ int s_base, s_cta_rows;
  if (threadIdx.x == 0) s_cta_rows = 0;
  const int nblocks = total;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_base = (int)atomicAdd(&a.work[0], (unsigned)kGrab);
    __syncthreads();
    const int base = s_base;
    if (base >= total) break;
    if (threadIdx.x == 0) s_cta_rows += min(total, base + kGrab) - base;
    // every grabbed row is pushed: a CTA with fewer than kGrab warps (rows wider than 25 KB) loops
    for (int r = base + warp; r < min(total, base + kGrab); r += nwarp) {
      {
    int lo = 0, hi = e_local - 1;  // largest el with s_pref[el] <= r
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_pref[mid] <= r) lo = mid; else hi = mid - 1;
    }
    const int el = lo, sl = r - s_pref[lo];
    const int idx = el * cap + sl;
    int s = 0;
    while (s + 1 < W && sl >= a.sc->layout_range[(el * W + s + 1) * 2]) ++s;
    if (s == a.me) continue;  // the reduce reads local rows straight from dev_y
    // the row never touches registers: TMA load (global -> this warp's stage), then TMA store (stage -> peer)
    if (lane == 0) {
      const int flat = __ldcg(src_info + idx);  // t*K + k on the source rank
      bulk_wait_read0();                        // the previous store has finished reading the stage
      mbar_expect_tx(&s_bar[warp], row_bytes);
      bulk_g2s(s_row, reinterpret_cast<const char*>(a.y) + (size_t)idx * row_bytes, row_bytes, &s_bar[warp]);
      mbar_wait(&s_bar[warp], uses & 1u);
      ++uses;
      bulk_s2g(a.peers.base[s] + a.slot_off + a.plan.off_stage + (size_t)flat * row_bytes, s_row, row_bytes);
      bulk_commit();
    }
      }
    }
  }
  if (blockIdx.x == 0) stamp(11);
  bulk_wait_all();
  if (blockIdx.x == 0) stamp(12);
  // all rows this CTA took have landed: add them to the completion count; the CTA that completes the
  // last row publishes (one system fence); the last CTA of the grid rewinds the queue for the next call
  __syncthreads();
  if (threadIdx.x == 0) {
    const int done_blocks = s_cta_rows;
    const unsigned prev = done_blocks ? atom_add_acq_rel_gpu(&a.work[1], (unsigned)done_blocks) : 0u;
    const bool last_block = done_blocks && prev + done_blocks == (unsigned)nblocks;
    const unsigned fin = atom_add_acq_rel_gpu(&a.work[2], 1u);  // CTAs that left the push loop
    if (last_block || (nblocks == 0 && fin == 0)) {
      fence_acq_rel_sys();
      for (int d = 0; d < W; ++d)
        st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(a.peers.base[d] + flag_off(kFlagLlComb, a.slot, a.me)), want);
      stamp(15);
    }
    if (fin == gridDim.x - 1) {
      // every CTA of the grid is past the queue (so each has read the old epoch): rewind the queue
      // and only now advance the epoch -- a CTA that is scheduled late must still compute `want` right
      a.work[0] = 0;
      a.work[1] = 0;
      a.work[2] = 0;
      *reinterpret_cast<volatile unsigned long long*>(a.epoch_ptr) = want;
    }
  }
  if (a.fused) {
    ll_reduce_items(a, [&]() {
      if ((int)threadIdx.x < W)
        wait_flag_sys(reinterpret_cast<const uint64_t*>(a.peers.base[a.me] + flag_off(kFlagLlComb, a.slot, threadIdx.x)), want,
                      kFlagLlComb, a.me, threadIdx.x);
      __syncthreads();
      if (blockIdx.x == 0) stamp(13);
    });
    if (blockIdx.x == 0) stamp(14);
  }
}

// out[t] = bf16( sum_k w[t,k] * row(t,k) ): fp32 multiply, fp32 add, k ascending.
// A warp owns (token, 512-byte column block) so that 128 tokens still fill the GPU.  Where the rows of an item are
// (ids, weights, slots: three dependent loads) is resolved BEFORE `wait` -- the wait for the peers' flags -- for a
// warp's first item, so that what follows the last flag is one batch of row loads and the stores.
template <typename Wait>
__device__ __forceinline__ void ll_reduce_items(const LlCombineArgs& a, Wait wait) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int K = a.plan.K, cap = a.plan.cap, e_local = a.plan.e_local;
  const int row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  const int nblk = (nvec + 31) / 32;
  const char* stage = a.peers.base[a.me] + a.slot_off + a.plan.off_stage;
  const int items = a.T * nblk;
  const char* rowp = nullptr;
  float w = 0.f;
  auto resolve = [&](int it) {
    const int t = it / nblk;
    rowp = nullptr;
    w = 0.f;
    if (lane < K) {
      const long long id = a.ids_i64 ? reinterpret_cast<const long long*>(a.ids)[(size_t)t * K + lane]
                                     : (long long)reinterpret_cast<const int*>(a.ids)[(size_t)t * K + lane];
      if (id >= 0 && id < a.E) {
        const int e = (int)id;
        w = a.weights[(size_t)t * K + lane];
        if (e / e_local == a.me)
          rowp = reinterpret_cast<const char*>(a.y) + ((size_t)(e % e_local) * cap + a.sc->send_slot[t * K + lane]) * row_bytes;
        else
          rowp = stage + ((size_t)t * K + lane) * row_bytes;
      }
    }
  };
  int it = blockIdx.x * nwarp + warp;
  if (it < items) resolve(it);
  wait();
  while (it < items) {
    const int t = it / nblk, v = (it % nblk) * 32 + lane;
    const unsigned kmask = __ballot_sync(0xffffffffu, rowp != nullptr);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k0 = 0; k0 < K; k0 += 8) {
      int4 raw[8];
      float wk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + j;
        const char* rp = reinterpret_cast<const char*>(__shfl_sync(0xffffffffu, (unsigned long long)rowp, k & 31));
        wk[j] = __shfl_sync(0xffffffffu, w, k & 31);
        if (k < K && ((kmask >> k) & 1u) && v < nvec) raw[j] = ld_cg_v4(rp + (size_t)v * 16);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + j;
        if (k < K && ((kmask >> k) & 1u)) {
          float f[8];
          unpack8(raw[j], f);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], __fmul_rn(wk[j], f[i]));
        }
      }
    }
    if (v < nvec) {
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        __nv_bfloat162 p = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
        o[i] = *reinterpret_cast<uint32_t*>(&p);
      }
      st_na_v4(reinterpret_cast<char*>(a.out) + (size_t)t * row_bytes + (size_t)v * 16,
               make_int4((int)o[0], (int)o[1], (int)o[2], (int)o[3]));
    }
    it += gridDim.x * nwarp;
    if (it < items) resolve(it);
  }
}

__global__ void __launch_bounds__(256) ll_combine_recv_kernel(const __grid_constant__ LlCombineArgs a) {
  ll_reduce_items(a, [&]() { cta_wait_all_sources(a.peers.base[a.me], a.W, kFlagLlComb, a.slot, a.me, a.epoch_ptr); });
}

__global__ void ll_wait_sources_kernel(const char* my_base, int W, int cat, int slot, int me,
                                       const unsigned long long* epoch_ptr) {
  cta_wait_all_sources(my_base, W, cat, slot, me, epoch_ptr);
}

}  // namespace b200ep

using namespace b200ep;

static int ll_check(b200ep_ctx* c, int slot, const char* who) {
  if (!c || slot < 0 || slot >= kNumSlots) {
    set_error("%s: bad context or slot", who);
    return B200EP_ERR_ARG;
  }
  if (!c->connected) {
    set_error("%s: peers not connected", who);
    return B200EP_ERR_STATE;
  }
  return B200EP_OK;
}

static int ll_grid(const b200ep_ctx* c, int want) {
  int sms = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
  if (want > sms) want = sms;
  return want < 1 ? 1 : want;
}

extern "C" int b200ep_ll_dispatch(b200ep_ctx* c, int slot, const void* dev_x_bf16, const void* dev_topk_ids,
                                  int ids_i64, int T, int H, int K, int E, int max_tokens_per_rank, int use_fp8,
                                  int ue8m0, int phases, void* stream) {
  int rc = ll_check(c, slot, "ll_dispatch");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  LlPlan plan;
  rc = make_ll_plan(c, H, K, E, max_tokens_per_rank, use_fp8, &plan);
  if (rc) return rc;
  if (T < 0 || T > max_tokens_per_rank || T > kLlMaxTokens || E > kLlMaxExperts) {
    set_error("ll_dispatch: T=%d must be <= max_tokens_per_rank=%d (<= %d), E=%d <= %d", T, max_tokens_per_rank,
              kLlMaxTokens, E, kLlMaxExperts);
    return B200EP_ERR_ARG;
  }
  if (T > 0 && (!dev_x_bf16 || !dev_topk_ids)) {
    set_error("ll_dispatch: null pointer");
    return B200EP_ERR_ARG;
  }
  LlScratch* sc = reinterpret_cast<LlScratch*>(c->ll_local) + slot;
  // COUNT and SEND requested together (the normal case): one launch does both.  Phase-by-phase callers
  // (single-GPU emulation of several ranks) get the separate count kernel.
  const bool fused = (phases & B200EP_LL_PHASE_COUNT) && (phases & B200EP_LL_PHASE_SEND);
  if (phases & B200EP_LL_PHASE_COUNT) c->ll_plan[slot] = plan;
  if ((phases & B200EP_LL_PHASE_COUNT) && !fused) {
    unsigned long long* ep = &c->local->epoch[kFlagLlCount][slot];
    const size_t smem = (size_t)E * ((T + 31) / 32) * 4;
    if (smem > 160 * 1024) {
      set_error("ll_dispatch: E=%d x T=%d needs %zu bytes of shared memory for the routing masks", E, T, smem);
      return B200EP_ERR_ARG;
    }
    if (ids_i64) {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ll_count_kernel<int64_t>, c->device, 160 * 1024));
      ll_count_kernel<int64_t><<<1, 1024, smem, st>>>(c->peers, c->world, c->rank, slot, (const int64_t*)dev_topk_ids,
                                                     T, K, E, sc, ep);
    } else {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ll_count_kernel<int32_t>, c->device, 160 * 1024));
      ll_count_kernel<int32_t><<<1, 1024, smem, st>>>(c->peers, c->world, c->rank, slot, (const int32_t*)dev_topk_ids,
                                                     T, K, E, sc, ep);
    }
    B200EP_CUDA(cudaGetLastError());
  }
  const bool wait_in_kernel = fused && (phases & B200EP_LL_PHASE_RECV);
  if (phases & B200EP_LL_PHASE_SEND) {
    LlDispatchArgs a;
    a.peers = c->peers;
    a.plan = plan;
    a.x = (const __nv_bfloat16*)dev_x_bf16;
    a.ids = dev_topk_ids;
    a.sc = sc;
    a.done = &c->local->done[4 + slot];
    a.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
    a.epoch_cnt_ptr = &c->local->epoch[kFlagLlCount][slot];
    a.epoch_data_ptr = &c->local->epoch[kFlagLlData][slot];
    a.T = T;
    a.W = c->world;
    a.me = c->rank;
    a.slot = slot;
    a.E = E;
    a.ue8m0 = ue8m0;
    a.ids_i64 = ids_i64;
    a.fused = fused ? 1 : 0;
    a.wait_data = wait_in_kernel ? 1 : 0;
    static const int trace_env = getenv("B200EP_TRACE") ? atoi(getenv("B200EP_TRACE")) : 0;
    a.trace = trace_env ? c->local->trace : nullptr;
    size_t smem = (size_t)(plan.row_bytes + plan.scale_bytes + 127) / 128 * 128;
    if (fused && smem < (size_t)E * 4) smem = (size_t)E * 4;  // the count CTA's histogram
    const int grid = ll_grid(c, T) + (fused ? 1 : 0);
    if (use_fp8) {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ll_dispatch_send_kernel<true>, c->device, (int)smem));
      ll_dispatch_send_kernel<true><<<grid, 256, smem, st>>>(a);
    } else {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ll_dispatch_send_kernel<false>, c->device, (int)smem));
      ll_dispatch_send_kernel<false><<<grid, 256, smem, st>>>(a);
    }
    B200EP_CUDA(cudaGetLastError());
  }
  if ((phases & B200EP_LL_PHASE_RECV) && !wait_in_kernel) {
    ll_wait_sources_kernel<<<1, 32, 0, st>>>(c->sym, c->world, kFlagLlData, slot, c->rank, &c->local->epoch[kFlagLlData][slot]);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}

extern "C" int b200ep_ll_views(b200ep_ctx* c, int slot, b200ep_ll_view* out) {
  if (!c || !out || slot < 0 || slot >= kNumSlots) {
    set_error("ll_views: bad argument");
    return B200EP_ERR_ARG;
  }
  const LlPlan& p = c->ll_plan[slot];
  if (p.H == 0) {
    set_error("ll_views: no low-latency dispatch has run on slot %d", slot);
    return B200EP_ERR_STATE;
  }
  char* b = slot_base(c, c->sym, slot);
  LlScratch* sc = reinterpret_cast<LlScratch*>(c->ll_local) + slot;
  out->packed_recv_x = b + p.off_x;
  out->packed_recv_scales = p.scale_bytes ? reinterpret_cast<float*>(b + p.off_scales) : nullptr;
  out->recv_count = sc->recv_count;
  out->src_info = reinterpret_cast<int32_t*>(b + p.off_src);
  out->layout_range = sc->layout_range;
  return B200EP_OK;
}

extern "C" int b200ep_ll_combine(b200ep_ctx* c, int slot, const void* dev_y, const void* dev_topk_ids, int ids_i64,
                                 const float* dev_topk_weights, void* dev_out, int T, int H, int K, int E,
                                 int max_tokens_per_rank, int phases, void* stream) {
  int rc = ll_check(c, slot, "ll_combine");
  if (rc) return rc;
  const LlPlan& plan = c->ll_plan[slot];
  if (plan.H != H || plan.K != K || plan.max_tok != max_tokens_per_rank || plan.e_local != E / c->world) {
    set_error("ll_combine: no matching low-latency dispatch on slot %d", slot);
    return B200EP_ERR_STATE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  LlCombineArgs a;
  a.peers = c->peers;
  a.plan = plan;
  a.y = dev_y;
  a.out = dev_out;
  a.ids = dev_topk_ids;
  a.weights = dev_topk_weights;
  a.sc = reinterpret_cast<LlScratch*>(c->ll_local) + slot;
  a.done = &c->local->done[6 + slot];
  a.work = &c->local->done[8 + slot * 4];
  a.slot_off = kCtrlBytes + (size_t)slot * c->slot_bytes;
  a.T = T;
  a.W = c->world;
  a.me = c->rank;
  a.slot = slot;
  a.E = E;
  a.ids_i64 = ids_i64;
  a.epoch_ptr = &c->local->epoch[kFlagLlComb][slot];
  a.fused = (phases == B200EP_PHASE_ALL) ? 1 : 0;
  static const int trace_env = getenv("B200EP_TRACE") ? atoi(getenv("B200EP_TRACE")) : 0;
  a.trace = trace_env ? c->local->trace : nullptr;
  const int row_bytes = 2 * H;
  if (phases & B200EP_PHASE_SEND) {
    if (!dev_y) {
      set_error("ll_combine: null dev_y");
      return B200EP_ERR_ARG;
    }
    int warps = 200 * 1024 / row_bytes;
    if (warps > 16) warps = 16;
    if (warps < 1) warps = 1;
    const size_t smem = (size_t)warps * row_bytes;
    const int grid = ll_grid(c, c->num_sms);
    B200EP_CUDA(ensure_dynamic_smem((const void*)ll_combine_send_kernel, c->device, (int)smem));
    ll_combine_send_kernel<<<grid, warps * 32, smem, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
  }
  if ((phases & B200EP_PHASE_RECV) && !a.fused) {
    if (T > 0 && (!dev_out || !dev_topk_ids || !dev_topk_weights)) {
      set_error("ll_combine: null pointer");
      return B200EP_ERR_ARG;
    }
    const int nblk = (row_bytes / 16 + 31) / 32;
    int grid = (T * nblk + 7) / 8;
    grid = ll_grid(c, grid) * (grid > c->num_sms ? 2 : 1);
    ll_combine_recv_kernel<<<grid, 256, 0, st>>>(a);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}

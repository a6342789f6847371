// Top-k routing layout kernel (north-star kernel 1).
//
// Replaces deep_ep.Buffer.get_dispatch_layout as vLLM calls it
// (vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ht.py:121-133):
// per-rank and per-expert token counts and the token x rank membership matrix,
// plus -- new here -- the exclusive prefix offset of every (token, rank) pair so
// that the dispatch can place rows deterministically.  CPU restatement:
// oracle/layout.py.
//
// One pass over topk_ids.  A block owns 256 consecutive tokens (one per thread):
//   * rank membership bits -> __ballot_sync / __popc warp scan -> block scan;
//   * block totals are published with a release store tagged by the call epoch;
//     block b sums the totals of blocks < b (all resident; a ticket makes the
//     logical block id follow scheduling order, so the wait cannot deadlock);
//   * expert histogram in shared memory, flushed with integer atomics (counts
//     only -- the result does not depend on arrival order) into private scratch;
//     the last block to finish writes the totals out, rewinds the scratch and,
//     on the production path, publishes the counts to the peers (notify.cuh).
// HBM traffic: T*K ids in, T*W*(1+4) bytes out; latency bound at prefill sizes.
#include "notify.cuh"

namespace b200ep {

// Spin on a published (tag << 32 | count) word until its tag matches; traps instead of hanging.
__device__ __forceinline__ uint64_t layout_wait_pub(const unsigned long long* p, unsigned tag) {
  uint64_t v = ld_acquire_gpu_u64(reinterpret_cast<const uint64_t*>(p));
  if ((unsigned)(v >> 32) == tag) return v;
  const long long t0 = clock64();
  do {
    if (clock64() - t0 > 20000000000LL) {  // ~10 s: the publishing block always runs (see ticket), so this is a bug or a dead GPU
      printf("b200ep: routing layout block %d timed out waiting for an earlier block's totals (tag %u, have %llu)\n",
             (int)blockIdx.x, tag, (unsigned long long)v);
      __trap();
    }
    v = ld_acquire_gpu_u64(reinterpret_cast<const uint64_t*>(p));
  } while ((unsigned)(v >> 32) != tag);
  return v;
}

// All per-call state (block ticket, finished-block count, call epoch, expert histogram) lives in the
// context's private device scratch and is rewound by the last block to finish: nothing is passed in
// from host counters and nothing is cleared by a memset, so the launch can be replayed.  Calls on ONE
// context must not overlap on the GPU (they share that scratch): issue them on one stream.
// The last block also opens the count exchange of the dispatch that follows (tail.mode, notify.cuh).
template <typename IdT, bool kSingleTile>
__global__ void __launch_bounds__(256)
    layout_kernel(const IdT* __restrict__ ids, int T, int K, int E, int W, int e_local,
                  int32_t* __restrict__ tok_per_rank, int32_t* __restrict__ tok_per_expert,
                  uint8_t* __restrict__ in_rank, int32_t* __restrict__ send_off, LocalScratch* ls,
                  int tiles_per_block, const __grid_constant__ NotifyTail tail) {
  extern __shared__ int s_hist[];  // [E]
  __shared__ int s_warp_tot[8][kMaxRanks];
  __shared__ int s_tile_base[kMaxRanks];
  __shared__ int s_base[kMaxRanks];
  __shared__ unsigned s_bid, s_tag, s_last;
  __shared__ unsigned s_mask[256];

  pdl_launch_dependents();  // the dispatch kernel may start loading its rows while this kernel runs
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    s_tag = (unsigned)(*reinterpret_cast<volatile unsigned long long*>(&ls->layout_epoch) + 1);
    s_bid = (unsigned)atomicAdd(&ls->layout_ticket, 1ull);  // logical block id follows scheduling order: look-back cannot deadlock
  }
  for (int e = tid; e < E; e += 256) s_hist[e] = 0;
  if (tid < kMaxRanks) {
    s_tile_base[tid] = 0;
    s_base[tid] = 0;
  }
  __syncthreads();
  const int bid = (int)s_bid;
  const unsigned tag = s_tag;

  int loc[kMaxRanks];       // single-tile path: block-local offsets kept in registers
  unsigned my_mask = 0;
  int my_t = -1;

  for (int tile = 0; tile < tiles_per_block; ++tile) {
    const int t_first = (bid * tiles_per_block + tile) * 256;
    const int t = t_first + tid;
    // the tile's ids as one coalesced stream of (token, k) entries, several loads in flight per thread;
    // a token's rank mask is OR-ed together in shared memory (order does not matter)
    s_mask[tid] = 0u;
    __syncthreads();
    const int nent = max(0, min(256, T - t_first)) * K;
    const IdT* tile_ids = ids + (size_t)t_first * K;
#pragma unroll 4
    for (int i = tid; i < nent; i += 256) {
      const long long id = (long long)tile_ids[i];
      if (id >= 0 && id < E) {
        atomicAdd(&s_hist[(int)id], 1);
        atomicOr(&s_mask[i / K], 1u << ((int)id / e_local));
      }
    }
    __syncthreads();
    const unsigned mask = s_mask[tid];
    int wpre[kMaxRanks];
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d) {
      if (d < W) {
        unsigned ball = __ballot_sync(0xffffffffu, (mask >> d) & 1u);
        wpre[d] = __popc(ball & ((1u << lane) - 1u));
        if (lane == 0) s_warp_tot[warp][d] = __popc(ball);
      }
    }
    __syncthreads();
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d) {
      if (d < W) {
        int p = s_tile_base[d];
        for (int w = 0; w < warp; ++w) p += s_warp_tot[w][d];
        loc[d] = p + wpre[d];
      }
    }
    if (kSingleTile) {
      my_mask = mask;
      my_t = t;
    } else if (t < T) {
      for (int d = 0; d < W; ++d) {
        in_rank[(size_t)t * W + d] = (mask >> d) & 1u;
        send_off[(size_t)t * W + d] = ((mask >> d) & 1u) ? loc[d] : -1;  // block base added below
      }
    }
    __syncthreads();
    if (tid < W) {
      int s = 0;
      for (int w = 0; w < 8; ++w) s += s_warp_tot[w][tid];
      s_tile_base[tid] += s;
    }
    __syncthreads();
  }

  // publish this block's totals, then sum the totals of all logically earlier blocks
  if (tid < W)
    st_release_gpu_u64(reinterpret_cast<uint64_t*>(&ls->layout_pub[bid][tid]),
                       ((uint64_t)tag << 32) | (uint32_t)s_tile_base[tid]);
  int part[kMaxRanks];
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) part[d] = 0;
  for (int b = tid; b < bid; b += 256) {
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d)
      if (d < W) part[d] += (int)(uint32_t)layout_wait_pub(&ls->layout_pub[b][d], tag);
  }
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d)
    if (d < W && part[d]) atomicAdd(&s_base[d], part[d]);
  __syncthreads();

  if (kSingleTile) {
    if (my_t >= 0 && my_t < T) {
#pragma unroll
      for (int d = 0; d < kMaxRanks; ++d) {
        if (d < W) {
          in_rank[(size_t)my_t * W + d] = (my_mask >> d) & 1u;
          send_off[(size_t)my_t * W + d] = ((my_mask >> d) & 1u) ? s_base[d] + loc[d] : -1;
        }
      }
    }
  } else {
    for (int tile = 0; tile < tiles_per_block; ++tile) {
      const int t = (bid * tiles_per_block + tile) * 256 + tid;
      if (t < T)
        for (int d = 0; d < W; ++d) {
          int v = send_off[(size_t)t * W + d];
          if (v >= 0) send_off[(size_t)t * W + d] = v + s_base[d];
        }
    }
  }
  if (bid == (int)gridDim.x - 1 && tid < W) tok_per_rank[tid] = s_base[tid] + s_tile_base[tid];
  for (int e = tid; e < E; e += 256) {
    int h = s_hist[e];
    if (h) atomicAdd(&ls->layout_hist[e], h);
  }

  // ---- the last block to finish writes the expert totals, rewinds the scratch and opens the count exchange
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    s_last = atom_add_acq_rel_gpu(&ls->layout_done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int e = tid; e < E; e += 256) {
    const int h = __ldcg(&ls->layout_hist[e]);
    tok_per_expert[e] = h;
    s_hist[e] = h;
    if (h) ls->layout_hist[e] = 0;
  }
  if (tid < W) s_base[tid] = __ldcg(&tok_per_rank[tid]);
  if (tid == 0) {
    ls->layout_ticket = 0;
    ls->layout_done = 0;
    volatile unsigned long long* ep = &ls->layout_epoch;  // every block of this call has read it (they have all finished)
    *ep = *ep + 1;
  }
  __syncthreads();
  if (tail.mode == kTailLocal) {
    const unsigned long long epoch = *reinterpret_cast<volatile unsigned long long*>(tail.epoch_ptr) + 1;
    notify_finish_local(s_base, s_hist, tail.slot, e_local, T, tail.meta, tail.host, epoch);
    __syncthreads();
    if (tid == 0) *reinterpret_cast<volatile unsigned long long*>(tail.epoch_ptr) = epoch;
  } else if (tail.mode == kTailPublish) {
    notify_publish(tail.peers, W, tail.me, tail.slot, e_local, T, s_base, s_hist, tail.epoch_ptr);
  }
}

}  // namespace b200ep

using namespace b200ep;

namespace b200ep {
int launch_layout(b200ep_ctx* c, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                  int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert, uint8_t* dev_is_token_in_rank,
                  int32_t* dev_send_off, const NotifyTail& tail, cudaStream_t st) {
  if (!c || T < 0 || K <= 0 || K > kMaxTopk || E <= 0 || E % c->world != 0 || E / c->world > kMaxLocalExperts) {
    set_error("layout: bad shape T=%d K=%d E=%d world=%d", T, K, E, c ? c->world : -1);
    return B200EP_ERR_ARG;
  }
  if (!dev_tokens_per_rank || !dev_tokens_per_expert || (T > 0 && (!dev_topk_ids || !dev_is_token_in_rank || !dev_send_off))) {
    set_error("layout: null pointer");
    return B200EP_ERR_ARG;
  }
  B200EP_CUDA(cudaSetDevice(c->device));
  // T == 0 still launches one block: it writes the (zero) counts and, with a tail, takes part in the exchange
  const int ntiles = T > 0 ? (T + 255) / 256 : 1;
  const int tpb = (ntiles + kMaxLayoutBlocks - 1) / kMaxLayoutBlocks;
  const int grid = (ntiles + tpb - 1) / tpb;
  const size_t smem = sizeof(int) * E;
  const int el = E / c->world;
#define LAUNCH(IDT, SINGLE)                                                                                      \
  layout_kernel<IDT, SINGLE><<<grid, 256, smem, st>>>((const IDT*)dev_topk_ids, T, K, E, c->world, el,            \
                                                      dev_tokens_per_rank, dev_tokens_per_expert,                 \
                                                      dev_is_token_in_rank, dev_send_off, c->local, tpb, tail)
  if (ids_i64) {
    if (tpb == 1) LAUNCH(int64_t, true); else LAUNCH(int64_t, false);
  } else {
    if (tpb == 1) LAUNCH(int32_t, true); else LAUNCH(int32_t, false);
  }
#undef LAUNCH
  B200EP_CUDA(cudaGetLastError());
  return B200EP_OK;
}
}  // namespace b200ep

extern "C" int b200ep_layout(b200ep_ctx* c, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                             int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                             uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, void* stream) {
  NotifyTail tail{};
  tail.mode = kTailNone;
  return launch_layout(c, dev_topk_ids, ids_i64, T, K, E, dev_tokens_per_rank, dev_tokens_per_expert,
                       dev_is_token_in_rank, dev_send_off, tail, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the layout kernel (entry point b200ep_layout_noalloc).
//
// b200ep_layout needs a context, and b200ep_create allocates (the symmetric region alone is
// gigabytes).  Some callers cannot allocate device memory at the point where they need a routing
// layout: a vLLM worker that has already handed the rest of the GPU to its KV-cache pool (a late
// cudaMalloc there fails with out-of-memory or eats into the pool's headroom), and small tools or
// tests that only want counts and offsets for a batch of top-k ids and have no reason to set up a
// communication context.  For them the scratch this kernel needs -- the block ticket and the
// per-block totals that blocks publish to each other -- lives in the fixed array below.
//
// Limits and sharing, for users of this entry point:
//   * at most kNoAllocMaxTokens tokens per call (the array has one row per 256-token block);
//     larger batches are rejected, use b200ep_layout for those;
//   * every GPU that loads libb200ep.so carries the array (about 16 KB) until the process exits;
//   * all calls on a device share the one array and each call begins by clearing it on the
//     caller's stream, so two calls whose kernels overlap on the GPU (different streams, or
//     different host threads) race with each other: serialise them;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress)
//     before it launches anything; make that first call before any stream capture begins.
// No cudaMalloc / cudaFree / allocating library routine is called here.
namespace b200ep {

constexpr int kNoAllocMaxTokens = 65536;
constexpr int kNoAllocMaxBlocks = kNoAllocMaxTokens / 256;
constexpr int kNoAllocMaxDevices = 64;

struct LayoutStaticScratch {
  unsigned long long ticket;                             // next logical block id; cleared per call
  unsigned long long pad[31];
  unsigned long long pub[kNoAllocMaxBlocks][kMaxRanks];  // (1 << 32 | count) once published, 0 before
};

// In host code this name is only a stand-in for the symbol; the device address is obtained once
// per device with cudaGetSymbolAddress and handed to the kernel as an ordinary pointer.
static __device__ __align__(256) unsigned char g_layout_static_scratch[sizeof(LayoutStaticScratch)];

template <typename IdT>
__global__ void __launch_bounds__(256)
    layout_noalloc_kernel(const IdT* __restrict__ ids, int T, int K, int E, int W, int e_local,
                          int32_t* __restrict__ tok_per_rank, int32_t* __restrict__ tok_per_expert,
                          uint8_t* __restrict__ in_rank, int32_t* __restrict__ send_off, LayoutStaticScratch* sc) {
  extern __shared__ int s_hist[];  // [E]
  __shared__ int s_warp_tot[8][kMaxRanks];
  __shared__ int s_total[kMaxRanks];
  __shared__ int s_base[kMaxRanks];
  __shared__ unsigned s_bid;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_bid = (unsigned)atomicAdd(&sc->ticket, 1ull);
  for (int e = tid; e < E; e += 256) s_hist[e] = 0;
  if (tid < kMaxRanks) s_base[tid] = 0;
  __syncthreads();
  const int bid = (int)s_bid;
  const int t = bid * 256 + tid;

  unsigned mask = 0;
  if (t < T) {
    const IdT* row = ids + (size_t)t * K;
    for (int k = 0; k < K; ++k) {
      long long id = (long long)row[k];
      if (id >= 0 && id < E) {
        atomicAdd(&s_hist[(int)id], 1);
        mask |= 1u << ((int)id / e_local);
      }
    }
  }
  int loc[kMaxRanks];
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) {
    if (d < W) {
      unsigned ball = __ballot_sync(0xffffffffu, (mask >> d) & 1u);
      loc[d] = __popc(ball & ((1u << lane) - 1u));
      if (lane == 0) s_warp_tot[warp][d] = __popc(ball);
    }
  }
  __syncthreads();
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) {
    if (d < W)
      for (int w = 0; w < warp; ++w) loc[d] += s_warp_tot[w][d];
  }
  if (tid < W) {
    int s = 0;
    for (int w = 0; w < 8; ++w) s += s_warp_tot[w][tid];
    s_total[tid] = s;
    st_release_gpu_u64(reinterpret_cast<uint64_t*>(&sc->pub[bid][tid]), (1ull << 32) | (uint32_t)s);
  }
  int part[kMaxRanks];
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) part[d] = 0;
  for (int b = tid; b < bid; b += 256) {
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d) {
      if (d < W) {
        part[d] += (int)(uint32_t)layout_wait_pub(&sc->pub[b][d], 1u);
      }
    }
  }
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d)
    if (d < W && part[d]) atomicAdd(&s_base[d], part[d]);
  __syncthreads();
  if (t < T) {
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d) {
      if (d < W) {
        in_rank[(size_t)t * W + d] = (mask >> d) & 1u;
        send_off[(size_t)t * W + d] = ((mask >> d) & 1u) ? s_base[d] + loc[d] : -1;
      }
    }
  }
  if (bid == (int)gridDim.x - 1 && tid < W) tok_per_rank[tid] = s_base[tid] + s_total[tid];
  for (int e = tid; e < E; e += 256) {
    int h = s_hist[e];
    if (h) atomicAdd(&tok_per_expert[e], h);
  }
}

}  // namespace b200ep

#include <mutex>

extern "C" int b200ep_layout_noalloc(int world, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                                     int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                                     uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, void* stream) {
  if (world < 1 || world > kMaxRanks || T < 0 || K <= 0 || K > kMaxTopk || E <= 0 || E % world != 0 ||
      E / world > kMaxLocalExperts) {
    set_error("layout_noalloc: bad shape T=%d K=%d E=%d world=%d", T, K, E, world);
    return B200EP_ERR_ARG;
  }
  if (T > kNoAllocMaxTokens) {
    set_error("layout_noalloc: T=%d is more than the %d tokens the fixed scratch array of this entry point covers; "
              "use b200ep_layout (which owns per-context scratch) for larger batches",
              T, kNoAllocMaxTokens);
    return B200EP_ERR_CAPACITY;
  }
  if (!dev_tokens_per_rank || !dev_tokens_per_expert || (T > 0 && (!dev_topk_ids || !dev_is_token_in_rank || !dev_send_off))) {
    set_error("layout_noalloc: null pointer");
    return B200EP_ERR_ARG;
  }
  // Set-up, once per device, with that device current: resolve the array's device address.
  static std::mutex mu;
  static LayoutStaticScratch* scratch_of[kNoAllocMaxDevices] = {};
  int dev = -1;
  B200EP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    set_error("layout_noalloc: device ordinal %d out of range", dev);
    return B200EP_ERR_ARG;
  }
  LayoutStaticScratch* sc = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!scratch_of[dev]) {
      void* p = nullptr;
      B200EP_CUDA(cudaGetSymbolAddress(&p, g_layout_static_scratch));
      scratch_of[dev] = reinterpret_cast<LayoutStaticScratch*>(p);
    }
    sc = scratch_of[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaMemsetAsync(dev_tokens_per_expert, 0, sizeof(int32_t) * E, st));
  if (T == 0) {
    B200EP_CUDA(cudaMemsetAsync(dev_tokens_per_rank, 0, sizeof(int32_t) * world, st));
    return B200EP_OK;
  }
  const int grid = (T + 255) / 256;
  // the array still holds the previous call's ticket and totals: clear the part this call uses
  B200EP_CUDA(cudaMemsetAsync(sc, 0, offsetof(LayoutStaticScratch, pub) + sizeof(unsigned long long) * kMaxRanks * grid, st));
  const size_t smem = sizeof(int) * E;
  const int el = E / world;
  if (ids_i64)
    layout_noalloc_kernel<int64_t><<<grid, 256, smem, st>>>((const int64_t*)dev_topk_ids, T, K, E, world, el,
                                                          dev_tokens_per_rank, dev_tokens_per_expert,
                                                          dev_is_token_in_rank, dev_send_off, sc);
  else
    layout_noalloc_kernel<int32_t><<<grid, 256, smem, st>>>((const int32_t*)dev_topk_ids, T, K, E, world, el,
                                                          dev_tokens_per_rank, dev_tokens_per_expert,
                                                          dev_is_token_in_rank, dev_send_off, sc);
  B200EP_CUDA(cudaGetLastError());
  return B200EP_OK;
}

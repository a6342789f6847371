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
//     only -- the result does not depend on arrival order).
// HBM traffic: T*K ids in, T*W*(1+4) bytes out; latency bound at prefill sizes.
#include "common.cuh"

namespace b200ep {

template <typename IdT, bool kSingleTile>
__global__ void __launch_bounds__(256)
    layout_kernel(const IdT* __restrict__ ids, int T, int K, int E, int W, int e_local,
                  int32_t* __restrict__ tok_per_rank, int32_t* __restrict__ tok_per_expert,
                  uint8_t* __restrict__ in_rank, int32_t* __restrict__ send_off, LocalScratch* ls,
                  unsigned long long ticket_base, unsigned epoch, int tiles_per_block) {
  extern __shared__ int s_hist[];  // [E]
  __shared__ int s_warp_tot[8][kMaxRanks];
  __shared__ int s_tile_base[kMaxRanks];
  __shared__ int s_base[kMaxRanks];
  __shared__ unsigned s_bid;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_bid = (unsigned)(atomicAdd(&ls->layout_ticket, 1ull) - ticket_base);
  for (int e = tid; e < E; e += 256) s_hist[e] = 0;
  if (tid < kMaxRanks) {
    s_tile_base[tid] = 0;
    s_base[tid] = 0;
  }
  __syncthreads();
  const int bid = (int)s_bid;

  int loc[kMaxRanks];       // single-tile path: block-local offsets kept in registers
  unsigned my_mask = 0;
  int my_t = -1;

  for (int tile = 0; tile < tiles_per_block; ++tile) {
    const int t = (bid * tiles_per_block + tile) * 256 + tid;
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
                       ((uint64_t)epoch << 32) | (uint32_t)s_tile_base[tid]);
  int part[kMaxRanks];
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) part[d] = 0;
  for (int b = tid; b < bid; b += 256) {
#pragma unroll
    for (int d = 0; d < kMaxRanks; ++d) {
      if (d < W) {
        uint64_t v;
        do {
          v = ld_acquire_gpu_u64(reinterpret_cast<const uint64_t*>(&ls->layout_pub[b][d]));
        } while ((unsigned)(v >> 32) != epoch);
        part[d] += (int)(uint32_t)v;
      }
    }
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
    if (h) atomicAdd(&tok_per_expert[e], h);
  }
}

}  // namespace b200ep

using namespace b200ep;

extern "C" int b200ep_layout(b200ep_ctx* c, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                             int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                             uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, void* stream) {
  if (!c || T < 0 || K <= 0 || K > kMaxTopk || E <= 0 || E % c->world != 0 || E / c->world > kMaxLocalExperts) {
    set_error("layout: bad shape T=%d K=%d E=%d world=%d", T, K, E, c ? c->world : -1);
    return B200EP_ERR_ARG;
  }
  if (!dev_tokens_per_rank || !dev_tokens_per_expert || (T > 0 && (!dev_topk_ids || !dev_is_token_in_rank || !dev_send_off))) {
    set_error("layout: null pointer");
    return B200EP_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  B200EP_CUDA(cudaMemsetAsync(dev_tokens_per_expert, 0, sizeof(int32_t) * E, st));
  if (T == 0) {
    B200EP_CUDA(cudaMemsetAsync(dev_tokens_per_rank, 0, sizeof(int32_t) * c->world, st));
    return B200EP_OK;
  }
  const int ntiles = (T + 255) / 256;
  const int tpb = (ntiles + kMaxLayoutBlocks - 1) / kMaxLayoutBlocks;
  const int grid = (ntiles + tpb - 1) / tpb;
  const unsigned epoch = (unsigned)(++c->layout_epoch);
  const unsigned long long tbase = c->layout_ticket_base;
  c->layout_ticket_base += (unsigned long long)grid;
  const size_t smem = sizeof(int) * E;
  const int el = E / c->world;
#define LAUNCH(IDT, SINGLE)                                                                                      \
  layout_kernel<IDT, SINGLE><<<grid, 256, smem, st>>>((const IDT*)dev_topk_ids, T, K, E, c->world, el,            \
                                                      dev_tokens_per_rank, dev_tokens_per_expert,                 \
                                                      dev_is_token_in_rank, dev_send_off, c->local, tbase, epoch, tpb)
  if (ids_i64) {
    if (tpb == 1) LAUNCH(int64_t, true); else LAUNCH(int64_t, false);
  } else {
    if (tpb == 1) LAUNCH(int32_t, true); else LAUNCH(int32_t, false);
  }
#undef LAUNCH
  B200EP_CUDA(cudaGetLastError());
  return B200EP_OK;
}

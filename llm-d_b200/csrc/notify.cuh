// Count exchange that opens a high-throughput dispatch (DeepEP: notify_dispatch inside
// deep_ep.Buffer.dispatch, vllm/.../prepare_finalize/deepep_ht.py:139-162).
//
// Every rank writes its per-destination-rank and per-expert token counts into every peer's control
// block and raises a flag; once all W flags are up a rank knows the whole W x W count matrix, hence
//   rank_prefix[s]  first row of its receive region that rank s fills (canonical order: source rank
//                   ascending, source token ascending),
//   send_base[d]    first row it owns in rank d's receive region,
//   recv_per_expert tokens each local expert receives (num_recv_tokens_per_expert_list).
// These device functions are shared by the stand-alone notify kernels (phase-by-phase callers), the
// tail of the routing-layout kernel (which publishes the counts it has just computed) and the
// prologue of the dispatch kernel (which waits for the peers' counts itself): the production path
// needs no separate notify launch at all.
#pragma once

#include "common.cuh"

namespace b200ep {

// Publish this rank's counts to every peer (one CTA, all threads), then one system fence and W flags.
__device__ __forceinline__ void notify_publish(const Peers& peers, int W, int me, int slot, int e_local,
                                               const int32_t* tok_per_rank, const int32_t* tok_per_expert,
                                               unsigned long long* epoch_ptr) {
  for (int d = 0; d < W; ++d) {
    char* ctrl = peers.base[d];
    int32_t* cr = reinterpret_cast<int32_t*>(ctrl + kCtrlHtCntRankOff) + ((size_t)slot * kMaxRanks + me) * kMaxRanks;
    int32_t* ce = reinterpret_cast<int32_t*>(ctrl + kCtrlHtCntExpOff) + ((size_t)slot * kMaxRanks + me) * kMaxLocalExperts;
    if ((int)threadIdx.x < W) cr[threadIdx.x] = tok_per_rank[threadIdx.x];
    for (int i = threadIdx.x; i < e_local; i += blockDim.x) ce[i] = tok_per_expert[d * e_local + i];
  }
  single_cta_signal(epoch_ptr, peers, W, kFlagHtNotify, slot, me);
}

// Derive this rank's receive offsets / totals from the count matrix and post them to the host
// (one CTA, >= 64 threads).  counts come from the control block (filled by every rank's publish) or,
// for a single rank, straight from the layout arrays (no exchange, no flags, no fence).
template <bool kPeerWritten = true>  // true: counts were written by other GPUs into the control block (read through L2)
__device__ __forceinline__ void notify_finish(const int32_t* cr, int cr_stride, const int32_t* ce, int ce_stride, int W,
                                              int me, int slot, int e_local, HtMeta* meta, HostMirror* host,
                                              unsigned long long epoch) {
  const int tid = threadIdx.x;
  volatile unsigned long long* hw = host->word[slot];
  auto ld = [](const int32_t* p) -> int { return kPeerWritten ? __ldcg(p) : *p; };  // (__ldcg needs a global address)
  if (tid == 0) {
    int p = 0;
    for (int s = 0; s < W; ++s) {
      meta->rank_prefix[s] = p;
      hw[1 + s] = host_word(epoch, p);
      p += ld(&cr[s * cr_stride + me]);
    }
    meta->rank_prefix[W] = p;
    hw[1 + W] = host_word(epoch, p);
    meta->num_recv = p;
    hw[0] = host_word(epoch, p);
  }
  if (tid >= 32 && tid < 32 + W) {
    const int d = tid - 32;
    int b = 0;
    for (int s = 0; s < me; ++s) b += ld(&cr[s * cr_stride + d]);
    meta->send_base[d] = b;
    meta->send_count[d] = ld(&cr[me * cr_stride + d]);
  }
  for (int i = tid; i < e_local; i += blockDim.x) {
    int n = 0;
    for (int s = 0; s < W; ++s) n += ld(&ce[(size_t)s * ce_stride + i]);
    meta->recv_per_expert[i] = n;
    hw[1 + kMaxRanks + 1 + i] = host_word(epoch, n);
  }
}

__device__ __forceinline__ const int32_t* ctrl_cnt_rank(const char* my_base, int slot) {
  return reinterpret_cast<const int32_t*>(my_base + kCtrlHtCntRankOff) + (size_t)slot * kMaxRanks * kMaxRanks;
}
__device__ __forceinline__ const int32_t* ctrl_cnt_expert(const char* my_base, int slot) {
  return reinterpret_cast<const int32_t*>(my_base + kCtrlHtCntExpOff) + (size_t)slot * kMaxRanks * kMaxLocalExperts;
}

// What the routing-layout kernel's last block does after the counts are final.
enum NotifyTailMode { kTailNone = 0, kTailLocal = 1, kTailPublish = 2 };
struct NotifyTail {
  Peers peers;
  HtMeta* meta;
  HostMirror* host;
  unsigned long long* epoch_ptr;  // LocalScratch::epoch[kFlagHtNotify][slot]
  int mode, me, slot;
};

}  // namespace b200ep

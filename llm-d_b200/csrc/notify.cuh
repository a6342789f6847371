// Count exchange that opens a high-throughput dispatch (DeepEP: notify_dispatch inside
// deep_ep.Buffer.dispatch, vllm/.../prepare_finalize/deepep_ht.py:139-162).
//
// Every rank writes its per-destination-rank and per-expert token counts into every peer's control
// block; once a rank has them all it knows the whole W x W count matrix, hence
//   rank_prefix[s]  first row of its receive region that rank s fills (canonical order: source rank
//                   ascending, source token ascending),
//   send_base[d]    first row it owns in rank d's receive region,
//   recv_per_expert tokens each local expert receives (num_recv_tokens_per_expert_list).
// Each count is ONE 8-byte word (call tag << 32 | count), so a word is valid by itself: the writer needs
// no system fence and no flag, the reader polls exactly the words it needs (a fence.acq_rel.sys costs
// ~5 us on B200 and the count exchange sits on the critical path of every dispatch).
// These device functions are shared by the stand-alone notify kernels (phase-by-phase callers), the
// tail of the routing-layout kernel (which publishes the counts it has just computed) and the
// prologue of the dispatch kernel (which waits for the peers' counts itself): the production path
// needs no separate notify launch at all.
#pragma once

#include "common.cuh"

namespace b200ep {

__device__ __forceinline__ uint64_t* ctrl_cnt_rank(char* base, int slot, int src) {
  return reinterpret_cast<uint64_t*>(base + kCtrlHtCntRankOff) + ((size_t)slot * kMaxRanks + src) * kMaxRanks;
}
__device__ __forceinline__ uint64_t* ctrl_cnt_expert(char* base, int slot, int src) {
  return reinterpret_cast<uint64_t*>(base + kCtrlHtCntExpOff) + ((size_t)slot * kMaxRanks + src) * kMaxLocalExperts;
}
__device__ __forceinline__ uint64_t count_word(unsigned long long tag, int n) {
  return ((tag & 0xffffffffull) << 32) | (uint32_t)n;
}

// Publish this rank's counts to every peer (one CTA, all threads) and advance the call tag.
__device__ __forceinline__ uint64_t* ctrl_tokens(char* base, int slot, int src) {
  return reinterpret_cast<uint64_t*>(base + kCtrlHtTokOff) + (size_t)slot * kMaxRanks + src;
}
__device__ __forceinline__ void notify_publish(const Peers& peers, int W, int me, int slot, int e_local, int T,
                                               const int32_t* tok_per_rank, const int32_t* tok_per_expert,
                                               unsigned long long* epoch_ptr) {
  const unsigned long long tag = *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) + 1;
  if ((int)threadIdx.x >= 64 && (int)threadIdx.x < 64 + W)
    st_relaxed_sys_u64(ctrl_tokens(peers.base[threadIdx.x - 64], slot, me), count_word(tag, T));
  // the W x W rank counts first: the peers' dispatch kernels wait for nothing else before they push
  for (int i = threadIdx.x; i < W * W; i += blockDim.x)
    st_relaxed_sys_u64(ctrl_cnt_rank(peers.base[i / W], slot, me) + i % W, count_word(tag, tok_per_rank[i % W]));
  for (int i = threadIdx.x; i < W * e_local; i += blockDim.x) {
    const int d = i / e_local, el = i % e_local;
    st_relaxed_sys_u64(ctrl_cnt_expert(peers.base[d], slot, me) + el, count_word(tag, tok_per_expert[d * e_local + el]));
  }
  __syncthreads();
  if (threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long*>(epoch_ptr) = tag;
}

// Tokens rank `src` sends to rank `dst` in the call tagged `tag` (waits for the word; traps on a timeout).
__device__ __forceinline__ int notify_rank_count(char* my_base, int slot, int src, int dst, unsigned tag, int me) {
  return wait_tagged_word(ctrl_cnt_rank(my_base, slot, src) + dst, tag, kFlagHtNotify, me, src);
}

// Derive this rank's receive offsets / totals from the peers' words and post them to the host
// (one CTA, >= 96 threads).  The rank-level results go out first: the host is usually waiting for them.
__device__ __forceinline__ void notify_finish(char* my_base, int W, int me, int slot, int e_local, HtMeta* meta,
                                              HostMirror* host, unsigned long long epoch) {
  const int tid = threadIdx.x;
  const unsigned tag = (unsigned)epoch;
  volatile unsigned long long* hw = host->word[slot];
  if (tid == 0) {
    int p = 0;
    for (int s = 0; s < W; ++s) {
      meta->rank_prefix[s] = p;
      hw[1 + s] = host_word(epoch, p);
      p += notify_rank_count(my_base, slot, s, me, tag, me);
    }
    meta->rank_prefix[W] = p;
    hw[1 + W] = host_word(epoch, p);
    meta->num_recv = p;
    hw[0] = host_word(epoch, p);
  }
  if (tid >= 32 && tid < 32 + W) {
    const int d = tid - 32;
    int b = 0;
    for (int s = 0; s < me; ++s) b += notify_rank_count(my_base, slot, s, d, tag, me);
    meta->send_base[d] = b;
    meta->send_count[d] = notify_rank_count(my_base, slot, me, d, tag, me);
  }
  if (tid >= 64 && tid < 64 + W)
    meta->src_tokens[tid - 64] = wait_tagged_word(ctrl_tokens(my_base, slot, tid - 64), tag, kFlagHtNotify, me, tid - 64);
  for (int i = tid; i < e_local; i += blockDim.x) {
    int n = 0;
    for (int s = 0; s < W; ++s) n += wait_tagged_word(ctrl_cnt_expert(my_base, slot, s) + i, tag, kFlagHtNotify, me, s);
    meta->recv_per_expert[i] = n;
    hw[1 + kMaxRanks + 1 + i] = host_word(epoch, n);
  }
}

// world == 1: straight from the layout arrays (generic loads: they may live in shared memory).
__device__ __forceinline__ void notify_finish_local(const int32_t* tok_per_rank, const int32_t* tok_per_expert, int slot,
                                                    int e_local, int T, HtMeta* meta, HostMirror* host,
                                                    unsigned long long epoch) {
  const int tid = threadIdx.x;
  volatile unsigned long long* hw = host->word[slot];
  if (tid == 0) {
    const int p = tok_per_rank[0];
    meta->rank_prefix[0] = 0;
    meta->rank_prefix[1] = p;
    meta->num_recv = p;
    meta->send_base[0] = 0;
    meta->send_count[0] = p;
    meta->src_tokens[0] = T;
    hw[1] = host_word(epoch, 0);
    hw[2] = host_word(epoch, p);
    hw[0] = host_word(epoch, p);
  }
  for (int i = tid; i < e_local; i += blockDim.x) {
    const int n = tok_per_expert[i];
    meta->recv_per_expert[i] = n;
    hw[1 + kMaxRanks + 1 + i] = host_word(epoch, n);
  }
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

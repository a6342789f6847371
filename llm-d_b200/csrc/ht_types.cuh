// Types shared by the high-throughput combine kernels (ht.cu) and the expert-grouped variants (permute.cu).
#pragma once

#include "common.cuh"

namespace b200ep {

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
  unsigned* red_next;    // [kCombChunks] next reduce item of each chunk, [kCombChunks] = warps finished
  unsigned* piece_next;  // [kCombChunks] next DMA piece of each chunk (pieces are handed out from a queue)
  int T, W, me, slot;
  unsigned long long* trace;  // B200EP_TRACE=1: %globaltimer stamps go here (measurement aid), else null
  int rows_per_piece;    // rows one DMA piece moves (<= stage size)
  unsigned stage_bytes;  // bytes of one DMA stage
  int self_staged;       // 1: this rank's own rows were written to its stage too (expert-grouped combine), 0: read from y
};

// Stage on the SOURCE rank: stage[dst rank d][p] = the row rank d returns for the p-th token this
// rank sent to d (p = send_off[t][d]).  Rank d holds those rows contiguously, in the same order, as
// rows [rank_prefix[s], rank_prefix[s+1]) of its receive order -- so the push is a block copy.
__host__ __device__ inline size_t stage_row_off(const HtPlan& p, int d, int pos, int row_bytes) {
  return p.off_stage + ((size_t)d * p.t_cap + pos) * row_bytes;
}

// Fill the fields every combine launch shares (ht.cu) and queue the RECV phase (wait for all peers, fp32 reduce).
int ht_combine_args(b200ep_ctx* c, int slot, const void* dev_y, void* dev_out, const uint8_t* dev_is_token_in_rank,
                    const int32_t* dev_send_off, int T, int H, HtCombineArgs* a);
int ht_combine_launch_recv(b200ep_ctx* c, const HtCombineArgs& a, cudaStream_t st);

}  // namespace b200ep

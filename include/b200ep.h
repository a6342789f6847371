/*
 * b200ep -- B200-native wide-expert-parallel MoE token dispatch / combine.
 *
 * C ABI of libb200ep.so.  Plain pointers, sizes and ints only: every pointer
 * argument named dev_* is a CUDA device pointer owned by the caller, `stream`
 * is a cudaStream_t passed as void*.  Every entry point returns 0 on success or
 * a B200EP_ERR_* code; b200ep_last_error() then holds a message (thread local).
 *
 * The entry points are exactly what a binding for this path would call in place
 * of DeepEP's pybind module (`deep_ep_cpp`), which the reference image builds at
 * docker/Dockerfile.cuda:62-63 (neuralmagic/DeepEP@38d21b7f) and which vLLM
 * drives from
 *   vllm/distributed/device_communicators/all2all.py:143-325        (managers)
 *   vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ht.py
 *   vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ll.py
 * (line numbers: vLLM 0.22.0; the reference pins v0.23.0, docker/common-versions:5-6).
 * The reference-side binding (ctypes) is shown in INTEGRATION.md.
 *
 * Process model: one process per GPU.  Each rank owns one "symmetric" device
 * region; ranks exchange 64-byte CUDA-IPC handles once (over any host channel,
 * e.g. a torch.distributed CPU group) and from then on move tokens with
 * one-sided NVLink stores into each other's regions.  Every *_dispatch /
 * *_combine / notify call is COLLECTIVE: all ranks of the context's world must
 * make the same sequence of calls with the same `slot`.
 *
 * Phases.  The collective operations are split into phases so that (a) a
 * caller can overlap other work between "send" and "receive" (DeepEP's
 * return_recv_hook / async_finish), and (b) several ranks can be emulated on one
 * GPU by running phase p for all ranks before phase p+1 for any.  `phases` is a
 * bit mask; phases of one call must be issued in ascending bit order.
 */
#ifndef B200EP_H_
#define B200EP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200EP_MAX_RANKS 8          /* one NVSwitch domain (8 x B200) */
#define B200EP_MAX_LOCAL_EXPERTS 1024
#define B200EP_MAX_TOPK 16
#define B200EP_NUM_SLOTS 2          /* operations in flight (DBO micro-batches) */
#define B200EP_IPC_HANDLE_BYTES 64

enum {
  B200EP_OK = 0,
  B200EP_ERR_ARG = 1,      /* bad argument / unsupported shape */
  B200EP_ERR_CUDA = 2,     /* a CUDA runtime call failed */
  B200EP_ERR_STATE = 3,    /* call out of order (peers not connected, bad phase) */
  B200EP_ERR_CAPACITY = 4, /* symmetric region too small for this call */
  B200EP_ERR_TIMEOUT = 5   /* host wait for the count exchange timed out */
};

/* Payload kinds of the high-throughput dispatch (deepep_ht.py:258-311):
 *   BF16      x is bf16 [T,H], rows are moved as they are        (non-block-quantised models)
 *   BF16_QUANT x is bf16 [T,H]; rows are quantised to e4m3 with one fp32 scale per
 *             128 elements INSIDE the dispatch kernel -- replaces the separate
 *             per_token_group_quant_fp8 pass of deepep_ht.py:282-289
 *   FP8       x is e4m3 [T,H] + fp32 scales [T,H/128] quantised by the caller
 *             (what deepep_ht.py hands to deep_ep.Buffer.dispatch today)      */
enum { B200EP_X_BF16 = 0, B200EP_X_BF16_QUANT = 1, B200EP_X_FP8 = 2 };

enum { B200EP_PHASE_SEND = 1, B200EP_PHASE_RECV = 2, B200EP_PHASE_ALL = 3 };
/* low-latency dispatch has three phases: counts, payload, wait */
enum { B200EP_LL_PHASE_COUNT = 1, B200EP_LL_PHASE_SEND = 2, B200EP_LL_PHASE_RECV = 4, B200EP_LL_PHASE_ALL = 7 };

typedef struct b200ep_ctx b200ep_ctx;

/* ------------------------------------------------------------------ runtime */

const char* b200ep_last_error(void);
int b200ep_version(void);

/* Replaces deep_ep.Buffer.__init__ (all2all.py:222-229,297-306): allocates the
 * symmetric region (cudaMalloc, zeroed) on `device` plus private scratch.
 * sym_bytes plays the role of num_nvl_bytes / num_rdma_bytes. */
int b200ep_create(b200ep_ctx** out, int rank, int world, int device, size_t sym_bytes);
/* Replaces deep_ep.Buffer.destroy (all2all.py:190-194): closes peer mappings,
 * frees everything.  The caller must have synchronised all ranks first. */
int b200ep_destroy(b200ep_ctx* ctx);
/* CUDA-IPC bootstrap (DeepEP: Buffer.sync / get_local_ipc_handle).  Export the
 * local handle, all-gather the 64-byte blobs on the host, import all of them
 * (entry `rank` is ignored).  Peer setup for ranks living in the SAME process
 * (tests, single-process emulation) uses b200ep_set_peer_ptrs with the bases
 * returned by b200ep_sym_base instead. */
int b200ep_export_handle(b200ep_ctx* ctx, void* out_handle64);
int b200ep_import_handles(b200ep_ctx* ctx, const void* handles, int n);
int b200ep_set_peer_ptrs(b200ep_ctx* ctx, void* const* dev_bases, int n);
int b200ep_sym_base(b200ep_ctx* ctx, void** dev_base, size_t* bytes);
/* Address at which rank `rank`'s symmetric region is mapped in THIS process (its own base for rank == ctx rank);
 * valid once the peers are connected.  DeepEP keeps the equivalent table private (buffer_ptrs); it is exposed here
 * for tools that move data between the regions themselves (bench.py --impl probe). */
int b200ep_peer_base(b200ep_ctx* ctx, int rank, void** dev_base);
/* Measurement aid, no DeepEP counterpart: with B200EP_TRACE=1 in the environment the kernels record
 * %globaltimer stamps (ns) of their phases in the context's private scratch; this copies the 128 slots out
 * (synchronous; indices are listed in tools/trace_report.py). */
int b200ep_debug_trace(b200ep_ctx* ctx, unsigned long long* out128);
/* Replaces deep_ep.Buffer.set_num_sms (all2all.py:247-255): upper bound on the
 * CTAs the data-moving kernels launch (0 = one wave over all SMs). */
int b200ep_set_num_sms(b200ep_ctx* ctx, int num_sms);

/* ------------------------------------------------------------------- layout */

/* Replaces deep_ep.Buffer.get_dispatch_layout (deepep_ht.py:121-133).
 *   dev_topk_ids        [T,K] int64 (ids_i64 != 0) or int32; negative = no expert
 *   dev_tokens_per_rank   [world] int32   out
 *   dev_tokens_per_expert [E]     int32   out
 *   dev_is_token_in_rank  [T,world] uint8 out (0/1; torch.bool compatible)
 *   dev_send_off          [T,world] int32 out: position of token t among the
 *                         tokens this rank sends to rank d, -1 if it does not
 *                         (exclusive prefix count; makes the receive order
 *                         canonical without atomics).  No DeepEP counterpart.
 * The kernel keeps its per-call state (block tickets, call epoch, expert histogram) in the context's
 * private device scratch and rewinds it itself, so the launch takes nothing from host counters and
 * may be replayed from a CUDA graph; but all calls on ONE context share that scratch: they must not
 * overlap on the GPU (issue them on one stream, as the deep_ep shim does, or order them with events). */
int b200ep_layout(b200ep_ctx* ctx, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                  int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                  uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, void* stream);

/* Same outputs as b200ep_layout, but no context and NO device allocation anywhere on the path.
 * This entry point has no counterpart in the reference (DeepEP's layout always runs on a Buffer).
 * It exists for callers that cannot allocate when they need a layout: a serving worker whose
 * remaining GPU memory already belongs to its KV-cache pool (a late cudaMalloc fails or shrinks
 * the pool's headroom), and tools/tests that only want counts and offsets and have no reason to
 * create a context (b200ep_create allocates gigabytes).  The kernel's scratch is a fixed array
 * inside the library (about 16 KB on every GPU that loads it, until the process exits), hence:
 *   - T <= 65536 tokens per call; more is rejected with B200EP_ERR_CAPACITY;
 *   - all calls on a device share that array and each call clears it first, so calls whose
 *     kernels overlap on the GPU race: serialise them (one stream, or external ordering);
 *   - the first call on a device resolves the array's address before launching; make that
 *     first call before any stream capture begins;
 *   - call it with the device that owns `stream` current (the address is looked up and cached per
 *     device ordinal); the cached address does not survive a cudaDeviceReset.
 * `world` is the number of ranks the experts are sharded over (1..B200EP_MAX_RANKS). */
int b200ep_layout_noalloc(int world, const void* dev_topk_ids, int ids_i64, int T, int K, int E,
                          int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                          uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, void* stream);

/* ------------------------------------------------- fp8 quantiser (stand-alone) */

/* Same arithmetic as vLLM per_token_group_quant_fp8 with group 128
 * (quantization/utils/fp8_utils.py:510-591, CUDA op per_token_group_quant_8bit):
 * s = max(amax,1e-10)/448 (IEEE division), q = sat_e4m3(x / s); ue8m0 rounds s up
 * to a power of two.  The dispatch kernels call the same device function. */
int b200ep_quant_fp8(const void* dev_x_bf16, void* dev_q, float* dev_scales, int T, int H, int ue8m0,
                     void* stream);

/* ------------------------------------------------ high-throughput (prefill) */

/* Largest T (tokens per rank) one slot can take for this shape, or < 0. */
int b200ep_ht_capacity(b200ep_ctx* ctx, int H, int K, int x_kind);

/* Count exchange that opens a dispatch (DeepEP notify_dispatch, inside
 * deep_ep.Buffer.dispatch, deepep_ht.py:139-162).
 * SEND: publish this rank's per-rank / per-expert counts and its token count T to every peer.
 * RECV: wait for all peers, derive receive offsets and per-expert totals and
 *       post them to pinned host memory.
 * b200ep_ht_notify_wait then blocks the HOST until RECV has run and returns
 *   *num_recv_tokens, per_expert[E/world] (num_recv_tokens_per_expert_list,
 *   deepep_ht.py:143,225-227) and rank_prefix[world+1] (rows
 *   [rank_prefix[s], rank_prefix[s+1]) of recv_x come from rank s). */
int b200ep_ht_notify(b200ep_ctx* ctx, int slot, const int32_t* dev_tokens_per_rank,
                     const int32_t* dev_tokens_per_expert, int T, int E, int phases, void* stream);
int b200ep_ht_notify_wait(b200ep_ctx* ctx, int slot, int* num_recv_tokens, int32_t* per_expert,
                          int32_t* rank_prefix, int timeout_ms);

/* Replaces deep_ep.Buffer.dispatch (deepep_ht.py:139-162) after the notify.
 * SEND: permute (+ quantise) + write every token once into each destination
 *       rank's receive region, with its top-k ids rewritten to destination-local
 *       ids (-1 elsewhere), weights (0 elsewhere) and source index.
 * RECV: wait until every source rank's rows have landed.
 * Results are read in place from the symmetric region, see b200ep_ht_views. */
int b200ep_ht_dispatch(b200ep_ctx* ctx, int slot, const void* dev_x, const float* dev_x_scales,
                       int x_kind, int ue8m0, const void* dev_topk_ids, int ids_i64,
                       const float* dev_topk_weights, const uint8_t* dev_is_token_in_rank,
                       const int32_t* dev_send_off, int T, int H, int K, int E, int phases,
                       void* stream);

/* b200ep_layout + b200ep_ht_notify + b200ep_ht_notify_wait + b200ep_ht_dispatch (all phases) in one
 * host call: the whole of deep_ep.Buffer.get_dispatch_layout + Buffer.dispatch (deepep_ht.py:121-162).
 * The four layout arrays are outputs the caller keeps for the combine.  Internally there is no notify launch: the
 * layout kernel's last block publishes the counts, the dispatch kernel waits for the peers' counts itself, and the
 * host waits for its row counts only after everything is queued (the wait runs under the dispatch kernel).
 * timeout_ms < 0 selects the HOST-SYNC-FREE form (DeepEP's num_worst_tokens mode): no host wait, the three
 * host outputs are not written; the number of received rows and the per-expert counts are read on the
 * device (b200ep_ht_view.rank_prefix[world], .recv_per_expert) and the caller sizes recv views for a worst case. */
int b200ep_ht_dispatch_all(b200ep_ctx* ctx, int slot, const void* dev_x, const float* dev_x_scales, int x_kind,
                           int ue8m0, const void* dev_topk_ids, int ids_i64, const float* dev_topk_weights,
                           int32_t* dev_tokens_per_rank, int32_t* dev_tokens_per_expert,
                           uint8_t* dev_is_token_in_rank, int32_t* dev_send_off, int T, int H, int K, int E,
                           int* num_recv_tokens, int32_t* per_expert, int32_t* rank_prefix, int timeout_ms,
                           void* stream);

typedef struct {
  void* recv_x;                /* [M, H] e4m3 or bf16 */
  float* recv_x_scales;        /* [M, H/128] fp32, NULL for B200EP_X_BF16 */
  int64_t* recv_topk_idx;      /* [M, K] local expert ids or -1 */
  float* recv_topk_weights;    /* [M, K] */
  int32_t* recv_src_idx;       /* [M] token index on the source rank */
  int32_t* rank_prefix;        /* [world+1] device copy */
  int32_t* recv_per_expert;    /* [E/world] device copy */
  int32_t capacity_rows;       /* rows the region can hold */
} b200ep_ht_view;
int b200ep_ht_views(b200ep_ctx* ctx, int slot, b200ep_ht_view* out);

/* Replaces deep_ep.Buffer.combine (deepep_ht.py:364-376), topk_weights=None.
 * SEND: push every received row's result dev_y[m] (bf16 [M,H]) back to its source.
 * RECV: wait for all peers; out[t] = bf16( sum over destination ranks, fp32,
 *       ascending rank order ) into dev_out [T,H] (the "output.copy_" of
 *       deepep_ht.py:387,398 is folded in).  Tokens sent nowhere give zeros. */
int b200ep_ht_combine(b200ep_ctx* ctx, int slot, const void* dev_y, void* dev_out,
                      const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H,
                      int phases, void* stream);

/* ------------------------------ expert-grouped layout for grouped GEMMs (after an HT dispatch) */

/* Replaces deepgemm_moe_permute -> ep_scatter (vLLM fused_moe/deep_gemm_utils.py:347-420, 56-249; called from
 * fused_moe/experts/deep_gemm_moe.py:279-291) on the rows the dispatch of `slot` received: every row is copied,
 * with its scales, to one output row per LOCAL expert in its top-k.  Expert e owns output rows
 * [start[e], start[e] + count[e]), start = exclusive prefix of the counts rounded up to `alignment`
 * (compute_aligned_M, deep_gemm_utils.py:26-44; 128 for DeepGEMM's contiguous layout).
 *   rows_bound        host-side upper bound on the rows received (num_recv_tokens, or the worst case after a
 *                     host-sync-free dispatch): sizes the grids, the true count is read on the device
 *   dev_out_x         [m_cap, H] e4m3 (or [m_cap, H] bf16 for B200EP_X_BF16)          out, 32-byte aligned
 *   dev_out_scales    [m_cap, H/128] fp32, row-major (NULL for B200EP_X_BF16)         out
 *   dev_expert_ids    [m_cap] int32: expert of every occupied row, -1 in the padding  out ("m_indices")
 *   dev_inv_perm      [rows_bound, K] int32: output row of entry (m, k), -1 if none   out
 *   dev_expert_start  [E/world + 1] int32: start[], last = rows the layout needs      out, may be NULL
 *   m_cap             rows the caller's buffers hold (the kernel traps if the layout needs more)
 * Rows inside an expert are in received order (source rank, then source token): deterministic, where the
 * reference uses an atomic counter.  A token's top-k ids must be distinct (top-k routing guarantees it).
 * Calls on one context must not overlap on the GPU (they share the plan kernel's scratch). */
int b200ep_ht_permute(b200ep_ctx* ctx, int slot, int alignment, int rows_bound, void* dev_out_x,
                      float* dev_out_scales, int32_t* dev_expert_ids, int32_t* dev_inv_perm,
                      int32_t* dev_expert_start, int m_cap, void* stream);

/* Replaces deepgemm_unpermute_and_reduce -> ep_gather (deep_gemm_utils.py:423-440, 252-344; called from
 * deep_gemm_moe.py:322-329) AND the deep_ep.Buffer.combine that follows it (deepep_ht.py:364-376) in one
 * operation.  dev_y_grouped is the grouped GEMM output, bf16 [m_cap, H], rows as laid out by b200ep_ht_permute.
 * SEND: for every received token m the expert rank forms z[m] = bf16( sum_k w[m,k] * y[inv_perm[m,k]] ) over the
 *       entries with a local expert (fp32 accumulator, one fused multiply-add per entry, k ascending: ep_gather's
 *       arithmetic) inside the push and sends it to the token's source rank; no [M, H] intermediate is written.
 * RECV: as b200ep_ht_combine (fp32 sum over the returned rows in rank order, one bf16 rounding).
 * dev_topk_weights: [rows, K] fp32, or NULL for the weights the dispatch delivered (recv_topk_weights); pass
 * ones when the router weights were already applied to the input (apply_router_weight_on_input). */
int b200ep_ht_combine_permuted(b200ep_ctx* ctx, int slot, const void* dev_y_grouped, const int32_t* dev_inv_perm,
                               const float* dev_topk_weights, void* dev_out, const uint8_t* dev_is_token_in_rank,
                               const int32_t* dev_send_off, int T, int H, int phases, void* stream);

/* ---------------------------------------------------- low-latency (decode) */

/* Replaces deep_ep.Buffer.get_low_latency_rdma_size_hint (all2all.py:287-292). */
size_t b200ep_ll_size_hint(int max_tokens_per_rank, int H, int world, int E);

/* Replaces deep_ep.Buffer.low_latency_dispatch (deepep_ll.py:292-314).
 * No host synchronisation, fixed shapes, safe to capture in a CUDA graph.
 * COUNT: per-expert counts of this rank's tokens -> all peers.
 * SEND : wait for counts, (quantise,) write each (token, expert) row straight to
 *        its final slot of packed_recv_x on the expert's owner.
 * RECV : wait for all sources (the "hook").
 * ue8m0: 0 = fp32 scales; 1 = fp32 scales rounded up to powers of two (DeepEP round_scale);
 *        2 = the same power-of-two scales as PACKED exponent bytes (DeepEP use_ue8m0; what DeepGEMM reads on
 *            sm_100, vLLM fp8_utils.py:643-714): b200ep_ll_view.packed_recv_scales is then an int32 array
 *            [E_local][ceil(H/512)][slots rounded up to 4], byte b of word j = biased exponent of group 4j+b,
 *            i.e. logical shape [E_local, slots, H/512] with the slot index fastest ("MN-major"). */
int b200ep_ll_dispatch(b200ep_ctx* ctx, int slot, const void* dev_x_bf16, const void* dev_topk_ids,
                       int ids_i64, int T, int H, int K, int E, int max_tokens_per_rank, int use_fp8,
                       int ue8m0, int phases, void* stream);

typedef struct {
  void* packed_recv_x;        /* [E_local, world*max_tok, H] e4m3 or bf16 */
  float* packed_recv_scales;  /* [E_local, world*max_tok, H/128] fp32 (NULL if bf16); int32 packed UE8M0 if ue8m0 == 2 */
  int32_t* recv_count;        /* [E_local] */
  int32_t* src_info;          /* [E_local, world*max_tok]: t*K+k on the source rank */
  int32_t* layout_range;      /* [E_local, world, 2]: (first slot, count) per source */
} b200ep_ll_view;
int b200ep_ll_views(b200ep_ctx* ctx, int slot, b200ep_ll_view* out);

/* Replaces deep_ep.Buffer.low_latency_combine (deepep_ll.py:399-408).
 * SEND: push every occupied slot's result row dev_y[e_local, slot, :] (bf16) back.
 * RECV: wait; out[t] = bf16( sum_k w[t,k] * y(t,k) ), fp32 multiply then add,
 *       k ascending. */
int b200ep_ll_combine(b200ep_ctx* ctx, int slot, const void* dev_y, const void* dev_topk_ids,
                      int ids_i64, const float* dev_topk_weights, void* dev_out, int T, int H, int K,
                      int E, int max_tokens_per_rank, int phases, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200EP_H_ */

// Expert-grouped permute of the rows a high-throughput dispatch received, and the combine that undoes it
// (SURVEY.md 8(f) f1).
//
// Replaces, on vLLM's DeepGEMM expert path (fused_moe/experts/deep_gemm_moe.py:279-329),
//   deepgemm_moe_permute          -> ep_scatter  (fused_moe/deep_gemm_utils.py:347-420, 56-249)
//   deepgemm_unpermute_and_reduce -> ep_gather   (deep_gemm_utils.py:252-344, 423-440)
// which are Triton passes between deep_ep.Buffer.dispatch / .combine and the grouped GEMMs.
// CPU restatement: oracle/permute.py.
//
//   permute  b200ep_ht_permute: two launches after the dispatch.
//     plan     one thread per received row; a block owns 256 consecutive rows.  Per-expert bit masks in
//              shared memory give every (row, k) entry its position among the block's rows for that expert
//              (popc below the row's bit: deterministic, no atomics for slots); block totals per expert are
//              published as epoch-tagged words and summed over the logically earlier blocks (ticketed
//              look-back, as in the routing-layout kernel).  Output row of an entry =
//              start[e] + rows of earlier blocks for e + position in the block, with start[] the exclusive
//              prefix of the per-expert counts rounded up to the GEMM's M alignment.  The same kernel writes
//              expert_ids (the expert of every occupied output row, -1 in the padding).
//              Order inside an expert = received-row order (source rank, then source token): the reference
//              hands rows out with an atomic counter, i.e. in no particular order.
//     scatter  a warp per (row, 1 KB segment): one 32-byte load per lane, one 32-byte store per local expert
//              of the row; scales likewise.  HBM-bound: M x row bytes in, (entries) x row bytes out.
//              A row crosses NVLink once per destination RANK and is replicated per local EXPERT here, on
//              the receiving GPU, at HBM speed (sending it once per expert would cost ~1.5x the NVLink
//              bytes at EP=8 and 4x at EP=2).
//   combine  b200ep_ht_combine_permuted: the expert rank forms the row it returns for a received token,
//              z[m] = bf16( sum_k fma(y[inv_perm[m,k]], w[m,k]) )   (fp32 accumulator, k ascending; ep_gather's
//              arithmetic), INSIDE the push: a warp gathers the <= K expert-output rows of a token straight
//              from the grouped GEMM output, reduces them in registers, writes the bf16 row into its shared
//              memory stage and sends it to the source rank with one TMA bulk copy.  The intermediate
//              [M, H] tensor ep_gather writes and Buffer.combine reads again never exists.  The source rank
//              then sums the <= W returned rows in fp32, rank order, exactly as b200ep_ht_combine does.
#include <stdlib.h>

#include "ht_types.cuh"

namespace b200ep {

constexpr int kPermTile = 256;

// Per-(block, expert) totals are self-validating words (call tag << 32 | count): a reader needs no acquire, it
// polls with relaxed loads until the tag is this call's.  Traps instead of hanging.
__device__ __forceinline__ uint64_t ld_relaxed_gpu_u64(const unsigned long long* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t perm_wait_pub(const unsigned long long* p, unsigned tag, uint64_t v) {
  if ((unsigned)(v >> 32) == tag) return v;
  const long long t0 = clock64();
  do {
    if (clock64() - t0 > 20000000000LL) {  // ~10 s: the publishing block always runs (ticket order)
      printf("b200ep: permute plan block %d timed out waiting for an earlier block's totals (tag %u, have %llu)\n",
             (int)blockIdx.x, tag, (unsigned long long)v);
      __trap();
    }
    v = ld_relaxed_gpu_u64(p);
  } while ((unsigned)(v >> 32) != tag);
  return v;
}

struct PermPlanArgs {
  const long long* recv_idx;  // [M, K] local expert ids or -1 (the slot's recv_topk_idx)
  const HtMeta* meta;         // num_recv and recv_per_expert of the dispatch
  PermScratch* ps;
  int32_t* inv_perm;          // [rows, K] out
  int32_t* expert_ids;        // [m_cap] out
  int32_t* expert_start;      // [e_local + 1] out, may be null
  int K, e_local, align, m_cap;
};

__global__ void __launch_bounds__(kPermTile) perm_plan_kernel(const __grid_constant__ PermPlanArgs a) {
  extern __shared__ int s_dyn[];
  const int E = a.e_local, K = a.K;
  unsigned* s_mask = reinterpret_cast<unsigned*>(s_dyn);  // [E][8]: bit r of word w = row 32w + r of this block routes to e
  int* s_start = s_dyn + E * 8;                           // [E + 1]
  int* s_cnt = s_start + E + 1;                           // [E]
  int* s_base = s_cnt + E;                                // [E]
  int* s_wpre = s_base + E;                               // [E][8]: rows of this block routed to e in the words below w
  __shared__ int s_wtot[8];
  __shared__ unsigned s_bid, s_tag, s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    s_tag = (unsigned)(*reinterpret_cast<volatile unsigned long long*>(&a.ps->epoch) + 1);
    s_bid = (unsigned)atomicAdd(&a.ps->ticket, 1ull);  // logical block id follows scheduling order
  }
  for (int i = tid; i < E * 8; i += kPermTile) s_mask[i] = 0u;
  for (int i = tid; i < E; i += kPermTile) s_base[i] = 0;
  // start[]: exclusive prefix of the aligned per-expert counts (every block computes it: E <= 1024)
  const int per = (E + kPermTile - 1) / kPermTile;
  const int lo = min(E, tid * per), hi = min(E, lo + per);
  int sum = 0;
  for (int i = lo; i < hi; ++i) {
    const int c = a.meta->recv_per_expert[i];
    s_cnt[i] = c;
    sum += (c + a.align - 1) / a.align * a.align;
  }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_wtot[warp] = incl;
  __syncthreads();
  int excl = incl - sum;
  for (int w = 0; w < warp; ++w) excl += s_wtot[w];
  {
    int run = excl;
    for (int i = lo; i < hi; ++i) {
      s_start[i] = run;
      run += (s_cnt[i] + a.align - 1) / a.align * a.align;
    }
    if (tid == kPermTile - 1) s_start[E] = excl + sum;
  }
  const int bid = (int)s_bid;
  const unsigned tag = s_tag;
  const int M = *reinterpret_cast<const volatile int32_t*>(&a.meta->num_recv);
  const int m = bid * kPermTile + tid;
  int my_e[kMaxTopk];
#pragma unroll
  for (int k = 0; k < kMaxTopk; ++k) my_e[k] = -1;
  if (m < M) {
#pragma unroll
    for (int k = 0; k < kMaxTopk; ++k) {
      if (k < K) {
        const long long id = __ldcg(a.recv_idx + (size_t)m * K + k);
        if (id >= 0 && id < E) {
          my_e[k] = (int)id;
          atomicOr(&s_mask[(int)id * 8 + warp], 1u << lane);
        }
      }
    }
  }
  __syncthreads();
  // this block's totals per expert -> published; everything needed from OTHER blocks comes from these words
  unsigned long long* pub = a.ps->pub;
  for (int e = tid; e < E; e += kPermTile) {
    int n = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      s_wpre[e * 8 + w] = n;
      n += __popc(s_mask[e * 8 + w]);
    }
    st_release_gpu_u64(reinterpret_cast<uint64_t*>(&pub[(size_t)bid * E + e]), ((uint64_t)tag << 32) | (uint32_t)n);
  }
  // rows of logically earlier blocks per expert: four independent loads in flight per thread (one dependent load
  // per word made this loop -- up to bid * E / 256 round trips -- most of the kernel's time)
  const long long nwords = (long long)bid * E;
  for (long long i0 = tid; i0 < nwords; i0 += 4 * kPermTile) {
    uint64_t v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + (long long)u * kPermTile;
      v[u] = i < nwords ? ld_relaxed_gpu_u64(&pub[i]) : 0;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + (long long)u * kPermTile;
      if (i < nwords) {
        const int n = (int)(uint32_t)perm_wait_pub(&pub[i], tag, v[u]);
        if (n) atomicAdd(&s_base[(int)(i % E)], n);
      }
    }
  }
  __syncthreads();
  if (m < M) {
#pragma unroll
    for (int k = 0; k < kMaxTopk; ++k) {
      if (k < K) {
        int dst = -1;
        const int e = my_e[k];
        if (e >= 0) {
          const int p = s_wpre[e * 8 + warp] + __popc(s_mask[e * 8 + warp] & ((1u << lane) - 1u));
          dst = s_start[e] + s_base[e] + p;
        }
        a.inv_perm[(size_t)m * K + k] = dst;
      }
    }
  }
  // expert_ids: the expert of every occupied output row, -1 in the alignment padding and beyond the total
  const int total = s_start[E];
  if (total > a.m_cap && bid == 0 && tid == 0) {
    printf("b200ep: ht_permute: the grouped layout needs %d rows, the caller's buffers hold %d\n", total, a.m_cap);
    __trap();
  }
  // every block fills one contiguous chunk of the output rows: one search for a thread's first row, then it walks forward
  // (start[] is non-decreasing; "the largest e with start[e] <= i" skips experts without rows)
  {
    const int chunk = (a.m_cap + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo_row = bid * chunk, hi_row = min(a.m_cap, lo_row + chunk);
    int i = lo_row + tid, e = 0;
    if (i < total) {
      int l = 0, h = E - 1;
      while (l < h) {
        const int mid = (l + h + 1) >> 1;
        if (s_start[mid] <= i) l = mid; else h = mid - 1;
      }
      e = l;
    }
    for (; i < hi_row; i += kPermTile) {
      int id = -1;
      if (i < total) {
        while (e + 1 < E && s_start[e + 1] <= i) ++e;
        if (i - s_start[e] < s_cnt[e]) id = e;
      }
      a.expert_ids[i] = id;
    }
  }
  if (a.expert_start && bid == 0)
    for (int i = tid; i <= E; i += kPermTile) a.expert_start[i] = s_start[i];
  // the last block to finish rewinds the ticket and advances the tag (all blocks have read both by then)
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    s_last = atom_add_acq_rel_gpu(&a.ps->done, 1u) == gridDim.x - 1;
    if (s_last) {
      a.ps->ticket = 0;
      a.ps->done = 0;
      volatile unsigned long long* ep = &a.ps->epoch;
      *ep = *ep + 1;
    }
  }
}

struct PermScatterArgs {
  const char* x;            // [M, row_bytes] received rows
  const float* scales;      // [M, ng] or null
  const int32_t* inv_perm;  // [M, K]
  const HtMeta* meta;
  char* out_x;              // [m_cap, row_bytes]
  float* out_scales;        // [m_cap, ng]
  int row_bytes, ng, K;
};

__device__ __forceinline__ void st_v8_u(void* p, const V8& v) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]),
               "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
               : "memory");
}

// A warp per (received row, 1 KB segment): the segment is loaded once and stored once per local expert of the row.
__global__ void __launch_bounds__(256) perm_scatter_kernel(const __grid_constant__ PermScatterArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = a.row_bytes / 32, nseg = (nv + 31) / 32, K = a.K;
  const long long it = (long long)blockIdx.x * 8 + warp;
  const int M = *reinterpret_cast<const volatile int32_t*>(&a.meta->num_recv);
  if (it >= (long long)M * nseg) return;
  const int m = (int)(it / nseg), seg = (int)(it % nseg);
  const int v = seg * 32 + lane;
  V8 raw;
  if (v < nv) raw = ld_nc_v8(a.x + (size_t)m * a.row_bytes + (size_t)v * 32);
  float sc[2] = {0.f, 0.f};
  const bool do_scales = seg == 0 && a.scales != nullptr;
  if (do_scales) {
#pragma unroll
    for (int j = 0; j < 2; ++j)
      if (lane + 32 * j < a.ng) sc[j] = __ldg(a.scales + (size_t)m * a.ng + lane + 32 * j);
  }
  int dst = -1;
  if (lane < K) dst = __ldg(a.inv_perm + (size_t)m * K + lane);
  unsigned kmask = __ballot_sync(0xffffffffu, dst >= 0);
  while (kmask) {
    const int k = __ffs(kmask) - 1;
    kmask &= kmask - 1;
    const int row = __shfl_sync(0xffffffffu, dst, k);
    if (v < nv) st_v8_u(a.out_x + (size_t)row * a.row_bytes + (size_t)v * 32, raw);
    if (do_scales) {
#pragma unroll
      for (int j = 0; j < 2; ++j)
        if (lane + 32 * j < a.ng) a.out_scales[(size_t)row * a.ng + lane + 32 * j] = sc[j];
      for (int g = lane + 64; g < a.ng; g += 32)  // rows wider than 8192 values
        a.out_scales[(size_t)row * a.ng + g] = __ldg(a.scales + (size_t)m * a.ng + g);
    }
  }
}

// ------------------------------------------------------------------ combine from the grouped layout
struct GatherArgs {
  HtCombineArgs c;            // c.y = grouped expert outputs, bf16 [m_cap, H]
  const int32_t* inv_perm;    // [M, K]
  const long long* recv_idx;  // [M, K] local ids or -1
  const float* recv_w;        // [M, K]
  int K;
  int nparts;                 // W > 1: a row is pushed in this many parts (one warp each)
};

constexpr int kGatU = 4;  // 16-byte vectors per lane and batch

__device__ __forceinline__ void fma8(float (&acc)[8], const int4& v, float w) {
  float f[8];
  unpack8(v, f);
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = __fmaf_rn(f[i], w, acc[i]);
}
__device__ __forceinline__ int4 pack8_bf16_rn(const float (&acc)[8]) {
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 p = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
    o[i] = *reinterpret_cast<uint32_t*>(&p);
  }
  return make_int4((int)o[0], (int)o[1], (int)o[2], (int)o[3]);
}

// One batch (kGatU x 32 vectors) of z[m]: all loads of up to two expert rows in flight, fma in k order.
__device__ __forceinline__ void gather_batch(const char* y, int row_bytes, unsigned kmask, int my_dst, float my_w, int v0,
                                             int nvec, int lane, float (&acc)[kGatU][8]) {
#pragma unroll
  for (int j = 0; j < kGatU; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
  unsigned m = kmask;
  while (m) {
    const int k0 = __ffs(m) - 1;
    m &= m - 1;
    const int k1 = m ? __ffs(m) - 1 : -1;
    if (k1 >= 0) m &= m - 1;
    const int r0 = __shfl_sync(0xffffffffu, my_dst, k0);
    const float w0 = __shfl_sync(0xffffffffu, my_w, k0);
    const int r1 = __shfl_sync(0xffffffffu, my_dst, k1 >= 0 ? k1 : 0);
    const float w1 = __shfl_sync(0xffffffffu, my_w, k1 >= 0 ? k1 : 0);
    int4 a0[kGatU], a1[kGatU];
#pragma unroll
    for (int j = 0; j < kGatU; ++j) {
      const int v = v0 + j * 32 + lane;
      if (v < nvec) {
        a0[j] = ld_nc_v4(y + (size_t)r0 * row_bytes + (size_t)v * 16);
        if (k1 >= 0) a1[j] = ld_nc_v4(y + (size_t)r1 * row_bytes + (size_t)v * 16);
      }
    }
#pragma unroll
    for (int j = 0; j < kGatU; ++j) {
      const int v = v0 + j * 32 + lane;
      if (v < nvec) {
        fma8(acc[j], a0[j], w0);
        if (k1 >= 0) fma8(acc[j], a1[j], w1);
      }
    }
  }
}

// W > 1, phase SEND.  A warp per (received row, part of the row), two shared-memory stages per warp: gather +
// reduce into the stage, one TMA bulk copy to stage[me][position] on the source rank (this rank's own rows go to
// its own stage the same way: the reduction then reads every rank's row from the stage).  Rows are cut into
// `nparts` parts of at most 8 KB so that more warps fit the shared memory.  A warp's work is a chain of memory
// round trips, and at EP=8 the fabric takes a part every ~20 us per warp, so the chain is kept short: the next
// item's ids / weights / rows are fetched while the current one is reduced, and the work of an item is a sequence of
// (batch, pair of experts) steps whose loads run one step ahead of the accumulation.
// Items interleave the destinations, starting with rank me+1, so that the ranks push along a permutation.
__global__ void __launch_bounds__(384, 1) ht_combine_gather_send_kernel(const __grid_constant__ GatherArgs g) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ int s_prefix[kMaxRanks + 1];
  const HtCombineArgs& a = g.c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int W = a.W, K = g.K, row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  const int nparts = g.nparts, part_vec = (nvec + nparts - 1) / nparts, part_bytes = part_vec * 16;
  char* s_stage = s_mem + (size_t)warp * 2 * part_bytes;
  if ((int)threadIdx.x <= W) s_prefix[threadIdx.x] = a.meta->rank_prefix[threadIdx.x];
  __syncthreads();
  int maxcnt = 0;
  for (int s = 0; s < W; ++s) maxcnt = max(maxcnt, s_prefix[s + 1] - s_prefix[s]);
  const long long total = (long long)maxcnt * W * nparts, stride = (long long)gridDim.x * nwarp;
  const char* y = reinterpret_cast<const char*>(a.y);

  struct Item {
    int s, j, part, dst;  // destination rank, position in its block, part of the row; lane k: grouped row of entry k or -1
    float w;
  };
  // first valid item at or after `it` (a source with fewer rows than the largest block leaves holes), its metadata loads issued
  auto fetch = [&](long long it, Item& q) -> long long {
    for (; it < total; it += stride) {
      const long long rit = it / nparts;
      q.s = (a.me + 1 + (int)(rit % W)) % W;
      q.j = (int)(rit / W);
      if (q.j < s_prefix[q.s + 1] - s_prefix[q.s]) break;
    }
    if (it >= total) return total;
    q.part = (int)(it % nparts);
    const int m = s_prefix[q.s] + q.j;
    q.dst = -1;
    q.w = 0.f;
    if (lane < K) {  // three independent loads; an entry counts if it has a local expert and a grouped row
      const long long id = __ldcg(g.recv_idx + (size_t)m * K + lane);
      const int d = __ldg(g.inv_perm + (size_t)m * K + lane);
      q.w = __ldcg(g.recv_w + (size_t)m * K + lane);
      q.dst = id >= 0 ? d : -1;
    }
    return it;
  };

  Item cur, nxt;
  long long it = fetch((long long)blockIdx.x * nwarp + warp, cur);
  int n = 0;  // parts this warp has staged
  while (it < total) {
    const long long it_next = fetch(it + stride, nxt);  // in flight while this item is reduced
    const unsigned kmask = __ballot_sync(0xffffffffu, cur.dst >= 0);
    const int npairs = max(1, (__popc(kmask) + 1) >> 1);
    char* st = s_stage + (size_t)(n & 1) * part_bytes;
    ++n;
    const int vbeg = cur.part * part_vec, vend = min(nvec, vbeg + part_vec);
    const int nsteps = ((vend - vbeg + 32 * kGatU - 1) / (32 * kGatU)) * npairs;
    // A step = (batch of kGatU x 32 vectors, pair of the row's experts in k order).  The loads of step i+1 are in
    // flight while step i is accumulated; a batch is rounded and stored after its last pair.
    float acc[kGatU][8];
    auto load = [&](int step, int4 (&r)[2][kGatU]) {
      const int v0 = vbeg + (step / npairs) * 32 * kGatU, p = step % npairs;
      const unsigned ka = __fns(kmask, 0, 2 * p + 1), kb = __fns(kmask, 0, 2 * p + 2);
      const int ra_ = __shfl_sync(0xffffffffu, cur.dst, ka & 31), rb_ = __shfl_sync(0xffffffffu, cur.dst, kb & 31);
#pragma unroll
      for (int jj = 0; jj < kGatU; ++jj) {
        const int v = v0 + jj * 32 + lane;
        if (v < vend) {
          if (ka < 32) r[0][jj] = ld_nc_v4(y + (size_t)ra_ * row_bytes + (size_t)v * 16);
          if (kb < 32) r[1][jj] = ld_nc_v4(y + (size_t)rb_ * row_bytes + (size_t)v * 16);
        }
      }
    };
    auto accumulate = [&](int step, const int4 (&r)[2][kGatU]) {
      const int v0 = vbeg + (step / npairs) * 32 * kGatU, p = step % npairs;
      const unsigned ka = __fns(kmask, 0, 2 * p + 1), kb = __fns(kmask, 0, 2 * p + 2);
      const float wa = __shfl_sync(0xffffffffu, cur.w, ka & 31), wb = __shfl_sync(0xffffffffu, cur.w, kb & 31);
#pragma unroll
      for (int jj = 0; jj < kGatU; ++jj) {
        if (p == 0) {
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[jj][i] = 0.f;
        }
        const int v = v0 + jj * 32 + lane;
        if (v < vend) {
          if (ka < 32) fma8(acc[jj], r[0][jj], wa);
          if (kb < 32) fma8(acc[jj], r[1][jj], wb);
          if (p == npairs - 1) reinterpret_cast<int4*>(st)[v - vbeg] = pack8_bf16_rn(acc[jj]);
        }
      }
    };
    int4 ra[2][kGatU], rb[2][kGatU];
    load(0, ra);
    if (lane == 0) bulk_wait_read1();  // the copy that last used this stage has finished reading it
    __syncwarp();
    for (int step = 0;;) {
      if (step + 1 < nsteps) load(step + 1, rb);
      accumulate(step, ra);
      if (step + 1 >= nsteps) break;
      if (step + 2 < nsteps) load(step + 2, ra);
      accumulate(step + 1, rb);
      if (step + 2 >= nsteps) break;
      step += 2;
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(a.peers.base[cur.s] + a.slot_off + stage_row_off(a.plan, a.me, cur.j, row_bytes) + (size_t)vbeg * 16, st,
               (uint32_t)(vend - vbeg) * 16);
      bulk_commit();
    }
    cur = nxt;
    it = it_next;
  }
  if (lane == 0) bulk_wait_all();
  cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtComb, a.slot, a.me);
}

// The same push without the register double-buffering: 14 warps per SM instead of 12 and fewer instructions per
// byte.  Used at EP <= 2, where every received row has ~4 local experts and the gather reads ~8x the bytes it pushes,
// i.e. it is bound by HBM bandwidth rather than by the length of a warp's chain of round trips.  (The EP=2 timings of
// the two kernels overlap -- 195-289 us for this one over three runs, 221-226 us pipelined over two: the event-timed
// figure includes the peer's skew -- so the split is by that argument, not by a measured gap.)
__global__ void __launch_bounds__(512, 1) ht_combine_gather_send_simple_kernel(const __grid_constant__ GatherArgs g) {
  extern __shared__ __align__(128) char s_mem[];
  const HtCombineArgs& a = g.c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int W = a.W, K = g.K, row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  const int nparts = g.nparts, part_vec = (nvec + nparts - 1) / nparts, part_bytes = part_vec * 16;
  char* s_stage = s_mem + (size_t)warp * 2 * part_bytes;
  int maxcnt = 0;
  for (int s = 0; s < W; ++s) maxcnt = max(maxcnt, a.meta->rank_prefix[s + 1] - a.meta->rank_prefix[s]);
  const long long total = (long long)maxcnt * W * nparts;
  const char* y = reinterpret_cast<const char*>(a.y);
  int n = 0;  // parts this warp has staged
  for (long long it = (long long)blockIdx.x * nwarp + warp; it < total; it += (long long)gridDim.x * nwarp) {
    const int part = (int)(it % nparts);
    const long long rit = it / nparts;
    const int s = (a.me + 1 + (int)(rit % W)) % W, j = (int)(rit / W);
    const int first = a.meta->rank_prefix[s];
    if (j >= a.meta->rank_prefix[s + 1] - first) continue;  // uniform over the warp
    const int m = first + j;
    int my_dst = -1;
    float my_w = 0.f;
    if (lane < K) {
      const long long id = __ldcg(g.recv_idx + (size_t)m * K + lane);
      if (id >= 0) {
        my_dst = __ldg(g.inv_perm + (size_t)m * K + lane);
        my_w = __ldcg(g.recv_w + (size_t)m * K + lane);
      }
    }
    const unsigned kmask = __ballot_sync(0xffffffffu, my_dst >= 0);
    char* st = s_stage + (size_t)(n & 1) * part_bytes;
    ++n;
    if (lane == 0) bulk_wait_read1();  // the copy that last used this stage has finished reading it
    __syncwarp();
    const int vbeg = part * part_vec, vend = min(nvec, vbeg + part_vec);
    for (int v0 = vbeg; v0 < vend; v0 += 32 * kGatU) {
      float acc[kGatU][8];
      gather_batch(y, row_bytes, kmask, my_dst, my_w, v0, vend, lane, acc);
#pragma unroll
      for (int jj = 0; jj < kGatU; ++jj) {
        const int v = v0 + jj * 32 + lane;
        if (v < vend) reinterpret_cast<int4*>(st)[v - vbeg] = pack8_bf16_rn(acc[jj]);
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(a.peers.base[s] + a.slot_off + stage_row_off(a.plan, a.me, j, row_bytes) + (size_t)vbeg * 16, st,
               (uint32_t)(vend - vbeg) * 16);
      bulk_commit();
    }
  }
  if (lane == 0) bulk_wait_all();
  cta_done_then_signal(a.done, a.epoch_ptr, a.peers, W, kFlagHtComb, a.slot, a.me);
}

// W == 1: nothing travels.  out[t] = bf16(0 + z[row(t)]) (the source-side fp32 accumulate with its single term),
// zeros for tokens that were sent nowhere.  A warp per (token, batch of kGatU x 32 vectors).
__global__ void __launch_bounds__(256) ht_combine_gather_local_kernel(const __grid_constant__ GatherArgs g) {
  const HtCombineArgs& a = g.c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int K = g.K, row_bytes = 2 * a.plan.H, nvec = row_bytes / 16;
  const int nb = (nvec + 32 * kGatU - 1) / (32 * kGatU);
  const long long it = (long long)blockIdx.x * 8 + warp;
  if (it >= (long long)a.T * nb) return;
  const int t = (int)(it / nb), v0 = (int)(it % nb) * 32 * kGatU;
  const int m = __ldg(a.send_off + t);
  int my_dst = -1;
  float my_w = 0.f;
  if (m >= 0 && lane < K) {
    const long long id = __ldcg(g.recv_idx + (size_t)m * K + lane);
    if (id >= 0) {
      my_dst = __ldg(g.inv_perm + (size_t)m * K + lane);
      my_w = __ldcg(g.recv_w + (size_t)m * K + lane);
    }
  }
  const unsigned kmask = __ballot_sync(0xffffffffu, my_dst >= 0);
  float acc[kGatU][8];
  gather_batch(reinterpret_cast<const char*>(a.y), row_bytes, kmask, my_dst, my_w, v0, nvec, lane, acc);
  int4* dst = reinterpret_cast<int4*>(reinterpret_cast<char*>(a.out) + (size_t)t * row_bytes);
#pragma unroll
  for (int jj = 0; jj < kGatU; ++jj) {
    const int v = v0 + jj * 32 + lane;
    if (v >= nvec) continue;
    const int4 z = pack8_bf16_rn(acc[jj]);
    float f[8], o[8];
    unpack8(z, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = __fadd_rn(0.f, f[i]);
    st_na_v4(dst + v, pack8_bf16_rn(o));
  }
}

}  // namespace b200ep

using namespace b200ep;

static int perm_check(b200ep_ctx* c, int slot, const char* who) {
  if (!c || slot < 0 || slot >= kNumSlots) {
    set_error("%s: bad context or slot", who);
    return B200EP_ERR_ARG;
  }
  if (!c->connected) {
    set_error("%s: peers not connected", who);
    return B200EP_ERR_STATE;
  }
  if (c->ht_plan[slot].H == 0) {
    set_error("%s: no high-throughput dispatch has run on slot %d", who, slot);
    return B200EP_ERR_STATE;
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_permute(b200ep_ctx* c, int slot, int alignment, int rows_bound, void* dev_out_x,
                                 float* dev_out_scales, int32_t* dev_expert_ids, int32_t* dev_inv_perm,
                                 int32_t* dev_expert_start, int m_cap, void* stream) {
  int rc = perm_check(c, slot, "ht_permute");
  if (rc) return rc;
  const HtPlan& plan = c->ht_plan[slot];
  const int el = c->ht_e_local[slot], K = plan.K;
  if (alignment < 1 || rows_bound < 0 || rows_bound > plan.cap_rows || m_cap < 0) {
    set_error("ht_permute: bad alignment %d / rows_bound %d (slot holds %d rows) / m_cap %d", alignment, rows_bound,
              plan.cap_rows, m_cap);
    return B200EP_ERR_ARG;
  }
  if ((m_cap > 0 && !dev_expert_ids) || (rows_bound > 0 && (!dev_inv_perm || !dev_out_x || (plan.scale_bytes && !dev_out_scales)))) {
    set_error("ht_permute: null pointer");
    return B200EP_ERR_ARG;
  }
  if ((uintptr_t)dev_out_x & 31) {
    set_error("ht_permute: out_x must be 32-byte aligned");
    return B200EP_ERR_ARG;
  }
  const int grid = rows_bound > 0 ? (rows_bound + kPermTile - 1) / kPermTile : 1;
  if ((size_t)grid * el > kPermPubWords) {
    set_error("ht_permute: %d rows x %d local experts exceed the plan scratch (%zu words)", rows_bound, el, kPermPubWords);
    return B200EP_ERR_CAPACITY;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  char* base = slot_base(c, c->sym, slot);
  PermPlanArgs p;
  p.recv_idx = reinterpret_cast<const long long*>(base + plan.off_idx);
  p.meta = &c->local->ht[slot];
  p.ps = c->perm;
  p.inv_perm = dev_inv_perm;
  p.expert_ids = dev_expert_ids;
  p.expert_start = dev_expert_start;
  p.K = K;
  p.e_local = el;
  p.align = alignment;
  p.m_cap = m_cap;
  const size_t smem = sizeof(int) * ((size_t)el * 8 + (el + 1) + el + el + (size_t)el * 8);
  B200EP_CUDA(ensure_dynamic_smem((const void*)perm_plan_kernel, c->device, (int)smem));
  perm_plan_kernel<<<grid, kPermTile, smem, st>>>(p);
  B200EP_CUDA(cudaGetLastError());
  if (rows_bound > 0) {
    PermScatterArgs s;
    s.x = base + plan.off_x;
    s.scales = plan.scale_bytes ? reinterpret_cast<const float*>(base + plan.off_scales) : nullptr;
    s.inv_perm = dev_inv_perm;
    s.meta = &c->local->ht[slot];
    s.out_x = reinterpret_cast<char*>(dev_out_x);
    s.out_scales = dev_out_scales;
    s.row_bytes = plan.row_bytes;
    s.ng = plan.scale_bytes / 4;
    s.K = K;
    const int nseg = (plan.row_bytes / 32 + 31) / 32;
    const long long items = (long long)rows_bound * nseg;
    perm_scatter_kernel<<<(unsigned)((items + 7) / 8), 256, 0, st>>>(s);
    B200EP_CUDA(cudaGetLastError());
  }
  return B200EP_OK;
}

extern "C" int b200ep_ht_combine_permuted(b200ep_ctx* c, int slot, const void* dev_y_grouped, const int32_t* dev_inv_perm,
                                          const float* dev_topk_weights, void* dev_out,
                                          const uint8_t* dev_is_token_in_rank, const int32_t* dev_send_off, int T, int H,
                                          int phases, void* stream) {
  int rc = perm_check(c, slot, "ht_combine_permuted");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  B200EP_CUDA(cudaSetDevice(c->device));
  GatherArgs g;
  rc = ht_combine_args(c, slot, dev_y_grouped, dev_out, dev_is_token_in_rank, dev_send_off, T, H, &g.c);
  if (rc) return rc;
  const HtPlan& plan = c->ht_plan[slot];
  char* base = slot_base(c, c->sym, slot);
  g.inv_perm = dev_inv_perm;
  g.recv_idx = reinterpret_cast<const long long*>(base + plan.off_idx);
  g.recv_w = dev_topk_weights ? dev_topk_weights : reinterpret_cast<const float*>(base + plan.off_w);
  g.K = plan.K;
  g.nparts = 1;
  g.c.self_staged = 1;
  if (T > 0 && (phases & B200EP_PHASE_RECV) && (!dev_out || !dev_send_off || (c->world > 1 && !dev_is_token_in_rank))) {
    set_error("ht_combine_permuted: null pointer");
    return B200EP_ERR_ARG;
  }
  if ((phases & B200EP_PHASE_SEND) && ((uintptr_t)dev_y_grouped & 15)) {
    set_error("ht_combine_permuted: y must be 16-byte aligned");
    return B200EP_ERR_ARG;
  }
  const int row_bytes = 2 * H;
  if (c->world == 1) {
    if (phases & B200EP_PHASE_RECV) {
      if (T > 0 && (!dev_y_grouped || !dev_inv_perm)) {
        set_error("ht_combine_permuted: null pointer");
        return B200EP_ERR_ARG;
      }
      const int nb = (row_bytes / 16 + 32 * kGatU - 1) / (32 * kGatU);
      const long long items = (long long)T * nb;
      ht_combine_gather_local_kernel<<<(unsigned)((items + 7) / 8 > 0 ? (items + 7) / 8 : 1), 256, 0, st>>>(g);
      B200EP_CUDA(cudaGetLastError());
    }
    return B200EP_OK;
  }
  if (phases & B200EP_PHASE_SEND) {
    const int nvec = row_bytes / 16;
    g.nparts = (row_bytes + 8191) / 8192;  // parts of at most 8 KB
    const int part_vec = (nvec + g.nparts - 1) / g.nparts;
    g.nparts = (nvec + part_vec - 1) / part_vec;  // no empty last part
    const int part_bytes = part_vec * 16;
    const bool pipelined = c->world >= 4;
    int warps = 200 * 1024 / (2 * part_bytes);
    if (warps > (pipelined ? 12 : 16)) warps = pipelined ? 12 : 16;  // the pipelined kernel double-buffers its loads in registers
    const size_t smem = (size_t)warps * 2 * part_bytes;
    const int grid = c->sm_limit > 0 && c->sm_limit < c->num_sms ? c->sm_limit : c->num_sms;
    if (pipelined) {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_gather_send_kernel, c->device, (int)smem));
      ht_combine_gather_send_kernel<<<grid, warps * 32, smem, st>>>(g);
    } else {
      B200EP_CUDA(ensure_dynamic_smem((const void*)ht_combine_gather_send_simple_kernel, c->device, (int)smem));
      ht_combine_gather_send_simple_kernel<<<grid, warps * 32, smem, st>>>(g);
    }
    B200EP_CUDA(cudaGetLastError());
  }
  if (phases & B200EP_PHASE_RECV) return ht_combine_launch_recv(c, g.c, st);
  return B200EP_OK;
}

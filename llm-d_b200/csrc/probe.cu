// Fabric probe: what does NVLink give SM-driven copies between the symmetric regions of the ranks?
// Measurement aid (bench.py --impl probe), built into its own library (libb200ep_probe.so) so that the
// product library carries no benchmark code.  Every mode moves `bytes_per_peer` between this rank and
// each of its W-1 peers, cut into pieces of `piece_bytes`; piece i goes to / comes from peer
// (me + 1 + i % (W-1)) % W, so that all ranks move along a permutation at any moment (the pattern of the
// combine push).  Modes:
//   0  push, TMA:        global(local) -> shared (cp.async.bulk + mbarrier) -> global(peer) (cp.async.bulk)
//   1  push, registers:  ld.global.v4 (local) -> st.global.v4 (peer), 8 x 16 B in flight per lane
//   2  pull, registers:  ld.global.v4 (peer)  -> st.global.v4 (local)
//   3  pull, TMA:        global(peer) -> shared -> global(local)
//   4  push, copy engines: one cudaMemcpyAsync per peer
//   5  push, TMA, RANDOM single destination per piece (hash of the piece index) instead of the rotation
//   6  push, TMA, each piece stored to a random SUBSET of the peers (each with probability 2/3) from one
//      shared-memory stage -- the fan-out of the dispatch kernel
//   7  as 6, plus three small bulk copies (224 + 64 + 32 bytes) per destination -- scales and top-k rows
// Region layout (caller allocates >= 2 * (W * bytes_per_peer) per rank): first half = source blocks, one
// per peer, second half = landing blocks, one per peer.
#include "common.cuh"

namespace b200ep {

constexpr int kProbeStages = 4;
constexpr int kProbeAhead = 2;

struct ProbeArgs {
  Peers peers;
  size_t bytes_per_peer, half;  // half = W * bytes_per_peer
  int piece_bytes, W, me, mode;
};

__device__ __forceinline__ void probe_addr(const ProbeArgs& a, long long i, const char** src, char** dst, unsigned* bytes) {
  const int W = a.W;
  const int k = (int)(i % (W - 1));
  const long long j = i / (W - 1);
  const int peer = (a.me + 1 + k) % W;
  const size_t off = (size_t)j * a.piece_bytes;
  *bytes = (unsigned)min((size_t)a.piece_bytes, a.bytes_per_peer - off);
  if (a.mode == 0 || a.mode == 1) {  // push: my source block for `peer` -> peer's landing block for me
    *src = a.peers.base[a.me] + (size_t)peer * a.bytes_per_peer + off;
    *dst = a.peers.base[peer] + a.half + (size_t)a.me * a.bytes_per_peer + off;
  } else {  // pull: peer's source block for me -> my landing block for `peer`
    *src = a.peers.base[peer] + (size_t)a.me * a.bytes_per_peer + off;
    *dst = a.peers.base[a.me] + a.half + (size_t)peer * a.bytes_per_peer + off;
  }
}

__device__ __forceinline__ unsigned probe_hash(unsigned x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

// modes 5-7: pieces come from this rank's source half in order; destinations are pseudo-random
__global__ void __launch_bounds__(128, 1) probe_fanout_kernel(const __grid_constant__ ProbeArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ __align__(8) uint64_t s_bar[kProbeStages];
  if (threadIdx.x == 0) {
    for (int i = 0; i < kProbeStages; ++i) mbar_init(&s_bar[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int W = a.W;
  // same total bytes leave the GPU as in the rotation modes: (W-1) * bytes_per_peer
  const long long per_peer = (long long)(a.bytes_per_peer / a.piece_bytes);
  const long long total = a.mode == 5 ? per_peer * (W - 1) : per_peer * (W - 1) * 3 / (2 * (W - 1)) + 1;  // mode 6/7: 2/3 of W-1 copies per piece
  long long idx[kProbeStages];
  int loads = 0, stores = 0;
  long long i = blockIdx.x;
  bool more = true;
  for (;;) {
    if (more && loads - stores < kProbeAhead) {
      if (i < total) {
        const int st = loads % kProbeStages;
        if (loads >= kProbeStages) bulk_wait_read<kProbeStages - kProbeAhead - 1>();
        const char* src = a.peers.base[a.me] + (size_t)((i * a.piece_bytes) % (long long)a.half);
        idx[st] = i;
        mbar_expect_tx(&s_bar[st], a.piece_bytes);
        bulk_g2s(s_mem + (size_t)st * a.piece_bytes, src, a.piece_bytes, &s_bar[st]);
        i += gridDim.x;
        ++loads;
        continue;
      }
      more = false;
    }
    if (stores < loads) {
      const int st = stores % kProbeStages;
      mbar_wait(&s_bar[st], (unsigned)(stores / kProbeStages) & 1u);
      const long long pi = idx[st];
      const unsigned h = probe_hash((unsigned)pi * 2654435761u + (unsigned)a.me * 97u);
      const size_t off = (size_t)((pi * a.piece_bytes) % (long long)(a.bytes_per_peer - a.piece_bytes + 16)) & ~(size_t)15;
      const char* stage = s_mem + (size_t)st * a.piece_bytes;
      if (a.mode == 5) {
        const int peer = (a.me + 1 + (int)(h % (unsigned)(W - 1))) % W;
        bulk_s2g(a.peers.base[peer] + a.half + (size_t)a.me * a.bytes_per_peer + off, stage, a.piece_bytes);
      } else {
        for (int k = 0; k < W - 1; ++k) {
          if (probe_hash(h + k * 0x9e3779b9u) % 3u == 0u) continue;  // each peer with probability 2/3
          const int peer = (a.me + 1 + k) % W;
          char* dst = a.peers.base[peer] + a.half + (size_t)a.me * a.bytes_per_peer + off;
          bulk_s2g(dst, stage, a.piece_bytes);
          if (a.mode == 7) {
            bulk_s2g(dst + 256, stage, 224);
            bulk_s2g(dst + 512, stage, 64);
            bulk_s2g(dst + 640, stage, 32);
          }
        }
      }
      bulk_commit();
      ++stores;
    } else if (!more) {
      break;
    }
  }
  bulk_wait_all();
}

__global__ void __launch_bounds__(128, 1) probe_tma_kernel(const __grid_constant__ ProbeArgs a) {
  extern __shared__ __align__(128) char s_mem[];
  __shared__ __align__(8) uint64_t s_bar[kProbeStages];
  if (threadIdx.x == 0) {
    for (int i = 0; i < kProbeStages; ++i) mbar_init(&s_bar[i], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const long long per_peer = (long long)((a.bytes_per_peer + a.piece_bytes - 1) / a.piece_bytes);
  const long long total = per_peer * (a.W - 1);
  char* dsts[kProbeStages];
  unsigned sizes[kProbeStages];
  int loads = 0, stores = 0;
  long long i = blockIdx.x;
  bool more = true;
  for (;;) {
    if (more && loads - stores < kProbeAhead) {
      if (i < total) {
        const char* src;
        char* dst;
        unsigned bytes;
        probe_addr(a, i, &src, &dst, &bytes);
        i += gridDim.x;
        const int st = loads % kProbeStages;
        if (loads >= kProbeStages) bulk_wait_read<kProbeStages - kProbeAhead - 1>();
        dsts[st] = dst;
        sizes[st] = bytes;
        mbar_expect_tx(&s_bar[st], bytes);
        bulk_g2s(s_mem + (size_t)st * a.piece_bytes, src, bytes, &s_bar[st]);
        ++loads;
        continue;
      }
      more = false;
    }
    if (stores < loads) {
      const int st = stores % kProbeStages;
      mbar_wait(&s_bar[st], (unsigned)(stores / kProbeStages) & 1u);
      bulk_s2g(dsts[st], s_mem + (size_t)st * a.piece_bytes, sizes[st]);
      bulk_commit();
      ++stores;
    } else if (!more) {
      break;
    }
  }
  bulk_wait_all();
}

// one warp per piece; lanes copy 16 B each, 8 vectors in flight per lane
__global__ void __launch_bounds__(512, 2) probe_reg_kernel(const __grid_constant__ ProbeArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const long long per_peer = (long long)((a.bytes_per_peer + a.piece_bytes - 1) / a.piece_bytes);
  const long long total = per_peer * (a.W - 1);
  for (long long i = (long long)blockIdx.x * nwarp + warp; i < total; i += (long long)gridDim.x * nwarp) {
    const char* src;
    char* dst;
    unsigned bytes;
    probe_addr(a, i, &src, &dst, &bytes);
    const int nvec = (int)(bytes / 16);
    for (int v0 = 0; v0 < nvec; v0 += 32 * 8) {
      int4 r[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) r[u] = ld_cg_v4(src + (size_t)v * 16);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) st_na_v4(dst + (size_t)v * 16, r[u]);
      }
    }
  }
}

}  // namespace b200ep

using namespace b200ep;

extern "C" int b200ep_probe_run(void* const* bases, int world, int me, int mode, size_t bytes_per_peer, int piece_bytes,
                                int grid, void* stream) {
  if (world < 2 || world > kMaxRanks || me < 0 || me >= world || piece_bytes <= 0 || piece_bytes % 16 || bytes_per_peer % 16)
    return B200EP_ERR_ARG;
  ProbeArgs a;
  for (int r = 0; r < world; ++r) a.peers.base[r] = (char*)bases[r];
  a.bytes_per_peer = bytes_per_peer;
  a.half = (size_t)world * bytes_per_peer;
  a.piece_bytes = piece_bytes;
  a.W = world;
  a.me = me;
  a.mode = mode;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0 || mode == 3) {
    const size_t smem = (size_t)kProbeStages * piece_bytes;
    if (smem > 200 * 1024) return B200EP_ERR_ARG;
    if (cudaFuncSetAttribute(probe_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return B200EP_ERR_CUDA;
    probe_tma_kernel<<<grid, 128, smem, st>>>(a);
  } else if (mode >= 5 && mode <= 7) {
    const size_t smem = (size_t)kProbeStages * piece_bytes;
    if (smem > 200 * 1024 || (size_t)piece_bytes * 2 > bytes_per_peer) return B200EP_ERR_ARG;
    if (cudaFuncSetAttribute(probe_fanout_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return B200EP_ERR_CUDA;
    probe_fanout_kernel<<<grid, 128, smem, st>>>(a);
  } else if (mode == 1 || mode == 2) {
    probe_reg_kernel<<<grid, 512, 0, st>>>(a);
  } else if (mode == 4) {
    // copy engines: one cudaMemcpyAsync per peer (push), each on a stream of its own, forked from / joined to `st`
    static cudaStream_t streams[kMaxRanks];
    static cudaEvent_t fork_ev, join_ev[kMaxRanks];
    static bool ready = false;
    if (!ready) {
      for (int k = 0; k < kMaxRanks; ++k) {
        if (cudaStreamCreateWithFlags(&streams[k], cudaStreamNonBlocking) != cudaSuccess) return B200EP_ERR_CUDA;
        if (cudaEventCreateWithFlags(&join_ev[k], cudaEventDisableTiming) != cudaSuccess) return B200EP_ERR_CUDA;
      }
      if (cudaEventCreateWithFlags(&fork_ev, cudaEventDisableTiming) != cudaSuccess) return B200EP_ERR_CUDA;
      ready = true;
    }
    cudaEventRecord(fork_ev, st);
    for (int k = 0; k < world - 1; ++k) {
      const int peer = (me + 1 + k) % world;
      cudaStreamWaitEvent(streams[k], fork_ev, 0);
      cudaMemcpyAsync(a.peers.base[peer] + a.half + (size_t)me * bytes_per_peer, a.peers.base[me] + (size_t)peer * bytes_per_peer,
                      bytes_per_peer, cudaMemcpyDeviceToDevice, streams[k]);
      cudaEventRecord(join_ev[k], streams[k]);
      cudaStreamWaitEvent(st, join_ev[k], 0);
    }
  } else {
    return B200EP_ERR_ARG;
  }
  return cudaGetLastError() == cudaSuccess ? B200EP_OK : B200EP_ERR_CUDA;
}

// Context, symmetric region, CUDA-IPC bootstrap and region planning for libb200ep.
//
// Replaces the runtime half of DeepEP's deep_ep_cpp.Buffer (buffer allocation,
// IPC handle exchange, destroy) that vLLM drives from
// vllm/distributed/device_communicators/all2all.py:190-194,222-245,297-315.
// No NVSHMEM: one cudaMalloc'd region per rank, mapped into every peer with
// cudaIpcOpenMemHandle; all later traffic is SM-issued stores over NVLink.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <chrono>
#include <new>
#include <thread>

#include "common.cuh"

namespace b200ep {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

cudaError_t ensure_dynamic_smem(const void* kernel, int device, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> have;
  std::lock_guard<std::mutex> lock(mu);
  int& cur = have[{kernel, device}];
  if (bytes <= cur) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) cur = bytes;
  return e;
}

// One slot of the high-throughput mode:
//   receive side  (written by peers' dispatch): x rows, scales, top-k ids, top-k weights, source index
//   combine stage (written by peers' combine) : [t_cap][world] bf16 rows
// sized for the worst case (every token of every rank lands here).
int make_ht_plan(const b200ep_ctx* c, int H, int K, int x_kind, HtPlan* p) {
  if (H <= 0 || H % kGroup != 0 || K <= 0 || K > kMaxTopk) {
    set_error("ht: unsupported shape H=%d (must be a positive multiple of 128) K=%d (1..%d)", H, K, kMaxTopk);
    return B200EP_ERR_ARG;
  }
  if (x_kind < B200EP_X_BF16 || x_kind > B200EP_X_FP8) {
    set_error("ht: bad x_kind %d", x_kind);
    return B200EP_ERR_ARG;
  }
  const int W = c->world;
  p->H = H;
  p->K = K;
  p->x_kind = x_kind;
  p->row_bytes = x_kind == B200EP_X_BF16 ? 2 * H : H;
  p->scale_bytes = x_kind == B200EP_X_BF16 ? 0 : H / kGroup * 4;
  p->side_bytes = (K * 12 + 4 + p->scale_bytes + 15) / 16 * 16;
  const size_t per_recv_row = (size_t)p->row_bytes + p->scale_bytes + (size_t)K * 12 + 4 + p->side_bytes;
  p->qrow_bytes = (p->row_bytes + p->scale_bytes + 15) / 16 * 16;
  const size_t per_token = (size_t)W * per_recv_row + (size_t)W * 2 * H + (size_t)p->qrow_bytes + (size_t)W * 4;
  const size_t slack = 13 * 4096;
  if (c->slot_bytes <= slack) {
    set_error("ht: symmetric region too small");
    return B200EP_ERR_CAPACITY;
  }
  long long t_cap = (long long)((c->slot_bytes - slack) / per_token);
  t_cap = t_cap / 8 * 8;
  if (t_cap > (1 << 20)) t_cap = 1 << 20;
  p->t_cap = (int)t_cap;
  p->cap_rows = W * p->t_cap;
  size_t off = 0;
  p->off_x = off;
  off = align_up(off + (size_t)p->cap_rows * p->row_bytes, 4096);
  p->off_scales = off;
  off = align_up(off + (size_t)p->cap_rows * p->scale_bytes, 4096);
  p->off_idx = off;
  off = align_up(off + (size_t)p->cap_rows * K * 8, 4096);
  p->off_w = off;
  off = align_up(off + (size_t)p->cap_rows * K * 4, 4096);
  p->off_src = off;
  off = align_up(off + (size_t)p->cap_rows * 4, 4096);
  p->off_side = off;
  off = align_up(off + (size_t)p->cap_rows * p->side_bytes, 4096);
  p->off_stage = off;
  off = align_up(off + (size_t)p->t_cap * W * 2 * H, 4096);
  p->off_qstage = off;
  off = align_up(off + (size_t)p->t_cap * p->qrow_bytes, 4096);
  p->off_sendlist = off;
  off = align_up(off + (size_t)p->t_cap * W * 4, 4096);
  if (off > c->slot_bytes) {
    set_error("ht: internal planning error (%zu > %zu)", off, c->slot_bytes);
    return B200EP_ERR_CAPACITY;
  }
  return B200EP_OK;
}

static size_t ll_slot_bytes(int max_tok, int H, int W, int E, int use_fp8, int K) {
  const int e_local = E / W;
  const size_t cap = (size_t)W * max_tok;
  const size_t row = use_fp8 ? H : 2 * (size_t)H;
  const size_t sc = use_fp8 ? (size_t)H / kGroup * 4 : 0;
  size_t off = 0;
  off = align_up(off + (size_t)e_local * cap * row, 4096);
  off = align_up(off + (size_t)e_local * (cap + 3) * (sc ? sc : 0), 4096);  // fp32 scales, or packed UE8M0 (never larger)
  off = align_up(off + (size_t)e_local * cap * 4, 4096);
  off = align_up(off + (size_t)max_tok * K * 2 * H, 4096);
  return off;
}

int make_ll_plan(const b200ep_ctx* c, int H, int K, int E, int max_tok, int use_fp8, LlPlan* p) {
  const int W = c->world;
  if (H <= 0 || H % kGroup != 0 || K <= 0 || K > kMaxTopk || E <= 0 || E % W != 0 || E / W > kMaxLocalExperts ||
      max_tok <= 0) {
    set_error("ll: unsupported shape H=%d K=%d E=%d world=%d max_tok=%d", H, K, E, W, max_tok);
    return B200EP_ERR_ARG;
  }
  p->H = H;
  p->K = K;
  p->max_tok = max_tok;
  p->use_fp8 = use_fp8;
  p->e_local = E / W;
  p->cap = W * max_tok;
  p->row_bytes = use_fp8 ? H : 2 * H;
  p->scale_bytes = use_fp8 ? H / kGroup * 4 : 0;
  size_t off = 0;
  p->off_x = off;
  off = align_up(off + (size_t)p->e_local * p->cap * p->row_bytes, 4096);
  p->g4 = (H / kGroup + 3) / 4;
  p->cap4 = (p->cap + 3) / 4 * 4;
  p->off_scales = off;
  off = align_up(off + (size_t)p->e_local * (p->cap + 3) * p->scale_bytes, 4096);  // fp32 rows, or packed UE8M0 (never larger)
  p->off_src = off;
  off = align_up(off + (size_t)p->e_local * p->cap * 4, 4096);
  p->off_stage = off;
  off = align_up(off + (size_t)max_tok * K * 2 * H, 4096);
  if (off > c->slot_bytes) {
    set_error("ll: symmetric region too small: slot needs %zu bytes, has %zu (size with b200ep_ll_size_hint)", off,
              c->slot_bytes);
    return B200EP_ERR_CAPACITY;
  }
  return B200EP_OK;
}

}  // namespace b200ep

using namespace b200ep;

extern "C" {

const char* b200ep_last_error(void) { return g_err; }
int b200ep_version(void) { return 100; }

size_t b200ep_ll_size_hint(int max_tokens_per_rank, int H, int world, int E) {
  if (world <= 0 || E <= 0 || E % world) return 0;
  // bf16 rows and the largest top-k the library accepts: an upper bound for every mode
  return kCtrlBytes + kNumSlots * (ll_slot_bytes(max_tokens_per_rank, H, world, E, 0, kMaxTopk) + 4096);
}

static void free_ctx(b200ep_ctx* c, cudaError_t* first_err) {
  auto note = [&](cudaError_t e) {
    if (e != cudaSuccess && first_err && *first_err == cudaSuccess) *first_err = e;
  };
  for (int r = 0; r < c->world; ++r)
    if (c->ipc_open[r]) note(cudaIpcCloseMemHandle(c->peers.base[r]));
  if (c->sym) note(cudaFree(c->sym));
  if (c->local) note(cudaFree(c->local));
  if (c->perm) note(cudaFree(c->perm));
  if (c->ll_local) note(cudaFree(c->ll_local));
  if (c->host) note(cudaFreeHost((void*)c->host));
  delete c;
}

int b200ep_create(b200ep_ctx** out, int rank, int world, int device, size_t sym_bytes) {
  if (out) *out = nullptr;
  if (!out || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) {
    set_error("create: bad rank/world %d/%d (world <= %d)", rank, world, kMaxRanks);
    return B200EP_ERR_ARG;
  }
  if (sym_bytes < kCtrlBytes + (size_t)kNumSlots * (1 << 20)) {
    set_error("create: sym_bytes %zu too small", sym_bytes);
    return B200EP_ERR_ARG;
  }
  B200EP_CUDA(cudaSetDevice(device));
  b200ep_ctx* c = new (std::nothrow) b200ep_ctx();
  if (!c) {
    set_error("create: out of host memory");
    return B200EP_ERR_ARG;
  }
  memset(c, 0, sizeof(*c));
  c->rank = rank;
  c->world = world;
  c->device = device;
  // any failure below frees what was allocated so far (a multi-GB cudaMalloc failing on a GPU whose
  // memory already belongs to a KV-cache pool is the likely case) and leaves *out null
#define CREATE_CUDA(call)                                                                            \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) {                                                                         \
      set_error("create: %s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      free_ctx(c, nullptr);                                                                          \
      (void)cudaGetLastError();                                                                      \
      return B200EP_ERR_CUDA;                                                                        \
    }                                                                                                \
  } while (0)
  CREATE_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, device));
  c->sm_limit = 0;
  c->sym_bytes = sym_bytes / 4096 * 4096;
  c->slot_bytes = (c->sym_bytes - kCtrlBytes) / kNumSlots / 4096 * 4096;
  CREATE_CUDA(cudaMalloc(&c->sym, c->sym_bytes));
  CREATE_CUDA(cudaMemset(c->sym, 0, kCtrlBytes));
  CREATE_CUDA(cudaMalloc(&c->local, sizeof(LocalScratch)));
  CREATE_CUDA(cudaMemset(c->local, 0, sizeof(LocalScratch)));
  CREATE_CUDA(cudaMalloc(&c->perm, sizeof(PermScratch)));
  CREATE_CUDA(cudaMemset(c->perm, 0, sizeof(PermScratch)));
  c->ll_local_bytes = 4 << 20;
  CREATE_CUDA(cudaMalloc(&c->ll_local, c->ll_local_bytes));
  CREATE_CUDA(cudaMemset(c->ll_local, 0, c->ll_local_bytes));
  CREATE_CUDA(cudaHostAlloc(&c->host, sizeof(HostMirror), cudaHostAllocMapped));
  memset((void*)c->host, 0, sizeof(HostMirror));
  CREATE_CUDA(cudaHostGetDevicePointer((void**)&c->host_dev, (void*)c->host, 0));
  CREATE_CUDA(cudaDeviceSynchronize());
#undef CREATE_CUDA
  c->peers.base[rank] = c->sym;
  c->connected = (world == 1);
  *out = c;
  return B200EP_OK;
}

int b200ep_destroy(b200ep_ctx* c) {
  if (!c) return B200EP_OK;
  cudaError_t first = cudaSetDevice(c->device);
  cudaError_t e = cudaDeviceSynchronize();  // surfaces a kernel that trapped (spin-wait timeout) instead of hiding it
  if (first == cudaSuccess) first = e;
  free_ctx(c, &first);
  if (first != cudaSuccess) {
    set_error("destroy: %s (everything that could be released was released)", cudaGetErrorString(first));
    return B200EP_ERR_CUDA;
  }
  return B200EP_OK;
}

int b200ep_export_handle(b200ep_ctx* c, void* out_handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == B200EP_IPC_HANDLE_BYTES, "handle size");
  if (!c || !out_handle64) {
    set_error("export_handle: null argument");
    return B200EP_ERR_ARG;
  }
  B200EP_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  B200EP_CUDA(cudaIpcGetMemHandle(&h, c->sym));
  memcpy(out_handle64, &h, sizeof(h));
  return B200EP_OK;
}

int b200ep_import_handles(b200ep_ctx* c, const void* handles, int n) {
  if (!c || !handles || n != c->world) {
    set_error("import_handles: need exactly world=%d handles", c ? c->world : -1);
    return B200EP_ERR_ARG;
  }
  B200EP_CUDA(cudaSetDevice(c->device));
  for (int r = 0; r < n; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + (size_t)r * sizeof(h), sizeof(h));
    void* p = nullptr;
    B200EP_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peers.base[r] = (char*)p;
    c->ipc_open[r] = true;
  }
  c->connected = true;
  return B200EP_OK;
}

int b200ep_set_peer_ptrs(b200ep_ctx* c, void* const* dev_bases, int n) {
  if (!c || !dev_bases || n != c->world) {
    set_error("set_peer_ptrs: need exactly world=%d bases", c ? c->world : -1);
    return B200EP_ERR_ARG;
  }
  for (int r = 0; r < n; ++r)
    if (r != c->rank) c->peers.base[r] = (char*)dev_bases[r];
  c->connected = true;
  return B200EP_OK;
}

int b200ep_sym_base(b200ep_ctx* c, void** dev_base, size_t* bytes) {
  if (!c) {
    set_error("sym_base: null context");
    return B200EP_ERR_ARG;
  }
  if (dev_base) *dev_base = c->sym;
  if (bytes) *bytes = c->sym_bytes;
  return B200EP_OK;
}

int b200ep_peer_base(b200ep_ctx* c, int rank, void** dev_base) {
  if (!c || !dev_base || rank < 0 || rank >= c->world) {
    set_error("peer_base: bad argument");
    return B200EP_ERR_ARG;
  }
  if (!c->connected) {
    set_error("peer_base: peers not connected");
    return B200EP_ERR_STATE;
  }
  *dev_base = c->peers.base[rank];
  return B200EP_OK;
}

int b200ep_debug_trace(b200ep_ctx* c, unsigned long long* out128) {
  if (!c || !out128) {
    set_error("debug_trace: bad argument");
    return B200EP_ERR_ARG;
  }
  B200EP_CUDA(cudaSetDevice(c->device));
  B200EP_CUDA(cudaMemcpy(out128, c->local->trace, sizeof(c->local->trace), cudaMemcpyDeviceToHost));
  return B200EP_OK;
}

int b200ep_set_num_sms(b200ep_ctx* c, int num_sms) {
  if (!c || num_sms < 0) {
    set_error("set_num_sms: bad argument");
    return B200EP_ERR_ARG;
  }
  c->sm_limit = num_sms;
  return B200EP_OK;
}

int b200ep_ht_capacity(b200ep_ctx* c, int H, int K, int x_kind) {
  HtPlan p;
  if (!c || make_ht_plan(c, H, K, x_kind, &p) != B200EP_OK) return -1;
  return p.t_cap;
}

int b200ep_ht_notify_wait(b200ep_ctx* c, int slot, int* num_recv_tokens, int32_t* per_expert, int32_t* rank_prefix,
                          int timeout_ms) {
  if (!c || slot < 0 || slot >= kNumSlots) {
    set_error("ht_notify_wait: bad slot");
    return B200EP_ERR_ARG;
  }
  const uint64_t want = c->epoch[kFlagHtNotify][slot] & 0xffffffffull;
  const int el = c->ht_e_local[slot];
  volatile unsigned long long* hw = c->host->word[slot];
  auto fresh = [&](int i) { return (hw[i] >> 32) == want; };
  auto all_fresh = [&]() {
    for (int i = 0; i <= c->world + 1; ++i)
      if (!fresh(i)) return false;
    for (int i = 0; i < el; ++i)
      if (!fresh(1 + kMaxRanks + 1 + i)) return false;
    return true;
  };
  // The answer normally arrives within microseconds (the kernels are already queued): spin on the mapped words
  // without giving up the core; look at the clock, the CUDA error state and the scheduler only every 4096 polls.
  auto t0 = std::chrono::steady_clock::now();
  for (unsigned spins = 1; !all_fresh(); ++spins) {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
    if ((spins & 4095u) != 0) continue;
    const long long ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    if (timeout_ms > 0 && ms > timeout_ms) {
      set_error("ht_notify_wait: rank %d timed out after %d ms waiting for the count exchange (epoch %llu)", c->rank,
                timeout_ms, (unsigned long long)want);
      return B200EP_ERR_TIMEOUT;
    }
    // also surface a dead stream instead of spinning on it
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
      set_error("ht_notify_wait: CUDA error while waiting: %s", cudaGetErrorString(e));
      return B200EP_ERR_CUDA;
    }
    if (ms >= 2) std::this_thread::yield();  // something is slow (a peer is late): stop hogging the core
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  c->host->num_recv_cache[slot] = (int32_t)(uint32_t)hw[0];
  if (num_recv_tokens) *num_recv_tokens = (int32_t)(uint32_t)hw[0];
  if (rank_prefix)
    for (int i = 0; i <= c->world; ++i) rank_prefix[i] = (int32_t)(uint32_t)hw[1 + i];
  if (per_expert)
    for (int i = 0; i < el; ++i) per_expert[i] = (int32_t)(uint32_t)hw[1 + kMaxRanks + 1 + i];
  return B200EP_OK;
}

}  // extern "C"

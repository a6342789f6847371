// Stand-alone per-token-group FP8 quantiser (group 128), same device functions as
// the fused dispatch kernels.  Parity target: vLLM per_token_group_quant_fp8
// (vllm/model_executor/layers/quantization/utils/fp8_utils.py:510-591) which runs
// the CUDA op per_token_group_quant_8bit on GPUs: s = max(amax,1e-10)/448 with an
// IEEE division, q = sat_e4m3(x / s); use_ue8m0 rounds s up to a power of two.
// CPU restatement: oracle/fp8.py formula "div".
// HBM traffic per token: 2H in, H + 4H/128 out.
#include "common.cuh"

namespace b200ep {

__global__ void __launch_bounds__(256)
    quant_fp8_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ s, int T, int H,
                     int ue8m0) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int nvec = H / 8, ng = H / kGroup;
  constexpr int U = 4;
  for (int t = blockIdx.x * nwarp + warp; t < T; t += gridDim.x * nwarp) {
    const int4* src = reinterpret_cast<const int4*>(x + (size_t)t * H);
    for (int v0 = 0; v0 < nvec; v0 += 32 * U) {
      int4 raw[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * 32 + lane;
        raw[u] = v < nvec ? ld_nc_v4(src + v) : make_int4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * 32 + lane;
        float f[8];
        unpack8(raw[u], f);
        float m = absmax8(f);
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        const float sc = group_scale(m, ue8m0 != 0);
        if (v < nvec) {
          *reinterpret_cast<uint2*>(q + (size_t)t * H + (size_t)v * 8) = quant8(f, sc);
          if ((lane & 15) == 0) s[(size_t)t * ng + (v >> 4)] = sc;
        }
      }
    }
  }
}

}  // namespace b200ep

using namespace b200ep;

extern "C" int b200ep_quant_fp8(const void* dev_x_bf16, void* dev_q, float* dev_scales, int T, int H, int ue8m0,
                                void* stream) {
  if (T < 0 || H <= 0 || H % kGroup) {
    set_error("quant_fp8: H=%d must be a positive multiple of 128", H);
    return B200EP_ERR_ARG;
  }
  if (T == 0) return B200EP_OK;
  if (!dev_x_bf16 || !dev_q || !dev_scales) {
    set_error("quant_fp8: null pointer");
    return B200EP_ERR_ARG;
  }
  int dev = 0, sms = 148;
  B200EP_CUDA(cudaGetDevice(&dev));
  B200EP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  int grid = (T + 7) / 8;
  if (grid > sms * 8) grid = sms * 8;
  quant_fp8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dev_x_bf16, (uint8_t*)dev_q,
                                                            dev_scales, T, H, ue8m0);
  B200EP_CUDA(cudaGetLastError());
  return B200EP_OK;
}

// DAC 44.1 kHz decoder (zonos/autoencoder.py:44-47 -> transformers DacModel.decode, modeling_dac.py:608-638).
//
// Every convolution is an implicit GEMM on tcgen05 (dac_conv.cuh) over channels-last fp16 activations
// [batch][time][channels] (pitch == channels: no pad columns in memory, the TMA unit zero-fills up to the 64-channel
// k-block):  out[t, co] = sum_tap sum_ci x[t + shift(tap), ci] * W[co][tap][ci]
//   * Conv1d k=7, dilation d : 7 taps, shift = (tap-3)*d, zero padding outside [0,T)     (modeling_dac.py:181-186,417,426)
//   * Conv1d k=1             : 1 tap
//   * ConvTranspose1d k=2s, stride s, padding s/2 (:243-249): each output phase r has exactly two taps, so it is a
//     2-tap conv with N = s*C_out columns (n = r*C_out + co) over rows j = 0..T_in, scattered to t = s*j + r - s/2.
// The A operand of tap `tap` is the same activation halo tile, read through a row-shifted UMMA descriptor.
// Fused epilogue: + bias, (+ residual), store raw fp16 and/or Snake(alpha of the CONSUMER layer) fp16, so Snake1d
// (:94-99) never makes its own pass over memory.  fp16 storage / fp32 accumulate mirrors the reference's CUDA autocast.
#include "../../include/zonos_b200.h"
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "dac_conv.cuh"

namespace zb {

// ------------------------------------------------------------------------------------------------ from_codes
// z[b,t,:] = sum_k table[k][codes[b,k,t]][:]  with table[k][c] = out_proj_k(codebook_k[c]) + bias_k precomputed in fp32
// at load time (modeling_dac.py:345-369: embedding -> 1x1 conv, summed over the 9 codebooks in order).
__global__ void dac_from_codes_kernel(const int64_t* __restrict__ codes, const float* __restrict__ table,
                                      f16* __restrict__ z, int T, int n_q, int n_codes, int latent) {
  const int t = blockIdx.x, b = blockIdx.y;
  __shared__ int sc[ZB200_MAX_CODEBOOKS];
  if (threadIdx.x < n_q) {
    const long long c = (long long)codes[((size_t)b * n_q + threadIdx.x) * T + t];
    sc[threadIdx.x] = (int)(c < 0 ? 0 : (c >= n_codes ? n_codes - 1 : c));
  }
  __syncthreads();
  for (int c = threadIdx.x; c < latent; c += blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < n_q; ++k) acc += __ldg(table + ((size_t)k * n_codes + sc[k]) * latent + c);
    z[((size_t)b * T + t) * latent + c] = __float2half_rn(acc);
  }
}

// ------------------------------------------------------------------------------------------------ last conv + tanh
// wav[b,t] = tanh( sum_{tap<7, ci<C} x[b, t+tap-3, ci] * w[tap][ci] + bias )   (modeling_dac.py:426-427,436-437)
constexpr int FIN_TILE = 256;
__global__ void __launch_bounds__(FIN_TILE) dac_final_conv_kernel(const f16* __restrict__ x, const float* __restrict__ w,
                                                                  float bias, float* __restrict__ wav, int T, int C,
                                                                  int pitch) {
  extern __shared__ uint8_t fin_smem[];
  const int sp = C / 2 + 1;  // row pitch in half2 words (+1: conflict-free column walks)
  __half2* xs = reinterpret_cast<__half2*>(fin_smem);
  float* ws = reinterpret_cast<float*>(xs + (FIN_TILE + 6) * sp);
  const int b = blockIdx.y, t0 = blockIdx.x * FIN_TILE;
  for (int i = threadIdx.x; i < 7 * C; i += FIN_TILE) ws[i] = w[i];
  const int words = C / 2;
  for (int i = threadIdx.x; i < (FIN_TILE + 6) * words; i += FIN_TILE) {
    const int r = i / words, cw = i - r * words, t = t0 + r - 3;
    __half2 v = __floats2half2_rn(0.f, 0.f);
    if (t >= 0 && t < T) v = *reinterpret_cast<const __half2*>(x + ((size_t)b * T + t) * pitch + 2 * cw);
    xs[r * sp + cw] = v;
  }
  __syncthreads();
  const int t = t0 + threadIdx.x;
  if (t >= T) return;
  float acc = bias;
  for (int tap = 0; tap < 7; ++tap) {
    const __half2* row = xs + (threadIdx.x + tap) * sp;
    const float* wr = ws + tap * C;
    for (int cw = 0; cw < words; ++cw) {
      const float2 f = __half22float2(row[cw]);
      acc = fmaf(f.x, wr[2 * cw], acc);
      acc = fmaf(f.y, wr[2 * cw + 1], acc);
    }
  }
  wav[(size_t)b * T + t] = tanhf(__half2float(__float2half_rn(acc)));
}

}  // namespace zb

using namespace zb;

// ------------------------------------------------------------------------------------------------ layer launch
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}
static int sm_count() {
  static int n = [] {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    return v > 0 ? v : 148;
  }();
  return n;
}
// Epilogue traffic through the TMA unit for the layers whose epilogue moves the most bytes per flop: the 1x1 convs that
// close a residual unit (residual in, raw + activated out).  ZB200_DAC_TE=0 switches back to direct loads / stores.
static bool dac_te_enabled() {
  const char* v = zb::tuning_env("ZB200_DAC_TE");  // read per call: the parity test compares both epilogues in one process
  return !(v && atoi(v) == 0);
}
static bool dac_wide_epilogue_enabled() {  // ZB200_DAC_EW12=0: 8-warp direct epilogue everywhere (A/B switch)
  const char* v = zb::tuning_env("ZB200_DAC_EW12");
  return !(v && atoi(v) == 0);
}
template <int NT, int MODE>
static int launch_dac_conv(const zb200_dac_layer& L, int batch, int T, void* const* regions, cudaStream_t st,
                           float* wav = nullptr) {
  constexpr bool TE = MODE == DC_TMA;
  const int T_in = L.t_mul_in * T;
  DacConvArgs a{};
  a.m_rows = L.up > 0 ? T_in + 1 : T_in;
  a.T_out = L.up > 0 ? T_in * L.up : T_in;
  a.n_total = L.c_out * (L.up > 0 ? L.up : 1);
  a.tiles_per_seq = ceil_div(a.m_rows, 128);
  a.Wp = (const uint8_t*)L.w;
  a.n_tiles = ceil_div(a.n_total, NT);
  const int max_tpc = 256 / NT;
  a.n_groups = ceil_div(a.n_tiles, max_tpc);
  a.tpc = ceil_div(a.n_tiles, a.n_groups);
  a.cblocks = ceil_div(L.pitch_in, 64), a.taps = L.taps, a.shift0 = L.shift0, a.shift_step = L.shift_step;
  const int s_last = L.shift0 + (L.taps - 1) * L.shift_step;
  a.smin = s_last < L.shift0 ? s_last : L.shift0;
  a.halo_rows = 128 + (s_last < L.shift0 ? L.shift0 - s_last : s_last - L.shift0);
  ZB_REQUIRE(a.halo_rows <= 256, "dac_decode: filter span %d too wide for one TMA box", a.halo_rows - 128);
  a.a_stage_bytes = (a.halo_rows * 128 + 1023) & ~1023;
  a.w_stage_bytes = a.tpc * NT * 128;
  const int budget = DC_SMEM_BYTES - DC_BAR_BYTES - 1024 - (TE ? DC_E_EXTRA : 0);
  const int kb = a.taps * a.cblocks;
  a.w_resident = a.n_groups == 1 && kb <= DC_MAX_SW && kb * a.w_stage_bytes + 2 * a.a_stage_bytes <= budget;
  if (a.w_resident) {
    a.sw = kb;
    a.sa = (budget - a.sw * a.w_stage_bytes) / a.a_stage_bytes;
  } else {
    a.sa = 3;
    a.sw = (budget - a.sa * a.a_stage_bytes) / a.w_stage_bytes;
    if (a.sw < 2) a.sa = 2, a.sw = (budget - a.sa * a.a_stage_bytes) / a.w_stage_bytes;
  }
  a.sa = a.sa > DC_MAX_SA ? DC_MAX_SA : a.sa;
  a.sw = a.sw > DC_MAX_SW ? DC_MAX_SW : a.sw;
  ZB_REQUIRE(a.sa >= 2 && a.sw >= (a.w_resident ? 1 : 2), "dac_decode: layer does not fit the shared-memory rings");
  a.total_tiles = batch * a.tiles_per_seq * a.n_groups;
  a.bias = L.bias;
  a.resid = L.resid_region >= 0 ? (const f16*)regions[L.resid_region] : nullptr;
  a.out_raw = L.raw_region >= 0 ? (f16*)regions[L.raw_region] : nullptr;
  a.out_act = L.act_region >= 0 ? (f16*)regions[L.act_region] : nullptr;
  a.wav = wav;
  a.alpha = L.alpha, a.inv_alpha = L.inv_alpha;
  a.pitch_out = L.pitch_out, a.C_out = L.c_out, a.up = L.up, a.pad = L.up / 2;
  ZB_REQUIRE(!(a.resid && a.up > 0), "dac_decode: residual input on a transposed conv");

  EncodeTiledFn enc = encode_tiled_fn();
  ZB_REQUIRE(enc != nullptr, "dac_decode: cuTensorMapEncodeTiled is not available from this driver");
  CUtensorMap xmap;
  const cuuint64_t gdim[3] = {(cuuint64_t)L.pitch_in, (cuuint64_t)T_in, (cuuint64_t)batch};
  const cuuint64_t gstr[2] = {(cuuint64_t)L.pitch_in * 2, (cuuint64_t)T_in * L.pitch_in * 2};
  const cuuint32_t box[3] = {64, (cuuint32_t)a.halo_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult cr = enc(&xmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, regions[L.in_region], gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ZB_REQUIRE(cr == CUDA_SUCCESS, "dac_decode: cuTensorMapEncodeTiled failed (%d)", (int)cr);
  CUtensorMap rmap = xmap, rawmap = xmap, actmap = xmap;
  if (TE) {  // [batch][T_out][C_out] fp16, 32-channel x 32-row boxes, 64-byte swizzle (conflict-free row-per-thread access)
    ZB_REQUIRE(a.up == 0 && a.pitch_out % 32 == 0 && a.n_total == a.C_out && a.C_out % 32 == 0 && a.n_tiles == a.tpc * a.n_groups,
               "dac_decode: TMA epilogue needs a plain conv with a multiple of 32 output channels and equal column groups");
    const cuuint64_t odim[3] = {(cuuint64_t)a.C_out, (cuuint64_t)a.T_out, (cuuint64_t)batch};
    const cuuint64_t ostr[2] = {(cuuint64_t)a.pitch_out * 2, (cuuint64_t)a.T_out * a.pitch_out * 2};
    const cuuint32_t obox[3] = {32, 32, 1};
    const void* ptrs[3] = {a.resid, a.out_raw, a.out_act};
    CUtensorMap* maps[3] = {&rmap, &rawmap, &actmap};
    for (int i = 0; i < 3; ++i) {
      if (!ptrs[i]) continue;
      const CUresult r2 = enc(maps[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(ptrs[i]), odim, ostr, obox, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      ZB_REQUIRE(r2 == CUDA_SUCCESS, "dac_decode: cuTensorMapEncodeTiled (epilogue map %d) failed (%d)", i, (int)r2);
    }
  }

  auto kern = dac_conv_kernel<NT, MODE>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    ZB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, DC_SMEM_BYTES));
    configured = true;
  }
  const int smem = a.sa * a.a_stage_bytes + a.sw * a.w_stage_bytes + DC_BAR_BYTES + 1024 + (TE ? DC_E_EXTRA : 0);
  const int grid = a.total_tiles < sm_count() ? a.total_tiles : sm_count();
#ifdef SK_TRACE
  static long long* trace_buf = nullptr;
  if (!trace_buf) cudaMalloc(&trace_buf, 256 * 8 * sizeof(long long));
  cudaMemsetAsync(trace_buf, 0, 256 * 8 * sizeof(long long), st);
  a.trace = trace_buf;
#endif
  kern<<<grid, DcCfg<MODE>::THREADS, smem, st>>>(xmap, rmap, rawmap, actmap, a);
  ZB_CHECK_LAUNCH("dac_conv_kernel");
#ifdef SK_TRACE
  {
    static long long host[256 * 8];
    cudaStreamSynchronize(st);
    cudaMemcpy(host, trace_buf, sizeof(host), cudaMemcpyDeviceToHost);
    double m[8] = {0};
    for (int c = 0; c < grid; ++c)
      for (int k = 0; k < 8; ++k) m[k] += (double)host[c * 8 + k] / grid;
    const double tiles = (double)a.total_tiles / grid;
    fprintf(stderr, "[dac trace] N=%d taps=%d cb=%d res=%d sa=%d sw=%d tiles/cta=%.1f | per tile (clk): mma wait t_empty %.0f a_full %.0f w_full %.0f total %.0f"
            " | epi wait t_full %.0f total %.0f | prodA wait a_empty %.0f total %.0f\n",
            a.n_total, a.taps, a.cblocks, a.w_resident, a.sa, a.sw, tiles, m[0] / tiles, m[1] / tiles, m[2] / tiles, m[3] / tiles,
            m[4] / tiles, m[5] / tiles, m[6] / tiles, m[7] / tiles);
  }
#endif
  return ZB200_OK;
}

extern "C" int zb200_dac_decode(const zb200_dac_plan* p, const int64_t* codes, int batch, int t, void* const* regions,
                                int64_t region_halves, float* wav, void* stream) {
  ZB_REQUIRE(p && codes && regions && wav && batch > 0 && t > 0, "dac_decode: bad arguments");
  ZB_REQUIRE(p->n_q > 0 && p->n_q <= ZB200_MAX_CODEBOOKS && p->latent % 64 == 0, "dac_decode: bad quantizer shape");
  cudaStream_t st = (cudaStream_t)stream;
  // capacity check of every region use
  for (int i = 0; i < p->n_layers; ++i) {
    const zb200_dac_layer& L = p->layers[i];
    ZB_REQUIRE(L.pitch_in % 8 == 0 && L.pitch_out % 8 == 0 && L.c_out % 8 == 0 && (L.tile_n == 64 || L.tile_n == 96),
               "dac_decode: layer %d has an unsupported shape", i);
    const int64_t T_in = (int64_t)L.t_mul_in * t, T_out = L.up > 0 ? T_in * L.up : T_in;
    ZB_REQUIRE(batch * T_in * L.pitch_in <= region_halves && batch * T_out * L.pitch_out <= region_halves,
               "dac_decode: layer %d does not fit the work regions (%lld halves)", i, (long long)region_halves);
    ZB_REQUIRE((int64_t)batch * ceil_div((int)(L.up > 0 ? T_in + 1 : T_in), 128) * ceil_div(L.c_out * (L.up > 0 ? L.up : 1), L.tile_n) < (1ll << 31),
               "dac_decode: layer %d grid too large", i);
  }
  {
    dim3 grid(t, batch);
    dac_from_codes_kernel<<<grid, 256, 0, st>>>(codes, p->code_table, (f16*)regions[p->z_region], t, p->n_q, p->n_codes, p->latent);
    ZB_CHECK_LAUNCH("dac_from_codes_kernel");
  }
  for (int i = 0; i < p->n_layers; ++i) {
    const zb200_dac_layer& L = p->layers[i];
    bool te = dac_te_enabled() && L.up == 0 && L.taps == 1 && L.c_out % 32 == 0 && L.pitch_out % 32 == 0;
    if (te) {  // all column groups of a tile must be alike (the epilogue warps walk a fixed block sequence)
      const int n_tiles = ceil_div(L.c_out, L.tile_n), max_tpc = 256 / L.tile_n, n_groups = ceil_div(n_tiles, max_tpc);
      te = n_tiles % n_groups == 0 && L.c_out % L.tile_n == 0;
    }
    // without a residual input the direct epilogue needs few registers and runs 12 warps (DC_DIRECT)
    const int mode = te ? DC_TMA : (L.resid_region < 0 && dac_wide_epilogue_enabled() ? DC_DIRECT : DC_DIRECT_RESID);
    int rc;
    if (L.tile_n == 96)
      rc = mode == DC_TMA      ? launch_dac_conv<96, DC_TMA>(L, batch, t, regions, st)
           : mode == DC_DIRECT ? launch_dac_conv<96, DC_DIRECT>(L, batch, t, regions, st)
                               : launch_dac_conv<96, DC_DIRECT_RESID>(L, batch, t, regions, st);
    else
      rc = mode == DC_TMA      ? launch_dac_conv<64, DC_TMA>(L, batch, t, regions, st)
           : mode == DC_DIRECT ? launch_dac_conv<64, DC_DIRECT>(L, batch, t, regions, st)
                               : launch_dac_conv<64, DC_DIRECT_RESID>(L, batch, t, regions, st);
    if (rc != ZB200_OK) return rc;
  }
  static const bool final_scalar = zb::tuning_env("ZB200_DAC_FINAL_SCALAR") != nullptr;  // A/B switch
  if (p->final_wp && p->final_bias8 && !final_scalar && p->final_t_mul % 8 == 0 && (8 * p->final_c) % 64 == 0 &&
      p->final_pitch == p->final_c) {
    // last conv on the tensor cores: rows of 8 time steps (8*final_c channels), 3 row taps, 8 output phases, tanh
    zb200_dac_layer L{};
    L.taps = 3, L.shift0 = -1, L.shift_step = 1, L.up = 0;
    L.pitch_in = 8 * p->final_c, L.c_out = 8, L.pitch_out = 8, L.tile_n = 16, L.t_mul_in = p->final_t_mul / 8;
    L.in_region = p->final_region, L.resid_region = L.raw_region = L.act_region = -1;
    L.w = p->final_wp, L.bias = p->final_bias8, L.alpha = nullptr, L.inv_alpha = nullptr;
    const int rc = launch_dac_conv<16, DC_DIRECT>(L, batch, t, regions, st, wav);
    if (rc != ZB200_OK) return rc;
  } else {
    const int T_out = p->final_t_mul * t;
    const size_t smem = (size_t)(FIN_TILE + 6) * (p->final_c / 2 + 1) * 4 + (size_t)7 * p->final_c * 4;
    static bool configured = false;
    if (!configured) {
      ZB_CHECK_CUDA(cudaFuncSetAttribute(dac_final_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      configured = true;
    }
    ZB_REQUIRE(smem <= 96 * 1024 && p->final_c % 2 == 0, "dac_decode: final conv too wide");
    dim3 grid(ceil_div(T_out, FIN_TILE), batch);
    dac_final_conv_kernel<<<grid, FIN_TILE, smem, st>>>((const f16*)regions[p->final_region], p->final_w, p->final_bias, wav,
                                                        T_out, p->final_c, p->final_pitch);
    ZB_CHECK_LAUNCH("dac_final_conv_kernel");
  }
  return ZB200_OK;
}

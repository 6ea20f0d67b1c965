// Fused MAT / SPADE-style conditional normalisation:
//   y = act( (x - mean) * rstd * (1 + gs + gI) + (bs + bI) )
// one read of x, one read of the spatial (gI,bI) map, one write of y; all 128-bit
// accesses, per-(n,c) parameters held in registers across the pixel loop.  The
// nearest x2 upsample in front of a block is folded into the x read (h/up, w/up).
// Statistics arrive as fixed-order partials from the producing conv's epilogue and
// are reduced by in_stats_finalize_kernel (fp64 combine, warp-shuffle reduced).
// HBM-bound: algorithmic bytes per output element = 2 (x, /up^2 when upsampled)
// + 4 (gI,bI) + 2 (y) in bf16.
#include "ptx.cuh"
#include <stdlib.h>

#include "s2p_internal.h"

namespace s2p {

// Statistics reduction shared by the two finalize kernels.  blockDim = (32 channels, 8 row
// slices): slice k sums partial rows k, k+8, ... in fp64, then the 8 slices are combined
// in a fixed order -> run-to-run bit-stable, and 8x the memory parallelism of a serial walk.
__device__ __forceinline__ bool reduce_partials(const float2* __restrict__ partial, int rows, int C, int n, int c,
                                                double& s_out, double& q_out) {
  __shared__ double sh[2][8][32];
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float2* src = partial + static_cast<size_t>(n) * rows * C + c;
    // four loads in flight per thread, summed in the same (ascending-row) order as a serial walk: at 100x100 a slice
    // is 40 rows, and 40 dependent L2 round trips made this a 20 us launch
    int r = threadIdx.y;
    for (; r + 24 < rows; r += 32) {
      const float2 v0 = __ldg(src + static_cast<size_t>(r) * C), v1 = __ldg(src + static_cast<size_t>(r + 8) * C);
      const float2 v2 = __ldg(src + static_cast<size_t>(r + 16) * C), v3 = __ldg(src + static_cast<size_t>(r + 24) * C);
      s += static_cast<double>(v0.x); q += static_cast<double>(v0.y);
      s += static_cast<double>(v1.x); q += static_cast<double>(v1.y);
      s += static_cast<double>(v2.x); q += static_cast<double>(v2.y);
      s += static_cast<double>(v3.x); q += static_cast<double>(v3.y);
    }
    for (; r < rows; r += 8) {
      const float2 v = __ldg(src + static_cast<size_t>(r) * C);
      s += static_cast<double>(v.x);
      q += static_cast<double>(v.y);
    }
  }
  sh[0][threadIdx.y][threadIdx.x] = s;
  sh[1][threadIdx.y][threadIdx.x] = q;
  __syncthreads();
  if (threadIdx.y != 0 || c >= C) return false;
  s = 0.0; q = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) { s += sh[0][k][threadIdx.x]; q += sh[1][k][threadIdx.x]; }
  s_out = s; q_out = q;
  return true;
}

// mean_rstd[n][c] = (mean, 1/sqrt(var+eps)).  grid = (ceil(C/32), B).
__global__ void __launch_bounds__(256)
in_stats_finalize_kernel(const float2* __restrict__ partial, int rows, float2* __restrict__ out,
                         int B, int C, double inv_hw, float eps) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int n = blockIdx.y;
  double s, q;
  if (!reduce_partials(partial, rows, C, n, c, s, q)) return;
  const double mean = s * inv_hw;
  double var = q * inv_hw - mean * mean;
  if (var < 0.0) var = 0.0;
  out[static_cast<size_t>(n) * C + c] =
      make_float2(static_cast<float>(mean), static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps))));
}

// Statistics finalize + per-(n,c) modulation table for the fused conv epilogue
// (s2p_conv_mat_fused_fwd):  tab[n][c] = (rstd, -mean*rstd, 1 + gs + bias_gamma, bs + bias_beta).
__global__ void __launch_bounds__(256)
mat_prepare_kernel(const float2* __restrict__ partial, int rows, const float* __restrict__ gs,
                   const float* __restrict__ bs, int gsbs_pitch, const float* __restrict__ bias_gb,
                   float4* __restrict__ tab, int B, int C, double inv_hw, float eps) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int n = blockIdx.y;
  double s, q;
  if (!reduce_partials(partial, rows, C, n, c, s, q)) return;
  const double mean = s * inv_hw;
  double var = q * inv_hw - mean * mean;
  if (var < 0.0) var = 0.0;
  const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
  const float fmean = static_cast<float>(mean);
  float g1 = 1.f, b0 = 0.f;
  if (gs != nullptr) {
    g1 += __ldg(gs + static_cast<size_t>(n) * gsbs_pitch + c);
    b0 += __ldg(bs + static_cast<size_t>(n) * gsbs_pitch + c);
  }
  if (bias_gb != nullptr) {
    g1 += __ldg(bias_gb + c);
    b0 += __ldg(bias_gb + C + c);
  }
  tab[static_cast<size_t>(n) * C + c] = make_float4(rstd, -fmean * rstd, g1, b0);
}

struct NormParams {
  const __nv_bfloat16* x;
  const __nv_bfloat16* x_lo;   // low halves of a wide (bf16 hi + lo pair) input, same pitch; or null
  __nv_bfloat16* y_lo;         // low halves of a wide output (same pitch as y) or null
  int y_f16;                   // y is stored as fp16 (an instance-normalised operand of an fp16 tensor-core conv)
  const float2* mean_rstd;   // [B][C]
  const float* gs;           // [B][gsbs_pitch] or null
  const float* bs;
  const __nv_bfloat16* gb;   // [B][H][W][gb_pitch]: gI in [0,C), bI in [C,2C); or null
  __nv_bfloat16* y;
  int x_pitch, gb_pitch, y_pitch, gsbs_pitch;
  int B, H, W, C, up, act;
  int pix_per_cta;
};

__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}

// blockDim = (C/8 channel groups, PY pixel slots).  grid = (chunks per image, B).
// (launch bounds: the plain-IN variants -- the image encoder's three norms -- are held to 3 CTAs per SM; at 96 registers
// the wide-input variant ran 2 per SM and its load / compute phases did not overlap enough to hold the HBM rate at the
// power-capped clock)
template <bool HAS_GB, bool UP2, bool WIDE, bool HAS_GS>
__global__ void __launch_bounds__(256, (HAS_GB || HAS_GS) ? 2 : 3) mat_norm_kernel(const NormParams p) {
  const int cg = threadIdx.x;              // channel group (8 channels)
  const int c0 = cg * 8;
  const int n = blockIdx.y;
  // per-(n,c) affine:  y = x * A + Bc + x*rstd... folded as
  //   xhat = x*rstd - mean*rstd ;  y = xhat*(1+gs+gI) + (bs+bI)
  float sc[8], sh[8], g1[8], b0[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float2 mr = __ldg(p.mean_rstd + static_cast<size_t>(n) * p.C + c0 + j);
    sc[j] = mr.y;
    sh[j] = -mr.x * mr.y;
    // (HAS_GS == false: plain instance norm -- no per-channel modulation registers at all)
    g1[j] = HAS_GS ? 1.f + __ldg(p.gs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 1.f;
    b0[j] = HAS_GS ? __ldg(p.bs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 0.f;
  }
  const int HW = p.H * p.W;
  const int Wx = UP2 ? (p.W >> 1) : p.W;
  const int Hx = UP2 ? (p.H >> 1) : p.H;
  const int pix_begin = blockIdx.x * p.pix_per_cta;
  const int pix_end = min(pix_begin + p.pix_per_cta, HW);
  const size_t img_pix = static_cast<size_t>(n) * HW;
  const size_t img_pix_x = static_cast<size_t>(n) * Hx * Wx;

  constexpr int UNROLL = 4;
  for (int base = pix_begin + threadIdx.y; base < pix_end; base += blockDim.y * UNROLL) {
    uint4 xv[UNROLL], xlv[UNROLL], gv[UNROLL], bv[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int pix = base + u * blockDim.y;
      if (pix < pix_end) {
        size_t xpix;
        if (UP2) {
          const int h = pix / p.W, w = pix - h * p.W;
          xpix = img_pix_x + static_cast<size_t>(h >> 1) * Wx + (w >> 1);
          xv[u] = __ldg(reinterpret_cast<const uint4*>(p.x + xpix * p.x_pitch + c0));
          if (WIDE) xlv[u] = __ldg(reinterpret_cast<const uint4*>(p.x_lo + xpix * p.x_pitch + c0));
        } else {
          xpix = img_pix + pix;
          xv[u] = ld_stream(p.x + xpix * p.x_pitch + c0);
          if (WIDE) xlv[u] = ld_stream(p.x_lo + xpix * p.x_pitch + c0);
        }
        if (HAS_GB) {
          const __nv_bfloat16* g = p.gb + (img_pix + pix) * p.gb_pitch + c0;
          gv[u] = ld_stream(g);
          bv[u] = ld_stream(g + p.C);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int pix = base + u * blockDim.y;
      if (pix < pix_end) {
        float xf[8], gf[8], bf[8], o[8];
        unpack8(xv[u], xf);
        if (WIDE) {
          float xl[8];
          unpack8(xlv[u], xl);
#pragma unroll
          for (int j = 0; j < 8; ++j) xf[j] += xl[j];
        }
        if (HAS_GB) { unpack8(gv[u], gf); unpack8(bv[u], bf); }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = fmaf(xf[j], sc[j], sh[j]);
          const float g = HAS_GB ? g1[j] + gf[j] : g1[j];
          const float b = HAS_GB ? b0[j] + bf[j] : b0[j];
          float v = (HAS_GB || HAS_GS) ? fmaf(xh, g, b) : xh;
          if (p.act == 2) v = v > 0.f ? v : 0.2f * v;
          else if (p.act == 1) v = fmaxf(v, 0.f);
          o[j] = v;
        }
        uint4 ov = p.y_f16 ? make_uint4(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3]), pack_f16x2(o[4], o[5]),
                                        pack_f16x2(o[6], o[7]))
                           : make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                                        pack_bf16x2(o[6], o[7]));
        st_stream(p.y + (img_pix + pix) * p.y_pitch + c0, ov);
        if (p.y_lo != nullptr) {
          const uint4 lv = make_uint4(pack_bf16x2(o[0] - bf16_lo(ov.x), o[1] - bf16_hi(ov.x)),
                                      pack_bf16x2(o[2] - bf16_lo(ov.y), o[3] - bf16_hi(ov.y)),
                                      pack_bf16x2(o[4] - bf16_lo(ov.z), o[5] - bf16_hi(ov.z)),
                                      pack_bf16x2(o[6] - bf16_lo(ov.w), o[7] - bf16_hi(ov.w)));
          st_stream(p.y_lo + (img_pix + pix) * p.y_pitch + c0, lv);
        }
      }
    }
  }
}

}  // namespace s2p

extern "C" int s2p_in_stats_finalize(const float* stats_partial, int rows, float* mean_rstd, int B, int C,
                                     int HW, float eps, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(stats_partial && mean_rstd, "stats_finalize: null pointer");
  S2P_CHECK_ARG(rows > 0 && B > 0 && C > 0 && HW > 0, "stats_finalize: bad dims");
  S2P_CHECK_ARG(B <= 65535, "stats_finalize: B=%d exceeds the grid's y extent (65535 images per call)", B);
  dim3 grid((C + 31) / 32, B), block(32, 8);
  in_stats_finalize_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float2*>(stats_partial), rows, reinterpret_cast<float2*>(mean_rstd), B, C,
      1.0 / static_cast<double>(HW), eps);
  count_launch();
  return check_launch("in_stats_finalize_kernel");
}

extern "C" int s2p_mat_prepare(const float* stats_partial, int rows, const float* gs, const float* bs,
                               int gsbs_pitch, const float* bias_gb, float* mod_table, int B, int C, int HW,
                               float eps, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(stats_partial && mod_table, "mat_prepare: null pointer");
  S2P_CHECK_ARG(rows > 0 && B > 0 && C > 0 && HW > 0, "mat_prepare: bad dims");
  S2P_CHECK_ARG((gs == nullptr) == (bs == nullptr), "mat_prepare: gs and bs must both be given or both null");
  S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(mod_table) & 15) == 0, "mat_prepare: table must be 16-byte aligned");
  S2P_CHECK_ARG(B <= 65535, "mat_prepare: B=%d exceeds the grid's y extent (65535 images per call)", B);
  dim3 grid((C + 31) / 32, B), block(32, 8);
  mat_prepare_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float2*>(stats_partial), rows, gs, bs, gsbs_pitch, bias_gb,
      reinterpret_cast<float4*>(mod_table), B, C, 1.0 / static_cast<double>(HW), eps);
  count_launch();
  return check_launch("mat_prepare_kernel");
}

extern "C" int s2p_mat_norm_fwd(const void* x, const void* x_lo, int x_pitch, int up, const float* mean_rstd,
                                const float* gs, const float* bs, int gsbs_pitch, const void* gb_spatial, int gb_pitch,
                                void* y, void* y_lo, int y_pitch, int B, int H, int W, int C, int act, int y_f16,
                                void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && mean_rstd && y, "mat_norm: null pointer");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && C <= 2048, "mat_norm: bad dims C=%d", C);
  S2P_CHECK_ARG(up == 1 || (up == 2 && H % 2 == 0 && W % 2 == 0), "mat_norm: up=%d H=%d W=%d", up, H, W);
  S2P_CHECK_ARG(x_pitch >= C && x_pitch % 8 == 0 && y_pitch >= C && y_pitch % 8 == 0, "mat_norm: pitch");
  S2P_CHECK_ARG(!gb_spatial || (gb_pitch >= 2 * C && gb_pitch % 8 == 0), "mat_norm: gb_pitch=%d", gb_pitch);
  S2P_CHECK_ARG((gs == nullptr) == (bs == nullptr), "mat_norm: gs and bs must both be given or both null");
  S2P_CHECK_ARG(act >= 0 && act <= 2, "mat_norm: act=%d", act);
  S2P_CHECK_ARG(!(y_f16 && y_lo), "mat_norm: an fp16 output has no low halves");
  S2P_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(x_lo) |
                  reinterpret_cast<uintptr_t>(y_lo) |
                  reinterpret_cast<uintptr_t>(gb_spatial)) & 15) == 0, "mat_norm: 16-byte alignment");
  NormParams p{};
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.x_lo = static_cast<const __nv_bfloat16*>(x_lo);
  p.y_lo = static_cast<__nv_bfloat16*>(y_lo);
  p.y_f16 = y_f16 != 0;
  p.mean_rstd = reinterpret_cast<const float2*>(mean_rstd);
  p.gs = gs; p.bs = bs;
  p.gb = static_cast<const __nv_bfloat16*>(gb_spatial);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.x_pitch = x_pitch; p.gb_pitch = gb_pitch; p.y_pitch = y_pitch; p.gsbs_pitch = gsbs_pitch;
  p.B = B; p.H = H; p.W = W; p.C = C; p.up = up; p.act = act;
  const int groups = C / 8;                    // threads along x
  int py = 256 / groups;                       // pixel slots per CTA
  if (py < 1) py = 1;
  const int HW = H * W;
  // ~8 CTAs per SM in flight at large batch; each CTA walks >= 4*py*4 pixels
  static const int pix_per_thread = [] {   // diagnostic override: S2P_NORM_PIX_PER_THREAD
    const char* e = getenv("S2P_NORM_PIX_PER_THREAD");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 16;
  }();
  int chunks = (HW + py * pix_per_thread - 1) / (py * pix_per_thread);
  const int max_chunks = (num_sms() * 16 + B - 1) / B;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  p.pix_per_cta = (HW + chunks - 1) / chunks;
  chunks = (HW + p.pix_per_cta - 1) / p.pix_per_cta;
  S2P_CHECK_ARG(B <= 65535, "mat_norm: B=%d exceeds the grid's y extent (65535 images per call)", B);
  dim3 block(groups, py), grid(chunks, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define S2P_NORM_LAUNCH(GB, UP)                                                      \
  do {                                                                               \
    if (x_lo) {                                                                      \
      if (gs) mat_norm_kernel<GB, UP, true, true><<<grid, block, 0, st>>>(p);        \
      else mat_norm_kernel<GB, UP, true, false><<<grid, block, 0, st>>>(p);          \
    } else {                                                                         \
      if (gs) mat_norm_kernel<GB, UP, false, true><<<grid, block, 0, st>>>(p);       \
      else mat_norm_kernel<GB, UP, false, false><<<grid, block, 0, st>>>(p);         \
    }                                                                                \
  } while (0)
  if (gb_spatial) {
    if (up == 2) S2P_NORM_LAUNCH(true, true);
    else S2P_NORM_LAUNCH(true, false);
  } else {
    if (up == 2) S2P_NORM_LAUNCH(false, true);
    else S2P_NORM_LAUNCH(false, false);
  }
#undef S2P_NORM_LAUNCH
  count_launch();
  return check_launch("mat_norm_kernel");
}

// =====================================================================================
// Backward of the fused IN + MAT modulation + activation (training path).
//   forward:  xh = (x_up - mean) * rstd ;  g = 1 + gs + gI ;  pre = xh * g + bs + bI ;  y = act(pre)
//   pass A (mat_norm_bwd_reduce_kernel), one read of dy / x / (gI,bI), per output pixel:
//       dpre = dy * act'(pre) ;  dgI = dpre * xh ;  dbI = dpre ;  dxh = dpre * g
//       writes the spatial (dgI | dbI) map and dxh (bf16), and per-(n,c) partial sums
//       (sum dpre*xh, sum dpre, sum dxh, sum dxh*xh) in a fixed order (no atomics)
//   sum4_finalize_kernel: partial rows -> sums[n][c][4]   (dgs, dbs, S1, S2)
//   pass B (mat_norm_bwd_dx_kernel), per INPUT pixel (the x2 nearest upsample's backward is the
//       sum over the 2x2 children):  dx = rstd * (sum_children dxh - S1/N - xh * S2/N) (+ add)
// =====================================================================================
namespace s2p {

struct NormBwdParams {
  const __nv_bfloat16* dy;
  const __nv_bfloat16* x;
  const float2* mean_rstd;
  const float* gs;
  const float* bs;
  const __nv_bfloat16* gb;
  __nv_bfloat16* dgb;
  __nv_bfloat16* dxh;
  float4* partial;
  int dy_pitch, x_pitch, gb_pitch, dgb_pitch, dxh_pitch, gsbs_pitch;
  int B, H, W, C, up, act;
  int pix_per_cta, chunks;
};

__device__ __forceinline__ uint4 pack8(const float (&o)[8]) {
  return make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                    pack_bf16x2(o[6], o[7]));
}

template <bool HAS_GB, bool UP2>
__global__ void __launch_bounds__(256) mat_norm_bwd_reduce_kernel(const NormBwdParams p) {
  extern __shared__ float4 red[];            // [blockDim.y][C]
  const int cg = threadIdx.x, c0 = cg * 8, n = blockIdx.y;
  float sc[8], sh[8], g1[8], b0[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float2 mr = __ldg(p.mean_rstd + static_cast<size_t>(n) * p.C + c0 + j);
    sc[j] = mr.y;
    sh[j] = -mr.x * mr.y;
    g1[j] = 1.f + (p.gs ? __ldg(p.gs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 0.f);
    b0[j] = p.bs ? __ldg(p.bs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 0.f;
  }
  float a0[8], a1[8], a2[8], a3[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a0[j] = 0.f; a1[j] = 0.f; a2[j] = 0.f; a3[j] = 0.f; }
  const int HW = p.H * p.W;
  const int Wx = UP2 ? (p.W >> 1) : p.W, Hx = UP2 ? (p.H >> 1) : p.H;
  const int pix_begin = blockIdx.x * p.pix_per_cta;
  const int pix_end = min(pix_begin + p.pix_per_cta, HW);
  const size_t img_pix = static_cast<size_t>(n) * HW;
  const size_t img_pix_x = static_cast<size_t>(n) * Hx * Wx;
  for (int pix = pix_begin + threadIdx.y; pix < pix_end; pix += blockDim.y) {
    size_t xpix;
    if (UP2) {
      const int h = pix / p.W, w = pix - h * p.W;
      xpix = img_pix_x + static_cast<size_t>(h >> 1) * Wx + (w >> 1);
    } else {
      xpix = img_pix + pix;
    }
    const uint4 xv = __ldg(reinterpret_cast<const uint4*>(p.x + xpix * p.x_pitch + c0));
    const uint4 dv = ld_stream(p.dy + (img_pix + pix) * p.dy_pitch + c0);
    float xf[8], df[8], gf[8], bf[8];
    unpack8(xv, xf);
    unpack8(dv, df);
    if (HAS_GB) {
      const __nv_bfloat16* g = p.gb + (img_pix + pix) * p.gb_pitch + c0;
      unpack8(ld_stream(g), gf);
      unpack8(ld_stream(g + p.C), bf);
    }
    float dg[8], db[8], dx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = fmaf(xf[j], sc[j], sh[j]);
      const float g = HAS_GB ? g1[j] + gf[j] : g1[j];
      const float b = HAS_GB ? b0[j] + bf[j] : b0[j];
      const float pre = fmaf(xh, g, b);
      float d = df[j];
      if (p.act == 2) d = pre > 0.f ? d : 0.2f * d;
      else if (p.act == 1) d = pre > 0.f ? d : 0.f;
      dg[j] = d * xh;
      db[j] = d;
      dx[j] = d * g;
      a0[j] += dg[j];
      a1[j] += d;
      a2[j] += dx[j];
      a3[j] += dx[j] * xh;
    }
    if (HAS_GB) {
      __nv_bfloat16* o = p.dgb + (img_pix + pix) * p.dgb_pitch + c0;
      st_stream(o, pack8(dg));
      st_stream(o + p.C, pack8(db));
    }
    st_stream(p.dxh + (img_pix + pix) * p.dxh_pitch + c0, pack8(dx));
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y * p.C + c0 + j] = make_float4(a0[j], a1[j], a2[j], a3[j]);
  __syncthreads();
  for (int c = threadIdx.y * blockDim.x + threadIdx.x; c < p.C; c += blockDim.x * blockDim.y) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < static_cast<int>(blockDim.y); ++k) {
      const float4 v = red[k * p.C + c];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    p.partial[(static_cast<size_t>(n) * p.chunks + blockIdx.x) * p.C + c] = s;
  }
}

// blockDim = (32 channels, 8 slices): slice k sums partial rows k, k+8, ... in fp64, the slices are combined in a fixed
// order (bit-stable); one thread per channel walking every row was a chain of dependent L2 round trips.
__global__ void __launch_bounds__(256)
sum4_finalize_kernel(const float4* __restrict__ partial, int rows, float4* __restrict__ out, int C) {
  __shared__ double red[8][32][4];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int n = blockIdx.y;
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  if (c < C)
    for (int r = threadIdx.y; r < rows; r += 8) {
      const float4 v = __ldg(partial + (static_cast<size_t>(n) * rows + r) * C + c);
      s0 += v.x; s1 += v.y; s2 += v.z; s3 += v.w;
    }
  double* mine = red[threadIdx.y][threadIdx.x];
  mine[0] = s0; mine[1] = s1; mine[2] = s2; mine[3] = s3;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    double t[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 8; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) t[j] += red[k][threadIdx.x][j];
    out[static_cast<size_t>(n) * C + c] = make_float4((float)t[0], (float)t[1], (float)t[2], (float)t[3]);
  }
}

// blockDim = (C/8 channel groups, PY pixel slots); grid = (chunks of INPUT pixels per image, B).  A thread keeps the
// per-(n,c) terms of its 8 channels (mean, rstd, S1/N, S2/N) in registers across its pixel loop: loading them per
// pixel (the first version) asked the LSU for 6x more bytes than the data itself and ran at 12 % of HBM peak.
template <bool UP2>
__global__ void __launch_bounds__(256)
mat_norm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dxh, int dxh_pitch, const __nv_bfloat16* __restrict__ x,
                       int x_pitch, const float2* __restrict__ mean_rstd, const float4* __restrict__ sums,
                       const __nv_bfloat16* __restrict__ add, int add_pitch, __nv_bfloat16* __restrict__ dx,
                       int dx_pitch, int Hx, int Wx, int C, int pix_per_cta) {
  const int c0 = threadIdx.x * 8;
  const int n = blockIdx.y;
  const int HWx = Hx * Wx;
  const float inv_n = 1.f / static_cast<float>(HWx);
  float mean[8], rstd[8], k1[8], k2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float2 mr = __ldg(mean_rstd + static_cast<size_t>(n) * C + c0 + j);
    const float4 sm = __ldg(sums + static_cast<size_t>(n) * C + c0 + j);
    mean[j] = mr.x; rstd[j] = mr.y; k1[j] = sm.z * inv_n; k2[j] = sm.w * inv_n;
  }
  const int p0 = blockIdx.x * pix_per_cta;
  const int p1 = min(HWx, p0 + pix_per_cta);
  for (int rem = p0 + threadIdx.y; rem < p1; rem += blockDim.y) {
    const size_t xpix = static_cast<size_t>(n) * HWx + rem;
    float D[8];
    if (UP2) {
      const int h = rem / Wx, w = rem - h * Wx;
      const int W = 2 * Wx;
      const size_t base = (static_cast<size_t>(n) * (2 * Hx) + 2 * h) * W + 2 * w;
      float t[8];
      unpack8(ld_stream(dxh + base * dxh_pitch + c0), D);
      unpack8(ld_stream(dxh + (base + 1) * dxh_pitch + c0), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) D[j] += t[j];
      unpack8(ld_stream(dxh + (base + W) * dxh_pitch + c0), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) D[j] += t[j];
      unpack8(ld_stream(dxh + (base + W + 1) * dxh_pitch + c0), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) D[j] += t[j];
    } else {
      unpack8(ld_stream(dxh + xpix * dxh_pitch + c0), D);
    }
    float xf[8], af[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(x + xpix * x_pitch + c0)), xf);
    if (add != nullptr) unpack8(__ldg(reinterpret_cast<const uint4*>(add + xpix * add_pitch + c0)), af);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = (xf[j] - mean[j]) * rstd[j];
      float v = rstd[j] * (D[j] - k1[j] - xh * k2[j]);
      if (add != nullptr) v += af[j];
      o[j] = v;
    }
    *reinterpret_cast<uint4*>(dx + xpix * dx_pitch + c0) = pack8(o);
  }
}

static void norm_bwd_geometry(int B, int H, int W, int C, int* py_out, int* chunks_out, int* ppc_out) {
  const int groups = C / 8;
  int py = 256 / groups;
  if (py < 1) py = 1;
  const int HW = H * W;
  static const int pix_per_thread = [] {   // diagnostic override: S2P_NORM_PIX_PER_THREAD
    const char* e = getenv("S2P_NORM_PIX_PER_THREAD");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 16;
  }();
  int chunks = (HW + py * pix_per_thread - 1) / (py * pix_per_thread);
  const int max_chunks = (num_sms() * 8 + B - 1) / B;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  const int ppc = (HW + chunks - 1) / chunks;
  *py_out = py;
  *ppc_out = ppc;
  *chunks_out = (HW + ppc - 1) / ppc;
}

}  // namespace s2p

extern "C" int s2p_mat_norm_bwd_chunks(int B, int H, int W, int C) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || C % 8) return -1;
  int py, chunks, ppc;
  s2p::norm_bwd_geometry(B, H, W, C, &py, &chunks, &ppc);
  return chunks;
}

extern "C" int s2p_mat_norm_bwd(const void* dy, int dy_pitch, const void* x, int x_pitch, int up,
                                const float* mean_rstd, const float* gs, const float* bs, int gsbs_pitch,
                                const void* gb_spatial, int gb_pitch, int act, void* d_gb, int dgb_pitch,
                                void* dxhat, int dxh_pitch, float* partial, float* sums, const void* add,
                                int add_pitch, void* dx, int dx_pitch, int B, int H, int W, int C, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy && x && mean_rstd && dxhat && partial && sums && dx, "mat_norm_bwd: null pointer");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && C <= 2048, "mat_norm_bwd: bad dims");
  S2P_CHECK_ARG(up == 1 || (up == 2 && H % 2 == 0 && W % 2 == 0), "mat_norm_bwd: up=%d", up);
  S2P_CHECK_ARG((gs == nullptr) == (bs == nullptr), "mat_norm_bwd: gs/bs");
  S2P_CHECK_ARG((gb_spatial == nullptr) == (d_gb == nullptr), "mat_norm_bwd: gb / d_gb must come together");
  S2P_CHECK_ARG(act >= 0 && act <= 2, "mat_norm_bwd: act=%d", act);
  S2P_CHECK_ARG(dy_pitch % 8 == 0 && x_pitch % 8 == 0 && dxh_pitch % 8 == 0 && dx_pitch % 8 == 0 &&
                    gb_pitch % 8 == 0 && dgb_pitch % 8 == 0 && add_pitch % 8 == 0, "mat_norm_bwd: pitch");
  NormBwdParams p{};
  p.dy = static_cast<const __nv_bfloat16*>(dy);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.mean_rstd = reinterpret_cast<const float2*>(mean_rstd);
  p.gs = gs; p.bs = bs;
  p.gb = static_cast<const __nv_bfloat16*>(gb_spatial);
  p.dgb = static_cast<__nv_bfloat16*>(d_gb);
  p.dxh = static_cast<__nv_bfloat16*>(dxhat);
  p.partial = reinterpret_cast<float4*>(partial);
  p.dy_pitch = dy_pitch; p.x_pitch = x_pitch; p.gb_pitch = gb_pitch; p.dgb_pitch = dgb_pitch;
  p.dxh_pitch = dxh_pitch; p.gsbs_pitch = gsbs_pitch;
  p.B = B; p.H = H; p.W = W; p.C = C; p.up = up; p.act = act;
  int py;
  norm_bwd_geometry(B, H, W, C, &py, &p.chunks, &p.pix_per_cta);
  S2P_CHECK_ARG(B <= 65535, "mat_norm_bwd: B=%d exceeds the grid's y extent (65535 images per call)", B);
  dim3 block(C / 8, py), grid(p.chunks, B);
  const size_t smem = static_cast<size_t>(py) * C * sizeof(float4);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  S2P_CHECK_ARG(smem <= 48 * 1024, "mat_norm_bwd: C=%d needs %zu bytes of shared memory", C, smem);
  if (gb_spatial) {
    if (up == 2) mat_norm_bwd_reduce_kernel<true, true><<<grid, block, smem, st>>>(p);
    else mat_norm_bwd_reduce_kernel<true, false><<<grid, block, smem, st>>>(p);
  } else {
    if (up == 2) mat_norm_bwd_reduce_kernel<false, true><<<grid, block, smem, st>>>(p);
    else mat_norm_bwd_reduce_kernel<false, false><<<grid, block, smem, st>>>(p);
  }
  count_launch();
  int rc = check_launch("mat_norm_bwd_reduce_kernel");
  if (rc != S2P_OK) return rc;
  sum4_finalize_kernel<<<dim3((C + 31) / 32, B), dim3(32, 8), 0, st>>>(reinterpret_cast<const float4*>(partial), p.chunks,
                                                               reinterpret_cast<float4*>(sums), C);
  count_launch();
  rc = check_launch("sum4_finalize_kernel");
  if (rc != S2P_OK) return rc;
  const int Hx = H / up, Wx = W / up;
  int dpy, dchunks, dppc;
  norm_bwd_geometry(B, Hx, Wx, C, &dpy, &dchunks, &dppc);
  const dim3 dgrid(dchunks, B), dblock(C / 8, dpy);
  if (up == 2)
    mat_norm_bwd_dx_kernel<true><<<dgrid, dblock, 0, st>>>(p.dxh, dxh_pitch, p.x, x_pitch, p.mean_rstd,
                                                           reinterpret_cast<const float4*>(sums),
                                                           static_cast<const __nv_bfloat16*>(add), add_pitch,
                                                           static_cast<__nv_bfloat16*>(dx), dx_pitch, Hx, Wx, C, dppc);
  else
    mat_norm_bwd_dx_kernel<false><<<dgrid, dblock, 0, st>>>(p.dxh, dxh_pitch, p.x, x_pitch, p.mean_rstd,
                                                            reinterpret_cast<const float4*>(sums),
                                                            static_cast<const __nv_bfloat16*>(add), add_pitch,
                                                            static_cast<__nv_bfloat16*>(dx), dx_pitch, Hx, Wx, C, dppc);
  count_launch();
  return check_launch("mat_norm_bwd_dx_kernel");
}

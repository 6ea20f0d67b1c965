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
    for (int r = threadIdx.y; r < rows; r += 8) {
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
template <bool HAS_GB, bool UP2>
__global__ void __launch_bounds__(256) mat_norm_kernel(const NormParams p) {
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
    g1[j] = 1.f + (p.gs ? __ldg(p.gs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 0.f);
    b0[j] = p.bs ? __ldg(p.bs + static_cast<size_t>(n) * p.gsbs_pitch + c0 + j) : 0.f;
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
    uint4 xv[UNROLL], gv[UNROLL], bv[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int pix = base + u * blockDim.y;
      if (pix < pix_end) {
        size_t xpix;
        if (UP2) {
          const int h = pix / p.W, w = pix - h * p.W;
          xpix = img_pix_x + static_cast<size_t>(h >> 1) * Wx + (w >> 1);
          xv[u] = __ldg(reinterpret_cast<const uint4*>(p.x + xpix * p.x_pitch + c0));
        } else {
          xpix = img_pix + pix;
          xv[u] = ld_stream(p.x + xpix * p.x_pitch + c0);
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
        if (HAS_GB) { unpack8(gv[u], gf); unpack8(bv[u], bf); }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = fmaf(xf[j], sc[j], sh[j]);
          const float g = HAS_GB ? g1[j] + gf[j] : g1[j];
          const float b = HAS_GB ? b0[j] + bf[j] : b0[j];
          float v = fmaf(xh, g, b);
          if (p.act == 2) v = v > 0.f ? v : 0.2f * v;
          else if (p.act == 1) v = fmaxf(v, 0.f);
          o[j] = v;
        }
        uint4 ov = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                              pack_bf16x2(o[6], o[7]));
        st_stream(p.y + (img_pix + pix) * p.y_pitch + c0, ov);
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
  dim3 grid((C + 31) / 32, B), block(32, 8);
  mat_prepare_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float2*>(stats_partial), rows, gs, bs, gsbs_pitch, bias_gb,
      reinterpret_cast<float4*>(mod_table), B, C, 1.0 / static_cast<double>(HW), eps);
  count_launch();
  return check_launch("mat_prepare_kernel");
}

extern "C" int s2p_mat_norm_fwd(const void* x, int x_pitch, int up, const float* mean_rstd, const float* gs,
                                const float* bs, int gsbs_pitch, const void* gb_spatial, int gb_pitch, void* y,
                                int y_pitch, int B, int H, int W, int C, int act, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && mean_rstd && y, "mat_norm: null pointer");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && C <= 2048, "mat_norm: bad dims C=%d", C);
  S2P_CHECK_ARG(up == 1 || (up == 2 && H % 2 == 0 && W % 2 == 0), "mat_norm: up=%d H=%d W=%d", up, H, W);
  S2P_CHECK_ARG(x_pitch >= C && x_pitch % 8 == 0 && y_pitch >= C && y_pitch % 8 == 0, "mat_norm: pitch");
  S2P_CHECK_ARG(!gb_spatial || (gb_pitch >= 2 * C && gb_pitch % 8 == 0), "mat_norm: gb_pitch=%d", gb_pitch);
  S2P_CHECK_ARG((gs == nullptr) == (bs == nullptr), "mat_norm: gs and bs must both be given or both null");
  S2P_CHECK_ARG(act >= 0 && act <= 2, "mat_norm: act=%d", act);
  S2P_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                  reinterpret_cast<uintptr_t>(gb_spatial)) & 15) == 0, "mat_norm: 16-byte alignment");
  NormParams p{};
  p.x = static_cast<const __nv_bfloat16*>(x);
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
  int chunks = (HW + py * 16 - 1) / (py * 16);
  const int max_chunks = (num_sms() * 16 + B - 1) / B;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  p.pix_per_cta = (HW + chunks - 1) / chunks;
  chunks = (HW + p.pix_per_cta - 1) / p.pix_per_cta;
  dim3 block(groups, py), grid(chunks, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (gb_spatial) {
    if (up == 2) mat_norm_kernel<true, true><<<grid, block, 0, st>>>(p);
    else mat_norm_kernel<true, false><<<grid, block, 0, st>>>(p);
  } else {
    if (up == 2) mat_norm_kernel<false, true><<<grid, block, 0, st>>>(p);
    else mat_norm_kernel<false, false><<<grid, block, 0, st>>>(p);
  }
  count_launch();
  return check_launch("mat_norm_kernel");
}

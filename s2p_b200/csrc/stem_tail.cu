// Boundary kernels of the generator:
//  stem: Cin=3 3x3 conv straight from the boundary image (fp32 NCHW in [-1,1] or the
//        dataset's uint8 HWC frames) to NHWC bf16 + instance-norm statistics partials.
//        K = 27 is too thin for UMMA: CUDA-core direct conv, 8 output channels per lane.
//  tail: LeakyReLU -> 3x3 conv Cin->3 -> tanh -> fp32 NCHW and/or uint8 HWC.  N = 3 is
//        too thin for UMMA: each pixel is owned by Cin/8 lanes (8 channels each, one
//        128-bit load per tap), reduced with warp shuffles.
#include <algorithm>
#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

constexpr int kStemRows = 8;      // output rows per CTA = one statistics partial row
constexpr int kStemMaxW = 256;

// CTA = kStemRows full-width output rows of one image.  The (rows+2) x (W+2) x 3 input patch is
// converted to fp32 once into shared memory; a thread owns 8 output channels x 8 consecutive
// pixels of a row, so every weight float4 read from shared memory feeds 32 FMAs.
template <bool U8>
__global__ void __launch_bounds__(256, 2)
stem_conv_kernel(const void* __restrict__ image, const float* __restrict__ w_oihw,
                 const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ y_lo,
                 int y_pitch, float2* __restrict__ stats, int H, int W, int Cout, int rows) {
  extern __shared__ __align__(16) float sm[];
  const int PW = (W + 2 + 8 + 3) & ~3;              // patch row pitch: halo + the last 8-pixel group's overhang, 16-B rows
  float* ws = sm;                                   // [27][Cout], tap = ci*9 + r*3 + s
  float* patch = ws + 27 * Cout;                    // [3][kStemRows+2][PW]
  float* red = patch + 3 * (kStemRows + 2) * PW;    // [slots][Cout][2]
  const int groups = Cout / 8;
  const int slots = blockDim.x / groups;
  const int g = threadIdx.x % groups, slot = threadIdx.x / groups;
  const int c0 = g * 8;
  const int n = blockIdx.y;
  const int h0 = blockIdx.x * kStemRows;
  const int nrows = min(kStemRows, H - h0);
  const int HW = H * W;
  // weights: coalesced global read, transposed into [tap][Cout]
  for (int i = threadIdx.x; i < 27 * Cout; i += blockDim.x) {
    const int co = i / 27, tap = i - co * 27;
    ws[tap * Cout + co] = __ldg(w_oihw + i);
  }
  // patch: one (channel, row) line per warp pass, lanes along the row -- no per-element divisions, and the
  // row's loads are independent of each other (they are issued together, not one L2 round trip at a time)
  {
    const int wid = threadIdx.x >> 5, ln = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int line = wid; line < 3 * (kStemRows + 2); line += nwarps) {
      const int ci = line / (kStemRows + 2), pr = line - ci * (kStemRows + 2);
      const int hh = h0 + pr - 1;
      const bool row_ok = hh >= 0 && hh < H;
      float* dst = patch + line * PW;
#pragma unroll 4
      for (int pw = ln; pw < PW; pw += 32) {
        const int ww = pw - 1;
        float v = 0.f;
        if (row_ok && ww >= 0 && ww < W) {
          if (U8) {
            const uint8_t* im = static_cast<const uint8_t*>(image);
            // two separately rounded steps (no FMA contraction): bit-identical to the driver-side
            // `u.float() * (1/127.5) - 1` conversion, so the uint8 and fp32 entries agree exactly
            v = __fadd_rn(__fmul_rn(static_cast<float>(__ldg(im + (static_cast<size_t>(n) * HW + hh * W + ww) * 3 + ci)),
                                    1.0f / 127.5f), -1.0f);
          } else {
            const float* im = static_cast<const float*>(image);
            v = __ldg(im + (static_cast<size_t>(n) * 3 + ci) * HW + hh * W + ww);
          }
        }
        dst[pw] = v;
      }
    }
  }
  __syncthreads();
  float bv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bv[j] = bias ? bias[c0 + j] : 0.f;
  float s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }

  const int gpr = (W + 7) / 8;                      // 8-pixel groups per row
  for (int pg = slot; pg < nrows * gpr; pg += slots) {
    const int lr = pg / gpr, w0 = (pg - lr * gpr) * 8;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = bv[j];
    // ci / r stay rolled: fully unrolled, the compiler keeps all 27 x 8 weights of the thread in registers
    // (249 registers, one CTA per SM, nothing to overlap the patch fill with); rolled, the 24 weights of a
    // filter row come from shared memory per iteration (each float4 feeds 32 FMAs) and two CTAs share an SM
#pragma unroll 1
    for (int ci = 0; ci < 3; ++ci) {
#pragma unroll 1
      for (int r = 0; r < 3; ++r) {
        const float* prow = patch + (ci * (kStemRows + 2) + lr + r) * PW + w0;
        float in[10];
        {
          const float4 p0 = *reinterpret_cast<const float4*>(prow);
          const float4 p1 = *reinterpret_cast<const float4*>(prow + 4);
          const float2 p2 = *reinterpret_cast<const float2*>(prow + 8);
          in[0] = p0.x; in[1] = p0.y; in[2] = p0.z; in[3] = p0.w;
          in[4] = p1.x; in[5] = p1.y; in[6] = p1.z; in[7] = p1.w;
          in[8] = p2.x; in[9] = p2.y;
        }
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const float4 wa = *reinterpret_cast<const float4*>(ws + (ci * 9 + r * 3 + s) * Cout + c0);
          const float4 wb = *reinterpret_cast<const float4*>(ws + (ci * 9 + r * 3 + s) * Cout + c0 + 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float v = in[i + s];
            acc[i][0] = fmaf(v, wa.x, acc[i][0]); acc[i][1] = fmaf(v, wa.y, acc[i][1]);
            acc[i][2] = fmaf(v, wa.z, acc[i][2]); acc[i][3] = fmaf(v, wa.w, acc[i][3]);
            acc[i][4] = fmaf(v, wb.x, acc[i][4]); acc[i][5] = fmaf(v, wb.y, acc[i][5]);
            acc[i][6] = fmaf(v, wb.z, acc[i][6]); acc[i][7] = fmaf(v, wb.w, acc[i][7]);
          }
        }
      }
    }
    const size_t pix0 = static_cast<size_t>(n) * HW + static_cast<size_t>(h0 + lr) * W + w0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (w0 + i >= W) break;
      uint4 ov = make_uint4(pack_bf16x2(acc[i][0], acc[i][1]), pack_bf16x2(acc[i][2], acc[i][3]),
                            pack_bf16x2(acc[i][4], acc[i][5]), pack_bf16x2(acc[i][6], acc[i][7]));
      *reinterpret_cast<uint4*>(y + (pix0 + i) * y_pitch + c0) = ov;
      const uint32_t pk[4] = {ov.x, ov.y, ov.z, ov.w};
      if (y_lo != nullptr) {
        // wide store (bf16 hi + lo pair): the statistics are those of the fp32 values
        uint32_t lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float a = acc[i][2 * j], b = acc[i][2 * j + 1];
          lo[j] = pack_bf16x2(a - bf16_lo(pk[j]), b - bf16_hi(pk[j]));
          s1[2 * j] += a; s1[2 * j + 1] += b;
          s2[2 * j] += a * a; s2[2 * j + 1] += b * b;
        }
        *reinterpret_cast<uint4*>(y_lo + (pix0 + i) * y_pitch + c0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        continue;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a = bf16_lo(pk[j]), b = bf16_hi(pk[j]);
        s1[2 * j] += a; s1[2 * j + 1] += b;
        s2[2 * j] += a * a; s2[2 * j + 1] += b * b;
      }
    }
  }
  if (stats == nullptr) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    red[(slot * Cout + c0 + j) * 2 + 0] = s1[j];
    red[(slot * Cout + c0 + j) * 2 + 1] = s2[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < Cout; c += blockDim.x) {
    float a = 0.f, b = 0.f;
    for (int sl = 0; sl < slots; ++sl) { a += red[(sl * Cout + c) * 2]; b += red[(sl * Cout + c) * 2 + 1]; }
    stats[(static_cast<size_t>(n) * rows + blockIdx.x) * Cout + c] = make_float2(a, b);
  }
}

// LPP = lanes per pixel = Cin/8
template <int LPP>
__global__ void __launch_bounds__(256)
torgb_tanh_kernel(const __nv_bfloat16* __restrict__ x, int x_pitch, const float* __restrict__ w_oihw,
                  const float* __restrict__ bias, float* __restrict__ out_f32, uint8_t* __restrict__ out_u8,
                  int B, int H, int W) {
  constexpr int Cin = LPP * 8;
  extern __shared__ float sm[];           // [9][Cin][3]
  for (int i = threadIdx.x; i < 9 * Cin * 3; i += blockDim.x) {
    const int o = i % 3, ci = (i / 3) % Cin, tap = i / (3 * Cin);
    sm[i] = w_oihw[(o * Cin + ci) * 9 + tap];
  }
  __syncthreads();
  const int g = threadIdx.x % LPP;
  const int c0 = g * 8;
  const int HW = H * W;
  const long long total = static_cast<long long>(B) * HW;
  const long long pix = static_cast<long long>(blockIdx.x) * (256 / LPP) + threadIdx.x / LPP;
  const bool active = pix < total;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f;
  int n = 0, h = 0, w = 0;
  if (active) {
    n = static_cast<int>(pix / HW);
    const int rem = static_cast<int>(pix - static_cast<long long>(n) * HW);
    h = rem / W; w = rem - h * W;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = h + r - 1;
      if (hh < 0 || hh >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = w + s - 1;
        if (ww < 0 || ww >= W) continue;
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(
            x + (static_cast<size_t>(n) * HW + hh * W + ww) * x_pitch + c0));
        float f[8];
        f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
        f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
        const float* wp = sm + ((r * 3 + s) * Cin + c0) * 3;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = f[j] > 0.f ? f[j] : 0.2f * f[j];
          a0 = fmaf(v, wp[j * 3 + 0], a0);
          a1 = fmaf(v, wp[j * 3 + 1], a1);
          a2 = fmaf(v, wp[j * 3 + 2], a2);
        }
      }
    }
  }
#pragma unroll
  for (int off = LPP / 2; off >= 1; off >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, off);
    a1 += __shfl_xor_sync(0xffffffffu, a1, off);
    a2 += __shfl_xor_sync(0xffffffffu, a2, off);
  }
  if (active && g == 0) {
    const float y0 = tanhf(a0 + (bias ? bias[0] : 0.f));
    const float y1 = tanhf(a1 + (bias ? bias[1] : 0.f));
    const float y2 = tanhf(a2 + (bias ? bias[2] : 0.f));
    const int rem = h * W + w;
    if (out_f32) {
      float* o = out_f32 + static_cast<size_t>(n) * 3 * HW + rem;
      o[0] = y0; o[HW] = y1; o[2 * HW] = y2;
    }
    if (out_u8) {
      uint8_t* o = out_u8 + (static_cast<size_t>(n) * HW + rem) * 3;
      o[0] = static_cast<uint8_t>(fminf(fmaxf(floorf((y0 + 1.f) * 127.5f + 0.5f), 0.f), 255.f));
      o[1] = static_cast<uint8_t>(fminf(fmaxf(floorf((y1 + 1.f) * 127.5f + 0.5f), 0.f), 255.f));
      o[2] = static_cast<uint8_t>(fminf(fmaxf(floorf((y2 + 1.f) * 127.5f + 0.5f), 0.f), 255.f));
    }
  }
}

}  // namespace s2p

extern "C" int s2p_stem_stats_rows(int H, int W) { (void)W; return (H + s2p::kStemRows - 1) / s2p::kStemRows; }

extern "C" int s2p_stem_conv_fwd(const void* image, int image_fmt, const float* w_oihw, const float* bias,
                                 void* y, void* y_lo, int y_pitch, float* stats_partial, int B, int H, int W, int Cout,
                                 void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(image && w_oihw && y, "stem: null pointer");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0 && W <= kStemMaxW, "stem: bad dims (W <= %d)", kStemMaxW);
  S2P_CHECK_ARG(Cout % 8 == 0 && Cout >= 8 && Cout <= 256 && 256 % (Cout / 8) == 0, "stem: Cout=%d", Cout);
  S2P_CHECK_ARG(y_pitch >= Cout && y_pitch % 8 == 0, "stem: y_pitch=%d", y_pitch);
  S2P_CHECK_ARG(image_fmt == S2P_IMG_F32_NCHW || image_fmt == S2P_IMG_U8_NHWC, "stem: image_fmt=%d", image_fmt);
  const int rows = s2p_stem_stats_rows(H, W);
  const int slots = 256 / (Cout / 8);
  const int pw_pitch = (W + 2 + 8 + 3) & ~3;      // the kernel's padded patch row
  const size_t smem = (27 * Cout + 3 * (kStemRows + 2) * pw_pitch + slots * Cout * 2) * sizeof(float);
  // threads: the fewest 8-channel x 8-pixel slots that finish a CTA's rows in the same number of passes, so the
  // last pass is not mostly idle (100 px wide, 8 rows: 104 groups = 4 passes of 26 slots, not 3.25 passes of 32)
  const int groups = Cout / 8;
  const int pgs = kStemRows * ((W + 7) / 8);
  const int passes = (pgs + slots - 1) / slots;
  const int slots_used = (pgs + passes - 1) / passes;
  const unsigned threads = static_cast<unsigned>(std::min(256, (slots_used * groups + 31) & ~31));
  S2P_CHECK_ARG(B <= 65535, "stem: B=%d exceeds the grid's y extent (65535 images per call)", B);
  dim3 grid(rows, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static size_t smem_configured_dev[kMaxDevices] = {};   // per device (0 = the 48 KB default)
  size_t& smem_configured = smem_configured_dev[current_device()];
  if (smem > 48 * 1024 && smem > smem_configured) {
    S2P_CHECK_ARG(smem <= 200 * 1024, "stem: W=%d Cout=%d needs %zu bytes of shared memory", W, Cout, smem);
    S2P_CHECK_CUDA(cudaFuncSetAttribute(stem_conv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    S2P_CHECK_CUDA(cudaFuncSetAttribute(stem_conv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_configured = smem;
  }
  if (image_fmt == S2P_IMG_U8_NHWC)
    stem_conv_kernel<true><<<grid, threads, smem, st>>>(image, w_oihw, bias, static_cast<__nv_bfloat16*>(y),
                                                    static_cast<__nv_bfloat16*>(y_lo), y_pitch,
                                                    reinterpret_cast<float2*>(stats_partial), H, W, Cout, rows);
  else
    stem_conv_kernel<false><<<grid, threads, smem, st>>>(image, w_oihw, bias, static_cast<__nv_bfloat16*>(y),
                                                     static_cast<__nv_bfloat16*>(y_lo), y_pitch,
                                                     reinterpret_cast<float2*>(stats_partial), H, W, Cout, rows);
  count_launch();
  return check_launch("stem_conv_kernel");
}

extern "C" int s2p_torgb_tanh_fwd(const void* x, int x_pitch, const float* w_oihw, const float* bias,
                                  float* out_f32_nchw, uint8_t* out_u8_nhwc, int B, int H, int W, int Cin,
                                  void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && w_oihw, "torgb: null pointer");
  S2P_CHECK_ARG(out_f32_nchw || out_u8_nhwc, "torgb: no output buffer given");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0, "torgb: bad dims");
  S2P_CHECK_ARG(x_pitch >= Cin && x_pitch % 8 == 0, "torgb: x_pitch=%d", x_pitch);
  const long long total = static_cast<long long>(B) * H * W;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(x);
  const size_t smem = 9 * Cin * 3 * sizeof(float);
#define S2P_TORGB_CASE(LPP)                                                                          \
  case LPP * 8: {                                                                                    \
    const long long blocks = (total + (256 / LPP) - 1) / (256 / LPP);                                \
    torgb_tanh_kernel<LPP><<<static_cast<unsigned>(blocks), 256, smem, st>>>(xb, x_pitch, w_oihw, bias, \
                                                                            out_f32_nchw, out_u8_nhwc, B, H, W); \
  } break;
  switch (Cin) {
    S2P_TORGB_CASE(4)
    S2P_TORGB_CASE(8)
    S2P_TORGB_CASE(16)
    default: return set_error(S2P_ERR_UNSUPPORTED, "torgb: Cin=%d (supported 32, 64, 128)", Cin);
  }
#undef S2P_TORGB_CASE
  count_launch();
  return check_launch("torgb_tanh_kernel");
}

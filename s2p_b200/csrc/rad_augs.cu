// RAD image augmentations (SURVEY.md section 8 f, rank 4): the "naive augmentation" comparator of S2P
// (/root/reference/examples/iql/data_augs.py; imported as `rad` by
// rlkit/launchers/experiments/awac/finetune_rl.py:34).  The reference runs the index-moving ones
// (random_crop :8-23, random_cutout :71-89, random_cutout_color :91-113, random_translate :244-254) as a
// per-image Python loop over numpy arrays on the host, and random_flip :117-146 / random_rotation :150-187 /
// random_grayscale :41-66 as 4-10 full-size torch passes (flip / rot90 copies of the whole batch, mask
// broadcasts, multiply-adds).  Here every function is ONE pass: an output element computes where it comes
// from out of the per-image draw (offset, box, flag, quarter turns), so the batch is read once and written once.
// Byte work, HBM-bound: algorithmic bytes per image = (C*H*W read + C*Ho*Wo written) * sizeof(element).
// The random draws stay on the host (np.random, same order as the reference: s2p_b200/data_augs.py) and arrive
// as int32 arrays, so a seeded run moves the same pixels.
#include <stdint.h>

#include "s2p_internal.h"
#include "ptx.cuh"

namespace s2p {
namespace rad {

constexpr int kThreads = 256;

template <typename T> struct Quad;
template <> struct Quad<uint8_t> { using type = uint32_t; };
template <> struct Quad<float> { using type = float4; };

__device__ __forceinline__ uint32_t pack4(const uint8_t* v) {
  return static_cast<uint32_t>(v[0]) | (static_cast<uint32_t>(v[1]) << 8) | (static_cast<uint32_t>(v[2]) << 16) |
         (static_cast<uint32_t>(v[3]) << 24);
}
__device__ __forceinline__ float4 pack4(const float* v) { return make_float4(v[0], v[1], v[2], v[3]); }

// One thread = kQuads groups of four consecutive output elements of one row of image blockIdx.y (the index
// arithmetic and the per-image draws are paid once per 8 elements; 4 groups per thread measured slower: the warp's accesses stop coalescing: the kernel is issue-bound, not load-bound).
constexpr int kQuads = 2;
template <typename T, int OP>
__global__ void __launch_bounds__(kThreads)
augment_kernel(const T* __restrict__ in, T* __restrict__ out, int C, int H, int W, int Ho, int Wo,
               const int* __restrict__ p0, const int* __restrict__ p1, const T* __restrict__ fill, int vec_store) {
  const unsigned quads_per_row = static_cast<unsigned>(Wo + 4 * kQuads - 1) / (4 * kQuads);   // thread slots per row
  const unsigned q = blockIdx.x * kThreads + threadIdx.x;   // < 2^31 (checked by the host): 32-bit divisions only
  if (q >= static_cast<unsigned>(C) * Ho * quads_per_row) return;
  const int xq = static_cast<int>(q % quads_per_row);
  const unsigned cy = q / quads_per_row;
  const int y = static_cast<int>(cy % Ho), c = static_cast<int>(cy / Ho);
  const long long n = blockIdx.y;
  int a = __ldg(p0 + n);
  int b = p1 ? __ldg(p1 + n) : 0;
  const T* src = in + (n * C + c) * static_cast<long long>(H) * W;
  T* dst = out + ((n * C + c) * static_cast<long long>(Ho) + y) * Wo;
#pragma unroll
  for (int hq = 0; hq < kQuads; ++hq) {
  const int x0 = (xq * kQuads + hq) << 2;
  if (x0 >= Wo) break;
  T v[4];
  if (OP == S2P_RAD_CROP) {            // out[y][x] = in[y + h1][x + w1]; offsets clamped: never read outside
    a = min(max(a, 0), H - Ho);
    b = min(max(b, 0), W - Wo);
    const T* row = src + static_cast<long long>(y + a) * W + b;
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = (x0 + i < Wo) ? __ldg(row + x0 + i) : T(0);
  } else if (OP == S2P_RAD_TRANSLATE) { // out[h1 + y][w1 + x] = in[y][x], zero elsewhere
    const int sy = y - a;
    const bool row_in = sy >= 0 && sy < H;
    const T* row = src + static_cast<long long>(row_in ? sy : 0) * W;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int sx = x0 + i - b;
      v[i] = (row_in && sx >= 0 && sx < W) ? __ldg(row + sx) : T(0);
    }
  } else if (OP == S2P_RAD_CUTOUT) {    // box rows [h1, 2*h1) x columns [w1, 2*w1) (the reference's slice), numpy-clipped
    const bool row_in = y >= a && y < 2 * a;
    const T f = fill ? __ldg(fill + n * C + c) : T(0);
    const T* row = src + static_cast<long long>(y) * W;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = x0 + i;
      const bool boxed = row_in && x >= b && x < 2 * b;
      v[i] = boxed ? f : ((x < W) ? __ldg(row + x) : T(0));
    }
  } else if (OP == S2P_RAD_FLIP) {      // flag ? in[y][W-1-x] : in[y][x]
    const T* row = src + static_cast<long long>(y) * W;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = x0 + i;
      v[i] = (x < W) ? __ldg(row + (a ? W - 1 - x : x)) : T(0);
    }
  } else {                              // S2P_RAD_ROT90: a quarter turns counter-clockwise (torch.rot90(k, [2, 3])), H == W
    a &= 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x = x0 + i;
      int sy = y, sx = x;
      if (a == 1) { sy = x; sx = W - 1 - y; }
      else if (a == 2) { sy = H - 1 - y; sx = W - 1 - x; }
      else if (a == 3) { sy = H - 1 - x; sx = y; }
      v[i] = (x < W) ? __ldg(src + static_cast<long long>(sy) * W + sx) : T(0);
    }
  }
  if (vec_store) {
    __stcs(reinterpret_cast<typename Quad<T>::type*>(dst + x0), pack4(v));   // written once, not re-read: evict first
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (x0 + i < Wo) dst[x0 + i] = v[i];
  }
  }
}

// random_grayscale (data_augs.py:26-66) in one pass, with the reference's roundings: u8 = trunc(x * 255);
// gray = trunc(u8_r * 0.2989f + u8_g * 0.587f + u8_b * 0.114f) (separately rounded fp32 products and sums);
// out = (flag ? gray : u8) / 255.
__device__ __forceinline__ void gray_px(float r, float g, float b, bool on, float& o0, float& o1, float& o2) {
  r = static_cast<float>(static_cast<uint8_t>(__fmul_rn(r, 255.f)));
  g = static_cast<float>(static_cast<uint8_t>(__fmul_rn(g, 255.f)));
  b = static_cast<float>(static_cast<uint8_t>(__fmul_rn(b, 255.f)));
  if (on) {
    const float s = __fadd_rn(__fadd_rn(__fmul_rn(r, 0.2989f), __fmul_rn(g, 0.587f)), __fmul_rn(b, 0.114f));
    r = g = b = static_cast<float>(static_cast<uint8_t>(s));
  }
  o0 = __fdiv_rn(r, 255.f); o1 = __fdiv_rn(g, 255.f); o2 = __fdiv_rn(b, 255.f);
}

// VEC = 4: H*W % 4 == 0 and 16-B aligned buffers, a thread owns four pixels of one frame (3 x 128-bit loads / stores)
template <int VEC>
__global__ void __launch_bounds__(kThreads)
grayscale_kernel(const float* __restrict__ in, float* __restrict__ out, int frames, int HW, const int* __restrict__ mask) {
  const unsigned per_frame = static_cast<unsigned>(HW) / VEC;
  const unsigned t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= static_cast<unsigned>(frames) * per_frame) return;
  const unsigned f = t / per_frame, p = (t - f * per_frame) * VEC;
  const long long n = blockIdx.y;
  const long long base = ((n * frames + f) * 3) * static_cast<long long>(HW) + p;
  const bool on = __ldg(mask + n) != 0;
  if (VEC == 4) {
    const float4 r = __ldcs(reinterpret_cast<const float4*>(in + base));
    const float4 g = __ldcs(reinterpret_cast<const float4*>(in + base + HW));
    const float4 b = __ldcs(reinterpret_cast<const float4*>(in + base + 2ll * HW));
    float4 o0, o1, o2;
    gray_px(r.x, g.x, b.x, on, o0.x, o1.x, o2.x);
    gray_px(r.y, g.y, b.y, on, o0.y, o1.y, o2.y);
    gray_px(r.z, g.z, b.z, on, o0.z, o1.z, o2.z);
    gray_px(r.w, g.w, b.w, on, o0.w, o1.w, o2.w);
    __stcs(reinterpret_cast<float4*>(out + base), o0);
    __stcs(reinterpret_cast<float4*>(out + base + HW), o1);
    __stcs(reinterpret_cast<float4*>(out + base + 2ll * HW), o2);
  } else {
    float o0, o1, o2;
    gray_px(__ldg(in + base), __ldg(in + base + HW), __ldg(in + base + 2ll * HW), on, o0, o1, o2);
    out[base] = o0; out[base + HW] = o1; out[base + 2ll * HW] = o2;
  }
}

// grayscale (data_augs.py:26-38), the public helper: values as they are (uint8 or float), per frame
// gray = float(uint8(r * 0.2989f + g * 0.587f + b * 0.114f)) tiled over the three channels of a NEW axis:
// out [N][frames][3][H][W] float32.  One thread = one pixel of one frame (three loads, three stores).
template <typename T>
__global__ void __launch_bounds__(kThreads)
grayscale_plain_kernel(const T* __restrict__ in, float* __restrict__ out, int frames, int HW) {
  const unsigned t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= static_cast<unsigned>(frames) * HW) return;
  const unsigned f = t / HW, p = t - f * HW;
  const long long base = ((static_cast<long long>(blockIdx.y) * frames + f) * 3) * HW + p;
  const float r = static_cast<float>(in[base]), g = static_cast<float>(in[base + HW]), b = static_cast<float>(in[base + 2ll * HW]);
  const float s = __fadd_rn(__fadd_rn(__fmul_rn(r, 0.2989f), __fmul_rn(g, 0.587f)), __fmul_rn(b, 0.114f));
  const float v = static_cast<float>(static_cast<uint8_t>(s));
  __stcs(out + base, v); __stcs(out + base + HW, v); __stcs(out + base + 2ll * HW, v);
}

// random_convolution (data_augs.py:193-223, "network randomization"): every image of the batch gets its OWN
// 3 -> 3 channel 3x3 convolution (padding 1, no bias), applied to each of its frames (3 channels per frame).  The
// reference runs B separate nn.Conv2d calls and concatenates their outputs one by one; here the B weight sets
// (drawn on the host side exactly as the reference draws them, s2p_b200/data_augs.py) arrive as one table and the
// whole batch is one launch.  One thread = 4 consecutive pixels of a row x 3 output channels of one frame.
__global__ void __launch_bounds__(kThreads)
random_conv_kernel(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ weights,
                   int frames, int H, int W) {
  __shared__ float ws[81];
  const long long n = blockIdx.y;
  if (threadIdx.x < 81) ws[threadIdx.x] = __ldg(weights + n * 81 + threadIdx.x);   // [co][ci][r][s]
  __syncthreads();
  const unsigned quads = static_cast<unsigned>(W + 3) / 4;
  const unsigned t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= static_cast<unsigned>(frames) * H * quads) return;
  const int xq = static_cast<int>(t % quads);
  const unsigned fy = t / quads;
  const int y = static_cast<int>(fy % H), f = static_cast<int>(fy / H);
  const int x0 = xq * 4;
  const long long HW = static_cast<long long>(H) * W;
  const float* src = in + (n * frames + f) * 3 * HW;
  float acc[3][4];
#pragma unroll
  for (int co = 0; co < 3; ++co)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[co][i] = 0.f;
#pragma unroll
  for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int yy = y + r - 1;
      float v[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const int xx = x0 + j - 1;
        v[j] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(src + ci * HW + static_cast<long long>(yy) * W + xx) : 0.f;
      }
#pragma unroll
      for (int s = 0; s < 3; ++s)
#pragma unroll
        for (int co = 0; co < 3; ++co) {
          const float wv = ws[((co * 3 + ci) * 3 + r) * 3 + s];
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[co][i] = fmaf(v[i + s], wv, acc[co][i]);
        }
    }
  }
  float* dst = out + (n * frames + f) * 3 * HW + static_cast<long long>(y) * W + x0;
#pragma unroll
  for (int co = 0; co < 3; ++co)
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (x0 + i < W) dst[co * HW + i] = acc[co][i];
}

// uint8 words path (rows of whole 32-bit words, 4-B aligned buffers; taken when the staged path below cannot be:
// input not 16-B aligned, or planes larger than the staging budget): one thread = 16 consecutive output bytes of a
// row.  The source run of a crop / translate / cutout / flip / half turn is contiguous in memory at an arbitrary byte
// offset: five ALIGNED 32-bit loads cover it and funnel shifts realign it (a byte-reversing permute for the mirrored
// cases), so a thread issues 5 loads + 1..4 stores per 16 bytes instead of 16 byte loads.  Words outside the input
// buffer are not loaded (they can only feed bytes the op masks or never stores).  Quarter turns walk a column: byte loads.
constexpr int kSegs = 1;   // 16-byte segments per thread (2 measured slower for crop / flip / quarter turns, 5-10 % faster for translate / cutout)
template <int OP>
__global__ void __launch_bounds__(kThreads)
augment_u8_words_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int C, int H, int W, int Ho, int Wo,
                        const int* __restrict__ p0, const int* __restrict__ p1, const uint8_t* __restrict__ fill,
                        long long in_bytes, long long n_base) {
  const unsigned segs = static_cast<unsigned>(Wo + 16 * kSegs - 1) / (16 * kSegs);   // thread slots per row
  const unsigned t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= static_cast<unsigned>(C) * Ho * segs) return;
  const unsigned seg = t % segs, cy = t / segs;   // (a 2-D launch geometry without these divisions measured 15 % slower: idle lanes)
  const int y = static_cast<int>(cy % Ho), c = static_cast<int>(cy / Ho);
  const long long n = n_base + blockIdx.y;
  int a = __ldg(p0 + n);
  int b = p1 ? __ldg(p1 + n) : 0;
  const long long plane = (n * C + c) * static_cast<long long>(H) * W;   // byte offset of this channel plane
  int srow = y, off = 0;
  bool reversed = false, column = false, row_in = true;
  if (OP == S2P_RAD_CROP) {
    a = min(max(a, 0), H - Ho);
    b = min(max(b, 0), W - Wo);
    srow = y + a; off = b;
  } else if (OP == S2P_RAD_TRANSLATE) {
    srow = y - a; off = -b;
    row_in = srow >= 0 && srow < H;
    if (!row_in) srow = 0;
  } else if (OP == S2P_RAD_FLIP) {
    reversed = a != 0;
  } else if (OP == S2P_RAD_ROT90) {
    a &= 3;
    if (a == 2) { srow = H - 1 - y; reversed = true; }
    column = (a & 1) != 0;
  }
#pragma unroll
  for (int hs = 0; hs < kSegs; ++hs) {
  const int x0 = static_cast<int>(seg * kSegs + hs) << 4;
  if (x0 >= Wo) break;
  const int nw = min(4, (Wo - x0) >> 2);
  uint32_t o[4];
  if (!column) {
    const long long s = plane + static_cast<long long>(srow) * W + (reversed ? (W - x0 - 16) : (off + x0));
    const int sh = static_cast<int>(s & 3) << 3;
    const long long sa = s - (s & 3);
    // a translate's segment outside the pasted image is all zero: no loads
    const bool live = OP != S2P_RAD_TRANSLATE || (row_in && x0 + 16 > b && x0 < b + W);
    uint32_t w[5] = {0u, 0u, 0u, 0u, 0u};
    if (live) {
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(in + sa);
      if (sa >= 0 && sa + 20 <= in_bytes) {   // the usual case: one range check for the five words
#pragma unroll
        for (int k = 0; k < 5; ++k) w[k] = __ldg(wp + k);
      } else {
#pragma unroll
        for (int k = 0; k < 5; ++k)
          if (sa + 4 * k >= 0 && sa + 4 * k + 4 <= in_bytes) w[k] = __ldg(wp + k);
      }
    }
    uint32_t run[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) run[j] = __funnelshift_r(w[j], w[j + 1], sh);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = reversed ? __byte_perm(run[3 - j], 0u, 0x0123) : run[j];
    if (OP == S2P_RAD_TRANSLATE || OP == S2P_RAD_CUTOUT) {
      // bytes [lo, hi) of the row: kept (translate: the pasted image) or replaced (cutout: the box)
      int lo, hi;
      uint32_t fw = 0u;
      if (OP == S2P_RAD_TRANSLATE) { lo = row_in ? b : 0; hi = row_in ? b + W : 0; }
      else {
        const bool boxed_row = y >= a && y < 2 * a;
        lo = boxed_row ? b : 0; hi = boxed_row ? 2 * b : 0;
        fw = fill ? static_cast<uint32_t>(__ldg(fill + n * C + c)) * 0x01010101u : 0u;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // bytes of word j inside [lo, hi): byte positions [s0, s1) of the word
        const int xj = x0 + 4 * j;
        const int s0 = min(max(lo - xj, 0), 4), s1 = min(max(hi - xj, 0), 4);
        const uint32_t m = s1 > s0 ? ((0xffffffffu >> (8 * (4 - (s1 - s0)))) << (8 * s0)) : 0u;
        o[j] = (OP == S2P_RAD_TRANSLATE) ? (o[j] & m) : ((o[j] & ~m) | (fw & m));
      }
    }
  } else {
    const uint8_t* src = in + plane;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t v = 0u;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int x = x0 + 4 * j + i;
        if (x < W) {
          const int sy = (a == 1) ? x : H - 1 - x, sx = (a == 1) ? W - 1 - y : y;
          v |= static_cast<uint32_t>(__ldg(src + static_cast<long long>(sy) * W + sx)) << (8 * i);
        }
      }
      o[j] = v;
    }
  }
  uint8_t* dst = out + ((n * C + c) * static_cast<long long>(Ho) + y) * Wo + x0;
  if (nw == 4 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    __stcs(reinterpret_cast<uint4*>(dst), make_uint4(o[0], o[1], o[2], o[3]));
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < nw) __stcs(reinterpret_cast<uint32_t*>(dst) + j, o[j]);
  }
  }
}

// uint8 staged path (the default for word-aligned rows and 16-B aligned buffers): a CTA owns a few channel planes of
// one image.  One thread bulk-copies (cp.async.bulk) the rows the op reads into shared memory -- whole 16-B lines, no
// per-thread global loads -- and then lane l of a warp produces output word l of a 128-byte run: two conflict-free
// shared loads + a funnel shift realign the source bytes, and the store is one coalesced 128-B line per warp.  The words
// kernel above spends ~160 warp instructions and 9 global memory instructions per 16 bytes and is issue bound at
// 0.4-0.6 of the HBM peak; this one keeps the global side at one bulk copy per plane and one STG per word.
constexpr int kStagePad = 16;        // slack before the data: a translate reads up to 3 bytes before a row's first byte
constexpr int kStageMaxBytes = 40 * 1024;

__device__ __forceinline__ uint32_t lds_u32_unaligned(uint32_t saddr) {
  const uint32_t al = saddr & ~3u;
  uint32_t lo, hi;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lo) : "r"(al));
  asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(hi) : "r"(al));
  return __funnelshift_r(lo, hi, (saddr & 3u) << 3);
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}

template <int OP>
__global__ void __launch_bounds__(kThreads)
augment_u8_staged_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int C, int H, int W, int Ho, int Wo,
                         const int* __restrict__ p0, const int* __restrict__ p1, const uint8_t* __restrict__ fill,
                         long long in_bytes, int planes, int slot, long long n_base) {
  extern __shared__ __align__(16) uint8_t stage[];
  __shared__ __align__(8) uint64_t bar;
  const int c0 = static_cast<int>(blockIdx.x) * planes;
  const int np = min(planes, C - c0);
  const long long n = n_base + blockIdx.y;
  int a = __ldg(p0 + n);
  int b = p1 ? __ldg(p1 + n) : 0;
  int r0 = 0, rows = H;           // source rows staged per plane
  if (OP == S2P_RAD_CROP) {
    a = min(max(a, 0), H - Ho);
    b = min(max(b, 0), W - Wo);
    r0 = a; rows = Ho;
  } else if (OP == S2P_RAD_ROT90) {
    a &= 3;
  }
  const long long HW = static_cast<long long>(H) * W;
  const long long g_first = (n * C + c0) * HW + static_cast<long long>(r0) * W;   // first staged byte of plane c0
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
    uint32_t total = 0;
    for (int p = 0; p < np; ++p) {
      const long long g0 = g_first + p * HW, g1 = g0 + static_cast<long long>(rows) * W;
      const long long ga = g0 & ~15ll, gb = min((g1 + 15) & ~15ll, in_bytes);
      total += static_cast<uint32_t>(gb - ga);
    }
    mbar_arrive_expect_tx(&bar, total);
    for (int p = 0; p < np; ++p) {
      const long long g0 = g_first + p * HW, g1 = g0 + static_cast<long long>(rows) * W;
      const long long ga = g0 & ~15ll, gb = min((g1 + 15) & ~15ll, in_bytes);
      bulk_load_1d(stage + p * slot + kStagePad, in + ga, static_cast<uint32_t>(gb - ga), &bar);
    }
  }
  // thread -> (row yr, word xw) of a group of R = kThreads / (Wo / 4) output rows; the rows of a plane are contiguous in
  // memory, so consecutive threads store consecutive words.  A thread keeps its column for the whole kernel (column
  // masks and source columns are loop invariants) and steps R rows per iteration: no divisions in the loop.
  const int wpr = Wo >> 2;                                   // words per output row (host: 4 <= Wo <= 4 * kThreads)
  const int R = kThreads / wpr;
  const int yr = static_cast<int>(threadIdx.x) / wpr, x = (static_cast<int>(threadIdx.x) - yr * wpr) << 2;
  const bool active = yr < R;
  // per-thread column state
  int scol = x;                    // source byte column of this thread's word
  uint32_t cmask = 0xffffffffu;    // translate: bytes inside the pasted image; cutout: bytes inside the box
  if (OP == S2P_RAD_CROP) scol = b + x;
  if (OP == S2P_RAD_TRANSLATE) {
    const int s0 = min(max(b - x, 0), 4), s1 = min(max(b + W - x, 0), 4);
    cmask = s1 > s0 ? ((0xffffffffu >> (8 * (4 - (s1 - s0)))) << (8 * s0)) : 0u;
    scol = min(max(x - b, -3), W - 1);                      // clamped: dead words never load, edge words stay in the pads
  }
  if (OP == S2P_RAD_CUTOUT) {
    const int s0 = min(max(b - x, 0), 4), s1 = min(max(2 * b - x, 0), 4);
    cmask = s1 > s0 ? ((0xffffffffu >> (8 * (4 - (s1 - s0)))) << (8 * s0)) : 0u;
  }
  const bool mirrored = (OP == S2P_RAD_FLIP && a != 0) || (OP == S2P_RAD_ROT90 && a == 2);
  if (mirrored) scol = W - x - 4;
  const int plane_out = Ho * Wo;
  __syncthreads();                 // barrier initialised before anyone polls it
  mbar_wait(&bar, 0);
  if (!active) return;
#pragma unroll 1
  for (int p = 0; p < np; ++p) {
    const uint32_t base = smem_u32(stage + p * slot + kStagePad) + static_cast<uint32_t>((g_first + p * HW) & 15);
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + (n * C + c0 + p) * static_cast<long long>(plane_out) + yr * Wo + x);
    const int dstep = R * wpr;     // words per row group
    uint32_t fw = 0u;
    if (OP == S2P_RAD_CUTOUT) fw = (fill ? static_cast<uint32_t>(__ldg(fill + n * C + c0 + p)) * 0x01010101u : 0u) & cmask;
    if (OP == S2P_RAD_ROT90 && (a & 1)) {
      // quarter turns walk a column of the staged plane: out(y, x + i) = in(x + i, W - 1 - y) or in(H - 1 - x - i, y)
      const int sstep = (a == 1) ? W : -W;
      const uint32_t s0 = base + ((a == 1) ? x : H - 1 - x) * W;
      for (int y = yr; y < Ho; y += R, dst += dstep) {
        const uint32_t s = s0 + ((a == 1) ? W - 1 - y : y);
        __stcs(dst, lds_u8(s) | (lds_u8(s + sstep) << 8) | (lds_u8(s + 2 * sstep) << 16) | (lds_u8(s + 3 * sstep) << 24));
      }
      continue;
    }
    const uint32_t sb = base + scol;
    for (int y = yr; y < Ho; y += R, dst += dstep) {
      uint32_t o;
      if (OP == S2P_RAD_CROP || OP == S2P_RAD_FLIP) {
        o = lds_u32_unaligned(sb + y * W);
        if (OP == S2P_RAD_FLIP && mirrored) o = __byte_perm(o, 0u, 0x0123);
      } else if (OP == S2P_RAD_TRANSLATE) {
        const int sy = y - a;
        o = 0u;
        if (cmask != 0u && sy >= 0 && sy < H) o = lds_u32_unaligned(sb + sy * W) & cmask;
      } else if (OP == S2P_RAD_CUTOUT) {
        o = lds_u32_unaligned(sb + y * W);
        if (y >= a && y < 2 * a) o = (o & ~cmask) | fw;
      } else {
        o = mirrored ? __byte_perm(lds_u32_unaligned(sb + (H - 1 - y) * W), 0u, 0x0123) : lds_u32_unaligned(sb + y * W);
      }
      __stcs(dst, o);
    }
  }
}

template <int OP>
int launch_u8_staged(const uint8_t* in, uint8_t* out, long long N, int C, int H, int W, int Ho, int Wo, const int* p0,
                     const int* p1, const uint8_t* fill, int slot, cudaStream_t st) {
  int planes = kStageMaxBytes / slot;
  planes = planes < 1 ? 1 : (planes > C ? C : planes);
  const unsigned gx = static_cast<unsigned>((C + planes - 1) / planes);
  const long long in_bytes = N * C * static_cast<long long>(H) * W;
  static bool configured[kMaxDevices] = {};
  const int dev = current_device();
  if (dev < 0 || dev >= kMaxDevices) return set_error(S2P_ERR_CUDA, "rad_augment: bad device %d", dev);
  if (!configured[dev]) {          // ask for the large shared-memory carve-out: 7 CTAs of 30 KB instead of the default's 5
    cudaFuncSetAttribute(augment_u8_staged_kernel<OP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    configured[dev] = true;
  }
  for (long long n0 = 0; n0 < N; n0 += 65535) {
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    augment_u8_staged_kernel<OP><<<dim3(gx, ny), kThreads, static_cast<size_t>(planes) * slot, st>>>(
        in, out, C, H, W, Ho, Wo, p0, p1, fill, in_bytes, planes, slot, n0);
    count_launch();
    const int rc = check_launch("rad augment_u8_staged_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

template <int OP>
int launch_u8_words(const uint8_t* in, uint8_t* out, long long N, int C, int H, int W, int Ho, int Wo, const int* p0,
                    const int* p1, const uint8_t* fill, cudaStream_t st) {
  const long long threads = static_cast<long long>(C) * Ho * ((Wo + 16 * kSegs - 1) / (16 * kSegs));
  const unsigned gx = static_cast<unsigned>((threads + kThreads - 1) / kThreads);
  const long long in_bytes = N * C * static_cast<long long>(H) * W;
  for (long long n0 = 0; n0 < N; n0 += 65535) {
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    augment_u8_words_kernel<OP><<<dim3(gx, ny), kThreads, 0, st>>>(in, out, C, H, W, Ho, Wo, p0, p1, fill, in_bytes, n0);
    count_launch();
    const int rc = check_launch("rad augment_u8_words_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

template <typename T>
int launch(const T* in, T* out, long long N, int C, int H, int W, int Ho, int Wo, int op, const int* p0, const int* p1,
           const T* fill, cudaStream_t st) {
  const long long quads = static_cast<long long>(C) * Ho * ((Wo + 4 * kQuads - 1) / (4 * kQuads));
  const unsigned gx = static_cast<unsigned>((quads + kThreads - 1) / kThreads);
  const int vec = (Wo % 4 == 0) && (reinterpret_cast<uintptr_t>(out) % (4 * sizeof(T)) == 0);
  for (long long n0 = 0; n0 < N; n0 += 65535) {   // gridDim.y limit
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    const dim3 grid(gx, ny);
    const T* i = in + n0 * C * static_cast<long long>(H) * W;
    T* o = out + n0 * C * static_cast<long long>(Ho) * Wo;
    const int* q0 = p0 + n0;
    const int* q1 = p1 ? p1 + n0 : nullptr;
    const T* fl = fill ? fill + n0 * C : nullptr;
    switch (op) {
      case S2P_RAD_CROP: augment_kernel<T, S2P_RAD_CROP><<<grid, kThreads, 0, st>>>(i, o, C, H, W, Ho, Wo, q0, q1, fl, vec); break;
      case S2P_RAD_TRANSLATE: augment_kernel<T, S2P_RAD_TRANSLATE><<<grid, kThreads, 0, st>>>(i, o, C, H, W, Ho, Wo, q0, q1, fl, vec); break;
      case S2P_RAD_CUTOUT: augment_kernel<T, S2P_RAD_CUTOUT><<<grid, kThreads, 0, st>>>(i, o, C, H, W, Ho, Wo, q0, q1, fl, vec); break;
      case S2P_RAD_FLIP: augment_kernel<T, S2P_RAD_FLIP><<<grid, kThreads, 0, st>>>(i, o, C, H, W, Ho, Wo, q0, q1, fl, vec); break;
      default: augment_kernel<T, S2P_RAD_ROT90><<<grid, kThreads, 0, st>>>(i, o, C, H, W, Ho, Wo, q0, q1, fl, vec); break;
    }
    count_launch();
    const int rc = check_launch("rad augment_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

}  // namespace rad
}  // namespace s2p

extern "C" int s2p_rad_augment(const void* in, void* out, int elem_bytes, long long N, int C, int H, int W, int Ho, int Wo,
                               int op, const int* p0, const int* p1, const void* fill, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(in && out && p0, "rad_augment: null pointer");
  S2P_CHECK_ARG(in != out, "rad_augment: in-place is not supported (every op reads other pixels than it writes)");
  S2P_CHECK_ARG(N > 0 && C > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0, "rad_augment: bad dims");
  S2P_CHECK_ARG(static_cast<long long>(C) * Ho * ((Wo + 3) / 4) < (1ll << 31), "rad_augment: image too large");
  if (elem_bytes != 1 && elem_bytes != 4) return set_error(S2P_ERR_UNSUPPORTED, "rad_augment: elem_bytes=%d (1: uint8, 4: float32)", elem_bytes);
  switch (op) {
    case S2P_RAD_CROP:
      S2P_CHECK_ARG(p1 && Ho <= H && Wo <= W, "rad_augment(crop): needs p1 and Ho <= H, Wo <= W");
      break;
    case S2P_RAD_TRANSLATE:
      S2P_CHECK_ARG(p1 && Ho >= H && Wo >= W, "rad_augment(translate): needs p1 and Ho >= H, Wo >= W");
      break;
    case S2P_RAD_CUTOUT:
      S2P_CHECK_ARG(p1 && Ho == H && Wo == W, "rad_augment(cutout): needs p1 and Ho == H, Wo == W");
      break;
    case S2P_RAD_FLIP:
      S2P_CHECK_ARG(Ho == H && Wo == W, "rad_augment(flip): needs Ho == H, Wo == W");
      break;
    case S2P_RAD_ROT90:
      S2P_CHECK_ARG(Ho == H && Wo == W && H == W, "rad_augment(rot90): needs square images and Ho == H, Wo == W");
      break;
    default:
      return set_error(S2P_ERR_UNSUPPORTED, "rad_augment: op=%d", op);
  }
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (elem_bytes == 1) {
    const uint8_t* i8 = static_cast<const uint8_t*>(in);
    uint8_t* o8 = static_cast<uint8_t*>(out);
    const uint8_t* f8 = static_cast<const uint8_t*>(fill);
    const bool words = Wo % 4 == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 3) == 0 &&
                       (N * C * static_cast<long long>(H) * W) % 4 == 0;
    // staged rows: 16-B aligned input (bulk copies), planes that fit the staging budget, rows of at least one word
    const long long in_total = N * C * static_cast<long long>(H) * W;
    const int slot = rad::kStagePad + static_cast<int>((static_cast<long long>(H) * W + 15) & ~15ll) + 32;
    const bool staged = words && (reinterpret_cast<uintptr_t>(in) & 15) == 0 && in_total % 16 == 0 && W >= 4 && Wo <= 4 * rad::kThreads &&
                        static_cast<long long>(H) * W + 64 <= rad::kStageMaxBytes && C <= 65535 * 8 &&
                        (static_cast<long long>(Ho) * Wo) < (1ll << 30);
    if (staged) {
      switch (op) {
        case S2P_RAD_CROP: return rad::launch_u8_staged<S2P_RAD_CROP>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, slot, st);
        case S2P_RAD_TRANSLATE: return rad::launch_u8_staged<S2P_RAD_TRANSLATE>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, slot, st);
        case S2P_RAD_CUTOUT: return rad::launch_u8_staged<S2P_RAD_CUTOUT>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, slot, st);
        case S2P_RAD_FLIP: return rad::launch_u8_staged<S2P_RAD_FLIP>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, slot, st);
        default: return rad::launch_u8_staged<S2P_RAD_ROT90>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, slot, st);
      }
    }
    if (words) {
      switch (op) {
        case S2P_RAD_CROP: return rad::launch_u8_words<S2P_RAD_CROP>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, st);
        case S2P_RAD_TRANSLATE: return rad::launch_u8_words<S2P_RAD_TRANSLATE>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, st);
        case S2P_RAD_CUTOUT: return rad::launch_u8_words<S2P_RAD_CUTOUT>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, st);
        case S2P_RAD_FLIP: return rad::launch_u8_words<S2P_RAD_FLIP>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, st);
        default: return rad::launch_u8_words<S2P_RAD_ROT90>(i8, o8, N, C, H, W, Ho, Wo, p0, p1, f8, st);
      }
    }
    return rad::launch<uint8_t>(static_cast<const uint8_t*>(in), static_cast<uint8_t*>(out), N, C, H, W, Ho, Wo, op, p0, p1,
                                static_cast<const uint8_t*>(fill), st);
  }
  S2P_CHECK_ARG(((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 3) == 0, "rad_augment: float buffers not 4-B aligned");
  return rad::launch<float>(static_cast<const float*>(in), static_cast<float*>(out), N, C, H, W, Ho, Wo, op, p0, p1,
                            static_cast<const float*>(fill), st);
}

extern "C" int s2p_rad_grayscale(const float* in, float* out, long long N, int frames, int H, int W, const int* mask,
                                 void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(in && out && mask, "rad_grayscale: null pointer");
  S2P_CHECK_ARG(N > 0 && frames > 0 && H > 0 && W > 0, "rad_grayscale: bad dims");
  const int HW = H * W;
  const long long per = static_cast<long long>(frames) * HW;
  S2P_CHECK_ARG(per < (1ll << 31), "rad_grayscale: image too large");
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = HW % 4 == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  const long long threads = vec ? per / 4 : per;
  const unsigned gx = static_cast<unsigned>((threads + rad::kThreads - 1) / rad::kThreads);
  for (long long n0 = 0; n0 < N; n0 += 65535) {
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    if (vec) rad::grayscale_kernel<4><<<dim3(gx, ny), rad::kThreads, 0, st>>>(in + n0 * per * 3, out + n0 * per * 3, frames, HW, mask + n0);
    else rad::grayscale_kernel<1><<<dim3(gx, ny), rad::kThreads, 0, st>>>(in + n0 * per * 3, out + n0 * per * 3, frames, HW, mask + n0);
    count_launch();
    const int rc = check_launch("rad grayscale_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

extern "C" int s2p_rad_grayscale_plain(const void* in, int elem_bytes, float* out, long long N, int frames, int H, int W,
                                       void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(in && out, "rad_grayscale_plain: null pointer");
  S2P_CHECK_ARG(elem_bytes == 1 || elem_bytes == 4, "rad_grayscale_plain: elem_bytes=%d (uint8 or float32)", elem_bytes);
  S2P_CHECK_ARG(N > 0 && frames > 0 && H > 0 && W > 0, "rad_grayscale_plain: bad dims");
  const int HW = H * W;
  const long long per = static_cast<long long>(frames) * HW;
  S2P_CHECK_ARG(per < (1ll << 31), "rad_grayscale_plain: image too large");
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned gx = static_cast<unsigned>((per + rad::kThreads - 1) / rad::kThreads);
  for (long long n0 = 0; n0 < N; n0 += 65535) {
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    if (elem_bytes == 1)
      rad::grayscale_plain_kernel<uint8_t><<<dim3(gx, ny), rad::kThreads, 0, st>>>(static_cast<const uint8_t*>(in) + n0 * per * 3,
                                                                                  out + n0 * per * 3, frames, HW);
    else
      rad::grayscale_plain_kernel<float><<<dim3(gx, ny), rad::kThreads, 0, st>>>(static_cast<const float*>(in) + n0 * per * 3,
                                                                                out + n0 * per * 3, frames, HW);
    count_launch();
    const int rc = check_launch("rad grayscale_plain_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

extern "C" int s2p_rad_random_convolution(const float* in, float* out, const float* weights, long long N, int frames, int H,
                                          int W, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(in && out && weights, "rad_random_convolution: null pointer");
  S2P_CHECK_ARG(N > 0 && frames > 0 && H > 0 && W > 0, "rad_random_convolution: bad dims");
  S2P_CHECK_ARG(in != out, "rad_random_convolution: out must not alias in");
  const long long slots = static_cast<long long>(frames) * H * ((W + 3) / 4);
  S2P_CHECK_ARG(slots < (1ll << 31), "rad_random_convolution: image too large");
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned gx = static_cast<unsigned>((slots + rad::kThreads - 1) / rad::kThreads);
  const long long per = static_cast<long long>(frames) * 3 * H * W;
  for (long long n0 = 0; n0 < N; n0 += 65535) {
    const unsigned ny = static_cast<unsigned>(N - n0 < 65535 ? N - n0 : 65535);
    rad::random_conv_kernel<<<dim3(gx, ny), rad::kThreads, 0, st>>>(in + n0 * per, out + n0 * per, weights + n0 * 81, frames, H, W);
    count_launch();
    const int rc = check_launch("rad random_conv_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

// Weight gradient of the implicit-GEMM convolution on tcgen05 (training path, SURVEY.md
// section 8 a8: backward of the nn.Conv2d layers of generator and discriminator).
//
//   dW[co, r, s, ci] = sum_{n,ho,wo} dY[n, ho, wo, co] * X[n, ho*st - pad + r, wo*st - pad + s, ci]
//
// GEMM view per filter tap: D_tap[co, ci] = sum_pix A[co, pix] * B_tap[ci, pix] with K = pixels.
// Both operands are NHWC in HBM, i.e. the GEMM's M (co) / N (ci) index is the contiguous one:
// "MN-major" UMMA operands.  A TMA box {64 channels, tile_w, tile_h, 1} lands in shared memory as
// [pixel rows][128 B] with the 128B swizzle, which IS the canonical MN-major SWIZZLE_128B layout
// (64 MN elements per row, 8 K rows per 1024 B atom, atoms along K at SBO = 1024 B, the next 64
// MN elements at LBO = rows * 128 B), so no transpose pass over the activations is needed.
//
// CTA = (128-wide co tile, BN-wide ci tile, one filter row r): S accumulators of BN columns in
// TMEM (S*BN <= 512), looping over its share of the pixel tiles (split-K over gridDim.z);
// partial results go to fp32 [split][Cout][R][S][Cin] and are combined in a fixed order by
// wgrad_reduce_kernel (bit-stable: no atomics), which also writes the OIHW layout of nn.Conv2d.
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2..5 epilogue.
#include <stdlib.h>

#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

struct WgradParams {
  int B, Ho, Wo;            // dY spatial size
  int Cin, Cout, R, S, pad, stride;
  int tile_w, tile_h, tiles_x, tiles_y;
  int kb_rows;              // tile_w * tile_h (multiple of 16)
  int num_pix_tiles;        // B * tiles_x * tiles_y
  int stages;               // depth of the TMA -> MMA ring (2..8, as many as fit in shared memory)
  float* partial;           // [splits][Cout][R][S][Cin]
};

// MN-major, SWIZZLE_128B shared-memory descriptor: LBO = byte distance between 64-element MN
// blocks, SBO = byte distance between 8-row K groups (1024 B).
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// kind::f16, D=f32, A=B=bf16, both MN-major (bits 15 / 16)
__host__ __device__ constexpr uint32_t umma_idesc_bf16_mn(uint32_t M, uint32_t N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

constexpr int kWgMaxStages = 8;

template <int BN, int TAPS>
__global__ void __launch_bounds__(192, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                  const WgradParams p) {
  constexpr int kAtomsA = 2;              // 128 co = two 64-wide MN blocks
  constexpr int kAtomsB = BN / 64;
  constexpr uint32_t kTmemCols = (TAPS * BN <= 128) ? 128 : (TAPS * BN <= 256) ? 256 : 512;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int atom_bytes = p.kb_rows * 128;                       // one 64-channel block of one tile
  const int atom_stride = ((p.kb_rows + 15) / 16) * 16 * 128;   // whole K steps per 64-channel block (1024-B aligned)
  const int stage_bytes = (kAtomsA + TAPS * kAtomsB) * atom_stride;
  const int kWgStages = p.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgStages * stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kWgMaxStages;
  uint64_t* done_bar = bars + 2 * kWgMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  if (p.kb_rows % 16) {
    // pixel tiles whose row count is not a multiple of the UMMA K (16), e.g. 25 x 5 = 125: the K loop runs to the next
    // multiple and the rows TMA never writes must read as zeros in BOTH operands -> clear the ring once (generic
    // proxy), then make it visible to the async proxy before the first TMA / MMA
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = kWgStages * stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
    for (int i = 0; i < kWgStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int co0 = blockIdx.x * 128;
  const int ci_tiles = p.Cin / BN;
  const int ci0 = (blockIdx.y % ci_tiles) * BN;
  // a CTA owns filter row r and TAPS consecutive taps of it starting at s_base (TAPS == S: the whole row)
  const int s_groups = p.S / TAPS;
  const int rs = blockIdx.y / ci_tiles;
  const int r = rs / s_groups;
  const int s_base = (rs % s_groups) * TAPS;
  const int split = blockIdx.z, splits = gridDim.z;
  const int t_begin = static_cast<int>(static_cast<long long>(p.num_pix_tiles) * split / splits);
  const int t_end = static_cast<int>(static_cast<long long>(p.num_pix_tiles) * (split + 1) / splits);
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const uint32_t stage_tx = static_cast<uint32_t>((kAtomsA + TAPS * kAtomsB) * atom_bytes);

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int t = t_begin; t < t_end; ++t) {
      const int img = t / tiles_per_img, tt = t % tiles_per_img;
      const int h0 = (tt / p.tiles_x) * p.tile_h, w0 = (tt % p.tiles_x) * p.tile_w;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        uint8_t* s0 = smem + stage * stage_bytes;
        mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
#pragma unroll
        for (int a = 0; a < kAtomsA; ++a)
          tma_load_4d(s0 + a * atom_stride, &tmap_dy, &full_bar[stage], co0 + 64 * a, w0, h0, img);
#pragma unroll
        for (int s = 0; s < TAPS; ++s)
#pragma unroll
          for (int b = 0; b < kAtomsB; ++b)
            tma_load_4d(s0 + (kAtomsA + s * kAtomsB + b) * atom_stride, &tmap_x, &full_bar[stage], ci0 + 64 * b,
                        w0 * p.stride - p.pad + s_base + s, h0 * p.stride - p.pad + r, img);
      }
      __syncwarp();
      if (++stage == kWgStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // The taps of a stage are ONE wide MMA where they fit (N = G * BN <= 256): their X tiles sit in consecutive
    // 64-channel blocks of the stage (LBO apart, exactly like the blocks of a wider ci tile) and their accumulators in
    // consecutive TMEM columns, so a group of G taps is an MN-major B operand of G * BN columns.  The dY slice (A,
    // 128 x 16) is then read from shared memory once per group instead of once per tap: 3 taps x 64 ci go from
    // 3 x (4 + 2) KB to 4 + 6 KB of operand reads per 96 MMA clocks (192 -> 107 B/clk against the 128 B/clk port).
    constexpr int G = (256 / BN) < TAPS ? (256 / BN) : TAPS;          // taps per MMA
    constexpr uint32_t idesc_full = umma_idesc_bf16_mn(128, G * BN);
    constexpr int kRest = TAPS % G;                                    // taps of a last, narrower MMA
    constexpr uint32_t idesc_rest = umma_idesc_bf16_mn(128, (kRest ? kRest : G) * BN);
    const uint32_t lbo = static_cast<uint32_t>(atom_stride), sbo = 1024u;
    int stage = 0;
    uint32_t phase = 0;
    const int ksteps = (p.kb_rows + 15) / 16;
    for (int t = t_begin; t < t_end; ++t) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t s0 = smem_u32(smem + stage * stage_bytes);
        const uint64_t da = umma_desc_mn_sw128(s0, lbo, sbo);
        for (int s = 0; s < TAPS; s += G) {
          const uint64_t db = umma_desc_mn_sw128(s0 + (kAtomsA + s * kAtomsB) * atom_stride, lbo, sbo);
          const uint32_t idesc = (s + G <= TAPS) ? idesc_full : idesc_rest;
          for (int k = 0; k < ksteps; ++k) {
            // K = 16 pixel rows = two 8-row groups = 2048 B further into every 64-channel block
            umma_bf16(tmem_base + s * BN, da + static_cast<uint64_t>(k * (2048 >> 4)),
                      db + static_cast<uint64_t>(k * (2048 >> 4)), idesc, (t != t_begin || k != 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[stage]);
        if (t == t_end - 1) umma_commit(done_bar);
      }
      __syncwarp();
      if (++stage == kWgStages) { stage = 0; phase ^= 1; }
    }
  } else {
    // ---------------------------------------------------------------- epilogue: TMEM -> fp32 partials
    const int q = warp & 3;
    const int co = co0 + q * 32 + lane;
    if (t_end > t_begin) {
      mbar_wait(done_bar, 0);
      tc_fence_after();
    }
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    for (int s = 0; s < TAPS; ++s) {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t raw[32];
        if (t_end > t_begin) {
          tmem_ld32(t_row + s * BN + c * 32, raw);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) raw[j] = 0u;
        }
        if (co < p.Cout) {
          float4* dst = reinterpret_cast<float4*>(
              p.partial + (((static_cast<size_t>(split) * p.Cout + co) * p.R + r) * p.S + s_base + s) * p.Cin + ci0 + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]),
                                 __uint_as_float(raw[4 * j + 2]), __uint_as_float(raw[4 * j + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kTmemCols>(tmem_base);
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-pair form (cta_group::2): one 256-wide co tile x one 128-wide ci tile x TAPS taps of a filter row.
// Each CTA of the pair loads ITS 128 co of the dY tile and HALF (64 ci) of every tap's X tile; the leader issues
// M = 256 MMAs whose B operand takes N/2 columns from each CTA.  Per CTA and 16-pixel K step that is 4 KB (A) + 2 KB
// per tap (B) of operand reads and (32 + 16 * TAPS) KB of TMA fill per pixel tile instead of 4 + 4 KB and
// (32 + 32 * TAPS) KB: the single-CTA kernel's 128 x 128 MMAs need 250 B/clk of the 128 B/clk shared-memory port
// (54-58 % tensor pipe, profiles/r02_train_kernel_table.txt); the pair with three taps needs ~125 B/clk.
// Pixel tiles are <= 64 rows here (8 KB per 64-channel block) so that a stage of 2 + TAPS blocks leaves a 5-deep ring.
// Accumulator columns of a group of G taps (N = G * 128): [g * 64, g * 64 + 64) = tap g, ci 0..63 (CTA 0's X tiles),
// [G * 64 + g * 64, ...) = tap g, ci 64..127 (CTA 1's).
template <int TAPS>
__global__ void __launch_bounds__(192, 1)
conv_wgrad_pair_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                       const WgradParams p) {
  constexpr int kAtomsA = 2;              // this CTA's 128 co = two 64-wide MN blocks
  constexpr int BNT = 128;                // ci tile of the pair; 64 per CTA
  constexpr uint32_t kTmemCols = (TAPS * BNT <= 128) ? 128 : (TAPS * BNT <= 256) ? 256 : 512;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int atom_bytes = p.kb_rows * 128;
  const int atom_stride = ((p.kb_rows + 15) / 16) * 16 * 128;
  const int stage_bytes = (kAtomsA + TAPS) * atom_stride;
  const int kWgStages = p.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgStages * stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kWgMaxStages;
  uint64_t* done_bar = bars + 2 * kWgMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  if (p.kb_rows % 16) {
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = kWgStages * stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
    for (int i = 0; i < kWgStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers and zero-filled ring exist before anything lands in them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int co0 = static_cast<int>(blockIdx.x >> 1) * 256 + static_cast<int>(rank) * 128;
  const int ci_tiles = p.Cin / BNT;
  const int ci_pair0 = (blockIdx.y % ci_tiles) * BNT;
  const int ci0 = ci_pair0 + static_cast<int>(rank) * 64;     // the 64 ci this CTA feeds
  const int s_groups = p.S / TAPS;
  const int rs = blockIdx.y / ci_tiles;
  const int r = rs / s_groups;
  const int s_base = (rs % s_groups) * TAPS;
  const int split = blockIdx.z, splits = gridDim.z;
  const int t_begin = static_cast<int>(static_cast<long long>(p.num_pix_tiles) * split / splits);
  const int t_end = static_cast<int>(static_cast<long long>(p.num_pix_tiles) * (split + 1) / splits);
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const uint32_t stage_tx = 2u * static_cast<uint32_t>((kAtomsA + TAPS) * atom_bytes);   // both CTAs' boxes

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int t = t_begin; t < t_end; ++t) {
      const int img = t / tiles_per_img, tt = t % tiles_per_img;
      const int h0 = (tt / p.tiles_x) * p.tile_h, w0 = (tt % p.tiles_x) * p.tile_w;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        uint8_t* s0 = smem + stage * stage_bytes;
        if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
        const uint32_t lead_bar = map_to_cta(&full_bar[stage], 0);
#pragma unroll
        for (int a = 0; a < kAtomsA; ++a)
          tma_load_4d_pair(s0 + a * atom_stride, &tmap_dy, lead_bar, co0 + 64 * a, w0, h0, img);
#pragma unroll
        for (int s = 0; s < TAPS; ++s)
          tma_load_4d_pair(s0 + (kAtomsA + s) * atom_stride, &tmap_x, lead_bar, ci0,
                           w0 * p.stride - p.pad + s_base + s, h0 * p.stride - p.pad + r, img);
      }
      __syncwarp();
      if (++stage == kWgStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr int G = (256 / BNT) < TAPS ? (256 / BNT) : TAPS;         // taps per MMA (N = G * 128 <= 256)
      constexpr uint32_t idesc_full = umma_idesc_bf16_mn(256, G * BNT);
      constexpr int kRest = TAPS % G;
      constexpr uint32_t idesc_rest = umma_idesc_bf16_mn(256, (kRest ? kRest : G) * BNT);
      const uint32_t lbo = static_cast<uint32_t>(atom_stride), sbo = 1024u;
      int stage = 0;
      uint32_t phase = 0;
      const int ksteps = (p.kb_rows + 15) / 16;
      for (int t = t_begin; t < t_end; ++t) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t s0 = smem_u32(smem + stage * stage_bytes);
          const uint64_t da = umma_desc_mn_sw128(s0, lbo, sbo);
          for (int s = 0; s < TAPS; s += G) {
            const uint64_t db = umma_desc_mn_sw128(s0 + (kAtomsA + s) * atom_stride, lbo, sbo);
            const uint32_t idesc = (s + G <= TAPS) ? idesc_full : idesc_rest;
            for (int k = 0; k < ksteps; ++k)
              umma_bf16_pair(tmem_base + s * BNT, da + static_cast<uint64_t>(k * (2048 >> 4)),
                             db + static_cast<uint64_t>(k * (2048 >> 4)), idesc, (t != t_begin || k != 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage], 3);
          if (t == t_end - 1) umma_commit_pair(done_bar, 3);
        }
        __syncwarp();
        if (++stage == kWgStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue: this CTA's 128 co rows -> fp32 partials
    constexpr int G = (256 / BNT) < TAPS ? (256 / BNT) : TAPS;
    const int q = warp & 3;
    const int co = co0 + q * 32 + lane;
    if (t_end > t_begin) {
      mbar_wait(done_bar, 0);
      tc_fence_after();
    }
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    for (int s = 0; s < TAPS; s += G) {
      const int g_taps = (s + G <= TAPS) ? G : (TAPS - s);      // taps in this group; its N is g_taps * 128
#pragma unroll 1
      for (int c = 0; c < g_taps * BNT / 32; ++c) {
        uint32_t raw[32];
        if (t_end > t_begin) {
          tmem_ld32(t_row + s * BNT + c * 32, raw);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) raw[j] = 0u;
        }
        // column c * 32 of the group: half = which CTA's X tiles, then tap, then ci inside the 64-wide block
        const int col = c * 32;
        const int half = col / (g_taps * 64), in_half = col - half * (g_taps * 64);
        const int tap = s + in_half / 64, ci = ci_pair0 + half * 64 + (in_half % 64);
        if (co < p.Cout) {
          float4* dst = reinterpret_cast<float4*>(
              p.partial + (((static_cast<size_t>(split) * p.Cout + co) * p.R + r) * p.S + s_base + tap) * p.Cin + ci);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]),
                                 __uint_as_float(raw[4 * j + 2]), __uint_as_float(raw[4 * j + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair<kTmemCols>(tmem_base);
}

// dW_oihw[co][ci][r][s] (+)= sum_split partial[split][co][r][s][ci]   (fixed order)
__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw_oihw, int splits,
                                    int Cout, int Cin, int R, int S, int accumulate) {
  const long long total = static_cast<long long>(Cout) * R * S * Cin;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < splits; ++k) acc += partial[k * total + i];
    const int ci = static_cast<int>(i % Cin);
    long long t = i / Cin;
    const int s = static_cast<int>(t % S); t /= S;
    const int r = static_cast<int>(t % R);
    const int co = static_cast<int>(t / R);
    float* d = dw_oihw + ((static_cast<long long>(co) * Cin + ci) * R + r) * S + s;
    *d = accumulate ? *d + acc : acc;
  }
}

// The same reduction with coalesced stores: a warp owns 32 input channels of one output channel and keeps their RS
// taps in registers (coalesced reads along ci for every tap and split, the same fixed split order), then transposes
// [RS][32] -> [32][RS] through shared memory: the 32 * RS floats of dW it produces are CONTIGUOUS in the OIHW layout.
// The scalar kernel above writes 4 bytes with a stride of RS * 4: 22 us for 22 MB (44 launches per training step).
template <int RS>
__global__ void __launch_bounds__(256)
wgrad_reduce_tile_kernel(const float* __restrict__ partial, float* __restrict__ dw_oihw, int splits, int Cout, int Cin,
                         int accumulate) {
  __shared__ float tile[8][RS * 32 + 1];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int ci_blocks = Cin / 32;
  const long long total = static_cast<long long>(Cout) * RS * Cin;
  const long long warps = static_cast<long long>(Cout) * ci_blocks;
  for (long long wv = blockIdx.x * 8ll + wid; wv < warps; wv += gridDim.x * 8ll) {
    const int co = static_cast<int>(wv / ci_blocks), cb = static_cast<int>(wv % ci_blocks);
    float acc[RS];
#pragma unroll
    for (int t = 0; t < RS; ++t) acc[t] = 0.f;
    const float* src = partial + (static_cast<long long>(co) * RS) * Cin + cb * 32 + lane;
    for (int k = 0; k < splits; ++k) {
#pragma unroll
      for (int t = 0; t < RS; ++t) acc[t] += __ldg(src + k * total + static_cast<long long>(t) * Cin);
    }
    float* ts = tile[wid];
#pragma unroll
    for (int t = 0; t < RS; ++t) ts[lane * RS + t] = acc[t];      // [ci][tap]: the OIHW order of this warp's block
    __syncwarp();
    float* dst = dw_oihw + (static_cast<long long>(co) * Cin + cb * 32) * RS;
#pragma unroll
    for (int t = 0; t < RS; ++t) {
      const int e = t * 32 + lane;
      dst[e] = accumulate ? dst[e] + ts[e] : ts[e];
    }
    __syncwarp();
  }
}

// pixel tile for the K dimension: tile_w * tile_h a multiple of 16, <= 128 rows, fewest wasted rows
static int wgrad_max_rows() {
  static int r = 0;
  if (r == 0) { const char* e = getenv("S2P_WGRAD_ROWS"); r = e ? atoi(e) : 128; if (r < 16) r = 16; if (r > 128) r = 128; }
  return r;
}

static int wgrad_tap_split() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("S2P_WGRAD_TAP_SPLIT"); v = e ? atoi(e) : 1; }
  return v;
}

static int wgrad_pad16() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("S2P_WGRAD_PAD16"); v = e ? atoi(e) : 1; }
  return v;
}

// Pixel tile (the GEMM K block): tile_w * tile_h <= 128 rows.  The MMA consumes K in steps of 16, so a tile costs
// ceil16(rows) rows of tensor time; rows need not be a multiple of 16 (the kernel zero-fills the slack), which is what
// lets 25 x 5 = 125 -> 128 tile 100^2, 50^2 and 25^2 maps at 97.7 % instead of 80-row tiles at 78-100 %.
static void wgrad_pick_tile(int H, int W, int* tw_out, int* th_out, int cap_rows = 128) {
  const int max_rows = wgrad_max_rows() < cap_rows ? wgrad_max_rows() : cap_rows;
  const bool pad_ok = wgrad_pad16() != 0;
  double best = -1.0;
  int btw = 16, bth = 1;
  for (int tw = 1; tw <= W && tw <= max_rows; ++tw) {
    for (int th = 1; th <= H && tw * th <= max_rows; ++th) {
      if (!pad_ok && (tw * th) % 16) continue;
      const int cost = (tw * th + 15) / 16 * 16;
      const long tiles = static_cast<long>((W + tw - 1) / tw) * ((H + th - 1) / th);
      const double eff = static_cast<double>(H) * W / (tiles * static_cast<double>(cost));
      // prefer efficiency, then bigger tiles (fewer pipeline turns).  Under a row cap (the pair kernel's 64) small
      // exact tiles would win on efficiency alone (4 x 4 = 16 rows at 100^2: one K step per ring slot); a slot's fixed
      // cost is worth about 8 rows of tensor time
      const double score = cap_rows < 128 ? eff * cost / (cost + 8.0) : eff + 1e-4 * cost;
      if (score > best) { best = score; btw = tw; bth = th; }
    }
  }
  *tw_out = btw;
  *th_out = bth;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn wg_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

struct WgradPlan {
  WgradParams p;
  int bn, taps, splits, smem_bytes;
  int pair;      // CTA-pair kernel (256 co x 128 ci per pair)
  dim3 grid;
};

static int wgrad_pairs() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("S2P_WGRAD_PAIRS"); v = e ? atoi(e) : 1; }
  return v;
}

static int wgrad_plan(int B, int Hin, int Win, int Cin, int Cout, int ksize, int stride, int pad, WgradPlan* pl) {
  WgradParams& p = pl->p;
  p = WgradParams{};
  p.B = B;
  p.Ho = (Hin + 2 * pad - ksize) / stride + 1;
  p.Wo = (Win + 2 * pad - ksize) / stride + 1;
  p.Cin = Cin; p.Cout = Cout; p.R = ksize; p.S = ksize; p.pad = pad; p.stride = stride;
  pl->pair = 0;
  if (wgrad_pairs() && Cout % 256 == 0 && Cin % 128 == 0 && (ksize == 3 || ksize == 1)) {
    // CTA pairs: 256 co x 128 ci x all taps of a filter row; pixel tiles of <= 64 rows keep a stage of 2 + taps
    // 64-channel blocks small enough for a deep ring
    pl->pair = 1;
    wgrad_pick_tile(p.Ho, p.Wo, &p.tile_w, &p.tile_h, ksize == 3 ? 64 : 128);
    p.tiles_x = (p.Wo + p.tile_w - 1) / p.tile_w;
    p.tiles_y = (p.Ho + p.tile_h - 1) / p.tile_h;
    p.kb_rows = p.tile_w * p.tile_h;
    p.num_pix_tiles = B * p.tiles_x * p.tiles_y;
    pl->taps = ksize;
    pl->bn = 128;
    const int ctas = 2 * (Cout / 256) * (Cin / 128) * ksize;
    int splits = num_sms() / ctas;
    if (splits > p.num_pix_tiles) splits = p.num_pix_tiles;
    if (splits < 1) splits = 1;
    pl->splits = splits;
    pl->grid = dim3(2 * (Cout / 256), (Cin / 128) * ksize, splits);
    const int atom_stride = ((p.kb_rows + 15) / 16) * 16 * 128;
    const int stage_bytes = (2 + ksize) * atom_stride;
    int stages = (200 * 1024) / stage_bytes;
    if (stages > kWgMaxStages) stages = kWgMaxStages;
    if (stages < 2) stages = 2;
    p.stages = stages;
    pl->smem_bytes = stages * stage_bytes + 1024 + 256;
    return S2P_OK;
  }
  wgrad_pick_tile(p.Ho, p.Wo, &p.tile_w, &p.tile_h);
  p.tiles_x = (p.Wo + p.tile_w - 1) / p.tile_w;
  p.tiles_y = (p.Ho + p.tile_h - 1) / p.tile_h;
  p.kb_rows = p.tile_w * p.tile_h;
  p.num_pix_tiles = B * p.tiles_x * p.tiles_y;
  pl->taps = ksize;
  const int atom = ((p.kb_rows + 15) / 16) * 16 * 128;
  pl->bn = (Cin % 128 == 0 && ksize * 128 <= 512 && 2 * (2 + ksize * 2) * atom <= 200 * 1024) ? 128 : 64;
  // big pixel tiles (25 x 5 = 125 rows, 16 KB per 64-channel block) leave no room for a 128-wide ci tile of all
  // three taps; a 128 x 64 MMA is operand-bound at ~46 % of the tensor pipe, so such layers run ONE tap per CTA
  // with the 128-wide tile instead (dY is then fetched once per tap, from L2)
  if (pl->bn == 64 && Cin % 128 == 0 && ksize == 3 && wgrad_tap_split()) { pl->bn = 128; pl->taps = 1; }
  const int ctas = ((Cout + 127) / 128) * (Cin / pl->bn) * ksize * (ksize / pl->taps);
  // split-K so that the grid is as close as possible to (but not above) ONE full wave: the kernel
  // runs one CTA per SM, so 2.03 waves cost 3
  int splits = num_sms() / ctas;
  if (splits > p.num_pix_tiles) splits = p.num_pix_tiles;
  if (splits < 1) splits = 1;
  pl->splits = splits;
  pl->grid = dim3((Cout + 127) / 128, (Cin / pl->bn) * ksize * (ksize / pl->taps), splits);
  const int atom_stride = ((p.kb_rows + 15) / 16) * 16 * 128;
  const int stage_bytes = (2 + pl->taps * (pl->bn / 64)) * atom_stride;
  int stages = (200 * 1024) / stage_bytes;
  if (stages > kWgMaxStages) stages = kWgMaxStages;
  if (stages < 2) stages = 2;
  p.stages = stages;
  pl->smem_bytes = stages * stage_bytes + 1024 + 256;
  return S2P_OK;
}

template <int BN, int TAPS>
static int launch_wgrad(const CUtensorMap& tdy, const CUtensorMap& tx, const WgradPlan& pl, cudaStream_t st) {
  static int configured[kMaxDevices] = {};   // per device
  const int dev = current_device();
  if (pl.smem_bytes > configured[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<BN, TAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        pl.smem_bytes));
    configured[dev] = pl.smem_bytes;
  }
  conv_wgrad_kernel<BN, TAPS><<<pl.grid, 192, pl.smem_bytes, st>>>(tdy, tx, pl.p);
  count_launch();
  return check_launch("conv_wgrad_kernel");
}

template <int TAPS>
static int launch_wgrad_pair(const CUtensorMap& tdy, const CUtensorMap& tx, const WgradPlan& pl, cudaStream_t st) {
  static int configured[kMaxDevices] = {};   // per device
  const int dev = current_device();
  if (pl.smem_bytes > configured[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_pair_kernel<TAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        pl.smem_bytes));
    configured[dev] = pl.smem_bytes;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = pl.grid;
  cfg.blockDim = dim3(192, 1, 1);
  cfg.dynamicSmemBytes = pl.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  S2P_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_wgrad_pair_kernel<TAPS>, tdy, tx, pl.p));
  count_launch();
  return check_launch("conv_wgrad_pair_kernel");
}

}  // namespace s2p

extern "C" long long s2p_conv_wgrad_workspace_bytes(int B, int Hin, int Win, int Cin, int Cout, int ksize,
                                                    int stride, int pad) {
  using namespace s2p;
  if (B <= 0 || Hin <= 0 || Win <= 0 || Cin <= 0 || Cin % 64 || Cout <= 0 || ksize < 1 || ksize > 4 || stride < 1)
    return -1;
  WgradPlan pl;
  wgrad_plan(B, Hin, Win, Cin, Cout, ksize, stride, pad, &pl);
  return static_cast<long long>(pl.splits) * Cout * ksize * ksize * Cin * 4;
}

extern "C" int s2p_conv_wgrad(const void* x, int x_pitch, const void* dy, int dy_pitch, float* dw_oihw,
                              void* workspace, long long workspace_bytes, int B, int Hin, int Win, int Cin,
                              int Cout, int ksize, int stride, int pad, int accumulate, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && dy && dw_oihw && workspace, "wgrad: null pointer");
  S2P_CHECK_ARG(B > 0 && Hin > 0 && Win > 0, "wgrad: bad dims");
  S2P_CHECK_ARG(Cin > 0 && Cin % 64 == 0, "wgrad: Cin=%d must be a multiple of 64", Cin);
  S2P_CHECK_ARG(Cout > 0 && Cout % 8 == 0, "wgrad: Cout=%d must be a multiple of 8", Cout);
  S2P_CHECK_ARG(ksize >= 1 && ksize <= 4 && (stride == 1 || stride == 2) && pad >= 0 && pad <= 3, "wgrad: k/stride/pad");
  S2P_CHECK_ARG(x_pitch >= Cin && x_pitch % 8 == 0 && dy_pitch >= Cout && dy_pitch % 8 == 0, "wgrad: pitch");
  S2P_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) |
                  reinterpret_cast<uintptr_t>(workspace)) & 15) == 0, "wgrad: 16-byte alignment");
  WgradPlan pl;
  wgrad_plan(B, Hin, Win, Cin, Cout, ksize, stride, pad, &pl);
  const long long need = static_cast<long long>(pl.splits) * Cout * ksize * ksize * Cin * 4;
  S2P_CHECK_ARG(workspace_bytes >= need, "wgrad: workspace %lld < %lld bytes", workspace_bytes, need);
  S2P_CHECK_ARG(pl.smem_bytes <= 227 * 1024, "wgrad: tile needs %d bytes of shared memory", pl.smem_bytes);
  EncodeTiledFn encode = wg_encode_fn();
  if (!encode) return set_error(S2P_ERR_CUDA, "cuTensorMapEncodeTiled unavailable");
  pl.p.partial = static_cast<float*>(workspace);
  const WgradParams& p = pl.p;
  CUtensorMap tdy, tx;
  {
    cuuint64_t dims[4] = {(cuuint64_t)Cout, (cuuint64_t)p.Wo, (cuuint64_t)p.Ho, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)dy_pitch * 2, (cuuint64_t)dy_pitch * 2 * p.Wo,
                             (cuuint64_t)dy_pitch * 2 * p.Wo * p.Ho};
    cuuint32_t box[4] = {64, (cuuint32_t)p.tile_w, (cuuint32_t)p.tile_h, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&tdy, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(dy), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "wgrad: dY tensor map encode failed (%d)", (int)r);
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)Cin, (cuuint64_t)Win, (cuuint64_t)Hin, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)x_pitch * 2, (cuuint64_t)x_pitch * 2 * Win,
                             (cuuint64_t)x_pitch * 2 * Win * Hin};
    cuuint32_t box[4] = {64, (cuuint32_t)(p.tile_w * stride), (cuuint32_t)(p.tile_h * stride), 1};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = encode(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "wgrad: x tensor map encode failed (%d)", (int)r);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc;
  if (pl.pair) {
    rc = pl.taps == 3 ? launch_wgrad_pair<3>(tdy, tx, pl, st) : launch_wgrad_pair<1>(tdy, tx, pl, st);
  } else if (pl.bn == 128) {
    switch (pl.taps) {
      case 1: rc = launch_wgrad<128, 1>(tdy, tx, pl, st); break;
      case 3: rc = launch_wgrad<128, 3>(tdy, tx, pl, st); break;
      case 4: rc = launch_wgrad<128, 4>(tdy, tx, pl, st); break;
      default: return set_error(S2P_ERR_UNSUPPORTED, "wgrad: ksize=%d", ksize);
    }
  } else {
    switch (pl.taps) {
      case 1: rc = launch_wgrad<64, 1>(tdy, tx, pl, st); break;
      case 3: rc = launch_wgrad<64, 3>(tdy, tx, pl, st); break;
      case 4: rc = launch_wgrad<64, 4>(tdy, tx, pl, st); break;
      default: return set_error(S2P_ERR_UNSUPPORTED, "wgrad: ksize=%d", ksize);
    }
  }
  if (rc != S2P_OK) return rc;
  const long long total = static_cast<long long>(Cout) * ksize * ksize * Cin;
  const int blocks = static_cast<int>((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048);
  const long long rwarps = static_cast<long long>(Cout) * (Cin / 32);
  const int rblocks = static_cast<int>((rwarps + 7) / 8 < 2048 ? (rwarps + 7) / 8 : 2048);
  const float* part = static_cast<const float*>(workspace);
  if (Cin % 32 == 0 && ksize == 3)
    wgrad_reduce_tile_kernel<9><<<rblocks, 256, 0, st>>>(part, dw_oihw, pl.splits, Cout, Cin, accumulate);
  else if (Cin % 32 == 0 && ksize == 4)
    wgrad_reduce_tile_kernel<16><<<rblocks, 256, 0, st>>>(part, dw_oihw, pl.splits, Cout, Cin, accumulate);
  else
    wgrad_reduce_kernel<<<blocks, 256, 0, st>>>(part, dw_oihw, pl.splits, Cout, Cin, ksize, ksize, accumulate);
  count_launch();
  return check_launch("wgrad_reduce_kernel");
}

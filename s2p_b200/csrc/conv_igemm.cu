// Implicit-GEMM convolution for sm_100a: TMA (4-D tiled tensor map over the NHWC
// activation, one box per filter tap) -> 128B-swizzled shared memory ->
// tcgen05.mma (M=128, N=BLOCK_N, K=16, bf16 x bf16 -> f32 in TMEM) -> tcgen05.ld
// epilogue (bias, residual add, activation, bf16 store, per-(n,c) statistics
// partials for the following instance norm).
//
// GEMM view:  D[m, co] = sum_{r,s,ci} X[n, ho*st - pad + r, wo*st - pad + s, ci] * Wp[co, (r,s,ci)]
//   m  = one output pixel of a (tile_h x tile_w) spatial tile of one image (<=128 rows)
//   K  = R*S*Cin walked as (tap, 64-channel chunk) blocks; padding = TMA out-of-bounds zero fill
//
// Replaces the cuDNN convs behind nn.Conv2d in the (absent) reference generator
// (oracle/s2p_oracle.py MATResBlk / MAT / ImageEncoder / head).
//
// Warp roles (384 threads): warps 0, 10, 11 = TMA producers (each fills every third ring stage), warp 1 = MMA issuer
// and owner of the TMEM allocation (both loops run warp-converged and elect one lane around the asynchronous issue),
// warps 2..9 = epilogue (TMEM lane quarter = warp % 4, two warps per quarter split the columns).
// CTA pairs (cta_group::2): two CTAs of a cluster walk two adjacent m-tiles of one n-tile as one
// M=256 UMMA, each loading its own activation rows and half of the weight tile.
// Persistent: grid = min(#tiles, #SMs); static round-robin tile schedule, n-tile fastest so
// CTAs running together share the activation tile in L2.  Accumulators are double
// buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1.
//
// Wide activations: a tensor the following instance norm / residual add reads (the residual stream x,
// the block-internal dx, the encoder's pre-norm e) can be stored as a bf16 PAIR hi + lo (lo = bf16(v - hi),
// ~16 mantissa bits) instead of one bf16: `out_lo`, `residual_lo`, `x_wide`.  Conv operands stay bf16.
#include <stdlib.h>

#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

struct ConvParams {
  int B, H, W;            // output spatial size
  int Cin, Cout;
  int R, S, pad, stride;
  int tile_w, tile_h, tiles_x, tiles_y;
  int mpitch;             // M rows per tile row: tile_w, or tile_w + 2 in the horizontal-halo mode
  int num_m_tiles, num_n_tiles;
  int out_pitch, res_pitch;
  int act;                // 0 none, 1 relu, 2 lrelu(0.2)
  const float* bias;
  const __nv_bfloat16* residual;
  const __nv_bfloat16* residual_lo;   // low halves of a wide residual (same pitch) or null
  __nv_bfloat16* out;
  __nv_bfloat16* out_lo;              // low halves of a wide output (same pitch) or null
  float* stats_partial;   // [B][tiles_per_img][Cout][2] (sum, sumsq) or null
  // fused MAT epilogue (conv_mat_fused): the accumulators are the spatial gamma|beta maps,
  // the stored value is act((x*A + Bc) * (G1 + gamma) + (B0 + beta)) with (A,Bc,G1,B0) = mat_tab[n][c]
  // the pre-norm activation x [B][H/up][W/up][..] arrives through tmap_x: per tile the producer warp TMA-loads the
  // tile's x pixels (64-channel boxes, 128B-swizzled; low halves of a wide x = the boxes at channel + mat_C) into
  // slots of the operand ring, and the epilogue reads them from shared memory
  const float4* mat_tab;        // [B][mat_C]
  int mat_up, mat_C;
  int x_wide;                   // x is a bf16 pair [hi | lo] (2 * mat_C channels)
  int x_bw;                     // pixels per row of an x box
  uint32_t x_box_bytes;         // bytes of one x box (x_bw * rows * 128)
  int cin_wrap;                 // A-operand channels >= cin_wrap re-read the tensor from channel 0 ([hi | lo] as [hi | lo | hi])
  int op_f16;                   // the activation and weight operands are fp16 (not bf16)
  // ToRGB epilogue (BLOCK_N = 16, 3 live columns): tanh(acc + bias) -> fp32 NCHW and/or uint8 NHWC
  float* rgb_f32;
  uint8_t* rgb_u8;
  // second operand (plain epilogue, stride 1): kb2 extra 64-channel k-blocks taken from a second NHWC tensor of the
  // output's spatial size through tmap_a2 -- a 1x1 convolution accumulated into the same tile, which is how a
  // res-block's learned shortcut rides along with its last 3x3 conv (weights concatenated along K)
  int kb2;
};

// kEpiSimple: the plain epilogue without residual, wide output and statistics (bias + activation + store only), as a
// kernel of its own: compiled into the general kernel as a second path it cost that path 3-9 % (register allocation).
enum { kEpiPlain = 0, kEpiMat = 1, kEpiToRGB = 2, kEpiSimple = 3 };

template <int BLOCK_N, int CL, int EPI, bool HH = false>
struct ConvCfg {
  static constexpr int kBlockM = 128;
  static constexpr int kBlockK = 64;
  static constexpr int kABytes = kBlockM * kBlockK * 2;        // 16 KB
  static constexpr int kBTapBytes = BLOCK_N / CL * kBlockK * 2;   // one tap's weight tile; a CTA of a pair holds half of it
  // HH (horizontal halo, thin tiles): a stage carries ONE activation box with a 1-pixel halo left and right and the
  // weight tiles of the three taps of a filter row that read it
  // (ToRGB folds the three taps of a filter row into the N dimension instead: one 48-row weight tile per stage)
  static constexpr int kBBytes = ((HH && EPI != kEpiToRGB) ? 3 : 1) * kBTapBytes;
  static constexpr int kStageBytes = kABytes + ((kBBytes + 1023) / 1024) * 1024;
  static constexpr int kStoreBytes = 8 * 2048;   // per-epilogue-warp 32 x 64 B transpose rows
  static constexpr int kTabBytes = 2 * 128 * 16;   // fused epilogue: double-buffered (A,Bc,G1,B0) rows
  // fused MAT epilogue: the tile's pre-norm activation x, TMA-loaded as 64-channel boxes of <= 16 KB
  // (BLOCK_N / 128 boxes of high halves + as many of low halves when x is stored wide)
  static constexpr int kXBoxes = EPI == kEpiMat ? 2 * (BLOCK_N / 128) : 0;
  static constexpr int kXBytes = kXBoxes * kABytes;
  // statistics: the four lane-quarter warps of a tile combine their (sum, sumsq) columns here before ONE partial row
  // per tile goes to global memory (double buffered by tile parity)
  static constexpr int kStatBytes = EPI == kEpiPlain ? 2 * 4 * BLOCK_N * 8 : 0;   // (kEpiSimple keeps no statistics)
  static constexpr int kFixedBytes = 1024 /*align*/ + 256 /*barriers*/ + kTabBytes + kStoreBytes + kXBytes + kStatBytes;
  static constexpr int kMaxStages = (227 * 1024 - kFixedBytes) / kStageBytes;
  // Ring depth.  Measured (profiles/r02_stage_sweep.txt): the wide tiles do not gain from more than 4-5 stages --
  // their operands come from L2 and 4 x 256 clk of look-ahead covers it -- while every KB of shared memory NOT
  // claimed is L1 for the epilogue's scattered accesses (BLOCK_N = 256: 6 -> 5 stages is +7 % on the store-bound
  // mlp_shared conv and neutral elsewhere).  The narrow tiles are operand-ingest bound and keep the deepest ring.
  static constexpr int kWant = BLOCK_N >= 256 ? (EPI == kEpiMat ? 4 : 5) : 8;
  static constexpr int kStages = kMaxStages > kWant ? kWant : kMaxStages;
  static constexpr int kAccStride = (BLOCK_N > 128) ? 256 : (BLOCK_N == 48 ? 64 : BLOCK_N);   // a power of two >= 32 per buffer pair
  static constexpr uint32_t kTmemCols = 2 * kAccStride;   // 32 / 128 / 256 / 512
  static constexpr int kSmemBytes = kStages * kStageBytes + kFixedBytes;
};

// transposed butterfly: on exit lane L holds sum over the warp's 32 rows of column L
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// Coalesced store of a warp's 32 accumulator rows x (NCH x 8) bf16 channels.  In the TMEM
// layout a thread owns one pixel row, so a direct store makes every warp-level st.global touch
// 32 different lines with 16 B each (LSU-bound on wide outputs).  The rows are transposed
// through a warp-private, XOR-swizzled (bank-conflict-free) shared-memory tile instead, and
// written back with NCH lanes per pixel: whole 32 B sectors, 32/NCH pixels per instruction.
// The reading side of warp_store_rows stores row rr = i * (32 / NCH) + lane / NCH in its i-th instruction: the
// destination pixel and validity of those rows are fetched from their owner lanes ONCE per tile (they are the same
// for every column chunk of the tile), not once per store instruction.
template <int NCH>
struct StoreRows {
  uint32_t pix[NCH];
  uint32_t valid_mask;
};
template <int NCH>
__device__ __forceinline__ StoreRows<NCH> store_rows_prepare(uint32_t pix, bool valid, int lane) {
  StoreRows<NCH> r;
  r.valid_mask = 0u;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int rr = i * (32 / NCH) + lane / NCH;
    r.pix[i] = __shfl_sync(0xffffffffu, pix, rr);
    r.valid_mask |= static_cast<uint32_t>(__shfl_sync(0xffffffffu, static_cast<int>(valid), rr)) << i;
  }
  return r;
}

template <int NCH>
__device__ __forceinline__ void warp_store_rows(uint32_t stg, const uint32_t* packed, __nv_bfloat16* out_col,
                                                int out_pitch, const StoreRows<NCH>& rows, int lane) {
  constexpr int kRow = NCH * 16;
  const int wkey = (NCH == 4) ? (lane >> 1) : (lane >> 2);
#pragma unroll
  for (int j = 0; j < NCH; ++j)
    sts128(stg + lane * kRow + ((j ^ wkey) & (NCH - 1)) * 16, packed[4 * j], packed[4 * j + 1], packed[4 * j + 2],
           packed[4 * j + 3]);
  __syncwarp();
  constexpr int kRowsPerInst = 32 / NCH;
  const int j = lane % NCH;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int rr = i * kRowsPerInst + lane / NCH;
    const int rkey = (NCH == 4) ? (rr >> 1) : (rr >> 2);
    const uint4 v = lds128(stg + rr * kRow + ((j ^ rkey) & (NCH - 1)) * 16);
    if ((rows.valid_mask >> i) & 1u)
      *reinterpret_cast<uint4*>(out_col + static_cast<size_t>(rows.pix[i]) * out_pitch + j * 8) = v;
  }
  __syncwarp();
}

// The store of the bias / activation-only epilogue (kEpiSimple), where the store phase is a third of a chunk's
// instructions: the destination of each of the NCH store instructions is kept as a 32-bit ELEMENT offset from the
// tile's image (pixel inside the image x pitch + this lane's 8-channel column; the host checks H * W * pitch < 2^32),
// so a store's address is one 64-bit add to a tile-uniform base instead of ten instructions of 64-bit index
// arithmetic, and the validity test is a predicate on the store itself (no divergent branch per instruction).
template <int NCH>
struct StoreOffs {
  uint32_t off[NCH];
  uint32_t valid_mask;
};
template <int NCH>
__device__ __forceinline__ StoreOffs<NCH> store_offs_prepare(uint32_t pix_in_image, bool valid, int lane, int out_pitch) {
  StoreOffs<NCH> r;
  r.valid_mask = 0u;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int rr = i * (32 / NCH) + lane / NCH;
    r.off[i] = __shfl_sync(0xffffffffu, pix_in_image, rr) * static_cast<uint32_t>(out_pitch) + static_cast<uint32_t>(lane % NCH) * 8u;
    r.valid_mask |= static_cast<uint32_t>(__shfl_sync(0xffffffffu, static_cast<int>(valid), rr)) << i;
  }
  return r;
}
__device__ __forceinline__ void stg128_if(void* ptr, const uint4& v, uint32_t pred) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\t@p st.global.v4.b32 [%0], {%1, %2, %3, %4};\n\t}\n" ::"l"(ptr),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(pred));   // (no memory clobber: nothing in the kernel reads `out` back)
}
template <int NCH>
__device__ __forceinline__ void warp_store_offs(uint32_t stg, const uint32_t* packed, __nv_bfloat16* out_col,
                                                const StoreOffs<NCH>& rows, int lane) {   // out_col: image base + channel
  constexpr int kRow = NCH * 16;
  const int wkey = (NCH == 4) ? (lane >> 1) : (lane >> 2);
#pragma unroll
  for (int j = 0; j < NCH; ++j)
    sts128(stg + lane * kRow + ((j ^ wkey) & (NCH - 1)) * 16, packed[4 * j], packed[4 * j + 1], packed[4 * j + 2],
           packed[4 * j + 3]);
  __syncwarp();
  constexpr int kRowsPerInst = 32 / NCH;
  const int j = lane % NCH;
  uint4 vv[NCH];             // all loads first, then the stores (both are volatile asm: program order is issue order)
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int rr = i * kRowsPerInst + lane / NCH;
    const int rkey = (NCH == 4) ? (rr >> 1) : (rr >> 2);
    vv[i] = lds128(stg + rr * kRow + ((j ^ rkey) & (NCH - 1)) * 16);
  }
#pragma unroll
  for (int i = 0; i < NCH; ++i) stg128_if(out_col + rows.off[i], vv[i], (rows.valid_mask >> i) & 1u);
  __syncwarp();
}

constexpr int kConvThreads = 384;   // 12 warps: TMA producer, MMA issuer, 8 epilogue warps, two more TMA producers

// Timeline probe (tools/conv_trace_build.sh builds a private copy of the library with -DS2P_CONV_TRACE; the product
// build has none of this): CTA 0 records clock64() at the pipeline hand-offs of its first kTraceN stages / tiles.
#ifdef S2P_CONV_TRACE
constexpr int kTraceN = 1024;
__device__ long long g_conv_trace[12][kTraceN];
#define S2P_TRACE(row, idx) do { if (blockIdx.x == 0 && (idx) < kTraceN) g_conv_trace[row][idx] = clock64(); } while (0)
#else
#define S2P_TRACE(row, idx) do { } while (0)
#endif

template <int BLOCK_N, int EPI, int CL, bool HH>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_a2, const __grid_constant__ CUtensorMap tmap_x,
                  const ConvParams p) {
  using Cfg = ConvCfg<BLOCK_N, CL, EPI, HH>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // [operand ring | x boxes (1024-aligned: the ring is a multiple of 8 KB) | barriers | modulation table | store staging]
  uint8_t* x_smem = smem + Cfg::kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(x_smem + Cfg::kXBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* x_full = tmem_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x_full + 1);
  float4* tab_smem = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(bars) + 256);
  uint8_t* store_smem = reinterpret_cast<uint8_t*>(bars) + 256 + Cfg::kTabBytes;
  float2* stat_smem = reinterpret_cast<float2*>(store_smem + Cfg::kStoreBytes);   // [2][4 quarters][BLOCK_N]

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    if (p.kb2 > 0) tma_prefetch_desc(&tmap_a2);
    if constexpr (EPI == kEpiMat) tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8 * CL);   // the leader's MMA thread waits for both CTAs' 8 epilogue warps
    }
    mbar_init(x_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    if constexpr (CL > 1) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    else tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();   // peers' barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // CL == 2: a CTA pair (cta_group::2) walks two adjacent m-tiles of the same n-tile as ONE
  // 256-row UMMA.  Each CTA loads its own 128 activation rows and HALF of the weight tile, so the
  // bytes entering each SM's shared memory per MMA drop from 48 KB to 32 KB per k-block (the
  // L2->SM ingest port, ~70 B/clk, is what bounds the single-CTA kernel).  The leader (rank 0)
  // issues the MMAs; its commits release the stages and publish the accumulators in both CTAs.
  // super tile = (pair of m-tiles, n-tile); an odd last m-tile gets a dummy partner (TMA zero
  // fill, nothing stored).
  const uint32_t cta_rank = (CL > 1) ? cluster_ctarank() : 0;
  const int num_tiles = ((p.num_m_tiles + CL - 1) / CL) * p.num_n_tiles;
  const int tile_first = blockIdx.x / CL, tile_step = gridDim.x / CL;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int cchunks = p.Cin / Cfg::kBlockK;
  const int num_kb_main = p.R * p.S * cchunks;
  const int num_kb = num_kb_main + p.kb2;      // the second operand's k-blocks follow the filter taps
  // HH: the M rows of a tile are laid out with pitch tile_w + 2 (the halo columns are rows whose results are
  // discarded), so that tap s of a filter row is the SAME activation box read s rows further down: a K-major
  // SWIZZLE_128B descriptor may start at any 128-byte row of a tile TMA wrote (profiles/r01_umma_rowshift_probe.txt).
  // One box per (filter row, channel chunk) feeds three taps: 2.8x fewer activation bytes enter the SM, which is what
  // bounds the thin (Cout <= 64) tiles; the price is 100 live rows of 128 instead of 125.
  // ToRGB in halo mode: the three taps of a filter row are folded into N (weight tile = 3 x 16 rows), the UMMAs read
  // the box un-shifted and the epilogue adds the row-shifted 16-column groups: 12 UMMAs per tile instead of 36, each of
  // which is bound by its 32-clk A-operand read whatever N is (profiles/r02_conv_trace.txt).
  constexpr bool kFold = (EPI == kEpiToRGB) && HH;
  const uint32_t a_bytes = static_cast<uint32_t>(p.mpitch * p.tile_h) * 128u;
  const uint32_t stage_tx = (a_bytes + Cfg::kBBytes) * CL;   // both CTAs' boxes signal the leader's barrier
  const int num_kb_hh = HH ? p.R * cchunks + p.kb2 : 0;      // HH: a k-block = (filter row, chunk): three taps

  // The producer and MMA loops run with the whole warp converged (uniform control flow, uniform
  // operands) and elect one lane only around the asynchronous issue itself: a loop written under
  // `lane == 0` forces the compiler to serialise every uniform-datapath operand through
  // elect/R2UR waterfall loops (~700 clk per k-block, which used to bound every layer).
  if (warp == 0 || warp >= 10) {
    // ------------------------------------------------------------ TMA producers (three warps)
    // The three warps walk the same (tile, k-block, ring stage) sequence and each fills every third stage.  One
    // thread issuing every stage took 330 clk per stage (barrier poll ~70, the uniform-datapath address arithmetic
    // and two UTMALDG ~170, loop ~90), 800 clk with the four loads of a halo stage: more than the 256 clk of MMA
    // work a BLOCK_N = 128 stage holds and 2-4x what a thin halo stage holds (tools/conv_trace.py,
    // profiles/r02_conv_trace.txt).  The stages still complete in ring order for the MMA thread: each has its own
    // full barrier, and a producer waits on its own stage's empty barrier only.
    const int my_turn = warp == 0 ? 0 : warp - 9;      // 0, 1, 2
    int turn = 0;
    int stage = 0;
    uint32_t phase = 0;
#ifdef S2P_CONV_TRACE
    int tr_p = 0;
#endif
    for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
      const int m_tile = (tile / p.num_n_tiles) * CL + cta_rank, n_tile = tile % p.num_n_tiles;
      const int img = m_tile / tiles_per_img, t = m_tile % tiles_per_img;   // img == B for a dummy tile
      const int h0 = (t / p.tiles_x) * p.tile_h, w0 = (t % p.tiles_x) * p.tile_w;
      const int hin0 = h0 * p.stride - p.pad, win0 = w0 * p.stride - p.pad;
      const int b_row = n_tile * BLOCK_N + (CL > 1 ? static_cast<int>(cta_rank) * (BLOCK_N / CL) : 0);   // first weight row of this CTA
      int r = 0, sx = 0, cc = 0;
      if constexpr (HH) {
        int rr = 0, ck = 0;          // (filter row, channel chunk) of the k-block; second-operand blocks follow
        for (int kb = 0; kb < num_kb_hh; ++kb) {
          if (turn == my_turn) {
            const bool main_kb = kb < p.R * cchunks;
            // everything the loads need is computed before the barrier poll
            uint8_t* sa = smem + stage * Cfg::kStageBytes;
            uint8_t* sb = sa + Cfg::kABytes;
            const int ntap = (main_kb && !kFold) ? 3 : 1;
            int ach = ck * Cfg::kBlockK;
            if (p.cin_wrap > 0 && ach >= p.cin_wrap) ach -= p.cin_wrap;
            const uint32_t tx = (a_bytes + static_cast<uint32_t>(ntap) * Cfg::kBTapBytes) * CL;
            const int kcol = kFold ? (rr * cchunks + ck) * Cfg::kBlockK
                                   : (main_kb ? (rr * 3 * cchunks + ck) * Cfg::kBlockK : (9 * cchunks + kb - p.R * cchunks) * Cfg::kBlockK);
            const int a2ch = (kb - p.R * cchunks) * Cfg::kBlockK;
            if (my_turn == 0) S2P_TRACE(0, tr_p);
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (my_turn == 0) S2P_TRACE(1, tr_p);
            if (elect_one()) {
              if constexpr (CL > 1) {
                if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], tx);
                const uint32_t lead_bar = map_to_cta(&full_bar[stage], 0);
                if (main_kb) tma_load_4d_pair(sa, &tmap_a, lead_bar, ach, w0 - 1, h0 - 1 + rr, img);
                else tma_load_4d_pair(sa, &tmap_a2, lead_bar, a2ch, w0 - 1, h0, img);
                for (int s = 0; s < ntap; ++s)
                  tma_load_2d_pair(sb + s * Cfg::kBTapBytes, &tmap_b, lead_bar, kcol + s * cchunks * Cfg::kBlockK, b_row);
              } else {
                mbar_arrive_expect_tx(&full_bar[stage], tx);
                if (main_kb) tma_load_4d(sa, &tmap_a, &full_bar[stage], ach, w0 - 1, h0 - 1 + rr, img);
                else tma_load_4d(sa, &tmap_a2, &full_bar[stage], a2ch, w0 - 1, h0, img);
                for (int s = 0; s < ntap; ++s)
                  tma_load_2d(sb + s * Cfg::kBTapBytes, &tmap_b, &full_bar[stage], kcol + s * cchunks * Cfg::kBlockK, b_row);
              }
            }
            __syncwarp();
            if (my_turn == 0) S2P_TRACE(2, tr_p);
#ifdef S2P_CONV_TRACE
            ++tr_p;
#endif
          }
          if (++turn == 3) turn = 0;
          if (++ck == cchunks) { ck = 0; ++rr; }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      } else {
        for (int kb = 0; kb < num_kb; ++kb) {
          if (turn == my_turn) {
            uint8_t* sa = smem + stage * Cfg::kStageBytes;
            uint8_t* sb = sa + Cfg::kABytes;
            const bool main_kb = kb < num_kb_main;
            int ach = cc * Cfg::kBlockK;
            if (p.cin_wrap > 0 && ach >= p.cin_wrap) ach -= p.cin_wrap;
            const int ca = main_kb ? ach : (kb - num_kb_main) * Cfg::kBlockK;
            const int cw = main_kb ? win0 + sx : w0, chh = main_kb ? hin0 + r : h0;
            if (my_turn == 0) S2P_TRACE(0, tr_p);
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (my_turn == 0) S2P_TRACE(1, tr_p);
            if (elect_one()) {
              if constexpr (CL > 1) {
                if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
                const uint32_t lead_bar = map_to_cta(&full_bar[stage], 0);
                if (main_kb) tma_load_4d_pair(sa, &tmap_a, lead_bar, ca, cw, chh, img);
                else tma_load_4d_pair(sa, &tmap_a2, lead_bar, ca, cw, chh, img);
                tma_load_2d_pair(sb, &tmap_b, lead_bar, kb * Cfg::kBlockK, b_row);
              } else {
                mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
                if (main_kb) tma_load_4d(sa, &tmap_a, &full_bar[stage], ca, cw, chh, img);
                else tma_load_4d(sa, &tmap_a2, &full_bar[stage], ca, cw, chh, img);
                tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * Cfg::kBlockK, b_row);
              }
            }
            __syncwarp();
            if (my_turn == 0) S2P_TRACE(2, tr_p);
#ifdef S2P_CONV_TRACE
            ++tr_p;
#endif
          }
          if (++turn == 3) turn = 0;
          if (++cc == cchunks) { cc = 0; if (++sx == p.S) { sx = 0; ++r; } }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (pair leader only)
    if (cta_rank == 0) {
      const uint32_t idesc = p.op_f16 ? umma_idesc_f16(128 * CL, BLOCK_N) : umma_idesc_bf16(128 * CL, BLOCK_N);
      const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
#ifdef S2P_CONV_TRACE
      int tr_m = 0;
#endif
      for (int tile = tile_first; tile < num_tiles; tile += tile_step, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        S2P_TRACE(6, it);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        S2P_TRACE(7, it);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * Cfg::kAccStride;
        // one ring stage's MMAs + the commit that frees the stage (run by the elected lane)
        auto issue_stage = [&](int stg, int kb, int nkb) {
          // descriptor start-address field is in 16-byte units: stage and operand offsets add directly
          const uint64_t da = desc0 + static_cast<uint64_t>(stg * (Cfg::kStageBytes >> 4));
          const uint64_t db = da + (Cfg::kABytes >> 4);
          if constexpr (HH) {
            const bool main_kb = kb < p.R * cchunks;
            // tap s reads the box s rows (s * 128 B = 8 descriptor units) further down; the second operand (a 1x1
            // convolution) reads its box at the centre column: one row down
            // (ToRGB fold: ONE UMMA group over the un-shifted box; the taps are 16-column groups of the accumulator)
            const int s_lo = (main_kb || kFold) ? 0 : 1, s_hi = kFold ? 1 : (main_kb ? 3 : 2);
            for (int s = s_lo; s < s_hi; ++s) {
              const uint64_t das = da + 8u * s, dbs = db + (main_kb ? s : 0) * (Cfg::kBTapBytes >> 4);
#pragma unroll
              for (int k = 0; k < Cfg::kBlockK / 16; ++k) {
                const uint32_t accum = (kb != 0 || s != s_lo || k != 0) ? 1u : 0u;
                if constexpr (CL > 1) umma_bf16_pair(d_tmem, das + 2 * k, dbs + 2 * k, idesc, accum);
                else umma_bf16(d_tmem, das + 2 * k, dbs + 2 * k, idesc, accum);
              }
            }
          } else {
#pragma unroll
            for (int k = 0; k < Cfg::kBlockK / 16; ++k) {
              // +32 B per K=16 step inside the 128 B swizzle row
              if constexpr (CL > 1) umma_bf16_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
              else umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
          }
          if constexpr (CL > 1) {
            umma_commit_pair(&empty_bar[stg], 3);
            if (kb == nkb - 1) umma_commit_pair(&tmem_full[acc], 3);
          } else {
            umma_commit(&empty_bar[stg]);
            if (kb == nkb - 1) umma_commit(&tmem_full[acc]);
          }
        };
        // Narrow tiles take the ring two stages at a time: a BLOCK_N = 128 stage holds 256 clk of tensor work, a halo
        // stage of a 64-wide tile 384, but this thread's serial path per stage was ~80 clk of barrier probe + the UMMA
        // issue + ~100 clk of elect / branch / loop (320 and 700 clk per stage: profiles/r02_conv_trace.txt).  One wait on
        // the barriers of a group and one elected region per group cut the fixed part: two stages per iteration for tiles
        // of at most 128 columns.  256-wide tiles (512 clk per stage, 4-5 ring stages) keep the one-stage loop.
        // (four stages per iteration measured 5 % SLOWER at BLOCK_N = 128: half of the 8-stage ring is then released at
        // once and the producers' refill shows up as a 500-clk wait per group)
        constexpr bool kPairStages = BLOCK_N <= 128;   // two ring stages per loop iteration
        const int nkb = HH ? num_kb_hh : num_kb;
        for (int kb = 0; kb < nkb;) {
          const bool two = kPairStages && kb + 1 < nkb;
          int stage_b = stage + 1;
          uint32_t phase_b = phase;
          if (stage_b == Cfg::kStages) { stage_b = 0; phase_b ^= 1; }
          S2P_TRACE(3, tr_m);
          if (two) mbar_wait2(&full_bar[stage], phase, &full_bar[stage_b], phase_b);
          else mbar_wait(&full_bar[stage], phase);
          S2P_TRACE(4, tr_m);
          tc_fence_after();
          if (elect_one()) {
            issue_stage(stage, kb, nkb);
            if (two) issue_stage(stage_b, kb + 1, nkb);
          }
          __syncwarp();
          S2P_TRACE(5, tr_m);
#ifdef S2P_CONV_TRACE
          ++tr_m;
#endif
          if (two) { stage = stage_b; phase = phase_b; kb += 2; } else { kb += 1; }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp < 10) {
    // ------------------------------------------------------------ epilogue (8 warps)
    // warp -> TMEM lane quarter q (hardware rule: warp%4) and one half of the tile's columns.
    // TMEM loads are software pipelined (the load of step i+1 is in flight while step i is
    // converted and stored), operands that do not depend on the accumulators are fetched before
    // the accumulator wait, and the accumulator buffer is released right after its last load.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int et = threadIdx.x - 64;  // 0..255 inside the epilogue group
    const int m = q * 32 + lane;      // accumulator row = pixel inside the tile
    const int my = m / p.mpitch, mx = m - my * p.mpitch;   // (HH: mpitch = tile_w + 2, columns >= tile_w are halo rows)
    const float* bias_s = p.bias;
    bool bias_staged = false;
    if constexpr (EPI == kEpiPlain || EPI == kEpiSimple) {
      // bias vector staged in shared memory once per CTA (broadcast LDS instead of global loads)
      if (p.bias != nullptr && p.Cout <= 1024) {
        float* bs = reinterpret_cast<float*>(tab_smem);
        for (int i = et; i < p.Cout; i += 256) bs[i] = __ldg(p.bias + i);
        bias_s = bs;
        bias_staged = true;
        asm volatile("bar.sync 1, 256;" ::: "memory");
      }
    }
    int it = 0;
    for (int tile = tile_first; tile < num_tiles; tile += tile_step, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      if (et == 0) S2P_TRACE(8, it);
      const int m_tile = (tile / p.num_n_tiles) * CL + cta_rank, n_tile = tile % p.num_n_tiles;
      const bool live = m_tile < p.num_m_tiles;     // false only for a pair's dummy tile
      const int img = m_tile / tiles_per_img, t = m_tile % tiles_per_img;
      const int h = (t / p.tiles_x) * p.tile_h + my, w = (t % p.tiles_x) * p.tile_w + mx;
      const bool valid = live && (mx < p.tile_w) && (my < p.tile_h) && (h < p.H) && (w < p.W);
      const size_t pix = (static_cast<size_t>(img) * p.H + h) * p.W + w;
      const int n0 = n_tile * BLOCK_N;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * Cfg::kAccStride;
      const uint32_t stg_s = smem_u32(store_smem + (warp - 2) * 2048);   // this warp's staging tile

      auto release_acc = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CL > 1) mbar_arrive_cluster(map_to_cta(&tmem_empty[acc], 0));
          else mbar_arrive(&tmem_empty[acc]);
        }
      };

      if constexpr (EPI == kEpiMat) {
        constexpr int CT = BLOCK_N / 2;        // channels per n-tile: columns [0,CT) = gamma, [CT,2CT) = beta
        constexpr int kSteps = CT / 32;        // 16-channel steps per warp (each warp owns CT/2 channels)
        const int cw = half * (CT / 2);        // first channel (inside the tile) of this warp
        const int ch0 = n_tile * CT;
        const StoreRows<2> rows2 = store_rows_prepare<2>(static_cast<uint32_t>(pix), valid, lane);
        float4* tab_s = tab_smem + acc * 128;
        if (et < CT && live) tab_s[et] = __ldg(p.mat_tab + static_cast<size_t>(img) * p.mat_C + ch0 + et);
        // x: the tile's pixels arrive as 64-channel boxes (128-byte rows, 128B-swizzled) TMA-loaded by this group
        // itself: once all eight warps are past the previous tile (the bar.sync below) its buffer is free, and the
        // load of THIS tile's boxes flies while the tile's MMAs finish.  Per-thread global loads of x (one 128-byte
        // line per pixel row, 32 lines per warp instruction) kept the SM's single L1 wavefront queue busy for ~90 %
        // of a tile's MMA time (profiles/r02_history.md).
        const bool wide_x = p.x_wide != 0;
        const int tw0 = (t % p.tiles_x) * p.tile_w, th0 = (t / p.tiles_x) * p.tile_h;
        const int rx = p.mat_up == 2 ? (((th0 + my) >> 1) - (th0 >> 1)) * p.x_bw + (((tw0 + mx) >> 1) - (tw0 >> 1)) : m;
        const int jbox = cw / 64;                          // 64-channel box of this warp's channels
        const int bidx = wide_x ? 2 * jbox : jbox;         // its (high-half) box; the low halves are the next box
        const uint32_t xs_hi = smem_u32(x_smem) + bidx * Cfg::kABytes + static_cast<uint32_t>(rx) * 128u;
        const uint32_t xs_lo = xs_hi + Cfg::kABytes;
        const uint32_t xkey = static_cast<uint32_t>(rx) & 7u;
        const uint32_t xc0 = static_cast<uint32_t>(cw % 64) / 8u;   // first 16-byte chunk of this warp's channels in the box
        asm volatile("bar.sync 1, 256;" ::: "memory");   // table rows visible to the 8 epilogue warps; previous tile fully read
        if (et == 0) {
          const int nbox = (BLOCK_N / 128) * (wide_x ? 2 : 1);
          const int xw0 = p.mat_up == 2 ? (tw0 >> 1) : tw0, xh0 = p.mat_up == 2 ? (th0 >> 1) : th0;
          mbar_arrive_expect_tx(x_full, static_cast<uint32_t>(nbox) * p.x_box_bytes);
          for (int j = 0; j < nbox; ++j) {
            const int jb = wide_x ? (j >> 1) : j, lo = wide_x ? (j & 1) : 0;
            tma_load_4d(x_smem + j * Cfg::kABytes, &tmap_x, x_full, lo * p.mat_C + ch0 + jb * 64, xw0, xh0, img);
          }
        }
        mbar_wait_relaxed(x_full, static_cast<uint32_t>(it) & 1u);
        mbar_wait_relaxed(&tmem_full[acc], acc_phase);
        tc_fence_after();
        const uint32_t tab_a = smem_u32(tab_s + cw);   // explicit ld.shared: a generic load of the table costs two wavefronts
        uint32_t rg[16], rb[16];
        tmem_ld16(t_row + cw, rg);
        tmem_ld16(t_row + CT + cw, rb);
#pragma unroll
        for (int sidx = 0; sidx < kSteps; ++sidx) {
          tmem_ld_wait();
          float g[16], bt[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) { g[k] = __uint_as_float(rg[k]); bt[k] = __uint_as_float(rb[k]); }
          uint4 xr0 = make_uint4(0u, 0u, 0u, 0u), xr1 = xr0, xq0 = xr0, xq1 = xr0;
          if (valid) {
            const uint32_t c0 = ((xc0 + 2u * sidx) ^ xkey) << 4, c1 = ((xc0 + 2u * sidx + 1u) ^ xkey) << 4;
            xr0 = lds128(xs_hi + c0);
            xr1 = lds128(xs_hi + c1);
            if (wide_x) { xq0 = lds128(xs_lo + c0); xq1 = lds128(xs_lo + c1); }
          }
          if (sidx + 1 < kSteps) {
            tmem_ld16(t_row + cw + (sidx + 1) * 16, rg);
            tmem_ld16(t_row + CT + cw + (sidx + 1) * 16, rb);
          } else {
            release_acc();
          }
          uint32_t ow_out[8] = {0, 0, 0, 0, 0, 0, 0, 0};
          if (valid) {
            const uint32_t xw[8] = {xr0.x, xr0.y, xr0.z, xr0.w, xr1.x, xr1.y, xr1.z, xr1.w};
            const uint32_t xlw[8] = {xq0.x, xq0.y, xq0.z, xq0.w, xq1.x, xq1.y, xq1.z, xq1.w};
            uint32_t ow[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              float o2[2];
#pragma unroll
              for (int hlf = 0; hlf < 2; ++hlf) {
                const int k = 2 * e + hlf;
                const float4 tb = lds128f(tab_a + static_cast<uint32_t>(sidx * 16 + k) * 16u);
                float xv = hlf ? bf16_hi(xw[e]) : bf16_lo(xw[e]);
                if (wide_x) xv += hlf ? bf16_hi(xlw[e]) : bf16_lo(xlw[e]);
                const float xh = fmaf(xv, tb.x, tb.y);
                float v = fmaf(xh, tb.z + g[k], tb.w + bt[k]);
                if (p.act == 2) v = v > 0.f ? v : 0.2f * v;
                else if (p.act == 1) v = fmaxf(v, 0.f);
                o2[hlf] = v;
              }
              ow[e] = pack_bf16x2(o2[0], o2[1]);
            }
            ow_out[0] = ow[0]; ow_out[1] = ow[1]; ow_out[2] = ow[2]; ow_out[3] = ow[3];
            ow_out[4] = ow[4]; ow_out[5] = ow[5]; ow_out[6] = ow[6]; ow_out[7] = ow[7];
          }
          warp_store_rows<2>(stg_s, ow_out, p.out + ch0 + cw + sidx * 16, p.out_pitch, rows2, lane);
        }
      } else if constexpr (EPI == kEpiToRGB) {
        mbar_wait_relaxed(&tmem_full[acc], acc_phase);
        if (et == 0) S2P_TRACE(9, it);
        tc_fence_after();
        static_assert(EPI != kEpiToRGB || (HH && BLOCK_N == 48), "ToRGB runs in the tap-folded halo mode only");
        // The two warp groups (column halves elsewhere) take the tiles in turn: a tile's epilogue is a chain of
        // latencies (wake-up, tcgen05.ld, exchange, tanh, stores: ~1300 clk) around very little work, and with the two
        // accumulator buffers two of them can be in flight.  The idle group still waits for the tile's accumulator
        // barrier before it releases, so no warp runs more than one tile ahead of the MMA thread.
        if (half == (it & 1)) {
          // accumulator columns [16 s, 16 s + 3) of row m hold tap s of the filter rows applied to the box row m:
          // out(m) = D(m)[0..2] + D(m + 1)[16..18] + D(m + 2)[32..34].  Rows m + 1, m + 2 are the next two lanes: the
          // host picks a row pitch (tile_w + 2) that divides 32, so the last two lanes of every warp are halo rows and
          // no valid row needs another warp's accumulators.  Fixed summation order.
          uint32_t g0[16], g1[16], g2[16];
          tmem_ld16(t_row, g0);
          tmem_ld16(t_row + 16, g1);
          tmem_ld16(t_row + 32, g2);
          tmem_ld_wait();
          release_acc();
          float o[3];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const float t1 = __shfl_down_sync(0xffffffffu, __uint_as_float(g1[k]), 1);
            const float t2 = __shfl_down_sync(0xffffffffu, __uint_as_float(g2[k]), 2);
            o[k] = tanhf(((__uint_as_float(g0[k]) + t1) + t2) + (p.bias ? __ldg(p.bias + k) : 0.f));
          }
          if (valid) {
            const int HW = p.H * p.W;
            const int rem = h * p.W + w;
            if (p.rgb_f32 != nullptr) {
              float* of = p.rgb_f32 + static_cast<size_t>(img) * 3 * HW + rem;
              of[0] = o[0]; of[HW] = o[1]; of[2 * HW] = o[2];
            }
            if (p.rgb_u8 != nullptr) {
              uint8_t* ou = p.rgb_u8 + pix * 3;
#pragma unroll
              for (int k = 0; k < 3; ++k)
                ou[k] = static_cast<uint8_t>(fminf(fmaxf(floorf((o[k] + 1.f) * 127.5f + 0.5f), 0.f), 255.f));
            }
          }
        } else {
          release_acc();
        }
      } else if constexpr (EPI == kEpiSimple) {
        // bias + activation + store only (the merged mlp_shared convs, the discriminator / VGG convs, every
        // data-gradient conv): the accumulators are converted in place and the next chunk is requested once they are
        // packed -- no register copies (the general epilogue below reloads `raw` under a live copy: 32 moves per
        // chunk), ReLU on the packed pairs (rounding to bf16 is monotonic and keeps 0: the same bits, 16 instructions
        // instead of 32), 32-bit store offsets.  310 -> ~200 instructions per chunk; hs1 930 -> 858 us.
        constexpr int kChunks = BLOCK_N / 64;          // 32-column chunks per warp
        const int cbeg = half * kChunks;
        const StoreOffs<4> offs4 = store_offs_prepare<4>(static_cast<uint32_t>(h * p.W + w), valid, lane, p.out_pitch);
        __nv_bfloat16* out_img = p.out + static_cast<size_t>(img) * p.H * p.W * p.out_pitch + n0;   // tile-uniform
        mbar_wait_relaxed(&tmem_full[acc], acc_phase);
        if (et == 0) S2P_TRACE(9, it);
        tc_fence_after();
        uint32_t raw[32];
        tmem_ld32(t_row + cbeg * 32, raw);
#pragma unroll 1
        for (int i = 0; i < kChunks; ++i) {
          const int c = cbeg + i;
          float4 bb[8];
          if (bias_staged) {
            const uint32_t b_a = smem_u32(bias_s + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) bb[j] = lds128f(b_a + 16u * j);
          } else if (bias_s != nullptr) {
            const float4* b4 = reinterpret_cast<const float4*>(bias_s + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) bb[j] = __ldg(b4 + j);
          }
          tmem_ld_wait();
          uint32_t packed[16];
          if (bias_s != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float a0 = __uint_as_float(raw[4 * j + 0]) + bb[j].x, a1 = __uint_as_float(raw[4 * j + 1]) + bb[j].y;
              float a2 = __uint_as_float(raw[4 * j + 2]) + bb[j].z, a3 = __uint_as_float(raw[4 * j + 3]) + bb[j].w;
              if (p.act == 2) {
                a0 = a0 > 0.f ? a0 : 0.2f * a0; a1 = a1 > 0.f ? a1 : 0.2f * a1;
                a2 = a2 > 0.f ? a2 : 0.2f * a2; a3 = a3 > 0.f ? a3 : 0.2f * a3;
              }
              packed[2 * j] = pack_bf16x2(a0, a1);
              packed[2 * j + 1] = pack_bf16x2(a2, a3);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float a0 = __uint_as_float(raw[2 * j]), a1 = __uint_as_float(raw[2 * j + 1]);
              if (p.act == 2) { a0 = a0 > 0.f ? a0 : 0.2f * a0; a1 = a1 > 0.f ? a1 : 0.2f * a1; }
              packed[j] = pack_bf16x2(a0, a1);
            }
          }
          if (p.act == 1) {
#pragma unroll
            for (int j = 0; j < 16; ++j) packed[j] = relu_bf16x2(packed[j]);
          }
          if (i + 1 < kChunks) tmem_ld32(t_row + (c + 1) * 32, raw);
          else release_acc();
          warp_store_offs<4>(stg_s, packed, out_img + c * 32, offs4, lane);
        }
      } else {
        constexpr int kChunks = BLOCK_N / 64;          // 32-column chunks per warp
        const int cbeg = half * kChunks;
        const bool has_res = (p.residual != nullptr) && valid;
        const StoreRows<4> rows4 = store_rows_prepare<4>(static_cast<uint32_t>(pix), valid, lane);
        const bool wide_res = has_res && (p.residual_lo != nullptr);
        const bool wide_out = p.out_lo != nullptr;
        const uint4* r4 = reinterpret_cast<const uint4*>(p.residual + pix * p.res_pitch + n0 + cbeg * 32);
        uint4 rn[4];
        if (has_res) {
#pragma unroll
          for (int j = 0; j < 4; ++j) rn[j] = __ldg(r4 + j);
        }
        mbar_wait_relaxed(&tmem_full[acc], acc_phase);
        if (et == 0) S2P_TRACE(9, it);
        tc_fence_after();
        uint32_t raw[32];
        tmem_ld32(t_row + cbeg * 32, raw);
#pragma unroll 1
        for (int i = 0; i < kChunks; ++i) {
          const int c = cbeg + i;
          // the chunk's bias values leave shared memory before the accumulator wait, not after it
          float4 bb[8];
          if (bias_staged) {       // explicit ld.shared (a generic load of a shared address costs two wavefronts)
            const uint32_t b_a = smem_u32(bias_s + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) bb[j] = lds128f(b_a + 16u * j);
          } else if (bias_s != nullptr) {
            const float4* b4 = reinterpret_cast<const float4*>(bias_s + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) bb[j] = __ldg(b4 + j);
          }
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
          uint4 rc[4];
          if (has_res) {
#pragma unroll
            for (int j = 0; j < 4; ++j) rc[j] = rn[j];
          }
          if (i + 1 < kChunks) {
            tmem_ld32(t_row + (c + 1) * 32, raw);
            if (has_res) {
#pragma unroll
              for (int j = 0; j < 4; ++j) rn[j] = __ldg(r4 + (i + 1) * 4 + j);
            }
          } else {
            release_acc();
          }
          const int ch = n0 + c * 32;
          if (bias_s != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[4 * j + 0] += bb[j].x; v[4 * j + 1] += bb[j].y; v[4 * j + 2] += bb[j].z; v[4 * j + 3] += bb[j].w;
            }
          }
          if (has_res) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 rr = rc[j];
              v[8 * j + 0] += bf16_lo(rr.x); v[8 * j + 1] += bf16_hi(rr.x);
              v[8 * j + 2] += bf16_lo(rr.y); v[8 * j + 3] += bf16_hi(rr.y);
              v[8 * j + 4] += bf16_lo(rr.z); v[8 * j + 5] += bf16_hi(rr.z);
              v[8 * j + 6] += bf16_lo(rr.w); v[8 * j + 7] += bf16_hi(rr.w);
            }
          }
          if (wide_res) {
            // low halves of a wide residual: loaded at the point of use (not pipelined like the high halves) -- the
            // only launches that take this branch are the identity-shortcut blocks at 25x25, K = 4608 per tile,
            // whose epilogue has slack, and the registers of a second prefetch set spill in the wide tiles
            // (address = the high halves' + a launch-uniform distance: no second per-thread pointer)
            const uint4* rl4 = r4 + (reinterpret_cast<const uint4*>(p.residual_lo) - reinterpret_cast<const uint4*>(p.residual));
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 rr = __ldg(rl4 + i * 4 + j);
              v[8 * j + 0] += bf16_lo(rr.x); v[8 * j + 1] += bf16_hi(rr.x);
              v[8 * j + 2] += bf16_lo(rr.y); v[8 * j + 3] += bf16_hi(rr.y);
              v[8 * j + 4] += bf16_lo(rr.z); v[8 * j + 5] += bf16_hi(rr.z);
              v[8 * j + 6] += bf16_lo(rr.w); v[8 * j + 7] += bf16_hi(rr.w);
            }
          }
          if (p.act == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          } else if (p.act == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : 0.2f * v[j];
          }
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) packed[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
          warp_store_rows<4>(stg_s, packed, p.out + ch, p.out_pitch, rows4, lane);
          if (wide_out) {
            // low halves of the wide store: v - hi is exact in fp32, its bf16 rounding leaves ~2^-17 relative
            uint32_t plo[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)
              plo[j] = pack_bf16x2(v[2 * j] - bf16_lo(packed[j]), v[2 * j + 1] - bf16_hi(packed[j]));
            warp_store_rows<4>(stg_s, plo, p.out_lo + ch, p.out_pitch, rows4, lane);
          }
          if (p.stats_partial != nullptr && live) {
            // statistics of the values as stored (bf16-rounded; the fp32 values when stored wide), fixed reduction order
            // (sum first, then sum of squares, each through one 32-register butterfly: both arrays live at once
            // next to the fp32 values of a wide store spill)
            if (!wide_out) {
#pragma unroll
              for (int j = 0; j < 16; ++j) { v[2 * j] = bf16_lo(packed[j]); v[2 * j + 1] = bf16_hi(packed[j]); }
            }
            float s1[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) s1[j] = valid ? v[j] : 0.f;
            const float cs = warp_colsum32(s1, lane);
#pragma unroll
            for (int j = 0; j < 32; ++j) s1[j] = valid ? v[j] * v[j] : 0.f;
            const float cq = warp_colsum32(s1, lane);
            stat_smem[((acc * 4 + q) * BLOCK_N) + c * 32 + lane] = make_float2(cs, cq);
          }
        }
        if (p.stats_partial != nullptr && live) {
          // one partial row per TILE: the quarter-0 warp of each column half adds the four quarters in a fixed order
          // (4x fewer partial rows for the finalize kernels to read: they were 50 MB per 100x100 layer and micro-batch)
          if (half == 0) asm volatile("bar.sync 2, 128;" ::: "memory");
          else asm volatile("bar.sync 3, 128;" ::: "memory");
          if (q == 0) {
            const size_t prow = static_cast<size_t>(img) * tiles_per_img + t;
            float2* dst = reinterpret_cast<float2*>(p.stats_partial) + prow * p.Cout + n0;
#pragma unroll 1
            for (int i = 0; i < kChunks; ++i) {
              const int col = (cbeg + i) * 32 + lane;
              const float2 a0 = stat_smem[(acc * 4 + 0) * BLOCK_N + col], a1 = stat_smem[(acc * 4 + 1) * BLOCK_N + col];
              const float2 a2 = stat_smem[(acc * 4 + 2) * BLOCK_N + col], a3 = stat_smem[(acc * 4 + 3) * BLOCK_N + col];
              dst[col] = make_float2(((a0.x + a1.x) + a2.x) + a3.x, ((a0.y + a1.y) + a2.y) + a3.y);
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();   // neither CTA leaves (or frees TMEM) while the pair is still working
  if (warp == 1) {
    if constexpr (CL > 1) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    else tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

}  // namespace s2p

#ifdef S2P_CONV_TRACE
extern "C" int s2p_conv_trace_read(long long* host) {
  return static_cast<int>(cudaMemcpyFromSymbol(host, s2p::g_conv_trace, sizeof(long long) * 12 * s2p::kTraceN));
}
#endif

// =============================================================== host side
namespace s2p {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
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

// spatial tile (tile_w x tile_h <= 128 pixels) that wastes the fewest MMA rows
void conv_pick_tile(int H, int W, int max_box_w, int* tw_out, int* th_out) {
  double best = -1.0;
  int btw = 1, bth = 1;
  for (int tw = 1; tw <= W && tw <= 128 && tw <= max_box_w; ++tw) {
    int th = 128 / tw;
    if (th > H) th = H;
    if (th < 1) continue;
    const long tiles = static_cast<long>((W + tw - 1) / tw) * ((H + th - 1) / th);
    const double eff = static_cast<double>(H) * W / (tiles * 128.0);
    if (eff > best + 1e-9 || (eff > best - 1e-9 && tw > btw)) { best = eff; btw = tw; bth = th; }
  }
  *tw_out = btw;
  *th_out = bth;
}

// horizontal-halo geometry: tile_w x tile_h output pixels laid out as tile_h rows of tile_w + 2 M-rows (<= 128 in all)
void conv_pick_tile_hh(int H, int W, int* tw_out, int* th_out) {
  double best = -1.0;
  int btw = 1, bth = 1;
  for (int tw = 1; tw <= W && tw + 2 <= 128; ++tw) {
    int th = 128 / (tw + 2);
    if (th > H) th = H;
    if (th < 1) continue;
    const long tiles = static_cast<long>((W + tw - 1) / tw) * ((H + th - 1) / th);
    const double eff = static_cast<double>(H) * W / (tiles * 128.0);
    if (eff > best + 1e-9 || (eff > best - 1e-9 && tw > btw)) { best = eff; btw = tw; bth = th; }
  }
  *tw_out = btw;
  *th_out = bth;
}

// ToRGB (tap-folded halo mode): the row pitch tile_w + 2 must divide 32, so that rows 30, 31 of every 32-row lane
// quarter are halo rows (the epilogue takes rows m + 1, m + 2 from the next lanes of the same warp).
void conv_pick_tile_torgb(int H, int W, int* tw_out, int* th_out) {
  double best = -1.0;
  int btw = 30, bth = 4;
  const int cand[3] = {30, 14, 6};
  for (int i = 0; i < 3; ++i) {
    const int tw = cand[i];
    int th = 128 / (tw + 2);
    if (th > H) th = H;
    const long tiles = static_cast<long>((W + tw - 1) / tw) * ((H + th - 1) / th);
    const double eff = static_cast<double>(H) * W / (tiles * 128.0);
    if (eff > best + 1e-9) { best = eff; btw = tw; bth = th; }
  }
  *tw_out = btw;
  *th_out = bth;
}

// which convolutions run in the horizontal-halo mode: 3x3 stride-1 'same' convs whose accumulator tile is at most 64
// columns wide (operand-ingest bound otherwise: 35 % tensor pipe at Cout = 64, 8 % for ToRGB)
static bool conv_uses_hh(int Cout_tile, int ksize, int stride, int pad, int epi) {
  return ksize == 3 && stride == 1 && pad == 1 && Cout_tile <= 64 && epi != kEpiMat;
}

template <int BLOCK_N, int EPI, int CL, bool HH = false>
static int launch_conv_cl(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& ta2, const CUtensorMap& tx,
                          const ConvParams& p, cudaStream_t st) {
  using Cfg = ConvCfg<BLOCK_N, CL, EPI, HH>;
  static bool configured[kMaxDevices] = {};   // the attribute is per device
  const int dev = current_device();
  if (!configured[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<BLOCK_N, EPI, CL, HH>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured[dev] = true;
  }
  const int super_tiles = ((p.num_m_tiles + CL - 1) / CL) * p.num_n_tiles;
  const int max_clusters = num_sms() / CL;
  const int grid = (super_tiles < max_clusters ? super_tiles : max_clusters) * CL;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(kConvThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  S2P_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_igemm_kernel<BLOCK_N, EPI, CL, HH>, ta, tb, ta2, tx, p));
  count_launch();
  return check_launch("conv_igemm_kernel");
}

// cluster size along M (CTAs sharing one multicast weight tile): 2 unless S2P_CONV_CLUSTER=1
int conv_cluster_size() {
  static int cl = 0;
  if (cl == 0) {
    const char* e = getenv("S2P_CONV_CLUSTER");
    cl = (e != nullptr && e[0] == '1') ? 1 : 2;
  }
  return cl;
}

template <int BLOCK_N, int EPI, bool HH = false>
static int launch_conv(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& ta2, const CUtensorMap& tx,
                       const ConvParams& p, cudaStream_t st) {
  if (conv_cluster_size() == 2) return launch_conv_cl<BLOCK_N, EPI, 2, HH>(ta, tb, ta2, tx, p, st);
  return launch_conv_cl<BLOCK_N, EPI, 1, HH>(ta, tb, ta2, tx, p, st);
}

}  // namespace s2p

extern "C" int s2p_conv_stats_rows(int Hout, int Wout, int Cout, int ksize, int stride) {
  int tw, th;
  const int block_n = (Cout % 256 == 0) ? 256 : (Cout % 192 == 0) ? 192 : (Cout % 128 == 0) ? 128 : 64;
  if (s2p::conv_uses_hh(block_n, ksize, stride, ksize / 2, s2p::kEpiPlain)) s2p::conv_pick_tile_hh(Hout, Wout, &tw, &th);
  else s2p::conv_pick_tile(Hout, Wout, 128, &tw, &th);
  return ((Wout + tw - 1) / tw) * ((Hout + th - 1) / th);   // one partial row per tile
}

namespace s2p {

// Shared host path of the two conv entry points: validates the GEMM geometry, builds the
// two tensor maps and launches.  `p` arrives with the epilogue fields already filled.
static int conv_dispatch(const void* x, int x_pitch, const void* w_packed, int B, int Hin, int Win, int Cin,
                         int Cout, int ksize, int stride, int pad, int block_n, int epi, ConvParams p, void* stream,
                         const void* x2 = nullptr, int x2_pitch = 0, int Cin2 = 0, const void* mat_x = nullptr,
                         int mat_x_pitch = 0) {
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return set_error(S2P_ERR_CUDA, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  p.B = B;
  p.H = (Hin + 2 * pad - ksize) / stride + 1;
  p.W = (Win + 2 * pad - ksize) / stride + 1;
  p.Cin = Cin; p.Cout = Cout; p.R = ksize; p.S = ksize; p.pad = pad; p.stride = stride;
  if (static_cast<long long>(B) * p.H * p.W >= (1LL << 32))
    return set_error(S2P_ERR_BAD_ARG, "conv: B*H*W must be below 2^32 pixels per call");
  // the bias / activation-only epilogue is a kernel of its own (32-bit store offsets inside an image)
  if (epi == kEpiPlain && p.residual == nullptr && p.out_lo == nullptr && p.stats_partial == nullptr &&
      static_cast<long long>(p.H) * p.W * p.out_pitch < (1LL << 32))
    epi = kEpiSimple;
  const bool hh = conv_uses_hh(block_n, ksize, stride, pad, epi);
  if (epi == kEpiToRGB) conv_pick_tile_torgb(p.H, p.W, &p.tile_w, &p.tile_h);
  else if (hh) conv_pick_tile_hh(p.H, p.W, &p.tile_w, &p.tile_h);
  else conv_pick_tile(p.H, p.W, 128, &p.tile_w, &p.tile_h);
  p.mpitch = hh ? p.tile_w + 2 : p.tile_w;
  p.tiles_x = (p.W + p.tile_w - 1) / p.tile_w;
  p.tiles_y = (p.H + p.tile_h - 1) / p.tile_h;
  p.num_m_tiles = B * p.tiles_x * p.tiles_y;
  p.num_n_tiles = Cout / block_n;

  p.kb2 = x2 != nullptr ? Cin2 / 64 : 0;
  CUtensorMap ta, tb, ta2;
  {
    // (with cin_wrap the logical Cin walks [hi | lo | hi] over a tensor that holds [hi | lo])
    cuuint64_t dims[4] = {(cuuint64_t)(p.cin_wrap > 0 ? p.cin_wrap : Cin), (cuuint64_t)Win, (cuuint64_t)Hin, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)x_pitch * 2, (cuuint64_t)x_pitch * 2 * Win,
                             (cuuint64_t)x_pitch * 2 * Win * Hin};
    cuuint32_t box[4] = {64, (cuuint32_t)(p.tile_w * stride), (cuuint32_t)(p.tile_h * stride), 1};
    if (hh) box[1] = (cuuint32_t)p.mpitch;      // one box per filter row: the tile's columns plus a halo column each side
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = encode(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box,
                        estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "conv: activation tensor map encode failed (CUresult %d)", (int)r);
  }
  {
    // (ToRGB: [48 = 3 taps x 16][3 filter rows x Cin], see s2p_torgb_igemm_fwd)
    const int K = epi == kEpiToRGB ? ksize * Cin : ksize * ksize * Cin + (x2 != nullptr ? Cin2 : 0);
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Cout};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(block_n / conv_cluster_size())};   // each CTA loads its share
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(&tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_packed), dims, strides,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "conv: weight tensor map encode failed (CUresult %d)", (int)r);
  }
  ta2 = ta;
  if (x2 != nullptr) {   // same spatial size as the output (stride 1 only), 1x1 taps
    cuuint64_t dims[4] = {(cuuint64_t)Cin2, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)x2_pitch * 2, (cuuint64_t)x2_pitch * 2 * p.W,
                             (cuuint64_t)x2_pitch * 2 * p.W * p.H};
    cuuint32_t box[4] = {64, (cuuint32_t)p.mpitch, (cuuint32_t)p.tile_h, 1};   // (halo mode: same row pitch as the main operand)
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&ta2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x2), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "conv: second operand tensor map encode failed (CUresult %d)", (int)r);
  }
  CUtensorMap tx = ta;
  if (epi == kEpiMat) {
    // the pre-norm activation of the fused MAT epilogue: [B][H/up][W/up][C or 2C] NHWC bf16, 64-channel boxes covering
    // one output tile's pixels (at half resolution: tile/2 + 1 per axis reaches every parent pixel of the tile)
    const int up = p.mat_up;
    const int Hx = p.H / up, Wx = p.W / up;
    const int bw = up == 2 ? p.tile_w / 2 + 1 : p.tile_w, bh = up == 2 ? p.tile_h / 2 + 1 : p.tile_h;
    if (bw * bh > 128 || bw > 256 || bh > 256) return set_error(S2P_ERR_UNSUPPORTED, "conv_mat_fused: x box %dx%d", bw, bh);
    p.x_bw = bw;
    p.x_box_bytes = static_cast<uint32_t>(bw) * bh * 128u;
    cuuint64_t dims[4] = {(cuuint64_t)(p.x_wide ? 2 * p.mat_C : p.mat_C), (cuuint64_t)Wx, (cuuint64_t)Hx, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)mat_x_pitch * 2, (cuuint64_t)mat_x_pitch * 2 * Wx,
                             (cuuint64_t)mat_x_pitch * 2 * Wx * Hx};
    cuuint32_t box[4] = {64, (cuuint32_t)bw, (cuuint32_t)bh, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(mat_x), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(S2P_ERR_CUDA, "conv_mat_fused: x tensor map encode failed (CUresult %d)", (int)r);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (epi == kEpiToRGB) {
    if (!hh || block_n != 48) return set_error(S2P_ERR_UNSUPPORTED, "torgb: the tap-folded kernel needs a 3x3 stride-1 'same' convolution");
    return launch_conv<48, kEpiToRGB, true>(ta, tb, ta2, tx, p, st);
  }
  if (epi == kEpiMat) {
    switch (block_n) {
      case 256: return launch_conv<256, kEpiMat>(ta, tb, ta2, tx, p, st);
      default: return launch_conv<128, kEpiMat>(ta, tb, ta2, tx, p, st);
    }
  }
  if (epi == kEpiSimple) {
    switch (block_n) {
      case 256: return launch_conv<256, kEpiSimple>(ta, tb, ta2, tx, p, st);
      case 192: return launch_conv<192, kEpiSimple>(ta, tb, ta2, tx, p, st);
      case 128: return launch_conv<128, kEpiSimple>(ta, tb, ta2, tx, p, st);
      default: return hh ? launch_conv<64, kEpiSimple, true>(ta, tb, ta2, tx, p, st) : launch_conv<64, kEpiSimple>(ta, tb, ta2, tx, p, st);
    }
  }
  switch (block_n) {
    case 256: return launch_conv<256, kEpiPlain>(ta, tb, ta2, tx, p, st);
    case 192: return launch_conv<192, kEpiPlain>(ta, tb, ta2, tx, p, st);
    case 128: return launch_conv<128, kEpiPlain>(ta, tb, ta2, tx, p, st);
    default: return hh ? launch_conv<64, kEpiPlain, true>(ta, tb, ta2, tx, p, st) : launch_conv<64, kEpiPlain>(ta, tb, ta2, tx, p, st);
  }
}

static inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

}  // namespace s2p

extern "C" int s2p_conv_igemm_fwd(const void* x, int x_pitch, const void* w_packed, const float* bias,
                                  const void* residual, int res_pitch, void* y, int y_pitch,
                                  float* stats_partial, int B, int Hin, int Win, int Cin, int Cout,
                                  int ksize, int stride, int pad, int act, void* stream) {
  return s2p_conv_igemm_fwd2(x, x_pitch, nullptr, 0, 0, w_packed, bias, residual, nullptr, res_pitch, y, nullptr, y_pitch,
                             stats_partial, B, Hin, Win, Cin, Cout, ksize, stride, pad, act, 0, 0, stream);
}

extern "C" int s2p_conv_igemm_fwd2(const void* x, int x_pitch, const void* x2, int x2_pitch, int Cin2,
                                   const void* w_packed, const float* bias, const void* residual,
                                   const void* residual_lo, int res_pitch, void* y, void* y_lo, int y_pitch,
                                   float* stats_partial, int B, int Hin, int Win, int Cin, int Cout, int ksize,
                                   int stride, int pad, int act, int cin_wrap, int operands_f16, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && w_packed && y, "conv: null pointer");
  S2P_CHECK_ARG(cin_wrap == 0 || (cin_wrap > 0 && cin_wrap % 64 == 0 && cin_wrap < Cin && Cin <= 2 * cin_wrap),
                "conv: cin_wrap=%d (Cin=%d)", cin_wrap, Cin);
  S2P_CHECK_ARG(!residual_lo || residual, "conv: residual_lo without residual");
  S2P_CHECK_ARG(aligned16(residual_lo) && aligned16(y_lo), "conv: pointers must be 16-byte aligned");
  if (x2 != nullptr) {
    S2P_CHECK_ARG(stride == 1 && 2 * pad == ksize - 1, "conv: a second operand needs a stride-1 'same' convolution");
    S2P_CHECK_ARG(Cin2 > 0 && Cin2 % 64 == 0, "conv: Cin2=%d must be a multiple of 64", Cin2);
    S2P_CHECK_ARG(x2_pitch >= Cin2 && x2_pitch % 8 == 0, "conv: x2_pitch=%d", x2_pitch);
    S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(x2) & 15) == 0, "conv: pointers must be 16-byte aligned");
  }
  S2P_CHECK_ARG(B > 0 && Hin > 0 && Win > 0, "conv: bad dims B=%d H=%d W=%d", B, Hin, Win);
  S2P_CHECK_ARG(Cin > 0 && Cin % 64 == 0, "conv: Cin=%d must be a multiple of 64", Cin);
  S2P_CHECK_ARG(Cout > 0 && Cout % 64 == 0, "conv: Cout=%d must be a multiple of 64", Cout);
  S2P_CHECK_ARG(ksize >= 1 && ksize <= 5, "conv: ksize=%d unsupported", ksize);
  S2P_CHECK_ARG(pad >= 0 && pad < ksize && Hin + 2 * pad >= ksize && Win + 2 * pad >= ksize, "conv: pad=%d", pad);
  S2P_CHECK_ARG(stride == 1 || stride == 2, "conv: stride=%d unsupported", stride);
  S2P_CHECK_ARG(act >= 0 && act <= 2, "conv: act=%d unsupported", act);
  S2P_CHECK_ARG(x_pitch >= (cin_wrap > 0 ? cin_wrap : Cin) && x_pitch % 8 == 0, "conv: x_pitch=%d", x_pitch);
  S2P_CHECK_ARG(y_pitch >= Cout && y_pitch % 8 == 0, "conv: y_pitch=%d", y_pitch);
  S2P_CHECK_ARG(!residual || (res_pitch >= Cout && res_pitch % 8 == 0), "conv: res_pitch=%d", res_pitch);
  S2P_CHECK_ARG(aligned16(x) && aligned16(y) && aligned16(w_packed) && aligned16(residual) && aligned16(bias),
                "conv: pointers must be 16-byte aligned");
  ConvParams p{};
  p.out_pitch = y_pitch; p.res_pitch = res_pitch; p.act = act;
  p.bias = bias;
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.residual_lo = static_cast<const __nv_bfloat16*>(residual_lo);
  p.out = static_cast<__nv_bfloat16*>(y);
  p.out_lo = static_cast<__nv_bfloat16*>(y_lo);
  p.stats_partial = stats_partial;
  p.cin_wrap = cin_wrap;
  p.op_f16 = operands_f16 != 0;
  const int block_n = (Cout % 256 == 0) ? 256 : (Cout % 192 == 0) ? 192 : (Cout % 128 == 0) ? 128 : 64;
  return conv_dispatch(x, x_pitch, w_packed, B, Hin, Win, Cin, Cout, ksize, stride, pad, block_n, kEpiPlain, p, stream,
                       x2, x2_pitch, Cin2);
}

extern "C" int s2p_mat_fused_ct(int C) { return C >= 128 ? 128 : C; }

extern "C" int s2p_conv_mat_fused_fwd(const void* h, int h_pitch, const void* w_packed_gb, const void* x,
                                      const void* x_lo, int x_pitch, int up, const float* mod_table, void* y,
                                      int y_pitch, int B, int H, int W, int Cin, int C, int ksize, int act,
                                      void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(h && w_packed_gb && x && mod_table && y, "conv_mat_fused: null pointer");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0, "conv_mat_fused: bad dims B=%d H=%d W=%d", B, H, W);
  S2P_CHECK_ARG(Cin > 0 && Cin % 64 == 0, "conv_mat_fused: Cin=%d must be a multiple of 64", Cin);
  S2P_CHECK_ARG(C == 64 || (C >= 128 && C % 128 == 0), "conv_mat_fused: C=%d must be 64 or a multiple of 128", C);
  S2P_CHECK_ARG(!x_lo || x_lo == static_cast<const __nv_bfloat16*>(x) + C,
                "conv_mat_fused: a wide x must be one [hi | lo] tensor (x_lo == x + C channels)");
  S2P_CHECK_ARG(ksize == 1 || ksize == 3, "conv_mat_fused: ksize=%d unsupported", ksize);
  S2P_CHECK_ARG(act >= 0 && act <= 2, "conv_mat_fused: act=%d unsupported", act);
  S2P_CHECK_ARG(up == 1 || (up == 2 && H % 2 == 0 && W % 2 == 0), "conv_mat_fused: up=%d H=%d W=%d", up, H, W);
  S2P_CHECK_ARG(h_pitch >= Cin && h_pitch % 8 == 0, "conv_mat_fused: h_pitch=%d", h_pitch);
  S2P_CHECK_ARG(x_pitch >= (x_lo ? 2 * C : C) && x_pitch % 8 == 0 && y_pitch >= C && y_pitch % 8 == 0, "conv_mat_fused: pitch");
  S2P_CHECK_ARG(aligned16(h) && aligned16(x) && aligned16(x_lo) && aligned16(y) && aligned16(w_packed_gb) &&
                    aligned16(mod_table), "conv_mat_fused: pointers must be 16-byte aligned");
  ConvParams p{};
  p.out_pitch = y_pitch; p.act = act;
  p.out = static_cast<__nv_bfloat16*>(y);
  p.mat_tab = reinterpret_cast<const float4*>(mod_table);
  p.mat_up = up; p.mat_C = C;
  p.x_wide = x_lo != nullptr;
  const int block_n = 2 * s2p_mat_fused_ct(C);
  return conv_dispatch(h, h_pitch, w_packed_gb, B, H, W, Cin, 2 * C, ksize, 1, ksize / 2, block_n, kEpiMat, p, stream,
                       nullptr, 0, 0, x, x_pitch);
}

extern "C" int s2p_torgb_igemm_fwd(const void* x, int x_pitch, const void* w_packed16, const float* bias,
                                   float* out_f32_nchw, uint8_t* out_u8_nhwc, int B, int H, int W, int Cin,
                                   void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && w_packed16, "torgb_igemm: null pointer");
  S2P_CHECK_ARG(out_f32_nchw || out_u8_nhwc, "torgb_igemm: no output buffer given");
  S2P_CHECK_ARG(B > 0 && H > 0 && W > 0, "torgb_igemm: bad dims");
  S2P_CHECK_ARG(Cin > 0 && Cin % 64 == 0, "torgb_igemm: Cin=%d must be a multiple of 64", Cin);
  S2P_CHECK_ARG(x_pitch >= Cin && x_pitch % 8 == 0, "torgb_igemm: x_pitch=%d", x_pitch);
  S2P_CHECK_ARG(aligned16(x) && aligned16(w_packed16), "torgb_igemm: pointers must be 16-byte aligned");
  ConvParams p{};
  p.bias = bias;
  p.rgb_f32 = out_f32_nchw;
  p.rgb_u8 = out_u8_nhwc;
  return conv_dispatch(x, x_pitch, w_packed16, B, H, W, Cin, 48, 3, 1, 1, 48, kEpiToRGB, p, stream);
}

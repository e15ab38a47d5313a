// Persistent warp-specialised tcgen05 GEMM for the CLIP vision tower:
//   out[M,N] = epilogue( A[M,K] (bf16, row-major) * W[N,K]^T (bf16, row-major) )
// replaces the nn.Linear / nn.Conv2d(patch) / `x @ proj` library GEMMs issued by
// openai/CLIP clip/model.py VisionTransformer.forward (SURVEY.md 2.2 K1,K4,K6-K9).
//
// One CTA per SM, 384 threads:
//   warp 0  : TMA producer   (one elected lane)   global -> 128B-swizzled smem ring
//   warp 1  : MMA issuer     (one elected lane)   tcgen05.mma kind::f16, fp32 accumulators in TMEM
//   warp 2  : TMEM allocator
//   warp 4-11: epilogue      tcgen05.ld -> registers -> fused bias/QuickGELU/residual -> staged, coalesced global
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace evr {

enum GemmEpilogue : int {
  EPI_BF16 = 0,       // out bf16 = acc + bias
  EPI_BF16_GELU = 1,  // out bf16 = quickgelu(acc + bias)
  EPI_F32_RESID = 2,  // out fp32 += acc + bias          (residual stream, in place, red.global.add.v4.f32)
  EPI_F32 = 3,        // out fp32 = acc + bias
  EPI_PATCH = 4,      // out fp32[token row] = acc + pos  (patch-embed: row remap b*P+p -> b*(P+1)+1+p)
};

struct GemmArgs {
  int M, N, K;
  const float* bias;  // [N] or nullptr
  void* out;          // bf16 or fp32, row stride ldo elements
  int ldo;
  const float* pos;   // EPI_PATCH: positional_embedding [(P+1), N]
  int patches;        // EPI_PATCH: P
  // Tail split (set by launch_gemm_bf16).  Tiles are walked round-robin by the persistent units; the tiles of the last,
  // partly filled round (tiles >= full_tiles) are cut into `tail_split` column strips of BLOCK_N / tail_split so that the
  // short round costs 1/tail_split of a tile instead of a whole one (600 tiles on 74 pairs: 8.1 rounds -> 8.25, not 9).
  int full_tiles;     // tiles [0, full_tiles) are whole BLOCK_N-wide tiles
  int tail_split;     // >= 1; strips per tail tile
  // evr_profile_enable(2): [0] = earliest %globaltimer at which a CTA got past pdl_wait, [1] = latest CTA exit; nullptr
  // otherwise.  Times the launch in place (PDL overlap intact), without a stream operation between the kernels.
  unsigned long long* prof;
  int reverse_m;      // 1: row blocks are walked from the last to the first (launch_gemm_bf16 decides)
};

// work item -> (tile, first column, width); items [0, full_tiles) are whole tiles, then tail_split strips per tile
struct GemmItem {
  int m_idx, n0, width;
};
template <int BLOCK_N>
__device__ __forceinline__ GemmItem gemm_item(int it, int num_n, int full_tiles, int tail_split, int num_m = 0,
                                              int reverse_m = 0) {
  GemmItem g;
  int tile = it, sub = 0;
  g.width = BLOCK_N;
  if (it >= full_tiles) {
    const int j = it - full_tiles;
    tile = full_tiles + j / tail_split;
    sub = j - (j / tail_split) * tail_split;
    g.width = BLOCK_N / tail_split;
  }
  const int m_walk = tile / num_n;  // n-fast: the N tiles of one row block run together, A is read from HBM once
  g.m_idx = reverse_m ? num_m - 1 - m_walk : m_walk;
  g.n0 = (tile - m_walk * num_n) * BLOCK_N + sub * g.width;
  return g;
}

constexpr int kGemmBlockM = 128;
constexpr int kGemmBlockK = 64;  // 64 x 16-bit = one 128-B swizzle atom row
constexpr int kGemmThreads = 384;  // 4 control warps + 8 epilogue warps

// Epilogue staging: every epilogue warp owns a [32 rows x 32 columns] tile in shared memory (row pitch padded by 16 B
// so the row-per-lane writes are bank-conflict free).  TMEM hands each lane one accumulator ROW; staging turns that
// into row-contiguous, fully coalesced 16-B global accesses (8 or 4 row segments of 64/128 B per warp instruction).
constexpr int kEpiCols = 32;
template <int EPI>
struct EpiCfg {
  static constexpr bool kOutBf16 = (EPI == 0 || EPI == 1);
  static constexpr int kElt = kOutBf16 ? 2 : 4;
  static constexpr int kRowBytes = kEpiCols * kElt;       // 64 | 128
  static constexpr int kPitch = kRowBytes + 16;           // 80 | 144
  static constexpr int kWarpBytes = 32 * kPitch;
  static constexpr int kBiasWarpBytes = 4 * 32 * 4;       // bias of this warp's <= 4 chunks x 32 columns
  static constexpr int kBytes = 8 * (kWarpBytes + kBiasWarpBytes);  // 24 576 | 40 960
  static constexpr int kUnitsPerRow = kRowBytes / 16;     // 4 | 8
};

// CG = 1: one CTA per 128 x BLOCK_N tile.  CG = 2: a CTA pair (cta_group::2) per 256 x BLOCK_N tile; each CTA stages its
// 128 rows of A and its half of the BLOCK_N weight rows, halving shared-memory and L2 operand traffic per FLOP.
template <int BLOCK_N, int EPI, int CG>
struct GemmCfg {
  static constexpr int kABytes = kGemmBlockM * kGemmBlockK * 2;
  static constexpr int kBRows = BLOCK_N / CG;  // weight rows staged by one CTA
  static constexpr int kBBytes = kBRows * kGemmBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = EpiCfg<EPI>::kBytes;
  static constexpr int kStages = (227 * 1024 - 1024 - 256 - kEpiBytes) / kStageBytes;
  static constexpr int kTmemCols = (2 * BLOCK_N <= 32)    ? 32
                                   : (2 * BLOCK_N <= 64)  ? 64
                                   : (2 * BLOCK_N <= 128) ? 128
                                   : (2 * BLOCK_N <= 256) ? 256
                                                          : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 32 && BLOCK_N <= 256, "invalid UMMA N");
  static_assert(kBBytes % 1024 == 0, "B stage must keep 1024-B alignment for SWIZZLE_128B");
  static_assert(kStages >= 3, "pipeline too shallow");
};

__device__ __forceinline__ float quick_gelu(float x) {
  // clip/model.py QuickGELU: x * sigmoid(1.702 x) = 0.5 x (1 + tanh(0.851 x)): one MUFU.TANH (rel. error 2^-11,
  // below the bf16 output's 2^-9) instead of EX2 + RCP
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.851f * x));
  const float h = 0.5f * x;
  return fmaf(h, t, h);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

template <int BLOCK_N, int EPI, int CG>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_b_tail, const GemmArgs args) {
  using Cfg = GemmCfg<BLOCK_N, EPI, CG>;
  using Epi = EpiCfg<EPI>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_epi = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;                     // [kStages]  TMA -> MMA
  uint64_t* empty_bar = bars + kStages;          // [kStages]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * kStages;      // [2]        MMA -> epilogue
  uint64_t* tmem_empty = bars + 2 * kStages + 2; // [2]        epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int kTileM = kGemmBlockM * CG;
  const int num_m = (args.M + kTileM - 1) / kTileM;
  uint32_t rank = 0;
  if constexpr (CG == 2) rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int unit = blockIdx.x / CG, num_units = gridDim.x / CG;
  if constexpr (CG == 2) cluster_sync_relaxed();  // both CTAs resident before the pair-wide TMEM allocation
  const int num_n = (args.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = num_m * num_n;
  const int num_items = args.full_tiles + (num_tiles - args.full_tiles) * args.tail_split;
  const int num_kb = args.K / kGemmBlockK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_b_tail);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], CG);  // one arrival per CTA of the unit, on the leader's barrier
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8 * CG);  // one arrive per epilogue warp of the unit
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (CG == 2) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    else tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();  // CTA-local visibility of the barrier inits and of the TMEM base address
  if constexpr (CG == 2) cluster_sync_relaxed();  // peer's barriers initialised (fence.mbarrier_init) before remote arrives
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; activations are touched only below
  if (args.prof != nullptr && threadIdx.x == 0) atomicMin(args.prof, globaltimer_ns());

  if (warp == 0) {
    if (elect_one()) {
      // ===== TMA producer =====
      int stage = 0;
      uint32_t phase = 0;
      for (int it = unit; it < num_items; it += num_units) {
        const GemmItem g = gemm_item<BLOCK_N>(it, num_n, args.full_tiles, args.tail_split, num_m, args.reverse_m);
        const int a_row = g.m_idx * kTileM + static_cast<int>(rank) * kGemmBlockM;
        const int b_rows = g.width / CG;  // weight rows staged by this CTA
        const int b_row = g.n0 + static_cast<int>(rank) * b_rows;
        const CUtensorMap* tb = g.width == BLOCK_N ? &tmap_b : &tmap_b_tail;  // box = b_rows x 64
        const uint32_t stage_tx = Cfg::kABytes + b_rows * kGemmBlockK * 2;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          if constexpr (CG == 2) {
            tma_load_2d_pair(sa, &tmap_a, &full_bar[stage], kb * kGemmBlockK, a_row, kEvictNormal);
            tma_load_2d_pair(sb, tb, &full_bar[stage], kb * kGemmBlockK, b_row, kEvictLast);
            if (leader) mbar_expect_tx(&full_bar[stage], stage_tx * 2);
            else mbar_arrive_cluster(&full_bar[stage], 0);
          } else {
            mbar_expect_tx(&full_bar[stage], stage_tx);
            tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kGemmBlockK, a_row, kEvictNormal);
            tma_load_2d(sb, tb, &full_bar[stage], kb * kGemmBlockK, b_row, kEvictLast);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && elect_one()) {
      // ===== MMA issuer (leader CTA only) =====
      constexpr uint32_t idesc_full = umma_idesc_f16(/*bf16*/ 1, kTileM, BLOCK_N);
      const uint32_t idesc_tail = umma_idesc_f16(/*bf16*/ 1, kTileM, BLOCK_N / args.tail_split);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int it = unit; it < num_items; it += num_units) {
        const uint32_t idesc = it < args.full_tiles ? idesc_full : idesc_tail;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t a_desc = umma_smem_desc_sw128(sa);
          const uint64_t b_desc = umma_smem_desc_sw128(sa + Cfg::kABytes);
#pragma unroll
          for (int k = 0; k < kGemmBlockK / 16; ++k) {
            // 16 bf16 = 32 B along K inside the swizzle atom -> +2 in the (addr >> 4) field
            if constexpr (CG == 2) umma_f16_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
            else umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
          }
          // frees the smem slot (in both CTAs) once these MMAs retire
          if constexpr (CG == 2) umma_commit_pair(&empty_bar[stage]);
          else umma_commit(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if constexpr (CG == 2) umma_commit_pair(&tmem_full[as]);  // accumulator complete
        else umma_commit(&tmem_full[as]);
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: 8 warps.  Warp w reads TMEM lanes [32 (w%4), +32) = 32 rows of the tile (the hardware's lane
    // quarter rule); the two warps of a quarter split every 64-column group into its two 32-column halves, so each
    // SM sub-partition has two epilogue warps to hide tcgen05.ld / shared / global latency behind each other.
    const int ew = warp - 4;
    const int q = ew & 3;
    const int hsel = ew >> 2;
    constexpr int kChunks = BLOCK_N / 64;  // 32-column chunks per warp
    uint8_t* stage = smem_epi + ew * Epi::kWarpBytes;
    uint8_t* my_row = stage + lane * Epi::kPitch;
    float* bias_s = reinterpret_cast<float*>(smem_epi + 8 * Epi::kWarpBytes + ew * Epi::kBiasWarpBytes);
    constexpr int kColsPerUnit = 16 / Epi::kElt;
    constexpr int kIters = Epi::kUnitsPerRow;  // 16-B units per lane per chunk: 4 (bf16) | 8 (fp32)
    int as = 0;
    uint32_t aphase = 0;
    for (int it = unit; it < num_items; it += num_units) {
      const GemmItem g = gemm_item<BLOCK_N>(it, num_n, args.full_tiles, args.tail_split, num_m, args.reverse_m);
      const int row0 = g.m_idx * kTileM + static_cast<int>(rank) * kGemmBlockM + q * 32;  // first row of this warp
      const int col_base = g.n0 + hsel * kEpiCols;
      const int n_lim = min(args.N, g.n0 + g.width);  // columns of this item (a tail strip is narrower than BLOCK_N)
      // this tile's bias values for the warp's columns go to shared memory while the MMAs are still running; the chunk
      // loop then reads them as broadcast ld.shared (a global/L2 round trip per chunk sat on the critical path: ncu
      // showed the bias FADDs as the top long_scoreboard stall of the epilogue)
      if (args.bias != nullptr) {
        if (lane < kChunks * 8) {
          const int col = col_base + (lane >> 3) * 64 + (lane & 7) * 4;
          float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
          if (col < n_lim) bb = __ldg(reinterpret_cast<const float4*>(args.bias + col));
          *reinterpret_cast<float4*>(bias_s + lane * 4) = bb;
        }
        __syncwarp();
      }
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      uint32_t v[2][32];
      const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + hsel * kEpiCols;
      tmem_ld_32x32(taddr0, v[0]);
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int col0 = col_base + c * 64;
        tmem_ld_wait();
        if (c + 1 < kChunks) {
          tmem_ld_32x32(taddr0 + (c + 1) * 64, v[(c + 1) & 1]);
        } else {
          // every accumulator value of this warp is in registers: hand the TMEM stage back to the MMA issuer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (leader) mbar_arrive(&tmem_empty[as]);
            else mbar_arrive_cluster(&tmem_empty[as], 0);
          }
        }
        if (col0 < n_lim) {  // warp-uniform (N and the strip width are multiples of 32)
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[c & 1][j]);
          if (args.bias != nullptr) {
            const float4* b4 = reinterpret_cast<const float4*>(bias_s + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 bb = b4[j];
              f[4 * j] += bb.x; f[4 * j + 1] += bb.y; f[4 * j + 2] += bb.z; f[4 * j + 3] += bb.w;
            }
          }
          if constexpr (EPI == EPI_BF16_GELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = quick_gelu(f[j]);
          }
          if constexpr (Epi::kOutBf16) {
            uint4* o = reinterpret_cast<uint4*>(my_row);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 w;
              w.x = pack_bf16x2(f[8 * j], f[8 * j + 1]);
              w.y = pack_bf16x2(f[8 * j + 2], f[8 * j + 3]);
              w.z = pack_bf16x2(f[8 * j + 4], f[8 * j + 5]);
              w.w = pack_bf16x2(f[8 * j + 6], f[8 * j + 7]);
              o[j] = w;
            }
          } else {
            float4* o = reinterpret_cast<float4*>(my_row);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
          }
          __syncwarp();
          // ---- coalesced copy-out: 16-B units, consecutive lanes walk along a row
#pragma unroll
          for (int it = 0; it < kIters; ++it) {
            const int u = it * 32 + lane;
            const int r = u / Epi::kUnitsPerRow, cu = u % Epi::kUnitsPerRow;
            const int row = row0 + r, col = col0 + cu * kColsPerUnit;
            if (row < args.M && col < n_lim) {
              const uint4 sv = *reinterpret_cast<const uint4*>(stage + r * Epi::kPitch + cu * 16);
              if constexpr (Epi::kOutBf16) {
                *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(args.out) + static_cast<size_t>(row) * args.ldo + col) = sv;
              } else {
                float4 x = *reinterpret_cast<const float4*>(&sv);
                size_t out_row = row;
                if constexpr (EPI == EPI_F32_RESID) {
                  // x += acc + bias as a fire-and-forget 16-B reduction performed in L2: every element receives exactly
                  // one add per launch, so the result is deterministic and no residual value ever travels to the SM
                  red_add_f32x4(static_cast<float*>(args.out) + out_row * args.ldo + col, x);
                  continue;
                }
                if constexpr (EPI == EPI_PATCH) {
                  const int bimg = row / args.patches, pp = row - bimg * args.patches;
                  out_row = static_cast<size_t>(bimg) * (args.patches + 1) + 1 + pp;
                  const float4 pe = __ldg(reinterpret_cast<const float4*>(args.pos + static_cast<size_t>(1 + pp) * args.N + col));
                  x.x += pe.x; x.y += pe.y; x.z += pe.z; x.w += pe.w;
                }
                *reinterpret_cast<float4*>(static_cast<float*>(args.out) + out_row * args.ldo + col) = x;
              }
            }
          }
          __syncwarp();  // the staging tile is rewritten by the next chunk
        }
      }
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all();
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (CG == 2) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    else tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
  if (args.prof != nullptr && threadIdx.x == 0) atomicMax(args.prof + 1, globaltimer_ns());
}

// host launcher (gemm_tc.cu)
int launch_gemm_bf16(const void* A, int lda, const void* W, int ldw, const GemmArgs& args, int epilogue,
                     cudaStream_t stream);

}  // namespace evr

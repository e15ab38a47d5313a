// Attention core on the 5th-generation tensor cores: softmax(q k^T / sqrt(64) [+ causal mask]) v per (sequence, head),
// S = Q K^T and O = P V as tcgen05.mma with both accumulators -- and P itself -- in tensor memory.
//
// Replaces the attention inside nn.MultiheadAttention of openai/CLIP's ResidualAttentionBlock (clip/model.py; call
// sites of the tower: Backend/services/embedding_service.py:41, 414, 490) for the sequence lengths between 65 and 257
// tokens: ViT-L/14 (257), ViT-B/16 (197), the text tower (77, causal).  SURVEY.md 2.2 row K5.
//
// One persistent CTA per SM walks (sequence, head) items.  Roles (384 threads):
//   warp 0      TMA producer: the item's Q, K and V head slices ([L, 64] bf16 each, 128-byte rows) land in a
//               shared-memory stage as 128B-swizzled tiles straight from the packed qkv activation [B L, 3 d].
//   warp 1      MMA issuer (one elected thread).  Per 128-query tile ("job"):
//                 S[128 x n] = Q K^T        4 x tcgen05.mma kind::f16, A = Q tile, B = K tile, both K-major (K = 64 dims)
//                 O[128 x 64] = P V         one tcgen05.mma per 16 keys with A = P read from TENSOR MEMORY (bf16, two
//                                           keys per 32-bit column) and B = the V tile used as it lies in memory (keys x
//                                           dims = MN-major operand: no transposed copy of V anywhere)
//                 l[128]     = P 1          the same P against a tile of ones: the softmax denominator comes out of the
//                                           tensor core, summed in fp32 over exactly the bf16 weights P V used
//   warps 2-9   two softmax groups of 128 threads, one per TMEM region; thread = query row (TMEM lane).  Two passes over
//               the row of S in TMEM (max, then exp2), 64 columns in flight while the previous 64 are processed; P is
//               written back over S in place with tcgen05.st; then O and l are read, O / l stored as 128-byte rows.
//   warps 10-11 L = 257 only: token 256, with warp-level MMA (mma.sync m16n8k16 + ldmatrix on the same swizzled tiles):
//               warp 10 its scores as a KEY against the 256 tiled rows (handed to the softmax threads through shared
//               memory), warp 11 its whole attention row as a QUERY, so the tcgen05 pipeline never runs a 128-row job
//               for one live row.
// TMEM (512 columns) = 2 regions x 256: S occupies a region's 256 columns; P (bf16) is written over its first 136, O
// lands on columns 136..199 and l on 200..215 -- columns of S the softmax has consumed by then.  While one group is in
// its softmax the issuer runs the other region's P V and the next job's Q K^T.
// L = 257 = 2 x 128 + 1: the 257th KEY's weight goes into P column 128, so P V and the denominator take it in through a
// 17th K-step over rows 256..271 of the V tile.
//
// Roofline: HBM/L2 -- every q, k, v element is read once and every output written once (8 d bytes per token per layer);
// the exponent unit (16 exp2 / clk / SM) is the second bound at L = 257.
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"
#include "vit_kernels.cuh"

namespace evr {

constexpr int kAtcThreads = 384;
constexpr int kAtcTileBytes = 128 * 128;  // 128 rows x 64 bf16
constexpr int kAtcLoneRows = 16;          // rows 256..271 of the sequence travel as one extra TMA box
constexpr int kAtcOCol = 136;             // O accumulator columns inside a region (P occupies 0..135)
constexpr int kAtcLCol = 200;             // softmax denominator (16 identical columns)
constexpr int kAtcOnesBytes = 2048;
constexpr int kAtcLoneBytes = 4 * 1024 + 1152 + 640;  // [4 stages][256] key scores, 272 row scores, 272 bf16 weights

struct AtcArgs {
  int B, L, d, heads, causal;
  int n_mtiles;    // 128-row query tiles = jobs per item (1 or 2)
  int lone;        // 1 when L == 257: token 256 is handled outside the 256-wide tiles
  int n_main;      // keys per S tile = round_up(min(L, 256), 16)
  int rows_alloc;  // smem rows per operand per stage = n_mtiles * 128 + 16 * lone
  int stages;
  int pack;        // 1 when L <= 64: two (sequence, head) items share a 128-row tile, rows 0..63 | 64..127
  float c1;        // softmax scale * log2(e)
  __nv_bfloat16* out;
  unsigned long long* prof;  // evr_profile_enable(2): device-timer slot of this launch, else nullptr
  long long* dbg;  // developer timeline (EVR_ATTN_DBG): [event][job] SM clock of CTA 0, else nullptr
};

// TMEM -> registers, 32 lanes x 16 columns
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM, 32 lanes x 8 columns
__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
      : "memory");
}
// registers -> TMEM, 32 lanes x 16 columns
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
// registers -> TMEM, 32 lanes x 32 columns
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
        "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
        "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem]; A: 128 lanes x K 16-bit elements packed two per column.
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// dot product / axpy of a 64-element bf16 row in a 128B-swizzled tile (row r of the tile, 8 x 16-byte chunks)
__device__ __forceinline__ const uint4* swz_chunk(const uint8_t* tile, int r, int c) {
  return reinterpret_cast<const uint4*>(tile + r * 128 + ((c ^ (r & 7)) << 4));
}
__device__ __forceinline__ float dot8(uint4 a, uint4 b) {
  const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 p = __bfloat1622float2(x[i]), q = __bfloat1622float2(y[i]);
    acc = fmaf(p.x, q.x, acc);
    acc = fmaf(p.y, q.y, acc);
  }
  return acc;
}

__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
  return v;
}
// 16-byte chunk c of row r of a 128B-swizzled [rows x 64 bf16] tile at shared address `tile`
__device__ __forceinline__ uint32_t swz_addr(uint32_t tile, int r, int c) {
  return tile + r * 128 + ((c ^ (r & 7)) << 4);
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// No-swizzle K-major descriptor over the tile of ones: every element is 1.0, so only the footprint matters
// (16 x 16 elements: core matrices 128 B apart in both directions, < 512 B in all).
__device__ __forceinline__ uint64_t ones_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(128 >> 4) << 16;
  d |= static_cast<uint64_t>(128 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

// ---- softmax passes over one 32-column chunk held in registers.  kMask: keys >= kmax are excluded (causal rows);
// the unmasked form is what the vision towers run -- their only partial chunk is the 16-column tail.
// (Code size matters here: the two softmax groups, the issuer and the lone-row warps all run different code at the
// same time, and ncu showed `no instruction` stalls once the unrolled variants outgrew the instruction cache.)
template <bool kMask>
__device__ __forceinline__ float chunk_max(const uint32_t (&v)[32], float m, int c, int kmax) {
  if constexpr (!kMask) {
    float m4[4] = {m, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
#pragma unroll
      for (int u = 0; u < 4; ++u) m4[u] = max3(m4[u], __uint_as_float(v[i + 2 * u]), __uint_as_float(v[i + 2 * u + 1]));
    }
    return fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (c + i < kmax) m = fmaxf(m, __uint_as_float(v[i]));
    return m;
  }
}
#ifndef ATC_EXPERIMENT
#define ATC_EXPERIMENT 0
#endif
__device__ __forceinline__ float exw(float t) {
#if ATC_EXPERIMENT == 1
  return t * 0.001f;   // timing experiment: no MUFU
#else
  return ex2(t);
#endif
}
template <bool kMask>
__device__ __forceinline__ void chunk_exp(const uint32_t (&v)[32], uint32_t (&pk)[16], float c1, float mc, int c,
                                          int kmax) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float p0 = exw(fmaf(__uint_as_float(v[2 * i]), c1, -mc));
    float p1 = exw(fmaf(__uint_as_float(v[2 * i + 1]), c1, -mc));
    if constexpr (kMask) {
      if (c + 2 * i >= kmax) p0 = 0.f;
      if (c + 2 * i + 1 >= kmax) p1 = 0.f;
    }
    pk[i] = pack_bf16x2(p0, p1);
  }
}

// masked maximum of a 32-column chunk, four independent chains (with a constant kmax the tests fold away)
__device__ __forceinline__ float chunk_max4(const uint32_t (&v)[32], int c, int kmax) {
  float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int i = 0; i < 32; ++i)
    if (c + i < kmax) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
  return fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
}
// masked exp2 of a 32-column chunk + the fp32 sum of the weights (pack mode: the thread owns its whole row, so the
// softmax denominator needs no pass through the tensor core)
__device__ __forceinline__ float chunk_exp_sum(const uint32_t (&v)[32], uint32_t (&pk)[16], float c1, float mc, int c,
                                               int kmax) {
  float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float p0 = c + 2 * i < kmax ? exw(fmaf(__uint_as_float(v[2 * i]), c1, -mc)) : 0.f;
    const float p1 = c + 2 * i + 1 < kmax ? exw(fmaf(__uint_as_float(v[2 * i + 1]), c1, -mc)) : 0.f;
    s4[i & 3] += p0 + p1;
    pk[i] = pack_bf16x2(p0, p1);
  }
  return (s4[0] + s4[1]) + (s4[2] + s4[3]);
}

// ---- warp-level MMA (m16n8k16, bf16 in, fp32 accumulate) for the lone warp: one warp, two small matrix products per
// item, operands straight out of the 128B-swizzled tiles with ldmatrix
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t saddr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}

// Token 256 of a 257-token sequence (L = 257 = 2 x 128 + 1) is handled by two warps with warp-level MMA.  Both of its
// roles reduce to "a tile of rows against one vector":  dst[r] = tile[r, :] . vec  is the matrix product
// [rows x 64] x [64 x 8] whose B operand carries `vec` in column 0 (fragments of lanes 0..3), so the result sits in column
// 0 of the accumulators (lanes with lane % 4 == 0).
//   as a KEY   (warp 10): tile = Q rows 0..255, vec = k_256  -> its score against every tiled query row
//   as a QUERY (warp 11): tile = K rows 0..271, vec = q_256  -> its own row of scores; softmax across the warp; then
//                         O^T[64 x 1] = V^T[64 x 272] p[272 x 1], again a product with the vector in column 0, the V^T
//                         fragments coming from the [key][dim] tile through ldmatrix.trans.
__device__ __forceinline__ void load_vec_frag(uint32_t tile, uint32_t (&vf)[4][2], int lane) {
  const int g = lane >> 2, t = lane & 3;   // row 256 of the operand: (256 & 7) == 0, its chunks are not permuted
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) vf[kk][hf] = g == 0 ? lds32(tile + 256 * 128 + (16 * kk + 8 * hf + 2 * t) * 2) : 0u;
  }
}
// dst[r] = scale * tile[r, :] . vec for r in [0, 16 * nblocks); four row blocks in flight (independent accumulators)
__device__ __forceinline__ void rows_dot_vec(uint32_t tile, int nblocks, const uint32_t (&vf)[4][2], float scale,
                                             uint32_t dst, int lane) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll 1
  for (int mb0 = 0; mb0 < nblocks; mb0 += 4) {
    float c[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) c[u][0] = c[u][1] = c[u][2] = c[u][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (mb0 + u < nblocks) {
          uint32_t af[4];
          ldsm_x4(af, swz_addr(tile, 16 * (mb0 + u) + ((lane >> 3) & 1) * 8 + (lane & 7), 2 * kk + (lane >> 4)));
          mma16816(c[u], af[0], af[1], af[2], af[3], vf[kk][0], vf[kk][1]);
        }
      }
    }
    if (t == 0) {   // column 0 of the product
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (mb0 + u < nblocks) {
          sts32(dst + (16 * (mb0 + u) + g) * 4, __float_as_uint(c[u][0] * scale));
          sts32(dst + (16 * (mb0 + u) + g + 8) * 4, __float_as_uint(c[u][2] * scale));
        }
      }
    }
  }
}

__device__ __noinline__ void lone_key_scores(uint32_t qs, int op_bytes, uint32_t col, int lane) {
  uint32_t kf[4][2];
  load_vec_frag(qs + op_bytes, kf, lane);
  rows_dot_vec(qs, 16, kf, 1.0f, col, lane);   // raw scores, like S in tensor memory
}

__device__ __noinline__ void lone_query_row(uint32_t qs, int op_bytes, float c1, uint32_t row_s, uint32_t row_p,
                                            __nv_bfloat16* out_row, int lane) {
  const uint32_t ks = qs + op_bytes, vs = ks + op_bytes;
  const int g = lane >> 2, t = lane & 3;
  uint32_t qf[4][2];
  load_vec_frag(qs, qf, lane);
  rows_dot_vec(ks, 17, qf, c1, row_s, lane);   // keys 0..271, in the exponent's units (257..271: other rows, masked)
  __syncwarp();
  float sc[9], m = -INFINITY;
#pragma unroll
  for (int i = 0; i < 9; ++i) {   // lane l takes keys l, l + 32, ...
    const int key = lane + 32 * i;
    sc[i] = key < 257 ? __uint_as_float(lds32(row_s + key * 4)) : -INFINITY;
    m = fmaxf(m, sc[i]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    const int key = lane + 32 * i;
    if (key < 272) {
      const __nv_bfloat16 pb = __float2bfloat16_rn(ex2(sc[i] - m));   // exp2(-inf) = 0 for the masked slots
      sum += __bfloat162float(pb);
      asm volatile("st.shared.b16 [%0], %1;" ::"r"(row_p + key * 2), "h"(__bfloat16_as_ushort(pb)) : "memory");
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  __syncwarp();
  // O^T = V^T p: A = V^T blocks ([16 dims x 16 keys], ldmatrix.trans of the [key][dim] tile), B = p in column 0
  float acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u) acc[u][0] = acc[u][1] = acc[u][2] = acc[u][3] = 0.f;
#pragma unroll 1
  for (int kk = 0; kk < 17; ++kk) {
    const uint32_t b0 = g == 0 ? lds32(row_p + (16 * kk + 2 * t) * 2) : 0u;
    const uint32_t b1 = g == 0 ? lds32(row_p + (16 * kk + 8 + 2 * t) * 2) : 0u;
    const int key = 16 * kk + (lane >> 4) * 8 + (lane & 7);
#pragma unroll
    for (int mb = 0; mb < 4; ++mb) {   // dims 16 mb .. 16 mb + 15
      uint32_t af[4];
      ldsm_x4_trans(af, swz_addr(vs, key, 2 * mb + ((lane >> 3) & 1)));
      mma16816(acc[mb], af[0], af[1], af[2], af[3], b0, b1);
    }
  }
  if (t == 0) {   // column 0: dims 16 mb + g and 16 mb + g + 8
    const float inv = 1.0f / sum;
#pragma unroll
    for (int mb = 0; mb < 4; ++mb) {
      out_row[16 * mb + g] = __float2bfloat16_rn(acc[mb][0] * inv);
      out_row[16 * mb + g + 8] = __float2bfloat16_rn(acc[mb][2] * inv);
    }
  }
}

// Waits of the softmax / lone warps: a spinning warp takes issue slots from the warps that share its scheduler (ncu:
// a third of all issued instructions were try_wait loops), so these back off between polls.
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
  uint32_t polls = 0;
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(64);
    if (++polls > (1u << 26)) mbar_timeout(smem_u32(bar), parity);   // > 4 s: a protocol bug traps instead of hanging
  }
}

constexpr int kDbgJobs = 64;
__device__ __forceinline__ void dbg_stamp(const AtcArgs& a, int event, int g) {
  if (a.dbg != nullptr && blockIdx.x == 0 && g < kDbgJobs) a.dbg[event * kDbgJobs + g] = clock64();
}

// kPack (L <= 64, ViT-B/32's 50 tokens): a job is a PAIR of items.  Q, K and V tiles hold item 0 in rows 0..63 and item
// 1 in rows 64..127 (two 64-row boxes of a [B][L][3 d] tensor map: rows past L arrive as zeros); S = Q K^T is the full
// 128 x 128 product of which each row uses its own item's 64 columns; P carries zeros in the other item's columns, so
// one P V over 128 keys serves both items.  3 072 items per ViT-B/32 layer become 1 536 tensor-core jobs whose loads
// run four stages (192 KB) ahead -- the warp-level MMA kernel spent 2 500 warp instructions per item.
// kPack: 0 = one item per job; 1 = pairs, sequence length read from the arguments; 50 = pairs, L = 50 folded into the
// code (every mask of the softmax becomes a constant: the generic pack code spent a third of its instructions on them)
template <bool kCausal, int kPack>
__global__ void __launch_bounds__(kAtcThreads, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap128, const __grid_constant__ CUtensorMap tmap_lone,
                    const AtcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int op_bytes = a.rows_alloc * 128;   // one operand (Q, K or V) of one stage
  const int stage_bytes = 3 * op_bytes;
  uint8_t* ones = smem + a.stages * stage_bytes;
  uint8_t* lone_mem = ones + kAtcOnesBytes;        // [4][256] fp32 key-256 scores | 272 fp32 | 272 bf16
  uint64_t* bars = reinterpret_cast<uint64_t*>(lone_mem + kAtcLoneBytes);
  uint64_t* full = bars;            // [4]   TMA landed
  uint64_t* empty = bars + 4;       // [4]   stage may be reloaded
  constexpr int kRegions = kPack != 0 ? 4 : 2;   // TMEM regions = jobs in flight (pack: 4 x 128 columns, else 2 x 256)
  uint64_t* s_full = bars + 8;      // [kRegions]  S of the region is complete
  uint64_t* p_ready = bars + 12;    // [kRegions]  P is in TMEM
  uint64_t* o_full = bars + 16;     // [kRegions]  O (and l) of the region are complete
  uint64_t* r_free = bars + 20;     // [kRegions]  the region's O has been read out
  uint64_t* col_ready = bars + 24;  // [4]   key 256's scores of the stage's item are in shared memory
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 28);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_jobs = a.n_mtiles;
  const int n_items = a.B * a.heads;
  const int items = kPack != 0 ? (n_items + 1) / 2 : n_items;   // units of work the CTAs walk

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap128);
    tma_prefetch_desc(&tmap_lone);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < a.stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], kPack != 0 ? 1 : 1 + 4 * n_jobs + 2 * a.lone);   // pack: only the P V issuer releases a stage
      mbar_init(&col_ready[i], 1);
    }
    for (int i = 0; i < kRegions; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&p_ready[i], 4); mbar_init(&o_full[i], 1); mbar_init(&r_free[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  for (int i = threadIdx.x; i < kAtcOnesBytes / 4; i += kAtcThreads) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;
  fence_proxy_async();   // the tile of ones is read by the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();  // qkv is the previous kernel's output; `out` may still be read by the one before it
  prof_stamp_begin(a.prof);

  if (warp == 0) {
    if (elect_one()) {
      // ===== TMA producer =====
      const uint32_t tx = 3u * (a.n_mtiles * kAtcTileBytes + a.lone * kAtcLoneRows * 128);
      int it = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
        const int b = item / a.heads, h = item - b * a.heads;
        const int st = it % a.stages;
        dbg_stamp(a, 9, it);
        mbar_wait(&empty[st], ((it / a.stages) & 1) ^ 1);
        dbg_stamp(a, 10, it);
        mbar_expect_tx(&full[st], tx);
        uint8_t* base = smem + st * stage_bytes;
        if constexpr (kPack != 0) {
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            const int single = min(2 * item + half, n_items - 1);   // an odd tail repeats its item (finite data, not stored)
            const int pb = single / a.heads, ph = single - pb * a.heads;
#pragma unroll 1
            for (int which = 0; which < 3; ++which)
              tma_load_3d(base + which * op_bytes + half * (kAtcTileBytes / 2), &tmap_lone, &full[st],
                          which * a.d + ph * 64, 0, pb, kEvictFirst);
          }
          continue;
        }
        const int row0 = b * a.L;
#pragma unroll 1
        for (int which = 0; which < 3; ++which) {
          uint8_t* dst = base + which * op_bytes;
          const int col = which * a.d + h * 64;
          for (int t = 0; t < a.n_mtiles; ++t)
            tma_load_2d(dst + t * kAtcTileBytes, &tmap128, &full[st], col, row0 + t * 128, kEvictFirst);
          if (a.lone)
            tma_load_2d(dst + a.n_mtiles * kAtcTileBytes, &tmap_lone, &full[st], col, row0 + 256, kEvictFirst);
        }
      }
    }
  } else if (kPack != 0 && (warp == 1 || warp == 10)) {
    // ===== pack mode: two issuers, so that neither product waits behind the other's barrier.  Warp 1: S = Q K^T of job g
    // as soon as its stage has landed and region g % 4 has been read out; warp 10: O = P V of job g as soon as its P
    // is in tensor memory (O lands on columns 64..127 of the region, over the half of S every row has consumed). =====
    if (elect_one()) {
      int g = 0;
      if (warp == 1) {
        const uint32_t idesc_qk = umma_idesc_f16(1, 128, 128);
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++g) {
          const int st = g % a.stages, r = g & 3;
          dbg_stamp(a, 11, g);
          mbar_wait(&full[st], (g / a.stages) & 1);
          dbg_stamp(a, 12, g);
          mbar_wait(&r_free[r], ((g >> 2) & 1) ^ 1);
          dbg_stamp(a, 0, g);
          tc_fence_after();
          const uint32_t q_addr = smem_u32(smem + st * stage_bytes);
          const uint64_t q_desc = umma_smem_desc_sw128(q_addr), k_desc = umma_smem_desc_sw128(q_addr + op_bytes);
#pragma unroll
          for (int kb = 0; kb < 4; ++kb)
            umma_f16(tmem_base + r * 128, q_desc + 2 * kb, k_desc + 2 * kb, idesc_qk, kb != 0);
          umma_commit(&s_full[r]);
          dbg_stamp(a, 1, g);
        }
      } else {
        const uint32_t idesc_pv = umma_idesc_f16(1, 128, 64) | (1u << 16);  // B (= V) is MN-major
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++g) {
          const int st = g % a.stages, r = g & 3;
          mbar_wait(&full[st], (g / a.stages) & 1);   // long complete; this thread observes the TMA writes of V itself
          mbar_wait_backoff(&p_ready[r], (g >> 2) & 1);
          dbg_stamp(a, 2, g);
          tc_fence_after();
          const uint32_t region = tmem_base + r * 128;
          const uint64_t v_desc = umma_smem_desc_sw128(smem_u32(smem + st * stage_bytes + 2 * op_bytes));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_f16_ts(region + 64, region + ks * 8, v_desc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_pv, ks != 0);
          umma_commit(&o_full[r]);
          umma_commit(&empty[st]);   // Q and K were consumed before P could exist: the stage is free once P V retires
          dbg_stamp(a, 3, g);
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // ===== MMA issuer =====
      const uint32_t idesc_qk = umma_idesc_f16(1, 128, a.n_main);
      const uint32_t idesc_pv = umma_idesc_f16(1, 128, 64) | (1u << 16);  // B (= V) is MN-major: [keys][dims] in smem
      const uint32_t idesc_l = umma_idesc_f16(1, 128, 16);
      const uint64_t one_desc = ones_desc(smem_u32(ones));
      const int ksteps = a.n_main / 16 + a.lone;
      int g = 0, it = 0;
      // the P V of job g - 1 is issued after the Q K^T of job g, so the softmax of g - 1 overlaps the latter
      int prev_stage = 0, prev_last = 0, have_prev = 0;
      auto issue_pv = [&](int gp, int stage, int last_of_item) {
        const int rp = gp & 1, np = gp >> 1;
        mbar_wait(&p_ready[rp], np & 1);
        dbg_stamp(a, 2, gp);
        tc_fence_after();
        const uint32_t region = tmem_base + rp * 256;
        const uint64_t v_desc = umma_smem_desc_sw128(smem_u32(smem + stage * stage_bytes + 2 * op_bytes));
        for (int ks = 0; ks < ksteps; ++ks) {  // 16 keys per step: 8 TMEM columns of P, two 8-key swizzle atoms of V
          umma_f16_ts(region + kAtcOCol, region + ks * 8, v_desc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_pv,
                      ks != 0);
          umma_f16_ts(region + kAtcLCol, region + ks * 8, one_desc, idesc_l, ks != 0);
        }
        umma_commit(&o_full[rp]);
        dbg_stamp(a, 3, gp);
        if (last_of_item) umma_commit(&empty[stage]);
      };
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
        const int st = it % a.stages;
        dbg_stamp(a, 11, it);
        mbar_wait(&full[st], (it / a.stages) & 1);
        dbg_stamp(a, 12, it);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(smem + st * stage_bytes);
        const uint64_t k_desc = umma_smem_desc_sw128(q_addr + op_bytes);
        for (int j = 0; j < n_jobs; ++j, ++g) {
          const int r = g & 1, n = g >> 1;
          dbg_stamp(a, 0, g);
          mbar_wait(&r_free[r], (n & 1) ^ 1);
          tc_fence_after();
          const uint64_t q_desc = umma_smem_desc_sw128(q_addr + j * kAtcTileBytes);
#pragma unroll
          for (int kb = 0; kb < 4; ++kb)
            umma_f16(tmem_base + r * 256, q_desc + 2 * kb, k_desc + 2 * kb, idesc_qk, kb != 0);
          umma_commit(&s_full[r]);
          dbg_stamp(a, 1, g);
          if (have_prev) issue_pv(g - 1, prev_stage, prev_last);
          prev_stage = st; prev_last = (j == n_jobs - 1); have_prev = 1;
        }
      }
      if (have_prev) issue_pv(g - 1, prev_stage, prev_last);
    }
  } else if (warp < 10) {
    // ===== softmax / epilogue: warps 2..9; group = TMEM region, quarter = TMEM lane quadrant (warp % 4) =====
    const int group = (warp - 2) >> 2, quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const int key_end = a.L < 256 ? a.L : 256;   // keys that live in the S tile
    // 32-column chunks without a masked key (every chunk of a causal row is masked); the rest goes 16 columns at a time
    const int nfull = kCausal ? a.n_main / 32 : key_end / 32;
    int g = 0, it = 0;
    if constexpr (kPack != 0) {
      // rows 0..63 belong to item 2 * pair, rows 64..127 to item 2 * pair + 1; a warp's 32 rows sit in one half.  The
      // group walks jobs group, group + 2, ...: softmax of job g, then the read-out of its previous job g - 2, whose
      // P V has had a whole softmax to finish.
      const int L = kPack > 1 ? kPack : a.L;
      const int half = quarter >> 1, r_in = row & 63;
      const bool live = (quarter & 1) * 32 < L;      // warp-uniform
      float inv_prev = 0.f;
      int pair_prev = -1, g_prev = 0;
      auto finish = [&](int gj, int pair, float inv) {
        const int r = gj & 3;
        if (quarter == 0 && lane == 0) dbg_stamp(a, 5, gj);
        mbar_wait_backoff(&o_full[r], (gj >> 2) & 1);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 7, gj);
        tc_fence_after();
        uint32_t o[64];
        if (live) {
          const uint32_t region = tmem_base + r * 128 + lane_addr;
          tmem_ld_32x32(region + 64, *reinterpret_cast<uint32_t(*)[32]>(&o[0]));
          tmem_ld_32x32(region + 96, *reinterpret_cast<uint32_t(*)[32]>(&o[32]));
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&r_free[r]);
        const int single = 2 * pair + half;
        if (live && r_in < L && single < n_items) {
          const int ob = single / a.heads, oh = single - ob * a.heads;
          uint4* dst = reinterpret_cast<uint4*>(a.out + (static_cast<size_t>(ob) * L + r_in) * a.d + oh * 64);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(o[c * 8 + 0]) * inv, __uint_as_float(o[c * 8 + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(o[c * 8 + 2]) * inv, __uint_as_float(o[c * 8 + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(o[c * 8 + 4]) * inv, __uint_as_float(o[c * 8 + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(o[c * 8 + 6]) * inv, __uint_as_float(o[c * 8 + 7]) * inv);
            dst[c] = w;
          }
        }
        if (quarter == 0 && lane == 0) dbg_stamp(a, 8, gj);
      };
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++g) {
        if ((g & 1) != group) continue;
        const int r = g & 3;
        const uint32_t region = tmem_base + r * 128 + lane_addr;
        mbar_wait_backoff(&s_full[r], (g >> 2) & 1);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 4, g);
        tc_fence_after();
        float inv = 0.f;
        if (live) {
          uint32_t va[32], vb[32], pk[16];
          tmem_ld_32x32(region + half * 64, va);
          tmem_ld_32x32(region + half * 64 + 32, vb);
          tmem_ld_wait();
          const float m = fmaxf(chunk_max4(va, 0, L), chunk_max4(vb, 32, L));
          const float mc = m * a.c1;
          // every score of the row is in registers: P (own 64 keys -> 32 columns) and the other item's zeros may
          // overwrite S
          float sum = chunk_exp_sum(va, pk, a.c1, mc, 0, L);
          tmem_st_32x16(region + half * 32, pk);
          sum += chunk_exp_sum(vb, pk, a.c1, mc, 32, L);
          tmem_st_32x16(region + half * 32 + 16, pk);
          const uint32_t zero[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll
          for (int c = 0; c < 32; c += 8) tmem_st_32x8(region + (1 - half) * 32 + c, zero);
          tmem_st_wait();
          inv = 1.0f / sum;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_ready[r]);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 6, g);
        if (pair_prev >= 0) finish(g_prev, pair_prev, inv_prev);
        pair_prev = item; g_prev = g; inv_prev = inv;
      }
      if (pair_prev >= 0) finish(g_prev, pair_prev, inv_prev);
    } else
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      const int b = item / a.heads, h = item - b * a.heads;
      const int st = it % a.stages;
      const uint32_t qs = smem_u32(smem + st * stage_bytes);
      for (int j = 0; j < n_jobs; ++j, ++g) {
        if ((g & 1) != group) continue;
        const int n = g >> 1;
        const uint32_t region = tmem_base + group * 256 + lane_addr;
        const int q_row = j * 128 + row;             // token index of this thread's query inside the sequence
        const bool valid = q_row < a.L;
        const bool warp_live = j * 128 + quarter * 32 < a.L;   // warp-uniform: any valid row in this warp
        int kmax = key_end;                           // exclusive bound of the keys this row may attend to
        if (kCausal && q_row + 1 < kmax) kmax = q_row + 1;
        const bool lone_key = a.lone && !kCausal;
        float s_lone = 0.f;
        if (warp_live && a.lone) {   // the 257th key's score against this row, from the lone warp
          mbar_wait_backoff(&col_ready[st], (it / a.stages) & 1);
          s_lone = __uint_as_float(lds32(smem_u32(lone_mem) + (st * 256 + q_row) * 4));
        }
        mbar_wait_backoff(&s_full[group], n & 1);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 4, g);
        tc_fence_after();
        if (warp_live) {
          uint32_t va[32], vb[32];
          // ---- pass 1: row maximum; the next 32 columns are in flight while these are reduced
          float m = lone_key ? s_lone : -INFINITY;
          if (nfull > 0) { tmem_ld_32x32(region, va); tmem_ld_wait(); }
          for (int c = 0; c < nfull; c += 2) {
            if (c + 1 < nfull) tmem_ld_32x32(region + (c + 1) * 32, vb);
            m = chunk_max<kCausal>(va, m, c * 32, kmax);
            if (c + 1 < nfull) {
              tmem_ld_wait();
              if (c + 2 < nfull) tmem_ld_32x32(region + (c + 2) * 32, va);
              m = chunk_max<kCausal>(vb, m, c * 32 + 32, kmax);
              if (c + 2 < nfull) tmem_ld_wait();
            }
          }
          for (int c = nfull * 32; c < a.n_main; c += 16) {
            uint32_t v[16];
            tmem_ld_32x16(region + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (c + i < kmax) m = fmaxf(m, __uint_as_float(v[i]));
          }
          // ---- pass 2: p = exp2((s - m) * scale * log2 e), written back over S as bf16 (two keys per column)
          const float mc = m * a.c1;
          if (quarter == 0 && lane == 0) dbg_stamp(a, 5, g);
          if (nfull > 0) { tmem_ld_32x32(region, va); tmem_ld_wait(); }
          for (int c = 0; c < nfull; c += 2) {
            uint32_t pk[16];
            if (c + 1 < nfull) tmem_ld_32x32(region + (c + 1) * 32, vb);
            chunk_exp<kCausal>(va, pk, a.c1, mc, c * 32, kmax);
            if (c + 1 < nfull) tmem_ld_wait();   // chunk c + 1 is in registers: its columns may be overwritten below
#if ATC_EXPERIMENT != 2
            tmem_st_32x16(region + c * 16, pk);
#endif
            if (c + 1 < nfull) {
              if (c + 2 < nfull) tmem_ld_32x32(region + (c + 2) * 32, va);
              chunk_exp<kCausal>(vb, pk, a.c1, mc, c * 32 + 32, kmax);
              if (c + 2 < nfull) tmem_ld_wait();
#if ATC_EXPERIMENT != 2
              tmem_st_32x16(region + (c + 1) * 16, pk);
#else
              asm volatile("" ::"r"(pk[0] ^ pk[7] ^ pk[15]));
#endif
            }
          }
          for (int c = nfull * 32; c < a.n_main; c += 16) {
            uint32_t v[16], pk[8];
            tmem_ld_32x16(region + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
              const float p0 = c + i < kmax ? ex2(fmaf(__uint_as_float(v[i]), a.c1, -mc)) : 0.f;
              const float p1 = c + i + 1 < kmax ? ex2(fmaf(__uint_as_float(v[i + 1]), a.c1, -mc)) : 0.f;
              pk[i >> 1] = pack_bf16x2(p0, p1);
            }
            tmem_st_32x8(region + (c >> 1), pk);
          }
          if (a.lone) {   // key 256 -> P column 128 (keys 257..271 get weight 0)
            uint32_t pk[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
            if (lone_key) pk[0] = pack_bf16x2(ex2(fmaf(s_lone, a.c1, -mc)), 0.f);
            tmem_st_32x8(region + 128, pk);
          }
          tmem_st_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_ready[group]);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 6, g);
        // ---- O = P V and l = P 1 are on their way; read them, release the region, normalise, store
        mbar_wait_backoff(&o_full[group], n & 1);
        if (quarter == 0 && lane == 0) dbg_stamp(a, 7, g);
        tc_fence_after();
        uint32_t o[64], l[16];
        if (warp_live) {
          tmem_ld_32x32(region + kAtcOCol, *reinterpret_cast<uint32_t(*)[32]>(&o[0]));
          tmem_ld_32x32(region + kAtcOCol + 32, *reinterpret_cast<uint32_t(*)[32]>(&o[32]));
          tmem_ld_32x16(region + kAtcLCol, l);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&r_free[group]);
          mbar_arrive(&empty[st]);   // this warp no longer reads the stage's shared memory
        }
        if (quarter == 0 && lane == 0) dbg_stamp(a, 8, g);
        if (warp_live && valid) {
          const float inv = 1.0f / __uint_as_float(l[0]);
          uint4* dst = reinterpret_cast<uint4*>(a.out + (static_cast<size_t>(b) * a.L + q_row) * a.d + h * 64);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(o[c * 8 + 0]) * inv, __uint_as_float(o[c * 8 + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(o[c * 8 + 2]) * inv, __uint_as_float(o[c * 8 + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(o[c * 8 + 4]) * inv, __uint_as_float(o[c * 8 + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(o[c * 8 + 6]) * inv, __uint_as_float(o[c * 8 + 7]) * inv);
            dst[c] = w;
          }
        }
      }
    }
  }

  if (warp >= 10 && a.lone) {
    // ===== warps 10, 11: token 256 of every item as a key / as a query =====
    const uint32_t lm = smem_u32(lone_mem);
    int it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      const int b = item / a.heads, h = item - b * a.heads;
      const int st = it % a.stages;
      const uint32_t qs = smem_u32(smem + st * stage_bytes);
      mbar_wait_backoff(&full[st], (it / a.stages) & 1);
      if (warp == 10) {
        lone_key_scores(qs, op_bytes, lm + st * 1024, lane);
        __syncwarp();
        if (lane == 0) mbar_arrive(&col_ready[st]);
      } else {
        if (lane == 0) dbg_stamp(a, 13, it);
        lone_query_row(qs, op_bytes, a.c1, lm + 4096, lm + 4096 + 1152,
                       a.out + (static_cast<size_t>(b) * a.L + 256) * a.d + h * 64, lane);
        if (lane == 0) dbg_stamp(a, 14, it);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);   // shared memory of the stage is no longer read
    }
  }

  tc_fence_before();
  __syncthreads();
  prof_stamp_end(a.prof);
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

bool attention_tc_supported(int L, int d, int heads) {
  return heads * 64 == d && L > 64 && L <= 257 && (d * 3) % 8 == 0;
}
bool attention_tc_pack_supported(int L, int d, int heads, int causal) {   // two items per tile
  return heads * 64 == d && L >= 1 && L <= 64 && !causal && (d * 3) % 8 == 0;
}

int launch_attention_tc(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                        cudaStream_t stream) {
  const bool pack = L <= 64;
  EVR_REQUIRE(pack ? attention_tc_pack_supported(L, d, heads, causal) : attention_tc_supported(L, d, heads),
              "attention_tc: unsupported shape L=%d d=%d heads=%d causal=%d", L, d, heads, causal);
  AtcArgs a{};
  a.B = B; a.L = L; a.d = d; a.heads = heads; a.causal = causal;
  a.lone = L == 257 ? 1 : 0;
  a.pack = pack ? 1 : 0;
  const int main_rows = pack ? 128 : (L < 256 ? L : 256);   // pack: 2 x 64 key slots
  a.n_mtiles = (main_rows + 127) / 128;
  a.n_main = (main_rows + 15) & ~15;
  a.rows_alloc = a.n_mtiles * 128 + kAtcLoneRows * a.lone;
  const int stage_bytes = 3 * a.rows_alloc * 128;
  int stages = (227 * 1024 - 1024 - 256 - kAtcOnesBytes - kAtcLoneBytes) / stage_bytes;
  if (stages > 4) stages = 4;
  EVR_REQUIRE(stages >= 1, "attention_tc: stage of %d bytes does not fit", stage_bytes);
  a.stages = stages;
  a.c1 = 0.125f * 1.4426950408889634f;
  a.out = out;
  a.dbg = nullptr;
  static long long* dbg_buf = nullptr;
  if (getenv("EVR_ATTN_DBG")) {
    if (!dbg_buf) EVR_CHECK_CUDA(cudaMalloc(&dbg_buf, 16 * kDbgJobs * sizeof(long long)));
    EVR_CHECK_CUDA(cudaMemsetAsync(dbg_buf, 0, 16 * kDbgJobs * sizeof(long long), stream));
    a.dbg = dbg_buf;
  }
  const int smem = stages * stage_bytes + kAtcOnesBytes + kAtcLoneBytes + 1024 + 256;
  CUtensorMap t128, t8;
  const uint64_t rows = static_cast<uint64_t>(B) * L;
  if (int r = make_tmap_2d_16b(&t128, qkv, rows, 3ull * d, 3ull * d * 2, 128, true)) return r;
  if (pack) {   // second map: [B][L][3 d], 64-row boxes that stop at the end of their sequence
    if (int r = make_tmap_3d_16b(&t8, qkv, B, L, 3ull * d, 3ull * d * 2, 64, true)) return r;
  } else {
    if (int r = make_tmap_2d_16b(&t8, qkv, rows, 3ull * d, 3ull * d * 2, kAtcLoneRows, true)) return r;
  }
  EVR_REQUIRE(!a.lone || (stages >= 2 && !causal), "attention_tc: L = 257 needs two stages and no causal mask");
  auto kern = pack ? (L == 50 ? attention_tc_kernel<false, 50> : attention_tc_kernel<false, 1>)
                   : (causal ? attention_tc_kernel<true, 0> : attention_tc_kernel<false, 0>);
  static PerDeviceMax configured[4];
  if (configured[pack ? (L == 50 ? 3 : 2) : (causal ? 1 : 0)].needs(smem)) {
    EVR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  const int items = pack ? (B * heads + 1) / 2 : B * heads, sms = sm_count();
  ProfScope ps(PF_ATTN, stream);
  a.prof = prof_device_slot(PF_ATTN);
  EVR_CHECK_CUDA(launch_pdl(kern, dim3(items < sms ? items : sms), dim3(kAtcThreads), static_cast<size_t>(smem), stream,
                            t128, t8, a));
  EVR_CHECK_LAUNCH();
  if (a.dbg) {   // developer timeline: SM clocks of CTA 0, relative to its first stamp
    static long long host[16 * kDbgJobs];
    EVR_CHECK_CUDA(cudaStreamSynchronize(stream));
    EVR_CHECK_CUDA(cudaMemcpy(host, a.dbg, sizeof(host), cudaMemcpyDeviceToHost));
    const long long t0 = host[0];
    const char* names[9] = {"qk_wait", "qk_issued", "p_ready_seen", "pv_issued", "s_full_seen", "pass1_done",
                            "p_written", "o_full_seen", "o_read"};   // pack mode: pass1_done = read-out begins
    for (int i = 0; i < 8; ++i)
      fprintf(stderr, "item %d: empty_wait %6lld tma_issue %6lld mma_full_wait %6lld mma_full_seen %6lld lone_start %6lld lone_end %6lld\n",
              i, host[9 * kDbgJobs + i] - t0, host[10 * kDbgJobs + i] - t0, host[11 * kDbgJobs + i] - t0,
              host[12 * kDbgJobs + i] - t0, host[13 * kDbgJobs + i] - t0, host[14 * kDbgJobs + i] - t0);
    for (int g = 0; g < 12; ++g) {
      fprintf(stderr, "job %2d:", g);
      for (int e = 0; e < 9; ++e) fprintf(stderr, " %s %6lld", names[e], host[e * kDbgJobs + g] ? host[e * kDbgJobs + g] - t0 : -1);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

}  // namespace evr

// Non-GEMM kernels of the CLIP vision tower: im2col(+normalise), cls rows, LayerNorm, attention core,
// final L2-normalise/cast and the one-off weight repacking.  Structure follows openai/CLIP
// clip/model.py VisionTransformer.forward / ResidualAttentionBlock.
#include <cstdlib>
#include <climits>

#include "vit_kernels.cuh"

#include <cmath>

namespace evr {

// ----------------------------------------------------------------------------- im2col
__constant__ float c_clip_mean[3] = {0.48145466f, 0.4578275f, 0.40821073f};
__constant__ float c_clip_std[3] = {0.26862954f, 0.26130258f, 0.27577711f};

// One warp per image line.  uint8 NHWC: the line (b, y) is S*3 contiguous bytes holding all three channels; fp32 NCHW:
// the line (b, c, y) is S contiguous floats.  A line feeds g patches: for every (patch px, channel c) it contributes
// the run out[row(b, py, px)][c*P*P + ky*P + 0..P-1].  Lanes walk the 3S (or S) outputs of the line two at a time, so
// global reads are whole lines and writes are bf16x2 along contiguous runs.  ToTensor + Normalize
// ((v/255 - mean)/std, clip._transform) come from a 3 x 256 table built per block with the same fp32 operations.
template <int P, bool U8>
__global__ void __launch_bounds__(256)
im2col_lines_kernel(const void* __restrict__ pixels, int lines, int S, int Kp, __nv_bfloat16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __nv_bfloat16 lut[3][256];
  __shared__ __align__(16) uint8_t line_buf[8][U8 ? 3 * 224 + 32 : 16];
  if constexpr (U8) {
    for (int i = threadIdx.x; i < 768; i += 256) {
      const int c = i >> 8, v = i & 255;
      lut[c][v] = __float2bfloat16_rn((static_cast<float>(v) / 255.0f - c_clip_mean[c]) / c_clip_std[c]);
    }
    __syncthreads();
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int line = blockIdx.x * 8 + warp;
  if (line >= lines) return;
  const int g = S / P;
  constexpr int PP = P * P;
  if constexpr (U8) {
    const int b = line / S, y = line - b * S;
    const int py = y / P, ky = y - py * P;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(pixels) +
                                                            static_cast<size_t>(line) * S * 3);  // S*3 % 4 == 0
    uint32_t* buf = reinterpret_cast<uint32_t*>(line_buf[warp]);
    for (int i = lane; i < S * 3 / 4; i += 32) buf[i] = __ldg(src + i);
    __syncwarp();
    const uint8_t* lb = line_buf[warp];
    __nv_bfloat16* obase = out + (static_cast<size_t>(b) * g * g + static_cast<size_t>(py) * g) * Kp + ky * P;
    for (int e = 2 * lane; e < 3 * S; e += 64) {       // e = (px*3 + c)*P + kx, kx even
      const int run = e / P, kx = e - run * P;
      const int px = run / 3, c = run - px * 3;
      const int x = px * P + kx;
      const __nv_bfloat162 v = __halves2bfloat162(lut[c][lb[x * 3 + c]], lut[c][lb[x * 3 + 3 + c]]);
      *reinterpret_cast<__nv_bfloat162*>(obase + static_cast<size_t>(px) * Kp + c * PP + kx) = v;
    }
  } else {
    const int b = line / (3 * S), rem = line - b * 3 * S;
    const int c = rem / S, y = rem - c * S;
    const int py = y / P, ky = y - py * P;
    const float2* src = reinterpret_cast<const float2*>(static_cast<const float*>(pixels) + static_cast<size_t>(line) * S);
    __nv_bfloat16* obase = out + (static_cast<size_t>(b) * g * g + static_cast<size_t>(py) * g) * Kp + c * PP + ky * P;
    for (int x = 2 * lane; x < S; x += 64) {
      const float2 t = __ldg(src + (x >> 1));
      const int px = x / P, kx = x - px * P;
      *reinterpret_cast<__nv_bfloat162*>(obase + static_cast<size_t>(px) * Kp + kx) = __floats2bfloat162_rn(t.x, t.y);
    }
  }
}

template <int P>
static int launch_im2col_p(const void* pixels, int pix_dtype, int B, int S, int Kp, __nv_bfloat16* out,
                           cudaStream_t stream) {
  const int lines = pix_dtype == 1 ? B * S : B * 3 * S;
  const unsigned blocks = static_cast<unsigned>((lines + 7) / 8);
  if (pix_dtype == 1)
    EVR_CHECK_CUDA(launch_pdl(im2col_lines_kernel<P, true>, dim3(blocks), dim3(256), 0, stream, pixels, lines, S, Kp, out));
  else
    EVR_CHECK_CUDA(launch_pdl(im2col_lines_kernel<P, false>, dim3(blocks), dim3(256), 0, stream, pixels, lines, S, Kp, out));
  return 0;
}

int launch_im2col(const void* pixels, int pix_dtype, int B, int S, int p, int Kp, __nv_bfloat16* out,
                  cudaStream_t stream) {
  EVR_REQUIRE(S % p == 0 && p % 2 == 0, "im2col: image %d / patch %d unsupported", S, p);
  EVR_REQUIRE(S == 224 && (p == 32 || p == 16 || p == 14), "im2col: geometry %d / %d not instantiated (224 / 32|16|14)", S, p);
  ProfScope ps(PF_VIT_MISC, stream);
  int r = p == 32 ? launch_im2col_p<32>(pixels, pix_dtype, B, S, Kp, out, stream)
          : p == 16 ? launch_im2col_p<16>(pixels, pix_dtype, B, S, Kp, out, stream)
                    : launch_im2col_p<14>(pixels, pix_dtype, B, S, Kp, out, stream);
  if (r) return r;
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- cls rows
__global__ void cls_rows_kernel(float* __restrict__ x, const float* __restrict__ cls, const float* __restrict__ pos,
                                int L, int d) {
  pdl_launch_dependents();
  pdl_wait();
  float* row = x + static_cast<size_t>(blockIdx.x) * L * d;
  for (int i = threadIdx.x; i < d; i += blockDim.x) row[i] = cls[i] + pos[i];
}

int launch_cls_rows(float* x, const float* cls, const float* pos, int B, int L, int d, cudaStream_t stream) {
  ProfScope ps(PF_VIT_MISC, stream);
  EVR_CHECK_CUDA(launch_pdl(cls_rows_kernel, dim3(B), dim3(256), 0, stream, x, cls, pos, L, d));
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- LayerNorm
template <int D4, bool OUT_BF16>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* in, size_t in_stride, void* out, size_t out_stride,  // in may alias out (ln_pre)
                 const float* __restrict__ gamma, const float* __restrict__ beta, int rows, int reverse,
                 unsigned long long* prof) {
  pdl_launch_dependents();
  pdl_wait();
  prof_stamp_begin(prof);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int row = blockIdx.x * 8 + warp;
  if (row >= rows) return;
  // A residual stream larger than L2 was last written front to back (the residual GEMM walks its row blocks in order),
  // so its END is what L2 still holds: walking it back to front turns those lines into hits, front to back every line
  // has been evicted by the time it is reached.
  if (reverse) row = rows - 1 - row;
  constexpr int d = D4 * 128;
  const float4* x4 = reinterpret_cast<const float4*>(in + static_cast<size_t>(row) * in_stride);
  float4 v[D4];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < D4; ++i) {
    v[i] = x4[lane + 32 * i];
    sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum * (1.0f / d);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < D4; ++i) {
    v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
    sq += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq * (1.0f / d) + 1e-5f);
  const float4* g4 = reinterpret_cast<const float4*>(gamma);
  const float4* b4 = reinterpret_cast<const float4*>(beta);
#pragma unroll
  for (int i = 0; i < D4; ++i) {
    const float4 gm = __ldg(g4 + lane + 32 * i), bt = __ldg(b4 + lane + 32 * i);
    float4 y;
    y.x = v[i].x * rstd * gm.x + bt.x;
    y.y = v[i].y * rstd * gm.y + bt.y;
    y.z = v[i].z * rstd * gm.z + bt.z;
    y.w = v[i].w * rstd * gm.w + bt.w;
    if constexpr (OUT_BF16) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(y.x, y.y), hi = __floats2bfloat162_rn(y.z, y.w);
      uint2 w;
      w.x = *reinterpret_cast<uint32_t*>(&lo);
      w.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(out) + static_cast<size_t>(row) * out_stride)[lane + 32 * i] = w;
    } else {
      reinterpret_cast<float4*>(static_cast<float*>(out) + static_cast<size_t>(row) * out_stride)[lane + 32 * i] = y;
    }
  }
  prof_stamp_end(prof);
}

int launch_layernorm(const float* in, size_t in_stride, void* out, size_t out_stride, int out_bf16,
                     const float* gamma, const float* beta, int rows, int d, cudaStream_t stream) {
  EVR_REQUIRE(d == 512 || d == 768 || d == 1024, "layernorm: width %d unsupported (512 | 768 | 1024)", d);
  const int blocks = (rows + 7) / 8;
  ProfScope ps(PF_LN, stream);
  unsigned long long* prof = prof_device_slot(PF_LN);
  static const int reverse_mode = [] { const char* e = getenv("EVR_LN_REVERSE"); return e ? atoi(e) : -1; }();   // -1: auto
  const int reverse = reverse_mode >= 0 ? reverse_mode
                                        : (static_cast<size_t>(rows) * d * 4 > (96u << 20) && in_stride == static_cast<size_t>(d));
  if (d == 512) {
    if (out_bf16) EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<4, true>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
    else EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<4, false>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
  } else if (d == 768) {
    if (out_bf16) EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<6, true>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
    else EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<6, false>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
  } else {
    if (out_bf16) EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<8, true>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
    else EVR_CHECK_CUDA(launch_pdl(layernorm_kernel<8, false>, dim3(blocks), dim3(256), 0, stream, in, in_stride, out, out_stride, gamma, beta, rows, reverse, prof));
  }
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- attention core
// One CTA per (head, image); K and V of the head staged in shared memory (V read with ldmatrix.trans); each warp owns 16-query
// tiles and runs a flash-style loop over 64-key blocks with mma.sync.m16n8k16 (bf16 in, fp32 acc).
// The whole sequence (50 or 257 tokens) is far below one tcgen05 tile, and the core is ~1 % of the
// tower's FLOPs, so it stays on the warp-level MMA path; no attention mask (vision tower).
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ float fast_exp2(float x) {  // ex2.approx: 2 ulp, exp2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int kAttnKStride = 72;  // bf16 elements per K row in smem (64 + 8 pad: conflict-free fragments)

// ---- short sequences (L <= 64: ViT-B/32 has 50 tokens): persistent CTAs, double-buffered cp.async staging ----------
// The one-item-per-CTA kernel above is latency-bound at L = 50 (every CTA pays its global round trips with 16 warps
// per SM resident: ncu long_scoreboard 5.1 of 6.4 stall cycles per issue).  Here a CTA walks a strided list of
// (image, head) items; Q, K, V of item i+1 stream into the second shared-memory stage with cp.async while item i is
// computed, all MMA fragments come from shared memory through ldmatrix, and O leaves through shared memory as whole
// 128-byte rows.  One key block, so the softmax is a single pass.
constexpr int kAttnTileBytes = 64 * kAttnKStride * 2;   // one [64][72] bf16 tile
constexpr int kAttnStageBytes = 3 * kAttnTileBytes;     // Q, K, V

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
               "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem_src) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_src))));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_src) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_src))));
}

// LC: compile-time sequence length (50 = ViT-B/32 at 224 px: the guards below fold away), 0 = use the runtime L.
template <int LC>
__global__ void __launch_bounds__(128)
attention_short_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int L_rt, int d, int heads,
                       int items, float scale_log2e, unsigned long long* prof) {
  pdl_launch_dependents();
  pdl_wait();
  prof_stamp_begin(prof);
  const int L = LC ? LC : L_rt;
  // key groups (8 keys) that hold a key / that are entirely valid: groups past nt_end cost nothing, only
  // [nt_full, nt_end) is masked.  ncu showed the loop bound by issue slots spent on ALU work, not by the tensor pipe.
  const int nt_end = (L + 7) >> 3, nt_full = L >> 3;
  extern __shared__ __align__(16) uint8_t attn_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ld = 3 * d;
  // rows >= L of both stages are never written by the loader: zero them once
  for (int i = threadIdx.x; i < 2 * 3 * (64 - L) * 9; i += 128) {
    const int tile = i / ((64 - L) * 9), rem = i % ((64 - L) * 9);
    *reinterpret_cast<uint4*>(attn_smem + tile * kAttnTileBytes + (L + rem / 9) * (kAttnKStride * 2) + (rem % 9) * 16) =
        make_uint4(0, 0, 0, 0);
  }
  auto prefetch = [&](int item, int stage) {
    const int b = item / heads, h = item - b * heads;
    const __nv_bfloat16* src = qkv + static_cast<size_t>(b) * L * ld + h * 64;
    uint8_t* dst = attn_smem + stage * kAttnStageBytes;
    // 16-B chunks: thread -> (row = tid / 8 + 16 j, chunk = tid % 8) of each of the three tiles; no index division
    const int c = threadIdx.x & 7;
    const __nv_bfloat16* s0 = src + c * 8;
    uint8_t* d0 = dst + c * 16;
    for (int r = threadIdx.x >> 3; r < L; r += 16) {
      const __nv_bfloat16* sr = s0 + static_cast<size_t>(r) * ld;
      uint8_t* dr = d0 + r * (kAttnKStride * 2);
      cp_async16(dr, sr);
      cp_async16(dr + kAttnTileBytes, sr + d);
      cp_async16(dr + 2 * kAttnTileBytes, sr + 2 * d);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  int item = blockIdx.x, stage = 0;
  if (item < items) prefetch(item, 0);
  for (; item < items; item += gridDim.x, stage ^= 1) {
    const int next = item + gridDim.x;
    if (next < items) {
      prefetch(next, stage ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(attn_smem + stage * kAttnStageBytes);
    const __nv_bfloat16* Ks = Qs + 64 * kAttnKStride;
    const __nv_bfloat16* Vs = Ks + 64 * kAttnKStride;
    const int m0 = warp * 16;  // this warp's 16 query rows
    if (m0 < L) {
      uint32_t qa[4][4];
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        ldmatrix_x4(qa[kk], Qs + (m0 + (lane & 7) + ((lane >> 3) & 1) * 8) * kAttnKStride + 16 * kk + (lane >> 4) * 8);
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; nt += 2) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        s[nt + 1][0] = s[nt + 1][1] = s[nt + 1][2] = s[nt + 1][3] = 0.f;
        if (nt < nt_end) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            uint32_t kb[4];  // K[key 8nt + g][dims 16kk + {0,8} + 2t..] for key tiles nt, nt+1
            ldmatrix_x4(kb, Ks + ((nt + (lane >> 4)) * 8 + (lane & 7)) * kAttnKStride + 16 * kk + ((lane >> 3) & 1) * 8);
            mma_bf16_16816(s[nt], qa[kk], kb[0], kb[1]);
            if (nt + 1 < nt_end) mma_bf16_16816(s[nt + 1], qa[kk], kb[2], kb[3]);
          }
        }
      }
      // raw scores: the positive scale commutes with the maximum and is folded into the exponent's FFMA below
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        if (nt >= nt_full) {
          const int col = nt * 8 + 2 * t;
          s[nt][0] = col < L ? s[nt][0] : -INFINITY;
          s[nt][1] = col + 1 < L ? s[nt][1] : -INFINITY;
          s[nt][2] = col < L ? s[nt][2] : -INFINITY;
          s[nt][3] = col + 1 < L ? s[nt][3] : -INFINITY;
        }
        mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
        mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      float l0 = 0.f, l1 = 0.f;
      const float ms0 = -mx0 * scale_log2e, ms1 = -mx1 * scale_log2e;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        if (nt < nt_end) {
          s[nt][0] = fast_exp2(fmaf(s[nt][0], scale_log2e, ms0));
          s[nt][1] = fast_exp2(fmaf(s[nt][1], scale_log2e, ms0));
          s[nt][2] = fast_exp2(fmaf(s[nt][2], scale_log2e, ms1));
          s[nt][3] = fast_exp2(fmaf(s[nt][3], scale_log2e, ms1));
          l0 += s[nt][0] + s[nt][1];
          l1 += s[nt][2] + s[nt][3];
        } else {
          s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        }
      }
      float o[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) { o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f; }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        if (2 * kk >= nt_end) break;  // 16-key groups past the last key carry p = 0
        uint32_t pa[4];
        pa[0] = pack2_bf16(s[2 * kk][0], s[2 * kk][1]);
        pa[1] = pack2_bf16(s[2 * kk][2], s[2 * kk][3]);
        pa[2] = pack2_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
        pa[3] = pack2_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
        for (int nt = 0; nt < 8; nt += 2) {
          uint32_t vb[4];
          ldmatrix_x4_trans(vb, Vs + (16 * kk + ((lane >> 3) & 1) * 8 + (lane & 7)) * kAttnKStride + (nt + (lane >> 4)) * 8);
          mma_bf16_16816(o[nt], pa, vb[0], vb[1]);
          mma_bf16_16816(o[nt + 1], pa, vb[2], vb[3]);
        }
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      const float i0 = 1.0f / l0, i1 = 1.0f / l1;
      // O through this warp's own 16 rows of the Q tile (its Q fragments are in registers), then whole rows to global
      __syncwarp();
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        *reinterpret_cast<uint32_t*>(Qs + (m0 + g) * kAttnKStride + nt * 8 + 2 * t) = pack2_bf16(o[nt][0] * i0, o[nt][1] * i0);
        *reinterpret_cast<uint32_t*>(Qs + (m0 + g + 8) * kAttnKStride + nt * 8 + 2 * t) = pack2_bf16(o[nt][2] * i1, o[nt][3] * i1);
      }
      __syncwarp();
      const int b = item / heads, h = item - b * heads;
      __nv_bfloat16* obase = out + static_cast<size_t>(b) * L * d + h * 64;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int u = i * 32 + lane, r = m0 + (u >> 3), c = u & 7;
        if (r < L)
          *reinterpret_cast<uint4*>(obase + static_cast<size_t>(r) * d + c * 8) =
              *reinterpret_cast<const uint4*>(Qs + r * kAttnKStride + c * 8);
      }
    }
    __syncthreads();  // this stage is refilled by the prefetch issued at the top of the next iteration
  }
  prof_stamp_end(prof);
}

// One 64-key block of the flash loop for a warp's 16-query tile.  Scores stay in raw (unscaled) units: the row maximum
// commutes with the positive scale, which is folded into the exponent's FFMA (p = 2^(s*c - m*c)); the running output
// is rescaled only when some row's maximum moved (warp vote).  ncu on L = 257 showed the loop issue-bound on ALU work
// (tensor pipe 37 %, issue slots 52 %, ~350 non-MMA instructions per 64 MMAs), hence the lean full-block path.
// PARTIAL: the block holds masked keys (sequence tail, causal diagonal): groups >= nt_end are skipped entirely.
template <bool PARTIAL>
__device__ __forceinline__ void attn_block(const __nv_bfloat16* Ks, const __nv_bfloat16* Vs, int kb, int lane, int t,
                                           const uint32_t (&qa)[4][4], float (&o)[8][4], float& mx0, float& mx1,
                                           float& l0, float& l1, float c, int lim0, int lim1, int nt_end) {
  float s[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; nt += 2) {
    s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
    s[nt + 1][0] = s[nt + 1][1] = s[nt + 1][2] = s[nt + 1][3] = 0.f;
    if (!PARTIAL || nt < nt_end) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t kf[4];
        ldmatrix_x4(kf, Ks + (kb * 64 + (nt + (lane >> 4)) * 8 + (lane & 7)) * kAttnKStride + 16 * kk +
                            ((lane >> 3) & 1) * 8);
        mma_bf16_16816(s[nt], qa[kk], kf[0], kf[1]);
        mma_bf16_16816(s[nt + 1], qa[kk], kf[2], kf[3]);
      }
    }
  }
  float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    if (PARTIAL) {
      const int col = kb * 64 + nt * 8 + 2 * t;
      s[nt][0] = col < lim0 ? s[nt][0] : -INFINITY;
      s[nt][1] = col + 1 < lim0 ? s[nt][1] : -INFINITY;
      s[nt][2] = col < lim1 ? s[nt][2] : -INFINITY;
      s[nt][3] = col + 1 < lim1 ? s[nt][3] : -INFINITY;
    }
    bm0 = fmaxf(bm0, fmaxf(s[nt][0], s[nt][1]));
    bm1 = fmaxf(bm1, fmaxf(s[nt][2], s[nt][3]));
  }
  bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1));
  bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
  bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1));
  bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
  const float mn0 = fmaxf(mx0, bm0), mn1 = fmaxf(mx1, bm1);  // finite: key block 0 always holds key 0
  if (__any_sync(0xffffffffu, mn0 > mx0 || mn1 > mx1)) {
    const float a0 = fast_exp2((mx0 - mn0) * c), a1 = fast_exp2((mx1 - mn1) * c);
    l0 *= a0; l1 *= a1;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= a0; o[nt][1] *= a0; o[nt][2] *= a1; o[nt][3] *= a1; }
    mx0 = mn0; mx1 = mn1;
  }
  const float ms0 = -mn0 * c, ms1 = -mn1 * c;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    s[nt][0] = fast_exp2(fmaf(s[nt][0], c, ms0));
    s[nt][1] = fast_exp2(fmaf(s[nt][1], c, ms0));
    s[nt][2] = fast_exp2(fmaf(s[nt][2], c, ms1));
    s[nt][3] = fast_exp2(fmaf(s[nt][3], c, ms1));
    l0 += s[nt][0] + s[nt][1];
    l1 += s[nt][2] + s[nt][3];
  }
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    if (PARTIAL && 2 * kk >= nt_end) break;  // 16-key groups past the last visible key carry p = 0
    uint32_t pa[4];
    pa[0] = pack2_bf16(s[2 * kk][0], s[2 * kk][1]);
    pa[1] = pack2_bf16(s[2 * kk][2], s[2 * kk][3]);
    pa[2] = pack2_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
    pa[3] = pack2_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
    for (int nt = 0; nt < 8; nt += 2) {
      uint32_t vb[4];
      ldmatrix_x4_trans(vb, Vs + (kb * 64 + 16 * kk + ((lane >> 3) & 1) * 8 + (lane & 7)) * kAttnKStride +
                                (nt + (lane >> 4)) * 8);
      mma_bf16_16816(o[nt], pa, vb[0], vb[1]);
      mma_bf16_16816(o[nt + 1], pa, vb[2], vb[3]);
    }
  }
}

// ---- long sequences (L > 64: 197 tokens ViT-B/16, 257 ViT-L/14, 77 text tower) --------------------------------------
// One CTA (8 warps) per (head, sequence): K and V of the head are staged once with cp.async ([Lp][72] bf16 each); every
// warp walks 16-query tiles: the Q tile goes through a per-warp [16][72] staging buffer (coalesced 16-B loads ->
// ldmatrix A fragments), the flash loop runs over 64-key blocks with K fragments from ldmatrix and V from
// ldmatrix.trans, the softmax works in the exp2 domain, and O leaves through the same per-warp buffer as whole rows.
// `causal`: row r attends keys <= r (clip/model.py build_attention_mask); key blocks entirely above the diagonal of a
// query tile are skipped.
constexpr int kAttnLongWarps = 8;

__global__ void __launch_bounds__(kAttnLongWarps * 32, 2)
attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int L, int d,
                 float scale_log2e, int causal, unsigned long long* prof) {
  pdl_launch_dependents();
  pdl_wait();
  prof_stamp_begin(prof);
  extern __shared__ __align__(16) uint8_t attn_smem[];
  const int h = blockIdx.x, b = blockIdx.y;
  const int Lp = (L + 63) & ~63;
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(attn_smem);  // [Lp][72]
  __nv_bfloat16* Vs = Ks + Lp * kAttnKStride;                       // [Lp][72]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  __nv_bfloat16* Qw = Vs + Lp * kAttnKStride + warp * 16 * kAttnKStride;  // this warp's [16][72] Q / O staging tile
  const size_t row0 = static_cast<size_t>(b) * L;
  const int ld = 3 * d;
  {
    const int c = threadIdx.x & 7;
    const __nv_bfloat16* src = qkv + row0 * ld + h * 64 + c * 8;
    for (int r = threadIdx.x >> 3; r < Lp; r += kAttnLongWarps * 4) {
      __nv_bfloat16* kd = Ks + r * kAttnKStride + c * 8;
      __nv_bfloat16* vd = Vs + r * kAttnKStride + c * 8;
      if (r < L) {
        cp_async16(kd, src + static_cast<size_t>(r) * ld + d);
        cp_async16(vd, src + static_cast<size_t>(r) * ld + 2 * d);
      } else {
        *reinterpret_cast<uint4*>(kd) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(vd) = make_uint4(0, 0, 0, 0);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncthreads();

  // L = 16 n + 1 (257 tokens, ViT-L/14): the 17th query tile would hold one row and cost one warp a third full round
  // while the other seven idle; that row is computed by the whole CTA on the FMA pipes after the tile loop instead.
  const bool lone_row = !causal && (L & 15) == 1 && L > 16;
  const int num_mt = lone_row ? (L >> 4) : ((L + 15) >> 4);
  for (int mt = warp; mt < num_mt; mt += kAttnLongWarps) {
    const int m0 = mt * 16;
    // ---- Q tile -> staging -> A fragments
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = i * 32 + lane, r = u >> 3, c = u & 7;
      uint4 q4 = make_uint4(0, 0, 0, 0);
      if (m0 + r < L) q4 = *reinterpret_cast<const uint4*>(qkv + (row0 + m0 + r) * ld + h * 64 + c * 8);
      *reinterpret_cast<uint4*>(Qw + r * kAttnKStride + c * 8) = q4;
    }
    __syncwarp();
    uint32_t qa[4][4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
      ldmatrix_x4(qa[kk], Qw + ((lane & 7) + ((lane >> 3) & 1) * 8) * kAttnKStride + 16 * kk + (lane >> 4) * 8);
    const int r0 = m0 + g, r1 = r0 + 8;
    const int lim0 = causal ? min(L, r0 + 1) : L, lim1 = causal ? min(L, r1 + 1) : L;  // keys visible to each row
    float o[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f; }
    float mx0 = -INFINITY, mx1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const int kb_end = causal ? min(Lp / 64, (m0 + 15) / 64 + 1) : Lp / 64;
    for (int kb = 0; kb < kb_end; ++kb) {
      // 8-key groups of this block with a key visible to the tile (warp-uniform): the last block of L = 257 holds one
      // key, of L = 197 five.  Full blocks take the lean path: no masking, no guards.
      const int vis = (causal ? min(L, m0 + 16) : L) - kb * 64;
      const bool partial = vis < 64 || (causal && kb * 64 + 63 > m0);
      if (partial)
        attn_block<true>(Ks, Vs, kb, lane, t, qa, o, mx0, mx1, l0, l1, scale_log2e, lim0, lim1, min(8, (vis + 7) >> 3));
      else
        attn_block<false>(Ks, Vs, kb, lane, t, qa, o, mx0, mx1, l0, l1, scale_log2e, lim0, lim1, 8);
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    __syncwarp();  // every lane's ldmatrix of the Q tile is done: the buffer now carries O
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      *reinterpret_cast<uint32_t*>(Qw + g * kAttnKStride + nt * 8 + 2 * t) = pack2_bf16(o[nt][0] * i0, o[nt][1] * i0);
      *reinterpret_cast<uint32_t*>(Qw + (g + 8) * kAttnKStride + nt * 8 + 2 * t) = pack2_bf16(o[nt][2] * i1, o[nt][3] * i1);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = i * 32 + lane, r = u >> 3, c = u & 7;
      if (m0 + r < L)
        *reinterpret_cast<uint4*>(out + (row0 + m0 + r) * d + h * 64 + c * 8) =
            *reinterpret_cast<const uint4*>(Qw + r * kAttnKStride + c * 8);
    }
    __syncwarp();  // the staging tile is rewritten by the next query tile
  }
  if (lone_row) {
    // ---- last query row, all 256 threads.  Scratch: the per-warp Q/O tiles (free after the barrier).
    __syncthreads();
    float* sc = reinterpret_cast<float*>(Vs + Lp * kAttnKStride);  // [Lp] scores -> probabilities
    float* red = sc + Lp;                                          // [8] warp maxima, [8] warp sums
    float* part = red + 16;                                        // [8][64] partial outputs
    const int r = L - 1;
    uint4 q4[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) q4[c] = *reinterpret_cast<const uint4*>(qkv + (row0 + r) * ld + h * 64 + c * 8);
    float tmax = -INFINITY;
    for (int j = threadIdx.x; j < L; j += kAttnLongWarps * 32) {
      float acc = 0.f;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint4 k4 = *reinterpret_cast<const uint4*>(Ks + j * kAttnKStride + c * 8);
        const __nv_bfloat162* qp = reinterpret_cast<const __nv_bfloat162*>(&q4[c]);
        const __nv_bfloat162* kp = reinterpret_cast<const __nv_bfloat162*>(&k4);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 a = __bfloat1622float2(qp[e]), bq = __bfloat1622float2(kp[e]);
          acc = fmaf(a.x, bq.x, acc);
          acc = fmaf(a.y, bq.y, acc);
        }
      }
      acc *= scale_log2e;
      sc[j] = acc;
      tmax = fmaxf(tmax, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
    if (lane == 0) red[warp] = tmax;
    __syncthreads();
    float rmax = red[0];
#pragma unroll
    for (int w = 1; w < kAttnLongWarps; ++w) rmax = fmaxf(rmax, red[w]);
    // warp w owns keys w, w + 8, ...; lane owns output dims 2 lane, 2 lane + 1
    float ox = 0.f, oy = 0.f, lsum = 0.f;
    for (int j = warp; j < L; j += kAttnLongWarps) {
      const float pj = fast_exp2(sc[j] - rmax);
      const float2 v2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(Vs + j * kAttnKStride + 2 * lane));
      ox = fmaf(pj, v2.x, ox);
      oy = fmaf(pj, v2.y, oy);
      lsum += pj;
    }
    part[warp * 64 + 2 * lane] = ox;
    part[warp * 64 + 2 * lane + 1] = oy;
    if (lane == 0) red[8 + warp] = lsum;
    __syncthreads();
    if (warp == 0) {
      float tx = 0.f, ty = 0.f, lt = 0.f;
#pragma unroll
      for (int w = 0; w < kAttnLongWarps; ++w) {
        tx += part[w * 64 + 2 * lane];
        ty += part[w * 64 + 2 * lane + 1];
        lt += red[8 + w];
      }
      const float inv = 1.0f / lt;
      *reinterpret_cast<uint32_t*>(out + (row0 + r) * d + h * 64 + 2 * lane) = pack2_bf16(tx * inv, ty * inv);
    }
  }
  prof_stamp_end(prof);
}

static bool attention_tc_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("EVR_ATTN_MMA_SYNC");  // A/B knob: keep the warp-level MMA core for every length
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}

int launch_attention(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                     cudaStream_t stream) {
  return launch_attention_impl(qkv, out, B, L, d, heads, causal, 0, stream);
}

int launch_attention_impl(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                          int impl, cudaStream_t stream) {
  EVR_REQUIRE(heads * 64 == d, "attention: head dim must be 64 (d=%d heads=%d)", d, heads);
  if (impl == 2 || (impl == 0 && attention_tc_enabled() &&
                    (attention_tc_supported(L, d, heads) || attention_tc_pack_supported(L, d, heads, causal))))
    return launch_attention_tc(qkv, out, B, L, d, heads, causal, stream);
  if (L <= 64 && !causal) {
    static PerDeviceMax configured_short;
    if (configured_short.needs(2 * kAttnStageBytes)) {
      EVR_CHECK_CUDA(cudaFuncSetAttribute(attention_short_kernel<50>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          2 * kAttnStageBytes));
      EVR_CHECK_CUDA(cudaFuncSetAttribute(attention_short_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          2 * kAttnStageBytes));
    }
    // 4 CTAs (2 x 27 KB stages each) per SM, all resident; item i goes to CTA i mod grid, so with grid = 4 x SMs every SM
    // walks items / SMs (+-1) of them (3072 items: 20-21 per SM; sizing the grid for equal items per CTA instead left
    // 512 CTAs of 6 items = 24 on the SMs that got four CTAs and 18 on the others)
    const int items = B * heads, slots = sm_count() * 4;
    const int grid = items < slots ? items : slots;
    ProfScope ps(PF_ATTN, stream);
    EVR_CHECK_CUDA(launch_pdl(L == 50 ? attention_short_kernel<50> : attention_short_kernel<0>, dim3(grid), dim3(128),
                              2 * kAttnStageBytes, stream, qkv, out, L, d, heads, items,
                              0.125f * 1.4426950408889634f, prof_device_slot(PF_ATTN)));
    EVR_CHECK_LAUNCH();
    return 0;
  }
  const int Lp = (L + 63) & ~63;
  const size_t smem = (2 * static_cast<size_t>(Lp) + kAttnLongWarps * 16) * kAttnKStride * 2;  // K, V, per-warp Q/O tiles
  EVR_REQUIRE(smem <= 200 * 1024, "attention: sequence %d too long for the single-CTA K/V staging", L);
  static PerDeviceMax configured;
  if (smem > 48 * 1024 && configured.needs(static_cast<int>(smem))) {
    EVR_CHECK_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        static_cast<int>(smem)));
  }
  ProfScope ps(PF_ATTN, stream);
  EVR_CHECK_CUDA(launch_pdl(attention_kernel, dim3(heads, B), dim3(kAttnLongWarps * 32), smem, stream, qkv, out, L, d,
                            0.125f * 1.4426950408889634f, causal, prof_device_slot(PF_ATTN)));
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- text tower helpers
// x[b*L + l, :] = token_embedding[tokens[b, l], :] + positional_embedding[l, :]   (clip/model.py CLIP.encode_text)
__global__ void __launch_bounds__(128)
embed_tokens_kernel(const int32_t* __restrict__ tokens, const float* __restrict__ table, const float* __restrict__ pos,
                    float* __restrict__ x, int rows, int L, int d, int vocab) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x;
  if (row >= rows) return;
  int tok = tokens[row];
  tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);  // ids are validated on the host; never read out of bounds
  const float4* e = reinterpret_cast<const float4*>(table + static_cast<size_t>(tok) * d);
  const float4* p = reinterpret_cast<const float4*>(pos + static_cast<size_t>(row % L) * d);
  float4* o = reinterpret_cast<float4*>(x + static_cast<size_t>(row) * d);
  for (int i = threadIdx.x; i < d / 4; i += blockDim.x) {
    const float4 a = __ldg(e + i), b = __ldg(p + i);
    o[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
}

int launch_embed_tokens(const int32_t* tokens, const float* table, const float* pos, float* x, int B, int L, int d,
                        int vocab, cudaStream_t stream) {
  ProfScope ps(PF_VIT_MISC, stream);
  EVR_CHECK_CUDA(launch_pdl(embed_tokens_kernel, dim3(B * L), dim3(128), 0, stream, tokens, table, pos, x, B * L, L, d,
                            vocab));
  EVR_CHECK_LAUNCH();
  return 0;
}

// out[b, :] = x[b*L + argmax_l tokens[b, l], :]  -- the EOT token carries the largest id (encode_text:
// x[arange(B), text.argmax(dim=-1)]); first occurrence on ties, as torch.argmax.
__global__ void __launch_bounds__(128)
gather_eot_kernel(const int32_t* __restrict__ tokens, const float* __restrict__ x, float* __restrict__ out, int L,
                  int d) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ int best_pos;
  const int b = blockIdx.x;
  if (threadIdx.x < 32) {
    int bv = INT_MIN, bp = 0;
    for (int l = threadIdx.x; l < L; l += 32) {
      const int v = tokens[b * L + l];
      if (v > bv) { bv = v; bp = l; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const int ov = __shfl_xor_sync(0xffffffffu, bv, o), op = __shfl_xor_sync(0xffffffffu, bp, o);
      if (ov > bv || (ov == bv && op < bp)) { bv = ov; bp = op; }
    }
    if (threadIdx.x == 0) best_pos = bp;
  }
  __syncthreads();
  const float4* src = reinterpret_cast<const float4*>(x + (static_cast<size_t>(b) * L + best_pos) * d);
  float4* dst = reinterpret_cast<float4*>(out + static_cast<size_t>(b) * d);
  for (int i = threadIdx.x; i < d / 4; i += blockDim.x) dst[i] = src[i];
}

int launch_gather_eot(const int32_t* tokens, const float* x, float* out, int B, int L, int d, cudaStream_t stream) {
  ProfScope ps(PF_VIT_MISC, stream);
  EVR_CHECK_CUDA(launch_pdl(gather_eot_kernel, dim3(B), dim3(128), 0, stream, tokens, x, out, L, d));
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- final normalise / cast
__global__ void __launch_bounds__(256)
l2norm_cast_kernel(const float* __restrict__ in, void* __restrict__ out, int out_dtype, int rows, int D,
                   int normalise) {
  pdl_launch_dependents();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= rows) return;
  const float* x = in + static_cast<size_t>(row) * D;
  float inv = 1.0f;
  if (normalise) {
    float sq = 0.f;
    for (int i = lane; i < D; i += 32) sq += x[i] * x[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    inv = 1.0f / sqrtf(sq);
  }
  for (int i = lane; i < D; i += 32) {
    const float y = normalise ? x[i] * inv : x[i];
    if (out_dtype == 0) static_cast<float*>(out)[static_cast<size_t>(row) * D + i] = y;
    else static_cast<__half*>(out)[static_cast<size_t>(row) * D + i] = __float2half_rn(y);
  }
}

int launch_l2norm_cast(const float* in, void* out, int out_dtype, int rows, int D, int normalise,
                       cudaStream_t stream) {
  ProfScope ps(PF_VIT_MISC, stream);
  EVR_CHECK_CUDA(launch_pdl(l2norm_cast_kernel, dim3((rows + 7) / 8), dim3(256), 0, stream, in, out, out_dtype, rows, D, normalise));
  EVR_CHECK_LAUNCH();
  return 0;
}

// ----------------------------------------------------------------------------- weight repacking
__global__ void cast_pad_bf16_kernel(const float* __restrict__ in, int rows, int cols, __nv_bfloat16* __restrict__ out,
                                     int out_cols) {
  const size_t total = static_cast<size_t>(rows) * out_cols;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / out_cols;
    const int c = static_cast<int>(i - r * out_cols);
    out[i] = __float2bfloat16_rn(c < cols ? in[r * cols + c] : 0.f);
  }
}
int launch_cast_pad_bf16(const float* in, int rows, int cols, __nv_bfloat16* out, int out_cols, cudaStream_t stream) {
  cast_pad_bf16_kernel<<<1024, 256, 0, stream>>>(in, rows, cols, out, out_cols);
  EVR_CHECK_LAUNCH();
  return 0;
}

// in [rows, cols] fp32 -> out [cols, rows] bf16
__global__ void transpose_cast_bf16_kernel(const float* __restrict__ in, int rows, int cols,
                                           __nv_bfloat16* __restrict__ out) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    tile[j][threadIdx.x] = (r < rows && c < cols) ? in[static_cast<size_t>(r) * cols + c] : 0.f;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) out[static_cast<size_t>(c) * rows + r] = __float2bfloat16_rn(tile[threadIdx.x][j]);
  }
}
int launch_transpose_cast_bf16(const float* in, int rows, int cols, __nv_bfloat16* out, cudaStream_t stream) {
  dim3 grid((cols + 31) / 32, (rows + 31) / 32);
  transpose_cast_bf16_kernel<<<grid, dim3(32, 8), 0, stream>>>(in, rows, cols, out);
  EVR_CHECK_LAUNCH();
  return 0;
}

}  // namespace evr

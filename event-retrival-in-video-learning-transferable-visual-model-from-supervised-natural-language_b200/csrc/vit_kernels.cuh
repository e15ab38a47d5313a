// Non-GEMM kernels of the CLIP vision tower (declarations; definitions in vit_kernels.cu).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace evr {

// pixels -> bf16 im2col matrix [B*g*g, Kp]; col = c*p*p + ky*p + kx (conv1.weight flattening).
// pix_dtype 0: fp32 NCHW already normalised; 1: uint8 NHWC, (x/255 - mean)/std fused (clip._transform).
int launch_im2col(const void* pixels, int pix_dtype, int B, int S, int p, int Kp, __nv_bfloat16* out,
                  cudaStream_t stream);

// x[b*L, :] = class_embedding + positional_embedding[0]
int launch_cls_rows(float* x, const float* cls, const float* pos, int B, int L, int d, cudaStream_t stream);

// LayerNorm over the last dim (fp32 statistics, eps 1e-5, clip/model.py LayerNorm).
// in: fp32 rows at in + r*in_stride; out: bf16 (out_bf16 != 0) or fp32 rows at out + r*out_stride.
int launch_layernorm(const float* in, size_t in_stride, void* out, size_t out_stride, int out_bf16,
                     const float* gamma, const float* beta, int rows, int d, cudaStream_t stream);

// softmax(q k^T / sqrt(hd) [+ causal mask]) v per (sequence, head); qkv bf16 [B*L, 3d] (q | k | v), out bf16 [B*L, d]. hd = 64.
int launch_attention(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                     cudaStream_t stream);
// the same on tcgen05 / TMEM (attention_tc.cu), 64 < L <= 257; launch_attention dispatches to it
bool attention_tc_supported(int L, int d, int heads);
bool attention_tc_pack_supported(int L, int d, int heads, int causal);  // L <= 64: two items per 128-row tile
int launch_attention_tc(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                        cudaStream_t stream);
// impl 0: dispatch (tcgen05 where supported), 1: warp-level MMA kernels, 2: tcgen05 or an error
int launch_attention_impl(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int d, int heads, int causal,
                          int impl, cudaStream_t stream);

// text tower: x = token_embedding[tokens] + positional_embedding ; EOT-row gather (argmax of the token ids)
int launch_embed_tokens(const int32_t* tokens, const float* table, const float* pos, float* x, int B, int L, int d,
                        int vocab, cudaStream_t stream);
int launch_gather_eot(const int32_t* tokens, const float* x, float* out, int B, int L, int d, cudaStream_t stream);

// out[r] = in[r] (/ ||in[r]||) ; out fp32 (out_dtype 0) or fp16 (1)
int launch_l2norm_cast(const float* in, void* out, int out_dtype, int rows, int D, int normalise,
                       cudaStream_t stream);

// weight repacking (once, at handle creation)
int launch_cast_pad_bf16(const float* in, int rows, int cols, __nv_bfloat16* out, int out_cols, cudaStream_t stream);
int launch_transpose_cast_bf16(const float* in, int rows, int cols, __nv_bfloat16* out, cudaStream_t stream);

}  // namespace evr

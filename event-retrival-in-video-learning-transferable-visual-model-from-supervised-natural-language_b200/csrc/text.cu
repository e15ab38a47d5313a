// CLIP text tower handle: encode_text of openai/CLIP (clip/model.py CLIP.encode_text: token embedding + positional
// embedding -> 12 pre-LN residual blocks under a causal mask -> ln_final -> EOT row -> @ text_projection).
// Replaces  model.encode_text(clip.tokenize([query]))  at Backend/services/embedding_service.py:169-177
// (SURVEY.md 8a row a8 / 8f row 1).  Same kernels as the vision tower: tcgen05 GEMMs with fused epilogues,
// LayerNorm, attention core (with the causal mask), fp32 residual stream.
#include <vector>

#include "../../include/evr.h"
#include "gemm_tc.cuh"
#include "vit_kernels.cuh"

using namespace evr;

namespace {
struct TextLayerW {
  const float *ln1_g, *ln1_b, *b_in, *b_out, *ln2_g, *ln2_b, *b_fc, *b_proj;  // fp32 params
  const __nv_bfloat16 *w_in, *w_out, *w_fc, *w_proj;                          // bf16 [N, K]
};
}  // namespace

struct evr_text {
  evr_text_cfg cfg;
  int max_batch;
  float* fparams = nullptr;          // token table, positional embedding, LayerNorm / bias vectors
  __nv_bfloat16* wparams = nullptr;  // GEMM weights
  const float *tok, *pos, *lnf_g, *lnf_b;
  const __nv_bfloat16* w_projT;      // text_projection^T [out, d]
  std::vector<TextLayerW> layers;
  __nv_bfloat16 *hbuf = nullptr, *qkv = nullptr, *attn = nullptr, *fbuf = nullptr, *eot_ln = nullptr;
  float *x = nullptr, *eot = nullptr, *emb = nullptr;
  int32_t* tokens = nullptr;         // [max_batch, ctx] copy of the caller's ids for the captured small-batch chain
  GraphCache graphs;
};

static int text_copy_f32(float*& cursor, const void* src, size_t n, const float** dst, cudaStream_t s) {
  EVR_CHECK_CUDA(cudaMemcpyAsync(cursor, src, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
  *dst = cursor;
  cursor += (n + 3) & ~size_t(3);
  return 0;
}

extern "C" int evr_text_create(const evr_text_cfg* cfg, const void* const* w, int n_weights, int max_batch,
                               evr_text** out) {
  EVR_REQUIRE(cfg && w && out, "evr_text_create: null argument");
  EVR_REQUIRE(cfg->width == 512 || cfg->width == 768 || cfg->width == 1024, "text width %d unsupported", cfg->width);
  EVR_REQUIRE(cfg->heads * 64 == cfg->width, "heads*64 must equal width");
  EVR_REQUIRE(cfg->out_dim % 32 == 0, "out_dim must be a multiple of 32");
  EVR_REQUIRE(cfg->context_length >= 1 && cfg->context_length <= 1024, "context_length %d", cfg->context_length);
  EVR_REQUIRE(cfg->vocab_size >= 1, "vocab_size %d", cfg->vocab_size);
  EVR_REQUIRE(n_weights == 2 + 12 * cfg->layers + 3, "expected %d weight pointers, got %d", 2 + 12 * cfg->layers + 3,
              n_weights);
  EVR_REQUIRE(max_batch >= 1, "max_batch must be >= 1");
  evr_text* h = new evr_text();
  h->cfg = *cfg;
  h->max_batch = max_batch;
  const int d = cfg->width, nl = cfg->layers, od = cfg->out_dim, L = cfg->context_length;
  cudaStream_t s = 0;
  const size_t nf = static_cast<size_t>(cfg->vocab_size) * d + static_cast<size_t>(L) * d + 2 * d + 64 +
                    static_cast<size_t>(nl) * (4 * d + 3 * d + d + 4 * d + d + 64);
  EVR_CHECK_CUDA(cudaMalloc(&h->fparams, nf * sizeof(float)));
  float* fc = h->fparams;
  if (text_copy_f32(fc, w[0], static_cast<size_t>(cfg->vocab_size) * d, &h->tok, s)) return 1;
  if (text_copy_f32(fc, w[1], static_cast<size_t>(L) * d, &h->pos, s)) return 1;
  const int tail = 2 + 12 * nl;
  if (text_copy_f32(fc, w[tail], d, &h->lnf_g, s)) return 1;
  if (text_copy_f32(fc, w[tail + 1], d, &h->lnf_b, s)) return 1;
  const size_t nw = static_cast<size_t>(od) * d + static_cast<size_t>(nl) * 12ull * d * d;
  EVR_CHECK_CUDA(cudaMalloc(&h->wparams, nw * sizeof(__nv_bfloat16)));
  __nv_bfloat16* wc = h->wparams;
  h->w_projT = wc;  // text_projection [d, out] -> [out, d]
  if (launch_transpose_cast_bf16(static_cast<const float*>(w[tail + 2]), d, od, wc, s)) return 1;
  wc += static_cast<size_t>(od) * d;
  h->layers.resize(nl);
  for (int i = 0; i < nl; ++i) {
    const void* const* lw = w + 2 + 12 * i;
    TextLayerW& l = h->layers[i];
    if (text_copy_f32(fc, lw[0], d, &l.ln1_g, s)) return 1;
    if (text_copy_f32(fc, lw[1], d, &l.ln1_b, s)) return 1;
    if (text_copy_f32(fc, lw[3], 3 * d, &l.b_in, s)) return 1;
    if (text_copy_f32(fc, lw[5], d, &l.b_out, s)) return 1;
    if (text_copy_f32(fc, lw[6], d, &l.ln2_g, s)) return 1;
    if (text_copy_f32(fc, lw[7], d, &l.ln2_b, s)) return 1;
    if (text_copy_f32(fc, lw[9], 4 * d, &l.b_fc, s)) return 1;
    if (text_copy_f32(fc, lw[11], d, &l.b_proj, s)) return 1;
    l.w_in = wc;
    if (launch_cast_pad_bf16(static_cast<const float*>(lw[2]), 3 * d, d, wc, d, s)) return 1;
    wc += 3ull * d * d;
    l.w_out = wc;
    if (launch_cast_pad_bf16(static_cast<const float*>(lw[4]), d, d, wc, d, s)) return 1;
    wc += 1ull * d * d;
    l.w_fc = wc;
    if (launch_cast_pad_bf16(static_cast<const float*>(lw[8]), 4 * d, d, wc, d, s)) return 1;
    wc += 4ull * d * d;
    l.w_proj = wc;
    if (launch_cast_pad_bf16(static_cast<const float*>(lw[10]), d, 4 * d, wc, 4 * d, s)) return 1;
    wc += 4ull * d * d;
  }
  const size_t rows = static_cast<size_t>(max_batch) * L;
  EVR_CHECK_CUDA(cudaMalloc(&h->x, rows * d * 4));
  EVR_CHECK_CUDA(cudaMalloc(&h->hbuf, rows * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->qkv, rows * 3 * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->attn, rows * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->fbuf, rows * 4 * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->eot, static_cast<size_t>(max_batch) * d * 4));
  EVR_CHECK_CUDA(cudaMalloc(&h->eot_ln, static_cast<size_t>(max_batch) * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->emb, static_cast<size_t>(max_batch) * od * 4));
  const size_t tok_rows = max_batch < GraphCache::kMaxBatch ? max_batch : GraphCache::kMaxBatch;
  EVR_CHECK_CUDA(cudaMalloc(&h->tokens, tok_rows * L * sizeof(int32_t)));
  EVR_CHECK_CUDA(cudaStreamSynchronize(s));
  *out = h;
  return 0;
}

extern "C" int evr_text_destroy(evr_text* h) {
  if (!h) return 0;
  h->graphs.destroy();
  cudaFree(h->fparams); cudaFree(h->wparams);
  cudaFree(h->x); cudaFree(h->hbuf); cudaFree(h->qkv); cudaFree(h->attn); cudaFree(h->fbuf);
  cudaFree(h->eot); cudaFree(h->eot_ln); cudaFree(h->emb); cudaFree(h->tokens);
  delete h;
  return 0;
}

static int tgemm(const void* A, int lda, const void* W, int K, int M, int N, const float* bias, void* out, int ldo,
                 int epi, cudaStream_t s) {
  GemmArgs a{};
  a.M = M; a.N = N; a.K = K;
  a.bias = bias; a.out = out; a.ldo = ldo;
  return launch_gemm_bf16(A, lda, W, K, a, epi, s);
}

// embed .. projection GEMM: reads `tokens` and handle buffers, writes h->emb
static int text_body(evr_text* h, const int32_t* tokens, int B, cudaStream_t s) {
  const int d = h->cfg.width, L = h->cfg.context_length, od = h->cfg.out_dim;
  const int M = B * L;
  if (launch_embed_tokens(tokens, h->tok, h->pos, h->x, B, L, d, h->cfg.vocab_size, s)) return 1;
  for (const TextLayerW& l : h->layers) {
    if (launch_layernorm(h->x, d, h->hbuf, d, 1, l.ln1_g, l.ln1_b, M, d, s)) return 1;
    if (tgemm(h->hbuf, d, l.w_in, d, M, 3 * d, l.b_in, h->qkv, 3 * d, EPI_BF16, s)) return 1;
    if (launch_attention(h->qkv, h->attn, B, L, d, h->cfg.heads, /*causal=*/1, s)) return 1;
    if (tgemm(h->attn, d, l.w_out, d, M, d, l.b_out, h->x, d, EPI_F32_RESID, s)) return 1;
    if (launch_layernorm(h->x, d, h->hbuf, d, 1, l.ln2_g, l.ln2_b, M, d, s)) return 1;
    if (tgemm(h->hbuf, d, l.w_fc, d, M, 4 * d, l.b_fc, h->fbuf, 4 * d, EPI_BF16_GELU, s)) return 1;
    if (tgemm(h->fbuf, 4 * d, l.w_proj, 4 * d, M, d, l.b_proj, h->x, d, EPI_F32_RESID, s)) return 1;
  }
  // LayerNorm is per row, so ln_final is applied to the EOT rows only (identical to ln_final(x)[eot])
  if (launch_gather_eot(tokens, h->x, h->eot, B, L, d, s)) return 1;
  if (launch_layernorm(h->eot, d, h->eot_ln, d, 1, h->lnf_g, h->lnf_b, B, d, s)) return 1;
  if (tgemm(h->eot_ln, d, h->w_projT, d, B, od, nullptr, h->emb, od, EPI_F32, s)) return 1;
  return 0;
}

static int text_chunk(evr_text* h, const int32_t* tokens, int B, void* out, int out_dtype, int l2_normalise,
                      cudaStream_t s) {
  const int L = h->cfg.context_length, od = h->cfg.out_dim;
  if (B <= GraphCache::kMaxBatch && graphs_enabled()) {
    // one query is the serving case (embedding_service.py:169): ids are staged so that the chain's pointers are fixed
    EVR_CHECK_CUDA(cudaMemcpyAsync(h->tokens, tokens, static_cast<size_t>(B) * L * sizeof(int32_t),
                                   cudaMemcpyDefault, s));
    if (int r = run_chain(h->graphs, B, s, [&](cudaStream_t cs) { return text_body(h, h->tokens, B, cs); })) return r;
  } else if (text_body(h, tokens, B, s)) {
    return 1;
  }
  if (launch_l2norm_cast(h->emb, out, out_dtype, B, od, l2_normalise, s)) return 1;
  return 0;
}

extern "C" int evr_text_encode(evr_text* h, const int32_t* tokens, int B, void* out, int out_dtype, int l2_normalise,
                               void* stream) {
  EVR_REQUIRE(h && tokens && out, "evr_text_encode: null argument");
  EVR_REQUIRE(out_dtype == 0 || out_dtype == 1, "out_dtype %d not in {0,1}", out_dtype);
  EVR_REQUIRE(B >= 0, "negative batch");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t out_bytes = static_cast<size_t>(h->cfg.out_dim) * (out_dtype == 0 ? 4 : 2);
  for (int b0 = 0; b0 < B; b0 += h->max_batch) {
    const int nb = (B - b0 < h->max_batch) ? (B - b0) : h->max_batch;
    if (int r = text_chunk(h, tokens + static_cast<size_t>(b0) * h->cfg.context_length, nb,
                           static_cast<uint8_t*>(out) + b0 * out_bytes, out_dtype, l2_normalise, s))
      return r;
  }
  return 0;
}

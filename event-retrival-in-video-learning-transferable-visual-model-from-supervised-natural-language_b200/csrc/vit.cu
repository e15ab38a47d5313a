// CLIP vision tower handle: weights repacked once to bf16, persistent workspaces, and the launch
// sequence of encode_image.  Replaces model.encode_image(batch) of openai/CLIP (call sites
// Backend/embedding.py:49, Backend/services/embedding_service.py:41,414,490); structure per
// clip/model.py VisionTransformer.forward.
#include <cstdlib>
#include <vector>

#include "../../include/evr.h"
#include "gemm_tc.cuh"
#include "vit_kernels.cuh"

using namespace evr;

struct LayerW {
  const float *ln1_g, *ln1_b, *b_in, *b_out, *ln2_g, *ln2_b, *b_fc, *b_proj;  // fp32 params
  const __nv_bfloat16 *w_in, *w_out, *w_fc, *w_proj;                          // bf16 [N, K]
};

struct evr_vit {
  evr_vit_cfg cfg;
  int g, L, K0, K0p, max_batch;
  // parameter arenas (owned)
  float* fparams = nullptr;
  __nv_bfloat16* wparams = nullptr;
  const float *cls, *pos, *ln_pre_g, *ln_pre_b, *ln_post_g, *ln_post_b;
  const __nv_bfloat16 *w_conv, *w_projT;
  std::vector<LayerW> layers;
  // workspaces (owned)
  __nv_bfloat16 *a0 = nullptr, *hbuf = nullptr, *qkv = nullptr, *attn = nullptr, *fbuf = nullptr, *cls_ln = nullptr;
  float *x = nullptr, *emb = nullptr;
  GraphCache graphs;  // captured small-batch chains (common.cuh)
  // L2 residency of the fp32 residual stream (touched 4x per layer, between which ~180 MB of activations stream by)
  size_t l2_persist = 0, l2_window_max = 0;
};

// Pins the residual stream `x` in L2 for the kernels that follow on `s` (access-policy window: hits persist, the rest of
// the traffic streams past it); bytes == 0 removes the window.  Best effort: a refusal only costs bandwidth.
static void set_residual_window(const evr_vit* h, size_t bytes, cudaStream_t s) {
  if (h->l2_persist == 0) return;
  cudaStreamAttrValue v{};
  if (bytes > h->l2_window_max) bytes = h->l2_window_max;
  v.accessPolicyWindow.base_ptr = bytes ? h->x : nullptr;
  v.accessPolicyWindow.num_bytes = bytes;
  v.accessPolicyWindow.hitRatio = bytes > h->l2_persist ? static_cast<float>(h->l2_persist) / bytes : 1.0f;
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  if (cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();
}

static int copy_f32(float*& cursor, const void* src, size_t n, const float** dst, cudaStream_t s) {
  EVR_CHECK_CUDA(cudaMemcpyAsync(cursor, src, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
  *dst = cursor;
  cursor += (n + 3) & ~size_t(3);  // keep float4 alignment
  return 0;
}

extern "C" int evr_vit_create(const evr_vit_cfg* cfg, const void* const* w, int n_weights, int max_batch,
                              evr_vit** out) {
  EVR_REQUIRE(cfg && w && out, "evr_vit_create: null argument");
  EVR_REQUIRE(cfg->width == 768 || cfg->width == 1024, "width %d unsupported", cfg->width);
  EVR_REQUIRE(cfg->heads * 64 == cfg->width, "heads*64 must equal width");
  EVR_REQUIRE(cfg->image_size % cfg->patch == 0 && cfg->patch % 2 == 0, "bad image/patch geometry");
  EVR_REQUIRE(cfg->out_dim % 32 == 0, "out_dim must be a multiple of 32");
  EVR_REQUIRE(n_weights == 5 + 12 * cfg->layers + 3, "expected %d weight pointers, got %d", 5 + 12 * cfg->layers + 3,
              n_weights);
  EVR_REQUIRE(max_batch >= 1, "max_batch must be >= 1");
  evr_vit* h = new evr_vit();
  h->cfg = *cfg;
  const int d = cfg->width, nl = cfg->layers, od = cfg->out_dim;
  h->g = cfg->image_size / cfg->patch;
  h->L = h->g * h->g + 1;
  h->K0 = 3 * cfg->patch * cfg->patch;
  h->K0p = (h->K0 + 63) & ~63;
  h->max_batch = max_batch;
  cudaStream_t s = 0;

  // ---- fp32 parameters
  size_t nf = d + static_cast<size_t>(h->L) * d + 4 * d + 64 + static_cast<size_t>(nl) * (4 * d + 3 * d + d + 4 * d + d + 64);
  EVR_CHECK_CUDA(cudaMalloc(&h->fparams, nf * sizeof(float)));
  float* fc = h->fparams;
  if (copy_f32(fc, w[1], d, &h->cls, s)) return 1;
  if (copy_f32(fc, w[2], static_cast<size_t>(h->L) * d, &h->pos, s)) return 1;
  if (copy_f32(fc, w[3], d, &h->ln_pre_g, s)) return 1;
  if (copy_f32(fc, w[4], d, &h->ln_pre_b, s)) return 1;
  const int tail = 5 + 12 * nl;
  if (copy_f32(fc, w[tail], d, &h->ln_post_g, s)) return 1;
  if (copy_f32(fc, w[tail + 1], d, &h->ln_post_b, s)) return 1;

  // ---- bf16 weights, [N, K] K-contiguous
  size_t nw = static_cast<size_t>(d) * h->K0p + static_cast<size_t>(od) * d +
              static_cast<size_t>(nl) * (3ull * d * d + 1ull * d * d + 4ull * d * d + 4ull * d * d);
  EVR_CHECK_CUDA(cudaMalloc(&h->wparams, nw * sizeof(__nv_bfloat16)));
  __nv_bfloat16* wc = h->wparams;
  h->w_conv = wc;
  if (launch_cast_pad_bf16(static_cast<const float*>(w[0]), d, h->K0, wc, h->K0p, s)) return 1;
  wc += static_cast<size_t>(d) * h->K0p;
  h->w_projT = wc;  // proj [d, out] -> [out, d]
  if (launch_transpose_cast_bf16(static_cast<const float*>(w[tail + 2]), d, od, wc, s)) return 1;
  wc += static_cast<size_t>(od) * d;
  h->layers.resize(nl);
  for (int i = 0; i < nl; ++i) {
    const void* const* lw = w + 5 + 12 * i;
    LayerW& l = h->layers[i];
    if (copy_f32(fc, lw[0], d, &l.ln1_g, s)) return 1;
    if (copy_f32(fc, lw[1], d, &l.ln1_b, s)) return 1;
    if (copy_f32(fc, lw[3], 3 * d, &l.b_in, s)) return 1;
    if (copy_f32(fc, lw[5], d, &l.b_out, s)) return 1;
    if (copy_f32(fc, lw[6], d, &l.ln2_g, s)) return 1;
    if (copy_f32(fc, lw[7], d, &l.ln2_b, s)) return 1;
    if (copy_f32(fc, lw[9], 4 * d, &l.b_fc, s)) return 1;
    if (copy_f32(fc, lw[11], d, &l.b_proj, s)) return 1;
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

  // ---- workspaces
  const size_t rows = static_cast<size_t>(max_batch) * h->L;
  const size_t prow = static_cast<size_t>(max_batch) * h->g * h->g;
  EVR_CHECK_CUDA(cudaMalloc(&h->a0, prow * h->K0p * 2));
  EVR_CHECK_CUDA(cudaMemsetAsync(h->a0, 0, prow * h->K0p * 2, s));  // K padding columns stay zero
  EVR_CHECK_CUDA(cudaMalloc(&h->x, rows * d * 4));
  EVR_CHECK_CUDA(cudaMalloc(&h->hbuf, rows * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->qkv, rows * 3 * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->attn, rows * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->fbuf, rows * 4 * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->cls_ln, static_cast<size_t>(max_batch) * d * 2));
  EVR_CHECK_CUDA(cudaMalloc(&h->emb, static_cast<size_t>(max_batch) * od * 4));
  {
    int dev = 0, max_persist = 0, max_window = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    size_t want = rows * d * 4;
    if (want > static_cast<size_t>(max_persist)) want = 0;   // does not fit: no carve-out at all (see encode_chunk)
    if (const char* e = getenv("EVR_L2_PERSIST_MB")) {   // experiments: cap (0 = no window at all)
      const size_t cap = static_cast<size_t>(atoi(e) > 0 ? atoi(e) : 0) << 20;
      if (want > cap) want = cap;
    }
    if (want > 0 && max_window > 0 && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
      h->l2_persist = want;
      h->l2_window_max = max_window;
    } else {
      cudaGetLastError();
    }
  }
  EVR_CHECK_CUDA(cudaStreamSynchronize(s));
  *out = h;
  return 0;
}

extern "C" int evr_vit_destroy(evr_vit* h) {
  if (!h) return 0;
  h->graphs.destroy();
  cudaFree(h->fparams); cudaFree(h->wparams);
  cudaFree(h->a0); cudaFree(h->x); cudaFree(h->hbuf); cudaFree(h->qkv); cudaFree(h->attn); cudaFree(h->fbuf);
  cudaFree(h->cls_ln); cudaFree(h->emb);
  delete h;
  return 0;
}

static int gemm(const void* A, int lda, const void* W, int K, int M, int N, const float* bias, void* out, int ldo,
                int epi, cudaStream_t s, const float* pos = nullptr, int patches = 0, int reverse_m = 0) {
  GemmArgs a{};
  a.reverse_m = reverse_m;
  a.M = M; a.N = N; a.K = K;
  a.bias = bias; a.out = out; a.ldo = ldo; a.pos = pos; a.patches = patches;
  return launch_gemm_bf16(A, lda, W, K, a, epi, s);
}

// patch GEMM .. projection GEMM: reads h->a0, writes h->emb, touches nothing outside the handle
static int encode_body(evr_vit* h, int B, cudaStream_t s) {
  const int d = h->cfg.width, L = h->L, P = h->g * h->g, od = h->cfg.out_dim;
  const int M = B * L;
  // K1/K2: patch embed (+pos) and class rows
  if (gemm(h->a0, h->K0p, h->w_conv, h->K0p, B * P, d, nullptr, h->x, d, EPI_PATCH, s, h->pos, P)) return 1;
  if (launch_cls_rows(h->x, h->cls, h->pos, B, L, d, s)) return 1;
  // ln_pre (fp32 in place: it is the start of the residual stream)
  if (launch_layernorm(h->x, d, h->x, d, 0, h->ln_pre_g, h->ln_pre_b, M, d, s)) return 1;
  // Activations that do not stay in L2 until their consumer runs are walked in alternating directions, so that the
  // consumer starts on the lines its producer touched last: the residual GEMMs go front to back and LayerNorm back to
  // front once the residual stream exceeds L2 (vit_kernels.cu: ViT-L/14, ViT-B/16: -7 % LayerNorm time), and c_fc writes
  // the 4 d-wide MLP buffer back to front so that c_proj finds its first row blocks hot (ViT-B/32 at B = 256, 79 MB
  // beside the pinned residual stream: 2.729 -> 2.697 ms per batch, three same-box A/B pairs; nothing to gain at 269 MB).
  // The same switch on the attention core (qkv back to front) measured no difference and is not kept.
  static const int fc_mode = [] { const char* e = getenv("EVR_FC_REVERSE"); return e ? atoi(e) : -1; }();
  const int fc_reverse = fc_mode >= 0 ? fc_mode : (static_cast<size_t>(M) * 4 * d * 2 > (48u << 20) ? 1 : 0);
  for (const LayerW& l : h->layers) {
    if (launch_layernorm(h->x, d, h->hbuf, d, 1, l.ln1_g, l.ln1_b, M, d, s)) return 1;
    if (gemm(h->hbuf, d, l.w_in, d, M, 3 * d, l.b_in, h->qkv, 3 * d, EPI_BF16, s)) return 1;
    if (launch_attention(h->qkv, h->attn, B, L, d, h->cfg.heads, 0, s)) return 1;
    if (gemm(h->attn, d, l.w_out, d, M, d, l.b_out, h->x, d, EPI_F32_RESID, s)) return 1;
    if (launch_layernorm(h->x, d, h->hbuf, d, 1, l.ln2_g, l.ln2_b, M, d, s)) return 1;
    if (gemm(h->hbuf, d, l.w_fc, d, M, 4 * d, l.b_fc, h->fbuf, 4 * d, EPI_BF16_GELU, s, nullptr, 0, fc_reverse)) return 1;
    if (gemm(h->fbuf, 4 * d, l.w_proj, 4 * d, M, d, l.b_proj, h->x, d, EPI_F32_RESID, s)) return 1;
  }
  // ln_post on the class token, projection, optional L2 normalise
  if (launch_layernorm(h->x, static_cast<size_t>(L) * d, h->cls_ln, d, 1, h->ln_post_g, h->ln_post_b, B, d, s)) return 1;
  if (gemm(h->cls_ln, d, h->w_projT, d, B, od, nullptr, h->emb, od, EPI_F32, s)) return 1;
  return 0;
}

static int encode_chunk(evr_vit* h, const void* pixels, int pix_dtype, int B, void* out, int out_dtype,
                        int l2_normalise, cudaStream_t s) {
  const int d = h->cfg.width, L = h->L, od = h->cfg.out_dim;
  const int M = B * L;
  // K1: im2col reads the caller's pixels; everything up to the projection lives in handle buffers (graph-able)
  if (launch_im2col(pixels, pix_dtype, B, h->cfg.image_size, h->cfg.patch, h->K0p, h->a0, s)) return 1;
  // small batches: x is a few MB, no window (and none in the graph).  A stream that does not fit the persisting
  // carve-out gets none either: a partial window (hitRatio < 1) takes that much of L2 away from the bf16 activations
  // and the weights for a fraction of x -- device-stamped A/B (scripts/gap_probe.py): ViT-B/16 at B = 256 (155 MB)
  // 14.5 -> 11.1 ms per batch without it, ViT-L/14 at B = 128 (135 MB) 23.1 -> 21.7 ms; ViT-B/32 (39 MB, fits)
  // 2.79 -> 2.76 ms with it.
  const bool window = B > GraphCache::kMaxBatch && static_cast<size_t>(M) * d * sizeof(float) <= h->l2_persist;
  if (window) set_residual_window(h, static_cast<size_t>(M) * d * sizeof(float), s);
  const int r = run_chain(h->graphs, B, s, [&](cudaStream_t cs) { return encode_body(h, B, cs); });
  if (r) return r;
  if (launch_l2norm_cast(h->emb, out, out_dtype, B, od, l2_normalise, s)) return 1;
  if (window) set_residual_window(h, 0, s);
  return 0;
}

extern "C" int evr_vit_encode(evr_vit* h, const void* pixels, int pix_dtype, int B, void* out, int out_dtype,
                              int l2_normalise, void* stream) {
  EVR_REQUIRE(h && pixels && out, "evr_vit_encode: null argument");
  EVR_REQUIRE(pix_dtype == 0 || pix_dtype == 1, "pix_dtype %d not in {0,1}", pix_dtype);
  EVR_REQUIRE(out_dtype == 0 || out_dtype == 1, "out_dtype %d not in {0,1}", out_dtype);
  EVR_REQUIRE(B >= 0, "negative batch");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t S = h->cfg.image_size;
  const size_t pix_bytes = 3 * S * S * (pix_dtype == 0 ? 4 : 1);
  const size_t out_bytes = static_cast<size_t>(h->cfg.out_dim) * (out_dtype == 0 ? 4 : 2);
  for (int b0 = 0; b0 < B; b0 += h->max_batch) {
    const int nb = (B - b0 < h->max_batch) ? (B - b0) : h->max_batch;
    if (int r = encode_chunk(h, static_cast<const uint8_t*>(pixels) + b0 * pix_bytes, pix_dtype, nb,
                             static_cast<uint8_t*>(out) + b0 * out_bytes, out_dtype, l2_normalise, s))
      return r;
  }
  return 0;
}

extern "C" int evr_layernorm(const float* in, int64_t in_stride, void* out, int64_t out_stride, int out_bf16,
                             const float* gamma, const float* beta, int rows, int d, void* stream) {
  EVR_REQUIRE(in && out && gamma && beta, "evr_layernorm: null pointer");
  EVR_REQUIRE(rows >= 1 && in_stride >= d && out_stride >= d && in_stride % 4 == 0 && out_stride % 4 == 0,
              "evr_layernorm: bad shape rows=%d d=%d strides %lld / %lld", rows, d, (long long)in_stride,
              (long long)out_stride);
  return launch_layernorm(in, static_cast<size_t>(in_stride), out, static_cast<size_t>(out_stride), out_bf16, gamma, beta,
                          rows, d, static_cast<cudaStream_t>(stream));
}

extern "C" int evr_attention(const void* qkv, void* out, int B, int L, int d, int heads, int causal, int impl,
                             void* stream) {
  EVR_REQUIRE(qkv && out, "evr_attention: null pointer");
  EVR_REQUIRE(B >= 1 && L >= 1 && impl >= 0 && impl <= 2, "evr_attention: bad arguments B=%d L=%d impl=%d", B, L, impl);
  return launch_attention_impl(static_cast<const __nv_bfloat16*>(qkv), static_cast<__nv_bfloat16*>(out), B, L, d, heads,
                               causal, impl, static_cast<cudaStream_t>(stream));
}

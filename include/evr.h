/* libevr.so -- C ABI of the B200-native retrieval hot path.
 *
 * The reference has no FFI layer; its operator API for this path is a set of Python
 * callables.  Each entry point below names the reference call it
 * replaces.  All pointers are DEVICE pointers unless the name ends in _host.  Every
 * function returns 0 on success, non-zero on error (message: evr_last_error()).
 * `stream` is a cudaStream_t passed as void*.  No C++ types cross this boundary.
 */
#ifndef EVR_H_
#define EVR_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* evr_last_error(void);
int evr_version(void);
/* number of kernels launched by this library since load (bench.py's gpu_launches) */
int64_t evr_launch_count(void);

/* Per-kernel-family device timing for the roofline report (bench.py): evr_profile_enable(1) clears the records and
 * makes every launch site bracket its kernel with CUDA events on the launching stream; evr_profile_read sums them.
 * Families: 0 tcgen05 GEMM, 1 attention core, 2 LayerNorm, 3 other ViT kernels, 4 fused scan+top-k,
 * 5 final select/merge, 6 store prepare / re-score.  n >= 7.
 * evr_profile_enable(2): no events (an event between two kernels costs ~5 us and breaks their programmatic overlap);
 * instead every GEMM, LayerNorm and attention launch records, on the device, the earliest %globaltimer at which one of
 * its CTAs got past griddepcontrol.wait and the latest CTA exit; evr_profile_read then reports families 0-2 from those
 * stamps, and evr_profile_timeline hands back the raw (start, end) pairs in launch order (start_end_ns[2 i], [2 i + 1],
 * family[i]; *count launches, at most `capacity`) so that the idle time between kernels can be measured too. */
int evr_profile_enable(int on);
int evr_profile_read(double* ms_by_family, int64_t* launches_by_family, int n);
int evr_profile_timeline(uint64_t* start_end_ns, int* family, int capacity, int* count);

/* ---- store preparation -------------------------------------------------------------
 * replaces  embeddings / np.linalg.norm(embeddings, axis=-1, keepdims=True)
 *           Backend/services/embedding_service.py:210 (get_embeddings)
 * src: [N, D] fp32 (src_dtype 0) or fp16 (src_dtype 1), row-major.
 * dst: [N, D] fp16 unit-norm rows (the HBM-resident store the scan streams).
 * fp16 sources follow numpy's fp16 semantics closely: norm accumulated in fp32, the
 * quotient rounded to fp16.  Zero rows stay zero (compare_models.py:1168-1171 guard). */
int evr_store_prepare(const void* src, int src_dtype, int64_t N, int D, void* dst_fp16, void* stream);

/* ---- fused cosine scan + top-k -------------------------------------------------------
 * replaces  np.dot(embeddings, q.T).flatten() ; np.argsort(sim)[::-1][:top_k]
 *           Backend/services/embedding_service.py:314-320 and :365-371
 * store   [N, D] fp16 (from evr_store_prepare), D % 64 == 0, D <= 768 (ViT-L/14); the row ring is 6 deep up to D = 512, 2 at 768
 * queries [Q, D] fp16
 * out_scores [Q, k] fp32 descending; out_idx [Q, k] int64 = row_offset + local row.
 * 1 <= k <= 1024.  If k > N the tail is filled with score -inf / idx -1.
 * For N >= 32768 the call first runs the same kernel over a sample of n0 rows (65536 at Q <= 128; at larger query
 * batches max(65536, min(3072 k, N/16)); never more than N/4; tiles strided evenly across the whole store) recording
 * only the maximum of every 64-row group; the k-th largest group maximum per query is the admission threshold the full
 * scan starts from (results are unchanged: it is a lower bound on the global k-th best).
 * workspace: evr_topk_workspace_bytes(N, D, Q, k) bytes of device memory. */
int evr_topk_workspace_bytes(int64_t N, int D, int Q, int k, size_t* out_bytes);
int evr_topk_search(const void* store, int64_t N, int D, int64_t row_offset, const void* queries, int Q, int k,
                    float* out_scores, int64_t* out_idx, void* workspace, size_t workspace_bytes, void* stream);

/* The same scan with its scratch in a static device array instead of a caller workspace.  No counterpart in the
 * reference: added for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  The call allocates nothing -- no cudaMalloc /
 * cudaFree, no allocating library routine; inputs and outputs are the caller's.  Limit: the request must fit
 * EVR_TOPK_STATIC_SCRATCH_BYTES (80 MiB: every query batch Q <= 128 with k <= 128, and Q <= 256 with k <= 100, at any N; e.g.
 * Q = 1024 or k = 1024 do not), otherwise the call fails with a message naming both sizes and evr_topk_search is the entry point to use.  Every
 * GPU that loads the library holds the array until the process exits.  All calls on a device share it: calls that
 * overlap on the GPU (two streams, two threads) race -- serialise them; results are bit-identical to evr_topk_search. */
#define EVR_TOPK_STATIC_SCRATCH_BYTES (80u << 20)
int evr_topk_search_static(const void* store, int64_t N, int D, int64_t row_offset, const void* queries, int Q, int k,
                           float* out_scores, int64_t* out_idx, void* stream);

/* P-way merge of per-shard sorted lists (after the NCCL all-gather of section 8e).
 * scores [P, Q, k] fp32, idx [P, Q, k] int64  ->  out [Q, k]. */
int evr_topk_merge(const float* scores, const int64_t* idx, int P, int Q, int k, float* out_scores,
                   int64_t* out_idx, void* stream);

/* Any top_k (the reference sorts all N scores, embedding_service.py:317-320, so it has no upper limit): requests
 * with k > 1024 run as a score pass (each store row read once per chunk of 8 queries, [8, N] fp32 written to the
 * workspace) followed by a per-query radix select + sort over global memory.  1 <= k <= N; same outputs and tie rule
 * (score desc, row asc) as evr_topk_search.  evr_topk_merge_large is the cross-shard merge for such k
 * (scratch >= Q * pow2ceil(k) * 8 bytes). */
int evr_topk_large_workspace_bytes(int64_t N, int Q, int64_t k, size_t* out_bytes);
int evr_topk_search_large(const void* store, int64_t N, int D, int64_t row_offset, const void* queries, int Q,
                          int64_t k, float* out_scores, int64_t* out_idx, void* workspace, size_t workspace_bytes,
                          void* stream);
int evr_topk_merge_large(const float* scores, const int64_t* idx, int P, int Q, int64_t k, float* out_scores,
                         int64_t* out_idx, void* scratch, size_t scratch_bytes, void* stream);

/* per-candidate re-score: replaces np.dot(embeddings[i:i+1], q.T)[0][0]
 * (embedding_service.py:274-279).  rows [R] int64 local row ids; out [R] fp32. */
int evr_rescore(const void* store, int64_t N, int D, const void* query_fp16, const int64_t* rows, int R,
                float* out, void* stream);

/* ---- CLIP vision tower (encode_image) ------------------------------------------------
 * replaces  model.encode_image(batch)  -> VisionTransformer.forward of openai/CLIP
 *           call sites Backend/embedding.py:49, services/embedding_service.py:41,414,490 */
typedef struct evr_vit evr_vit;
typedef struct {
  int image_size; /* 224 */
  int patch;      /* 32 | 16 | 14 */
  int width;      /* 768 | 1024 */
  int layers;     /* 12 | 24 */
  int heads;      /* 12 | 16 */
  int out_dim;    /* 512 | 768 */
} evr_vit_cfg;

/* Weight order (fp32 device pointers, openai state_dict keys):
 *   0 conv1.weight [d,3,p,p]  1 class_embedding [d]  2 positional_embedding [L,d]
 *   3 ln_pre.weight 4 ln_pre.bias
 *   per layer i, base 5 + 12*i:
 *     +0 ln_1.weight +1 ln_1.bias +2 attn.in_proj_weight [3d,d] +3 attn.in_proj_bias
 *     +4 attn.out_proj.weight [d,d] +5 attn.out_proj.bias +6 ln_2.weight +7 ln_2.bias
 *     +8 mlp.c_fc.weight [4d,d] +9 mlp.c_fc.bias +10 mlp.c_proj.weight [d,4d] +11 mlp.c_proj.bias
 *   then ln_post.weight, ln_post.bias, proj [d,out]
 * Weights are repacked once (bf16, K-major) into memory owned by the handle. */
int evr_vit_create(const evr_vit_cfg* cfg, const void* const* weights_fp32, int n_weights, int max_batch,
                   evr_vit** out);
int evr_vit_destroy(evr_vit* h);
/* pixels: [B,3,S,S] fp32 already normalised (pix_dtype 0), or [B,S,S,3] uint8 RGB
 * (pix_dtype 1: /255 + CLIP mean/std fused on device, clip._transform's ToTensor+Normalize).
 * out: [B, out_dim] fp32 (out_dtype 0) or fp16 (out_dtype 1); l2_normalise != 0 applies
 * x / ||x|| (embedding_service.py:45,502). */
int evr_vit_encode(evr_vit* h, const void* pixels, int pix_dtype, int B, void* out, int out_dtype,
                   int l2_normalise, void* stream);

/* ---- frame preprocessing ---------------------------------------------------------------
 * replaces  Resize(n_px, BICUBIC) + CenterCrop(n_px) of clip._transform (PIL ImagingResample, 8 bits per channel)
 *           inside self.preprocess(Image.open(p).convert('RGB'))   Backend/services/embedding_service.py:475
 * src [B, H, W, 3] uint8 RGB (decoded frames of one size) -> dst [B, S, S, 3] uint8, bit-identical to PIL.
 * hbounds [S, 2] (xmin, taps) and hk [S, hksize]: horizontal coefficients of the S cropped output columns in PIL's
 * 22-bit fixed point; vbounds / vk likewise for the S cropped output rows (ymin in source rows).  Source rows
 * [y0, y0 + rows) are the ones the vertical pass reads; tmp [B, rows, S, 3] uint8 is the intermediate image.
 * preprocess.py builds the tables (float64, as Resample.c precompute_coeffs / normalize_coeffs_8bpc). */
int evr_resize_crop_u8(const uint8_t* src, int B, int H, int W, int S, const int32_t* hbounds, const int32_t* hk,
                       int hksize, const int32_t* vbounds, const int32_t* vk, int vksize, int y0, int rows,
                       uint8_t* tmp, uint8_t* dst, void* stream);

/* ---- scene-cut scoring -----------------------------------------------------------------
 * replaces  the per-frame work of PySceneDetect's ContentDetector behind
 *           scene_manager.detect_scenes(frame_source=video_manager)   Backend/segment_video.py:9-13
 *           (cv2.cvtColor(frame, COLOR_BGR2HSV); per channel np.sum(np.abs(curr - last)))
 * frames [T, H, W, 3] uint8 decoded frames (bgr != 0: OpenCV's B,G,R order; 0: R,G,B), consecutive in time.
 * sums [T-1, 3] (zeroed by the call): for the pair (t, t+1) the sums over the pixels of frame[::downscale, ::downscale]
 * of |dH|, |dS|, |dV| of OpenCV's 8-bit HSV (H in 0..179), exact integers.  sdiv_table / hdiv_table: OpenCV's
 * 12-bit fixed-point division tables (256 int32 each, device memory; scenecut.py builds them as color_hsv does).
 * The threshold / min_scene_len rule and the scene list are host logic (scenecut.py). */
int evr_scene_deltas(const uint8_t* frames, int T, int H, int W, int bgr, int downscale, const int32_t* sdiv_table,
                     const int32_t* hdiv_table, unsigned long long* sums, void* stream);

/* ---- CLIP text tower (encode_text) ---------------------------------------------------
 * replaces  model.encode_text(clip.tokenize([query]))   Backend/services/embedding_service.py:169-177
 *           (openai/CLIP CLIP.encode_text: causal transformer, EOT row, @ text_projection)
 * Weight order (fp32 device pointers, openai state_dict keys of the whole CLIP):
 *   0 token_embedding.weight [vocab,d]   1 positional_embedding [ctx,d]
 *   per layer i, base 2 + 12*i: transformer.resblocks.i.{ln_1.weight, ln_1.bias, attn.in_proj_weight [3d,d],
 *     attn.in_proj_bias, attn.out_proj.weight, attn.out_proj.bias, ln_2.weight, ln_2.bias, mlp.c_fc.weight [4d,d],
 *     mlp.c_fc.bias, mlp.c_proj.weight [d,4d], mlp.c_proj.bias}
 *   then ln_final.weight, ln_final.bias, text_projection [d,out]
 * tokens: [B, context_length] int32 device ids as clip.tokenize emits them (the EOT id is the largest of a row).
 * out: [B, out_dim] fp32 (out_dtype 0) or fp16 (1); l2_normalise != 0 applies x / ||x|| (embedding_service.py:180). */
typedef struct evr_text evr_text;
typedef struct {
  int context_length; /* 77 */
  int vocab_size;     /* 49408 */
  int width;          /* 512 | 768 */
  int layers;         /* 12 */
  int heads;          /* width / 64 */
  int out_dim;        /* 512 | 768 */
} evr_text_cfg;
int evr_text_create(const evr_text_cfg* cfg, const void* const* weights_fp32, int n_weights, int max_batch,
                    evr_text** out);
int evr_text_destroy(evr_text* h);
int evr_text_encode(evr_text* h, const int32_t* tokens, int B, void* out, int out_dtype, int l2_normalise,
                    void* stream);

/* building-block GEMM exported for the parity tests:  C[M,N] = A[M,K] * W[N,K]^T (+bias)
 * A, W bf16 row-major; epilogue 0: bf16 out (+bias), 1: bias+QuickGELU bf16,
 * 2: fp32 out = acc + bias + residual (in place on out). */
int evr_gemm_bf16(const void* A, const void* W, const float* bias, void* out, int M, int N, int K, int epilogue,
                  void* stream);

/* ---- JPEG key-frames decoded on the device (nvJPEG batched decode; hardware engines when the stream allows) ------
 * replaces  Image.open(path).convert('RGB')   Backend/services/embedding_service.py:472-475, Backend/embedding.py:45
 * data[i] / lengths[i]: the i-th file's bytes in HOST memory; every image of a batch has the same H x W;
 * dst: DEVICE [n, H, W, 3] uint8 RGB (what evr_resize_crop_u8 takes).  Not bit-identical to libjpeg-turbo / PIL
 * (IDCT and chroma up-sampling differ): opt-in, parity stated on the embeddings (cosine >= 0.999). */
typedef struct evr_jpeg evr_jpeg;
int evr_jpeg_create(int prefer_hardware, evr_jpeg** out);
int evr_jpeg_destroy(evr_jpeg* h);
int evr_jpeg_backend(const evr_jpeg* h); /* nvjpegBackend_t in use: 3 = hardware engine, 2 = GPU hybrid, 0 = default */
int evr_jpeg_info(evr_jpeg* h, const uint8_t* data, size_t length, int* width, int* height, int* components);
int evr_jpeg_decode_batch(evr_jpeg* h, const uint8_t* const* data, const size_t* lengths, int n, int H, int W,
                          uint8_t* dst, void* stream);

/* Building block exported for the parity tests: LayerNorm of the residual stream as the towers run it
 * (openai/CLIP clip/model.py `LayerNorm`: fp32 statistics, eps 1e-5): out[r, :] = (in[r, :] - mean) * rstd * gamma + beta
 * for `rows` rows of width d (512 | 768 | 1024), strides in elements, output bf16 (out_bf16 = 1) or fp32 (may be in
 * place).  One row per warp; 59 MB per launch at 12 800 x 768 runs at 5.4 TB/s whether the rows come from HBM or L2. */
int evr_layernorm(const float* in, int64_t in_stride, void* out, int64_t out_stride, int out_bf16, const float* gamma,
                  const float* beta, int rows, int d, void* stream);

/* building-block attention core exported for the parity tests: softmax(q k^T / 8 [+ causal mask]) v per
 * (sequence, head), head dim 64.  qkv bf16 [B L, 3 d] (q | k | v as the in_proj GEMM writes it), out bf16 [B L, d].
 * impl 0: what the towers use (tcgen05 / TMEM for 64 < L <= 257, warp-level MMA otherwise), 1: warp-level MMA kernels,
 * 2: tcgen05 / TMEM kernel.  Replaces the core of nn.MultiheadAttention in clip/model.py ResidualAttentionBlock. */
int evr_attention(const void* qkv_bf16, void* out_bf16, int B, int L, int d, int heads, int causal, int impl,
                  void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EVR_H_ */

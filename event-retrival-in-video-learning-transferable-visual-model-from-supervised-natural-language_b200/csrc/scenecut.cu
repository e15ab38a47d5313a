// Scene-cut scoring on the device: the per-frame arithmetic of PySceneDetect's ContentDetector, which the reference
// runs on the CPU for every frame of an uploaded video (Backend/segment_video.py:9-13 -> SceneManager.detect_scenes ->
// ContentDetector.process_frame: cv2.cvtColor(frame, COLOR_BGR2HSV), then per channel sum |current - previous|).
// SURVEY.md 8f row 4.  Integer work end to end: OpenCV's 8-bit RGB2HSV_b (12-bit fixed-point division tables, hue
// range 180) is restated exactly, the kernel returns the three integer sums per consecutive frame pair and the host
// (scenecut.py) applies the threshold / min_scene_len rule.  HBM-bound: 3 bytes per pixel per frame; a frame is read
// by two pairs, the second time from L2 (a 1280x720 frame is 2.8 MB).
#include "../../include/evr.h"
#include "common.cuh"

namespace evr {
namespace {

struct Hsv {
  int h, s, v;
};

// c0, c1, c2 in memory order; bgr != 0: (B, G, R), else (R, G, B)
__device__ __forceinline__ Hsv hsv_u8(int c0, int c1, int c2, int bgr, const int* sdiv, const int* hdiv) {
  const int b = bgr ? c0 : c2, g = c1, r = bgr ? c2 : c0;
  const int v = max(max(b, g), r), vmin = min(min(b, g), r), diff = v - vmin;
  Hsv o;
  o.v = v;
  o.s = (diff * sdiv[v] + (1 << 11)) >> 12;
  int h = (v == r) ? (g - b) : (v == g) ? (b - r + 2 * diff) : (r - g + 4 * diff);
  h = (h * hdiv[diff] + (1 << 11)) >> 12;  // arithmetic shift, as OpenCV's int arithmetic
  o.h = h < 0 ? h + 180 : h;
  return o;
}

__device__ __forceinline__ void accumulate(int a0, int a1, int a2, int b0, int b1, int b2, int bgr, const int* sdiv,
                                           const int* hdiv, int& dh, int& ds, int& dv) {
  const Hsv p = hsv_u8(a0, a1, a2, bgr, sdiv, hdiv), q = hsv_u8(b0, b1, b2, bgr, sdiv, hdiv);
  dh += abs(p.h - q.h);  // no hue wrap-around: ContentDetector takes the plain difference
  ds += abs(p.s - q.s);
  dv += abs(p.v - q.v);
}

// grid (x, T-1): pair t = frames t, t+1.  vec (downscale 1, H W a multiple of 4): 4 pixels = three 32-bit words per
// thread and step; otherwise one (sub-sampled) pixel (y f, x f) per thread and step, byte loads.
__global__ void __launch_bounds__(256)
scene_delta_kernel(const uint8_t* __restrict__ frames, int H, int W, int bgr, int f, int vec, const int* __restrict__ sdiv_g,
                   const int* __restrict__ hdiv_g, unsigned long long* __restrict__ sums) {
  __shared__ int sdiv[256], hdiv[256];
  __shared__ int red[3][8];
  sdiv[threadIdx.x] = sdiv_g[threadIdx.x];
  hdiv[threadIdx.x] = hdiv_g[threadIdx.x];
  __syncthreads();
  const size_t frame_bytes = static_cast<size_t>(H) * W * 3;
  const uint8_t* fa = frames + blockIdx.y * frame_bytes;
  const uint8_t* fb = fa + frame_bytes;
  int dh = 0, ds = 0, dv = 0;
  const int stride = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {  // f == 1 and every frame starts on a 4-byte boundary (H W % 4 == 0)
    const int pixels = H * W, groups = pixels >> 2;
    const uint32_t* wa = reinterpret_cast<const uint32_t*>(fa);
    const uint32_t* wb = reinterpret_cast<const uint32_t*>(fb);
    for (int gi = tid; gi < groups; gi += stride) {
      const uint32_t a0 = __ldg(wa + 3 * gi), a1 = __ldg(wa + 3 * gi + 1), a2 = __ldg(wa + 3 * gi + 2);
      const uint32_t b0 = __ldg(wb + 3 * gi), b1 = __ldg(wb + 3 * gi + 1), b2 = __ldg(wb + 3 * gi + 2);
      accumulate(a0 & 255, (a0 >> 8) & 255, (a0 >> 16) & 255, b0 & 255, (b0 >> 8) & 255, (b0 >> 16) & 255, bgr, sdiv,
                 hdiv, dh, ds, dv);
      accumulate(a0 >> 24, a1 & 255, (a1 >> 8) & 255, b0 >> 24, b1 & 255, (b1 >> 8) & 255, bgr, sdiv, hdiv, dh, ds, dv);
      accumulate((a1 >> 16) & 255, a1 >> 24, a2 & 255, (b1 >> 16) & 255, b1 >> 24, b2 & 255, bgr, sdiv, hdiv, dh, ds,
                 dv);
      accumulate((a2 >> 8) & 255, (a2 >> 16) & 255, a2 >> 24, (b2 >> 8) & 255, (b2 >> 16) & 255, b2 >> 24, bgr, sdiv,
                 hdiv, dh, ds, dv);
    }
  } else {
    const int Hs = (H + f - 1) / f, Ws = (W + f - 1) / f;  // frame[::f, ::f]
    for (int p = tid; p < Hs * Ws; p += stride) {
      const int y = p / Ws, x = p - y * Ws;
      const size_t o = (static_cast<size_t>(y) * f * W + static_cast<size_t>(x) * f) * 3;
      accumulate(fa[o], fa[o + 1], fa[o + 2], fb[o], fb[o + 1], fb[o + 2], bgr, sdiv, hdiv, dh, ds, dv);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dh += __shfl_xor_sync(0xffffffffu, dh, o);
    ds += __shfl_xor_sync(0xffffffffu, ds, o);
    dv += __shfl_xor_sync(0xffffffffu, dv, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { red[0][warp] = dh; red[1][warp] = ds; red[2][warp] = dv; }
  __syncthreads();
  if (threadIdx.x < 3) {
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += static_cast<unsigned long long>(red[threadIdx.x][w]);
    atomicAdd(sums + 3 * blockIdx.y + threadIdx.x, t);
  }
}

// Aligned full-resolution path: a thread keeps its 4-pixel groups and walks them through kChunk consecutive frame
// pairs, so every frame is read once and converted to HSV once per chunk (the pair kernel converts each frame twice);
// the HSV of the previous frame stays packed in registers (one word per channel for the 4 pixels).  grid (x, ceil(pairs / kChunk)).
constexpr int kChunk = 8;

struct Words3 {
  uint32_t a0, a1, a2;  // 4 pixels = 12 bytes
};
__device__ __forceinline__ Words3 load_group(const uint32_t* w, int gi) {
  Words3 r;
  r.a0 = __ldg(w + 3 * gi);
  r.a1 = __ldg(w + 3 * gi + 1);
  r.a2 = __ldg(w + 3 * gi + 2);
  return r;
}
// 4 pixels -> three words, one per channel (byte k = pixel k), so that one vabsdiff4.add per channel sums the four
// differences; the packing multiplies run on the FMA pipe, which the conversion leaves idle
__device__ __forceinline__ void hsv_group(const Words3& w, int bgr, const int* sdiv, const int* hdiv, uint32_t (&o)[3]) {
  const Hsv p0 = hsv_u8(w.a0 & 255, (w.a0 >> 8) & 255, (w.a0 >> 16) & 255, bgr, sdiv, hdiv);
  const Hsv p1 = hsv_u8(w.a0 >> 24, w.a1 & 255, (w.a1 >> 8) & 255, bgr, sdiv, hdiv);
  const Hsv p2 = hsv_u8((w.a1 >> 16) & 255, w.a1 >> 24, w.a2 & 255, bgr, sdiv, hdiv);
  const Hsv p3 = hsv_u8((w.a2 >> 8) & 255, (w.a2 >> 16) & 255, w.a2 >> 24, bgr, sdiv, hdiv);
  o[0] = p0.h + p1.h * 256u + p2.h * 65536u + p3.h * 16777216u;
  o[1] = p0.s + p1.s * 256u + p2.s * 65536u + p3.s * 16777216u;
  o[2] = p0.v + p1.v * 256u + p2.v * 65536u + p3.v * 16777216u;
}

__global__ void __launch_bounds__(256, 4)
scene_walk_kernel(const uint8_t* __restrict__ frames, int H, int W, int pairs, int bgr, const int* __restrict__ sdiv_g,
                  const int* __restrict__ hdiv_g, unsigned long long* __restrict__ sums) {
  __shared__ int sdiv[256], hdiv[256];
  __shared__ unsigned long long red[kChunk][3];
  sdiv[threadIdx.x] = sdiv_g[threadIdx.x];
  hdiv[threadIdx.x] = hdiv_g[threadIdx.x];
  if (threadIdx.x < kChunk * 3) (&red[0][0])[threadIdx.x] = 0ull;
  __syncthreads();
  const size_t frame_words = static_cast<size_t>(H) * W * 3 / 4;
  const int t0 = blockIdx.y * kChunk, nt = min(kChunk, pairs - t0);
  const uint32_t* base = reinterpret_cast<const uint32_t*>(frames) + t0 * frame_words;
  const int groups = (H * W) >> 2;
  int acc[kChunk][3];
#pragma unroll
  for (int t = 0; t < kChunk; ++t) acc[t][0] = acc[t][1] = acc[t][2] = 0;
  for (int gi = blockIdx.x * blockDim.x + threadIdx.x; gi < groups; gi += gridDim.x * blockDim.x) {
    uint32_t prev[3], cur[3];
    hsv_group(load_group(base, gi), bgr, sdiv, hdiv, prev);
    Words3 nxt = load_group(base + frame_words, gi);  // nt >= 1; the loads run one frame ahead of the arithmetic
#pragma unroll
    for (int t = 0; t < kChunk; ++t) {
      if (t < nt) {
        const Words3 w = nxt;
        if (t + 1 < nt) nxt = load_group(base + (t + 2) * frame_words, gi);
        hsv_group(w, bgr, sdiv, hdiv, cur);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          acc[t][c] += __vsadu4(prev[c], cur[c]);  // sum over the 4 pixels of |d| of this channel
          prev[c] = cur[c];
        }
      }
    }
  }
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t < kChunk; ++t) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      int v = acc[t][c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && t < nt) atomicAdd(&red[t][c], static_cast<unsigned long long>(v));
    }
  }
  __syncthreads();
  if (threadIdx.x < nt * 3)
    atomicAdd(sums + 3 * static_cast<size_t>(t0) + threadIdx.x, (&red[0][0])[threadIdx.x]);
}

}  // namespace
}  // namespace evr

extern "C" int evr_scene_deltas(const uint8_t* frames, int T, int H, int W, int bgr, int downscale,
                                const int32_t* sdiv_table, const int32_t* hdiv_table, unsigned long long* sums,
                                void* stream) {
  using namespace evr;
  EVR_REQUIRE(frames && sdiv_table && hdiv_table && sums, "evr_scene_deltas: null pointer");
  EVR_REQUIRE(T >= 1 && H >= 1 && W >= 1, "evr_scene_deltas: bad geometry %d x %d x %d", T, H, W);
  EVR_REQUIRE(downscale >= 1, "evr_scene_deltas: downscale %d", downscale);
  EVR_REQUIRE(static_cast<long long>(H) * W < (1ll << 30), "evr_scene_deltas: frame too large");
  EVR_REQUIRE(T - 1 <= 65535, "evr_scene_deltas: at most 65536 frames per call");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (T < 2) return 0;
  EVR_CHECK_CUDA(cudaMemsetAsync(sums, 0, static_cast<size_t>(T - 1) * 3 * sizeof(unsigned long long), s));
  const int vec = downscale == 1 && (static_cast<long long>(H) * W) % 4 == 0 && (reinterpret_cast<uintptr_t>(frames) & 3) == 0;
  const long long work = vec ? (static_cast<long long>(H) * W) / 4
                                        : static_cast<long long>((H + downscale - 1) / downscale) *
                                              ((W + downscale - 1) / downscale);
  long long bx = (work + 255) / 256;
  const long long cap = static_cast<long long>(sm_count()) * 8;  // 8 resident 256-thread blocks per SM, one wave per pair row
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  ProfScope ps(PF_STORE, s);
  if (vec && T - 1 >= 4) {
    // one wave: x blocks per chunk row so that x * rows ~ 8 blocks per SM
    const int rows = (T - 1 + kChunk - 1) / kChunk;
    long long wx = (cap + rows - 1) / rows;
    const long long need = (work + 255) / 256;
    if (wx > need) wx = need;
    if (wx < 1) wx = 1;
    scene_walk_kernel<<<dim3(static_cast<unsigned>(wx), rows), 256, 0, s>>>(frames, H, W, T - 1, bgr ? 1 : 0, sdiv_table,
                                                                           hdiv_table, sums);
  } else {
    scene_delta_kernel<<<dim3(static_cast<unsigned>(bx), T - 1), 256, 0, s>>>(frames, H, W, bgr ? 1 : 0, downscale, vec,
                                                                              sdiv_table, hdiv_table, sums);
  }
  EVR_CHECK_LAUNCH();
  return 0;
}

// Bicubic resize + centre crop of decoded frames on the device, bit-identical to PIL's 8-bit ImagingResample
// (libImaging/Resample.c: ImagingResampleHorizontal_8bpc then ImagingResampleVertical_8bpc), which is what
// clip._transform's Resize(n_px, BICUBIC) + CenterCrop(n_px) runs on the host in the reference
// (Backend/services/embedding_service.py:475, Backend/embedding.py:46).
// The 22-bit fixed-point coefficient tables come from the host (preprocess.py builds them in float64 exactly as
// precompute_coeffs / normalize_coeffs_8bpc do); both passes are integer MACs + a rounding shift + clip to [0, 255].
// Only the cropped window is computed: S output columns in pass 1, S x S pixels in pass 2.
#include "../../include/evr.h"
#include "common.cuh"

namespace evr {

constexpr int kPrecisionBits = 32 - 8 - 2;

__device__ __forceinline__ uint8_t clip8(int v) {
  v >>= kPrecisionBits;  // arithmetic shift, as the C code's lookup index
  return static_cast<uint8_t>(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// tmp[b][r][x][c] = clip8(2^21 + sum_i src[b][y0 + r][xmin(x) + i][c] * hk[x][i])
__global__ void __launch_bounds__(256)
resample_h_kernel(const uint8_t* __restrict__ src, int H, int W, int S, const int2* __restrict__ hb,
                  const int32_t* __restrict__ hk, int hksize, int y0, int rows, uint8_t* __restrict__ tmp) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y, b = blockIdx.z;
  if (x >= S) return;
  const int2 bound = hb[x];
  const uint8_t* p = src + ((static_cast<size_t>(b) * H + y0 + r) * W + bound.x) * 3;
  const int32_t* k = hk + static_cast<size_t>(x) * hksize;
  int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
  for (int i = 0; i < bound.y; ++i) {
    const int c = __ldg(k + i);
    s0 += p[3 * i] * c;
    s1 += p[3 * i + 1] * c;
    s2 += p[3 * i + 2] * c;
  }
  uint8_t* o = tmp + ((static_cast<size_t>(b) * rows + r) * S + x) * 3;
  o[0] = clip8(s0); o[1] = clip8(s1); o[2] = clip8(s2);
}

// dst[b][y][x][c] = clip8(2^21 + sum_i tmp[b][ymin(y) - y0 + i][x][c] * vk[y][i])
__global__ void __launch_bounds__(256)
resample_v_kernel(const uint8_t* __restrict__ tmp, int S, const int2* __restrict__ vb, const int32_t* __restrict__ vk,
                  int vksize, int y0, int rows, uint8_t* __restrict__ dst) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;  // element of the S*3 bytes of one output row
  const int y = blockIdx.y, b = blockIdx.z;
  if (e >= S * 3) return;
  const int2 bound = vb[y];
  const uint8_t* p = tmp + (static_cast<size_t>(b) * rows + (bound.x - y0)) * S * 3 + e;
  const int32_t* k = vk + static_cast<size_t>(y) * vksize;
  int s = 1 << (kPrecisionBits - 1);
  for (int i = 0; i < bound.y; ++i) s += p[static_cast<size_t>(i) * S * 3] * __ldg(k + i);
  dst[(static_cast<size_t>(b) * S + y) * S * 3 + e] = clip8(s);
}

}  // namespace evr

extern "C" int evr_resize_crop_u8(const uint8_t* src, int B, int H, int W, int S, const int32_t* hbounds,
                                  const int32_t* hk, int hksize, const int32_t* vbounds, const int32_t* vk,
                                  int vksize, int y0, int rows, uint8_t* tmp, uint8_t* dst, void* stream) {
  using namespace evr;
  EVR_REQUIRE(src && hbounds && hk && vbounds && vk && tmp && dst, "evr_resize_crop_u8: null argument");
  EVR_REQUIRE(B >= 1 && H >= 1 && W >= 1 && S >= 1, "evr_resize_crop_u8: bad geometry B=%d H=%d W=%d S=%d", B, H, W, S);
  EVR_REQUIRE(hksize >= 1 && vksize >= 1 && y0 >= 0 && rows >= 1 && y0 + rows <= H,
              "evr_resize_crop_u8: source rows [%d, %d) outside the image (H=%d)", y0, y0 + rows, H);
  EVR_REQUIRE(B <= 65535 && rows <= 65535 && S <= 65535, "evr_resize_crop_u8: grid too large");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  ProfScope ps(PF_VIT_MISC, s);
  resample_h_kernel<<<dim3((S + 255) / 256, rows, B), 256, 0, s>>>(src, H, W, S, reinterpret_cast<const int2*>(hbounds),
                                                                   hk, hksize, y0, rows, tmp);
  EVR_CHECK_LAUNCH();
  resample_v_kernel<<<dim3((S * 3 + 255) / 256, S, B), 256, 0, s>>>(tmp, S, reinterpret_cast<const int2*>(vbounds), vk,
                                                                    vksize, y0, rows, dst);
  EVR_CHECK_LAUNCH();
  return 0;
}

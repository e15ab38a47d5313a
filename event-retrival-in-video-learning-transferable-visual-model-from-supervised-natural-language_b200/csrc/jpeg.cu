// JPEG key-frames decoded on the device: the `.jpg` files of a frame folder go from host bytes to
// interleaved RGB in HBM through nvJPEG's batched decoder (the B200's hardware JPEG engines when the stream is a
// single-scan baseline JPEG, nvJPEG's GPU Huffman path otherwise), where `evr_resize_crop_u8` and the tower take over.
//
// Replaces  Image.open(path).convert('RGB')  of the ingest loop, Backend/services/embedding_service.py:472-475 (and
// Backend/embedding.py:45).  nvJPEG is a library (CUDA toolkit); the pixels it produces are NOT bit-identical to
// libjpeg-turbo's (different IDCT and chroma up-sampling), so this path is opt-in and its parity is stated on the
// embeddings (cosine >= 0.999 against the PIL-decoded path, tests/test_jpeg_gpu.py); PIL stays the exact mode.
#include <nvjpeg.h>

#include <vector>

#include "../../include/evr.h"
#include "common.cuh"

using namespace evr;

struct evr_jpeg {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  int backend = 0;     // nvjpegBackend_t actually in use
  int batch = 0;       // batch size the state is initialised for
};

#define EVR_CHECK_NVJPEG(expr)                                                          \
  do {                                                                                  \
    nvjpegStatus_t s_ = (expr);                                                         \
    if (s_ != NVJPEG_STATUS_SUCCESS) {                                                  \
      evr::set_error("%s:%d: %s -> nvjpeg status %d", __FILE__, __LINE__, #expr, (int)s_); \
      return 1;                                                                         \
    }                                                                                   \
  } while (0)

extern "C" int evr_jpeg_create(int prefer_hardware, evr_jpeg** out) {
  EVR_REQUIRE(out, "evr_jpeg_create: null out");
  evr_jpeg* h = new evr_jpeg();
  nvjpegStatus_t st = NVJPEG_STATUS_NOT_INITIALIZED;
  if (prefer_hardware) {
    st = nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &h->handle);
    if (st == NVJPEG_STATUS_SUCCESS) h->backend = NVJPEG_BACKEND_HARDWARE;
  }
  if (st != NVJPEG_STATUS_SUCCESS) {
    st = nvjpegCreateEx(NVJPEG_BACKEND_GPU_HYBRID, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &h->handle);
    if (st == NVJPEG_STATUS_SUCCESS) h->backend = NVJPEG_BACKEND_GPU_HYBRID;
  }
  if (st != NVJPEG_STATUS_SUCCESS) {
    st = nvjpegCreateSimple(&h->handle);
    h->backend = NVJPEG_BACKEND_DEFAULT;
  }
  if (st != NVJPEG_STATUS_SUCCESS) {
    delete h;
    evr::set_error("evr_jpeg_create: nvjpegCreate failed with status %d", (int)st);
    return 1;
  }
  if (nvjpegJpegStateCreate(h->handle, &h->state) != NVJPEG_STATUS_SUCCESS) {
    nvjpegDestroy(h->handle);
    delete h;
    evr::set_error("evr_jpeg_create: nvjpegJpegStateCreate failed");
    return 1;
  }
  *out = h;
  return 0;
}

extern "C" int evr_jpeg_backend(const evr_jpeg* h) { return h ? h->backend : -1; }

extern "C" int evr_jpeg_destroy(evr_jpeg* h) {
  if (!h) return 0;
  if (h->state) nvjpegJpegStateDestroy(h->state);
  if (h->handle) nvjpegDestroy(h->handle);
  delete h;
  return 0;
}

extern "C" int evr_jpeg_info(evr_jpeg* h, const uint8_t* data, size_t length, int* width, int* height,
                             int* components) {
  EVR_REQUIRE(h && data && width && height, "evr_jpeg_info: null argument");
  int nc = 0, w[NVJPEG_MAX_COMPONENT] = {}, hh[NVJPEG_MAX_COMPONENT] = {};
  nvjpegChromaSubsampling_t ss;
  EVR_CHECK_NVJPEG(nvjpegGetImageInfo(h->handle, data, length, &nc, &ss, w, hh));
  *width = w[0];
  *height = hh[0];
  if (components) *components = nc;
  return 0;
}

extern "C" int evr_jpeg_decode_batch(evr_jpeg* h, const uint8_t* const* data, const size_t* lengths, int n, int H,
                                     int W, uint8_t* dst, void* stream) {
  EVR_REQUIRE(h && data && lengths && dst, "evr_jpeg_decode_batch: null argument");
  EVR_REQUIRE(n >= 1 && H >= 1 && W >= 1, "evr_jpeg_decode_batch: bad shape n=%d H=%d W=%d", n, H, W);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int i = 0; i < n; ++i) {   // every frame of the batch must have the announced size (rows of one tensor)
    int nc = 0, w[NVJPEG_MAX_COMPONENT] = {}, hh[NVJPEG_MAX_COMPONENT] = {};
    nvjpegChromaSubsampling_t ss;
    EVR_CHECK_NVJPEG(nvjpegGetImageInfo(h->handle, data[i], lengths[i], &nc, &ss, w, hh));
    EVR_REQUIRE(w[0] == W && hh[0] == H, "evr_jpeg_decode_batch: image %d is %dx%d, batch is %dx%d", i, w[0], hh[0], W, H);
  }
  if (h->batch != n) {
    EVR_CHECK_NVJPEG(nvjpegDecodeBatchedInitialize(h->handle, h->state, n, 1, NVJPEG_OUTPUT_RGBI));
    h->batch = n;
  }
  std::vector<nvjpegImage_t> outs(n);
  const size_t frame = static_cast<size_t>(H) * W * 3;
  for (int i = 0; i < n; ++i) {
    for (int c = 0; c < NVJPEG_MAX_COMPONENT; ++c) { outs[i].channel[c] = nullptr; outs[i].pitch[c] = 0; }
    outs[i].channel[0] = dst + i * frame;
    outs[i].pitch[0] = static_cast<size_t>(W) * 3;
  }
  nvjpegStatus_t st = nvjpegDecodeBatched(h->handle, h->state, data, lengths, outs.data(), s);
  if (st != NVJPEG_STATUS_SUCCESS) {
    h->batch = 0;   // a failed batch leaves the state to be re-initialised (nvjpeg.h)
    evr::set_error("evr_jpeg_decode_batch: nvjpegDecodeBatched -> status %d (backend %d)", (int)st, h->backend);
    return 1;
  }
  count_launch();
  return 0;
}

// JPEG key-frames decoded on the device: the `.jpg` files of a frame folder go from host bytes to
// interleaved RGB in HBM through nvJPEG's batched decoder (the B200's hardware JPEG engines when the stream is a
// single-scan baseline JPEG, nvJPEG's GPU Huffman path otherwise), where `evr_resize_crop_u8` and the tower take over.
//
// Replaces  Image.open(path).convert('RGB')  of the ingest loop, Backend/services/embedding_service.py:472-475 (and
// Backend/embedding.py:45).  nvJPEG is a library (CUDA toolkit); the pixels it produces are NOT bit-identical to
// libjpeg-turbo's (different IDCT and chroma up-sampling), so this path is opt-in and its parity is stated on the
// embeddings (cosine >= 0.999 against the PIL-decoded path, tests/test_jpeg_gpu.py); PIL stays the exact mode.
#include <nvjpeg.h>

#include <atomic>
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "../../include/evr.h"
#include "common.cuh"

using namespace evr;

// nvJPEG allocates and frees its work buffers around every batched call; with the default allocators that is a
// cudaMalloc / cudaFree (a device-wide synchronisation) per call, ~20 ms per batch whatever its size.  These pools
// hand the same blocks back: power-of-two buckets, kept until the last decoder is destroyed.
namespace {
struct BlockPool {
  std::mutex mu;
  std::unordered_map<void*, size_t> bucket_of;             // every block ever handed out -> its bucket size
  std::unordered_map<size_t, std::vector<void*>> free_list;
  bool pinned;
  explicit BlockPool(bool pin) : pinned(pin) {}
  static size_t bucket(size_t n) {
    size_t b = 4096;
    while (b < n) b <<= 1;
    return b;
  }
  int get(void** p, size_t n) {
    const size_t b = bucket(n);
    std::lock_guard<std::mutex> g(mu);
    auto& fl = free_list[b];
    if (!fl.empty()) {
      *p = fl.back();
      fl.pop_back();
      return 0;
    }
    cudaError_t e = pinned ? cudaMallocHost(p, b) : cudaMalloc(p, b);
    if (e != cudaSuccess) return 1;
    bucket_of[*p] = b;
    return 0;
  }
  int put(void* p) {
    if (!p) return 0;
    std::lock_guard<std::mutex> g(mu);
    auto it = bucket_of.find(p);
    if (it == bucket_of.end()) return 1;
    free_list[it->second].push_back(p);
    return 0;
  }
  void release() {
    std::lock_guard<std::mutex> g(mu);
    for (auto& kv : free_list)
      for (void* p : kv.second) {
        if (pinned) cudaFreeHost(p); else cudaFree(p);
        bucket_of.erase(p);
      }
    free_list.clear();
  }
};
BlockPool g_dev_pool(false), g_pin_pool(true);
std::atomic<int> g_decoders{0};
int dev_malloc(void** p, size_t n) { return g_dev_pool.get(p, n); }
int dev_free(void* p) { return g_dev_pool.put(p); }
int pin_malloc(void** p, size_t n, unsigned int) { return g_pin_pool.get(p, n); }
int pin_free(void* p) { return g_pin_pool.put(p); }
}  // namespace

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
  static nvjpegDevAllocator_t dev_alloc = {dev_malloc, dev_free};       // nvJPEG may keep the pointers
  static nvjpegPinnedAllocator_t pin_alloc = {pin_malloc, pin_free};
  nvjpegDevAllocator_t* da = getenv("EVR_JPEG_NO_POOL") ? nullptr : &dev_alloc;
  nvjpegPinnedAllocator_t* pa = getenv("EVR_JPEG_NO_POOL") ? nullptr : &pin_alloc;
  if (prefer_hardware) {
    st = nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE, da, pa, NVJPEG_FLAGS_DEFAULT, &h->handle);
    if (st == NVJPEG_STATUS_SUCCESS) h->backend = NVJPEG_BACKEND_HARDWARE;
  }
  if (st != NVJPEG_STATUS_SUCCESS) {
    st = nvjpegCreateEx(NVJPEG_BACKEND_GPU_HYBRID, da, pa, NVJPEG_FLAGS_DEFAULT, &h->handle);
    if (st == NVJPEG_STATUS_SUCCESS) h->backend = NVJPEG_BACKEND_GPU_HYBRID;
  }
  if (st != NVJPEG_STATUS_SUCCESS) {
    st = nvjpegCreateEx(NVJPEG_BACKEND_DEFAULT, da, pa, NVJPEG_FLAGS_DEFAULT, &h->handle);
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
  g_decoders.fetch_add(1);
  *out = h;
  return 0;
}

extern "C" int evr_jpeg_backend(const evr_jpeg* h) { return h ? h->backend : -1; }

extern "C" int evr_jpeg_destroy(evr_jpeg* h) {
  if (!h) return 0;
  if (h->state) nvjpegJpegStateDestroy(h->state);
  if (h->handle) nvjpegDestroy(h->handle);
  delete h;
  if (g_decoders.fetch_sub(1) == 1) {   // last decoder gone: give the pooled blocks back
    g_dev_pool.release();
    g_pin_pool.release();
  }
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
    int threads = 1;   // host threads nvJPEG may use for the parts it keeps on the CPU (EVR_JPEG_THREADS: experiments)
    if (const char* e = getenv("EVR_JPEG_THREADS")) threads = atoi(e) > 0 ? atoi(e) : 1;
    EVR_CHECK_NVJPEG(nvjpegDecodeBatchedInitialize(h->handle, h->state, n, threads, NVJPEG_OUTPUT_RGBI));
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

// libevr.so: error string, version, launch counter, tensor-map encoding.
#include <mutex>
#include <vector>

#include "../../include/evr.h"
#include <cstdlib>

#include "common.cuh"

namespace evr {

static thread_local char g_err[1024] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode() {
  static encode_tiled_fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_tiled_fn>(p);
  });
  return fn;
}

int make_tmap_2d_16b(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_bytes,
                     uint32_t box_rows, bool is_bf16) {
  encode_tiled_fn enc = get_encode();
  EVR_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  EVR_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tensor base %p not 16-B aligned", base);
  EVR_REQUIRE(row_stride_bytes % 16 == 0, "row stride %llu not a multiple of 16 B",
              (unsigned long long)row_stride_bytes);
  EVR_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows %u out of range", box_rows);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  EVR_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed: CUresult %d (rows %llu cols %llu stride %llu box %u)",
              (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)row_stride_bytes,
              box_rows);
  return 0;
}

// [batches][rows][cols] view of a packed activation: a box of `box_rows` rows never crosses into the next batch entry --
// rows past `rows` are out of bounds in their own dimension and arrive as zeros.
int make_tmap_3d_16b(CUtensorMap* out, const void* base, uint64_t batches, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_bytes, uint32_t box_rows, bool is_bf16) {
  encode_tiled_fn enc = get_encode();
  EVR_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  EVR_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tensor base %p not 16-B aligned", base);
  EVR_REQUIRE(row_stride_bytes % 16 == 0, "row stride %llu not a multiple of 16 B",
              (unsigned long long)row_stride_bytes);
  EVR_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows %u out of range", box_rows);
  cuuint64_t dims[3] = {cols, rows, batches};
  cuuint64_t strides[2] = {row_stride_bytes, row_stride_bytes * rows};
  cuuint32_t box[3] = {64, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  EVR_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (3-D) failed: CUresult %d (batches %llu rows %llu cols %llu)",
              (int)r, (unsigned long long)batches, (unsigned long long)rows, (unsigned long long)cols);
  return 0;
}

int sm_count() {
  static int cached[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  int& n = cached[dev & 63];
  if (n == 0 && (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)) n = 148;
  return n;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("EVR_NO_PDL");
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}

bool tail_split_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("EVR_NO_TAILSPLIT");
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}

bool graphs_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("EVR_NO_GRAPH");
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}

int graph_begin(GraphCache& g, int B, cudaStream_t* cap_stream) {
  (void)B;
  if (!g.cap && cudaStreamCreateWithFlags(&g.cap, cudaStreamNonBlocking) != cudaSuccess) {
    cudaGetLastError();
    return 1;
  }
  if (cudaStreamBeginCapture(g.cap, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
    cudaGetLastError();
    return 1;
  }
  *cap_stream = g.cap;
  return 0;
}

int graph_end(GraphCache& g, int B, int body_status, int64_t launches_before) {
  cudaGraph_t graph = nullptr;
  const cudaError_t e = cudaStreamEndCapture(g.cap, &graph);
  const int64_t captured = g_launches.load(std::memory_order_relaxed) - launches_before;
  g_launches.fetch_sub(captured, std::memory_order_relaxed);  // recorded, not run
  if (e != cudaSuccess || body_status != 0 || !graph) {
    cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    return 1;
  }
  const cudaError_t ei = cudaGraphInstantiate(&g.exec[B], graph, 0);
  cudaGraphDestroy(graph);
  if (ei != cudaSuccess) {
    cudaGetLastError();
    g.exec[B] = nullptr;
    return 1;
  }
  g.kernels[B] = static_cast<int>(captured);
  g.state[B] = 2;
  return 0;
}

int graph_replay(GraphCache& g, int B, cudaStream_t s) {
  EVR_CHECK_CUDA(cudaGraphLaunch(g.exec[B], s));
  count_launch(g.kernels[B]);
  return 0;
}

void GraphCache::destroy() {
  for (auto& e : exec)
    if (e) { cudaGraphExecDestroy(e); e = nullptr; }
  if (cap) { cudaStreamDestroy(cap); cap = nullptr; }
  for (auto& st : state) st = 0;
}

// ---- per-family kernel timing ---------------------------------------------------------------------
std::atomic<int> g_profile_on{0};
struct ProfRec {
  int family;
  cudaEvent_t e0, e1;
};
static std::mutex g_prof_mu;
static std::vector<ProfRec*> g_prof;

void prof_begin(int family, cudaStream_t s, void** token) {
  ProfRec* r = new ProfRec();
  r->family = family;
  if (cudaEventCreate(&r->e0) != cudaSuccess || cudaEventCreate(&r->e1) != cudaSuccess) {
    delete r;
    return;
  }
  cudaEventRecord(r->e0, s);
  *token = r;
}
void prof_end(void* token, cudaStream_t s) {
  ProfRec* r = static_cast<ProfRec*>(token);
  cudaEventRecord(r->e1, s);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof.push_back(r);
}
// device-timer slots: pool[2 i] = min start, pool[2 i + 1] = max end of launch i
static constexpr int kProfSlots = 16384;
static unsigned long long* g_dt_pool = nullptr;
static std::atomic<int> g_dt_used{0};
static uint8_t g_dt_family[kProfSlots];
unsigned long long* prof_device_slot(int family) {
  if (g_profile_on.load(std::memory_order_relaxed) != 2 || g_dt_pool == nullptr) return nullptr;
  const int i = g_dt_used.fetch_add(1, std::memory_order_relaxed);
  if (i >= kProfSlots) return nullptr;
  g_dt_family[i] = static_cast<uint8_t>(family);
  return g_dt_pool + 2 * i;
}
static int prof_device_reset() {
  if (g_dt_pool == nullptr) EVR_CHECK_CUDA(cudaMalloc(&g_dt_pool, 2 * kProfSlots * sizeof(unsigned long long)));
  std::vector<unsigned long long> init(2 * kProfSlots);
  for (int i = 0; i < kProfSlots; ++i) { init[2 * i] = ~0ull; init[2 * i + 1] = 0ull; }
  EVR_CHECK_CUDA(cudaMemcpy(g_dt_pool, init.data(), init.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  g_dt_used.store(0);
  return 0;
}

static void prof_clear() {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (ProfRec* r : g_prof) {
    cudaEventDestroy(r->e0);
    cudaEventDestroy(r->e1);
    delete r;
  }
  g_prof.clear();
  g_dt_used.store(0);
}

}  // namespace evr

extern "C" {
int evr_profile_enable(int on) {
  evr::prof_clear();
  if (on == 2) {
    if (int r = evr::prof_device_reset()) return r;
  }
  evr::g_profile_on.store(on == 2 ? 2 : (on ? 1 : 0));
  return 0;
}
int evr_profile_timeline(uint64_t* start_end_ns, int* family, int capacity, int* count) {
  EVR_REQUIRE(start_end_ns && family && count, "evr_profile_timeline: null argument");
  int used = evr::g_dt_used.load() < evr::kProfSlots ? evr::g_dt_used.load() : evr::kProfSlots;
  if (used > capacity) used = capacity;
  *count = 0;
  if (evr::g_dt_pool == nullptr || used <= 0) return 0;
  EVR_CHECK_CUDA(cudaDeviceSynchronize());
  EVR_CHECK_CUDA(cudaMemcpy(start_end_ns, evr::g_dt_pool, 2 * static_cast<size_t>(used) * sizeof(uint64_t),
                            cudaMemcpyDeviceToHost));
  for (int i = 0; i < used; ++i) family[i] = evr::g_dt_family[i];
  *count = used;
  return 0;
}
int evr_profile_read(double* ms_by_family, int64_t* launches_by_family, int n) {
  EVR_REQUIRE(ms_by_family && launches_by_family && n >= evr::PF_COUNT, "evr_profile_read: need %d slots",
              (int)evr::PF_COUNT);
  for (int i = 0; i < n; ++i) { ms_by_family[i] = 0.0; launches_by_family[i] = 0; }
  std::lock_guard<std::mutex> lk(evr::g_prof_mu);
  for (evr::ProfRec* r : evr::g_prof) {
    EVR_CHECK_CUDA(cudaEventSynchronize(r->e1));
    float ms = 0.f;
    EVR_CHECK_CUDA(cudaEventElapsedTime(&ms, r->e0, r->e1));
    ms_by_family[r->family] += ms;
    launches_by_family[r->family] += 1;
  }
  // device-timer mode: launches stamped by the kernels themselves
  const int used = evr::g_dt_used.load() < evr::kProfSlots ? evr::g_dt_used.load() : evr::kProfSlots;
  if (evr::g_dt_pool != nullptr && used > 0) {
    EVR_CHECK_CUDA(cudaDeviceSynchronize());
    std::vector<unsigned long long> t(2 * static_cast<size_t>(used));
    EVR_CHECK_CUDA(cudaMemcpy(t.data(), evr::g_dt_pool, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    for (int i = 0; i < used; ++i) {
      if (t[2 * i + 1] > t[2 * i]) {
        ms_by_family[evr::g_dt_family[i]] += static_cast<double>(t[2 * i + 1] - t[2 * i]) * 1e-6;
        launches_by_family[evr::g_dt_family[i]] += 1;
      }
    }
  }
  return 0;
}
const char* evr_last_error(void) { return evr::g_err; }
int evr_version(void) { return 100; }
int64_t evr_launch_count(void) { return evr::g_launches.load(); }
}

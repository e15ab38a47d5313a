// Host-side plumbing shared by every translation unit of libevr.so:
// error string, launch counter, TMA tensor-map encoding through the driver entry point.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

namespace evr {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define EVR_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) {                                                                   \
      evr::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_));    \
      return 1;                                                                                \
    }                                                                                          \
  } while (0)

#define EVR_REQUIRE(cond, ...)        \
  do {                                \
    if (!(cond)) {                    \
      evr::set_error(__VA_ARGS__);    \
      return 2;                       \
    }                                 \
  } while (0)

// after a kernel launch: catches launch-configuration errors without synchronising
#define EVR_CHECK_LAUNCH()                                                                     \
  do {                                                                                         \
    cudaError_t e_ = cudaGetLastError();                                                       \
    if (e_ != cudaSuccess) {                                                                   \
      evr::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return 1;                                                                                \
    }                                                                                          \
    evr::count_launch();                                                                       \
  } while (0)

// 2-D row-major [rows, cols] 16-bit tensor, box = [box_rows, 64 cols] (=128 B inner), 128-B swizzle.
// Out-of-bounds box elements are filled with zeros.
int make_tmap_2d_16b(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_bytes,
                     uint32_t box_rows, bool is_bf16);

int make_tmap_3d_16b(CUtensorMap* out, const void* base, uint64_t batches, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_bytes, uint32_t box_rows, bool is_bf16);

int sm_count();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: a call site remembers the largest size it has
// configured on each device (one process per GPU is the design, but a process that touches several GPUs must work).
struct PerDeviceMax {
  int v[64] = {};
  bool needs(int want) {
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    if (want > v[dev]) { v[dev] = want; return true; }
    return false;
  }
};

// Programmatic dependent launch (PDL): kernels of one encode are launched with programmatic stream serialisation, so
// a kernel's CTAs become resident -- and run their prologue (barrier init, TMEM allocation, descriptor prefetch) --
// while the previous kernel drains; `pdl_wait()` blocks until that previous grid has completed and flushed its
// writes.  Every kernel launched this way calls pdl_wait() before its first access to activations (which also keeps
// completion transitive along the chain).  EVR_NO_PDL=1 turns the attribute off.
bool pdl_enabled();
bool tail_split_enabled();  // gemm_tc.cu: strips for the last round's tiles; EVR_NO_TAILSPLIT=1 turns it off (A/B runs)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

// Optional per-kernel-family timing (bench.py's roofline leg): when enabled through evr_profile_enable(1), every
// launch site brackets its kernel with a pair of CUDA events on the launching stream.  Off by default (no cost).
enum ProfFamily : int {
  PF_GEMM = 0,      // gemm_tc_kernel (all epilogues)
  PF_ATTN = 1,      // attention_kernel
  PF_LN = 2,        // layernorm_kernel
  PF_VIT_MISC = 3,  // im2col, cls rows, final normalise/cast
  PF_SCAN = 4,      // topk_scan_kernel
  PF_FINAL = 5,     // topk_final_kernel (select / merge)
  PF_STORE = 6,     // store_prepare_kernel, rescore_kernel
  PF_COUNT = 7
};
extern std::atomic<int> g_profile_on;
void prof_begin(int family, cudaStream_t s, void** token);
void prof_end(void* token, cudaStream_t s);
// evr_profile_enable(2): no events; the GEMM, LayerNorm and attention kernels stamp %globaltimer themselves (first CTA
// past griddepcontrol.wait, last CTA exit).  Returns the slot of the next launch, or nullptr when that mode is off /
// the pool is used up.
unsigned long long* prof_device_slot(int family);
#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long prof_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// thread 0 of every CTA: [0] = earliest time a CTA got past griddepcontrol.wait, [1] = latest CTA exit
__device__ __forceinline__ void prof_stamp_begin(unsigned long long* p) {
  if (p != nullptr && threadIdx.x == 0) atomicMin(p, prof_now_ns());
}
__device__ __forceinline__ void prof_stamp_end(unsigned long long* p) {
  if (p != nullptr && threadIdx.x == 0) atomicMax(p + 1, prof_now_ns());
}
#endif
struct ProfScope {
  void* token = nullptr;
  cudaStream_t s;
  ProfScope(int family, cudaStream_t stream) : s(stream) {
    if (g_profile_on.load(std::memory_order_relaxed) == 1) prof_begin(family, stream, &token);
  }
  ~ProfScope() {
    if (token) prof_end(token, s);
  }
};

// Small-batch launch path.  Between the first kernel that reads caller memory and the last one that writes it, an
// encode touches only handle-owned buffers, so for a given batch size the chain (~90 kernels) is a fixed DAG.  At
// batch <= kMaxBatch the step is bound by the host's launch rate (one query = 0.57 ms of launches for ~0.2 ms of GPU
// work), so the chain is captured once per batch size -- on a private stream, after one plain pass has done every
// lazy cudaFuncSetAttribute -- and replayed as one cudaGraphLaunch on the caller's stream.  PDL edges are kept by the
// capture.  Off while per-family profiling is on, or with EVR_NO_GRAPH=1.
struct GraphCache {
  static constexpr int kMaxBatch = 32;
  cudaStream_t cap = nullptr;
  cudaGraphExec_t exec[kMaxBatch + 1] = {};
  int kernels[kMaxBatch + 1] = {};
  uint8_t state[kMaxBatch + 1] = {};  // 0 never run, 1 ran plain, 2 captured, 3 capture failed (stay plain)
  void destroy();
};
bool graphs_enabled();
// body(stream) enqueues the chain; returns body's status (0 = ok)
int graph_begin(GraphCache& g, int B, cudaStream_t* cap_stream);   // 0: capture into *cap_stream; <0: run plain; >0: replay
int graph_end(GraphCache& g, int B, int body_status, int64_t launches_before);
int graph_replay(GraphCache& g, int B, cudaStream_t s);

template <typename F>
inline int run_chain(GraphCache& g, int B, cudaStream_t s, F&& body) {
  if (B > GraphCache::kMaxBatch || !graphs_enabled() || g_profile_on.load(std::memory_order_relaxed)) return body(s);
  if (g.state[B] == 0) { g.state[B] = 1; return body(s); }
  if (g.state[B] == 3) return body(s);
  if (g.state[B] == 1) {
    cudaStream_t cs = nullptr;
    if (graph_begin(g, B, &cs) != 0) { g.state[B] = 3; return body(s); }
    const int64_t before = g_launches.load(std::memory_order_relaxed);
    const int st = body(cs);
    if (graph_end(g, B, st, before) != 0) { g.state[B] = 3; return body(s); }
  }
  return graph_replay(g, B, s);
}

}  // namespace evr

// Host side of the tcgen05 GEMM: tile-shape choice, tensor maps, launch; plus the exported
// building-block entry point evr_gemm_bf16 (include/evr.h) used by the parity tests.
#include <cstdlib>

#include "gemm_tc.cuh"

#include "../../include/evr.h"

namespace evr {

template <int BLOCK_N, int EPI, int CG>
static int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tbt, const GemmArgs& args,
                      int grid, cudaStream_t stream) {
  using Cfg = GemmCfg<BLOCK_N, EPI, CG>;
  auto kern = gemm_tc_kernel<BLOCK_N, EPI, CG>;
  static PerDeviceMax configured;  // per instantiation
  if (configured.needs(Cfg::kSmemBytes))
    EVR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in the kernel
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  ProfScope ps(PF_GEMM, stream);
  EVR_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, tbt, args));
  EVR_CHECK_LAUNCH();
  return 0;
}

template <int BLOCK_N, int CG>
static int launch_epi(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tbt, const GemmArgs& args,
                      int epilogue, int grid, cudaStream_t stream) {
  switch (epilogue) {
    case EPI_BF16: return launch_one<BLOCK_N, EPI_BF16, CG>(ta, tb, tbt, args, grid, stream);
    case EPI_BF16_GELU: return launch_one<BLOCK_N, EPI_BF16_GELU, CG>(ta, tb, tbt, args, grid, stream);
    case EPI_F32_RESID: return launch_one<BLOCK_N, EPI_F32_RESID, CG>(ta, tb, tbt, args, grid, stream);
    case EPI_F32: return launch_one<BLOCK_N, EPI_F32, CG>(ta, tb, tbt, args, grid, stream);
    case EPI_PATCH: return launch_one<BLOCK_N, EPI_PATCH, CG>(ta, tb, tbt, args, grid, stream);
  }
  set_error("unknown GEMM epilogue %d", epilogue);
  return 2;
}

// Strips per tile of the last, partly filled round: the largest divisor that keeps the strips a multiple of 32 columns
// and the tail round within one pass over the units.  (tail == 0: the rounds are full, nothing to split.)
static int choose_tail_split(int tiles, int units, int bn) {
  const int tail = tiles % units;
  if (tail == 0 || !tail_split_enabled()) return 1;
  int best = 1;
  for (int sp = 2; sp <= 8; ++sp)
    if (bn % (32 * sp) == 0 && tail * sp <= units) best = sp;
  return best;
}

// Pick the N tile: time ~ columns walked by the busiest unit (full rounds x BLOCK_N + the tail strip), with the
// operand-traffic bonus of wide tiles (the main loop is bound by L2 -> SM operand delivery, ncu lts2xbar ~9.5 TB/s;
// bytes per FLOP fall with the tile width).  Measured on M=12800 (ncu, warm, whole tail tiles): N=3072 -> 55.1 /
// 50.9 / 52.2 us for BLOCK_N 128 / 192 / 256, N=2304 -> 43.3 / 41.3 / 41.9, N=768,K=3072 -> 51.0 / 46.1 / 56.7:
// per column walked 25.3 / 24.1 / 22.7 ns, i.e. the 256-wide tile is the fastest once its tail round is cut in strips.
static int choose_block_n(int M, int N, int units, int tile_m) {
  const int num_m = (M + tile_m - 1) / tile_m;
  const int cands[3] = {256, 192, 128};
  const double per_col[3] = {0.90, 0.955, 1.0};
  int best = 128;
  double best_cost = 1e30;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    const int tiles = num_m * ((N + bn - 1) / bn);
    const int sp = choose_tail_split(tiles, units, bn);
    const int tail = tiles % units;
    const double cols = static_cast<double>(tiles / units) * bn + (tail ? static_cast<double>(bn) / sp + 16.0 : 0.0);
    const double cost = cols * per_col[i];
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

int launch_gemm_bf16(const void* A, int lda, const void* W, int ldw, const GemmArgs& args, int epilogue,
                     cudaStream_t stream) {
  EVR_REQUIRE(args.M > 0 && args.N > 0 && args.K > 0, "GEMM: empty problem %d x %d x %d", args.M, args.N, args.K);
  EVR_REQUIRE(args.K % kGemmBlockK == 0, "GEMM: K=%d must be a multiple of %d", args.K, kGemmBlockK);
  EVR_REQUIRE(args.N % 32 == 0, "GEMM: N=%d must be a multiple of 32", args.N);
  EVR_REQUIRE(args.ldo % 8 == 0, "GEMM: ldo=%d must be a multiple of 8", args.ldo);
  const int sms = sm_count();
  // CTA pairs (256-row tiles) whenever there is more than one 128-row tile of work
  const int cg = (args.M > kGemmBlockM && sms % 2 == 0) ? 2 : 1;
  const int units = sms / cg, tile_m = kGemmBlockM * cg;
  int bn = choose_block_n(args.M, args.N, units, tile_m);
  if (const char* e = getenv("EVR_GEMM_BN")) {  // developer override for tile-shape experiments
    const int v = atoi(e);
    if (v == 128 || v == 192 || v == 256) bn = v;
  }
  const int num_m = (args.M + tile_m - 1) / tile_m;
  const int tiles = num_m * ((args.N + bn - 1) / bn);
  GemmArgs ga = args;
  ga.prof = prof_device_slot(PF_GEMM);
  ga.reverse_m = args.reverse_m;
  ga.tail_split = choose_tail_split(tiles, units, bn);
  ga.full_tiles = ga.tail_split > 1 ? tiles - tiles % units : tiles;
  const int items = ga.full_tiles + (tiles - ga.full_tiles) * ga.tail_split;
  CUtensorMap ta, tb, tbt;
  if (int r = make_tmap_2d_16b(&ta, A, args.M, args.K, static_cast<uint64_t>(lda) * 2, kGemmBlockM, true)) return r;
  if (int r = make_tmap_2d_16b(&tb, W, args.N, args.K, static_cast<uint64_t>(ldw) * 2, bn / cg, true)) return r;
  if (int r = make_tmap_2d_16b(&tbt, W, args.N, args.K, static_cast<uint64_t>(ldw) * 2, bn / ga.tail_split / cg, true))
    return r;
  const int grid = (items < units ? items : units) * cg;
  if (cg == 2) {
    switch (bn) {
      case 256: return launch_epi<256, 2>(ta, tb, tbt, ga, epilogue, grid, stream);
      case 192: return launch_epi<192, 2>(ta, tb, tbt, ga, epilogue, grid, stream);
      default: return launch_epi<128, 2>(ta, tb, tbt, ga, epilogue, grid, stream);
    }
  }
  switch (bn) {
    case 256: return launch_epi<256, 1>(ta, tb, tbt, ga, epilogue, grid, stream);
    case 192: return launch_epi<192, 1>(ta, tb, tbt, ga, epilogue, grid, stream);
    default: return launch_epi<128, 1>(ta, tb, tbt, ga, epilogue, grid, stream);
  }
}

}  // namespace evr

extern "C" int evr_gemm_bf16(const void* A, const void* W, const float* bias, void* out, int M, int N, int K,
                             int epilogue, void* stream) {
  EVR_REQUIRE(epilogue >= 0 && epilogue <= 3, "evr_gemm_bf16: epilogue %d not in 0..3", epilogue);
  evr::GemmArgs a{};
  a.M = M; a.N = N; a.K = K;
  a.bias = bias;
  a.out = out;
  a.ldo = N;
  return evr::launch_gemm_bf16(A, K, W, K, a, epilogue, static_cast<cudaStream_t>(stream));
}

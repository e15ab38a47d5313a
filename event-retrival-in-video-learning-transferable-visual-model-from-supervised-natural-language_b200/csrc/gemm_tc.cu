// Host side of the tcgen05 GEMM: tile-shape choice, tensor maps, launch; plus the exported
// building-block entry point evr_gemm_bf16 (include/evr.h) used by the parity tests.
#include "gemm_tc.cuh"

#include "../../include/evr.h"

namespace evr {

template <int BLOCK_N, int EPI>
static int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& args, int grid,
                      cudaStream_t stream) {
  using Cfg = GemmCfg<BLOCK_N, EPI>;
  auto kern = gemm_tc_kernel<BLOCK_N, EPI>;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    EVR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  ProfScope ps(PF_GEMM, stream);
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(ta, tb, args);
  EVR_CHECK_LAUNCH();
  return 0;
}

template <int BLOCK_N>
static int launch_epi(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& args, int epilogue, int grid,
                      cudaStream_t stream) {
  switch (epilogue) {
    case EPI_BF16: return launch_one<BLOCK_N, EPI_BF16>(ta, tb, args, grid, stream);
    case EPI_BF16_GELU: return launch_one<BLOCK_N, EPI_BF16_GELU>(ta, tb, args, grid, stream);
    case EPI_F32_RESID: return launch_one<BLOCK_N, EPI_F32_RESID>(ta, tb, args, grid, stream);
    case EPI_F32: return launch_one<BLOCK_N, EPI_F32>(ta, tb, args, grid, stream);
    case EPI_PATCH: return launch_one<BLOCK_N, EPI_PATCH>(ta, tb, args, grid, stream);
  }
  set_error("unknown GEMM epilogue %d", epilogue);
  return 2;
}

// Pick the N tile that wastes the fewest SM-rounds (wave quantisation over the persistent grid).
static int choose_block_n(int M, int N, int sms) {
  const int num_m = (M + kGemmBlockM - 1) / kGemmBlockM;
  const int cands[3] = {256, 192, 128};
  int best = 128;
  double best_cost = 1e30;
  for (int bn : cands) {
    const int tiles = num_m * ((N + bn - 1) / bn);
    const int rounds = (tiles + sms - 1) / sms;
    // time per tile ~ BLOCK_N; narrow tiles pay more smem bandwidth per MMA
    double cost = static_cast<double>(rounds) * bn * (bn == 128 ? 1.06 : 1.0);
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
  const int bn = choose_block_n(args.M, args.N, sms);
  CUtensorMap ta, tb;
  if (int r = make_tmap_2d_16b(&ta, A, args.M, args.K, static_cast<uint64_t>(lda) * 2, kGemmBlockM, true)) return r;
  if (int r = make_tmap_2d_16b(&tb, W, args.N, args.K, static_cast<uint64_t>(ldw) * 2, bn, true)) return r;
  const int num_m = (args.M + kGemmBlockM - 1) / kGemmBlockM;
  const int tiles = num_m * ((args.N + bn - 1) / bn);
  const int grid = tiles < sms ? tiles : sms;
  switch (bn) {
    case 256: return launch_epi<256>(ta, tb, args, epilogue, grid, stream);
    case 192: return launch_epi<192>(ta, tb, args, epilogue, grid, stream);
    default: return launch_epi<128>(ta, tb, args, epilogue, grid, stream);
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

// Query-side hot path: store preparation (L2-normalise -> fp16), the fused cosine scan + top-k,
// the final per-query select / cross-shard merge and the per-candidate re-score.
//
// Replaces (Backend/services/embedding_service.py):
//   :210       embeddings / np.linalg.norm(embeddings, axis=-1, keepdims=True)     -> store_prepare
//   :314 :365  similarities = np.dot(embeddings, q.T).flatten()                    -> topk_scan_kernel (MMA)
//   :317-320   np.argsort(similarities)[::-1][:top_k]                              -> topk_scan_kernel epilogue
//                                                                                     + topk_select_kernel
//   :274-279   np.dot(embeddings[index:index+1], q.T)[0][0]                        -> rescore_kernel
//
// Scan kernel (one persistent CTA per SM, 192 threads):
//   warp 0   TMA producer: the 128-query tile is made resident in smem (A operand); store rows stream
//            through a 6-stage ring of [128 rows x 64 dims] 128B-swizzled tiles (B operand).
//   warp 1   MMA issuer: tcgen05.mma kind::f16 (fp16 x fp16 -> fp32), D[128 queries x 128 rows] in TMEM,
//            4 accumulator stages.
//   warp 2-5 epilogue: TMEM lane = query, so every thread owns one query and sees 128 row scores per
//            tile in registers.  A running threshold (the current k-th best) filters them; survivors are
//            appended to a per-query candidate list in global memory (L2 resident).  When a list is
//            nearly full the warp prunes it cooperatively to the exact top-k (bitwise binary search for
//            the k-th key + stable ballot compaction) and raises the threshold.  The [Q, N] score matrix
//            never exists in memory.
#include <cuda_fp16.h>

#include <atomic>
#include <cstdlib>
#include <mutex>

#include "../../include/evr.h"
#include "common.cuh"
#include "ptx.cuh"

namespace evr {

// ============================================================================ store preparation
template <typename SrcT>
__global__ void __launch_bounds__(256)
store_prepare_kernel(const SrcT* __restrict__ src, int64_t N, int D, __half* __restrict__ dst) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * 8 + warp;
  if (row >= N) return;
  const SrcT* x = src + row * D;
  float sq = 0.f;
  for (int i = lane * 2; i < D; i += 64) {
    float a, b;
    if constexpr (sizeof(SrcT) == 4) {
      const float2 t = *reinterpret_cast<const float2*>(x + i);
      a = t.x; b = t.y;
    } else {
      const float2 t = __half22float2(*reinterpret_cast<const __half2*>(x + i));
      a = t.x; b = t.y;
    }
    sq += a * a + b * b;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  // zero rows stay zero (the reference would emit NaN: 0/0; compare_models.py:1168-1171 guards norm > 1e-8)
  const float inv = sq > 0.f ? 1.0f / sqrtf(sq) : 0.f;
  for (int i = lane * 2; i < D; i += 64) {
    float a, b;
    if constexpr (sizeof(SrcT) == 4) {
      const float2 t = *reinterpret_cast<const float2*>(x + i);
      a = t.x; b = t.y;
    } else {
      const float2 t = __half22float2(*reinterpret_cast<const __half2*>(x + i));
      a = t.x; b = t.y;
    }
    *reinterpret_cast<__half2*>(dst + row * D + i) = __floats2half2_rn(a * inv, b * inv);
  }
}

// ============================================================================ key mapping
__device__ __forceinline__ uint32_t float_to_key(uint32_t bits) {  // monotone: larger float <=> larger key
  return bits ^ (static_cast<uint32_t>(static_cast<int32_t>(bits) >> 31) | 0x80000000u);
}
__device__ __forceinline__ uint32_t key_to_float_bits(uint32_t key) {
  return key ^ ((key >> 31) ? 0x80000000u : 0xFFFFFFFFu);
}
constexpr uint32_t kNegInfBits = 0xFF800000u;

__device__ __forceinline__ float fmax3(float a, float b, float c) {  // one FMNMX3 on sm_100
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// ============================================================================ fused scan + top-k
constexpr int kScanQTile = 128;     // queries per tile  (UMMA M, TMEM lanes)
constexpr int kScanRowTile = 128;   // store rows per tile (UMMA N, TMEM columns per stage)
constexpr int kScanStages = 6;      // smem ring depth, 16 KB per stage
constexpr int kScanAccStages = 4;   // TMEM accumulator stages (4 x 128 = 512 columns)
constexpr int kScanThreads = 192;
constexpr int kScanStageBytes = kScanRowTile * 64 * 2;
constexpr int kMaxFusedK = 1024;   // candidate lists of 32*EPL entries must hold k + one tile of admissions

struct ScanArgs {
  int64_t N;           // store rows scanned by this launch
  int D, Q, k;
  int num_qt;          // query tiles
  int num_slices;      // store slices
  int tiles_per_slice; // 128-row tiles per slice
  int q_pad;           // num_qt * 128
  int cap;             // candidate-list capacity per (slice, query) = 32 * EPL
  const float* init_thresh;  // per-query admission threshold from the sample pass (nullptr: -inf)
  int ring;            // candidate lists per query: slice s appends to list s % ring (ring == num_slices: one each)
  int2* lists;         // [ring][q_pad][cap]  (score bits, local row): candidates of one (list, query)
  int* counts;         // [ring][q_pad]       entries used in each list
  float* list_thresh;  // [ring][q_pad]       admission threshold a list's last item ended with
  unsigned* done;      // [q_pad / 32][num_slices]  == epoch once the warp owning those 32 queries has finished the slice
  unsigned epoch;      // unique per call (the workspace is reused)
  float* group_max;    // sample pass only (else nullptr): [q_pad][groups] max score of every 64-row group
  int groups;          // 64-row groups in the sample = N / 64
  int stages;          // depth of the store-row smem ring actually used (<= kScanStages; 2 at D = 768)
  int row_evict_last;  // A/B knob (EVR_SCAN_EVICT_LAST=1): store rows stay in L2 as long as possible
  int tile_stride;     // sample pass: tile t of the launch reads store tile t * tile_stride (1 in the full scan)
  int* sched_counter;  // dynamic item distribution (nullptr: static round-robin): next unclaimed item, zeroed per call
  unsigned long long* sched_tickets;  // [units][kSchedDepth]: (k + 1) << 32 | item claimed for the unit's k-th step
  int qt_group;        // query tiles that walk the store together (a pass); num_qt = one pass (items slice-major)
  int rotate;          // 1: the query tiles of a pass start a slice at different tiles (and wrap around)
};

// item -> (query tile, slice).  The query tiles are taken `qt_group` at a time; inside such a pass items are
// slice-major, so the tiles of a pass read a slice together and the passes walk the store one after the other.
struct ScanItem { int qt, slice, lane_in_group, group_size; };
__device__ __forceinline__ ScanItem scan_item(const ScanArgs& a, int item) {
  ScanItem r;
  const int G = a.qt_group;
  const int per_pass = G * a.num_slices;
  int pass = item / per_pass;
  const int last = (a.num_qt - 1) / G;
  if (pass > last) pass = last;
  const int rem = item - pass * per_pass;
  r.group_size = min(G, a.num_qt - pass * G);
  r.slice = rem / r.group_size;
  r.lane_in_group = rem - r.slice * r.group_size;
  r.qt = pass * G + r.lane_in_group;
  return r;
}
// Dynamic item distribution.  With a static round-robin the units' positions drift apart for good (SMs differ by a
// per cent or so in speed: 170 slices over a 100 M-row scan), fast units then stand at the ring flags of slow ones and
// the scan loses 8-10 %; a deeper ring only moves the point where that starts.  Here the producer thread of a unit's
// leader CTA claims the next item with an atomic add -- one step ahead, so the round trip is off the critical path --
// and publishes it as a (step, item) ticket in global memory that the unit's other roles (peer producer, MMA issuer,
// epilogue warps) read when they get to that step.  Items are claimed in increasing order, so every unit stays within
// a round of the others, the slices being read stay the same few for all units, and the ring stays shallow.
constexpr int kSchedDepth = 16;  // tickets per unit.  A slot is reused 16 steps later; the producer is at most 6 shared-memory
                                 // stages + 4 accumulators = 10 row tiles = 5 two-tile items (D = 64) ahead of the epilogue warps
__device__ __forceinline__ int sched_claim(const ScanArgs& a, int unit, int k, int num_items) {
  int item = atomicAdd(a.sched_counter, 1);
  if (item >= num_items) item = -1;
  const unsigned long long word = (static_cast<unsigned long long>(k + 1) << 32) | static_cast<unsigned>(item);
  unsigned long long* t = a.sched_tickets + static_cast<size_t>(unit) * kSchedDepth + (k & (kSchedDepth - 1));
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(t), "l"(word) : "memory");
  return item;
}
__device__ __forceinline__ int sched_read(const ScanArgs& a, int unit, int k) {
  const unsigned long long* t = a.sched_tickets + static_cast<size_t>(unit) * kSchedDepth + (k & (kSchedDepth - 1));
  unsigned long long w;
  unsigned spins = 0;
  while (true) {
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(t) : "memory");
    if (static_cast<unsigned>(w >> 32) == static_cast<unsigned>(k + 1)) break;
    __nanosleep(100);
    if (++spins > (1u << 25)) asm volatile("trap;");   // seconds: a protocol bug traps instead of hanging
  }
  return static_cast<int>(static_cast<unsigned>(w));
}
// step k of a unit's walk -> item (or -1: done)
__device__ __forceinline__ int sched_item(const ScanArgs& a, int unit, int num_units, int k, int num_items) {
  if (a.sched_counter != nullptr) return sched_read(a, unit, k);
  const long long it = unit + static_cast<long long>(k) * num_units;
  return it < num_items ? static_cast<int>(it) : -1;
}

// tile visited at step i of an item's walk over [t0, t1)
__device__ __forceinline__ int64_t scan_tile(const ScanArgs& a, const ScanItem& it, int64_t t0, int64_t t1, int64_t i) {
  if (!a.rotate) return t0 + i;
  const int64_t n = t1 - t0;
  const int64_t rot = (static_cast<int64_t>(it.lane_in_group) * n) / it.group_size;
  int64_t j = i + rot;
  if (j >= n) j -= n;
  return t0 + j;
}

// Warp-cooperative exact prune of one query's candidate list to its top-k.  All 32 lanes call with the
// same arguments.  Returns the key of the k-th best entry (new admission threshold).
template <int EPL>
__device__ __forceinline__ uint32_t warp_prune(int2* buf, int cnt, int k, int lane) {
  uint32_t key[EPL];
#pragma unroll
  for (int i = 0; i < EPL; ++i) {
    const int e = lane + 32 * i;
    key[i] = e < cnt ? float_to_key(static_cast<uint32_t>(__ldcg(buf + e).x)) : 0u;
  }
  uint32_t T = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t c = T | (1u << bit);
    int n = 0;
#pragma unroll
    for (int i = 0; i < EPL; ++i) n += key[i] >= c;
    n = __reduce_add_sync(0xffffffffu, n);
    if (n >= k) T = c;
  }
  int n_gt = 0;
#pragma unroll
  for (int i = 0; i < EPL; ++i) n_gt += key[i] > T;
  n_gt = __reduce_add_sync(0xffffffffu, n_gt);
  const int need_eq = k - n_gt;
  const uint32_t lt_mask = (1u << lane) - 1u;
  int base = 0, eq_seen = 0;
#pragma unroll
  for (int i = 0; i < EPL; ++i) {
    const int e = lane + 32 * i;
    int2 ent = make_int2(0, 0);
    if (e < cnt) ent = __ldcg(buf + e);
    __syncwarp();  // every lane's load of this round lands before any lane's store below
    const bool gt = key[i] > T, eq = key[i] == T;
    const uint32_t m_eq = __ballot_sync(0xffffffffu, eq);
    const bool take = gt || (eq && (eq_seen + __popc(m_eq & lt_mask) < need_eq));
    const uint32_t m = __ballot_sync(0xffffffffu, take);  // also orders this round's loads before its stores
    if (take) __stcg(buf + base + __popc(m & lt_mask), ent);
    base += __popc(m);
    eq_seen += __popc(m_eq);
  }
  __syncwarp();
  return T;
}

// CG = 1: one CTA per SM, 128 queries x 128-row tiles.
// CG = 2: CTA pairs (cta_group::2): 256 queries (128 per CTA) x 256-row tiles; each CTA streams only its 128 rows of
//         every tile, so shared-memory and L2 traffic per FLOP are halved.  The leader (cluster rank 0) issues the
//         MMAs and owns the TMA-arrival barriers; commits are multicast to both CTAs.
template <int EPL, int CG>
__global__ void __launch_bounds__(kScanThreads, 1)
topk_scan_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_s,
                 const ScanArgs args) {
  constexpr int kRowTile = kScanRowTile * CG;    // store rows per tile (UMMA N)
  constexpr int kAcc = kScanAccStages / CG;      // TMEM accumulator stages (512 columns in all)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int num_kb = args.D / 64;
  const int q_bytes = kScanQTile * args.D * 2;
  uint8_t* smem_q = smem;                 // [num_kb][128 x 64] swizzled, resident
  uint8_t* smem_s = smem + q_bytes;       // [kScanStages][128 x 64]
  const int nst = args.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_s + nst * kScanStageBytes);
  uint64_t* full_bar = bars;                                         // [kScanStages]
  uint64_t* empty_bar = bars + kScanStages;                          // [kScanStages]
  uint64_t* acc_full = bars + 2 * kScanStages;                       // [kAcc] (room for 4)
  uint64_t* acc_empty = bars + 2 * kScanStages + kScanAccStages;     // [kAcc]
  uint64_t* q_full = bars + 2 * kScanStages + 2 * kScanAccStages;    // [1]
  uint64_t* q_empty = q_full + 1;                                    // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_items = args.num_qt * args.num_slices;
  const int64_t total_tiles = (args.N + kRowTile - 1) / kRowTile;
  uint32_t rank = 0;
  if constexpr (CG == 2) rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int unit = blockIdx.x / CG, num_units = gridDim.x / CG;  // a unit = one CTA or one CTA pair
  if constexpr (CG == 2) cluster_sync_all();  // both CTAs resident before the pair-wide TMEM allocation

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_s);
  }
  if (warp == 1 && lane == 0) {
    // full / q_full / acc_empty collect one arrival per CTA of the unit (the peer arrives remotely on the leader's)
    for (int i = 0; i < nst; ++i) { mbar_init(&full_bar[i], CG); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < kAcc; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4 * CG); }
    mbar_init(q_full, CG);
    mbar_init(q_empty, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (CG == 2) tmem_alloc_pair<512>(tmem_slot);
    else tmem_alloc<512>(tmem_slot);
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      // ===== TMA producer =====
      int stage = 0;
      uint32_t phase = 0, qphase = 0;
      uint64_t row_policy = args.num_qt > 1 ? kEvictNormal : kEvictFirst;
      if (args.row_evict_last) row_policy = kEvictLast;
      const bool claims = args.sched_counter != nullptr && leader;   // this thread distributes the unit's items
      int next_item = claims ? sched_claim(args, unit, 0, num_items) : 0;
      for (int k = 0;; ++k) {
        int item;
        if (claims) {
          item = next_item;
          if (item >= 0) next_item = sched_claim(args, unit, k + 1, num_items);   // one step ahead
        } else {
          item = sched_item(args, unit, num_units, k, num_items);
        }
        if (item < 0) break;
        const ScanItem si = scan_item(args, item);
        const int qt = si.qt, slice = si.slice;
        mbar_wait(q_empty, qphase ^ 1);  // MMAs of the previous item no longer read the query tile
        const int q_row = (qt * CG + static_cast<int>(rank)) * kScanQTile;
        if constexpr (CG == 2) {
          for (int kb = 0; kb < num_kb; ++kb)
            tma_load_2d_pair(smem_q + kb * kScanStageBytes, &tmap_q, q_full, kb * 64, q_row, kEvictLast);
          if (leader) mbar_expect_tx(q_full, q_bytes * 2);
          else mbar_arrive_cluster(q_full, 0);
        } else {
          mbar_expect_tx(q_full, q_bytes);
          for (int kb = 0; kb < num_kb; ++kb)
            tma_load_2d(smem_q + kb * kScanStageBytes, &tmap_q, q_full, kb * 64, q_row, kEvictLast);
        }
        qphase ^= 1;
        const int64_t t0 = static_cast<int64_t>(slice) * args.tiles_per_slice;
        int64_t t1 = t0 + args.tiles_per_slice;
        if (t1 > total_tiles) t1 = total_tiles;
        for (int64_t ti = 0; ti < t1 - t0; ++ti) {
          const int64_t t = scan_tile(args, si, t0, t1, ti);
          const int row = static_cast<int>(t * args.tile_stride * kRowTile) +
                          static_cast<int>(rank) * kScanRowTile;  // this CTA's half of the (strided) tile
          for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if constexpr (CG == 2) {
              tma_load_2d_pair(smem_s + stage * kScanStageBytes, &tmap_s, &full_bar[stage], kb * 64, row, row_policy);
              if (leader) mbar_expect_tx(&full_bar[stage], kScanStageBytes * 2);
              else mbar_arrive_cluster(&full_bar[stage], 0);
            } else {
              mbar_expect_tx(&full_bar[stage], kScanStageBytes);
              tma_load_2d(smem_s + stage * kScanStageBytes, &tmap_s, &full_bar[stage], kb * 64, row, row_policy);
            }
            if (++stage == nst) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && elect_one()) {
      // ===== MMA issuer (leader CTA only) =====
      constexpr uint32_t idesc = umma_idesc_f16(/*fp16*/ 0, kScanQTile * CG, kRowTile);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0, qphase = 0;
      const uint32_t q_addr = smem_u32(smem_q);
      for (int k = 0;; ++k) {
        const int item = sched_item(args, unit, num_units, k, num_items);
        if (item < 0) break;
        const int slice = scan_item(args, item).slice;
        mbar_wait(q_full, qphase);
        qphase ^= 1;
        tc_fence_after();
        const int64_t t0 = static_cast<int64_t>(slice) * args.tiles_per_slice;
        int64_t t1 = t0 + args.tiles_per_slice;
        if (t1 > total_tiles) t1 = total_tiles;
        for (int64_t t = t0; t < t1; ++t) {
          mbar_wait(&acc_empty[as], aphase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * kRowTile;
          for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint64_t a_desc = umma_smem_desc_sw128(q_addr + kb * kScanStageBytes);
            const uint64_t b_desc = umma_smem_desc_sw128(smem_u32(smem_s + stage * kScanStageBytes));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if constexpr (CG == 2) umma_f16_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
              else umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
            }
            if constexpr (CG == 2) umma_commit_pair(&empty_bar[stage]);
            else umma_commit(&empty_bar[stage]);
            if (++stage == nst) { stage = 0; phase ^= 1; }
          }
          if constexpr (CG == 2) umma_commit_pair(&acc_full[as]);
          else umma_commit(&acc_full[as]);
          if (++as == kAcc) { as = 0; aphase ^= 1; }
        }
        if constexpr (CG == 2) umma_commit_pair(q_empty);
        else umma_commit(q_empty);
      }
    }
  } else {
    // ===== epilogue / selection: warps 2..5, TMEM lane quadrant = warp % 4 =====
    const int quad = warp & 3;
    const int k = args.k, cap = args.cap;
    const int trigger = cap - kRowTile;
    int as = 0;
    uint32_t aphase = 0;
    for (int step = 0;; ++step) {
      int item = 0;
      if (lane == 0) item = sched_item(args, unit, num_units, step, num_items);
      item = __shfl_sync(0xffffffffu, item, 0);
      if (item < 0) break;
      const ScanItem si = scan_item(args, item);
      const int qt = si.qt, slice = si.slice;
      const int q0 = (qt * CG + static_cast<int>(rank)) * kScanQTile + quad * 32;  // first query of this warp
      const bool q_valid = q0 + lane < args.Q;
      // the candidate list of (slice, query) is filled in place; the final select kernel reads it with its count
      const int ring_slot = slice % args.ring;
      int2* const cand_warp = args.lists + (static_cast<size_t>(ring_slot) * args.q_pad + q0) * cap;
      int2* const buf = cand_warp + static_cast<size_t>(lane) * cap;
      // Admission is `score >= thresh`.  Any lower bound on the query's global k-th best score is a valid
      // threshold (at least k rows reach it): the sample pre-pass supplies one (init_thresh), and every local prune
      // raises it to the k-th best of the rows this item has seen.  NaN never admits (padding lanes).
      float thresh = q_valid ? (args.init_thresh ? __ldg(args.init_thresh + q0 + lane) : -INFINITY)
                             : __int_as_float(0x7fc00000);
      int cnt = 0;
      if (args.group_max == nullptr && slice >= args.ring) {
        // This list already holds the candidates of slice - ring (same queries, scanned by some other unit a few
        // slices back): wait until that warp has published them, then go on from its count and its threshold.  Items
        // run in slice-major order on resident CTAs, so the item waited for never depends on this one.
        if (lane == 0) {
          const volatile unsigned* f = args.done + static_cast<size_t>(q0 >> 5) * args.num_slices + (slice - args.ring);
          while (*f != args.epoch) __nanosleep(200);
        }
        __syncwarp();
        __threadfence();
        if (q_valid) {
          cnt = __ldcg(args.counts + static_cast<size_t>(ring_slot) * args.q_pad + q0 + lane);
          thresh = fmaxf(thresh, __ldcg(args.list_thresh + static_cast<size_t>(ring_slot) * args.q_pad + q0 + lane));
        }
      }
      const int64_t t0 = static_cast<int64_t>(slice) * args.tiles_per_slice;
      int64_t t1 = t0 + args.tiles_per_slice;
      if (t1 > total_tiles) t1 = total_tiles;
      for (int64_t ti = 0; ti < t1 - t0; ++ti) {
        const int64_t t = scan_tile(args, si, t0, t1, ti);
        mbar_wait(&acc_full[as], aphase);
        tc_fence_after();
        const int row_base = static_cast<int>(t * kRowTile);
        const int rows_left = static_cast<int>(args.N - row_base);  // > 0
#pragma unroll 1
        for (int c = 0; c < kRowTile / 64; ++c) {
          uint32_t v[2][32];
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * kRowTile + c * 64;
          tmem_ld_32x32(taddr, v[0]);
          tmem_ld_32x32(taddr + 32, v[1]);
          tmem_ld_wait();
          float gmx = -INFINITY;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float m4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              m4[u] = fmax3(__uint_as_float(v[h][8 * u]), __uint_as_float(v[h][8 * u + 1]), __uint_as_float(v[h][8 * u + 2]));
              m4[u] = fmax3(m4[u], __uint_as_float(v[h][8 * u + 3]), __uint_as_float(v[h][8 * u + 4]));
              m4[u] = fmax3(m4[u], __uint_as_float(v[h][8 * u + 5]), __uint_as_float(v[h][8 * u + 6]));
              m4[u] = fmaxf(m4[u], __uint_as_float(v[h][8 * u + 7]));
            }
            const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
            gmx = fmaxf(gmx, mx);
            if (args.group_max == nullptr && mx >= thresh) {  // rare once the threshold has settled
              // only the 8-column groups whose own maximum passes are walked (the warp pays for a group when any
              // of its 32 queries admits from it: ~4x fewer predicated stores than walking all 32 columns)
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                if (m4[u] >= thresh) {
#pragma unroll
                  for (int j = 8 * u; j < 8 * u + 8; ++j) {
                    const int r = c * 64 + h * 32 + j;
                    if (__uint_as_float(v[h][j]) >= thresh && r < rows_left) {
                      __stcg(buf + cnt, make_int2(static_cast<int>(v[h][j]), row_base + r));
                      ++cnt;
                    }
                  }
                }
              }
            }
          }
          if (args.group_max != nullptr && q_valid)  // sample pass: N is a multiple of the tile, no padding rows
            args.group_max[static_cast<size_t>(q0 + lane) * args.groups + (t * (kRowTile / 64) + c)] = gmx;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {  // release TMEM before any pruning (the leader's barrier gates the MMA issuer)
          if (leader) mbar_arrive(&acc_empty[as]);
          else mbar_arrive_cluster(&acc_empty[as], 0);
        }
        if (++as == kAcc) { as = 0; aphase ^= 1; }

        uint32_t need = __ballot_sync(0xffffffffu, cnt > trigger);
        while (need) {
          const int src = __ffs(need) - 1;
          need &= need - 1;
          const int c_src = __shfl_sync(0xffffffffu, cnt, src);
          const uint32_t T = warp_prune<EPL>(cand_warp + static_cast<size_t>(src) * cap, c_src, k, lane);
          if (lane == src) {
            cnt = k;
            thresh = __uint_as_float(key_to_float_bits(T));
          }
        }
      }
      // ---- item done: the list stays where it is; publish its length
      if (args.counts != nullptr) {
        if (q_valid) {
          __stcg(args.counts + static_cast<size_t>(ring_slot) * args.q_pad + q0 + lane, cnt);
          __stcg(args.list_thresh + static_cast<size_t>(ring_slot) * args.q_pad + q0 + lane, thresh);
        }
        __threadfence();   // list entries, count and threshold before the flag
        __syncwarp();
        if (lane == 0) {
          volatile unsigned* f = args.done + static_cast<size_t>(q0 >> 5) * args.num_slices + slice;
          *f = args.epoch;
        }
      }
    }
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all();  // the peer's remote arrivals and MMA reads of our smem are done
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (CG == 2) tmem_dealloc_pair<512>(tmem_base);
    else tmem_dealloc<512>(tmem_base);
  }
}

// ============================================================================ sample threshold
// k-th largest of a query's 64-row group maxima: k distinct rows score at least that much, so it is a lower bound
// on the query's global k-th best score (the admission threshold of the full scan).  One warp per query.
__global__ void __launch_bounds__(256)
sample_kth_kernel(const float* __restrict__ group_max, int groups, int Q, int k, float* __restrict__ thresh) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * 8 + warp;
  if (q >= Q) return;
  const float* g = group_max + static_cast<size_t>(q) * groups;
  uint32_t T = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t c = T | (1u << bit);
    int n = 0;
    for (int i = lane; i < groups; i += 32) n += float_to_key(__float_as_uint(g[i])) >= c;  // L1-resident after pass 1
    n = __reduce_add_sync(0xffffffffu, n);
    if (n >= k) T = c;
  }
  if (lane == 0) thresh[q] = __uint_as_float(key_to_float_bits(T));
}

// ============================================================================ final select
// One CTA per query: exact top-k of the candidates its slices admitted (lists[slice][q][0..counts)), sorted by
// (score desc, row asc).  Candidates are gathered once into shared memory as 64-bit keys
// (score key << 32 | ~row); an 8-pass MSB radix select finds the k-th key, the k winners are bitonic-sorted.
// If a query has more candidates than fit (adversarial ties), the passes re-read the lists from global memory.
struct SelectArgs {
  const int2* lists;
  const int* counts;
  int P, q_pad, cap, k, smem_cap;
  int64_t row_offset;
  float* out_scores;
  int64_t* out_idx;
};

__global__ void __launch_bounds__(256)
topk_select_kernel(const SelectArgs a) {
  extern __shared__ __align__(16) uint8_t ssm[];
  __shared__ unsigned hist[256];
  __shared__ int s_off[257];
  __shared__ unsigned long long s_prefix;
  __shared__ int s_remaining, s_count, s_done;
  const int q = blockIdx.x, tid = threadIdx.x;
  int kp = 1;
  while (kp < a.k) kp <<= 1;
  unsigned long long* win = reinterpret_cast<unsigned long long*>(ssm);  // [kp] winners
  unsigned long long* keys = win + kp;                                   // [smem_cap] gathered candidates

  // slice offsets (P <= 256)
  if (tid < a.P) s_off[tid + 1] = a.counts[static_cast<size_t>(tid) * a.q_pad + q];
  if (tid == 0) s_off[0] = 0;
  __syncthreads();
  if (tid == 0) {
    int acc = 0;
    for (int i = 1; i <= a.P; ++i) { acc += s_off[i]; s_off[i] = acc; }
    s_prefix = 0ull; s_remaining = a.k; s_count = 0; s_done = 0;
  }
  __syncthreads();
  const int total = s_off[a.P];
  const bool in_smem = total <= a.smem_cap;
  auto make_key = [](int2 e) -> unsigned long long {
    return (static_cast<unsigned long long>(float_to_key(static_cast<uint32_t>(e.x))) << 32) |
           static_cast<uint32_t>(~static_cast<uint32_t>(e.y));
  };
  const int warp = tid >> 5, lane = tid & 31;
  if (in_smem) {
    for (int sl = warp; sl < a.P; sl += 8) {
      const int o = s_off[sl], c = s_off[sl + 1] - o;
      const int2* src = a.lists + (static_cast<size_t>(sl) * a.q_pad + q) * a.cap;
      for (int j = lane; j < c; j += 32) keys[o + j] = make_key(__ldcg(src + j));
    }
  }
  __syncthreads();
  // visit every candidate key
  auto for_each = [&](auto&& f) {
    if (in_smem) {
      for (int i = tid; i < total; i += 256) f(keys[i]);
    } else {
      for (int sl = warp; sl < a.P; sl += 8) {
        const int c = s_off[sl + 1] - s_off[sl];
        const int2* src = a.lists + (static_cast<size_t>(sl) * a.q_pad + q) * a.cap;
        for (int j = lane; j < c; j += 32) f(make_key(__ldcg(src + j)));
      }
    }
  };
  for (int i = tid; i < kp; i += 256) win[i] = 0ull;  // key 0 sorts last: score -inf / row -1 padding
  unsigned long long T = 0ull;
  if (total > a.k) {
    unsigned long long himask = 0ull;
    for (int shift = 56; shift >= 0; shift -= 8) {
      hist[tid] = 0;
      __syncthreads();
      const unsigned long long prefix = s_prefix;
      for_each([&](unsigned long long key) {
        if ((key & himask) == prefix) atomicAdd(&hist[(key >> shift) & 255ull], 1u);
      });
      __syncthreads();
      if (warp == 0) {
        // descending scan of the 256 bins: lane l owns bins 255-8l .. 248-8l
        const int rem = s_remaining;
        int loc[8], sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { loc[j] = static_cast<int>(hist[255 - 8 * lane - j]); sum += loc[j]; }
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const int excl = incl - sum;
        const bool mine = excl < rem && incl >= rem;  // the k-th key falls into one of this lane's bins
        if (mine) {
          int acc = excl, j = 0;
          for (; j < 7; ++j) {
            if (acc + loc[j] >= rem) break;
            acc += loc[j];
          }
          s_remaining = rem - acc;
          s_prefix = prefix | (static_cast<unsigned long long>(255 - 8 * lane - j) << shift);
          s_done = (loc[j] == rem - acc);  // every key of that bin is a winner: lower digits need no pass
        }
      }
      himask |= 255ull << shift;
      __syncthreads();
      if (s_done) break;
    }
    T = s_prefix;  // winners are exactly the keys >= T (keys are unique: the row id is part of the key)
  }
  __syncthreads();
  for_each([&](unsigned long long key) {
    if (key >= T) {
      const int pos = atomicAdd(&s_count, 1);
      if (pos < kp) win[pos] = key;
    }
  });
  __syncthreads();
  for (int size = 2; size <= kp; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < kp; i += 256) {
        const int j = i ^ stride;
        if (j > i) {
          const bool desc = (i & size) == 0;
          const unsigned long long x = win[i], y = win[j];
          if ((x < y) == desc) { win[i] = y; win[j] = x; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = tid; i < a.k; i += 256) {
    const unsigned long long key = win[i];
    const bool pad = key == 0ull;
    a.out_scores[static_cast<size_t>(q) * a.k + i] =
        pad ? -INFINITY : __uint_as_float(key_to_float_bits(static_cast<uint32_t>(key >> 32)));
    a.out_idx[static_cast<size_t>(q) * a.k + i] =
        pad ? -1 : static_cast<int64_t>(~static_cast<uint32_t>(key)) + a.row_offset;
  }
}

// ============================================================================ cross-shard merge
// One CTA per query: exact top-k of n = P*k candidates (P partial lists of k), sorted by
// (score desc, index asc).  8-pass radix select on the 64-bit key (score key << 32 | ~idx),
// then a bitonic sort of the k survivors in shared memory.
template <typename IdxT>
__global__ void __launch_bounds__(256)
topk_final_kernel(const float* __restrict__ scores, const IdxT* __restrict__ idx, int P, int q_stride, int k,
                  int64_t row_offset, float* __restrict__ out_scores, int64_t* __restrict__ out_idx) {
  extern __shared__ __align__(16) uint8_t fsm[];
  __shared__ unsigned hist[256];
  __shared__ unsigned long long s_prefix;
  __shared__ int s_remaining, s_count;
  const int q = blockIdx.x;
  const int n = P * k;
  int kp = 1;
  while (kp < k) kp <<= 1;
  unsigned long long* skey = reinterpret_cast<unsigned long long*>(fsm);  // [kp]
  long long* sidx = reinterpret_cast<long long*>(skey + kp);              // [kp]

  auto load_key = [&](int e, long long& id) -> unsigned long long {
    const int p = e / k, j = e - p * k;
    const size_t off = (static_cast<size_t>(p) * q_stride + q) * k + j;
    id = static_cast<long long>(idx[off]);
    const uint32_t sk = float_to_key(__float_as_uint(scores[off]));
    return (static_cast<unsigned long long>(sk) << 32) | static_cast<uint32_t>(~static_cast<uint32_t>(id));
  };

  if (threadIdx.x == 0) { s_prefix = 0ull; s_remaining = k; s_count = 0; }
  __syncthreads();
  unsigned long long himask = 0ull;
  for (int shift = 56; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const unsigned long long prefix = s_prefix;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
      long long id;
      const unsigned long long key = load_key(e, id);
      if ((key & himask) == prefix) atomicAdd(&hist[(key >> shift) & 255ull], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int rem = s_remaining, acc = 0, b = 255;
      for (; b > 0; --b) {
        if (acc + static_cast<int>(hist[b]) >= rem) break;
        acc += hist[b];
      }
      s_remaining = rem - acc;
      s_prefix = prefix | (static_cast<unsigned long long>(b) << shift);
    }
    himask |= 255ull << shift;
    __syncthreads();
  }
  const unsigned long long T = s_prefix;  // k-th largest 64-bit key
  for (int i = threadIdx.x; i < kp; i += blockDim.x) { skey[i] = 0ull; sidx[i] = -1; }
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    long long id;
    const unsigned long long key = load_key(e, id);
    if (key >= T) {
      const int pos = atomicAdd(&s_count, 1);
      if (pos < k) { skey[pos] = key; sidx[pos] = id; }
    }
  }
  __syncthreads();
  // bitonic sort, descending by key
  for (int size = 2; size <= kp; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < kp; i += blockDim.x) {
        const int j = i ^ stride;
        if (j > i) {
          const bool desc = (i & size) == 0;
          const unsigned long long a = skey[i], b = skey[j];
          if ((a < b) == desc) {
            skey[i] = b; skey[j] = a;
            const long long t = sidx[i]; sidx[i] = sidx[j]; sidx[j] = t;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    const long long id = sidx[i];
    out_scores[static_cast<size_t>(q) * k + i] = __uint_as_float(key_to_float_bits(static_cast<uint32_t>(skey[i] >> 32)));
    out_idx[static_cast<size_t>(q) * k + i] = id < 0 ? -1 : id + row_offset;
  }
}

// ============================================================================ large k (k > kMaxFusedK)
// The reference accepts any top_k (embedding_service.py:317-320 sorts all N scores).  Requests beyond the fused
// kernel's list capacity take this two-step path: (1) all scores of a few queries are written out once,
// `scores_rows_kernel`: a warp per store row, the row read once and dotted with up to 8 resident queries (HBM-bound,
// N*D*2 bytes per chunk of 8 queries + N*4*QC written); (2) `topk_large_kernel`, one CTA per query: 8-pass radix
// select of the k-th 64-bit key (score key << 32 | ~row: unique, ties broken by ascending row) straight over global
// memory, compaction of the k winners, bitonic sort of them in a global scratch area.  Also the merge of gathered
// per-shard lists when k > 1024 (element e of query q lives at ((e / kin) * q_stride + q) * kin + e % kin).
constexpr int kLargeQChunk = 8;

__global__ void __launch_bounds__(256)
scores_rows_kernel(const __half* __restrict__ store, int64_t N, int D, const __half* __restrict__ queries, int QC,
                   float* __restrict__ out /* [QC][N] */) {
  extern __shared__ __align__(16) uint8_t qsm[];
  __half2* q2 = reinterpret_cast<__half2*>(qsm);  // [QC][D/2]
  const int half_d = D / 2;
  for (int i = threadIdx.x; i < QC * half_d; i += blockDim.x)
    q2[i] = reinterpret_cast<const __half2*>(queries)[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * 8 + warp; row < N; row += static_cast<int64_t>(gridDim.x) * 8) {
    const __half2* x = reinterpret_cast<const __half2*>(store + row * D);
    float acc[kLargeQChunk];
#pragma unroll
    for (int j = 0; j < kLargeQChunk; ++j) acc[j] = 0.f;
    for (int i = lane; i < half_d; i += 32) {
      const float2 a = __half22float2(x[i]);
#pragma unroll
      for (int j = 0; j < kLargeQChunk; ++j) {
        if (j < QC) {
          const float2 b = __half22float2(q2[j * half_d + i]);
          acc[j] += a.x * b.x + a.y * b.y;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kLargeQChunk; ++j) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
    }
    if (lane == 0) {
      for (int j = 0; j < QC; ++j) out[static_cast<size_t>(j) * N + row] = acc[j];
    }
  }
}

__global__ void __launch_bounds__(1024)
topk_large_kernel(const float* __restrict__ scores, const int64_t* __restrict__ idx /* nullptr: idx = position */,
                  int P, int q_stride, int64_t kin, int64_t kout, int64_t row_offset,
                  unsigned long long* __restrict__ scratch /* [gridDim.x][kp] */, int64_t kp,
                  float* __restrict__ out_scores, int64_t* __restrict__ out_idx) {
  __shared__ unsigned hist[256];
  __shared__ unsigned long long s_prefix;
  __shared__ long long s_remaining;
  __shared__ unsigned long long s_count;
  const int q = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
  const int64_t n = static_cast<int64_t>(P) * kin;
  unsigned long long* win = scratch + static_cast<size_t>(q) * kp;
  auto key_at = [&](int64_t e) -> unsigned long long {
    const int64_t p = e / kin, j = e - p * kin;
    const size_t off = (static_cast<size_t>(p) * q_stride + q) * kin + j;
    const uint32_t id = idx ? static_cast<uint32_t>(idx[off]) : static_cast<uint32_t>(j);
    return (static_cast<unsigned long long>(float_to_key(__float_as_uint(scores[off]))) << 32) |
           static_cast<uint32_t>(~id);
  };
  if (tid == 0) { s_prefix = 0ull; s_remaining = kout; s_count = 0ull; }
  __syncthreads();
  unsigned long long T = 0ull;
  if (n > kout) {
    unsigned long long himask = 0ull;
    for (int shift = 56; shift >= 0; shift -= 8) {
      if (tid < 256) hist[tid] = 0;
      __syncthreads();
      const unsigned long long prefix = s_prefix;
      for (int64_t e = tid; e < n; e += nt) {
        const unsigned long long key = key_at(e);
        if ((key & himask) == prefix) atomicAdd(&hist[(key >> shift) & 255ull], 1u);
      }
      __syncthreads();
      if (tid == 0) {
        long long rem = s_remaining, acc = 0;
        int b = 255;
        for (; b > 0; --b) {
          if (acc + static_cast<long long>(hist[b]) >= rem) break;
          acc += hist[b];
        }
        s_remaining = rem - acc;
        s_prefix = prefix | (static_cast<unsigned long long>(b) << shift);
      }
      himask |= 255ull << shift;
      __syncthreads();
    }
    T = s_prefix;
  }
  for (int64_t i = tid; i < kp; i += nt) win[i] = 0ull;  // key 0 sorts last (score -inf / row -1 padding)
  __syncthreads();
  for (int64_t e = tid; e < n; e += nt) {
    const unsigned long long key = key_at(e);
    if (key >= T) {
      const unsigned long long pos = atomicAdd(&s_count, 1ull);
      if (pos < static_cast<unsigned long long>(kp)) win[pos] = key;
    }
  }
  __syncthreads();
  for (int64_t size = 2; size <= kp; size <<= 1) {
    for (int64_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (int64_t i = tid; i < kp; i += nt) {
        const int64_t j = i ^ stride;
        if (j > i) {
          const bool desc = (i & size) == 0;
          const unsigned long long x = win[i], y = win[j];
          if ((x < y) == desc) { win[i] = y; win[j] = x; }
        }
      }
      __syncthreads();
    }
  }
  for (int64_t i = tid; i < kout; i += nt) {
    const unsigned long long key = win[i];
    const uint32_t sbits = key_to_float_bits(static_cast<uint32_t>(key >> 32));
    const bool pad = key == 0ull || sbits == kNegInfBits;
    out_scores[static_cast<size_t>(q) * kout + i] = pad ? -INFINITY : __uint_as_float(sbits);
    out_idx[static_cast<size_t>(q) * kout + i] =
        pad ? -1 : static_cast<int64_t>(~static_cast<uint32_t>(key)) + row_offset;
  }
}

// ============================================================================ re-score
__global__ void __launch_bounds__(256)
rescore_kernel(const __half* __restrict__ store, int64_t N, int D, const __half* __restrict__ query,
               const int64_t* __restrict__ rows, int R, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + warp;
  if (r >= R) return;
  const int64_t row = rows[r];
  float acc = 0.f;
  if (row >= 0 && row < N) {
    const __half* x = store + row * D;
    for (int i = lane * 2; i < D; i += 64) {
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(x + i));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(query + i));
      acc += a.x * b.x + a.y * b.y;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[r] = acc;
}

// ============================================================================ host side
struct ScanCore {  // one launch of topk_scan_kernel over N rows
  int cg;  // 1: one CTA per unit, 2: CTA pairs (cta_group::2)
  int epl, cap, num_qt, q_pad, num_slices, tiles_per_slice, grid, ring;
  int qt_group, rotate;  // see ScanArgs
};
struct ScanPlan {
  ScanCore main, pre;
  int64_t n0;   // rows of the sample pass (0: none)
  int groups;   // n0 / 64
  int select_cap;  // candidate keys the select kernel can hold in shared memory
  size_t lists_bytes, counts_bytes, gmax_bytes, thresh_bytes, lthresh_bytes, done_bytes, sched_bytes, total_bytes;
};

static size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

static void plan_core(int64_t N, int D, int Q, int k, int sm_total, ScanCore* p) {
  // more than one 128-query tile: CTA pairs (256 queries x 256-row tiles) halve the operand traffic per FLOP;
  // a single tile is HBM-bound and keeps one CTA per SM so every SM streams its own slice
  p->cg = (Q > kScanQTile && sm_total % 2 == 0) ? 2 : 1;
  const int sms = sm_total / p->cg;  // units
  const int qtile = kScanQTile * p->cg, rowtile = kScanRowTile * p->cg;
  p->epl = k <= 128 ? 16 : 64;
  p->cap = 32 * p->epl;
  p->num_qt = (Q + qtile - 1) / qtile;
  p->q_pad = p->num_qt * qtile;
  p->qt_group = p->num_qt;
  // the query tiles that share a slice start it at different tiles (and wrap around): a row tile is then fetched by one
  // unit and found in L2 by the others over the length of an item, instead of being requested by all of them at once
  // (config 3, ncu: DRAM reads 106 -> 73 GB; 1-3 % less time at 20-40 query tiles, neutral below 4)
  p->rotate = p->num_qt >= 4 ? 1 : 0;
  if (const char* e = getenv("EVR_SCAN_QT_GROUP")) {   // experiment: passes of g query tiles (measured: 2.5 x MORE DRAM reads)
    if (atoi(e) > 0 && atoi(e) < p->num_qt) p->qt_group = atoi(e);
  }
  if (const char* e = getenv("EVR_SCAN_ROTATE")) p->rotate = e[0] == '1' ? 1 : 0;
  const int64_t tiles = (N + rowtile - 1) / rowtile;
  int S;
  if (p->num_qt == 1) {
    S = static_cast<int>(tiles < sms ? tiles : sms);
  } else {
    // several query tiles: pick the slice count whose item count quantises best onto the SMs
    int best = 1;
    double best_eff = -1.0;
    const int max_s = static_cast<int>(tiles < 64 ? tiles : 64);
    for (int s = 1; s <= max_s; ++s) {
      const long long items = static_cast<long long>(p->num_qt) * s;
      const long long rounds = (items + sms - 1) / sms;
      double eff = static_cast<double>(items) / static_cast<double>(rounds * sms);
      if (rounds < 2 && items < sms) eff *= 0.5;
      if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
    }
    // the smallest slice count within 3 % of the best: fewer restarts of the per-item selection, fewer candidates
    for (int s = 1; s <= best; ++s) {
      const long long items = static_cast<long long>(p->num_qt) * s;
      const long long rounds = (items + sms - 1) / sms;
      double eff = static_cast<double>(items) / static_cast<double>(rounds * sms);
      if (rounds < 2 && items < sms) eff *= 0.5;
      if (eff >= 0.97 * best_eff) { best = s; break; }
    }
    S = best;
    // A slice is read by all num_qt query tiles, on different units.  Over round 1's 277 MB slices (10 M rows / 37) the
    // units drifted further apart than L2 can hold and most of them fetched their rows from HBM on their own: ncu
    // measured 136.6 GB of DRAM reads for a 10.24 GB store at Q = 10 000.  Slices are therefore small -- everything in
    // flight at one time, ceil(units / num_qt) consecutive slices, is a few MB -- and handed out in slice-major order
    // (dynamically, see sched_claim), so the units reading a slice are never more than a round apart: 11 GB of DRAM
    // reads (profiles/r02_i_scan_c3_l2.txt).  6-96 MB in flight all measure the same.
    const int conc = (sms + p->qt_group - 1) / p->qt_group;
    int64_t inflight_mb = 12;
    if (const char* e = getenv("EVR_SCAN_INFLIGHT_MB")) {   // tuning knob
      if (atoi(e) > 0) inflight_mb = atoi(e);
    }
    int64_t target = (inflight_mb << 20) / conc;
    if (target > (32ll << 20)) target = 32ll << 20;
    if (target < (1ll << 20)) target = 1ll << 20;
    int64_t tps = target / (static_cast<int64_t>(rowtile) * D * 2);
    if (tps < 2) tps = 2;
    const int64_t s_l2 = (tiles + tps - 1) / tps;
    if (s_l2 > S) S = static_cast<int>(s_l2 < (1 << 20) ? s_l2 : (1 << 20));
  }
  p->tiles_per_slice = static_cast<int>((tiles + S - 1) / S);
  p->num_slices = static_cast<int>((tiles + p->tiles_per_slice - 1) / p->tiles_per_slice);
  // candidate lists: one per slice while there are few; otherwise a ring that slice s shares with s - ring, s - 2 ring,
  // ... (the kernel orders them with a flag per (32 queries, slice)); deep enough for the slices in flight plus drift
  const int conc = p->num_qt > 1 ? (sms + p->qt_group - 1) / p->qt_group : p->num_slices;
  // (2 conc + 6 was too shallow: at config 3, conc = 2, the epilogue warps stood at the flag of slice - 10 often enough
  // to cost 8 % -- scripts/scan_plan_probe.py: 89.8 ms at ring 10, 82-84 ms at 16-40 -- and, by holding units back at
  // different moments, to pull the units that share a slice apart: DRAM reads 80-113 GB at ring 10, 21-29 GB at 20,
  // 15-20 GB at 32, 14.5 GB at 48 for the 10.24 GB store; 32 lists of 4 KB per query = 1.3 GB of workspace at Q = 10 000)
  int ring = 2 * conc + 28;
  // (With the static round-robin of round 2's first half the depth had to grow with the scan -- 50 M rows: 447 ms at
  // ring 32, 391 at 96; 100 M rows: 933 ms at 96, 845 at 224 -- because the units drift apart for good; with the items
  // claimed dynamically, see sched_claim, 32 is enough at every size: 50 M rows 384 ms, 100 M rows 758 ms.)
  if (const char* e = getenv("EVR_SCAN_RING")) {   // experiments: how far the units may drift apart, in slices
    if (atoi(e) > 0) ring = atoi(e);
  }
  if (p->num_qt == 1 || ring > p->num_slices) ring = p->num_slices;
  p->ring = ring;
  const long long items = static_cast<long long>(p->num_qt) * p->num_slices;
  p->grid = static_cast<int>(items < sms ? items : sms) * p->cg;
}

// Sample pass: the scan kernel runs over n0 rows (whole tiles spread evenly over the store) recording only the maximum of every 64-row group; the
// k-th largest group maximum of a query is a lower bound on its global k-th best score, so the full scan starts
// with a pass rate of ~k/n0 instead of admitting (and pruning) everything until its own threshold settles.
// n0 ~ the rows one SM sees at small query batch, <= 1/64 of larger stores: < 2 % of the scan.
static int64_t sample_rows(int64_t N, int k, int Q) {
  if (N < 32768) return 0;
  int64_t n0 = 65536;
  if (Q > kScanQTile) {
    // tensor-bound regime: the epilogue's cost is the admissions, so buy a tighter threshold with a larger sample
    // (pass rate k/n0 ~ 3e-4; the sample pass itself admits nothing and runs at full MMA rate)
    n0 = 3072ll * k;
    if (n0 > N / 16) n0 = N / 16;
    if (n0 < 65536) n0 = 65536;
  }
  if (n0 > N / 4) n0 = N / 4;
  n0 &= ~int64_t(2 * kScanRowTile - 1);  // whole tiles for CTA pairs too: the sample pass sees no padding rows
  if (n0 / 64 < 2 * static_cast<int64_t>(k)) return 0;
  return n0;
}

static int plan_scan(int64_t N, int D, int Q, int k, ScanPlan* p) {
  EVR_REQUIRE(N >= 1 && N < (int64_t(1) << 31) - 256, "top-k scan: N=%lld out of range", (long long)N);
  EVR_REQUIRE(D % 64 == 0 && D >= 64 && D <= 768, "top-k scan: D=%d must be a multiple of 64 in [64, 768]", D);
  EVR_REQUIRE(Q >= 1, "top-k scan: Q=%d", Q);
  EVR_REQUIRE(k >= 1 && k <= kMaxFusedK, "top-k scan: k=%d must be in [1, %d]", k, kMaxFusedK);
  const int sms = sm_count();
  plan_core(N, D, Q, k, sms, &p->main);
  EVR_REQUIRE(p->main.ring <= 256, "top-k scan: %d candidate lists per query exceed the select kernel's limit",
              p->main.ring);
  p->n0 = sample_rows(N, k, Q);
  p->groups = static_cast<int>(p->n0 / 64);
  p->pre = ScanCore{};
  if (p->n0 > 0) plan_core(p->n0, D, Q, k, sms, &p->pre);
  const long long worst = static_cast<long long>(p->main.ring) * p->main.cap;
  // shared-memory budget of the select kernel: everything at small query batch (one CTA per query, ~15 k
  // candidates), a few CTAs per SM at large batch; longer lists fall back to re-reading global memory
  const long long budget = Q <= 256 ? 24576 : 6144;
  p->select_cap = static_cast<int>(worst < budget ? worst : budget);
  p->lists_bytes = up256(static_cast<size_t>(p->main.ring) * p->main.q_pad * p->main.cap * sizeof(int2));
  p->counts_bytes = up256(static_cast<size_t>(p->main.ring) * p->main.q_pad * sizeof(int));
  p->lthresh_bytes = up256(static_cast<size_t>(p->main.ring) * p->main.q_pad * sizeof(float));
  p->done_bytes = up256(static_cast<size_t>(p->main.q_pad / 32) * p->main.num_slices * sizeof(unsigned));
  p->gmax_bytes = p->n0 > 0 ? up256(static_cast<size_t>(p->main.q_pad) * p->groups * sizeof(float)) : 0;
  p->thresh_bytes = up256(static_cast<size_t>(p->main.q_pad) * sizeof(float));
  p->sched_bytes = up256(256 + 160 * kSchedDepth * sizeof(unsigned long long));   // item counter + tickets of <= 160 units
  p->total_bytes = p->lists_bytes + p->counts_bytes + p->gmax_bytes + p->thresh_bytes + p->lthresh_bytes +
                   p->done_bytes + p->sched_bytes + 256;
  return 0;
}

template <typename IdxT>
static int launch_final(const float* scores, const IdxT* idx, int P, int q_stride, int Q, int k, int64_t row_offset,
                        float* out_scores, int64_t* out_idx, cudaStream_t stream) {
  int kp = 1;
  while (kp < k) kp <<= 1;
  const size_t smem = static_cast<size_t>(kp) * 16;
  ProfScope ps(PF_FINAL, stream);
  topk_final_kernel<IdxT><<<Q, 256, smem, stream>>>(scores, idx, P, q_stride, k, row_offset, out_scores, out_idx);
  EVR_CHECK_LAUNCH();
  return 0;
}

template <int EPL, int CG>
static int launch_scan(int grid, int smem, cudaStream_t s, const CUtensorMap& tq, const CUtensorMap& ts,
                       const ScanArgs& a) {
  auto kern = topk_scan_kernel<EPL, CG>;
  static PerDeviceMax configured;  // per instantiation
  if (configured.needs(smem)) {
    EVR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kScanThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  EVR_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tq, ts, a));
  return 0;
}

}  // namespace evr

using namespace evr;

extern "C" int evr_store_prepare(const void* src, int src_dtype, int64_t N, int D, void* dst_fp16, void* stream) {
  EVR_REQUIRE(src && dst_fp16, "evr_store_prepare: null pointer");
  EVR_REQUIRE(src_dtype == 0 || src_dtype == 1, "evr_store_prepare: src_dtype %d not in {0,1}", src_dtype);
  EVR_REQUIRE(D >= 2 && D % 2 == 0, "evr_store_prepare: D=%d must be even", D);
  if (N <= 0) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const unsigned blocks = static_cast<unsigned>((N + 7) / 8);
  ProfScope ps(PF_STORE, s);
  if (src_dtype == 0)
    store_prepare_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(src), N, D, static_cast<__half*>(dst_fp16));
  else
    store_prepare_kernel<__half><<<blocks, 256, 0, s>>>(static_cast<const __half*>(src), N, D, static_cast<__half*>(dst_fp16));
  EVR_CHECK_LAUNCH();
  return 0;
}

extern "C" int evr_topk_workspace_bytes(int64_t N, int D, int Q, int k, size_t* out_bytes) {
  EVR_REQUIRE(out_bytes, "evr_topk_workspace_bytes: null out");
  ScanPlan p;
  if (int r = plan_scan(N, D, Q, k, &p)) return r;
  *out_bytes = p.total_bytes;
  return 0;
}

// N rows are scanned; the store itself holds N_store >= N rows and tile t of the launch reads store tile
// t * tile_stride (the sample pass spreads its n0 rows evenly over the whole store, so a store laid out video by video
// in temporal order still yields a representative threshold; tile_stride = 1 and N_store = N for the full scan).
static int run_scan(const ScanCore& c, const void* store, int64_t N, int64_t N_store, int tile_stride, int D,
                    const void* queries, int Q, int k, const float* init_thresh, int2* lists, int* counts,
                    float* list_thresh, unsigned* done, unsigned epoch, float* group_max, int groups, cudaStream_t s,
                    int* sched_counter = nullptr, unsigned long long* sched_tickets = nullptr) {
  ScanArgs a{};
  a.sched_counter = sched_counter; a.sched_tickets = sched_tickets;
  a.N = N; a.D = D; a.Q = Q; a.k = k; a.tile_stride = tile_stride;
  { const char* e = getenv("EVR_SCAN_EVICT_LAST"); a.row_evict_last = (e && e[0] == '1' && c.num_qt > 1) ? 1 : 0; }
  a.num_qt = c.num_qt; a.num_slices = c.num_slices; a.tiles_per_slice = c.tiles_per_slice; a.q_pad = c.q_pad;
  a.cap = c.cap;
  a.init_thresh = init_thresh;
  a.lists = lists; a.counts = counts; a.group_max = group_max; a.groups = groups;
  a.ring = c.ring; a.list_thresh = list_thresh; a.done = done; a.epoch = epoch;
  a.qt_group = c.qt_group > 0 ? c.qt_group : c.num_qt;
  a.rotate = c.rotate;
  CUtensorMap tq, ts;
  if (int r = make_tmap_2d_16b(&tq, queries, Q, D, static_cast<uint64_t>(D) * 2, kScanQTile, false)) return r;
  if (int r = make_tmap_2d_16b(&ts, store, N_store, D, static_cast<uint64_t>(D) * 2, kScanRowTile, false)) return r;
  // the resident query tile takes 128 * D * 2 bytes; the ring gets what is left of the 227 KB (6 stages up to D = 512,
  // 2 at D = 768: ViT-L/14 stores scan correctly, with a shallower pipeline)
  int stages = (227 * 1024 - 1024 - 256 - kScanQTile * D * 2) / kScanStageBytes;
  if (stages > kScanStages) stages = kScanStages;
  a.stages = stages;
  const int smem = kScanQTile * D * 2 + stages * kScanStageBytes + 1024 + 256;
  {
    ProfScope ps(PF_SCAN, s);
    int r;
    if (c.cg == 2) r = c.epl == 16 ? launch_scan<16, 2>(c.grid, smem, s, tq, ts, a) : launch_scan<64, 2>(c.grid, smem, s, tq, ts, a);
    else r = c.epl == 16 ? launch_scan<16, 1>(c.grid, smem, s, tq, ts, a) : launch_scan<64, 1>(c.grid, smem, s, tq, ts, a);
    if (r) return r;
  }
  EVR_CHECK_LAUNCH();
  return 0;
}

extern "C" int evr_topk_search(const void* store, int64_t N, int D, int64_t row_offset, const void* queries, int Q,
                               int k, float* out_scores, int64_t* out_idx, void* workspace, size_t workspace_bytes,
                               void* stream) {
  EVR_REQUIRE(store && queries && out_scores && out_idx && workspace, "evr_topk_search: null pointer");
  ScanPlan p;
  if (int r = plan_scan(N, D, Q, k, &p)) return r;
  EVR_REQUIRE(workspace_bytes >= p.total_bytes, "evr_topk_search: workspace %zu < required %zu", workspace_bytes,
              p.total_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  int2* lists = reinterpret_cast<int2*>(ws);
  int* counts = reinterpret_cast<int*>(ws + p.lists_bytes);
  float* gmax = reinterpret_cast<float*>(ws + p.lists_bytes + p.counts_bytes);
  float* thresh = reinterpret_cast<float*>(ws + p.lists_bytes + p.counts_bytes + p.gmax_bytes);
  const float* init = nullptr;
  if (p.n0 > 0) {
    const int rowtile = kScanRowTile * p.pre.cg;
    int64_t stride = (N / rowtile) / (p.n0 / rowtile);  // >= 4: n0 <= N / 4, whole tiles only
    if (const char* e = getenv("EVR_SAMPLE_PREFIX")) {  // A/B knob: sample the first n0 rows instead
      if (e[0] == '1') stride = 1;
    }
    if (int r = run_scan(p.pre, store, p.n0, N, static_cast<int>(stride), D, queries, Q, k, nullptr, nullptr, nullptr,
                         nullptr, nullptr, 0u, gmax, p.groups, s)) return r;
    {
      ProfScope ps(PF_FINAL, s);
      sample_kth_kernel<<<(Q + 7) / 8, 256, 0, s>>>(gmax, p.groups, Q, k, thresh);
    }
    EVR_CHECK_LAUNCH();
    init = thresh;
  }
  float* lthresh = reinterpret_cast<float*>(ws + p.lists_bytes + p.counts_bytes + p.gmax_bytes + p.thresh_bytes);
  unsigned* done = reinterpret_cast<unsigned*>(ws + p.lists_bytes + p.counts_bytes + p.gmax_bytes + p.thresh_bytes +
                                               p.lthresh_bytes);
  static std::atomic<unsigned> epoch_counter{0};
  unsigned epoch = epoch_counter.fetch_add(1, std::memory_order_relaxed) + 1;
  if (epoch == 0) epoch = epoch_counter.fetch_add(1, std::memory_order_relaxed) + 1;   // 0 is what fresh memory holds
  // dynamic item distribution for the multi-tile plans (a single query tile has one slice per SM: nothing to balance)
  int* sched_counter = nullptr;
  unsigned long long* sched_tickets = nullptr;
  static const bool static_items = [] { const char* e = getenv("EVR_SCAN_STATIC"); return e && e[0] == '1'; }();
  if (p.main.num_qt > 1 && !static_items && p.main.grid / p.main.cg <= 160) {
    uint8_t* sched = reinterpret_cast<uint8_t*>(done) + p.done_bytes;
    EVR_CHECK_CUDA(cudaMemsetAsync(sched, 0, p.sched_bytes, s));
    sched_counter = reinterpret_cast<int*>(sched);
    sched_tickets = reinterpret_cast<unsigned long long*>(sched + 256);
  }
  if (int r = run_scan(p.main, store, N, N, 1, D, queries, Q, k, init, lists, counts, lthresh, done, epoch, nullptr, 0,
                       s, sched_counter, sched_tickets)) return r;
  SelectArgs sa{};
  sa.lists = lists; sa.counts = counts;
  sa.P = p.main.ring; sa.q_pad = p.main.q_pad; sa.cap = p.main.cap; sa.k = k; sa.smem_cap = p.select_cap;
  sa.row_offset = row_offset; sa.out_scores = out_scores; sa.out_idx = out_idx;
  int kp = 1;
  while (kp < k) kp <<= 1;
  const int sel_smem = (kp + p.select_cap) * 8;
  static PerDeviceMax sel_configured;
  if (sel_smem > 48 * 1024 && sel_configured.needs(sel_smem)) {
    EVR_CHECK_CUDA(cudaFuncSetAttribute(topk_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sel_smem));
  }
  {
    ProfScope ps(PF_FINAL, s);
    topk_select_kernel<<<Q, 256, sel_smem, s>>>(sa);
  }
  EVR_CHECK_LAUNCH();
  return 0;
}

// ---- the scan with its scratch in a static device array -------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The array belongs to this
// module: every GPU that loads the library holds its EVR_TOPK_STATIC_SCRATCH_BYTES until the process exits.  In host code
// `g_static_scan_scratch` is only a stand-in for the symbol -- the device address comes from cudaGetSymbolAddress, once
// per device with that device current.
static __device__ __align__(256) unsigned char g_static_scan_scratch[EVR_TOPK_STATIC_SCRATCH_BYTES];

static int static_scan_scratch(void** out) {
  static std::mutex mu;
  static void* address[64] = {};
  int dev = 0;
  EVR_CHECK_CUDA(cudaGetDevice(&dev));
  EVR_REQUIRE(dev >= 0 && dev < 64, "evr_topk_search_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (address[dev] == nullptr) {
    void* p = nullptr;
    EVR_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_scan_scratch));
    EVR_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0,
                "evr_topk_search_static: bad address %p for the static scratch array", p);
    address[dev] = p;
  }
  *out = address[dev];
  return 0;
}

extern "C" int evr_topk_search_static(const void* store, int64_t N, int D, int64_t row_offset, const void* queries,
                                      int Q, int k, float* out_scores, int64_t* out_idx, void* stream) {
  EVR_REQUIRE(store && queries && out_scores && out_idx, "evr_topk_search_static: null pointer");
  ScanPlan p;
  if (int r = plan_scan(N, D, Q, k, &p)) return r;
  EVR_REQUIRE(p.total_bytes <= static_cast<size_t>(EVR_TOPK_STATIC_SCRATCH_BYTES),
              "evr_topk_search_static: N=%lld D=%d Q=%d k=%d needs %zu bytes of scratch, the static array holds %zu "
              "(use evr_topk_search with a caller-owned workspace)", (long long)N, D, Q, k, p.total_bytes,
              static_cast<size_t>(EVR_TOPK_STATIC_SCRATCH_BYTES));
  void* ws = nullptr;
  if (int r = static_scan_scratch(&ws)) return r;
  // The array keeps the previous call's contents.  The candidate lists need no clearing (a list is written before it is
  // read); the counters, thresholds and slice flags behind them start every call from zero.
  const size_t tail = p.total_bytes > p.lists_bytes ? p.total_bytes - p.lists_bytes : 0;
  if (tail) EVR_CHECK_CUDA(cudaMemsetAsync(static_cast<uint8_t*>(ws) + p.lists_bytes, 0, tail,
                                           static_cast<cudaStream_t>(stream)));
  return evr_topk_search(store, N, D, row_offset, queries, Q, k, out_scores, out_idx, ws,
                         static_cast<size_t>(EVR_TOPK_STATIC_SCRATCH_BYTES), stream);
}

extern "C" int evr_topk_merge(const float* scores, const int64_t* idx, int P, int Q, int k, float* out_scores,
                              int64_t* out_idx, void* stream) {
  EVR_REQUIRE(scores && idx && out_scores && out_idx, "evr_topk_merge: null pointer");
  EVR_REQUIRE(P >= 1 && Q >= 1 && k >= 1 && k <= 1024, "evr_topk_merge: bad shape P=%d Q=%d k=%d", P, Q, k);
  return launch_final<int64_t>(scores, idx, P, Q, Q, k, 0, out_scores, out_idx, static_cast<cudaStream_t>(stream));
}

static int64_t pow2_at_least(int64_t k) {
  int64_t kp = 1;
  while (kp < k) kp <<= 1;
  return kp;
}

extern "C" int evr_topk_large_workspace_bytes(int64_t N, int Q, int64_t k, size_t* out_bytes) {
  EVR_REQUIRE(out_bytes && N >= 1 && Q >= 1 && k >= 1, "evr_topk_large_workspace_bytes: bad arguments");
  const int qc = Q < kLargeQChunk ? Q : kLargeQChunk;
  *out_bytes = up256(static_cast<size_t>(qc) * N * sizeof(float)) +
               up256(static_cast<size_t>(qc) * pow2_at_least(k < N ? k : N) * 8) + 256;
  return 0;
}

extern "C" int evr_topk_search_large(const void* store, int64_t N, int D, int64_t row_offset, const void* queries,
                                     int Q, int64_t k, float* out_scores, int64_t* out_idx, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  EVR_REQUIRE(store && queries && out_scores && out_idx && workspace, "evr_topk_search_large: null pointer");
  EVR_REQUIRE(N >= 1 && N < (int64_t(1) << 31) && D >= 2 && D % 2 == 0 && Q >= 1 && k >= 1 && k <= N,
              "evr_topk_search_large: bad shape N=%lld D=%d Q=%d k=%lld (need 1 <= k <= N)", (long long)N, D, Q,
              (long long)k);
  size_t need = 0;
  if (int r = evr_topk_large_workspace_bytes(N, Q, k, &need)) return r;
  EVR_REQUIRE(workspace_bytes >= need, "evr_topk_search_large: workspace %zu < required %zu", workspace_bytes, need);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int qc_max = Q < kLargeQChunk ? Q : kLargeQChunk;
  const int64_t kp = pow2_at_least(k);
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  float* scores = reinterpret_cast<float*>(ws);
  unsigned long long* scratch =
      reinterpret_cast<unsigned long long*>(ws + up256(static_cast<size_t>(qc_max) * N * sizeof(float)));
  const size_t qsm = static_cast<size_t>(kLargeQChunk) * D * 2;
  EVR_REQUIRE(qsm <= 48 * 1024, "evr_topk_search_large: D=%d too wide", D);
  const int64_t want_blocks = (N + 7) / 8;
  const int blocks = static_cast<int>(want_blocks < 8LL * sm_count() ? want_blocks : 8LL * sm_count());
  for (int q0 = 0; q0 < Q; q0 += kLargeQChunk) {
    const int qc = Q - q0 < kLargeQChunk ? Q - q0 : kLargeQChunk;
    {
      ProfScope ps(PF_SCAN, s);
      scores_rows_kernel<<<blocks, 256, qsm, s>>>(static_cast<const __half*>(store), N, D,
                                                  static_cast<const __half*>(queries) + static_cast<size_t>(q0) * D, qc,
                                                  scores);
    }
    EVR_CHECK_LAUNCH();
    {
      ProfScope ps(PF_FINAL, s);
      topk_large_kernel<<<qc, 1024, 0, s>>>(scores, nullptr, 1, qc, N, k, row_offset, scratch, kp,
                                            out_scores + static_cast<size_t>(q0) * k, out_idx + static_cast<size_t>(q0) * k);
    }
    EVR_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int evr_topk_merge_large(const float* scores, const int64_t* idx, int P, int Q, int64_t k, float* out_scores,
                                    int64_t* out_idx, void* scratch, size_t scratch_bytes, void* stream) {
  EVR_REQUIRE(scores && idx && out_scores && out_idx && scratch, "evr_topk_merge_large: null pointer");
  EVR_REQUIRE(P >= 1 && Q >= 1 && k >= 1, "evr_topk_merge_large: bad shape P=%d Q=%d k=%lld", P, Q, (long long)k);
  const int64_t kp = pow2_at_least(k);
  EVR_REQUIRE(scratch_bytes >= static_cast<size_t>(Q) * kp * 8, "evr_topk_merge_large: scratch %zu < %zu",
              scratch_bytes, static_cast<size_t>(Q) * kp * 8);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  ProfScope ps(PF_FINAL, s);
  topk_large_kernel<<<Q, 1024, 0, s>>>(scores, idx, P, Q, k, k, 0, static_cast<unsigned long long*>(scratch), kp,
                                       out_scores, out_idx);
  EVR_CHECK_LAUNCH();
  return 0;
}

extern "C" int evr_rescore(const void* store, int64_t N, int D, const void* query_fp16, const int64_t* rows, int R,
                           float* out, void* stream) {
  EVR_REQUIRE(store && query_fp16 && rows && out, "evr_rescore: null pointer");
  EVR_REQUIRE(D % 2 == 0, "evr_rescore: D must be even");
  if (R <= 0) return 0;
  rescore_kernel<<<(R + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __half*>(store), N, D, static_cast<const __half*>(query_fp16), rows, R, out);
  EVR_CHECK_LAUNCH();
  return 0;
}

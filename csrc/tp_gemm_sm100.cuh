// Persistent warp-specialised tcgen05 GEMM for sm_100a with the tensor-parallel
// collectives fused into its tile schedule.
//
//   mode PLAIN      C[M,N]      = A[M,K] * B                        (world = 1 / compute ceiling)
//   mode AG_BEFORE  C[m,N]      = allgather_rows(A)[m,K] * B        (tp_columnwise, gather inputs)
//   mode AG_AFTER   C[m,N]      = allgather_rows(A_local * B)       (tp_columnwise, gather outputs)
//   mode RS         C[m/d,N]    = reduce_scatter_rows(A_local * B)  (tp_rowwise)
//
// Replaces the reference's library call pairs: ncclAllGather + cuBLAS GEMM
// (ddlb/primitives/TPColumnwise/pytorch.py:96-102) and cuBLAS GEMM +
// ncclReduceScatter (ddlb/primitives/TPRowwise/pytorch.py:82-83), and the
// stream-pipelined nvFuser/TE variants (TPColumnwise/fuser.py:16-146,
// TPRowwise/fuser.py:15-169): their "algorithm/s/order" knobs are tile orders here, their
// transports (nccl / symmetric-memory copies / multimem) are the data paths below.
//
// CTA roles: warp 0 = TMA producer, warp 1 = MMA issuer (one lane), warp 2 = TMEM allocator,
// warps 4..7 = epilogue (one TMEM lane quadrant each). Warps 2, 3 and 8..11 ("aux" warps; the
// last four exist only in launches of 384 threads) carry the collective:
//   AG_BEFORE  they move A shards over NVLink (128-bit loads/stores through registers, or
//              multimem.st through the NVSwitch) and publish per-chunk arrival counters that
//              the TMA producers wait on;
//   RS         they reduce: every partial tile is stored (by the epilogue warps) into a
//              staging slot - at the owner (push), or locally (pull) - and the owner's aux
//              warps sum the d slots of a 32-row x 128-byte unit as soon as all of them have
//              landed: local loads (push), peer loads (pull) or ONE multimem.ld_reduce per
//              16 bytes, summed inside the NVSwitch (pull over multicast).
//   fp32       (3xTF32) warps 8..11 split the fp32 operands into tf32 hi + lo parts.
//
// Cross-GPU protocol: all flags are monotonic (value = epoch * arrivals) so
// back-to-back launches never need a clearing pass; data that can be overwritten by
// a peer running one launch ahead is double-buffered by epoch parity.
#pragma once

#include "ptx.cuh"

namespace ddlb {

constexpr int kBlockM = 128;   // rows per CTA (a cta_group::2 pair covers 256)
constexpr int kBlockK = 64;    // 64 x 16-bit = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kBaseThreads = 256;               // warps 0..7
constexpr int kMaxThreads = 384;                // + aux warps 8..11
constexpr int kMaxWorld = 8;
constexpr int kEpiBufBytes = 32 * 128;          // 32 rows x 128 B per warp buffer
constexpr int kEpiBytes = 4 * 2 * kEpiBufBytes;  // 4 warps x double buffer
constexpr int kSmemBudget = 193 * 1024;         // operand ring

enum : int { kModePlain = 0, kModeAGBefore = 1, kModeAGAfter = 2, kModeRS = 3 };
// who moves the A shards in AG_BEFORE
enum : int { kAgPull = 0, kAgPush = 1, kAgPushMulticast = 2 };
// how AG_AFTER's epilogue writes a C tile into every rank's C
enum : int { kEpiTma = 0, kEpiSt128 = 1, kEpiMulticast = 2 };
// where RS partial tiles are staged and how the owner sums them
enum : int { kRsPush = 0, kRsPull = 1, kRsPullMulticast = 2 };
// activation fused into the epilogue (applied to the fp32 accumulator before rounding)
enum : int { kActNone = 0, kActRelu = 1, kActSilu = 2, kActGeluTanh = 3 };

// word offsets inside a rank's signal block
constexpr int kSigReady = 0;    // [src]   epoch written by src when its kernel started
constexpr int kSigDone = 8;     // [src]   AG_AFTER: += 1 per CTA of src whose stores landed
                                //         AG_BEFORE (pull): += 1 per aux warp of src that has
                                //         finished reading this rank's shard
constexpr int kSigClaim = 16;   // [2]     RS: reducer work queue heads, by epoch parity
constexpr int kSigChunk = 32;   // [chunk] AG_BEFORE: += 1 per unit copied;  RS: [cta_tile] += 1
                                //         per epilogue warp of every rank whose store landed

struct TpGemmParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  CUtensorMap tmOut[kMaxWorld];
  int mode;
  int rank, world;
  int num_m_blocks;       // cluster tiles along M (of the A operand map)
  int num_n_blocks;
  int sched_n_blocks;     // n-blocks the tile order sees: num_n_blocks / pairs (cluster of 2 MMA pairs)
  int num_k_blocks;
  int group_m;            // rasterisation: m-blocks per group (L2 reuse)
  int mblks_per_part;     // cluster tiles per rank shard (AG_BEFORE / RS), else num_m_blocks
  int chunk_mblks;        // cluster tiles per chunk (divides mblks_per_part)
  int part0;              // first shard in visiting order (AG: own; RS push: next rank; else 0)
  int part_rows;          // rows per shard (m / world)
  int out_n;              // N (elements per output row)
  int block_n;            // BLOCK_N of this kernel (the reducers are not templated on it)
  int elem;               // bytes per element
  uint32_t idesc;
  uint32_t epoch;
  int stages;             // operand ring depth actually used (<= Cfg::kStages)
  int act;                // epilogue activation on the final output (PLAIN / AG_*): kAct*
  int pdl;                // launched with programmatic stream serialization (see griddep_wait)
  // AG_BEFORE
  int ag_transport;
  int comm_warps;                     // aux warps per CTA that move data (0 = all of them)
  const uint8_t* peer_a[kMaxWorld];   // every rank's gathered-A buffer (peer mapped)
  uint8_t* local_a;                   // this rank's gathered-A buffer
  uint32_t chunk_bytes;               // bytes of one chunk (chunk rows * K * elem)
  // AG_AFTER
  int epi_transport;
  // RS
  int rs_transport;
  uint8_t* stage_base[kMaxWorld];     // every rank's staging [2][world][part_rows][N]
  int rs_tiles;                       // (part_rows / 128) * num_n_blocks counters
  // multicast alias of: AG_BEFORE the gathered A, AG_AFTER C, RS the staging (or null)
  uint8_t* mc_ptr;
  // signals
  uint32_t* sig_local;
  uint32_t* sig_peer[kMaxWorld];
  // optional device trace (kTrace kernels only): [cta][role][kTraceSlots] x {tag, arg, time ns}
  uint4* trace;
  // raw output addresses: AG_AFTER every rank's C, otherwise [rank] = this rank's output
  uint8_t* out_raw[kMaxWorld];
};

// kWide: one scheduled tile = TWO neighbouring n-blocks (256 x 2*BLOCK_N per CTA pair) that share
// every A stage; both accumulators live in TMEM at once (no double buffering, the epilogue is
// exposed once per tile), and a k-block moves 48 KB instead of 2 x 32 KB per CTA.
template <int kCtaGroup, int kBlockN, bool kSplit = false, bool kWide = false>
struct TpGemmCfg {
  static_assert(!kWide || (kCtaGroup == 2 && kBlockN == 256 && !kSplit), "wide tiles: cta_group 2, BLOCK_N 256");
  static constexpr int kNSub = kWide ? 2 : 1;                    // n-blocks per scheduled tile
  static constexpr int kTileM = kBlockM * kCtaGroup;
  static constexpr int kLoadN = kBlockN / kCtaGroup;            // B columns staged per CTA
  static constexpr int kStageA = kBlockM * kBlockK * 2;
  static constexpr int kStageB = kLoadN * kBlockK * 2;
  static constexpr int kStageOps = kStageA + kNSub * kStageB;
  // 3xTF32: every stage also holds the low-order parts of A and B
  static constexpr int kStageBytes = kSplit ? 2 * kStageOps : kStageOps;
  static constexpr int kStagesRaw = kSmemBudget / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = 2 * kBlockN;
  static constexpr int kBarBytes = 512;
  static constexpr int kFixedBytes = 1024 + kEpiBytes + kBarBytes;
  static constexpr int kSmemBytes = kFixedBytes + kStages * kStageBytes;
  static_assert(kStages >= 3, "operand ring too shallow");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget exceeded");
};

struct TileCoord {
  int m_blk;   // cluster tile index along M in the A operand map
  int n_blk;
  int part;    // owning / source rank of these rows
  int chunk;   // chunk index inside the part
};

// Linear tile index -> tile, identical in every role. Grouped rasterisation over an
// arrival-ordered sequence of m-blocks.
__host__ __device__ __forceinline__ TileCoord tile_coord(const TpGemmParams& p, int t) {
  const int per_group = p.group_m * p.sched_n_blocks;
  const int group = t / per_group;
  const int within = t - group * per_group;
  const int first = group * p.group_m;
  const int left = p.num_m_blocks - first;
  const int rows_in_group = p.group_m < left ? p.group_m : left;
  const int n_blk = within / rows_in_group;
  const int ord = first + (within - n_blk * rows_in_group);
  TileCoord c;
  c.n_blk = n_blk;
  if (p.mode == kModePlain || p.mode == kModeAGAfter) {
    c.m_blk = ord;
    c.part = p.rank;
    c.chunk = 0;
    return c;
  }
  const int per_stage = p.world * p.chunk_mblks;
  const int stage = ord / per_stage;
  const int rem = ord - stage * per_stage;
  const int pi = rem / p.chunk_mblks;
  const int w = rem - pi * p.chunk_mblks;
  int part = p.part0 + pi;
  part = part >= p.world ? part - p.world : part;
  c.part = part;
  c.chunk = stage;
  c.m_blk = part * p.mblks_per_part + stage * p.chunk_mblks + w;
  return c;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi, bool bf16) {
  uint32_t out;
  if (bf16) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    memcpy(&out, &v, 4);
  } else {
    const __half2 v = __floats2half2_rn(lo, hi);
    memcpy(&out, &v, 4);
  }
  return out;
}

__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == kActRelu) return fmaxf(x, 0.f);
  if (act == kActSilu) return x * __frcp_rn(1.f + __expf(-x));
  // gelu, tanh form: 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3))) = x * sigmoid(2 u)
  const float u = 0.7978845608f * (x + 0.044715f * x * x * x);
  return x * __frcp_rn(1.f + __expf(-2.f * u));
}

__device__ __forceinline__ float2 unpack2(uint32_t w, bool bf16) {
  if (bf16) {
    __nv_bfloat162 h;
    memcpy(&h, &w, 4);
    return __bfloat1622float2(h);
  }
  __half2 h;
  memcpy(&h, &w, 4);
  return __half22float2(h);
}

// ---------------------------------------------------------------------------
// Opt-in device trace: a timeline of one launch without a profiler (ncu cannot replay
// kernels that wait for other ranks). Compiled in only for kTrace instantiations, so the
// production kernels carry no trace code at all.
// ---------------------------------------------------------------------------
constexpr int kTraceSlots = 128;
constexpr int kTraceRoles = 4;   // 0 producer, 1 mma, 2 epilogue (warp 4), 3 aux (warp 2)
enum : uint32_t {
  kEvStart = 1, kEvTileBegin = 2, kEvFlagWaited = 3, kEvTileIssued = 4, kEvAccReady = 5,
  kEvUnitReady = 6, kEvTileStored = 7, kEvUnitDone = 8, kEvEnd = 9,
};

template <bool kTrace>
struct Tracer {
  uint4* base;
  int n;
  __device__ __forceinline__ Tracer(const TpGemmParams& p, int cta, int role) : base(nullptr), n(0) {
    if constexpr (kTrace) {
      if (p.trace != nullptr && cta >= 0)
        base = p.trace + (static_cast<size_t>(cta) * kTraceRoles + role) * kTraceSlots;
    }
  }
  __device__ __forceinline__ void ev(uint32_t tag, uint32_t arg) {
    if constexpr (kTrace) {
      if (base != nullptr && n < kTraceSlots) {
        const uint64_t t = globaltimer_ns();
        base[n++] = make_uint4(tag, arg, static_cast<uint32_t>(t), static_cast<uint32_t>(t >> 32));
      }
    }
  }
};

// ---------------------------------------------------------------------------
// AG_BEFORE: aux warps of EVERY CTA move the A shards (register path).
// Measured on B200/NVSwitch: a 128-bit load/store copy reaches ~700 GB/s to or from a peer
// while cp.async.bulk through shared memory stays near 400 GB/s, and a handful of dedicated
// comm CTAs cannot fill the links (one SM sustains ~20 GB/s over NVLink). Spread over all
// CTAs the per-SM demand is ~5 GB/s and the GEMM keeps every SM.
// The shards are cut into 32 KiB units; the units of all chunks, in arrival order, are dealt
// round-robin to the aux warps of the whole grid.
// ---------------------------------------------------------------------------
constexpr int kPushUnit = 32 * 1024;
constexpr int kCopyGroup = 8;                         // 16-byte vectors per lane per group
constexpr uint32_t kCopyStep = 32 * 16 * kCopyGroup;  // 4 KiB per warp per group

// One unit of work of a comm warp: `kPushUnit` bytes [begin, end) of chunk `chunk` of rank
// `src`'s shard; `base` is the byte offset of that chunk in a gathered buffer.
// Host-callable so the CPU tests can check that the units tile the shards exactly.
struct CommUnit {
  int src;
  int chunk;
  uint32_t begin, end;
  size_t base;
};

__host__ __device__ __forceinline__ CommUnit pull_unit(const TpGemmParams& p, int u, int upc,
                                                       size_t part_bytes) {
  CommUnit cu;
  const int c = u / upc;                               // chunk in fetch order
  const int stage = c / (p.world - 1);
  const int f = c - stage * (p.world - 1);             // f-th foreign shard in visiting order
  int own_at = p.rank - p.part0;                       // position of the own shard in that order
  own_at = own_at < 0 ? own_at + p.world : own_at;
  const int pi = f < own_at ? f : f + 1;
  int src = p.part0 + pi;
  src = src >= p.world ? src - p.world : src;
  cu.src = src;
  cu.chunk = stage;
  cu.begin = static_cast<uint32_t>(u - c * upc) * kPushUnit;
  const uint32_t lim = cu.begin + static_cast<uint32_t>(kPushUnit);
  cu.end = lim < p.chunk_bytes ? lim : p.chunk_bytes;
  cu.base = static_cast<size_t>(src) * part_bytes + static_cast<size_t>(stage) * p.chunk_bytes;
  return cu;
}

// Copy [begin, end) (multiples of 16 bytes) with two register groups in flight: the loads of
// group i+1 are issued before the stores of group i, so a warp always has 4..8 KiB outstanding
// (one round trip to a peer is ~2-4 us; the round-1 loop drained between groups and ran the
// links at ~60 %).
__device__ __forceinline__ void copy_pipelined(uint8_t* dst, const uint8_t* src, uint32_t begin,
                                               uint32_t end, int lane) {
  uint4 va[kCopyGroup], vb[kCopyGroup];
  const uint32_t l16 = static_cast<uint32_t>(lane) * 16u;
#define DDLB_LD_GROUP(v, off)                                                          \
  _Pragma("unroll") for (int i = 0; i < kCopyGroup; ++i) {                             \
    const uint32_t o = (off) + l16 + static_cast<uint32_t>(i) * 512u;                  \
    if (o < end) v[i] = __ldcs(reinterpret_cast<const uint4*>(src + o));               \
  }
#define DDLB_ST_GROUP(v, off)                                                          \
  _Pragma("unroll") for (int i = 0; i < kCopyGroup; ++i) {                             \
    const uint32_t o = (off) + l16 + static_cast<uint32_t>(i) * 512u;                  \
    if (o < end) __stcg(reinterpret_cast<uint4*>(dst + o), v[i]);                      \
  }
  uint32_t off = begin;
  DDLB_LD_GROUP(va, off)
  while (off < end) {
    DDLB_LD_GROUP(vb, off + kCopyStep)
    DDLB_ST_GROUP(va, off)
    DDLB_LD_GROUP(va, off + 2 * kCopyStep)
    DDLB_ST_GROUP(vb, off + kCopyStep)
    off += 2 * kCopyStep;
  }
#undef DDLB_LD_GROUP
#undef DDLB_ST_GROUP
}

// PULL (default): every unit of every peer's shard is read from the peer's buffer into
// registers and stored locally, then the local chunk counter is bumped. Units are CLAIMED from
// a queue head in fetch order (a static deal left the slowest warp 20 us behind the fastest
// after seven units, and the GEMM waits for the last chunk), and the first 4 KiB of the next
// unit are already in flight when the fence of the current one is issued, so a warp always has
// 4..8 KiB outstanding. Measured on B200 (device trace, 2 ranks): 2 warps per CTA move a unit in
// ~12.7 us = ~590 GB/s in total, the rate of a plain copy kernel with both directions busy; 6
// warps per CTA oversubscribe the fabric (28 us per unit, lower total). At the end each warp
// tells every peer that it will not read that peer's shard again in this launch.
template <bool kTrace>
__device__ __forceinline__ void comm_warp_pull(const TpGemmParams& p, int lane, Tracer<kTrace>& tr) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int upc = static_cast<int>((p.chunk_bytes + kPushUnit - 1) / kPushUnit);
  const int total = chunks_per_part * (p.world - 1) * upc;
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;
  uint32_t* head = p.sig_local + kSigClaim + (p.epoch & 1u);
  const uint32_t l16 = static_cast<uint32_t>(lane) * 16u;

  // a peer's shard is final once that peer has started this launch
  if (lane < p.world && lane != p.rank) wait_flag_ge<true>(p.sig_local + kSigReady + lane, p.epoch);
  __syncwarp();
  tr.ev(kEvFlagWaited, 0);

  auto claim = [&]() {
    int u = 0;
    if (lane == 0) u = static_cast<int>(atomicAdd(head, 1u));
    return __shfl_sync(0xffffffffu, u, 0);
  };
#define DDLB_LD_GROUP(v, from, off, lim)                                               \
  _Pragma("unroll") for (int i = 0; i < kCopyGroup; ++i) {                             \
    const uint32_t o = (off) + l16 + static_cast<uint32_t>(i) * 512u;                  \
    if (o < (lim)) v[i] = __ldcs(reinterpret_cast<const uint4*>((from) + o));          \
  }
#define DDLB_ST_GROUP(v, to, off, lim)                                                 \
  _Pragma("unroll") for (int i = 0; i < kCopyGroup; ++i) {                             \
    const uint32_t o = (off) + l16 + static_cast<uint32_t>(i) * 512u;                  \
    if (o < (lim)) __stcg(reinterpret_cast<uint4*>((to) + o), v[i]);                   \
  }
  uint4 va[kCopyGroup], vb[kCopyGroup];
  int u = claim();
  CommUnit cu = pull_unit(p, u < total ? u : 0, upc, part_bytes);
  if (u < total) {
    const uint8_t* s0 = p.peer_a[cu.src] + cu.base;
    DDLB_LD_GROUP(va, s0, cu.begin, cu.end)
  }
  while (u < total) {
    const uint8_t* s0 = p.peer_a[cu.src] + cu.base;
    uint8_t* d0 = p.local_a + cu.base;
    const int un = claim();
    const CommUnit cn = pull_unit(p, un < total ? un : 0, upc, part_bytes);
    for (uint32_t off = cu.begin; off < cu.end; off += 2 * kCopyStep) {
      DDLB_LD_GROUP(vb, s0, off + kCopyStep, cu.end)
      DDLB_ST_GROUP(va, d0, off, cu.end)
      if (off + 2 * kCopyStep < cu.end) {
        DDLB_LD_GROUP(va, s0, off + 2 * kCopyStep, cu.end)
      } else if (un < total) {
        const uint8_t* s1 = p.peer_a[cn.src] + cn.base;     // next unit: first group in flight
        DDLB_LD_GROUP(va, s1, cn.begin, cn.end)
      }
      DDLB_ST_GROUP(vb, d0, off + kCopyStep, cu.end)
    }
    tr.ev(kEvUnitReady, static_cast<uint32_t>(u));      // stores issued
    __threadfence();
    __syncwarp();
    if (lane == 0)
      red_add_release_gpu(p.sig_local + kSigChunk + cu.src * chunks_per_part + cu.chunk, 1u);
    tr.ev(kEvUnitDone, static_cast<uint32_t>(u));       // fenced and announced
    u = un;
    cu = cn;
  }
#undef DDLB_LD_GROUP
#undef DDLB_ST_GROUP
  __syncwarp();
  // every byte this warp read has been stored already (data dependency): no fence needed to say
  // "I will not read your shard again"
  if (lane < p.world && lane != p.rank)
    red_add_relaxed_sys(p.sig_peer[lane] + kSigDone + p.rank, 1u);
}

// PUSH: every unit of the own shard is read once from local memory and stored to all d-1
// peers' gathered buffers (posted writes), fenced at system scope and announced with one
// remote counter bump per peer. kMulticast: ONE multimem.st per 16 bytes instead, replicated
// to every rank by the NVSwitch (egress 1x instead of (d-1)x).
template <bool kMulticast>
__device__ __forceinline__ void comm_warp_push(const TpGemmParams& p, int gw, int nw, int lane) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int upc = static_cast<int>((p.chunk_bytes + kPushUnit - 1) / kPushUnit);
  const int total = chunks_per_part * upc;
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;

  // a peer's gathered buffer may be overwritten once that peer has started this launch
  if (lane < p.world && lane != p.rank) wait_flag_ge<true>(p.sig_local + kSigReady + lane, p.epoch);
  __syncwarp();

  for (int u = gw; u < total; u += nw) {
    const int chunk = u / upc;
    const uint32_t begin = static_cast<uint32_t>(u - chunk * upc) * kPushUnit;
    const uint32_t end = min(begin + static_cast<uint32_t>(kPushUnit), p.chunk_bytes);
    const size_t base = static_cast<size_t>(p.rank) * part_bytes + static_cast<size_t>(chunk) * p.chunk_bytes;
    for (uint32_t off = begin + lane * 16; off < end; off += kCopyStep * 2) {
      uint4 v[2 * kCopyGroup];
#pragma unroll
      for (int i = 0; i < 2 * kCopyGroup; ++i) {
        const uint32_t o = off + i * 512;
        if (o < end) v[i] = __ldcs(reinterpret_cast<const uint4*>(p.local_a + base + o));
      }
      if constexpr (kMulticast) {
#pragma unroll
        for (int i = 0; i < 2 * kCopyGroup; ++i) {
          const uint32_t o = off + i * 512;
          if (o < end) multimem_st_v4(p.mc_ptr + base + o, v[i]);
        }
      } else {
        for (int d = 1; d < p.world; ++d) {
          int peer = p.rank + d;
          peer = peer >= p.world ? peer - p.world : peer;
          uint8_t* dst = const_cast<uint8_t*>(p.peer_a[peer]) + base;
#pragma unroll
          for (int i = 0; i < 2 * kCopyGroup; ++i) {
            const uint32_t o = off + i * 512;
            if (o < end) __stcs(reinterpret_cast<uint4*>(dst + o), v[i]);
          }
        }
      }
    }
    __threadfence_system();          // this lane's remote stores are performed
    __syncwarp();
    if (lane >= 1 && lane < p.world) {
      int peer = p.rank + lane;
      peer = peer >= p.world ? peer - p.world : peer;
      red_add_relaxed_sys(p.sig_peer[peer] + kSigChunk + p.rank * chunks_per_part + chunk, 1u);
    }
  }
}

// Store one staged epilogue chunk (32 rows x 128 bytes, 128-byte swizzled in shared memory)
// with 128-bit stores: per instruction the warp writes 4 rows x 128 contiguous bytes. Plain
// stores move ~690 GB/s to a peer over NVLink; TMA stores were measured at ~400-530 GB/s.
// kMulticast: multimem.st, one store reaches every rank's copy.
template <bool kMulticast>
__device__ __forceinline__ void store_chunk_st128(uint8_t* dst_base, size_t pitch_bytes, int row0,
                                                  int col_byte0, int col_limit_bytes,
                                                  uint32_t sbase, int lane) {
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int r = it * 4 + (lane >> 3);
    const int j = lane & 7;
    uint4 v;
    const uint32_t src = sbase + r * 128 + (static_cast<uint32_t>(j ^ (r & 7)) << 4);
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(src)
                 : "memory");
    const int colb = col_byte0 + j * 16;
    if (colb < col_limit_bytes) {
      uint8_t* dst = dst_base + static_cast<size_t>(row0 + r) * pitch_bytes + colb;
      if constexpr (kMulticast) multimem_st_v4(dst, v);
      else st_global_v4(dst, v);
    }
  }
}

// Sum the d staged copies of this lane's share of one unit (8 row groups x 16 bytes) and write
// the result. Sixteen independent 16-byte loads are in flight per lane whatever d is: the
// loads are numbered i = t * d + s (row group t, source s), so a running sum that is flushed
// whenever s wraps needs no register array indexed by a run-time value (one round at d = 2, two
// at d = 4, four at d = 8). ONE instantiation on purpose: several inlined (or called) variants
// make ptxas spill kilobytes in a kernel that sits at its 168-register cap.
// push: the d partials sit in the local staging slots [parity][src];
// pull: in slot [parity][rank] of every rank's staging (peer loads).
__device__ __forceinline__ void rs_sum_unit(const TpGemmParams& p, uint32_t parity, size_t slot_bytes,
                                            size_t unit_off, size_t pitch, int row0, int r_in, int colb,
                                            uint8_t* out, bool bf16) {
  const bool pull = p.rs_transport == kRsPull;
  const int d = p.world;
  const int groups = 16 / d;                       // row groups per round (d <= 8 -> >= 2)
  const int per_round = groups * d;
#pragma unroll 1
  for (int t0 = 0; t0 < 8; t0 += groups) {
    uint4 raw[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int t = i / d, s = i - t * d;
      if (i < per_round && t0 + t < 8) {
        const uint8_t* src = pull
            ? p.stage_base[s] + (static_cast<size_t>(parity) * d + p.rank) * slot_bytes
            : p.stage_base[p.rank] + (static_cast<size_t>(parity) * d + s) * slot_bytes;
        raw[i] = __ldcg(reinterpret_cast<const uint4*>(src + unit_off + ((t0 + t) * 4 + r_in) * 128));
      }
    }
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int t = i / d, s = i - t * d;
      if (i < per_round && t0 + t < 8) {
        if (p.elem == 4) {
          acc[0] += __uint_as_float(raw[i].x);
          acc[1] += __uint_as_float(raw[i].y);
          acc[2] += __uint_as_float(raw[i].z);
          acc[3] += __uint_as_float(raw[i].w);
        } else {
          const uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = unpack2(w[j], bf16);
            acc[2 * j] += f.x;
            acc[2 * j + 1] += f.y;
          }
        }
        if (s == d - 1) {                           // last source of row group t: flush
          uint4 res;
          if (p.elem == 4)
            res = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]),
                             __float_as_uint(acc[3]));
          else
            res = make_uint4(pack2(acc[0], acc[1], bf16), pack2(acc[2], acc[3], bf16),
                             pack2(acc[4], acc[5], bf16), pack2(acc[6], acc[7], bf16));
          const size_t o = static_cast<size_t>(row0 + (t0 + t) * 4 + r_in) * pitch + colb;
          st_global_v4(out + o, res);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// RS: the reducers. The own shard [part_rows, N] is cut into units of 32 rows x 128 bytes
// (the epilogue's staging granule). Unit u belongs to counter tile u / units_per_tile; a
// unit is ready when all 4 epilogue warps of all d ranks have announced that tile. Warps
// claim units from a queue head in the signal block (one head per epoch parity; the other
// one is reset for the next launch).
// ---------------------------------------------------------------------------
template <bool kTrace>
__device__ __forceinline__ void rs_reduce_loop(const TpGemmParams& p, int lane, bool bf16,
                                               Tracer<kTrace>& tr) {
  const uint32_t parity = p.epoch & 1u;
  const int segs = p.block_n * p.elem / 128;            // 128-byte column segments per n-block
  const int units_per_tile = 4 * segs;
  const int total = p.rs_tiles * units_per_tile;
  const size_t pitch = static_cast<size_t>(p.out_n) * p.elem;
  // staging is UNIT-major: unit u (32 rows x 128 bytes of one tile) is one contiguous 4 KiB block,
  // so a partial tile crosses NVLink as 512 contiguous bytes per warp instruction instead of four
  // 128-byte row segments 2 KiB apart; only the final output is row-major
  const size_t slot_bytes = static_cast<size_t>(total) * kEpiBufBytes;
  const uint32_t target = 4u * static_cast<uint32_t>(p.world) * p.epoch;
  uint32_t* head = p.sig_local + kSigClaim + parity;
  const int r_in = lane >> 3, j = lane & 7;
  uint8_t* out = p.out_raw[p.rank];

  while (true) {
    int u = 0;
    if (lane == 0) u = static_cast<int>(atomicAdd(head, 1u));
    u = __shfl_sync(0xffffffffu, u, 0);
    if (u >= total) break;
    const int tile = u / units_per_tile;
    const int sub = u - tile * units_per_tile;
    const int q = sub / segs, c = sub - q * segs;
    const int rb = tile / p.num_n_blocks, nb = tile - rb * p.num_n_blocks;
    const int row0 = rb * kBlockM + q * 32;
    const int colb = (nb * p.block_n) * p.elem + c * 128 + j * 16;
    if (lane == 0) wait_flag_ge<true>(p.sig_local + kSigChunk + tile, target);
    __syncwarp();
    tr.ev(kEvUnitReady, static_cast<uint32_t>(u));
    const bool active = colb < static_cast<int>(pitch);   // n-blocks may overhang N

    if (!active) {
      // nothing in this 16-byte column for this lane; stay converged for the next claim
    } else if (p.rs_transport == kRsPullMulticast) {
      // one load per 16 bytes returns the sum over all ranks, added in the NVSwitch
      const uint8_t* src = p.mc_ptr + (static_cast<size_t>(parity) * p.world + p.rank) * slot_bytes +
                           static_cast<size_t>(u) * kEpiBufBytes + j * 16;
      uint4 v[8];
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const size_t o = static_cast<size_t>(it * 4 + r_in) * 128;
        if (p.elem == 4) v[it] = multimem_ld_reduce_f32x4(src + o);
        else if (bf16) v[it] = multimem_ld_reduce_bf16x8(src + o);
        else v[it] = multimem_ld_reduce_f16x8(src + o);
      }
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const size_t o = static_cast<size_t>(row0 + it * 4 + r_in) * pitch + colb;
        st_global_v4(out + o, v[it]);
      }
    } else {
      rs_sum_unit(p, parity, slot_bytes, static_cast<size_t>(u) * kEpiBufBytes + j * 16, pitch, row0, r_in,
                  colb, out, bf16);
    }
    __syncwarp();
    tr.ev(kEvUnitDone, static_cast<uint32_t>(u));
  }
}

// ---------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------
// kF32: fp32 inputs and outputs through kind::tf32. A 128-byte swizzle row then holds 32
// elements and one MMA covers 8 of them, so the byte geometry of the operand ring is
// identical to the 16-bit case. kSplit (with kF32): 3xTF32 - warps 8..11 write the low-order
// parts lo = x - tf32(x) of both operand tiles next to them and the issuer runs three MMAs
// per k-step (hi*hi + hi*lo + lo*hi), which restores ~fp32 accuracy
// (BASELINE.json north star; the reference's default dtype is float32,
// ddlb/primitives/TPColumnwise/tp_columnwise.py:32).
// kPairs = 2: a cluster of FOUR CTAs = two MMA pairs that compute two neighbouring n-blocks of the
// same 256 rows; the A tile is loaded once from L2 and multicast to both pairs. Measured (ncu):
// with one pair per cluster the operand stream from L2 runs at the fabric's cap (~6.6 KB/clk,
// 1.07 GB per 100 us launch) and the tensor pipe idles 22 % of its active cycles.
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, bool kSplit = false,
          bool kTrace = false, int kPairs = 1, bool kWide = false>
__device__ __forceinline__ void tp_gemm_body(const TpGemmParams& p, const int cluster,
                                             const int nclusters) {
  using Cfg = TpGemmCfg<kCtaGroup, kBlockN, kSplit, kWide>;
  constexpr int kNSub = Cfg::kNSub;
  static_assert(!kWide || (kPairs == 1 && !kF32), "wide tiles: one pair per cluster, 16-bit operands");
  static_assert(!kSplit || kF32, "the hi/lo split exists for fp32 only");
  static_assert(kPairs == 1 || (kPairs == 2 && kCtaGroup == 2 && !kSplit), "two pairs need cta_group 2");
  constexpr int kClusterCtas = kCtaGroup * kPairs;
  constexpr int kElem = kF32 ? 4 : 2;                 // bytes per element
  constexpr int kRowElems = 128 / kElem;              // elements per 128-byte swizzle row
  constexpr int kKPerStage = kRowElems;               // k elements per pipeline stage
  constexpr int kKPerMma = 32 / kElem;                // 16 (f16/bf16) or 8 (tf32)
  const int kStages = p.stages;                       // runtime ring depth

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_ops = smem;                                   // [stage]{A, B[, A_lo, B_lo]}
  uint8_t* smem_epi = smem + kStages * Cfg::kStageBytes;      // [warp][2][4 KB]
  uint8_t* smem_bar = smem_epi + kEpiBytes;
  const uint32_t bar_full = smem_u32(smem_bar);               // [kStages] TMA landed
  const uint32_t bar_empty = bar_full + 8 * 8;                // [kStages] MMAs retired
  const uint32_t bar_split = bar_empty + 8 * 8;               // [kStages] lo parts written (kSplit)
  const uint32_t bar_tfull = bar_split + 8 * 8;               // [2]
  const uint32_t bar_tempty = bar_tfull + 2 * 8;              // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem_bar + 28 * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = static_cast<int>(blockDim.x >> 5);
  const int n_aux = 2 + (nwarps - 8);                         // warps 2, 3, 8..
  const int aux = warp == 2 ? 0 : warp == 3 ? 1 : warp >= 8 ? warp - 6 : -1;
  const uint32_t cluster_rank = kClusterCtas > 1 ? cluster_ctarank() : 0u;   // 0 .. kClusterCtas-1
  const uint32_t cta_rank = cluster_rank & static_cast<uint32_t>(kCtaGroup - 1);   // rank in the MMA pair
  const uint32_t pair = kCtaGroup == 2 ? cluster_rank >> 1 : 0u;             // which pair of the cluster
  const uint32_t leader_rank = cluster_rank & ~1u;                           // even CTA of this pair
  const int cta = cluster * kClusterCtas + static_cast<int>(cluster_rank);
  const int nctas = nclusters * kClusterCtas;
  const bool bf16 = ((p.idesc >> 7) & 7u) == 1u;

  // ------------------------------------------------------------------ setup
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA);
    prefetch_tmap(&p.tmB);
    if (p.mode != kModeRS) prefetch_tmap(&p.tmOut[p.rank]);   // RS stores without TMA
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < 8; ++s) {
      mbar_init(bar_full + s * 8, 1);
      mbar_init(bar_empty + s * 8, kPairs);                   // one commit per pair that read it
      mbar_init(bar_split + s * 8, 4 * kCtaGroup);            // one arrival per split warp
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_tfull + a * 8, 1);
      mbar_init(bar_tempty + a * 8, 4 * kCtaGroup);
    }
    fence_mbar_init();
  }
  __syncwarp();
  if (warp == 2) {
    tmem_alloc<kCtaGroup>(smem_u32(tmem_ptr_smem), Cfg::kTmemCols);
    tmem_relinquish<kCtaGroup>();
  }
  tc_fence_before();
  if constexpr (kClusterCtas > 1) cluster_sync(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (p.pdl) {
    // everything above touched only shared memory, TMEM and the kernel parameters; from here on
    // global memory is read and written, so the previous kernel of the stream must be complete
    griddep_launch_dependents();
    griddep_wait();
  }

  const int num_tiles = p.num_m_blocks * p.sched_n_blocks;

  if (warp == 0 && lane == 0) {
    // -------------------------------------------------------- TMA producer
    const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
    const uint32_t chunk_target = p.epoch * ((p.chunk_bytes + kPushUnit - 1) / kPushUnit);
    const uint32_t full_leader = kCtaGroup == 2 ? mapa(bar_full, leader_rank) : bar_full;
    int stage = 0;
    uint32_t phase = 0;
    Tracer<kTrace> tr(p, cta, 0);
    tr.ev(kEvStart, 0);
    for (int t = cluster; t < num_tiles; t += nclusters) {
      const TileCoord tc = tile_coord(p, t);
      tr.ev(kEvTileBegin, static_cast<uint32_t>(t));
      if (p.mode == kModeAGBefore && tc.part != p.rank) {
        const uint32_t* flag = p.sig_local + kSigChunk + tc.part * chunks_per_part + tc.chunk;
        if (p.ag_transport == kAgPull) wait_flag_ge<false>(flag, chunk_target);
        else wait_flag_ge<true>(flag, chunk_target);       // counter bumped over NVLink
        fence_proxy_async();      // flag acquired by the generic proxy, data read by TMA
        tr.ev(kEvFlagWaited, static_cast<uint32_t>(t));
      }
      const int m0 = tc.m_blk * Cfg::kTileM + static_cast<int>(cta_rank) * kBlockM;
      const int n0 = (tc.n_blk * kPairs + static_cast<int>(pair)) * kNSub * kBlockN +
                     static_cast<int>(cta_rank) * Cfg::kLoadN;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(bar_empty + stage * 8, phase ^ 1u);
        uint8_t* sa = smem_ops + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kStageA;
        // 3xTF32: each CTA watches its own half land (its split warps transform it); the
        // issuer then waits for the split barrier only. Otherwise both halves signal the
        // leader's barrier.
        if constexpr (kSplit) mbar_arrive_expect_tx(bar_full + stage * 8, Cfg::kStageOps);
        else if (kCtaGroup == 1 || cta_rank == 0)
          mbar_arrive_expect_tx(bar_full + stage * 8, Cfg::kStageOps * kCtaGroup);
        const uint32_t fb = (kSplit ? bar_full : full_leader) + stage * 8;
        if constexpr (kPairs == 2) {
          // this CTA fetches HALF of its 128 A rows and multicasts them to the CTA with the same
          // rank in the other pair (the other half comes from there): A crosses L2 -> SM once
          const uint16_t mask = static_cast<uint16_t>((1u << cta_rank) | (1u << (2u + cta_rank)));
          tma_load_2d_multicast(smem_u32(sa) + pair * (Cfg::kStageA / 2), &p.tmA, fb, kb * kKPerStage,
                                m0 + static_cast<int>(pair) * (kBlockM / 2), mask);
        } else {
          tma_load_2d<kCtaGroup>(smem_u32(sa), &p.tmA, fb, kb * kKPerStage, m0);
        }
#pragma unroll
        for (int sub = 0; sub < kNSub; ++sub) {
          const int ns = n0 + sub * kBlockN;
          const uint32_t sbs = smem_u32(sb) + sub * Cfg::kStageB;
          if constexpr (kBNMajor) tma_load_3d<kCtaGroup>(sbs, &p.tmB, fb, 0, kb * kKPerStage, ns / kRowElems);
          else tma_load_2d<kCtaGroup>(sbs, &p.tmB, fb, kb * kKPerStage, ns);
        }
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1 && lane == 0 && cta_rank == 0) {
    // -------------------------------------------------------- MMA issuer
    // K-major operand: 8-row groups 1024 B apart, k-step = 32 B inside the swizzle row.
    // N-major B: 64-column blocks 8 KB apart (LBO), 8-k-row groups 1024 B apart (SBO),
    //            k-step = 16 rows * 128 B.
    constexpr uint32_t kBLbo = kBNMajor ? kKPerStage * 128 : 16;   // next 128-byte column block
    constexpr uint32_t kBStep = kBNMajor ? (kKPerMma * 128) >> 4 : 32 >> 4;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    Tracer<kTrace> tr(p, cta, 1);
    for (int t = cluster; t < num_tiles; t += nclusters) {
      mbar_wait(bar_tempty + acc * 8, acc_phase ^ 1u);
      if constexpr (kWide) mbar_wait(bar_tempty + 8, acc_phase ^ 1u);    // both accumulators drained
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kBlockN);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        if constexpr (kSplit) mbar_wait(bar_split + stage * 8, phase);   // implies both halves landed
        else mbar_wait(bar_full + stage * 8, phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem_ops + stage * Cfg::kStageBytes);
        const uint64_t adesc = make_smem_desc_sw128(sa, 16, 1024);
        const uint64_t bdesc = make_smem_desc_sw128(sa + Cfg::kStageA, kBLbo, 1024);
#pragma unroll
        for (int k = 0; k < kKPerStage / kKPerMma; ++k) {
          const uint64_t ak = adesc + static_cast<uint64_t>(k * (32 >> 4));
          const uint64_t bk = bdesc + static_cast<uint64_t>(k * kBStep);
          if constexpr (kF32) {
            umma_tf32<kCtaGroup>(tmem_d, ak, bk, p.idesc, static_cast<uint32_t>((kb | k) != 0));
            if constexpr (kSplit) {
              constexpr uint64_t kLo = static_cast<uint64_t>(Cfg::kStageOps >> 4);
              umma_tf32<kCtaGroup>(tmem_d, ak, bk + kLo, p.idesc, 1u);   // hi(A) * lo(B)
              umma_tf32<kCtaGroup>(tmem_d, ak + kLo, bk, p.idesc, 1u);   // lo(A) * hi(B)
            }
          } else {
            umma_f16<kCtaGroup>(tmem_d, ak, bk, p.idesc, static_cast<uint32_t>((kb | k) != 0));
            if constexpr (kWide)          // second n-block: same A, next B tile, next accumulator
              umma_f16<kCtaGroup>(tmem_d + kBlockN, ak, bk + static_cast<uint64_t>(Cfg::kStageB >> 4),
                                  p.idesc, static_cast<uint32_t>((kb | k) != 0));
          }
        }
        // smem slot reusable when the MMAs retire: every CTA of the cluster hears it (with two
        // pairs a CTA's slot is also written by the other pair's multicast)
        umma_commit<kCtaGroup>(bar_empty + stage * 8, static_cast<uint16_t>((1u << kClusterCtas) - 1u));
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      umma_commit<kCtaGroup>(bar_tfull + acc * 8,          // accumulator ready: this pair's epilogues
                             static_cast<uint16_t>(0x3u << leader_rank));
      if constexpr (kWide)
        umma_commit<kCtaGroup>(bar_tfull + 8, static_cast<uint16_t>(0x3u << leader_rank));
      tr.ev(kEvTileIssued, static_cast<uint32_t>(t));
      if constexpr (kWide) acc_phase ^= 1u;                  // acc stays 0: both slots used per tile
      else if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (kSplit && warp >= 8) {
    // -------------------------------------------------------- 3xTF32 split warps
    // The tensor core reads the upper 19 bits of an fp32 operand. lo = x - that value is
    // exact in fp32 and itself fits a tf32 mantissa to ~2^-21 relative to x. The lo tiles use
    // the same swizzled byte positions as the hi tiles, so the transform is elementwise.
    if constexpr (kSplit) {
      const int sw = warp - 8;                               // 0..3
      const uint32_t split_leader = kCtaGroup == 2 ? mapa(bar_split, 0) : bar_split;
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster; t < num_tiles; t += nclusters) {
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(bar_full + stage * 8, phase);            // this CTA's half has landed
          const uint32_t base = smem_u32(smem_ops + stage * Cfg::kStageBytes);
          constexpr int kVecs = Cfg::kStageOps / 16;         // 16-byte vectors in A + B
          for (int v = sw * 32 + lane; v < kVecs; v += 128) {
            // shared-state-space accesses (a generic float4* compiles to LD.E / ST.E here)
            uint32_t x0, x1, x2, x3;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(base + v * 16) : "memory");
            const uint32_t l0 = __float_as_uint(__uint_as_float(x0) - __uint_as_float(x0 & 0xffffe000u));
            const uint32_t l1 = __float_as_uint(__uint_as_float(x1) - __uint_as_float(x1 & 0xffffe000u));
            const uint32_t l2 = __float_as_uint(__uint_as_float(x2) - __uint_as_float(x2 & 0xffffe000u));
            const uint32_t l3 = __float_as_uint(__uint_as_float(x3) - __uint_as_float(x3 & 0xffffe000u));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                         :: "r"(base + Cfg::kStageOps + v * 16), "r"(l0), "r"(l1), "r"(l2), "r"(l3) : "memory");
          }
          fence_proxy_async_smem();                          // generic writes -> tensor-core reads
          __syncwarp();
          if (lane == 0) {
            if (kCtaGroup == 2 && cta_rank != 0) mbar_arrive_cluster(split_leader + stage * 8);
            else mbar_arrive(bar_split + stage * 8);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (aux >= 0 && p.mode == kModeAGBefore && p.world > 1) {
    // -------------------------------------------------------- comm warps (every CTA)
    if (cta == 0 && aux == 1 && lane < p.world && lane != p.rank)
      st_release_sys(p.sig_peer[lane] + kSigReady + p.rank, p.epoch);   // my previous launch is over
    // 2 warps per CTA fill the links, more oversubscribe them (see comm_warp_pull; push and
    // multimem.st behave the same: 213 vs 240 us and 320 vs 352 us at 2 ranks); warps 8.. split
    // operands in 3xTF32
    int na = 2;                                              // every transport: measured, see above
    if (p.comm_warps > 0) na = p.comm_warps < n_aux ? p.comm_warps : n_aux;
    if (kSplit && na > 2) na = 2;
    if (cta == 0 && aux == 0 && lane == 0) p.sig_local[kSigClaim + ((p.epoch & 1u) ^ 1u)] = 0u;
    if (aux < na) {
      Tracer<kTrace> tr(p, (warp == 2 && lane == 0) ? cta : -1, 3);
      tr.ev(kEvStart, 0);
      const int gw = cta * na + aux, nw = nctas * na;
      if (p.ag_transport == kAgPull) comm_warp_pull<kTrace>(p, lane, tr);
      else if (p.ag_transport == kAgPushMulticast) comm_warp_push<true>(p, gw, nw, lane);
      else comm_warp_push<false>(p, gw, nw, lane);
      tr.ev(kEvEnd, 0);
      if (p.ag_transport == kAgPull && cta == 0 && aux == 0) {
        // every peer has finished reading this rank's shard: the caller may overwrite it as
        // soon as this kernel is complete (same guarantee as ncclAllGather's send buffer)
        if (lane < p.world && lane != p.rank)
          wait_flag_ge<true>(p.sig_local + kSigDone + lane, p.epoch * static_cast<uint32_t>(nw));
      }
    }
  } else if (aux >= 0 && p.mode == kModeRS && !(kSplit && warp >= 8)) {
    // -------------------------------------------------------- reducers (every CTA)
    if (cta == 0 && aux == 0 && lane == 0) p.sig_local[kSigClaim + ((p.epoch & 1u) ^ 1u)] = 0u;
    Tracer<kTrace> tr(p, (warp == 2 && lane == 0) ? cta : -1, 3);
    tr.ev(kEvStart, 0);
    rs_reduce_loop<kTrace>(p, lane, bf16, tr);
    tr.ev(kEvEnd, 0);
  } else if (warp >= 4 && warp < 8) {
    // -------------------------------------------------------- epilogue
    const int q = warp & 3;                                 // TMEM lane quadrant
    uint8_t* ebuf = smem_epi + q * 2 * kEpiBufBytes;
    const uint32_t tempty_leader = kCtaGroup == 2 ? mapa(bar_tempty, leader_rank) : bar_tempty;
    int buf = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    const uint32_t parity = p.epoch & 1u;
    const size_t pitch = static_cast<size_t>(p.out_n) * kElem;
    Tracer<kTrace> tr(p, (warp == 4 && lane == 0) ? cta : -1, 2);
    for (int t = cluster; t < num_tiles; t += nclusters) {
      const TileCoord tc = tile_coord(p, t);
      const int row_cta = tc.m_blk * Cfg::kTileM + static_cast<int>(cta_rank) * kBlockM;
      const int row0 = row_cta + q * 32;
      // wide tiles: the two n-blocks of the tile are drained one after the other (accumulator 0, 1)
#pragma unroll 1
      for (int sub = 0; sub < kNSub; ++sub) {
        const int n_blk = (tc.n_blk * kPairs + static_cast<int>(pair)) * kNSub + sub;
        const int n0 = n_blk * kBlockN;
        mbar_wait(bar_tfull + acc * 8, acc_phase);
        tc_fence_after();
        tr.ev(kEvAccReady, static_cast<uint32_t>(t));

        int out_row = row0;
        int cta_tile = 0;
        uint8_t* rs_dst = nullptr;
        if (p.mode == kModeRS) {
          // row inside the owner's shard; staging slot [parity][src] at the owner (push) or
          // [parity][owner] locally (pull)
          out_row = row0 - tc.part * p.part_rows;
          cta_tile = ((row_cta - tc.part * p.part_rows) / kBlockM) * p.num_n_blocks + n_blk;
          const size_t slot_bytes = static_cast<size_t>(p.rs_tiles) * (4 * (kBlockN / kRowElems)) * kEpiBufBytes;
          rs_dst = p.rs_transport == kRsPush
                       ? p.stage_base[tc.part] + (static_cast<size_t>(parity) * p.world + p.rank) * slot_bytes
                       : p.stage_base[p.rank] + (static_cast<size_t>(parity) * p.world + tc.part) * slot_bytes;
        } else if (p.mode == kModeAGAfter) {
          out_row = p.rank * p.part_rows + row0;
        }

#pragma unroll 1
        for (int c = 0; c < kBlockN / kRowElems; ++c) {
          // 32 rows x 128 bytes of output per step: this lane's row, as 32 words
          uint32_t packed[32];
          if constexpr (kF32) {
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                   static_cast<uint32_t>(acc * kBlockN + c * 32);
            tmem_ld_32x32b_x32(taddr, &packed[0]);
            tmem_ld_wait();
            if (p.act != kActNone) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                packed[i] = __float_as_uint(apply_act(__uint_as_float(packed[i]), p.act));
            }
          } else {
            uint32_t v[64];
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                   static_cast<uint32_t>(acc * kBlockN + c * 64);
            tmem_ld_32x32b_x32(taddr, &v[0]);
            tmem_ld_32x32b_x32(taddr + 32, &v[32]);
            tmem_ld_wait();
            if (p.act != kActNone) {
#pragma unroll
              for (int i = 0; i < 64; ++i) v[i] = __float_as_uint(apply_act(__uint_as_float(v[i]), p.act));
            }
#pragma unroll
            for (int i = 0; i < 32; ++i)
              packed[i] = pack2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), bf16);
          }

          if (lane == 0) bulk_wait_read<1>();     // the TMA store that last used this buffer has read it
          __syncwarp();
          const uint32_t sbase = smem_u32(ebuf + buf * kEpiBufBytes);
          const uint32_t srow = sbase + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t dst = srow + (static_cast<uint32_t>(j ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(packed[4 * j]),
                         "r"(packed[4 * j + 1]), "r"(packed[4 * j + 2]), "r"(packed[4 * j + 3])
                         : "memory");
          }
          const int col0 = n0 + c * kRowElems;
          const int colb = col0 * kElem;
          if (p.mode == kModeRS) {
            __syncwarp();
            // unit-major staging: this chunk is unit (cta_tile, q, c), 4 KiB contiguous
            store_chunk_st128<false>(rs_dst + (static_cast<size_t>(cta_tile * 4 + q) * (kBlockN / kRowElems) + c) *
                                                  kEpiBufBytes,
                                     128, 0, 0, 128, sbase, lane);
            __syncwarp();            // every lane has read the staged rows before they are reused
          } else if (p.mode == kModeAGAfter && p.epi_transport == kEpiMulticast) {
            __syncwarp();
            store_chunk_st128<true>(p.mc_ptr, pitch, out_row, colb, static_cast<int>(pitch), sbase, lane);
            __syncwarp();
          } else if (p.mode == kModeAGAfter && p.epi_transport == kEpiSt128) {
            fence_proxy_async_smem();
            __syncwarp();
            for (int d = 0; d < p.world; ++d)
              if (d != p.rank)
                store_chunk_st128<false>(p.out_raw[d], pitch, out_row, colb, static_cast<int>(pitch), sbase, lane);
            if (lane == 0) {
              tma_store_2d(&p.tmOut[p.rank], sbase, col0, out_row);
              bulk_commit();
            }
            __syncwarp();
          } else {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (p.mode == kModeAGAfter) {
                for (int d = 0; d < p.world; ++d) tma_store_2d(&p.tmOut[d], sbase, col0, out_row);
              } else {
                tma_store_2d(&p.tmOut[p.rank], sbase, col0, out_row);
              }
              bulk_commit();
            }
          }
          buf ^= 1;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kCtaGroup == 2 && cta_rank != 0) mbar_arrive_cluster(tempty_leader + acc * 8);
          else mbar_arrive(bar_tempty + acc * 8);
        }
        tr.ev(kEvTileStored, static_cast<uint32_t>(t));
        if (p.mode == kModeRS) {
          // announce the tile to its owner as soon as this warp's stores have been performed
          // (the accumulator is already released, only this warp waits for the round trip)
          // push to a peer: the stores crossed NVLink, only a system-scope fence knows when they
          // have landed. Stores into LOCAL staging (own tiles; every tile of the pull transports)
          // are complete for every reader, local or remote, once they are in this GPU's L2: a
          // GPU-scope fence (~1 us instead of ~5 us).
          if (p.rs_transport == kRsPush && tc.part != p.rank) __threadfence_system();
          else __threadfence();
          __syncwarp();
          if (lane == 0) red_add_relaxed_sys(p.sig_peer[tc.part] + kSigChunk + cta_tile, 1u);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
    if (lane == 0) bulk_wait<0>();
    tr.ev(kEvEnd, 0);
    if (p.mode == kModeAGAfter && p.world > 1) {
      // every CTA tells every peer that its share of C has landed, then waits for
      // all CTAs of all peers so that C is complete when this kernel ends
      if (p.epi_transport != kEpiTma) __threadfence_system();
      if (lane == 0) { fence_proxy_async(); fence_acq_rel_sys(); }
      named_bar_sync(1, 128);
      if (warp == 4 && lane < p.world && lane != p.rank) {
        red_add_relaxed_sys(p.sig_peer[lane] + kSigDone + p.rank, 1u);
        wait_flag_ge<true>(p.sig_local + kSigDone + lane, p.epoch * static_cast<uint32_t>(nctas));
      }
    }
    // RS: help the aux warps with whatever is left of the reduction
    if (p.mode == kModeRS) {
      Tracer<false> none(p, -1, 3);
      rs_reduce_loop<false>(p, lane, bf16, none);
    }
  }

  // ---------------------------------------------------------------- teardown
  __syncwarp();
  tc_fence_before();
  if constexpr (kClusterCtas > 1) cluster_sync(); else __syncthreads();
  if (warp == 2) tmem_dealloc<kCtaGroup>(tmem_base, Cfg::kTmemCols);
}

// One rank per process / GPU: parameters in the kernel parameter space.
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, bool kSplit = false,
          bool kTrace = false, int kPairs = 1, bool kWide = false>
__global__ void __launch_bounds__(kMaxThreads, 1)
tp_gemm_kernel(const __grid_constant__ TpGemmParams p) {
  constexpr bool kClustered = kCtaGroup * kPairs > 1;
  const int cluster = kClustered ? static_cast<int>(cluster_id_x()) : static_cast<int>(blockIdx.x);
  const int nclusters = kClustered ? static_cast<int>(num_clusters_x()) : static_cast<int>(gridDim.x);
  tp_gemm_body<kCtaGroup, kBlockN, kBNMajor, kF32, kSplit, kTrace, kPairs, kWide>(p, cluster, nclusters);
}

// Several ranks emulated by ONE launch on one GPU (tests): the grid is split into
// `world` equal groups of clusters, group r runs rank r's parameter block. All
// CTAs of one launch are co-resident (grid <= SM count), so ranks that wait on
// each other make progress, which separate launches on one GPU cannot guarantee.
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, bool kSplit = false,
          int kPairs = 1, bool kWide = false>
__global__ void __launch_bounds__(kMaxThreads, 1)
tp_gemm_emulated_kernel(const TpGemmParams* __restrict__ plist, int clusters_per_rank) {
  constexpr bool kClustered = kCtaGroup * kPairs > 1;
  const int cluster = kClustered ? static_cast<int>(cluster_id_x()) : static_cast<int>(blockIdx.x);
  const int vr = cluster / clusters_per_rank;
  tp_gemm_body<kCtaGroup, kBlockN, kBNMajor, kF32, kSplit, false, kPairs, kWide>(
      plist[vr], cluster - vr * clusters_per_rank, clusters_per_rank);
}

}  // namespace ddlb

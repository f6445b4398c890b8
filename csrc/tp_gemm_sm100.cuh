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
// TPRowwise/fuser.py:15-169): their "algorithm/s/order" knobs are tile orders here.
//
// CTA roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one lane),
// warp 2 = TMEM allocator, warps 4..7 = epilogue (one TMEM lane quadrant each).
// In AG_BEFORE the first `n_comm_clusters` clusters do not compute: they pull the
// peers' row shards over NVLink with bulk-async copies (global -> smem ring -> local
// global) and publish per-chunk arrival counters that the GEMM producers wait on.
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
constexpr int kNumThreads = 256;
constexpr int kMaxWorld = 8;
constexpr int kEpiBufBytes = 32 * 128;          // 32 rows x 64 cols x 2 B per warp buffer
constexpr int kEpiBytes = 4 * 2 * kEpiBufBytes;  // 4 warps x double buffer
constexpr int kSmemBudget = 193 * 1024;         // operand ring
constexpr int kCommPiece = 16 * 1024;           // bytes per bulk copy of a comm CTA
constexpr int kCommRing = 12;
constexpr int kCommLag = 4;
constexpr int kWarpPiece = 8 * 1024;         // comm warp in every CTA: piece, ring depth, store lag
constexpr int kWarpRing = 4;
constexpr int kWarpLag = 2;
constexpr int kWarpRingBytes = kWarpRing * kWarpPiece;

enum : int { kModePlain = 0, kModeAGBefore = 1, kModeAGAfter = 2, kModeRS = 3 };

// word offsets inside a rank's signal block
constexpr int kSigReady = 0;    // [src]   epoch written by src when its kernel started
constexpr int kSigDone = 8;     // [src]   AG_AFTER: += 1 per CTA of src whose stores landed
constexpr int kSigChunk = 16;   // [chunk] AG_BEFORE: += 1 per comm CTA slice copied
// RS tile counters follow the chunk flags: [src][cta_tile] += 1 per epilogue warp

struct TpGemmParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  CUtensorMap tmOut[kMaxWorld];
  int mode;
  int rank, world;
  int num_m_blocks;       // cluster tiles along M (of the A operand map)
  int num_n_blocks;
  int num_k_blocks;
  int group_m;            // rasterisation: m-blocks per group (L2 reuse)
  int mblks_per_part;     // cluster tiles per rank shard (AG_BEFORE / RS), else num_m_blocks
  int chunk_mblks;        // cluster tiles per chunk (divides mblks_per_part)
  int rot0;               // 0: own part first (AG), 1: own part last (RS)
  int n_comm_clusters;    // AG_BEFORE only
  int part_rows;          // rows per shard (m / world)
  int out_n;              // N (elements per output row)
  uint32_t idesc;
  uint32_t epoch;
  // AG_BEFORE
  const uint8_t* peer_a[kMaxWorld];   // peers' gathered-A buffers (peer mapped)
  uint8_t* local_a;                   // this rank's gathered-A buffer
  uint32_t chunk_bytes;               // bytes of one chunk (chunk rows * K * 2)
  uint32_t slice_bytes;               // chunk_bytes / comm CTAs
  int comm_push;                      // dedicated comm CTAs: 1 = push own shard, 0 = pull peers' shards
  int comm_in_cta;                    // 1: no dedicated CTAs, warp 3 of EVERY CTA pulls pieces
  int stages;                         // operand ring depth actually used (<= Cfg::kStages)
  // RS
  const uint8_t* stage_local;         // [2][world][part_rows][N] partials pushed by peers
  int rs_sig_off;                     // word offset of the RS counters in a signal block
  int rs_tiles_per_src;               // (part_rows / 128) * num_n_blocks
  // signals
  uint32_t* sig_local;
  uint32_t* sig_peer[kMaxWorld];
  // optional device trace (kTrace kernels only): [cta][role][kTraceSlots] x {tag, arg, time ns}
  uint4* trace;
  // raw output addresses (same buffers as tmOut[]), used by the register-store push variant
  uint8_t* out_raw[kMaxWorld];
};

template <int kCtaGroup, int kBlockN>
struct TpGemmCfg {
  static constexpr int kTileM = kBlockM * kCtaGroup;
  static constexpr int kLoadN = kBlockN / kCtaGroup;            // B columns staged per CTA
  static constexpr int kStageA = kBlockM * kBlockK * 2;
  static constexpr int kStageB = kLoadN * kBlockK * 2;
  static constexpr int kStageBytes = kStageA + kStageB;
  static constexpr int kStagesRaw = kSmemBudget / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = 2 * kBlockN;
  static constexpr int kBarBytes = 512;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kEpiBytes + kBarBytes;
  static_assert(kStages >= 3, "operand ring too shallow");
  static_assert(kCommRing * kCommPiece <= kStages * kStageBytes, "comm ring does not fit");
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
  const int per_group = p.group_m * p.num_n_blocks;
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
  int part = p.rank + p.rot0 + pi;
  part = part >= p.world ? part - p.world : part;
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

// v[0..8) hold fp32 bit patterns; add 8 packed 16-bit values to them in fp32.
__device__ __forceinline__ void add_packed8(uint32_t* v, uint4 raw, bool bf16) {
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f;
    if (bf16) {
      __nv_bfloat162 h;
      memcpy(&h, &w[i], 4);
      f = __bfloat1622float2(h);
    } else {
      __half2 h;
      memcpy(&h, &w[i], 4);
      f = __half22float2(h);
    }
    v[2 * i] = __float_as_uint(__uint_as_float(v[2 * i]) + f.x);
    v[2 * i + 1] = __float_as_uint(__uint_as_float(v[2 * i + 1]) + f.y);
  }
}

// ---------------------------------------------------------------------------
// communication CTAs (AG_BEFORE): pull peers' chunks into the local gathered-A
// ---------------------------------------------------------------------------
__device__ __forceinline__ void comm_cta_loop(const TpGemmParams& p, uint8_t* ring, uint32_t bar_base,
                                              int ci, int nci) {
  // bar_base: kCommRing mbarriers (8 B each), already initialised with count 1
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int n_chunks = chunks_per_part * (p.world - 1);
  const uint32_t pieces_per_slice = (p.slice_bytes + kCommPiece - 1) / kCommPiece;
  const int total = n_chunks * static_cast<int>(pieces_per_slice);
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;

  if (ci == 0 && threadIdx.x < static_cast<unsigned>(p.world) &&
      static_cast<int>(threadIdx.x) != p.rank) {
    // tell every peer that this rank's shard is final for this epoch
    st_release_sys(p.sig_peer[threadIdx.x] + kSigReady + p.rank, p.epoch);
  }
  if (threadIdx.x != 0) return;

  uint32_t ready_mask = 1u << p.rank;
  auto piece_addr = [&](int i, size_t& off, uint32_t& bytes, int& part, int& chunk_in_part,
                        bool& last) {
    const int c = i / static_cast<int>(pieces_per_slice);        // chunk in fetch order
    const uint32_t pc = i - c * pieces_per_slice;
    const int stage = c / (p.world - 1);
    const int pi = c - stage * (p.world - 1);
    int src = p.rank + 1 + pi;
    src = src >= p.world ? src - p.world : src;
    part = src;
    chunk_in_part = stage;
    const uint32_t in_slice = pc * kCommPiece;
    bytes = min(static_cast<uint32_t>(kCommPiece), p.slice_bytes - in_slice);
    off = static_cast<size_t>(src) * part_bytes + static_cast<size_t>(stage) * p.chunk_bytes +
          static_cast<size_t>(ci) * p.slice_bytes + in_slice;
    last = (pc + 1 == pieces_per_slice);
  };
  auto issue_load = [&](int i) {
    size_t off; uint32_t bytes; int part, cip; bool last;
    piece_addr(i, off, bytes, part, cip, last);
    if (!(ready_mask & (1u << part))) {
      wait_flag_ge<true>(p.sig_local + kSigReady + part, p.epoch);
      ready_mask |= 1u << part;
    }
    const int slot = i % kCommRing;
    const uint32_t bar = bar_base + slot * 8;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_load(smem_u32(ring + slot * kCommPiece), p.peer_a[part] + off, bytes, bar);
  };

  const int prefill = total < kCommRing ? total : kCommRing;
  for (int i = 0; i < prefill; ++i) issue_load(i);
  // Store pipeline kCommLag groups deep. Group i is piece i's store (an empty group in the
  // tail), so after committing group i "at most kCommLag pending" means piece i-kCommLag is
  // in local memory: publish its chunk counter if it closed a chunk, recycle its ring slot.
  for (int i = 0; i < total + kCommLag; ++i) {
    if (i < total) {
      size_t off; uint32_t bytes; int part, cip; bool last;
      piece_addr(i, off, bytes, part, cip, last);
      const int slot = i % kCommRing;
      mbar_wait(bar_base + slot * 8, (i / kCommRing) & 1);
      bulk_store(p.local_a + off, smem_u32(ring + slot * kCommPiece), bytes);
    }
    bulk_commit();
    const int j = i - kCommLag;
    if (j >= 0) {
      bulk_wait<kCommLag>();
      size_t off; uint32_t bytes; int part, cip; bool last;
      piece_addr(j, off, bytes, part, cip, last);
      if (last) {
        fence_proxy_async();
        red_add_release_gpu(p.sig_local + kSigChunk + part * chunks_per_part + cip, 1u);
      }
      if (j + kCommRing < total) issue_load(j + kCommRing);
    }
  }
  bulk_wait<0>();
}

// ---------------------------------------------------------------------------
// communication CTAs (AG_BEFORE, push transport): every rank writes its own shard
// into all peers' gathered-A buffers. A piece is read ONCE from local memory into
// shared memory and stored to the d-1 peers from there; remote stores are posted,
// so a slot is recycled as soon as the stores have read it (short lag) while the
// chunk counters on the peers are bumped only when the stores have completed
// (long lag) - neither waits for a round trip on the critical path.
// ---------------------------------------------------------------------------
constexpr int kPushReadLag = 3;
constexpr int kPushDoneLag = 8;

__device__ __forceinline__ void comm_cta_push_loop(const TpGemmParams& p, uint8_t* ring,
                                                   uint32_t bar_base, int ci, int nci) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const uint32_t pieces_per_slice = (p.slice_bytes + kCommPiece - 1) / kCommPiece;
  const int total = chunks_per_part * static_cast<int>(pieces_per_slice);
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;

  if (ci == 0 && threadIdx.x < static_cast<unsigned>(p.world) &&
      static_cast<int>(threadIdx.x) != p.rank) {
    // this rank's previous launch is over: peers may overwrite its gathered buffer
    st_release_sys(p.sig_peer[threadIdx.x] + kSigReady + p.rank, p.epoch);
  }
  if (threadIdx.x != 0) return;

  auto piece_addr = [&](int i, size_t& off, uint32_t& bytes, int& chunk, bool& last) {
    chunk = i / static_cast<int>(pieces_per_slice);
    const uint32_t pc = i - chunk * pieces_per_slice;
    const uint32_t in_slice = pc * kCommPiece;
    bytes = min(static_cast<uint32_t>(kCommPiece), p.slice_bytes - in_slice);
    off = static_cast<size_t>(p.rank) * part_bytes + static_cast<size_t>(chunk) * p.chunk_bytes +
          static_cast<size_t>(ci) * p.slice_bytes + in_slice;
    last = (pc + 1 == pieces_per_slice);
  };
  auto issue_load = [&](int i) {
    size_t off; uint32_t bytes; int chunk; bool last;
    piece_addr(i, off, bytes, chunk, last);
    const int slot = i % kCommRing;
    const uint32_t bar = bar_base + slot * 8;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_load(smem_u32(ring + slot * kCommPiece), p.local_a + off, bytes, bar);
  };

  const int prefill = total < kCommRing ? total : kCommRing;
  for (int i = 0; i < prefill; ++i) issue_load(i);
  for (int d = 1; d < p.world; ++d) {
    int peer = p.rank + d;
    peer = peer >= p.world ? peer - p.world : peer;
    wait_flag_ge<true>(p.sig_local + kSigReady + peer, p.epoch);
  }
  for (int i = 0; i < total + kPushDoneLag; ++i) {
    if (i < total) {
      size_t off; uint32_t bytes; int chunk; bool last;
      piece_addr(i, off, bytes, chunk, last);
      const int slot = i % kCommRing;
      mbar_wait(bar_base + slot * 8, (i / kCommRing) & 1);
      const uint32_t src = smem_u32(ring + slot * kCommPiece);
      for (int d = 1; d < p.world; ++d) {
        int peer = p.rank + d;
        peer = peer >= p.world ? peer - p.world : peer;
        bulk_store(const_cast<uint8_t*>(p.peer_a[peer]) + off, src, bytes);
      }
    }
    bulk_commit();
    const int jr = i - kPushReadLag;             // stores of piece jr have read their slot
    if (jr >= 0 && jr + kCommRing < total) {
      bulk_wait_read<kPushReadLag>();
      issue_load(jr + kCommRing);
    }
    const int jd = i - kPushDoneLag;             // stores of piece jd have completed
    if (jd >= 0) {
      size_t off; uint32_t bytes; int chunk; bool last;
      piece_addr(jd, off, bytes, chunk, last);
      if (last) {
        bulk_wait<kPushDoneLag>();
        fence_proxy_async();
        for (int d = 1; d < p.world; ++d) {
          int peer = p.rank + d;
          peer = peer >= p.world ? peer - p.world : peer;
          red_add_release_sys(p.sig_peer[peer] + kSigChunk + p.rank * chunks_per_part + chunk, 1u);
        }
      }
    }
  }
  bulk_wait<0>();
}

// ---------------------------------------------------------------------------
// communication warp in EVERY CTA (AG_BEFORE, default transport). One SM sustains only
// ~20 GB/s over NVLink whichever way the bytes move, so a handful of dedicated CTAs
// cannot fill the links; spread over all CTAs the per-SM demand is ~5 GB/s and the
// GEMM keeps every SM. The pieces (8 KiB) of all chunks, in arrival order, are dealt
// round-robin to the CTAs; lane 0 of warp 3 pulls its pieces peer -> smem -> local
// gathered buffer and bumps the chunk's counter once per piece.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void comm_warp_loop(const TpGemmParams& p, uint8_t* ring, uint32_t bar_base,
                                               int ci, int nci) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int n_chunks = chunks_per_part * (p.world - 1);
  const int ppc = static_cast<int>((p.chunk_bytes + kWarpPiece - 1) / kWarpPiece);   // pieces per chunk
  const int total = n_chunks * ppc;
  const int n_my = ci < total ? (total - ci + nci - 1) / nci : 0;
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;

  uint32_t ready_mask = 1u << p.rank;
  auto piece_addr = [&](int i, size_t& off, uint32_t& bytes, int& part, int& cip) {
    const int g = ci + i * nci;                    // global piece index
    const int c = g / ppc;                         // chunk in fetch order (stage-major, peers rotated)
    const int pc = g - c * ppc;
    const int stage = c / (p.world - 1);
    const int pi = c - stage * (p.world - 1);
    int src = p.rank + 1 + pi;
    src = src >= p.world ? src - p.world : src;
    part = src;
    cip = stage;
    const uint32_t in_chunk = static_cast<uint32_t>(pc) * kWarpPiece;
    bytes = min(static_cast<uint32_t>(kWarpPiece), p.chunk_bytes - in_chunk);
    off = static_cast<size_t>(src) * part_bytes + static_cast<size_t>(stage) * p.chunk_bytes + in_chunk;
  };
  auto issue_load = [&](int i) {
    size_t off; uint32_t bytes; int part, cip;
    piece_addr(i, off, bytes, part, cip);
    if (!(ready_mask & (1u << part))) {
      wait_flag_ge<true>(p.sig_local + kSigReady + part, p.epoch);
      ready_mask |= 1u << part;
    }
    const int slot = i % kWarpRing;
    const uint32_t bar = bar_base + slot * 8;
    mbar_arrive_expect_tx(bar, bytes);
    bulk_load(smem_u32(ring + slot * kWarpPiece), p.peer_a[part] + off, bytes, bar);
  };

  const int prefill = n_my < kWarpRing ? n_my : kWarpRing;
  for (int i = 0; i < prefill; ++i) issue_load(i);
  for (int i = 0; i < n_my + kWarpLag; ++i) {
    if (i < n_my) {
      size_t off; uint32_t bytes; int part, cip;
      piece_addr(i, off, bytes, part, cip);
      const int slot = i % kWarpRing;
      mbar_wait(bar_base + slot * 8, (i / kWarpRing) & 1);
      bulk_store(p.local_a + off, smem_u32(ring + slot * kWarpPiece), bytes);
    }
    bulk_commit();
    const int j = i - kWarpLag;
    if (j >= 0 && j < n_my) {
      bulk_wait<kWarpLag>();                       // piece j is in local memory
      size_t off; uint32_t bytes; int part, cip;
      piece_addr(j, off, bytes, part, cip);
      fence_proxy_async();
      red_add_release_gpu(p.sig_local + kSigChunk + part * chunks_per_part + cip, 1u);
      if (j + kWarpRing < n_my) issue_load(j + kWarpRing);
    }
  }
  bulk_wait<0>();
}

// ---------------------------------------------------------------------------
// communication warps in EVERY CTA, register path (AG_BEFORE, default transport).
// Measured on B200/NVSwitch: a 128-bit load/store copy moves ~700 GB/s to a peer (both
// directions busy) while cp.async.bulk through shared memory stays near 400 GB/s. So warps
// 2 and 3 of every CTA PUSH this rank's shard: each 32 KiB unit is read once from local
// memory into registers and stored to all d-1 peers' gathered buffers (posted writes),
// then fenced at system scope and announced with one remote counter bump per peer. No
// shared memory is used, all SMs keep computing, per-SM NVLink demand is ~5 GB/s.
// ---------------------------------------------------------------------------
constexpr int kPushUnit = 32 * 1024;
constexpr int kPushUnroll = 16;      // 16 x 16 B per lane in flight: 8 KiB per warp

__device__ __forceinline__ void comm_warp_push_regs(const TpGemmParams& p, int gw, int nw, int lane) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int upc = static_cast<int>((p.chunk_bytes + kPushUnit - 1) / kPushUnit);   // units per chunk
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
    for (uint32_t off = begin + lane * 16; off < end; off += 32 * 16 * kPushUnroll) {
      uint4 v[kPushUnroll];
#pragma unroll
      for (int i = 0; i < kPushUnroll; ++i) {
        const uint32_t o = off + i * 512;
        if (o < end) v[i] = __ldcs(reinterpret_cast<const uint4*>(p.local_a + base + o));
      }
      for (int d = 1; d < p.world; ++d) {
        int peer = p.rank + d;
        peer = peer >= p.world ? peer - p.world : peer;
        uint8_t* dst = const_cast<uint8_t*>(p.peer_a[peer]) + base;
#pragma unroll
        for (int i = 0; i < kPushUnroll; ++i) {
          const uint32_t o = off + i * 512;
          if (o < end) __stcs(reinterpret_cast<uint4*>(dst + o), v[i]);
        }
      }
    }
    __threadfence_system();          // this lane's remote stores are performed
    __syncwarp();
    if (lane >= 1 && lane < p.world) {
      int peer = p.rank + lane;
      peer = peer >= p.world ? peer - p.world : peer;
      red_add_release_sys(p.sig_peer[peer] + kSigChunk + p.rank * chunks_per_part + chunk, 1u);
    }
  }
}

// One unit of work of a register-path comm warp: `kPushUnit` bytes [begin, end) of chunk
// `chunk` of rank `src`'s shard; `base` is the byte offset of that chunk in a gathered buffer.
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
  const int pi = c - stage * (p.world - 1);
  int src = p.rank + 1 + pi;
  src = src >= p.world ? src - p.world : src;
  cu.src = src;
  cu.chunk = stage;
  cu.begin = static_cast<uint32_t>(u - c * upc) * kPushUnit;
  const uint32_t lim = cu.begin + static_cast<uint32_t>(kPushUnit);
  cu.end = lim < p.chunk_bytes ? lim : p.chunk_bytes;
  cu.base = static_cast<size_t>(src) * part_bytes + static_cast<size_t>(stage) * p.chunk_bytes;
  return cu;
}

// Register-path PULL counterpart of comm_warp_push_regs: the units of all peers' chunks, in
// arrival order, are dealt round-robin to the comm warps; each is read from the peer's
// gathered buffer into registers and stored locally, then the local chunk counter is bumped.
__device__ __forceinline__ void comm_warp_pull_regs(const TpGemmParams& p, int gw, int nw, int lane) {
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int upc = static_cast<int>((p.chunk_bytes + kPushUnit - 1) / kPushUnit);
  const int total = chunks_per_part * (p.world - 1) * upc;
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;

  if (lane < p.world && lane != p.rank) wait_flag_ge<true>(p.sig_local + kSigReady + lane, p.epoch);
  __syncwarp();

  for (int u = gw; u < total; u += nw) {
    const CommUnit cu = pull_unit(p, u, upc, part_bytes);
    const int src = cu.src;
    const int stage = cu.chunk;
    const uint32_t begin = cu.begin;
    const uint32_t end = cu.end;
    const size_t base = cu.base;
    const uint8_t* from = p.peer_a[src] + base;
    for (uint32_t off = begin + lane * 16; off < end; off += 32 * 16 * kPushUnroll) {
      uint4 v[kPushUnroll];
#pragma unroll
      for (int i = 0; i < kPushUnroll; ++i) {
        const uint32_t o = off + i * 512;
        if (o < end) v[i] = __ldcs(reinterpret_cast<const uint4*>(from + o));
      }
#pragma unroll
      for (int i = 0; i < kPushUnroll; ++i) {
        const uint32_t o = off + i * 512;
        if (o < end) __stcg(reinterpret_cast<uint4*>(p.local_a + base + o), v[i]);
      }
    }
    __threadfence();
    __syncwarp();
    if (lane == 0)
      red_add_release_gpu(p.sig_local + kSigChunk + src * chunks_per_part + stage, 1u);
  }
}

// ---------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Opt-in device trace: a timeline of one launch without a profiler (ncu cannot replay
// kernels that wait for other ranks). Compiled in only for kTrace instantiations, so the
// production kernels carry no trace code at all.
// ---------------------------------------------------------------------------
constexpr int kTraceSlots = 128;
constexpr int kTraceRoles = 4;   // 0 producer, 1 mma, 2 epilogue (warp 4), 3 comm (warp 2)
enum : uint32_t {
  kEvStart = 1, kEvTileBegin = 2, kEvFlagWaited = 3, kEvTileIssued = 4, kEvAccReady = 5,
  kEvOwnWaited = 6, kEvTileStored = 7, kEvCommUnit = 8, kEvEnd = 9,
};

template <bool kTrace>
struct Tracer {
  uint4* base;
  int n;
  __device__ __forceinline__ Tracer(const TpGemmParams& p, int cta, int role) : base(nullptr), n(0) {
    if constexpr (kTrace) {
      if (p.trace != nullptr)
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

// Store one staged epilogue chunk (32 rows x 128 bytes, 128-byte swizzled in shared memory)
// with 128-bit stores: per instruction the warp writes 4 rows x 128 contiguous bytes. Used by
// the push variant for PEER destinations, where plain stores move ~690 GB/s over NVLink and
// TMA stores were measured at ~400-530 GB/s.
__device__ __forceinline__ void store_chunk_st128(uint8_t* dst_base, size_t pitch_bytes, int row0,
                                                  int col_byte0, int row_limit, int col_limit_bytes,
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
    if (row0 + r < row_limit && colb < col_limit_bytes) {
      uint4* dst = reinterpret_cast<uint4*>(dst_base + static_cast<size_t>(row0 + r) * pitch_bytes + colb);
      asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(v.x), "r"(v.y), "r"(v.z),
                   "r"(v.w)
                   : "memory");
    }
  }
}

// kF32: fp32 inputs and outputs through kind::tf32 (operands are read as TF32 by the tensor
// core, fp32 accumulation). A 128-byte swizzle row then holds 32 elements and one MMA covers
// 8 of them, so the byte geometry of the operand ring is identical to the 16-bit case.
// kVariant: experimental schedule variants compiled as separate instantiations (bit 0: RS
// announces a pushed tile as soon as its stores have completed instead of at the next tile;
// bit 1: RS own-tile epilogue issues the staged-partial loads of two sources before adding;
// bit 2: tiles pushed to PEERS (RS, AG_AFTER) are written with 128-bit stores from the staged
// shared-memory rows instead of TMA stores).
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, bool kTrace = false,
          int kVariant = 0>
__device__ __forceinline__ void tp_gemm_body(const TpGemmParams& p, const int cluster,
                                             const int nclusters) {
  using Cfg = TpGemmCfg<kCtaGroup, kBlockN>;
  constexpr int kElem = kF32 ? 4 : 2;                 // bytes per element
  constexpr int kRowElems = 128 / kElem;              // elements per 128-byte swizzle row
  constexpr int kKPerStage = kRowElems;               // k elements per pipeline stage
  constexpr int kKPerMma = 32 / kElem;                // 16 (f16/bf16) or 8 (tf32)
  const int kStages = p.stages;                               // runtime ring depth
  const bool comm_in_cta = p.mode == kModeAGBefore && p.comm_in_cta == 1;   // TMA pull ring in warp 3
  const bool comm_regs = p.mode == kModeAGBefore && p.comm_in_cta >= 2;     // register path, warps 2+3
  const bool comm_regs_pull = p.comm_in_cta == 3;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_ops = smem;                                   // [stage]{A, B}
  uint8_t* smem_epi = smem + kStages * Cfg::kStageBytes;      // [warp][2][4 KB]
  uint8_t* smem_comm = smem_epi + kEpiBytes;                  // comm warp ring (comm_in_cta only)
  uint8_t* smem_bar = smem_comm + (comm_in_cta ? kWarpRingBytes : 0);
  const uint32_t bar_full = smem_u32(smem_bar);               // [kStages] (also comm ring, <= 12)
  const uint32_t bar_empty = bar_full + 12 * 8;               // [kStages]
  const uint32_t bar_tfull = bar_empty + 12 * 8;              // [2]
  const uint32_t bar_tempty = bar_tfull + 2 * 8;              // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem_bar + 28 * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = kCtaGroup == 2 ? cluster_ctarank() : 0u;
  const bool is_comm = (p.mode == kModeAGBefore) && cluster < p.n_comm_clusters;
  const bool bf16 = ((p.idesc >> 7) & 7u) == 1u;

  // ------------------------------------------------------------------ setup
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA);
    prefetch_tmap(&p.tmB);
    prefetch_tmap(&p.tmOut[p.rank]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < 12; ++s) {
      mbar_init(bar_full + s * 8, 1);
      mbar_init(bar_empty + s * 8, 1);
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
  if constexpr (kCtaGroup == 2) cluster_sync(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int num_tiles = p.num_m_blocks * p.num_n_blocks;
  const int gemm_cluster = cluster - (p.mode == kModeAGBefore ? p.n_comm_clusters : 0);
  const int gemm_clusters = nclusters - (p.mode == kModeAGBefore ? p.n_comm_clusters : 0);

  if (is_comm) {
    // -------------------------------------------------------- comm role
    if (p.comm_push)
      comm_cta_push_loop(p, smem_ops, bar_full, cluster * kCtaGroup + static_cast<int>(cta_rank),
                         p.n_comm_clusters * kCtaGroup);
    else
      comm_cta_loop(p, smem_ops, bar_full, cluster * kCtaGroup + static_cast<int>(cta_rank),
                    p.n_comm_clusters * kCtaGroup);
  } else if (warp == 0 && lane == 0) {
    // -------------------------------------------------------- TMA producer
    const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
    const uint32_t chunk_target =
        p.epoch * (comm_regs ? (p.chunk_bytes + kPushUnit - 1) / kPushUnit
                   : comm_in_cta ? (p.chunk_bytes + kWarpPiece - 1) / kWarpPiece
                                 : static_cast<uint32_t>(p.n_comm_clusters * kCtaGroup));
    const uint32_t full_leader = kCtaGroup == 2 ? mapa(bar_full, 0) : bar_full;
    int stage = 0;
    uint32_t phase = 0;
    Tracer<kTrace> tr(p, cluster * kCtaGroup + static_cast<int>(cta_rank), 0);
    tr.ev(kEvStart, 0);
    for (int t = gemm_cluster; t < num_tiles; t += gemm_clusters) {
      const TileCoord tc = tile_coord(p, t);
      tr.ev(kEvTileBegin, static_cast<uint32_t>(t));
      if (p.mode == kModeAGBefore && tc.part != p.rank) {
        if (p.comm_push || (comm_regs && !comm_regs_pull))   // counter bumped over NVLink
          wait_flag_ge<true>(p.sig_local + kSigChunk + tc.part * chunks_per_part + tc.chunk,
                             chunk_target);
        else
          wait_flag_ge<false>(p.sig_local + kSigChunk + tc.part * chunks_per_part + tc.chunk,
                              chunk_target);
        fence_proxy_async();      // flag acquired by the generic proxy, data read by TMA
        tr.ev(kEvFlagWaited, static_cast<uint32_t>(t));
      }
      const int m0 = tc.m_blk * Cfg::kTileM + static_cast<int>(cta_rank) * kBlockM;
      const int n0 = tc.n_blk * kBlockN + static_cast<int>(cta_rank) * Cfg::kLoadN;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(bar_empty + stage * 8, phase ^ 1u);
        uint8_t* sa = smem_ops + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kStageA;
        if (kCtaGroup == 1 || cta_rank == 0)
          mbar_arrive_expect_tx(bar_full + stage * 8, Cfg::kStageBytes * kCtaGroup);
        const uint32_t fb = full_leader + stage * 8;
        tma_load_2d<kCtaGroup>(smem_u32(sa), &p.tmA, fb, kb * kKPerStage, m0);
        if constexpr (kBNMajor)
          tma_load_3d<kCtaGroup>(smem_u32(sb), &p.tmB, fb, 0, kb * kKPerStage, n0 / kRowElems);
        else
          tma_load_2d<kCtaGroup>(smem_u32(sb), &p.tmB, fb, kb * kKPerStage, n0);
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if ((warp == 2 || warp == 3) && comm_regs) {
    // -------------------------------------------------------- comm warps (every CTA, registers)
    const int ci = cluster * kCtaGroup + static_cast<int>(cta_rank);
    if (ci == 0 && warp == 3 && lane < p.world && lane != p.rank)
      st_release_sys(p.sig_peer[lane] + kSigReady + p.rank, p.epoch);   // my previous launch is over
    if (comm_regs_pull) comm_warp_pull_regs(p, ci * 2 + (warp - 2), nclusters * kCtaGroup * 2, lane);
    else comm_warp_push_regs(p, ci * 2 + (warp - 2), nclusters * kCtaGroup * 2, lane);
  } else if (warp == 3 && comm_in_cta) {
    // -------------------------------------------------------- comm warp (every CTA)
    const int ci = cluster * kCtaGroup + static_cast<int>(cta_rank);
    if (ci == 0 && lane < p.world && lane != p.rank)
      st_release_sys(p.sig_peer[lane] + kSigReady + p.rank, p.epoch);   // my shard is final
    if (lane == 0)
      comm_warp_loop(p, smem_comm, bar_full + 8 * 8, ci, nclusters * kCtaGroup);
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
    Tracer<kTrace> tr(p, cluster * kCtaGroup + static_cast<int>(cta_rank), 1);
    for (int t = gemm_cluster; t < num_tiles; t += gemm_clusters) {
      mbar_wait(bar_tempty + acc * 8, acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kBlockN);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(bar_full + stage * 8, phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem_ops + stage * Cfg::kStageBytes);
        const uint64_t adesc = make_smem_desc_sw128(sa, 16, 1024);
        const uint64_t bdesc = make_smem_desc_sw128(sa + Cfg::kStageA, kBLbo, 1024);
#pragma unroll
        for (int k = 0; k < kKPerStage / kKPerMma; ++k) {
          if constexpr (kF32)
            umma_tf32<kCtaGroup>(tmem_d, adesc + static_cast<uint64_t>(k * (32 >> 4)),
                                 bdesc + static_cast<uint64_t>(k * kBStep), p.idesc,
                                 static_cast<uint32_t>((kb | k) != 0));
          else
            umma_f16<kCtaGroup>(tmem_d, adesc + static_cast<uint64_t>(k * (32 >> 4)),
                                bdesc + static_cast<uint64_t>(k * kBStep), p.idesc,
                                static_cast<uint32_t>((kb | k) != 0));
        }
        umma_commit<kCtaGroup>(bar_empty + stage * 8);     // smem slot reusable when MMAs retire
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      umma_commit<kCtaGroup>(bar_tfull + acc * 8);         // accumulator ready for the epilogue
      tr.ev(kEvTileIssued, static_cast<uint32_t>(t));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------- epilogue
    const int q = warp & 3;                                 // TMEM lane quadrant
    uint8_t* ebuf = smem_epi + q * 2 * kEpiBufBytes;
    const uint32_t tempty_leader = kCtaGroup == 2 ? mapa(bar_tempty, 0) : bar_tempty;
    int buf = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t* pending_sig = nullptr;                        // RS: counter of the previous pushed tile
    const uint32_t parity = p.epoch & 1u;
    Tracer<kTrace> tr(p, (warp == 4 && lane == 0) ? cluster * kCtaGroup + static_cast<int>(cta_rank) : -1, 2);
    if (!(warp == 4 && lane == 0)) tr.base = nullptr;
    for (int t = gemm_cluster; t < num_tiles; t += gemm_clusters) {
      const TileCoord tc = tile_coord(p, t);
      const int row_cta = tc.m_blk * Cfg::kTileM + static_cast<int>(cta_rank) * kBlockM;
      const int row0 = row_cta + q * 32;
      const int n0 = tc.n_blk * kBlockN;
      mbar_wait(bar_tfull + acc * 8, acc_phase);
      tc_fence_after();
      tr.ev(kEvAccReady, static_cast<uint32_t>(t));

      const bool rs_push = p.mode == kModeRS && tc.part != p.rank;
      const bool rs_own = p.mode == kModeRS && tc.part == p.rank;
      int out_row = row0;
      int local_row = 0;
      int cta_tile = 0;
      if (p.mode == kModeRS) {
        local_row = row0 - tc.part * p.part_rows;           // row inside the owner's shard
        cta_tile = ((row_cta - tc.part * p.part_rows) / kBlockM) * p.num_n_blocks + tc.n_blk;
        out_row = rs_push ? (static_cast<int>(parity) * p.world + p.rank) * p.part_rows + local_row
                          : local_row;
      } else if (p.mode == kModeAGAfter) {
        out_row = p.rank * p.part_rows + row0;
      }
      if (pending_sig != nullptr) {
        // previous tile's remote stores: complete them, then publish its counter
        if (lane == 0) {
          bulk_wait<0>();
          fence_proxy_async();
          red_add_release_sys(pending_sig, 1u);
        }
        pending_sig = nullptr;
      }
      if (rs_own) {
        if (lane < p.world && lane != p.rank)
          wait_flag_ge<true>(p.sig_local + p.rs_sig_off + lane * p.rs_tiles_per_src + cta_tile,
                             4u * p.epoch);
        __syncwarp();
        tr.ev(kEvOwnWaited, static_cast<uint32_t>(t));
      }

      if constexpr (kF32) {
        // fp32 output: 32 columns (one 128-byte row segment) per step, no packing
#pragma unroll 1
        for (int c = 0; c < kBlockN / 32; ++c) {
          uint32_t v[32];
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                 static_cast<uint32_t>(acc * kBlockN + c * 32);
          tmem_ld_32x32b_x32(taddr, &v[0]);
          tmem_ld_wait();

          if (rs_own) {
            const int lr = local_row + lane;
            for (int s = 0; s < p.world; ++s) {
              if (s == p.rank) continue;
              const size_t row = static_cast<size_t>(parity * p.world + s) * p.part_rows + lr;
              const uint4* src = reinterpret_cast<const uint4*>(
                  p.stage_local + (row * p.out_n + n0 + c * 32) * 4);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint4 raw = __ldcg(src + j);
                v[4 * j] = __float_as_uint(__uint_as_float(v[4 * j]) + __uint_as_float(raw.x));
                v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + __uint_as_float(raw.y));
                v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + __uint_as_float(raw.z));
                v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + __uint_as_float(raw.w));
              }
            }
          }

          if (lane == 0) bulk_wait_read<1>();
          __syncwarp();
          const uint32_t sbase = smem_u32(ebuf + buf * kEpiBufBytes);
          const uint32_t srow = sbase + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t dst = srow + (static_cast<uint32_t>(j ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v[4 * j]),
                         "r"(v[4 * j + 1]), "r"(v[4 * j + 2]), "r"(v[4 * j + 3])
                         : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (p.mode == kModeAGAfter) {
              for (int d = 0; d < p.world; ++d) tma_store_2d(&p.tmOut[d], sbase, n0 + c * 32, out_row);
            } else {
              tma_store_2d(&p.tmOut[p.mode == kModeRS ? tc.part : p.rank], sbase, n0 + c * 32, out_row);
            }
            bulk_commit();
          }
          buf ^= 1;
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < kBlockN / 64; ++c) {
        uint32_t v[64];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(acc * kBlockN + c * 64);
        tmem_ld_32x32b_x32(taddr, &v[0]);
        tmem_ld_32x32b_x32(taddr + 32, &v[32]);
        tmem_ld_wait();

        if (rs_own) {
          const int lr = local_row + lane;
          if constexpr ((kVariant & 2) != 0) {
            // two sources in flight: 16 independent 16-byte loads before the first add
            const size_t col = static_cast<size_t>(n0 + c * 64);
            for (int i = 1; i < p.world; i += 2) {
              int s0 = p.rank + i, s1 = p.rank + i + 1;
              s0 = s0 >= p.world ? s0 - p.world : s0;
              s1 = s1 >= p.world ? s1 - p.world : s1;
              const bool two = i + 1 < p.world;
              const size_t row0 = static_cast<size_t>(parity * p.world + s0) * p.part_rows + lr;
              const size_t row1 = static_cast<size_t>(parity * p.world + (two ? s1 : s0)) * p.part_rows + lr;
              const uint4* src0 = reinterpret_cast<const uint4*>(p.stage_local + (row0 * p.out_n + col) * 2);
              const uint4* src1 = reinterpret_cast<const uint4*>(p.stage_local + (row1 * p.out_n + col) * 2);
              uint4 r0[8], r1[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) r0[j] = __ldcg(src0 + j);
#pragma unroll
              for (int j = 0; j < 8; ++j) r1[j] = __ldcg(src1 + j);
#pragma unroll
              for (int j = 0; j < 8; ++j) add_packed8(&v[8 * j], r0[j], bf16);
              if (two) {
#pragma unroll
                for (int j = 0; j < 8; ++j) add_packed8(&v[8 * j], r1[j], bf16);
              }
            }
          } else {
          for (int s = 0; s < p.world; ++s) {
            if (s == p.rank) continue;
            const size_t row = static_cast<size_t>(parity * p.world + s) * p.part_rows + lr;
            const uint4* src = reinterpret_cast<const uint4*>(
                p.stage_local + (row * p.out_n + n0 + c * 64) * 2);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 raw = __ldcg(src + j);
              add_packed8(&v[8 * j], raw, bf16);
            }
          }
          }
        }

        uint32_t packed[32];
#pragma unroll
        for (int i = 0; i < 32; ++i)
          packed[i] = pack2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), bf16);

        if (lane == 0) bulk_wait_read<1>();     // the store that last used this buffer has read it
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
        fence_proxy_async_smem();
        __syncwarp();
        if constexpr ((kVariant & 4) != 0) {
          // push variant: peers are written with 128-bit stores from the staged rows, only
          // local destinations go through TMA
          const size_t pitch = static_cast<size_t>(p.out_n) * 2;
          const int colb = (n0 + c * 64) * 2;
          if (rs_push) {
            store_chunk_st128(p.out_raw[tc.part], pitch, out_row, colb, 2 * p.world * p.part_rows,
                              p.out_n * 2, sbase, lane);
          } else if (p.mode == kModeAGAfter) {
            for (int d = 0; d < p.world; ++d)
              if (d != p.rank)
                store_chunk_st128(p.out_raw[d], pitch, out_row, colb, p.world * p.part_rows,
                                  p.out_n * 2, sbase, lane);
            if (lane == 0) {
              tma_store_2d(&p.tmOut[p.rank], sbase, n0 + c * 64, out_row);
              bulk_commit();
            }
          } else if (lane == 0) {
            tma_store_2d(&p.tmOut[p.mode == kModeRS ? tc.part : p.rank], sbase, n0 + c * 64, out_row);
            bulk_commit();
          }
          __syncwarp();            // every lane has read the staged rows before they are reused
        } else {
        if (lane == 0) {
          if (p.mode == kModeAGAfter) {
            for (int d = 0; d < p.world; ++d) tma_store_2d(&p.tmOut[d], sbase, n0 + c * 64, out_row);
          } else {
            // RS: the owner's staging (or own shard) map; otherwise this rank's C
            tma_store_2d(&p.tmOut[p.mode == kModeRS ? tc.part : p.rank], sbase, n0 + c * 64, out_row);
          }
          bulk_commit();
        }
        }
        buf ^= 1;
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kCtaGroup == 2 && cta_rank != 0) mbar_arrive_cluster(tempty_leader + acc * 8);
        else mbar_arrive(bar_tempty + acc * 8);
      }
      tr.ev(kEvTileStored, static_cast<uint32_t>(t));
      if constexpr ((kVariant & 4) != 0) {
        if (rs_push) {
          __threadfence_system();          // this lane's peer stores are performed
          __syncwarp();
          if (lane == 0)
            red_add_release_sys(p.sig_peer[tc.part] + p.rs_sig_off + p.rank * p.rs_tiles_per_src + cta_tile, 1u);
        }
      } else if constexpr ((kVariant & 1) != 0) {
        if (rs_push && lane == 0) {
          // announce now: the accumulator is already released, only this warp waits
          bulk_wait<0>();
          fence_proxy_async();
          red_add_release_sys(p.sig_peer[tc.part] + p.rs_sig_off + p.rank * p.rs_tiles_per_src + cta_tile, 1u);
        }
      } else {
      if (rs_push)
        pending_sig = p.sig_peer[tc.part] + p.rs_sig_off + p.rank * p.rs_tiles_per_src + cta_tile;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) {
      bulk_wait<0>();
      if (pending_sig != nullptr) {
        fence_proxy_async();
        red_add_release_sys(pending_sig, 1u);
      }
    }
    tr.ev(kEvEnd, 0);
    if (p.mode == kModeAGAfter && p.world > 1) {
      // every CTA tells every peer that its share of C has landed, then waits for
      // all CTAs of all peers so that C is complete when this kernel ends
      if constexpr ((kVariant & 4) != 0) __threadfence_system();
      if (lane == 0) { fence_proxy_async(); fence_acq_rel_sys(); }
      named_bar_sync(1, 128);
      if (warp == 4 && lane < p.world && lane != p.rank) {
        red_add_release_sys(p.sig_peer[lane] + kSigDone + p.rank, 1u);
        wait_flag_ge<true>(p.sig_local + kSigDone + lane,
                           p.epoch * static_cast<uint32_t>(nclusters * kCtaGroup));
      }
    }
  }

  // ---------------------------------------------------------------- teardown
  __syncwarp();
  tc_fence_before();
  if constexpr (kCtaGroup == 2) cluster_sync(); else __syncthreads();
  if (warp == 2) tmem_dealloc<kCtaGroup>(tmem_base, Cfg::kTmemCols);
}

// One rank per process / GPU: parameters in the kernel parameter space.
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, bool kTrace = false,
          int kVariant = 0>
__global__ void __launch_bounds__(kNumThreads, 1)
tp_gemm_kernel(const __grid_constant__ TpGemmParams p) {
  const int cluster = kCtaGroup == 2 ? static_cast<int>(cluster_id_x()) : static_cast<int>(blockIdx.x);
  const int nclusters = kCtaGroup == 2 ? static_cast<int>(num_clusters_x()) : static_cast<int>(gridDim.x);
  tp_gemm_body<kCtaGroup, kBlockN, kBNMajor, kF32, kTrace, kVariant>(p, cluster, nclusters);
}

// Several ranks emulated by ONE launch on one GPU (tests): the grid is split into
// `world` equal groups of clusters, group r runs rank r's parameter block. All
// CTAs of one launch are co-resident (grid <= SM count), so ranks that wait on
// each other make progress, which separate launches on one GPU cannot guarantee.
template <int kCtaGroup, int kBlockN, bool kBNMajor, bool kF32 = false, int kVariant = 0>
__global__ void __launch_bounds__(kNumThreads, 1)
tp_gemm_emulated_kernel(const TpGemmParams* __restrict__ plist, int clusters_per_rank) {
  const int cluster = kCtaGroup == 2 ? static_cast<int>(cluster_id_x()) : static_cast<int>(blockIdx.x);
  const int vr = cluster / clusters_per_rank;
  tp_gemm_body<kCtaGroup, kBlockN, kBNMajor, kF32, false, kVariant>(
      plist[vr], cluster - vr * clusters_per_rank, clusters_per_rank);
}

}  // namespace ddlb

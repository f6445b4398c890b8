/* C ABI of libddlb_b200.so — the native layer of the B200 tensor-parallel
 * primitive library. Every function returns 0 on success and a non-zero code on
 * failure; ddlb_last_error() then describes the failure (thread-local string).
 *
 * Pointers named "peer" are device addresses of ANOTHER rank's memory mapped into
 * this process (see ddlb_symm_*). Streams are cudaStream_t handles passed as void*.
 */
#ifndef DDLB_B200_H_
#define DDLB_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDLB_MAX_WORLD 8

const char* ddlb_last_error(void);
int ddlb_version(void);

/* ---------------------------------------------------------------- topology
 * JSON description of the visible devices: name, SM count, memory, peer-access
 * matrix, P2P attributes, multicast (NVLS) support. No counterpart in the
 * reference beyond torch.cuda.set_device (ddlb/communicator.py:56-60).
 * Writes at most `cap` bytes (NUL terminated); returns bytes needed in *needed. */
int ddlb_topology_json(char* out, int64_t cap, int64_t* needed);

/* ----------------------------------------------------------- symmetric heap
 * One cudaMalloc arena per rank, exported with cudaIpc and mapped by every peer
 * (replaces NCCL communicator setup, ddlb/primitives/TPColumnwise/pytorch.py:53-59).
 * The first DDLB_SIGNAL_BYTES of the arena are zero-initialised signal pads. */
#define DDLB_IPC_HANDLE_BYTES 64
typedef struct ddlb_symm ddlb_symm;

int ddlb_symm_create(int device, int rank, int world, int64_t bytes, ddlb_symm** out);
int ddlb_symm_export(ddlb_symm* s, void* handle_out /* DDLB_IPC_HANDLE_BYTES */);
/* handles: world * DDLB_IPC_HANDLE_BYTES, indexed by rank (own entry ignored) */
int ddlb_symm_open_peers(ddlb_symm* s, const void* handles);
/* single-process emulation (tests): adopt another heap of the same process as peer */
int ddlb_symm_adopt_peer(ddlb_symm* s, int peer_rank, ddlb_symm* peer);
void* ddlb_symm_base(ddlb_symm* s, int rank);   /* base address of rank's arena as seen here */
int64_t ddlb_symm_bytes(ddlb_symm* s);
int ddlb_symm_destroy(ddlb_symm* s);

/* Device-side barrier over the signal pads (replaces dist.barrier /
 * the dummy all_reduce of ddlb/benchmark.py:134-137, ddlb/communicator.py:74).
 * Asynchronous on `stream`; every rank must call it the same number of times. */
int ddlb_symm_barrier(ddlb_symm* s, void* stream);
/* the same barrier on caller-provided pads: pad_bases[r] = rank r's pad as mapped here; epoch
 * = 1, 2, 3, ... per call, identical on all ranks */
int ddlb_barrier_raw(void* const* pad_bases, int rank, int world, uint32_t epoch, void* stream);

/* ------------------------------------------- VMM heap + NVLS multicast
 * Same role as ddlb_symm_* but on cuMemCreate/cuMemMap with POSIX file-descriptor handles and,
 * where the devices support it, a multicast object (multimem.st / multimem.ld_reduce through
 * the NVSwitch). The default heap of the Python layer on NVSwitch systems (runs at 2, 4 and 8
 * GPUs, tests/test_vmm_multigpu.py). Protocol:
 *   every rank: ddlb_vmm_create; ddlb_vmm_export_fd -> send fd to peers; ddlb_vmm_import_peer
 *   rank 0: ddlb_vmm_mc_create -> send fd; others: ddlb_vmm_mc_import
 *   every rank: ddlb_vmm_mc_add_device; HOST BARRIER; ddlb_vmm_mc_bind
 * Counterpart in the reference: nvFuser's "cuda" backend with multicast_protocol=multimem
 * (ddlb/primitives/TPColumnwise/fuser.py:171-206), which lives outside the reference repo. */
typedef struct ddlb_vmm ddlb_vmm;
int ddlb_vmm_multicast_supported(int device, int* supported);
int ddlb_vmm_create(int device, int rank, int world, int64_t bytes, ddlb_vmm** out);
int ddlb_vmm_export_fd(ddlb_vmm* s, int* fd);
int ddlb_vmm_import_peer(ddlb_vmm* s, int peer, int fd);
int ddlb_vmm_mc_create(ddlb_vmm* s, int* fd);
int ddlb_vmm_mc_import(ddlb_vmm* s, int fd);
int ddlb_vmm_mc_add_device(ddlb_vmm* s);
int ddlb_vmm_mc_bind(ddlb_vmm* s);
void* ddlb_vmm_base(ddlb_vmm* s, int rank);
void* ddlb_vmm_mc_base(ddlb_vmm* s);
int64_t ddlb_vmm_bytes(ddlb_vmm* s);
int ddlb_vmm_selftest_store(ddlb_vmm* s, int64_t offset, int n_per_rank, void* stream);
int ddlb_vmm_selftest_reduce(ddlb_vmm* s, int64_t offset, float* out, int n, void* stream);
int ddlb_vmm_has_multicast(ddlb_vmm* s);
int ddlb_vmm_destroy(ddlb_vmm* s);

/* --------------------------------------------------------------- fused GEMM */
enum { DDLB_MODE_PLAIN = 0, DDLB_MODE_AG_BEFORE = 1, DDLB_MODE_AG_AFTER = 2, DDLB_MODE_RS = 3 };
/* F32: fp32 inputs and outputs on the tensor cores. Default 3xTF32 (hi/lo split of both
 * operands, three kind::tf32 MMAs per k-step, ~fp32 accuracy); DDLB_FP32_TF32 runs the single
 * pass (10-bit mantissa inputs). fp32 needs B as [n,k] (DDLB_B_ROW_MAJOR_NK). */
enum { DDLB_DTYPE_F16 = 0, DDLB_DTYPE_BF16 = 1, DDLB_DTYPE_F32 = 2 };
enum { DDLB_FP32_3XTF32 = 0, DDLB_FP32_TF32 = 1 };
enum { DDLB_B_ROW_MAJOR_KN = 0, DDLB_B_ROW_MAJOR_NK = 1 };
/* Who moves the A shards in AG_BEFORE (aux warps of every CTA, 128-bit accesses through
 * registers). PULL (default): read the peers' shards, write the local gathered buffer.
 * PUSH: read the own shard once, write it into every peer's gathered buffer.
 * PUSH_MULTICAST: one multimem.st per 16 bytes, replicated by the NVSwitch (needs mc_ptr).
 * Reference counterparts: ncclAllGather (TPColumnwise/pytorch.py:96), nvFuser's p2p / multicast
 * protocols (TPColumnwise/fuser.py:171-206). */
enum { DDLB_AG_PULL = 0, DDLB_AG_PUSH = 1, DDLB_AG_PUSH_MULTICAST = 2 };
/* How AG_AFTER's epilogue writes a C tile into every rank's C: TMA stores to all d ranks,
 * 128-bit stores to the peers (+ TMA locally), or one multimem.st (needs mc_ptr). */
enum { DDLB_EPI_TMA = 0, DDLB_EPI_ST128 = 1, DDLB_EPI_MULTICAST = 2 };
/* RS: PUSH partial tiles into the owner's staging and let the owner sum its local slots;
 * PULL: keep them in the local staging, the owner loads all d copies over NVLink;
 * PULL_MULTICAST: the owner issues multimem.ld_reduce, the NVSwitch adds the d copies
 * (needs mc_ptr). Reference counterparts: ncclReduceScatter (TPRowwise/pytorch.py:83),
 * nvFuser p2p pipeline and multimem protocol (TPRowwise/fuser.py:116-169,186-197). */
enum { DDLB_RS_PUSH = 0, DDLB_RS_PULL = 1, DDLB_RS_PULL_MULTICAST = 2 };
/* activation fused into the epilogue (no reference counterpart: the reference has no model
 * code; used by ddlb_b200.models.TPMLP so that the block needs no elementwise kernel) */
enum { DDLB_ACT_NONE = 0, DDLB_ACT_RELU = 1, DDLB_ACT_SILU = 2, DDLB_ACT_GELU_TANH = 3 };

typedef struct ddlb_tp_gemm_config {
  int32_t mode;          /* DDLB_MODE_* */
  int32_t dtype;         /* DDLB_DTYPE_* (inputs and output; fp32 accumulate) */
  int32_t b_layout;      /* DDLB_B_ROW_MAJOR_KN: B is [k,n]; _NK: B is [n,k] (pre-transposed) */
  int32_t cta_group;     /* 1 or 2 (tcgen05 cta_group), 0 = auto */
  int32_t block_n;       /* 128 or 256, 0 = auto */
  int32_t group_m;       /* rasterisation group, 0 = auto */
  int32_t num_sms;       /* CTAs to launch, 0 = all SMs */
  int32_t rank, world;
  int32_t chunk_rows;    /* rows per pipeline chunk inside a shard, 0 = auto */
  int32_t ag_transport;  /* AG_BEFORE: DDLB_AG_* */
  int32_t epi_transport; /* AG_AFTER: DDLB_EPI_* */
  int32_t rs_transport;  /* RS: DDLB_RS_* */
  int32_t unrotated;     /* 1: shards are visited 0..d-1 on every rank instead of starting at the
                          * own / next rank (nvFuser offset_stream_indexing_by_rank=False,
                          * TPColumnwise/fuser.py:165,250) */
  int32_t fp32_mode;     /* DDLB_FP32_* */
  int32_t epilogue_act;  /* DDLB_ACT_*: applied to the fp32 accumulator before the output is
                          * rounded (PLAIN / AG_*; not on tp_rowwise partial sums) */
  int32_t comm_warps;    /* AG_BEFORE: aux warps per CTA that move data, 0 = auto (2 for pull) */
  int32_t cluster_pairs; /* 2: clusters of four CTAs = two MMA pairs on neighbouring n-blocks, the A
                          * tile multicast to both (fp16/bf16, cta_group 2, block_n 256, even
                          * number of n-blocks); 0 / 1: one pair per cluster. `a` must then be
                          * memory of the current device itself (checked): TMA multicast loads
                          * here read only this GPU's memory */
  int32_t pdl;           /* programmatic dependent launch: the kernel is launched with stream
                          * serialization relaxed, so its prologue (barrier init, TMEM allocation,
                          * tensor-map prefetch) runs while the previous kernel of the stream drains;
                          * every global access still waits for that kernel (griddepcontrol.wait).
                          * 0 = auto (on when world == 1), 1 = on, 2 = off; DDLB_PDL=0 in the
                          * environment turns it off everywhere. No reference counterpart. */
  int32_t wide_n;        /* wide tiles: a CTA pair computes 256 x 512 (two neighbouring n-blocks share
                          * every A stage, both accumulators fill TMEM): 25 % fewer operand bytes per
                          * MMA, epilogue not overlapped. fp16/bf16, cta_group 2, block_n 256, even
                          * number of n-blocks. 1 = on; 0 (auto) and 2 = off: measured 0.80 - 1.03x of
                          * the 256 x 256 tiles, so auto never picks it. No reference counterpart. */
  int64_t m, n, k;       /* GLOBAL problem C[m,n] = A[m,k] B[k,n] of the primitive */
  /* PLAIN:     a = A[m,k]                 out[rank] = C[m,n]
   * AG_BEFORE: a = gathered A[m,k] (own shard rows already in place), peer_a[r] = rank r's
   *            gathered buffer, out[rank] = C[m,n]; mc_ptr = multicast alias of the gathered A
   * AG_AFTER:  a = A_local[m/d,k]         out[r] = rank r's C[m,n]; mc_ptr = alias of C
   * RS:        a = A_local[m,k/d], b = B_local[k/d,n], out[rank] = C_shard[m/d,n],
   *            stage[r] = rank r's staging [2][d][m/d][n]; mc_ptr = alias of the staging */
  void* a;
  void* b;
  void* out[DDLB_MAX_WORLD];
  void* peer_a[DDLB_MAX_WORLD];
  void* stage[DDLB_MAX_WORLD];
  void* mc_ptr;
  uint32_t* sig_local;                  /* this op's signal words (zero-initialised) */
  uint32_t* sig_peer[DDLB_MAX_WORLD];   /* the same block on every peer */
  /* optional device trace buffer (NULL = off): grid * DDLB_TRACE_ROLES * DDLB_TRACE_SLOTS
   * records of 16 bytes {tag, arg, time_lo, time_hi (globaltimer ns)}, zero-initialised by the
   * caller. Only for fp16/bf16, cta_group 2, block_n 256, B [k,n]. No reference counterpart. */
  void* trace;
} ddlb_tp_gemm_config;
#define DDLB_TRACE_ROLES 4
#define DDLB_TRACE_SLOTS 128

typedef struct ddlb_tp_gemm_plan ddlb_tp_gemm_plan;

/* number of uint32 signal words the op needs for this config */
int64_t ddlb_tp_gemm_sig_words(const ddlb_tp_gemm_config* cfg);
int ddlb_tp_gemm_create(const ddlb_tp_gemm_config* cfg, ddlb_tp_gemm_plan** out);
/* epoch: 1, 2, 3, ... incremented by the caller on every launch, identical on all ranks */
int ddlb_tp_gemm_run(ddlb_tp_gemm_plan* plan, uint32_t epoch, void* stream);
/* resolved launch shape: {grid, CTAs per cluster, smem_bytes, cta_group, block_n, stages, threads,
 * chunk_rows, pairs per cluster, wide tiles (0/1), programmatic dependent launch (0/1), 0} */
int ddlb_tp_gemm_info(const ddlb_tp_gemm_plan* plan, int32_t out[12]);
int ddlb_tp_gemm_destroy(ddlb_tp_gemm_plan* plan);
/* Tests on one GPU: run `world` plans (one per emulated rank, all created on the current
 * device, same kernel variant) as ONE launch whose CTAs are split between the ranks, so
 * ranks that wait on each other are co-resident. No counterpart in the reference. */
int ddlb_tp_gemm_run_emulated(ddlb_tp_gemm_plan** plans, int world, uint32_t epoch, void* stream);

/* Same launch as ddlb_tp_gemm_run_emulated, but it allocates nothing (this entry point has
 * no counterpart in the reference). The per-rank parameter blocks the kernel reads, which
 * the call above takes from cudaMallocAsync each time, sit in a fixed array inside the
 * library image: DDLB_EMULATED_STATIC_MAX_WORLD blocks, about 13 KiB, resident on every
 * device that loads the library until the process ends. Meant for hosts that cannot
 * allocate device memory once they are running: processes that gave the whole device to
 * their own arena at start-up (nothing is left for the driver's pool), and loops that
 * must never take an allocator lock. It calls no cudaMalloc / cudaFree / pool routine.
 *  - world > DDLB_EMULATED_STATIC_MAX_WORLD is refused with an error.
 *  - The first call on each device looks up the array's device address (with that device
 *    current); make that first call before any stream capture begins. The per-device address
 *    cache is a plain array: two host threads making the first call on one device at the same
 *    time both store the same value, which is harmless.
 *  - All calls share the one array. Each call rewrites every block it uses, so nothing
 *    has to be cleared beforehand; but two calls whose work overlaps on the GPU (different
 *    streams, same device) overwrite each other's blocks: keep calls on one stream, or
 *    order them with events.
 * Inputs, outputs, staging and signal words are still the caller's, through the plans. */
#define DDLB_EMULATED_STATIC_MAX_WORLD 8
int ddlb_tp_gemm_run_emulated_static(ddlb_tp_gemm_plan** plans, int world, uint32_t epoch,
                                     void* stream);

/* Introspection of the tile schedule on the host (no GPU needed when cfg->num_sms > 0): the
 * resolved launch/schedule parameters {num_m_blocks, num_n_blocks, mblks_per_part, chunk_mblks,
 * group_m, tile_m, block_n, cta_group, chunk_rows, part0, grid, threads}, and the
 * kernel's own tile_coord() evaluated for tiles [first, first+count). No reference counterpart. */
int ddlb_debug_schedule(const ddlb_tp_gemm_config* cfg, int32_t out[12]);
int ddlb_debug_tile_coords(const ddlb_tp_gemm_config* cfg, int32_t first, int32_t count, int32_t* out);
/* units {src, chunk, begin, end, base} moved by the default AG_BEFORE transport of cfg->rank;
 * pass out = NULL to get the count */
int ddlb_debug_pull_units(const ddlb_tp_gemm_config* cfg, int64_t cap, int64_t* out, int64_t* count);

/* Diagnostic: how many clusters of `cluster_size` CTAs (flagship kernel, `threads` per CTA) fit
 * on the current device at once (clusters are confined to a GPC). */
int ddlb_max_active_clusters(int cluster_size, int threads, int* out);

/* Diagnostic: copy `bytes` (multiple of 16) with a 128-bit load/store kernel on `blocks` CTAs
 * (0 = 4 per SM). With a peer-mapped src it measures pull bandwidth, with a peer-mapped dst push. */
int ddlb_copy(void* dst, const void* src, int64_t bytes, int blocks, void* stream);

/* Diagnostic: the same copy with the access patterns the fused kernels can use (0 dense
 * grid-stride, 1 per-warp units of unit_bytes, 2/3 warps interleaved at 4 KiB with 8/16 KiB in
 * flight per warp) on blocks x threads. See csrc/runtime.inl.
 * smem_bytes of dynamic shared memory are requested only to shrink the L1 like the GEMM does;
 * flags: 1 = loads bypass L1, 2 = fence + counter bump per unit, 4 = next unit prefetched. */
int ddlb_copy_pattern(void* dst, const void* src, int64_t bytes, int blocks, int threads,
                      int pattern, int unit_bytes, int smem_bytes, int flags, void* stream);

/* device-side error word set by a kernel that timed out waiting ({code, block, aux0, aux1}) */
int ddlb_read_error_word(uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif /* DDLB_B200_H_ */

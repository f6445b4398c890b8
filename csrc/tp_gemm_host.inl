// Host side of the fused GEMM: config resolution, TMA tensor maps, launch.
// Included by ddlb_b200.cu (unity build).

namespace ddlb {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) !=
            cudaSuccess || qres != cudaDriverEntryPointSuccess || sym == nullptr) {
      (void)cudaGetLastError();
      return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// rank-2/3 tiled map of 16-bit (elem = 2) or 32-bit (elem = 4) elements, 128-byte swizzle
static int make_tmap(CUtensorMap* map, void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, int elem = 2) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
  cuuint64_t gdims[3];
  cuuint64_t gstrides[2];
  cuuint32_t gbox[3];
  cuuint32_t estr[3] = {1, 1, 1};
  for (int i = 0; i < rank; ++i) { gdims[i] = dims[i]; gbox[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) gstrides[i] = strides_bytes[i];
  CUresult r = fn(map, elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT16,
                  static_cast<cuuint32_t>(rank), base, gdims,
                  gstrides, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error("cuTensorMapEncodeTiled failed with CUresult %d (base=%p rank=%d dims=%llu,%llu)",
                     static_cast<int>(r), base, rank, static_cast<unsigned long long>(dims[0]),
                     static_cast<unsigned long long>(dims[1]));
  return 0;
}

struct Resolved {
  int cta_group, block_n, group_m, grid, chunk_rows;
  int tile_m, part_rows, a_rows, kk;   // kk: contraction length seen by this kernel
  int elem;      // bytes per element: 2 (fp16/bf16) or 4 (fp32 through TF32)
  int split;     // 1: 3xTF32 (fp32 only)
  int threads;   // 256, or 384 when the aux warps 8..11 have work (collectives, 3xTF32)
  int part0;     // first shard in visiting order
  int pairs;     // MMA pairs per cluster: 2 = four-CTA clusters with the A tile multicast
  int wide;      // 1: a scheduled tile is two neighbouring n-blocks sharing every A stage
};

static int resolve(const ddlb_tp_gemm_config& c, Resolved& r) {
  if (c.world < 1 || c.world > DDLB_MAX_WORLD || c.rank < 0 || c.rank >= c.world)
    return set_error("bad rank/world %d/%d", c.rank, c.world);
  if (c.mode < 0 || c.mode > 3) return set_error("bad mode %d", c.mode);
  if (c.dtype != DDLB_DTYPE_F16 && c.dtype != DDLB_DTYPE_BF16 && c.dtype != DDLB_DTYPE_F32)
    return set_error("bad dtype %d", c.dtype);
  r.elem = c.dtype == DDLB_DTYPE_F32 ? 4 : 2;
  if (r.elem == 4 && c.b_layout != DDLB_B_ROW_MAJOR_NK)
    return set_error("float32 needs B as [n,k] (DDLB_B_ROW_MAJOR_NK): the tensor cores read an "
                     "N-major tf32 operand in a different shared-memory layout; transpose the "
                     "weight once");
  if (c.fp32_mode != DDLB_FP32_3XTF32 && c.fp32_mode != DDLB_FP32_TF32) return set_error("bad fp32_mode");
  r.split = (r.elem == 4 && c.fp32_mode == DDLB_FP32_3XTF32) ? 1 : 0;
  const int row_elems = 128 / r.elem;          // elements per 128-byte swizzle row
  if (c.m <= 0 || c.n <= 0 || c.k <= 0) return set_error("bad shape");
  if (c.m % c.world != 0) return set_error("m (%lld) must be divisible by world (%d)", (long long)c.m, c.world);
  r.part_rows = static_cast<int>(c.m / c.world);
  r.kk = static_cast<int>(c.mode == DDLB_MODE_RS ? c.k / c.world : c.k);
  if (c.mode == DDLB_MODE_RS && c.k % c.world != 0)
    return set_error("k (%lld) must be divisible by world (%d)", (long long)c.k, c.world);
  if (r.kk % (16 / r.elem) != 0)
    return set_error("contraction length %d must be a multiple of %d (16-byte rows)", r.kk, 16 / r.elem);
  if (c.n % (16 / r.elem) != 0) return set_error("n must be a multiple of %d (16-byte rows)", 16 / r.elem);
  if (c.b_layout == DDLB_B_ROW_MAJOR_KN && c.n % row_elems != 0)
    return set_error("n (%lld) must be a multiple of %d when B is [k,n] row-major", (long long)c.n, row_elems);
  r.a_rows = static_cast<int>(c.mode == DDLB_MODE_AG_AFTER ? r.part_rows : c.m);

  const bool sharded_rows = c.mode != DDLB_MODE_PLAIN;   // tiles must not straddle shards
  r.cta_group = c.cta_group;
  if (r.cta_group == 0) r.cta_group = (!sharded_rows || r.part_rows % 256 == 0) ? 2 : 1;
  if (r.cta_group != 1 && r.cta_group != 2) return set_error("cta_group must be 1 or 2");
  r.tile_m = kBlockM * r.cta_group;
  if (sharded_rows && r.part_rows % r.tile_m != 0)
    return set_error("rows per rank (%d) must be a multiple of %d for the fused kernel", r.part_rows, r.tile_m);
  r.block_n = c.block_n != 0 ? c.block_n : (c.n <= 128 ? 128 : 256);
  if (r.split && r.cta_group == 1 && c.block_n == 0) r.block_n = 128;   // hi+lo ring must fit
  if (r.block_n != 128 && r.block_n != 256) return set_error("block_n must be 128 or 256");
  if (r.split && r.cta_group == 1 && r.block_n == 256)
    return set_error("3xTF32 with cta_group=1 needs block_n=128 (shared memory)");

  // two MMA pairs per cluster (A multicast): 16-bit, cta_group 2, BLOCK_N 256, an even number
  // of n-blocks. Explicit request only for now (cluster_pairs = 2); 0 and 1 mean one pair.
  if (c.cluster_pairs < 0 || c.cluster_pairs > 2) return set_error("cluster_pairs must be 0, 1 or 2");
  r.pairs = c.cluster_pairs == 2 ? 2 : 1;
  if (r.pairs == 2) {
    const int64_t nb = (c.n + r.block_n - 1) / r.block_n;
    if (r.elem != 2 || r.cta_group != 2 || r.block_n != 256 || nb % 2 != 0)
      return set_error("cluster_pairs=2 needs fp16/bf16, cta_group=2, block_n=256 and an even number "
                       "of 256-column blocks (n = %lld)", (long long)c.n);
  }
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    (void)cudaGetLastError();
    // no device (schedule introspection on a CPU box): an explicit num_sms is taken as is
    if (c.num_sms <= 0) return set_error("cannot query the current CUDA device");
    sms = c.num_sms;
  }
  r.grid = c.num_sms > 0 ? (c.num_sms < sms ? c.num_sms : sms) : sms;
  // Small single-rank problems with the tile shape left on auto: a 256x256 tile per CTA pair
  // leaves most SMs idle when there are far fewer tiles than clusters, and both the main loop and
  // the (exposed) epilogue of the only wave scale with the tile. Halve it (BLOCK_N 128: every CTA
  // of a pair computes 128x128) while one wave still holds all tiles. Measured at 1024^3: 16 tiles
  // on 32 SMs took 13.8 us against cuBLAS' 8.5 us.
  if (c.mode == DDLB_MODE_PLAIN && c.cta_group == 0 && c.block_n == 0 && r.pairs == 1 && r.elem == 2 &&
      c.n > 128 && c.wide_n != 1) {                      // a wide-tile request implies BLOCK_N 256
    const int64_t mb256 = (c.m + 255) / 256, nb128 = (c.n + 127) / 128;
    if (mb256 * nb128 <= r.grid / 2) r.block_n = 128;
  }
  // Wide tiles (256 x 512 per CTA pair, both accumulators in TMEM at once): 25 % fewer operand
  // bytes per MMA, but a 4-deep operand ring and an epilogue that is no longer hidden behind the
  // next tile's main loop. 16-bit, cta_group 2, BLOCK_N 256, an even number of n-blocks.
  // Explicit request only: measured against the 256x256 tiles it is 0.80x at k = 1024, 0.91 -
  // 0.99x at k = 4096 .. 16384 and 1.03x only where it removes a partial wave (8192x1024x8192);
  // sustained large GEMMs are not bound by the L2 -> SM operand stream.
  if (c.wide_n < 0 || c.wide_n > 2) return set_error("wide_n must be 0 (auto), 1 (on) or 2 (off)");
  {
    const int64_t nb = (c.n + r.block_n - 1) / r.block_n;
    const bool can = r.elem == 2 && r.cta_group == 2 && r.block_n == 256 && r.pairs == 1 && nb % 2 == 0 &&
                     c.trace == nullptr;
    if (c.wide_n == 1 && !can)
      return set_error("wide_n=1 needs fp16/bf16, cta_group=2, block_n=256, one pair per cluster, no "
                       "trace and an even number of 256-column blocks (n = %lld)", (long long)c.n);
    r.wide = c.wide_n == 1 ? 1 : 0;
  }
  const int cluster_ctas = r.cta_group * r.pairs;
  r.grid -= r.grid % cluster_ctas;
  if (r.grid < cluster_ctas) return set_error("no SMs to launch on");

  r.group_m = c.group_m > 0 ? c.group_m : 8;

  r.chunk_rows = r.part_rows;
  if (c.mode == DDLB_MODE_AG_BEFORE || c.mode == DDLB_MODE_RS) {
    int cr = c.chunk_rows;
    if (cr <= 0) {
      // auto: about 1 MiB of A per chunk (AG) / the whole shard (RS), dividing the shard evenly
      const int mblks = r.part_rows / r.tile_m;
      int want = c.mode == DDLB_MODE_AG_BEFORE
                     ? static_cast<int>((1 << 20) / (static_cast<int64_t>(r.kk) * r.elem * r.tile_m))
                     : mblks;
      if (want < 1) want = 1;
      if (want > mblks) want = mblks;
      while (mblks % want != 0) --want;
      cr = want * r.tile_m;
    }
    if (cr % r.tile_m != 0 || r.part_rows % cr != 0)
      return set_error("chunk_rows (%d) must be a multiple of %d and divide rows per rank (%d)", cr,
                       r.tile_m, r.part_rows);
    r.chunk_rows = cr;
    const int64_t chunk_bytes = static_cast<int64_t>(r.chunk_rows) * r.kk * r.elem;
    if (c.mode == DDLB_MODE_AG_BEFORE && chunk_bytes > 0xffffffffll) return set_error("chunk too large");
  }
  if (c.ag_transport < 0 || c.ag_transport > 2) return set_error("bad ag_transport %d", c.ag_transport);
  if (c.epi_transport < 0 || c.epi_transport > 2) return set_error("bad epi_transport %d", c.epi_transport);
  if (c.rs_transport < 0 || c.rs_transport > 2) return set_error("bad rs_transport %d", c.rs_transport);
  if (c.epilogue_act < 0 || c.epilogue_act > 3) return set_error("bad epilogue_act %d", c.epilogue_act);
  if (c.pdl < 0 || c.pdl > 2) return set_error("pdl must be 0 (auto), 1 (on) or 2 (off), got %d", c.pdl);
  if (c.epilogue_act != 0 && c.mode == DDLB_MODE_RS && c.world > 1)
    return set_error("an epilogue activation cannot be applied to partial sums (tp_rowwise)");

  // visiting order of the shards: AG starts with the own shard (it is already here), RS push
  // ends with it (its partials need no NVLink); the pull reductions and the un-rotated
  // order visit 0..d-1 on every rank so that a shard completes everywhere at the same time
  r.part0 = 0;
  if (!c.unrotated) {
    if (c.mode == DDLB_MODE_AG_BEFORE) r.part0 = c.rank;
    if (c.mode == DDLB_MODE_RS && c.rs_transport == DDLB_RS_PUSH) r.part0 = (c.rank + 1) % c.world;
  }
  const bool collective = c.world > 1 && (c.mode == DDLB_MODE_AG_BEFORE || c.mode == DDLB_MODE_RS);
  r.threads = (collective || r.split) ? kMaxThreads : kBaseThreads;
  return 0;
}

}  // namespace ddlb

struct ddlb_tp_gemm_plan {
  ddlb::TpGemmParams params;
  ddlb::Resolved res;
  ddlb_tp_gemm_config cfg;
  void (*kernel)(ddlb::TpGemmParams);
  void (*emu_kernel)(const ddlb::TpGemmParams*, int);
  int smem_bytes;
  int stages;
};

namespace ddlb {

static void fill_launch(cudaLaunchConfig_t& lc, cudaLaunchAttribute* attr, unsigned grid, int threads,
                        int smem_bytes, int cta_group, cudaStream_t st);

template <int G, int BN, bool NM, bool F32 = false, bool SPLIT = false, int PAIRS = 1, bool WIDE = false>
static int bind_kernel(ddlb_tp_gemm_plan* plan) {
  using Cfg = TpGemmCfg<G, BN, SPLIT, WIDE>;
  auto k = tp_gemm_kernel<G, BN, NM, F32, SPLIT, false, PAIRS, WIDE>;
  plan->kernel = reinterpret_cast<void (*)(TpGemmParams)>(k);
  auto ke = tp_gemm_emulated_kernel<G, BN, NM, F32, SPLIT, PAIRS, WIDE>;
  plan->emu_kernel = ke;
  plan->stages = Cfg::kStages;
  plan->smem_bytes = Cfg::kSmemBytes;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(ke, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error("cudaFuncSetAttribute(smem=%d): %s", Cfg::kSmemBytes, cudaGetErrorString(e));
  }
  if (G == 2) {
    e = cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) (void)cudaGetLastError();
  }
  if (PAIRS == 2) {
    // clusters are confined to a GPC: launch as many four-CTA clusters as can be resident
    cudaLaunchConfig_t lc;
    cudaLaunchAttribute attr[1];
    fill_launch(lc, attr, static_cast<unsigned>(plan->res.grid), plan->res.threads, Cfg::kSmemBytes,
                G * PAIRS, nullptr);
    int fit = 0;
    e = cudaOccupancyMaxActiveClusters(&fit, reinterpret_cast<const void*>(k), &lc);
    if (e != cudaSuccess || fit < 1) {
      (void)cudaGetLastError();
      return set_error("cudaOccupancyMaxActiveClusters(cluster of 4): %s", cudaGetErrorString(e));
    }
    if (fit * G * PAIRS < plan->res.grid) plan->res.grid = fit * G * PAIRS;
  }
  return 0;
}

static int select_kernel(ddlb_tp_gemm_plan* plan) {
  const int g = plan->res.cta_group, bn = plan->res.block_n;
  const bool nm = plan->cfg.b_layout == DDLB_B_ROW_MAJOR_KN;
  if (plan->cfg.trace != nullptr) {
    // traced variant exists for the flagship configuration only (16-bit, cta_group 2, BN 256)
    if (plan->res.elem == 2 && g == 2 && bn == 256 && nm) {
      using Cfg = TpGemmCfg<2, 256>;
      auto k = tp_gemm_kernel<2, 256, true, false, false, true>;
      plan->kernel = reinterpret_cast<void (*)(TpGemmParams)>(k);
      plan->emu_kernel = nullptr;
      cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
      if (e != cudaSuccess) { (void)cudaGetLastError(); return set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
      (void)cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
      (void)cudaGetLastError();
      plan->stages = Cfg::kStages;
      plan->smem_bytes = Cfg::kSmemBytes;
      return 0;
    }
    return set_error("tracing is compiled for fp16/bf16, cta_group=2, block_n=256, B [k,n] only");
  }
  if (plan->res.elem == 4) {      // fp32 on kind::tf32, B [n,k]
    if (plan->res.split) {
      if (g == 2 && bn == 256) return bind_kernel<2, 256, false, true, true>(plan);
      if (g == 1 && bn == 128) return bind_kernel<1, 128, false, true, true>(plan);
      if (g == 2 && bn == 128) return bind_kernel<2, 128, false, true, true>(plan);
      return set_error("no 3xTF32 kernel for cta_group=%d block_n=%d", g, bn);
    }
    if (g == 1 && bn == 256) return bind_kernel<1, 256, false, true>(plan);
    if (g == 2 && bn == 256) return bind_kernel<2, 256, false, true>(plan);
    if (g == 1 && bn == 128) return bind_kernel<1, 128, false, true>(plan);
    if (g == 2 && bn == 128) return bind_kernel<2, 128, false, true>(plan);
    return set_error("no fp32 kernel for cta_group=%d block_n=%d", g, bn);
  }
  if (plan->res.pairs == 2)
    return nm ? bind_kernel<2, 256, true, false, false, 2>(plan) : bind_kernel<2, 256, false, false, false, 2>(plan);
  if (plan->res.wide)
    return nm ? bind_kernel<2, 256, true, false, false, 1, true>(plan)
              : bind_kernel<2, 256, false, false, false, 1, true>(plan);
  if (g == 1 && bn == 256) return nm ? bind_kernel<1, 256, true>(plan) : bind_kernel<1, 256, false>(plan);
  if (g == 2 && bn == 256) return nm ? bind_kernel<2, 256, true>(plan) : bind_kernel<2, 256, false>(plan);
  if (g == 1 && bn == 128) return nm ? bind_kernel<1, 128, true>(plan) : bind_kernel<1, 128, false>(plan);
  if (g == 2 && bn == 128) return nm ? bind_kernel<2, 128, true>(plan) : bind_kernel<2, 128, false>(plan);
  return set_error("no kernel for cta_group=%d block_n=%d", g, bn);
}

}  // namespace ddlb

namespace ddlb {
// The part of the kernel parameters that defines the tile schedule (shared by plan creation
// and by the host-side introspection entry points below).
static void fill_schedule(TpGemmParams& p, const ddlb_tp_gemm_config& c, const Resolved& r) {
  p.mode = c.mode;
  p.rank = c.rank;
  p.world = c.world;
  p.num_m_blocks = (r.a_rows + r.tile_m - 1) / r.tile_m;
  p.num_n_blocks = static_cast<int>((c.n + r.block_n - 1) / r.block_n);
  p.sched_n_blocks = p.num_n_blocks / (r.pairs * (r.wide ? 2 : 1));
  p.group_m = r.group_m;
  p.part_rows = r.part_rows;
  p.out_n = static_cast<int>(c.n);
  p.block_n = r.block_n;
  p.elem = r.elem;
  if (c.mode == DDLB_MODE_AG_BEFORE || c.mode == DDLB_MODE_RS) {
    p.mblks_per_part = r.part_rows / r.tile_m;
    p.chunk_mblks = r.chunk_rows / r.tile_m;
  } else {
    p.mblks_per_part = p.num_m_blocks;
    p.chunk_mblks = p.num_m_blocks;
  }
  p.part0 = r.part0;
  p.ag_transport = c.ag_transport;
  p.epi_transport = c.epi_transport;
  p.rs_transport = c.rs_transport;
  p.act = c.epilogue_act;
  {
    // programmatic dependent launch: auto = single-rank plans only (the multi-rank kernels wait for
    // peers, and their overlap with the previous launch has not been measured on real NVLink)
    const char* env = getenv("DDLB_PDL");
    const bool env_off = env != nullptr && env[0] == '0';
    p.pdl = (!env_off && (c.pdl == 1 || (c.pdl == 0 && c.world == 1))) ? 1 : 0;
  }
  p.comm_warps = c.comm_warps;
}
}  // namespace ddlb

extern "C" int64_t ddlb_tp_gemm_sig_words(const ddlb_tp_gemm_config* cfg) {
  if (cfg == nullptr) return -1;
  ddlb::Resolved r;
  if (ddlb::resolve(*cfg, r) != 0) return -1;
  int64_t words = ddlb::kSigChunk;
  if (cfg->mode == DDLB_MODE_AG_BEFORE) words += static_cast<int64_t>(cfg->world) * (r.part_rows / r.chunk_rows);
  if (cfg->mode == DDLB_MODE_RS) {
    const int64_t nb = (cfg->n + r.block_n - 1) / r.block_n;
    words += static_cast<int64_t>(r.part_rows / ddlb::kBlockM) * nb;
  }
  return (words + 63) / 64 * 64;
}

extern "C" int ddlb_tp_gemm_create(const ddlb_tp_gemm_config* cfg, ddlb_tp_gemm_plan** out) {
  using namespace ddlb;
  if (cfg == nullptr || out == nullptr) return set_error("null argument");
  *out = nullptr;
  ddlb_tp_gemm_config normalized = *cfg;
  if (normalized.world == 1) normalized.mode = DDLB_MODE_PLAIN;   // one rank: nothing to gather or reduce
  Resolved r;
  if (int rc = resolve(normalized, r)) return rc;
  const ddlb_tp_gemm_config& c = normalized;
  if (c.a == nullptr || c.b == nullptr || c.out[c.rank] == nullptr) return set_error("null a/b/out pointer");
  if (r.pairs == 2) {
    // The A tile is fetched with a TMA multicast load. TMA multicast loads here read only this
    // GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU:
    // refuse an A operand that is not resident on the current device (first and last byte).
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return set_error("cluster_pairs=2: no current device"); }
    const size_t a_bytes = static_cast<size_t>(r.a_rows) * static_cast<size_t>(r.kk) * r.elem;
    const char* ends[2] = {static_cast<const char*>(c.a), static_cast<const char*>(c.a) + a_bytes - 1};
    for (const char* q : ends) {
      cudaPointerAttributes attr;
      cudaError_t e = cudaPointerGetAttributes(&attr, q);
      if (e != cudaSuccess || attr.type != cudaMemoryTypeDevice || attr.device != dev) {
        (void)cudaGetLastError();
        return set_error("cluster_pairs=2 multicasts the A tile and needs A in the current device's own "
                         "memory (device %d); got memory of type %d on device %d", dev,
                         e == cudaSuccess ? static_cast<int>(attr.type) : -1,
                         e == cudaSuccess ? attr.device : -1);
      }
    }
  }
  if (c.world > 1 && c.mode != DDLB_MODE_PLAIN) {
    if (c.sig_local == nullptr) return set_error("sig_local is required for fused modes");
    for (int p = 0; p < c.world; ++p) {
      if (p == c.rank) continue;
      if (c.sig_peer[p] == nullptr) return set_error("sig_peer[%d] is null", p);
      if (c.mode == DDLB_MODE_AG_BEFORE && c.peer_a[p] == nullptr) return set_error("peer_a[%d] is null", p);
      if (c.mode == DDLB_MODE_AG_AFTER && c.out[p] == nullptr) return set_error("out[%d] is null", p);
      if (c.mode == DDLB_MODE_RS && c.stage[p] == nullptr) return set_error("stage[%d] is null", p);
    }
    if (c.mode == DDLB_MODE_RS && c.stage[c.rank] == nullptr) return set_error("stage[rank] is null");
    const bool wants_mc = (c.mode == DDLB_MODE_AG_BEFORE && c.ag_transport == DDLB_AG_PUSH_MULTICAST) ||
                          (c.mode == DDLB_MODE_AG_AFTER && c.epi_transport == DDLB_EPI_MULTICAST) ||
                          (c.mode == DDLB_MODE_RS && c.rs_transport == DDLB_RS_PULL_MULTICAST);
    if (wants_mc && c.mc_ptr == nullptr)
      return set_error("a multicast transport was selected but mc_ptr is null (no NVLS window)");
  }

  auto* plan = new ddlb_tp_gemm_plan();
  plan->cfg = c;
  plan->res = r;
  if (int rc = select_kernel(plan)) { delete plan; return rc; }

  TpGemmParams& p = plan->params;
  memset(&p, 0, sizeof(p));
  const uint64_t N = static_cast<uint64_t>(c.n), K = static_cast<uint64_t>(r.kk);
  const int load_n = r.block_n / r.cta_group;
  const uint64_t E = static_cast<uint64_t>(r.elem);
  const uint32_t row_elems = static_cast<uint32_t>(128 / r.elem);   // 64 (16-bit) or 32 (fp32)
  int rc = 0;
  {  // A: [a_rows, K] row-major, box = one 128-byte row segment x 128 rows
    const uint64_t dims[2] = {K, static_cast<uint64_t>(r.a_rows)};
    const uint64_t strides[1] = {K * E};
    // two pairs per cluster: every CTA loads half of its 128 rows and multicasts them
    const uint32_t box[2] = {row_elems, static_cast<uint32_t>(kBlockM / r.pairs)};
    rc = make_tmap(&p.tmA, c.a, 2, dims, strides, box, r.elem);
  }
  if (rc == 0) {
    if (c.b_layout == DDLB_B_ROW_MAJOR_KN) {  // [K, N]: 3-D view {128 B of n, K, N / that}
      const uint64_t dims[3] = {row_elems, K, N / row_elems};
      const uint64_t strides[2] = {N * E, 128};
      const uint32_t box[3] = {row_elems, row_elems, static_cast<uint32_t>(load_n) / row_elems};
      rc = make_tmap(&p.tmB, c.b, 3, dims, strides, box, r.elem);
    } else {                                  // [N, K]: box = 128 B of k x load_n rows
      const uint64_t dims[2] = {K, N};
      const uint64_t strides[1] = {K * E};
      const uint32_t box[2] = {row_elems, static_cast<uint32_t>(load_n)};
      rc = make_tmap(&p.tmB, c.b, 2, dims, strides, box, r.elem);
    }
  }
  // output maps (TMA stores): C of this rank; in AG_AFTER C of every rank. RS stores with
  // plain 128-bit stores and needs none.
  for (int d = 0; d < c.world && rc == 0; ++d) {
    if (c.mode == DDLB_MODE_RS) break;
    if (c.out[d] == nullptr) continue;
    if (c.mode != DDLB_MODE_AG_AFTER && d != c.rank) continue;
    const uint64_t dims[2] = {N, static_cast<uint64_t>(c.m)};
    const uint64_t strides[1] = {N * E};
    const uint32_t box[2] = {row_elems, 32};
    rc = make_tmap(&p.tmOut[d], c.out[d], 2, dims, strides, box, r.elem);
  }
  if (rc != 0) { delete plan; return rc; }

  fill_schedule(p, c, r);
  p.num_k_blocks = (r.kk + static_cast<int>(row_elems) - 1) / static_cast<int>(row_elems);
  p.idesc = make_idesc(c.dtype == DDLB_DTYPE_F32 ? 2u : (c.dtype == DDLB_DTYPE_BF16 ? 1u : 0u),
                       static_cast<uint32_t>(r.tile_m),
                       static_cast<uint32_t>(r.block_n), false, c.b_layout == DDLB_B_ROW_MAJOR_KN);
  p.epoch = 0;
  p.local_a = static_cast<uint8_t*>(c.a);
  for (int d = 0; d < c.world; ++d) {
    p.peer_a[d] = static_cast<const uint8_t*>(c.peer_a[d]);
    p.sig_peer[d] = c.sig_peer[d];
    p.stage_base[d] = static_cast<uint8_t*>(c.stage[d]);
    p.out_raw[d] = static_cast<uint8_t*>(c.out[d]);
  }
  p.peer_a[c.rank] = static_cast<const uint8_t*>(c.a);
  p.chunk_bytes = static_cast<uint32_t>(static_cast<int64_t>(r.chunk_rows) * r.kk * r.elem);
  p.stages = plan->stages;
  p.rs_tiles = (r.part_rows / kBlockM) * p.num_n_blocks;
  p.mc_ptr = static_cast<uint8_t*>(c.mc_ptr);
  p.sig_local = c.sig_local;
  p.sig_peer[c.rank] = c.sig_local;
  p.trace = static_cast<uint4*>(c.trace);
  *out = plan;
  return 0;
}

namespace ddlb {
static void fill_launch(cudaLaunchConfig_t& lc, cudaLaunchAttribute* attr, unsigned grid, int threads,
                        int smem_bytes, int cta_group, cudaStream_t st) {
  memset(&lc, 0, sizeof(lc));
  lc.gridDim = dim3(grid, 1, 1);
  lc.blockDim = dim3(static_cast<unsigned>(threads), 1, 1);
  lc.dynamicSmemBytes = static_cast<size_t>(smem_bytes);
  lc.stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned>(cta_group);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  lc.attrs = attr;
  lc.numAttrs = 1;
}
}  // namespace ddlb

extern "C" int ddlb_tp_gemm_run(ddlb_tp_gemm_plan* plan, uint32_t epoch, void* stream) {
  using namespace ddlb;
  if (plan == nullptr) return set_error("null plan");
  plan->params.epoch = epoch;
  cudaLaunchConfig_t lc;
  cudaLaunchAttribute attr[2];
  fill_launch(lc, attr, static_cast<unsigned>(plan->res.grid), plan->res.threads, plan->smem_bytes,
              plan->res.cta_group * plan->res.pairs, static_cast<cudaStream_t>(stream));
  if (plan->params.pdl) {
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    lc.numAttrs = 2;
  }
  void* args[1] = {&plan->params};
  cudaError_t e = cudaLaunchKernelExC(&lc, reinterpret_cast<const void*>(plan->kernel), args);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error("kernel launch failed: %s", cudaGetErrorString(e));
  }
  return 0;
}

extern "C" int ddlb_tp_gemm_info(const ddlb_tp_gemm_plan* plan, int32_t out[12]) {
  if (plan == nullptr || out == nullptr) return ddlb::set_error("null argument");
  out[0] = plan->res.grid;
  out[1] = plan->res.cta_group * plan->res.pairs;
  out[2] = plan->smem_bytes;
  out[3] = plan->res.cta_group;
  out[4] = plan->res.block_n;
  out[5] = plan->stages;
  out[6] = plan->res.threads;
  out[7] = plan->res.chunk_rows;
  out[8] = plan->res.pairs;
  out[9] = plan->res.wide;
  out[10] = plan->params.pdl;
  out[11] = 0;
  return 0;
}

extern "C" int ddlb_tp_gemm_destroy(ddlb_tp_gemm_plan* plan) {
  delete plan;
  return 0;
}

// Test helper: run `world` plans (one per emulated rank, all on the current device)
// as ONE launch; see tp_gemm_emulated_kernel.
extern "C" int ddlb_tp_gemm_run_emulated(ddlb_tp_gemm_plan** plans, int world, uint32_t epoch,
                                         void* stream) {
  using namespace ddlb;
  if (plans == nullptr || world < 1 || world > DDLB_MAX_WORLD) return set_error("bad arguments");
  const ddlb_tp_gemm_plan* first = plans[0];
  for (int r = 0; r < world; ++r) {
    if (plans[r] == nullptr) return set_error("null plan %d", r);
    if (plans[r]->emu_kernel == nullptr) return set_error("traced plans cannot be emulated");
    if (plans[r]->emu_kernel != first->emu_kernel || plans[r]->res.cta_group != first->res.cta_group ||
        plans[r]->res.threads != first->res.threads || plans[r]->res.pairs != first->res.pairs ||
        plans[r]->res.wide != first->res.wide)
      return set_error("emulated ranks must use the same kernel variant");
  }
  int dev = 0, sms = 0;
  DDLB_CUDA(cudaGetDevice(&dev));
  DDLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int g = first->res.cta_group * first->res.pairs;   // CTAs per cluster
  int clusters_per_rank = (sms / world) / g;
  const int want = first->res.grid / g;            // honour a smaller num_sms request
  if (first->cfg.num_sms > 0 && want < clusters_per_rank) clusters_per_rank = want;
  if (first->res.pairs == 2) {
    // four-CTA clusters are confined to a GPC: fewer fit than sms / 4, and an emulated launch
    // must be fully resident
    cudaLaunchConfig_t probe;
    cudaLaunchAttribute pattr[1];
    fill_launch(probe, pattr, static_cast<unsigned>(sms - sms % g), first->res.threads, first->smem_bytes, g,
                nullptr);
    int fit = 0;
    DDLB_CUDA(cudaOccupancyMaxActiveClusters(&fit, reinterpret_cast<const void*>(first->emu_kernel), &probe));
    if (fit / world < clusters_per_rank) clusters_per_rank = fit / world;
  }
  if (clusters_per_rank < 1) return set_error("no cluster left per emulated rank");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  TpGemmParams* dparams = nullptr;
  DDLB_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&dparams), sizeof(TpGemmParams) * world, st));
  for (int r = 0; r < world; ++r) {
    plans[r]->params.epoch = epoch;
    DDLB_CUDA(cudaMemcpyAsync(dparams + r, &plans[r]->params, sizeof(TpGemmParams),
                              cudaMemcpyHostToDevice, st));
  }
  cudaLaunchConfig_t lc;
  cudaLaunchAttribute attr[1];
  fill_launch(lc, attr, static_cast<unsigned>(clusters_per_rank * g * world), first->res.threads,
              first->smem_bytes, g, st);
  const TpGemmParams* carg = dparams;
  void* args[2] = {&carg, &clusters_per_rank};
  cudaError_t e = cudaLaunchKernelExC(&lc, reinterpret_cast<const void*>(first->emu_kernel), args);
  cudaFreeAsync(dparams, st);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error("emulated launch failed: %s", cudaGetErrorString(e));
  }
  return 0;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of ddlb_tp_gemm_run_emulated.
//
// The emulated kernel reads one parameter block per rank (tensor maps, flag and
// buffer addresses) from global memory. ddlb_tp_gemm_run_emulated obtains that
// scratch from cudaMallocAsync on every call. Some hosts of this library cannot do
// that: a process that handed every remaining byte of the device to its own pool at
// start-up (a serving engine's KV-cache arena, a framework allocator with a hard
// cap) has nothing left for the driver to allocate from, and control loops that
// must not block forbid calls that may take the allocator's lock or trim a pool.
// For them the scratch lives in the array below: fixed size, part of the module
// image, held by every device that loads the library until the process exits.
//
// Capacity: DDLB_EMULATED_STATIC_MAX_WORLD parameter blocks (about 1.6 KiB each).
// ---------------------------------------------------------------------------
namespace ddlb {

static __device__ __align__(256) unsigned char
    g_emulated_param_scratch[DDLB_EMULATED_STATIC_MAX_WORLD * sizeof(TpGemmParams)];

constexpr int kMaxTrackedDevices = 64;
// Device address of the array above, per device ordinal. In host code the symbol name
// is only a handle: it must go through cudaGetSymbolAddress with the device current.
static TpGemmParams* g_emulated_scratch_addr[kMaxTrackedDevices] = {};

static int emulated_scratch_for_current_device(TpGemmParams** out) {
  int dev = 0;
  DDLB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxTrackedDevices) return set_error("device ordinal %d is not supported", dev);
  if (g_emulated_scratch_addr[dev] == nullptr) {
    void* addr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&addr, g_emulated_param_scratch);
    if (e != cudaSuccess || addr == nullptr) {
      (void)cudaGetLastError();
      return set_error("cudaGetSymbolAddress(static scratch) on device %d: %s", dev,
                       cudaGetErrorString(e));
    }
    g_emulated_scratch_addr[dev] = static_cast<TpGemmParams*>(addr);
  }
  *out = g_emulated_scratch_addr[dev];
  return 0;
}

}  // namespace ddlb

extern "C" int ddlb_tp_gemm_run_emulated_static(ddlb_tp_gemm_plan** plans, int world,
                                                uint32_t epoch, void* stream) {
  using namespace ddlb;
  if (plans == nullptr || world < 1) return set_error("bad arguments");
  if (world > DDLB_EMULATED_STATIC_MAX_WORLD)
    return set_error("world %d exceeds the static scratch of this entry point (%d ranks); use "
                     "ddlb_tp_gemm_run_emulated, which allocates", world,
                     DDLB_EMULATED_STATIC_MAX_WORLD);
  const ddlb_tp_gemm_plan* first = plans[0];
  for (int r = 0; r < world; ++r) {
    if (plans[r] == nullptr) return set_error("null plan %d", r);
    if (plans[r]->emu_kernel == nullptr) return set_error("traced plans cannot be emulated");
    if (plans[r]->emu_kernel != first->emu_kernel || plans[r]->res.cta_group != first->res.cta_group ||
        plans[r]->res.threads != first->res.threads || plans[r]->res.pairs != first->res.pairs ||
        plans[r]->res.wide != first->res.wide)
      return set_error("emulated ranks must use the same kernel variant");
  }
  int dev = 0, sms = 0;
  DDLB_CUDA(cudaGetDevice(&dev));
  DDLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int g = first->res.cta_group * first->res.pairs;   // CTAs per cluster
  int clusters_per_rank = (sms / world) / g;
  const int want = first->res.grid / g;
  if (first->cfg.num_sms > 0 && want < clusters_per_rank) clusters_per_rank = want;
  if (first->res.pairs == 2) {
    // four-CTA clusters are confined to a GPC: fewer fit than sms / 4, and an emulated launch
    // must be fully resident
    cudaLaunchConfig_t probe;
    cudaLaunchAttribute pattr[1];
    fill_launch(probe, pattr, static_cast<unsigned>(sms - sms % g), first->res.threads, first->smem_bytes, g,
                nullptr);
    int fit = 0;
    DDLB_CUDA(cudaOccupancyMaxActiveClusters(&fit, reinterpret_cast<const void*>(first->emu_kernel), &probe));
    if (fit / world < clusters_per_rank) clusters_per_rank = fit / world;
  }
  if (clusters_per_rank < 1) return set_error("no cluster left per emulated rank");

  TpGemmParams* dparams = nullptr;
  if (int rc = emulated_scratch_for_current_device(&dparams)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Every block the kernel will read is rewritten here, so stale contents of earlier
  // calls are never observed and nothing in this scratch has to start at zero.
  for (int r = 0; r < world; ++r) {
    plans[r]->params.epoch = epoch;
    DDLB_CUDA(cudaMemcpyAsync(dparams + r, &plans[r]->params, sizeof(TpGemmParams),
                              cudaMemcpyHostToDevice, st));
  }
  cudaLaunchConfig_t lc;
  cudaLaunchAttribute attr[1];
  fill_launch(lc, attr, static_cast<unsigned>(clusters_per_rank * g * world), first->res.threads,
              first->smem_bytes, g, st);
  const TpGemmParams* carg = dparams;
  void* args[2] = {&carg, &clusters_per_rank};
  cudaError_t e = cudaLaunchKernelExC(&lc, reinterpret_cast<const void*>(first->emu_kernel), args);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error("emulated launch failed: %s", cudaGetErrorString(e));
  }
  return 0;
}

// How many clusters of `cluster_size` CTAs of the flagship kernel (cta_group 2, BLOCK_N 256,
// full shared memory) can be resident at once on the current device. Clusters must sit inside
// one GPC, so 148 SMs do not always yield 148 / cluster_size clusters; this number decides
// whether operand multicast across more than one CTA pair can pay. No reference counterpart.
extern "C" int ddlb_max_active_clusters(int cluster_size, int threads, int* out) {
  using namespace ddlb;
  if (out == nullptr || cluster_size < 1 || cluster_size > 16) return set_error("bad arguments");
  using Cfg = TpGemmCfg<2, 256>;
  auto k = tp_gemm_kernel<2, 256, true>;
  DDLB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  DDLB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  int dev = 0, sms = 0;
  DDLB_CUDA(cudaGetDevice(&dev));
  DDLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  cudaLaunchConfig_t lc;
  cudaLaunchAttribute attr[1];
  fill_launch(lc, attr, static_cast<unsigned>(sms - sms % cluster_size), threads, Cfg::kSmemBytes,
              cluster_size, nullptr);
  DDLB_CUDA(cudaOccupancyMaxActiveClusters(out, reinterpret_cast<const void*>(k), &lc));
  return 0;
}

// ---------------------------------------------------------------------------
// Host-side introspection of the tile schedule (runs the kernel's own tile_coord on the
// CPU). Works without a GPU when cfg->num_sms is given. Used by the CPU property tests.
// ---------------------------------------------------------------------------
extern "C" int ddlb_debug_schedule(const ddlb_tp_gemm_config* cfg, int32_t out[12]) {
  using namespace ddlb;
  if (cfg == nullptr || out == nullptr) return set_error("null argument");
  Resolved r;
  if (int rc = resolve(*cfg, r)) return rc;
  TpGemmParams p;
  memset(&p, 0, sizeof(p));
  fill_schedule(p, *cfg, r);
  out[0] = p.num_m_blocks; out[1] = p.num_n_blocks; out[2] = p.mblks_per_part; out[3] = p.chunk_mblks;
  out[4] = p.group_m; out[5] = r.tile_m; out[6] = r.block_n; out[7] = r.cta_group;
  out[8] = r.chunk_rows; out[9] = r.part0; out[10] = r.grid; out[11] = r.threads;
  (void)p.sched_n_blocks;
  return 0;
}

extern "C" int ddlb_debug_tile_coords(const ddlb_tp_gemm_config* cfg, int32_t first, int32_t count,
                                      int32_t* out /* count x {m_blk, n_blk, part, chunk} */) {
  using namespace ddlb;
  if (cfg == nullptr || out == nullptr || first < 0 || count < 0) return set_error("bad arguments");
  Resolved r;
  if (int rc = resolve(*cfg, r)) return rc;
  TpGemmParams p;
  memset(&p, 0, sizeof(p));
  fill_schedule(p, *cfg, r);
  const int total = p.num_m_blocks * p.sched_n_blocks;
  if (first + count > total) return set_error("tiles %d..%d out of range (%d tiles)", first, first + count, total);
  for (int i = 0; i < count; ++i) {
    const TileCoord tc = tile_coord(p, first + i);
    out[4 * i] = tc.m_blk; out[4 * i + 1] = tc.n_blk; out[4 * i + 2] = tc.part; out[4 * i + 3] = tc.chunk;
  }
  return 0;
}

// All units the register-path comm warps of `cfg->rank` process in AG_BEFORE (pull transport),
// produced by the kernel's own pull_unit(): out = count x {src, chunk, begin, end, base}.
extern "C" int ddlb_debug_pull_units(const ddlb_tp_gemm_config* cfg, int64_t cap, int64_t* out,
                                     int64_t* count) {
  using namespace ddlb;
  if (cfg == nullptr || count == nullptr) return set_error("null argument");
  if (cfg->mode != DDLB_MODE_AG_BEFORE || cfg->world < 2) return set_error("AG_BEFORE with world >= 2 only");
  Resolved r;
  if (int rc = resolve(*cfg, r)) return rc;
  TpGemmParams p;
  memset(&p, 0, sizeof(p));
  fill_schedule(p, *cfg, r);
  p.chunk_bytes = static_cast<uint32_t>(static_cast<int64_t>(r.chunk_rows) * r.kk * r.elem);
  const int chunks_per_part = p.mblks_per_part / p.chunk_mblks;
  const int upc = static_cast<int>((p.chunk_bytes + kPushUnit - 1) / kPushUnit);
  const int64_t total = static_cast<int64_t>(chunks_per_part) * (p.world - 1) * upc;
  const size_t part_bytes = static_cast<size_t>(p.chunk_bytes) * chunks_per_part;
  *count = total;
  if (out == nullptr) return 0;
  if (cap < total) return set_error("buffer holds %lld units, %lld needed", (long long)cap, (long long)total);
  for (int64_t u = 0; u < total; ++u) {
    const CommUnit cu = pull_unit(p, static_cast<int>(u), upc, part_bytes);
    out[5 * u] = cu.src; out[5 * u + 1] = cu.chunk; out[5 * u + 2] = cu.begin; out[5 * u + 3] = cu.end;
    out[5 * u + 4] = static_cast<int64_t>(cu.base);
  }
  return 0;
}

// Symmetric heap (cudaIpc), device barrier, topology report. Included by ddlb_b200.cu.

namespace ddlb {

constexpr int64_t kSignalBytes = 1 << 20;   // zeroed signal pads at the start of every arena
constexpr int kBarrierSlot = 0;             // words [0, 8) of the pads: barrier arrivals per source

__global__ void barrier_kernel(uint32_t* sig_local, uint32_t* const* sig_peer_tbl_unused,
                               uint32_t* p0, uint32_t* p1, uint32_t* p2, uint32_t* p3, uint32_t* p4,
                               uint32_t* p5, uint32_t* p6, uint32_t* p7, int rank, int world,
                               uint32_t epoch) {
  uint32_t* peers[8] = {p0, p1, p2, p3, p4, p5, p6, p7};
  const int t = threadIdx.x;
  if (t < world && t != rank) {
    // everything this stream did before the barrier is visible before the arrival
    fence_acq_rel_sys();
    st_release_sys(peers[t] + kBarrierSlot + rank, epoch);
    wait_flag_ge<true>(sig_local + kBarrierSlot + t, epoch);
  }
}

}  // namespace ddlb

struct ddlb_symm {
  int device = 0, rank = 0, world = 1;
  int64_t bytes = 0;
  void* base[DDLB_MAX_WORLD] = {};
  bool opened[DDLB_MAX_WORLD] = {};
  uint32_t barrier_epoch = 0;
};

extern "C" int ddlb_symm_create(int device, int rank, int world, int64_t bytes, ddlb_symm** out) {
  using namespace ddlb;
  if (out == nullptr) return set_error("null argument");
  *out = nullptr;
  if (world < 1 || world > DDLB_MAX_WORLD || rank < 0 || rank >= world)
    return set_error("bad rank/world %d/%d", rank, world);
  if (bytes < kSignalBytes * 2) bytes = kSignalBytes * 2;
  bytes = (bytes + (2 << 20) - 1) / (2 << 20) * (2 << 20);
  DDLB_CUDA(cudaSetDevice(device));
  auto* s = new ddlb_symm();
  s->device = device; s->rank = rank; s->world = world; s->bytes = bytes;
  cudaError_t e = cudaMalloc(&s->base[rank], static_cast<size_t>(bytes));
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    delete s;
    return set_error("cudaMalloc(%lld) for the symmetric heap: %s", (long long)bytes, cudaGetErrorString(e));
  }
  e = cudaMemset(s->base[rank], 0, static_cast<size_t>(kSignalBytes));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(s->base[rank]);
    delete s;
    return set_error("clearing signal pads: %s", cudaGetErrorString(e));
  }
  *out = s;
  return 0;
}

extern "C" int ddlb_symm_export(ddlb_symm* s, void* handle_out) {
  using namespace ddlb;
  if (s == nullptr || handle_out == nullptr) return set_error("null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == DDLB_IPC_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  DDLB_CUDA(cudaSetDevice(s->device));
  DDLB_CUDA(cudaIpcGetMemHandle(&h, s->base[s->rank]));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int ddlb_symm_open_peers(ddlb_symm* s, const void* handles) {
  using namespace ddlb;
  if (s == nullptr || handles == nullptr) return set_error("null argument");
  DDLB_CUDA(cudaSetDevice(s->device));
  for (int p = 0; p < s->world; ++p) {
    if (p == s->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(handles) + static_cast<size_t>(p) * DDLB_IPC_HANDLE_BYTES, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle(&s->base[p], h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return set_error("cudaIpcOpenMemHandle(rank %d): %s", p, cudaGetErrorString(e));
    }
    s->opened[p] = true;
  }
  return 0;
}

extern "C" int ddlb_symm_adopt_peer(ddlb_symm* s, int peer_rank, ddlb_symm* peer) {
  using namespace ddlb;
  if (s == nullptr || peer == nullptr || peer_rank < 0 || peer_rank >= s->world || peer_rank == s->rank)
    return set_error("bad adopt_peer arguments");
  s->base[peer_rank] = peer->base[peer->rank];
  return 0;
}

extern "C" void* ddlb_symm_base(ddlb_symm* s, int rank) {
  if (s == nullptr || rank < 0 || rank >= s->world) return nullptr;
  return s->base[rank];
}

extern "C" int64_t ddlb_symm_bytes(ddlb_symm* s) { return s ? s->bytes : 0; }

extern "C" int ddlb_symm_barrier(ddlb_symm* s, void* stream) {
  using namespace ddlb;
  if (s == nullptr) return set_error("null heap");
  if (s->world == 1) return 0;
  for (int p = 0; p < s->world; ++p)
    if (s->base[p] == nullptr) return set_error("peer %d is not mapped", p);
  s->barrier_epoch += 1;
  uint32_t* sig[8] = {};
  for (int p = 0; p < s->world; ++p) sig[p] = static_cast<uint32_t*>(s->base[p]);
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      sig[s->rank], nullptr, sig[0], sig[1], sig[2], sig[3], sig[4], sig[5], sig[6], sig[7], s->rank,
      s->world, s->barrier_epoch);
  DDLB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddlb_symm_destroy(ddlb_symm* s) {
  if (s == nullptr) return 0;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  for (int p = 0; p < s->world; ++p)
    if (s->opened[p] && s->base[p] != nullptr) cudaIpcCloseMemHandle(s->base[p]);
  if (s->base[s->rank] != nullptr) cudaFree(s->base[s->rank]);
  (void)cudaGetLastError();
  delete s;
  return 0;
}

extern "C" int ddlb_read_error_word(uint32_t out[4]) {
  using namespace ddlb;
  if (out == nullptr) return set_error("null argument");
  DDLB_CUDA(cudaMemcpyFromSymbol(out, g_ddlb_error_word, 4 * sizeof(uint32_t)));
  return 0;
}

// ------------------------------------------------------------------ topology
extern "C" int ddlb_topology_json(char* out, int64_t cap, int64_t* needed) {
  using namespace ddlb;
  std::string js = "{";
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess) { (void)cudaGetLastError(); ndev = 0; }
  int drv = 0, rt = 0;
  cudaDriverGetVersion(&drv);
  cudaRuntimeGetVersion(&rt);
  (void)cudaGetLastError();
  js += "\"driver_version\": " + std::to_string(drv) + ", \"runtime_version\": " + std::to_string(rt);
  js += ", \"device_count\": " + std::to_string(ndev) + ", \"devices\": [";
  for (int d = 0; d < ndev; ++d) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, d) != cudaSuccess) { (void)cudaGetLastError(); continue; }
    int mc = 0;
    {
      // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED through the driver entry point (no -lcuda link)
      using AttrFn = CUresult (*)(int*, CUdevice_attribute, CUdevice);
      void* sym = nullptr;
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &sym, cudaEnableDefault, &q) == cudaSuccess &&
          q == cudaDriverEntryPointSuccess && sym != nullptr)
        reinterpret_cast<AttrFn>(sym)(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d);
      (void)cudaGetLastError();
    }
    if (d) js += ", ";
    js += "{\"index\": " + std::to_string(d) + ", \"name\": \"" + prop.name + "\"";
    js += ", \"cc\": \"" + std::to_string(prop.major) + "." + std::to_string(prop.minor) + "\"";
    js += ", \"sms\": " + std::to_string(prop.multiProcessorCount);
    js += ", \"mem_bytes\": " + std::to_string(static_cast<long long>(prop.totalGlobalMem));
    js += ", \"l2_bytes\": " + std::to_string(prop.l2CacheSize);
    js += ", \"smem_per_block_optin\": " + std::to_string(static_cast<long long>(prop.sharedMemPerBlockOptin));
    js += ", \"multicast_supported\": " + std::to_string(mc);
    js += ", \"peers\": [";
    bool first = true;
    for (int q = 0; q < ndev; ++q) {
      if (q == d) continue;
      int can = 0, perf = -1, atom = 0;
      cudaDeviceCanAccessPeer(&can, d, q);
      cudaDeviceGetP2PAttribute(&perf, cudaDevP2PAttrPerformanceRank, d, q);
      cudaDeviceGetP2PAttribute(&atom, cudaDevP2PAttrNativeAtomicSupported, d, q);
      (void)cudaGetLastError();
      if (!first) js += ", ";
      first = false;
      js += "{\"peer\": " + std::to_string(q) + ", \"access\": " + std::to_string(can) +
            ", \"perf_rank\": " + std::to_string(perf) + ", \"native_atomics\": " + std::to_string(atom) + "}";
    }
    js += "]}";
  }
  js += "]}";
  if (needed != nullptr) *needed = static_cast<int64_t>(js.size()) + 1;
  if (out != nullptr && cap > 0) {
    const size_t n = js.size() < static_cast<size_t>(cap - 1) ? js.size() : static_cast<size_t>(cap - 1);
    memcpy(out, js.data(), n);
    out[n] = '\0';
  }
  return 0;
}

// ------------------------------------------------------------------ copy diagnostic
// Plain 128-bit load/store copy over all SMs; used by `python -m ddlb_b200.parallel.p2p_bench`
// to measure what the NVLink path to a peer delivers (pull: src is peer memory, push: dst is).
namespace ddlb {
__global__ void __launch_bounds__(512) copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                   size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 7 * stride < n; i += 8 * stride) {
    uint4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 8; ++u) __stcs(dst + i + u * stride, v[u]);
  }
  for (; i < n; i += stride) dst[i] = src[i];
}
}  // namespace ddlb

extern "C" int ddlb_copy(void* dst, const void* src, int64_t bytes, int blocks, void* stream) {
  using namespace ddlb;
  if (bytes % 16 != 0) return set_error("bytes must be a multiple of 16");
  if (blocks <= 0) blocks = 148 * 4;
  copy_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4*>(dst), static_cast<const uint4*>(src), static_cast<size_t>(bytes / 16));
  DDLB_CUDA(cudaGetLastError());
  return 0;
}

// Access-pattern diagnostic for the in-kernel collectives: the same bytes copied with the
// patterns the fused kernels can use, on a chosen grid x block.
//   pattern 0: dense grid-stride, 16 B per thread, 8 loads in flight (the copy_kernel above)
//   pattern 1: every warp streams its own `unit_bytes` units (round-robin), two 4 KiB register
//              groups in flight (copy_pipelined) - the round-2 comm warps
//   pattern 2: warps interleave at 4 KiB: warp w moves pieces w, w + nw, ... so the whole grid
//              sweeps the buffer densely; two pieces in flight per warp
//   pattern 3: like 2 with 16 KiB in flight per warp (four pieces)
namespace ddlb {
// flags: bit 0 loads bypass L1 (ld.global.cg instead of .cs); bit 1 pattern 1 fences and bumps a
// counter after every unit like the comm warps do; bit 2 (with bit 1) the first group of the
// NEXT unit is already in flight when the fence is issued.
template <bool kCg>
__device__ __forceinline__ uint4 ld_peer(const uint8_t* p) {
  return kCg ? __ldcg(reinterpret_cast<const uint4*>(p)) : __ldcs(reinterpret_cast<const uint4*>(p));
}

template <bool kCg>
__device__ __forceinline__ void copy_units(uint8_t* dst, const uint8_t* src, size_t bytes, uint32_t unit_bytes,
                                           int flags, uint32_t* counter, size_t gw, size_t nw, int lane) {
  const size_t units = bytes / unit_bytes;
  const uint32_t l16 = static_cast<uint32_t>(lane) * 16u;
  const bool fence = (flags & 2) != 0, prefetch = (flags & 4) != 0;
  uint4 va[kCopyGroup], vb[kCopyGroup];
  bool have_a = false;
  for (size_t u = gw; u < units; u += nw) {
    const uint8_t* s0 = src + u * unit_bytes;
    uint8_t* d0 = dst + u * unit_bytes;
    if (!have_a) {
#pragma unroll
      for (int i = 0; i < kCopyGroup; ++i) va[i] = ld_peer<kCg>(s0 + l16 + i * 512);
    }
    have_a = false;
    for (uint32_t off = 0; off < unit_bytes; off += 2 * kCopyStep) {
#pragma unroll
      for (int i = 0; i < kCopyGroup; ++i) vb[i] = ld_peer<kCg>(s0 + off + kCopyStep + l16 + i * 512);
#pragma unroll
      for (int i = 0; i < kCopyGroup; ++i) __stcg(reinterpret_cast<uint4*>(d0 + off + l16 + i * 512), va[i]);
      if (off + 2 * kCopyStep < unit_bytes) {
#pragma unroll
        for (int i = 0; i < kCopyGroup; ++i) va[i] = ld_peer<kCg>(s0 + off + 2 * kCopyStep + l16 + i * 512);
      } else if (prefetch && u + nw < units) {
        const uint8_t* s1 = src + (u + nw) * unit_bytes;
#pragma unroll
        for (int i = 0; i < kCopyGroup; ++i) va[i] = ld_peer<kCg>(s1 + l16 + i * 512);
        have_a = true;
      }
#pragma unroll
      for (int i = 0; i < kCopyGroup; ++i)
        __stcg(reinterpret_cast<uint4*>(d0 + off + kCopyStep + l16 + i * 512), vb[i]);
    }
    if (fence) {
      __threadfence();
      __syncwarp();
      if (lane == 0) red_add_release_gpu(counter + (u & 63), 1u);
    }
  }
}

__global__ void __launch_bounds__(512) copy_pattern_kernel(uint8_t* __restrict__ dst,
                                                           const uint8_t* __restrict__ src, size_t bytes,
                                                           int pattern, uint32_t unit_bytes, int flags,
                                                           uint32_t* counter) {
  extern __shared__ uint8_t carve_out[];      // only to shrink the L1 like the GEMM's operand ring does
  if (threadIdx.x == 0 && bytes == 1) carve_out[0] = 0;
  const int lane = threadIdx.x & 31;
  const size_t gw = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const size_t nw = (static_cast<size_t>(gridDim.x) * blockDim.x) >> 5;
  if (pattern == 1) {
    if (flags & 1) copy_units<true>(dst, src, bytes, unit_bytes, flags, counter, gw, nw, lane);
    else copy_units<false>(dst, src, bytes, unit_bytes, flags, counter, gw, nw, lane);
    return;
  }
  const size_t pieces = bytes / kCopyStep;
  const int depth = pattern == 3 ? 4 : 2;
  // `depth` pieces in flight: load all, then store all
  for (size_t p0 = gw; p0 < pieces; p0 += nw * depth) {
    uint4 v[4][kCopyGroup];
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      const size_t pc = p0 + static_cast<size_t>(d) * nw;
      if (d < depth && pc < pieces) {
#pragma unroll
        for (int i = 0; i < kCopyGroup; ++i)
          v[d][i] = __ldcs(reinterpret_cast<const uint4*>(src + pc * kCopyStep + lane * 16 + i * 512));
      }
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      const size_t pc = p0 + static_cast<size_t>(d) * nw;
      if (d < depth && pc < pieces) {
#pragma unroll
        for (int i = 0; i < kCopyGroup; ++i)
          __stcg(reinterpret_cast<uint4*>(dst + pc * kCopyStep + lane * 16 + i * 512), v[d][i]);
      }
    }
  }
}
__device__ uint32_t g_copy_pattern_counter[64];
}  // namespace ddlb

extern "C" int ddlb_copy_pattern(void* dst, const void* src, int64_t bytes, int blocks, int threads,
                                 int pattern, int unit_bytes, int smem_bytes, int flags, void* stream) {
  using namespace ddlb;
  if (bytes % (32 * 1024) != 0 || unit_bytes <= 0 || unit_bytes % 8192 != 0 || bytes % unit_bytes != 0)
    return set_error("bytes must be a multiple of 32 KiB and of unit_bytes (a multiple of 8192)");
  if (threads <= 0 || threads > 512 || threads % 32 != 0 || blocks <= 0) return set_error("bad launch shape");
  if (pattern == 0) {
    copy_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<uint4*>(dst), static_cast<const uint4*>(src), static_cast<size_t>(bytes / 16));
  } else {
    uint32_t* counter = nullptr;
    DDLB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&counter), g_copy_pattern_counter));
    DDLB_CUDA(cudaFuncSetAttribute(copy_pattern_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    copy_pattern_kernel<<<blocks, threads, static_cast<size_t>(smem_bytes), static_cast<cudaStream_t>(stream)>>>(
        static_cast<uint8_t*>(dst), static_cast<const uint8_t*>(src), static_cast<size_t>(bytes), pattern,
        static_cast<uint32_t>(unit_bytes), flags, counter);
  }
  DDLB_CUDA(cudaGetLastError());
  return 0;
}

// Barrier on caller-provided signal pads (one pad base per rank, as seen from this process):
// the same kernel as ddlb_symm_barrier for heaps that are not ddlb_symm objects (VMM heap).
extern "C" int ddlb_barrier_raw(void* const* pad_bases, int rank, int world, uint32_t epoch, void* stream) {
  using namespace ddlb;
  if (pad_bases == nullptr || world < 1 || world > DDLB_MAX_WORLD || rank < 0 || rank >= world)
    return set_error("bad arguments");
  if (world == 1) return 0;
  uint32_t* sig[8] = {};
  for (int p = 0; p < world; ++p) {
    if (pad_bases[p] == nullptr) return set_error("pad of rank %d is null", p);
    sig[p] = static_cast<uint32_t*>(pad_bases[p]);
  }
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      sig[rank], nullptr, sig[0], sig[1], sig[2], sig[3], sig[4], sig[5], sig[6], sig[7], rank, world, epoch);
  DDLB_CUDA(cudaGetLastError());
  return 0;
}

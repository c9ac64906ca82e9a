// kvx_vmm_host.inl — pools whose memory comes from CUDA virtual memory management (cuMemCreate / cuMemMap: what
// engines running with expandable segments hold their KV cache in), shared between processes as POSIX file
// descriptors.  Included at the end of kvx.cu.  The driver API is reached through cudaGetDriverEntryPoint, so the
// library still links against the runtime only.
#include <unistd.h>

namespace {

constexpr uint32_t kPoolMagicFd = 0x4658564bu;  // "KVXF"

struct VmmApi {
  CUresult (*create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*release)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*addr_free)(CUdeviceptr, size_t) = nullptr;
  CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*unmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*set_access)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*export_fd)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*import_fd)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*retain)(CUmemGenericAllocationHandle*, void*) = nullptr;
  CUresult (*granularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*range)(CUdeviceptr*, size_t*, CUdeviceptr) = nullptr;
  bool ok = false;
};

const VmmApi* vmm_api() {
  static VmmApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    auto get = [](const char* name, void** out) {
      cudaDriverEntryPointQueryResult q;
      return cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *out;
    };
    bool ok = true;
    ok = get("cuMemCreate", reinterpret_cast<void**>(&api.create)) && ok;
    ok = get("cuMemRelease", reinterpret_cast<void**>(&api.release)) && ok;
    ok = get("cuMemAddressReserve", reinterpret_cast<void**>(&api.reserve)) && ok;
    ok = get("cuMemAddressFree", reinterpret_cast<void**>(&api.addr_free)) && ok;
    ok = get("cuMemMap", reinterpret_cast<void**>(&api.map)) && ok;
    ok = get("cuMemUnmap", reinterpret_cast<void**>(&api.unmap)) && ok;
    ok = get("cuMemSetAccess", reinterpret_cast<void**>(&api.set_access)) && ok;
    ok = get("cuMemExportToShareableHandle", reinterpret_cast<void**>(&api.export_fd)) && ok;
    ok = get("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&api.import_fd)) && ok;
    ok = get("cuMemRetainAllocationHandle", reinterpret_cast<void**>(&api.retain)) && ok;
    ok = get("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&api.granularity)) && ok;
    ok = get("cuMemGetAddressRange", reinterpret_cast<void**>(&api.range)) && ok;
    (void)cudaGetLastError();
    api.ok = ok;
  });
  return api.ok ? &api : nullptr;
}

int32_t vmm_fail(kvx_ctx* ctx, CUresult r, const char* what) {
  ctx->last_error = std::string(what) + " failed with CUresult " + std::to_string(static_cast<int>(r));
  return KVX_EIPC;
}

struct FdAlloc {  // one physical allocation under an exported pool, as the exporter maps it
  uint64_t va, size;
};

struct FdPoolHeader {  // wire format: header, n_allocs x FdAlloc, regions x uint64 (region base in the exporter's address space)
  uint32_t magic, regions;
  uint32_t block_tokens, heads, head_dim, elem_bytes, layout, tp_rank, tp_size, head_stride;
  uint64_t num_blocks, block_stride;
  uint32_t n_allocs, reserved;
  unsigned char device_uuid[16];
};

uint64_t pool_span(const kvx_pool* pool) {
  const kvx_pool_desc& d = pool->desc;
  const uint64_t row_bytes = static_cast<uint64_t>(d.head_dim) * d.elem_bytes;
  const uint64_t chunk_extent = d.head_stride ? static_cast<uint64_t>(d.heads - 1) * d.head_stride + row_bytes * d.block_tokens : pool->chunk_bytes;
  return (d.num_blocks - 1) * d.block_stride + chunk_extent;
}

// The physical allocations that lie under every region of the pool, sorted by address, each once.
int32_t pool_allocations(kvx_ctx* ctx, const kvx_pool* pool, std::vector<FdAlloc>* out) {
  const VmmApi* api = vmm_api();
  if (!api) return KVX_EUNSUPPORTED;
  std::map<uint64_t, uint64_t> seen;
  const uint64_t span = pool_span(pool);
  for (void* b : pool->base) {
    uint64_t at = reinterpret_cast<uint64_t>(b);
    const uint64_t end = at + span;
    while (at < end) {
      CUdeviceptr ab = 0;
      size_t asz = 0;
      CUresult r = api->range(&ab, &asz, static_cast<CUdeviceptr>(at));
      if (r != CUDA_SUCCESS || asz == 0) return vmm_fail(ctx, r, "cuMemGetAddressRange");
      seen[static_cast<uint64_t>(ab)] = asz;
      at = static_cast<uint64_t>(ab) + asz;
    }
  }
  out->clear();
  for (const auto& kv : seen) out->push_back(FdAlloc{kv.first, kv.second});
  return KVX_OK;
}

}  // namespace

extern "C" {

int32_t kvx_mem_alloc_vmm(kvx_ctx* ctx, uint64_t bytes, uint64_t piece_bytes, void** out) try {
  if (!ctx || !out || bytes == 0) return KVX_EINVAL;
  const VmmApi* api = vmm_api();
  if (!api) return KVX_EUNSUPPORTED;
  DeviceGuard guard(ctx->device);
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ctx->device;
  size_t gran = 0;
  CUresult r = api->granularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
  if (r != CUDA_SUCCESS || gran == 0) return vmm_fail(ctx, r, "cuMemGetAllocationGranularity");
  const uint64_t piece = piece_bytes ? (piece_bytes + gran - 1) / gran * gran : (bytes + gran - 1) / gran * gran;
  const uint64_t total = (bytes + piece - 1) / piece * piece;
  CUdeviceptr va = 0;
  r = api->reserve(&va, total, gran, 0, 0);
  if (r != CUDA_SUCCESS) return r == CUDA_ERROR_OUT_OF_MEMORY ? KVX_ENOMEM : vmm_fail(ctx, r, "cuMemAddressReserve");
  for (uint64_t off = 0; off < total; off += piece) {
    CUmemGenericAllocationHandle h = 0;
    r = api->create(&h, piece, &prop, 0);
    if (r == CUDA_SUCCESS) {
      r = api->map(va + off, piece, 0, h, 0);
      api->release(h);  // the mapping keeps the memory alive
    }
    if (r != CUDA_SUCCESS) {
      if (off) api->unmap(va, off);
      api->addr_free(va, total);
      return r == CUDA_ERROR_OUT_OF_MEMORY ? KVX_ENOMEM : vmm_fail(ctx, r, "cuMemCreate / cuMemMap");
    }
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ctx->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = api->set_access(va, total, &acc, 1);
  if (r != CUDA_SUCCESS) {
    api->unmap(va, total);
    api->addr_free(va, total);
    return vmm_fail(ctx, r, "cuMemSetAccess");
  }
  ctx->vmm_allocs[reinterpret_cast<void*>(va)] = total;
  *out = reinterpret_cast<void*>(va);
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_mem_free_vmm(kvx_ctx* ctx, void* ptr) try {
  if (!ctx || !ptr) return KVX_EINVAL;
  const VmmApi* api = vmm_api();
  auto it = ctx->vmm_allocs.find(ptr);
  if (!api || it == ctx->vmm_allocs.end()) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  cudaDeviceSynchronize();
  api->unmap(reinterpret_cast<CUdeviceptr>(ptr), it->second);
  api->addr_free(reinterpret_cast<CUdeviceptr>(ptr), it->second);
  ctx->vmm_allocs.erase(it);
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_export_fds(kvx_pool* pool, void* buf, size_t* len, int32_t* fds, uint32_t* count) try {
  if (!pool || !len || !count) return KVX_EINVAL;
  kvx_ctx* ctx = pool->ctx;
  const VmmApi* api = vmm_api();
  if (!api) return KVX_EUNSUPPORTED;
  DeviceGuard guard(ctx->device);
  std::vector<FdAlloc> allocs;
  int32_t rc = pool_allocations(ctx, pool, &allocs);
  if (rc) return rc;
  const size_t need = sizeof(FdPoolHeader) + allocs.size() * sizeof(FdAlloc) + pool->base.size() * sizeof(uint64_t);
  if (!buf || *len < need || !fds || *count < allocs.size()) {
    *len = need;
    *count = static_cast<uint32_t>(allocs.size());
    return KVX_ERANGE;
  }
  FdPoolHeader h{};
  const kvx_pool_desc& d = pool->desc;
  h.magic = kPoolMagicFd; h.regions = d.regions; h.block_tokens = d.block_tokens; h.heads = d.heads; h.head_dim = d.head_dim;
  h.elem_bytes = d.elem_bytes; h.layout = d.layout; h.tp_rank = d.tp_rank; h.tp_size = d.tp_size; h.head_stride = d.head_stride;
  h.num_blocks = d.num_blocks; h.block_stride = d.block_stride; h.n_allocs = static_cast<uint32_t>(allocs.size());
  memcpy(h.device_uuid, ctx->uuid, 16);
  for (size_t i = 0; i < allocs.size(); ++i) {
    CUmemGenericAllocationHandle mh = 0;
    CUresult r = api->retain(&mh, reinterpret_cast<void*>(allocs[i].va));
    int fd = -1;
    if (r == CUDA_SUCCESS) {
      r = api->export_fd(&fd, mh, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      api->release(mh);
    }
    if (r != CUDA_SUCCESS) {
      for (size_t j = 0; j < i; ++j) close(fds[j]);
      ctx->last_error = "the pool's memory cannot be exported as a file descriptor either (CUresult " + std::to_string(static_cast<int>(r)) +
                        "): it was not created with CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR";
      return KVX_EIPC;
    }
    fds[i] = fd;
  }
  char* w = static_cast<char*>(buf);
  memcpy(w, &h, sizeof(h));
  memcpy(w + sizeof(h), allocs.data(), allocs.size() * sizeof(FdAlloc));
  uint64_t* rb = reinterpret_cast<uint64_t*>(w + sizeof(h) + allocs.size() * sizeof(FdAlloc));
  for (size_t r = 0; r < pool->base.size(); ++r) rb[r] = reinterpret_cast<uint64_t>(pool->base[r]);
  *len = need;
  *count = static_cast<uint32_t>(allocs.size());
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_import_fds(kvx_ctx* ctx, const void* buf, size_t len, const int32_t* fds, uint32_t count, kvx_pool** out) try {
  if (!ctx || !buf || !out || !fds || len < sizeof(FdPoolHeader)) return KVX_EINVAL;
  const VmmApi* api = vmm_api();
  if (!api) return KVX_EUNSUPPORTED;
  FdPoolHeader h;
  memcpy(&h, buf, sizeof(h));
  if (h.magic != kPoolMagicFd || h.n_allocs != count || h.n_allocs == 0) return KVX_EINVAL;
  if (len < sizeof(h) + static_cast<size_t>(h.n_allocs) * sizeof(FdAlloc) + static_cast<size_t>(h.regions) * sizeof(uint64_t)) return KVX_EINVAL;
  std::vector<FdAlloc> allocs(h.n_allocs);
  memcpy(allocs.data(), static_cast<const char*>(buf) + sizeof(h), allocs.size() * sizeof(FdAlloc));
  std::vector<uint64_t> region_va(h.regions);
  memcpy(region_va.data(), static_cast<const char*>(buf) + sizeof(h) + allocs.size() * sizeof(FdAlloc), region_va.size() * sizeof(uint64_t));
  DeviceGuard guard(ctx->device);
  // Allocations that were contiguous in the exporter's address space stay contiguous here: one reservation per run.
  struct Run { uint64_t old_va, size, new_va; };
  std::vector<Run> runs;
  std::vector<kvx_pool::VmmRange> mapped;
  auto undo = [&]() {
    for (const auto& m : mapped) {
      api->unmap(m.va, m.size);
      api->addr_free(m.va, m.size);
    }
  };
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ctx->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  for (size_t i = 0; i < allocs.size();) {
    size_t j = i;
    uint64_t size = 0;
    while (j < allocs.size() && allocs[j].va == allocs[i].va + size) size += allocs[j++].size;
    CUdeviceptr va = 0;
    CUresult r = api->reserve(&va, size, 0, 0, 0);
    if (r != CUDA_SUCCESS) { undo(); return vmm_fail(ctx, r, "cuMemAddressReserve"); }
    uint64_t off = 0;
    for (size_t k = i; k < j && r == CUDA_SUCCESS; ++k) {
      CUmemGenericAllocationHandle mh = 0;
      r = api->import_fd(&mh, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[k])), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r == CUDA_SUCCESS) {
        r = api->map(va + off, allocs[k].size, 0, mh, 0);
        api->release(mh);
        if (r == CUDA_SUCCESS) off += allocs[k].size;
      }
    }
    if (r == CUDA_SUCCESS) r = api->set_access(va, size, &acc, 1);
    if (r != CUDA_SUCCESS) {
      if (off) api->unmap(va, off);
      api->addr_free(va, size);
      undo();
      return vmm_fail(ctx, r, "cuMemImportFromShareableHandle / cuMemMap / cuMemSetAccess");
    }
    mapped.push_back(kvx_pool::VmmRange{va, size});
    runs.push_back(Run{allocs[i].va, size, static_cast<uint64_t>(va)});
    i = j;
  }
  std::vector<void*> bases(h.regions, nullptr);
  for (uint32_t r = 0; r < h.regions; ++r) {
    for (const Run& run : runs)
      if (region_va[r] >= run.old_va && region_va[r] < run.old_va + run.size) bases[r] = reinterpret_cast<void*>(run.new_va + (region_va[r] - run.old_va));
    if (!bases[r]) { undo(); return KVX_EINVAL; }
  }
  kvx_pool_desc d{};
  d.struct_size = sizeof(d); d.regions = h.regions; d.block_tokens = h.block_tokens; d.heads = h.heads; d.head_dim = h.head_dim;
  d.elem_bytes = h.elem_bytes; d.layout = h.layout; d.tp_rank = h.tp_rank; d.tp_size = h.tp_size; d.head_stride = h.head_stride;
  d.num_blocks = h.num_blocks; d.block_stride = h.block_stride; d.base = bases.data();
  kvx_pool* pool = nullptr;
  int32_t rc = kvx_pool_register(ctx, &d, &pool);
  if (rc) { undo(); return rc; }
  pool->vmm_ranges = mapped;
  pool->local = memcmp(h.device_uuid, ctx->uuid, 16) == 0;  // where the bytes live, from the exporter's record
  *out = pool;
  return KVX_OK;
} KVX_CATCH_ALL

}  // extern "C"

namespace {
void pool_unmap_vmm(kvx_pool* pool) {
  const VmmApi* api = vmm_api();
  if (!api) return;
  for (const auto& m : pool->vmm_ranges) {
    api->unmap(m.va, m.size);
    api->addr_free(m.va, m.size);
  }
  pool->vmm_ranges.clear();
}
}  // namespace

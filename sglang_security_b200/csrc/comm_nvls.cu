// NVLS: all-reduce whose additions happen INSIDE the NVSwitch (multimem.ld_reduce over a multicast
// mapping) — SURVEY §8f row 3.  Reference analogue: the one-shot kernel of the custom all-reduce
// (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-225, policy custom_all_reduce.py:414-437),
// which pulls every peer's buffer over NVLink and adds on the SM; NCCL's NVLS algorithms are the
// library form of what is done here.
//
// Memory: every rank creates one physical allocation (cuMemCreate) and binds it to a shared
// MULTICAST object (cuMulticastCreate on rank 0, exported as a POSIX file descriptor and imported
// by the peers; cuMulticastAddDevice / cuMulticastBindMem on every rank).  The allocation is mapped
// twice: `uc` = ordinary local view, `mc` = multicast view.  A multimem.ld_reduce on mc + x returns
// the SUM over all ranks of their uc + x, computed by the switch: one NVLink round trip, the bytes
// cross each link once, no SM adds.
//
// Protocol of call e (flags and epochs are the peer-memory communicator's, comm.cuh):
//   copy    CTA c copies its chunk of the input into uc[e & 1] (local stores);
//   signal  bar.sync, then one thread per peer: st.release.sys flags[e&1][my rank][c] = e;
//   wait    one thread per peer spins until its local flags[e&1][peer][c] == e;
//   reduce  multimem.ld_reduce (fp32 accumulation for 16-bit types) over the same chunk -> output.
// CTA c only ever reads addresses that CTA c of every rank wrote, so the per-CTA flags suffice.
#include <cuda.h>

#include <cstdlib>

#include "comm.cuh"

namespace b200 {

// ---- driver entry points (no link-time dependency on libcuda: the library must load on CPU boxes) --
template <typename Fn>
static Fn drv(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess) {
    (void)cudaGetLastError();
    return nullptr;
  }
  return reinterpret_cast<Fn>(p);
}
#define B200_DRV(var, name, ...)                                   \
  typedef CUresult (*var##_t)(__VA_ARGS__);                        \
  static var##_t var = drv<var##_t>(name);                         \
  B200_REQUIRE(var != nullptr, "NVLS: driver entry point %s is not available", name)
#define B200_CU_OK(expr)                                                                  \
  do {                                                                                    \
    CUresult r__ = (expr);                                                                \
    if (r__ != CUDA_SUCCESS) {                                                            \
      ::b200::set_error("NVLS: %s failed with CUresult %d", #expr, static_cast<int>(r__)); \
      return 4;                                                                           \
    }                                                                                     \
  } while (0)

struct NvlsBuf {
  CUmemGenericAllocationHandle mc = 0;   // multicast object
  CUmemGenericAllocationHandle mem = 0;  // this rank's physical allocation
  CUdeviceptr uc_ptr = 0, mc_ptr = 0;
  size_t size = 0;
  int world = 0, dev = 0;
  bool bound = false;
};

static int nvls_props(int world, size_t size, CUmulticastObjectProp* prop) {
  *prop = {};
  prop->numDevices = static_cast<unsigned>(world);
  prop->size = size;
  prop->handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop->flags = 0;
  return 0;
}

}  // namespace b200

namespace b200 {

// ---- kernel -----------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ uint4 multimem_ld_reduce(const void* mc_addr);
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce<__nv_bfloat16>(const void* a) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(a) : "memory");
  return r;
}
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce<__half>(const void* a) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(a) : "memory");
  return r;
}
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce<float>(const void* a) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(a) : "memory");
  return r;
}

template <typename T>
__global__ void __launch_bounds__(256)
all_reduce_nvls_kernel(const CommDev c, uint8_t* __restrict__ uc, const uint8_t* __restrict__ mc,
                       uint64_t slot_bytes, const uint4* __restrict__ in, uint4* __restrict__ out,
                       int64_t num_vec, int vec_per_cta) {
  griddep_wait();
  griddep_launch_dependents();
  const uint32_t e = comm_begin(c);
  const int cta = blockIdx.x;
  const int64_t v0 = static_cast<int64_t>(cta) * vec_per_cta;
  const int64_t v1 = min(num_vec, v0 + vec_per_cta);
  uint4* mine = reinterpret_cast<uint4*>(uc + (e & 1u) * slot_bytes);
  const uint4* sum = reinterpret_cast<const uint4*>(mc + (e & 1u) * slot_bytes);
  for (int64_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) mine[v] = in[v];
  __syncthreads();
  if (threadIdx.x < c.world) {
    comm_signal(c, e, cta, threadIdx.x);
    comm_wait(c, e, cta, threadIdx.x);
  }
  __syncthreads();
  for (int64_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) out[v] = multimem_ld_reduce<T>(sum + v);
  __syncthreads();
  if (threadIdx.x == 0) comm_end(c, e, gridDim.x);
}

static int64_t nvls_min_bytes() {
  static int64_t v = [] {
    const char* e = std::getenv("B200_COMM_NVLS_MIN");
    // 128 KiB: below that the LL protocol's single NVLink trip wins (TP 8: 9.4 vs 10.7 us at 64 KiB,
    // 14.9 vs 12.9 us at 256 KiB; profiles/r02_bench_allreduce_tp8.json)
    return (e && *e) ? std::atoll(e) : static_cast<int64_t>(128 << 10);
  }();
  return v;
}

int comm_all_reduce_nvls(const Comm* c, const void* in, void* out, int64_t bytes, int dtype,
                         cudaStream_t st) {
  if (!c->nvls_mc || bytes < nvls_min_bytes() || static_cast<uint64_t>(bytes) > c->nvls_slot) return -1;
  const int64_t num_vec = bytes / 16;
  // 4 KiB per CTA (one vector per thread) up to 2 CTAs per SM, then grow the per-CTA share
  int64_t vec_per_cta = 256;
  const int64_t sm2 = 2 * static_cast<int64_t>(device_sm_count());
  const int64_t max_ctas = sm2 < kCommMaxBlocks ? sm2 : kCommMaxBlocks;
  if ((num_vec + vec_per_cta - 1) / vec_per_cta > max_ctas)
    vec_per_cta = (num_vec + max_ctas - 1) / max_ctas;
  const int ctas = static_cast<int>((num_vec + vec_per_cta - 1) / vec_per_cta);
  const uint4* i4 = static_cast<const uint4*>(in);
  uint4* o4 = static_cast<uint4*>(out);
  {
    ScopedKernelTimer timer("all_reduce_nvls", st);
    cudaError_t err;
    if (dtype == B200_BF16)
      err = launch_kernel(all_reduce_nvls_kernel<__nv_bfloat16>, dim3(ctas), dim3(256), 0, st, 1,
                          kPdlOther, c->dev, c->nvls_uc, c->nvls_mc, static_cast<uint64_t>(c->nvls_slot),
                          i4, o4, num_vec, static_cast<int>(vec_per_cta));
    else if (dtype == B200_F16)
      err = launch_kernel(all_reduce_nvls_kernel<__half>, dim3(ctas), dim3(256), 0, st, 1, kPdlOther,
                          c->dev, c->nvls_uc, c->nvls_mc, static_cast<uint64_t>(c->nvls_slot), i4, o4,
                          num_vec, static_cast<int>(vec_per_cta));
    else
      err = launch_kernel(all_reduce_nvls_kernel<float>, dim3(ctas), dim3(256), 0, st, 1, kPdlOther,
                          c->dev, c->nvls_uc, c->nvls_mc, static_cast<uint64_t>(c->nvls_slot), i4, o4,
                          num_vec, static_cast<int>(vec_per_cta));
    B200_CUDA_OK(err);
  }
  B200_CHECK_LAUNCH("all_reduce_nvls_kernel");
  return 0;
}

}  // namespace b200

extern "C" {

/* 1 when the current device and driver support multicast objects (NVSwitch + NVLS), else 0. */
int b200_nvls_supported(void) {
  using namespace b200;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  typedef CUresult (*attr_t)(int*, CUdevice_attribute, CUdevice);
  static attr_t get_attr = drv<attr_t>("cuDeviceGetAttribute");
  typedef CUresult (*gran_t)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  static gran_t get_gran = drv<gran_t>("cuMulticastGetGranularity");
  if (!get_attr || !get_gran) return 0;
  int ok = 0;
  if (get_attr(&ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return ok ? 1 : 0;
}

/* Size a multicast-bound buffer must have to hold `bytes` (multiple of the multicast and allocation
 * granularities); 0 on error. */
size_t b200_nvls_buffer_bytes(int world, size_t bytes) {
  using namespace b200;
  typedef CUresult (*gran_t)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  static gran_t get_gran = drv<gran_t>("cuMulticastGetGranularity");
  typedef CUresult (*agran_t)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  static agran_t get_agran = drv<agran_t>("cuMemGetAllocationGranularity");
  if (!get_gran || !get_agran || world < 2 || bytes == 0) return 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  CUmulticastObjectProp mp;
  nvls_props(world, bytes, &mp);
  size_t g1 = 0, g2 = 0;
  if (get_gran(&g1, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g1 == 0) return 0;
  CUmemAllocationProp ap = {};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = dev;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (get_agran(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g2 == 0) return 0;
  const size_t g = g1 > g2 ? g1 : g2;
  return (bytes + g - 1) / g * g;
}

/* Rank 0: create the multicast object for `world` devices and export it as a file descriptor the
 * peers import (the descriptor is passed between processes by the host side). */
int b200_nvls_create(int world, size_t size, void** buf_out, int* fd_out) {
  using namespace b200;
  B200_REQUIRE(buf_out && fd_out && world >= 2 && size > 0, "b200_nvls_create: bad arguments");
  B200_DRV(mc_create, "cuMulticastCreate", CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  B200_DRV(mem_export, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle,
           CUmemAllocationHandleType, unsigned long long);
  CUmulticastObjectProp mp;
  nvls_props(world, size, &mp);
  NvlsBuf* b = new NvlsBuf();
  b->size = size;
  b->world = world;
  B200_CUDA_OK(cudaGetDevice(&b->dev));
  B200_CUDA_OK(cudaFree(nullptr));  // make sure the primary context exists
  B200_CU_OK(mc_create(&b->mc, &mp));
  int fd = -1;
  B200_CU_OK(mem_export(&fd, b->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  *buf_out = b;
  return 0;
}

/* Other ranks: import the multicast object from a descriptor valid in THIS process. */
int b200_nvls_import(int fd, int world, size_t size, void** buf_out) {
  using namespace b200;
  B200_REQUIRE(buf_out && fd >= 0 && world >= 2 && size > 0, "b200_nvls_import: bad arguments");
  B200_DRV(mem_import, "cuMemImportFromShareableHandle", CUmemGenericAllocationHandle*, void*,
           CUmemAllocationHandleType);
  NvlsBuf* b = new NvlsBuf();
  b->size = size;
  b->world = world;
  B200_CUDA_OK(cudaGetDevice(&b->dev));
  B200_CUDA_OK(cudaFree(nullptr));
  B200_CU_OK(mem_import(&b->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                        CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *buf_out = b;
  return 0;
}

/* Every rank, BEFORE any rank binds memory: add this process's device to the object. */
int b200_nvls_add_device(void* buf) {
  using namespace b200;
  B200_REQUIRE(buf, "b200_nvls_add_device: null buffer");
  B200_DRV(mc_add, "cuMulticastAddDevice", CUmemGenericAllocationHandle, CUdevice);
  NvlsBuf* b = static_cast<NvlsBuf*>(buf);
  B200_CU_OK(mc_add(b->mc, b->dev));
  return 0;
}

/* Every rank, AFTER all ranks have added their device (host-side barrier in between): allocate
 * this rank's physical memory, bind it to the object, map the local and the multicast view, zero
 * the memory.  Returns both pointers. */
int b200_nvls_bind(void* buf, void** uc_ptr, void** mc_ptr) {
  using namespace b200;
  B200_REQUIRE(buf && uc_ptr && mc_ptr, "b200_nvls_bind: null pointer");
  B200_DRV(mem_create, "cuMemCreate", CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
           unsigned long long);
  B200_DRV(mc_bind, "cuMulticastBindMem", CUmemGenericAllocationHandle, size_t,
           CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  B200_DRV(va_reserve, "cuMemAddressReserve", CUdeviceptr*, size_t, size_t, CUdeviceptr,
           unsigned long long);
  B200_DRV(mem_map, "cuMemMap", CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
           unsigned long long);
  B200_DRV(set_access, "cuMemSetAccess", CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  NvlsBuf* b = static_cast<NvlsBuf*>(buf);
  B200_REQUIRE(!b->bound, "b200_nvls_bind: already bound");
  CUmemAllocationProp ap = {};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = b->dev;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  B200_CU_OK(mem_create(&b->mem, b->size, &ap, 0));
  B200_CU_OK(mc_bind(b->mc, 0, b->mem, 0, b->size, 0));
  CUmemAccessDesc ad = {};
  ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ad.location.id = b->dev;
  ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B200_CU_OK(va_reserve(&b->uc_ptr, b->size, 0, 0, 0));
  B200_CU_OK(mem_map(b->uc_ptr, b->size, 0, b->mem, 0));
  B200_CU_OK(set_access(b->uc_ptr, b->size, &ad, 1));
  B200_CU_OK(va_reserve(&b->mc_ptr, b->size, 0, 0, 0));
  B200_CU_OK(mem_map(b->mc_ptr, b->size, 0, b->mc, 0));
  B200_CU_OK(set_access(b->mc_ptr, b->size, &ad, 1));
  B200_CUDA_OK(cudaMemset(reinterpret_cast<void*>(b->uc_ptr), 0, b->size));
  B200_CUDA_OK(cudaDeviceSynchronize());
  b->bound = true;
  *uc_ptr = reinterpret_cast<void*>(b->uc_ptr);
  *mc_ptr = reinterpret_cast<void*>(b->mc_ptr);
  return 0;
}

int b200_nvls_free(void* buf) {
  using namespace b200;
  if (!buf) return 0;
  NvlsBuf* b = static_cast<NvlsBuf*>(buf);
  typedef CUresult (*unmap_t)(CUdeviceptr, size_t);
  typedef CUresult (*rel_t)(CUmemGenericAllocationHandle);
  static unmap_t mem_unmap = drv<unmap_t>("cuMemUnmap");
  static unmap_t va_free = drv<unmap_t>("cuMemAddressFree");
  static rel_t mem_release = drv<rel_t>("cuMemRelease");
  if (mem_unmap && va_free && mem_release) {
    if (b->mc_ptr) { mem_unmap(b->mc_ptr, b->size); va_free(b->mc_ptr, b->size); }
    if (b->uc_ptr) { mem_unmap(b->uc_ptr, b->size); va_free(b->uc_ptr, b->size); }
    if (b->mem) mem_release(b->mem);
    if (b->mc) mem_release(b->mc);
  }
  delete b;
  return 0;
}

/* Hand a bound NVLS buffer to a communicator: from now on b200_comm_all_reduce reduces messages of
 * B200_COMM_NVLS_MIN (default 128 KiB) .. size / 2 bytes inside the switch. */
int b200_comm_attach_nvls(void* comm, void* nvls_buf) {
  using namespace b200;
  B200_REQUIRE(comm, "b200_comm_attach_nvls: null communicator");
  Comm* c = static_cast<Comm*>(comm);
  if (!nvls_buf) {
    c->nvls_uc = c->nvls_mc = nullptr;
    c->nvls_slot = 0;
    return 0;
  }
  NvlsBuf* b = static_cast<NvlsBuf*>(nvls_buf);
  B200_REQUIRE(b->bound && b->world == c->dev.world, "b200_comm_attach_nvls: buffer not bound or world mismatch");
  c->nvls_uc = reinterpret_cast<uint8_t*>(b->uc_ptr);
  c->nvls_mc = reinterpret_cast<uint8_t*>(b->mc_ptr);
  c->nvls_slot = (b->size / 2) / 1024 * 1024;
  return 0;
}

}  // extern "C"

// Symmetric-memory allocator: cuMem* virtual memory management + POSIX-fd export/import + NVSwitch
// multicast objects.  This is the control-plane half of the NVLink data path (SURVEY §5.9 "memory
// model"); the reference reaches the equivalent functionality through NCCL's P2P/IPC registration
// (notebooks/2_pytorch_dist_smddp_gpu.ipynb:400).  The driver API is resolved at run time through
// cudaGetDriverEntryPoint so the library imports on a machine without libcuda (the CPU dev box).
#include <cuda.h>
#include <stdarg.h>
#include <unistd.h>

#include "common.cuh"
#include <stdlib.h>

static thread_local char g_err[1024] = "";

bool b200_pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B200DDP_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

void b200_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_version(void) { return 100; }

namespace {

struct DriverApi {
  bool loaded = false;
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemGetAllocationGranularity) MemGetAllocationGranularity = nullptr;
  decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle = nullptr;
  decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle = nullptr;
  decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
  decltype(&cuMemAddressFree) MemAddressFree = nullptr;
  decltype(&cuMemMap) MemMap = nullptr;
  decltype(&cuMemUnmap) MemUnmap = nullptr;
  decltype(&cuMemSetAccess) MemSetAccess = nullptr;
  decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
  decltype(&cuMulticastCreate) MulticastCreate = nullptr;
  decltype(&cuMulticastAddDevice) MulticastAddDevice = nullptr;
  decltype(&cuMulticastBindMem) MulticastBindMem = nullptr;
  decltype(&cuMulticastGetGranularity) MulticastGetGranularity = nullptr;
  decltype(&cuMulticastUnbind) MulticastUnbind = nullptr;
  decltype(&cuGetErrorString) GetErrorString = nullptr;
};

DriverApi g_drv;

template <typename F>
bool load_sym(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult st;
  cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &st);
  if (e != cudaSuccess || fn == nullptr || st != cudaDriverEntryPointSuccess) {
    cudaGetLastError();
    b200_set_error("driver entry point %s unavailable (%s)", name, cudaGetErrorString(e));
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}

const char* drv_err(CUresult r) {
  const char* s = nullptr;
  if (g_drv.GetErrorString && g_drv.GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUresult";
}

#define DRV(call)                                                                         \
  do {                                                                                    \
    CUresult _r = (call);                                                                 \
    if (_r != CUDA_SUCCESS) {                                                             \
      b200_set_error("%s:%d %s -> %d %s", __FILE__, __LINE__, #call, (int)_r, drv_err(_r)); \
      return B200_ERR_DRIVER;                                                             \
    }                                                                                     \
  } while (0)

CUmemAllocationProp make_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_handle(int device, CUmemGenericAllocationHandle h, size_t bytes, size_t align, void** ptr_out) {
  CUdeviceptr va = 0;
  DRV(g_drv.MemAddressReserve(&va, bytes, align, 0, 0));
  CUresult r = g_drv.MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.MemAddressFree(va, bytes);
    b200_set_error("cuMemMap(%zu bytes) -> %d %s", bytes, (int)r, drv_err(r));
    return B200_ERR_DRIVER;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_drv.MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    g_drv.MemUnmap(va, bytes);
    g_drv.MemAddressFree(va, bytes);
    b200_set_error("cuMemSetAccess(dev %d, %zu bytes) -> %d %s", device, bytes, (int)r, drv_err(r));
    return B200_ERR_DRIVER;
  }
  *ptr_out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

extern "C" int b200_symm_init(int device) {
  B200_CHECK_CUDA(cudaSetDevice(device));
  B200_CHECK_CUDA(cudaFree(0));  // make sure the primary context exists
  if (g_drv.loaded) return 0;
  bool ok = load_sym("cuGetErrorString", &g_drv.GetErrorString) && load_sym("cuMemCreate", &g_drv.MemCreate) &&
            load_sym("cuMemRelease", &g_drv.MemRelease) &&
            load_sym("cuMemGetAllocationGranularity", &g_drv.MemGetAllocationGranularity) &&
            load_sym("cuMemExportToShareableHandle", &g_drv.MemExportToShareableHandle) &&
            load_sym("cuMemImportFromShareableHandle", &g_drv.MemImportFromShareableHandle) &&
            load_sym("cuMemAddressReserve", &g_drv.MemAddressReserve) &&
            load_sym("cuMemAddressFree", &g_drv.MemAddressFree) && load_sym("cuMemMap", &g_drv.MemMap) &&
            load_sym("cuMemUnmap", &g_drv.MemUnmap) && load_sym("cuMemSetAccess", &g_drv.MemSetAccess) &&
            load_sym("cuDeviceGetAttribute", &g_drv.DeviceGetAttribute);
  if (!ok) return B200_ERR_DRIVER;
  // multicast is optional: absence only disables the NVLS algorithm
  bool mc = load_sym("cuMulticastCreate", &g_drv.MulticastCreate) &&
            load_sym("cuMulticastAddDevice", &g_drv.MulticastAddDevice) &&
            load_sym("cuMulticastBindMem", &g_drv.MulticastBindMem) &&
            load_sym("cuMulticastGetGranularity", &g_drv.MulticastGetGranularity) &&
            load_sym("cuMulticastUnbind", &g_drv.MulticastUnbind);
  if (!mc) g_drv.MulticastCreate = nullptr;
  g_drv.loaded = true;
  return 0;
}

extern "C" int b200_symm_granularity(int device, size_t* gran_out, size_t* mc_gran_out) {
  B200_REQUIRE(g_drv.loaded, "b200_symm_init not called");
  CUmemAllocationProp prop = make_prop(device);
  size_t gran = 0;
  DRV(g_drv.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  *gran_out = gran;
  *mc_gran_out = 0;
  if (g_drv.MulticastCreate && b200_mc_supported(device) == 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = 2;
    mp.size = gran;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (g_drv.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) *mc_gran_out = mg;
  }
  return 0;
}

extern "C" int b200_symm_alloc(int device, size_t bytes, void** ptr_out, int* fd_out, uint64_t* handle_out) {
  B200_REQUIRE(g_drv.loaded, "b200_symm_init not called");
  CUmemAllocationProp prop = make_prop(device);
  CUmemGenericAllocationHandle h = 0;
  DRV(g_drv.MemCreate(&h, bytes, &prop, 0));
  int fd = -1;
  CUresult r = g_drv.MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.MemRelease(h);
    b200_set_error("cuMemExportToShareableHandle -> %d %s", (int)r, drv_err(r));
    return B200_ERR_DRIVER;
  }
  size_t gran = 0, mcg = 0;
  b200_symm_granularity(device, &gran, &mcg);
  int rc = map_handle(device, h, bytes, mcg > gran ? mcg : gran, ptr_out);
  if (rc != 0) {
    close(fd);
    g_drv.MemRelease(h);
    return rc;
  }
  *fd_out = fd;
  *handle_out = (uint64_t)h;
  return 0;
}

extern "C" int b200_symm_import(int device, int fd, size_t bytes, void** ptr_out, uint64_t* handle_out) {
  B200_REQUIRE(g_drv.loaded, "b200_symm_init not called");
  CUmemGenericAllocationHandle h = 0;
  DRV(g_drv.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  size_t gran = 0, mcg = 0;
  b200_symm_granularity(device, &gran, &mcg);
  int rc = map_handle(device, h, bytes, mcg > gran ? mcg : gran, ptr_out);
  if (rc != 0) {
    g_drv.MemRelease(h);
    return rc;
  }
  *handle_out = (uint64_t)h;
  return 0;
}

extern "C" int b200_symm_unmap(void* ptr, size_t bytes, uint64_t handle) {
  if (!g_drv.loaded) return 0;
  if (ptr) {
    g_drv.MemUnmap((CUdeviceptr)ptr, bytes);
    g_drv.MemAddressFree((CUdeviceptr)ptr, bytes);
  }
  if (handle) g_drv.MemRelease((CUmemGenericAllocationHandle)handle);
  return 0;
}

extern "C" int b200_mc_supported(int device) {
  if (!g_drv.loaded || !g_drv.MulticastCreate) return 0;
  int v = 0;
  if (g_drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  return v ? 1 : 0;
}

extern "C" int b200_mc_create(int ndev, size_t bytes, uint64_t* mc_handle_out, int* fd_out) {
  B200_REQUIRE(g_drv.loaded && g_drv.MulticastCreate, "multicast API unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = ndev;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mc = 0;
  DRV(g_drv.MulticastCreate(&mc, &mp));
  int fd = -1;
  CUresult r = g_drv.MemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.MemRelease(mc);
    b200_set_error("export multicast handle -> %d %s", (int)r, drv_err(r));
    return B200_ERR_DRIVER;
  }
  *mc_handle_out = (uint64_t)mc;
  *fd_out = fd;
  return 0;
}

extern "C" int b200_mc_import(int fd, uint64_t* mc_handle_out) {
  B200_REQUIRE(g_drv.loaded && g_drv.MulticastCreate, "multicast API unavailable");
  CUmemGenericAllocationHandle mc = 0;
  DRV(g_drv.MemImportFromShareableHandle(&mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_handle_out = (uint64_t)mc;
  return 0;
}

extern "C" int b200_mc_add_device(uint64_t mc_handle, int device) {
  B200_REQUIRE(g_drv.loaded && g_drv.MulticastCreate, "multicast API unavailable");
  DRV(g_drv.MulticastAddDevice((CUmemGenericAllocationHandle)mc_handle, (CUdevice)device));
  return 0;
}

extern "C" int b200_mc_bind_map(int device, uint64_t mc_handle, uint64_t mem_handle, size_t bytes, void** mc_ptr_out) {
  B200_REQUIRE(g_drv.loaded && g_drv.MulticastCreate, "multicast API unavailable");
  DRV(g_drv.MulticastBindMem((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle, 0,
                             bytes, 0));
  size_t gran = 0, mcg = 0;
  b200_symm_granularity(device, &gran, &mcg);
  return map_handle(device, (CUmemGenericAllocationHandle)mc_handle, bytes, mcg > gran ? mcg : gran, mc_ptr_out);
}

extern "C" int b200_mc_release(int device, uint64_t mc_handle, void* mc_ptr, size_t bytes) {
  if (!g_drv.loaded || !g_drv.MulticastCreate) return 0;
  if (mc_ptr) {
    g_drv.MemUnmap((CUdeviceptr)mc_ptr, bytes);
    g_drv.MemAddressFree((CUdeviceptr)mc_ptr, bytes);
  }
  if (mc_handle) {
    g_drv.MulticastUnbind((CUmemGenericAllocationHandle)mc_handle, (CUdevice)device, 0, bytes);
    g_drv.MemRelease((CUmemGenericAllocationHandle)mc_handle);
  }
  return 0;
}

extern "C" int b200_close_fd(int fd) { return fd >= 0 ? close(fd) : 0; }

extern "C" int b200_comm_alloc_errflag(int** host_ptr_out, int** dev_ptr_out) {
  int* h = nullptr;
  B200_CHECK_CUDA(cudaHostAlloc((void**)&h, sizeof(int) * 4, cudaHostAllocMapped));
  h[0] = h[1] = h[2] = h[3] = 0;
  int* d = nullptr;
  B200_CHECK_CUDA(cudaHostGetDevicePointer((void**)&d, h, 0));
  *host_ptr_out = h;
  *dev_ptr_out = d;
  return 0;
}

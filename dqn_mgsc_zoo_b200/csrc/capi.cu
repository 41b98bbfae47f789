// Library-wide pieces of the C ABI: error string, version, raw device buffers.
#include <stdarg.h>

#include "common.cuh"

namespace dqnb {
static thread_local char g_error[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
}  // namespace dqnb

using namespace dqnb;

extern "C" int dqnb_version(void) { return 100; }
extern "C" const char* dqnb_last_error(void) { return g_error; }

extern "C" int dqnb_device_count(int* count) {
  DQNB_REQUIRE(count, "null argument");
  DQNB_CUDA(cudaGetDeviceCount(count));
  return DQNB_OK;
}
extern "C" int dqnb_dev_alloc(int device, size_t bytes, void** dev_ptr) {
  DQNB_REQUIRE(dev_ptr, "null argument");
  DeviceGuard guard(device);
  DQNB_CUDA(cudaMalloc(dev_ptr, bytes ? bytes : 16));
  return DQNB_OK;
}
extern "C" int dqnb_dev_free(int device, void* dev_ptr) {
  DeviceGuard guard(device);
  DQNB_CUDA(cudaFree(dev_ptr));
  return DQNB_OK;
}
extern "C" int dqnb_copy_h2d(void* dev_dst, const void* host_src, size_t bytes, uintptr_t stream) {
  DQNB_CUDA(cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
  return DQNB_OK;
}
extern "C" int dqnb_copy_d2h(void* host_dst, const void* dev_src, size_t bytes, uintptr_t stream) {
  DQNB_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
  DQNB_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return DQNB_OK;
}
extern "C" int dqnb_stream_sync(uintptr_t stream) {
  DQNB_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return DQNB_OK;
}

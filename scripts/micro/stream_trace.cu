// Standalone harness for the stream kernel: conv1 double-Q forward at batch B (argv[1], default 1024) on
// random frames, timed with CUDA events, plus the event trace of CTA 0 (see t32s_trace).
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -I dqn_mgsc_zoo_b200/csrc -o /tmp/stream_trace scripts/micro/stream_trace.cu
#define DQNB_STREAM_KNOBS 1
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "nature_tf32.cuh"
#include "gemm_tf32_stream.cuh"
using namespace dqnb;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
int main(int argc, char** argv) {
  const int B = argc > 1 ? atoi(argv[1]) : 1024;
  const int which = argc > 2 ? atoi(argv[2]) : 0;   // 0: conv1 dq, 1: conv2 fwd
  uint8_t *x0, *x1; float *w0, *w1, *b0, *y[3]; long long* trace;
  size_t nx = (size_t)B * Conv1::IN_PER_IMG;
  CK(cudaMalloc(&x0, nx)); CK(cudaMalloc(&x1, nx)); CK(cudaMemset(x0, 3, nx)); CK(cudaMemset(x1, 5, nx));
  CK(cudaMalloc(&w0, 1 << 20)); CK(cudaMalloc(&w1, 1 << 20)); CK(cudaMalloc(&b0, 4096));
  CK(cudaMemset(w0, 0, 1 << 20)); CK(cudaMemset(w1, 0, 1 << 20)); CK(cudaMemset(b0, 0, 4096));
  for (auto& p : y) CK(cudaMalloc(&p, (size_t)B * Conv1::OUT_PER_IMG * 4));
  CK(cudaMalloc(&trace, 8 * 256 * 8)); CK(cudaMemset(trace, 0, 8 * 256 * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  SplitK sk{}; sk.dbg = trace;
  float ms = 0;
  if (which == 0) {
    T32ConvFwdDoubleQ<Conv1, uint8_t> p;
    p.Mh = B * Conv1::HOUT * Conv1::WOUT; p.M = 2 * p.Mh; p.N = 2 * Conv1::COUT; p.K = Conv1::KCOL;
    p.x[0] = x0; p.x[1] = x1; p.w[0] = w0; p.w[1] = w1; p.bias[0] = b0; p.bias[1] = b0;
    for (int z = 0; z < 3; ++z) p.y[z] = y[z];
    for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); CK(launch_gemm_tf32_stream(p, 1, 1, sk, 0)); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); cudaEventElapsedTime(&ms, e0, e1); }
    printf("conv1 dq B=%d: %.1f us, %d tiles -> %.2f us per tile per SM\n", B, ms * 1e3, p.M / 128, ms * 1e3 / ((p.M / 128 + 147) / 148));
  } else {
    T32ConvFwd<Conv2, float> p;
    p.M = B * Conv2::HOUT * Conv2::WOUT; p.N = Conv2::COUT; p.K = Conv2::KCOL;
    for (int z = 0; z < 3; ++z) { p.x[z] = y[z]; p.w[z] = w0; p.bias[z] = b0; p.y[z] = reinterpret_cast<float*>(z == 0 ? (void*)x0 : (void*)x1); }
    // outputs alias the (large enough?) frame buffers only for timing: B*81*64*4 <= B*28224 bytes
    for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); CK(launch_gemm_tf32_stream(p, 2, 1, sk, 0)); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); cudaEventElapsedTime(&ms, e0, e1); }
    printf("conv2 fwd (2 nets) B=%d: %.1f us\n", B, ms * 1e3);
  }
  std::vector<long long> h(8 * 256);
  CK(cudaMemcpy(h.data(), trace, h.size() * 8, cudaMemcpyDeviceToHost));
  long long t0 = h[0 * 256 + 0];
  const char* names[6] = {"prod_slot", "acc_avail", "mma_full", "tile_issued", "epi_begin", "epi_done"};
  for (int sidx = 0; sidx < 6; ++sidx) {
    printf("%-11s", names[sidx]);
    for (int i = 0; i < 40; ++i) printf(" %6lld", h[sidx * 256 + i] ? h[sidx * 256 + i] - t0 : -1);
    printf("\n");
  }
  return 0;
}

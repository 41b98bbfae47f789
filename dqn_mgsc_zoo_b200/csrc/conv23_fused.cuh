// conv2 + conv3 forward of the Nature CNN (networks.py:199-211) in ONE kernel for the small-batch learner step:
// one CTA per (image, net).  An image's conv3 input is exactly that image's conv2 output (81 positions x 64
// channels = 10 KB in fp16), so the second convolution needs no other CTA's data: the conv2 accumulators leave TMEM
// as fp16 into shared memory, the conv3 im2col tile is rebuilt from there with ld.shared / st.shared, and the conv3
// filters were pulled by TMA while conv2 was still running.  Against the two separate slab-GEMM launches this drops
// one kernel boundary and conv3's whole operand-ingest phase (147 KB per tile) from the dependent chain.
//
// Operands are fp16 on tcgen05.mma.kind::f16 with fp32 accumulation, exactly as in H16ConvFwd (nature_tf32.cuh):
//   conv2: A[81 x 512]  = im2col of the image's fp16 conv1 output a1h: the raw 25.6 KB image arrives with ONE bulk
//          copy and the 81 KB tile is built from it in shared memory (a k-block of 64 halves = two horizontally
//          adjacent input pixels x 32 channels = 128 contiguous bytes) - an SM ingests only ~50 bytes per clock
//          from L2, so the 4x im2col expansion must not cross that port,
//          B[64 x 512]  = fp16 K-major filter copy, 8 TMA boxes requested ahead of griddepcontrol.wait;
//          D in TMEM columns [0, 64)
//   conv3: A[49 x 576]  from the shared-memory copy of conv2's output (a k-block = one filter tap x 64 channels),
//          B[64 x 576]  = 9 TMA boxes, requested once the conv2 tile is built (they land under conv2's MMAs and
//          epilogue; the 9th goes into conv2's filter space when its MMAs are done);   D in TMEM columns [64, 128)
// fp32 outputs leave through a shared-memory tile as coalesced 512-byte warp stores (an accumulator row lives in one
// thread: stored directly, every st.global would touch 32 lines).
// The MMAs are M = 128: rows past 81 / 49 read whatever follows in shared memory and produce accumulator lanes
// nobody reads; the A tiles are therefore packed at 88 / 56 rows per k-block (11 / 7 KB instead of 16).
// Outputs: a2 fp32 (tf32-rounded, only where the backward pass reads it: null otherwise) and a3 fp32 (tf32-rounded,
// fc1's A operand) - the same values H16ConvFwd<Conv2> / <Conv3> store.
#pragma once
#include "nature_tf32.cuh"

namespace dqnb {

struct Conv23Args {
  const __half* a1h[kMaxNets];    // [batch, 20, 20, 32] fp16 conv1 output of net z
  const void* map2[kMaxNets];     // CUtensorMap of the fp16 conv2 filters [64][512] used by net z (box 64 x 64)
  const void* map3[kMaxNets];     // ... conv3 filters [64][576]
  const float* bias2[kMaxNets];
  const float* bias3[kMaxNets];
  float* a2[kMaxNets];            // [batch, 9, 9, 64] fp32 or null
  float* a3[kMaxNets];            // [batch, 7, 7, 64] fp32
  __half* a3h[kMaxNets];          // optional fp16 copy of a3 (A operand of the fp16 fc1 forward), or null
  unsigned long long* tl;         // optional timeline slot
  long long* dbg;                 // optional: CTA (0, 0) stamps clock64() at its phase boundaries (profiling)
};

constexpr int kC23Threads = 288;                     // 8 staging / epilogue warps + the MMA warp
constexpr int kC23A2Pitch = 88 * 128;                // bytes per conv2 k-block tile (81 rows -> 11 groups of 8)
constexpr int kC23A3Pitch = 56 * 128;                // bytes per conv3 k-block tile (49 rows -> 7 groups of 8)
constexpr int kC23ARegion = 7 * kC23A2Pitch + 16384; // the last k-block is read as a full 128-row tile
constexpr int kC23WRegion = 8 * 8192;                // 8 filter k-blocks of 64 rows x 128 bytes
constexpr int kC23Smem = kC23ARegion + 2 * kC23WRegion + 1024;
static_assert(8 * kC23A3Pitch + 16384 <= kC23ARegion, "the conv3 tile reuses the conv2 tile's space");
static_assert(kC23Smem <= 227 * 1024, "shared memory budget");
static_assert(Conv2::HOUT * Conv2::WOUT == 81 && Conv3::HOUT * Conv3::WOUT == 49 && Conv2::KCOL == 512 &&
              Conv3::KCOL == 576 && Conv2::COUT == 64 && Conv3::COUT == 64 && Conv2::CIN == 32, "layer shapes");

template <class T> __device__ __forceinline__ T c23_pick(const T (&a)[kMaxNets], int z) { return z == 0 ? a[0] : z == 1 ? a[1] : a[2]; }
__device__ __forceinline__ uint32_t c23_swz(int row, int chunk) {   // K-major SWIZZLE_128B position of a 16-byte chunk
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}

// 32 consecutive fp32 columns of this thread's TMEM lane: two loads in flight, one wait
__device__ __forceinline__ void c23_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
#pragma unroll
  for (int h = 0; h < 2; ++h)
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[16 * h + 0]), "=r"(r[16 * h + 1]), "=r"(r[16 * h + 2]), "=r"(r[16 * h + 3]), "=r"(r[16 * h + 4]),
          "=r"(r[16 * h + 5]), "=r"(r[16 * h + 6]), "=r"(r[16 * h + 7]), "=r"(r[16 * h + 8]), "=r"(r[16 * h + 9]),
          "=r"(r[16 * h + 10]), "=r"(r[16 * h + 11]), "=r"(r[16 * h + 12]), "=r"(r[16 * h + 13]), "=r"(r[16 * h + 14]),
          "=r"(r[16 * h + 15])
        : "r"(taddr + 16 * h)
        : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__global__ void __launch_bounds__(kC23Threads, 1) conv23_fwd_kernel(Conv23Args g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full2[8];     // conv2 k-block i: A built (8 warps) + B landed (TMA)
  __shared__ uint64_t a3full[9];    // conv3 tap i: A rebuilt from conv2's output (8 warps)
  __shared__ uint64_t rawfull, done2, w3a, w3b, done3;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[128];   // conv2 | conv3 biases of this net
  const int t = threadIdx.x, warp = t >> 5;
  const int img = blockIdx.x, z = blockIdx.y;
  const unsigned long long tl_t0 = tl_begin(g.tl);
  const bool dbg = g.dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && t == 0;
  if (dbg) g.dbg[0] = clock64();
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t A = smem, W2 = smem + kC23ARegion, W3 = W2 + kC23WRegion;
  const uint32_t RAW = W3;          // the raw fp16 conv1 image [400][64 bytes], until the conv2 tile is built
  const uint32_t A2H = W2 + 8192;   // conv2's fp16 output [81][128 bytes] (chunks swizzled by row & 7); lives in
                                    // conv2's filter space once its MMAs are done, behind the 9th conv3 filter box
  const uint32_t OUT = A + 8 * kC23A3Pitch + 16384;   // fp32 output staging [rows][256 bytes], chunks swizzled; behind
                                                      // everything the conv3 MMAs read
  static_assert(8 * kC23A3Pitch + 16384 + 81 * 256 <= kC23ARegion, "output staging fits behind the conv3 tile");
  static_assert(Conv2::IN_PER_IMG * 2 <= kC23WRegion, "the raw image borrows the conv3 filter space");
  if (t == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&full2[i], 8 + 1);
    for (int i = 0; i < 9; ++i) mbar_init(&a3full[i], 8);
    mbar_init(&rawfull, 1); mbar_init(&done2, 1); mbar_init(&w3a, 1); mbar_init(&w3b, 1); mbar_init(&done3, 1);
    fence_mbar_init();
  }
  if (warp == 8) t32_alloc<128>(&tmem_slot);
  if (t < 32)   // parameters: final before the update starts
    reinterpret_cast<float4*>(bias_s)[t] =
        __ldg(reinterpret_cast<const float4*>(t < 16 ? c23_pick(g.bias2, z) : c23_pick(g.bias3, z)) + (t & 15));
  t32_fence_before();
  __syncthreads();
  t32_fence_after();
  const uint32_t tmem = tmem_slot;
  const void* map3 = c23_pick(g.map3, z);
  if (warp == 8 && elect_one()) {   // the filters are final before the update starts: request conv2's ahead of the dependency wait
    const void* map2 = c23_pick(g.map2, z);
    for (int i = 0; i < 8; ++i) {
      mbar_expect_tx(&full2[i], 8192);
      tma_load_2d(W2 + i * 8192, map2, i * 64, 0, smem_u32(&full2[i]));
    }
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;");   // fc1's CTAs set up and pull their weights on the idle SMs
  if (warp == 8 && elect_one()) {
    mbar_expect_tx(&rawfull, Conv2::IN_PER_IMG * 2);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(RAW),
                 "l"(c23_pick(g.a1h, z) + (size_t)img * Conv2::IN_PER_IMG), "r"(Conv2::IN_PER_IMG * 2),
                 "r"(smem_u32(&rawfull))
                 : "memory");
  }
  if (dbg) g.dbg[1] = clock64();   // setup done

  const uint32_t idesc = t32_idesc_f16(64);
  if (warp < 8) {
    // ---- conv2 A tile from the raw image: thread t moves chunk c = t & 7 of rows (t >> 3) + 32 u, k-block by k-block
    const int c = t & 7;
    int off[3]; uint32_t dst[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int r = (t >> 3) + 32 * u;
      const int oy = r / 9, ox = r - oy * 9;
      off[u] = r < 81 ? ((2 * oy) * Conv2::WIN + 2 * ox) * (Conv2::CIN * 2) + c * 16 : -1;   // bytes into RAW
      dst[u] = c23_swz(r, c);
    }
    mbar_wait(&rawfull, 0);
    if (dbg) g.dbg[2] = clock64();   // raw image landed
#pragma unroll
    for (int kp = 0; kp < 8; kp += 4) {   // four k-blocks of loads in flight, then their stores
      float4 mv[4][3];
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) {
        const int kb = kp + k4;
        const int koff = ((kb >> 1) * Conv2::WIN + 2 * (kb & 1)) * (Conv2::CIN * 2);
#pragma unroll
        for (int u = 0; u < 3; ++u)
          if (off[u] >= 0) mv[k4][u] = ld_shared_f4(RAW + off[u] + koff);
      }
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) {
        const uint32_t tile = A + (kp + k4) * kC23A2Pitch;
#pragma unroll
        for (int u = 0; u < 3; ++u)
          if (off[u] >= 0) st16(tile + dst[u], mv[k4][u].x, mv[k4][u].y, mv[k4][u].z, mv[k4][u].w);
        fence_proxy_async_smem();
        mbar_arrive_warp(&full2[kp + k4]);
      }
    }
    if (dbg) g.dbg[3] = clock64();   // conv2 tile built
    // ---- epilogues: accumulator row = TMEM lane = t mod 128; warps 0-3 take columns [0, 32), warps 4-7 [32, 64)
    const int row = t & 127, half = t >> 7;
    const uint32_t lane_addr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + half * 32;
    // ---- conv2 epilogue: bias + relu; fp16 to smem (conv3's input), fp32 (tf32-rounded) to the staging tile
    mbar_wait(&done2, 0);
    t32_fence_after();
    if (dbg) g.dbg[4] = clock64();   // conv2 MMAs complete
    float* out2 = c23_pick(g.a2, z);
    {
      float v[32];
      c23_ld32(lane_addr, v);
      if (row < 81) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 bb = *reinterpret_cast<const float4*>(bias_s + half * 32 + i);
          v[i] = fmaxf(v[i] + bb.x, 0.f); v[i + 1] = fmaxf(v[i + 1] + bb.y, 0.f);
          v[i + 2] = fmaxf(v[i + 2] + bb.z, 0.f); v[i + 3] = fmaxf(v[i + 3] + bb.w, 0.f);
        }
        if (out2 != nullptr) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 o = tf32_rn4(make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
            st16(OUT + row * 256 + (((half * 8 + (i >> 2)) ^ (row & 7)) << 4), o.x, o.y, o.z, o.w);
          }
        }
#pragma unroll
        for (int i = 0; i < 32; i += 8) {   // 8 halves = one 16-byte chunk
          const __half2 h0 = __floats2half2_rn(v[i], v[i + 1]), h1 = __floats2half2_rn(v[i + 2], v[i + 3]);
          const __half2 h2 = __floats2half2_rn(v[i + 4], v[i + 5]), h3 = __floats2half2_rn(v[i + 6], v[i + 7]);
          const int chunk = half * 4 + (i >> 3);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(A2H + row * 128 + ((chunk ^ (row & 7)) << 4)),
                       "r"(*reinterpret_cast<const uint32_t*>(&h0)), "r"(*reinterpret_cast<const uint32_t*>(&h1)),
                       "r"(*reinterpret_cast<const uint32_t*>(&h2)), "r"(*reinterpret_cast<const uint32_t*>(&h3))
                       : "memory");
        }
      }
    }
    t32_fence_before();
    asm volatile("bar.sync 1, 256;" ::: "memory");   // the 8 staging warps: conv2's output is complete in smem
    if (dbg) g.dbg[5] = clock64();   // conv2 epilogue done
    // ---- conv3 A tile from smem: thread t moves chunk c = t & 7 of rows (t >> 3) and (t >> 3) + 32 for every tap;
    // all loads first, then tap by tap: store, make visible to the tensor core, publish the tap
    {
      int r2base[2]; uint32_t d3[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int r3 = (t >> 3) + 32 * u;
        const int oy = r3 / 7, ox = r3 - oy * 7;
        r2base[u] = r3 < 49 ? oy * 9 + ox : -1;
        d3[u] = c23_swz(r3, c);
      }
      float4 mv[9][2];
#pragma unroll
      for (int kb = 0; kb < 9; ++kb) {
        const int tap = (kb / 3) * 9 + kb % 3;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (r2base[u] >= 0) {
            const int r2 = r2base[u] + tap;
            mv[kb][u] = ld_shared_f4(A2H + r2 * 128 + ((c ^ (r2 & 7)) << 4));
          }
        }
      }
#pragma unroll
      for (int kb = 0; kb < 9; ++kb) {
#pragma unroll
        for (int u = 0; u < 2; ++u)
          if (r2base[u] >= 0) st16(A + kb * kC23A3Pitch + d3[u], mv[kb][u].x, mv[kb][u].y, mv[kb][u].z, mv[kb][u].w);
        fence_proxy_async_smem();
        mbar_arrive_warp(&a3full[kb]);
      }
    }
    if (dbg) g.dbg[7] = clock64();   // conv3 tile rebuilt  (reported as "sync1": dbg[7] - dbg[5])
    // conv2's fp32 output leaves while the conv3 MMAs run: 81 rows x 16 chunks, a warp writes 512 contiguous bytes
    if (out2 != nullptr) {
      float4* dst2 = reinterpret_cast<float4*>(out2 + (size_t)img * 81 * 64);
      for (int i = t; i < 81 * 16; i += 256) {
        const int r = i >> 4, ch = i & 15;
        dst2[i] = ld_shared_f4(OUT + r * 256 + ((ch ^ (r & 7)) << 4));
      }
    }
    // ---- conv3 epilogue, through the same staging tile
    mbar_wait(&done3, 0);
    t32_fence_after();
    asm volatile("bar.sync 1, 256;" ::: "memory");   // (everybody has read conv2's rows out of the staging tile)
    {
      float v[32];
      c23_ld32(lane_addr + 64, v);
      if (row < 49) {
        __half* oh = c23_pick(g.a3h, z);
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 bb = *reinterpret_cast<const float4*>(bias_s + 64 + half * 32 + i);
          const float4 r = make_float4(fmaxf(v[i] + bb.x, 0.f), fmaxf(v[i + 1] + bb.y, 0.f), fmaxf(v[i + 2] + bb.z, 0.f),
                                       fmaxf(v[i + 3] + bb.w, 0.f));
          const float4 o = tf32_rn4(r);
          st16(OUT + row * 256 + (((half * 8 + (i >> 2)) ^ (row & 7)) << 4), o.x, o.y, o.z, o.w);
          if (oh != nullptr) t32_store_half4(oh + ((size_t)img * 49 + row) * 64 + half * 32 + i, r);
        }
      }
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    {
      float4* dst3 = reinterpret_cast<float4*>(c23_pick(g.a3, z) + (size_t)img * 49 * 64);
      for (int i = t; i < 49 * 16; i += 256) {
        const int r = i >> 4, ch = i & 15;
        dst3[i] = ld_shared_f4(OUT + r * 256 + ((ch ^ (r & 7)) << 4));
      }
    }
  } else {
    // ---- MMA issuer: the whole warp walks the barriers, one elected lane issues (descriptors stay warp-uniform:
    // a base descriptor plus an immediate per k-step)
    const bool leader = elect_one();
    const uint64_t a_d = t32_desc(A, 16, 1024), w2_d = t32_desc(W2, 16, 1024), w3_d = t32_desc(W3, 16, 1024);
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) {
      mbar_wait(&full2[kb], 0);
      t32_fence_after();
      if (kb == 7 && leader) {   // every warp is done reading the raw image: its space is the conv3 filters' from here on
        mbar_expect_tx(&w3a, 8 * 8192);
        for (int i = 0; i < 8; ++i) tma_load_2d(W3 + i * 8192, map3, i * 64, 0, smem_u32(&w3a));
      }
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          t32_mma_f16(tmem, a_d + ((kb * kC23A2Pitch + ks * 32) >> 4), w2_d + ((kb * 8192 + ks * 32) >> 4), idesc,
                      kb > 0 || ks > 0);
      }
    }
    if (leader) t32_commit(&done2);
    mbar_wait(&done2, 0);   // conv2's filter space is free: the 9th conv3 filter box goes to its first 8 KB
    if (leader) {
      mbar_expect_tx(&w3b, 8192);
      tma_load_2d(W2, map3, 8 * 64, 0, smem_u32(&w3b));
    }
    mbar_wait(&w3a, 0);
#pragma unroll
    for (int kb = 0; kb < 9; ++kb) {
      if (kb == 8) mbar_wait(&w3b, 0);
      mbar_wait(&a3full[kb], 0);
      t32_fence_after();
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          t32_mma_f16(tmem + 64, a_d + ((kb * kC23A3Pitch + ks * 32) >> 4),
                      (kb < 8 ? w3_d + ((kb * 8192 + ks * 32) >> 4) : w2_d + ((ks * 32) >> 4)), idesc, kb > 0 || ks > 0);
      }
    }
    if (leader) t32_commit(&done3);
  }
  if (dbg) g.dbg[6] = clock64();   // conv3 MMAs + epilogue done  (reported as "fixup": dbg[6] - dbg[5])
  t32_fence_before();
  __syncthreads();
  if (warp == 8) t32_dealloc<128>(tmem);
  tl_end(g.tl, tl_t0);
}

inline cudaError_t launch_conv23_fwd(const Conv23Args& g, int batch, int nets, cudaStream_t st, int priority, bool pdl) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(conv23_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kC23Smem);
    if (e != cudaSuccess) { cudaGetLastError(); return e; }
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(batch, nets); cfg.blockDim = dim3(kC23Threads); cfg.dynamicSmemBytes = kC23Smem; cfg.stream = st;
  cudaLaunchAttribute attr[2];
  unsigned na = 0;
  attr[na].id = cudaLaunchAttributePriority;
  attr[na++].val.priority = priority;
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na++].val.programmaticStreamSerializationAllowed = 1;
  }
  cfg.attrs = attr; cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, conv23_fwd_kernel, g);
}

}  // namespace dqnb

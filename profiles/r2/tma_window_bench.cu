// "measure the TMA option once, with numbers".  Stand-alone micro-benchmark of the reference-window stage of
// k_sad_table (thor_b200/csrc/sad.cu: a (64 + 2R) x pitch box of the padded reference luma per 64x64 tile, double buffered):
//   A  what the kernel does today: 16-byte cp.async (LDGSTS) per thread, commit / wait_group
//   B  one cp.async.bulk.tensor.2d (UTMALDG) per tile issued by one thread, completion on an mbarrier
// Both stage the same boxes of the same pictures through two shared-memory buffers, persistent CTAs, and "consume" a tile
// with one pass of packed byte sums (so the loads cannot be dropped and the two variants can be compared word for word).
// The real kernel spends ~330 instructions of VABSDIFF4 work per 16 bytes staged; this measures the staging alone.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tma_window_bench profiles/r2/tma_window_bench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s -> %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int R = 16, BOX_H = 64 + 2 * R, BOX_W = 112;          // SadCfg<16>: 96 rows, pitch 112 (64 + 32 + 16 of alignment slack)
constexpr int BOX_BYTES = BOX_H * BOX_W, THREADS = 256, PAD = 96;

struct Plane { const uint8_t *origin; int stride, w, h; };      // origin = visible (0,0); PAD samples of border on every side

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// one pass over a staged box: every thread sums its share of 4-byte words (dp4a with ones = byte sum)
__device__ __forceinline__ uint32_t consume(const uint8_t *box, int tid) {
  uint32_t s = 0;
  const uint32_t *w = reinterpret_cast<const uint32_t *>(box);
  for (int i = tid; i < BOX_BYTES / 4; i += THREADS) s = __dp4a(w[i], 0x01010101u, s);
  return s;
}
__device__ __forceinline__ uint32_t block_sum(uint32_t v, uint32_t *red) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  uint32_t t = 0;
  for (int k = 0; k < THREADS / 32; k++) t += red[k];
  return t;
}

// tile L -> (picture z, box origin): the window of superblock (sx, sy) around displacement 0, x aligned down to 16
__device__ __forceinline__ void tile_of(int L, int gx, int gy, int &z, int &ax, int &wy) {
  z = L / (gx * gy);
  const int rem = L - z * gx * gy, sy = rem / gx, sx = rem - sy * gx;
  ax = (sx * 64 - R) & ~15;
  wy = sy * 64 - R;
}

// ---- A: cp.async, 16 bytes per thread -----------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, 2) k_stage_cpasync(const Plane *planes, int gx, int gy, int ntiles, uint32_t *sums) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t red[THREADS / 32];
  const int tid = threadIdx.x;
  auto issue = [&](int L, int st) {
    int z, ax, wy;
    tile_of(L, gx, gy, z, ax, wy);
    const Plane P = planes[z];
    constexpr int cpr = BOX_W / 16;
    for (int i = tid; i < BOX_H * cpr; i += THREADS) {
      const int r = i / cpr, c = i - r * cpr;
      const uint8_t *src = P.origin + (ptrdiff_t)(wy + r) * P.stride + ax + 16 * c;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem + st * BOX_BYTES + r * BOX_W + 16 * c)), "l"(src));
    }
  };
  int L = blockIdx.x, st = 0;
  if (L < ntiles) issue(L, 0);
  asm volatile("cp.async.commit_group;");
  while (L < ntiles) {
    const int Ln = L + gridDim.x;
    if (Ln < ntiles) issue(Ln, st ^ 1);
    asm volatile("cp.async.commit_group;");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    const uint32_t s = block_sum(consume(smem + st * BOX_BYTES, tid), red);
    if (tid == 0) sums[L] = s;
    __syncthreads();                                     // the buffer is free for the tile after next
    L = Ln;
    st ^= 1;
  }
}

// ---- B: TMA, one bulk tensor copy per tile ------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, 2) k_stage_tma(const CUtensorMap *maps, int gx, int gy, int ntiles, uint32_t *sums) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t red[THREADS / 32];
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int k = 0; k < 2; k++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[k])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int L, int st) {                      // thread 0 only
    int z, ax, wy;
    tile_of(L, gx, gy, z, ax, wy);
    const uint32_t b = smem_u32(&bar[st]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(BOX_BYTES) : "memory");
    // tensor coordinates are relative to the allocation's first byte: the visible origin sits at (PAD, PAD)
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(smem + st * BOX_BYTES)),
                 "l"(maps + z), "r"(ax + PAD), "r"(wy + PAD), "r"(b)
                 : "memory");
  };
  int L = blockIdx.x, st = 0;
  uint32_t phase[2] = {0, 0};
  if (tid == 0 && L < ntiles) issue(L, 0);
  while (L < ntiles) {
    const int Ln = L + gridDim.x;
    if (tid == 0 && Ln < ntiles) issue(Ln, st ^ 1);
    {
      const uint32_t b = smem_u32(&bar[st]);
      uint32_t done = 0;
      while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b), "r"(phase[st]) : "memory");
      phase[st] ^= 1;
    }
    const uint32_t s = block_sum(consume(smem + st * BOX_BYTES, tid), red);
    if (tid == 0) sums[L] = s;
    __syncthreads();                                     // all reads of the buffer are done before thread 0 re-arms it
    L = Ln;
    st ^= 1;
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                             const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char **argv) {
  const int W = 1920, H = 1080, NPIC = argc > 1 ? atoi(argv[1]) : 32, REPS = 20;
  const int stride = (W + 2 * PAD + 15) & ~15, rows = H + 2 * PAD;
  const int gx = (W + 63) / 64, gy = (H + 63) / 64, ntiles = gx * gy * NPIC;
  CK(cudaSetDevice(0));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  uint8_t *pics;
  CK(cudaMalloc(&pics, (size_t)NPIC * stride * rows + 4096));
  std::vector<uint8_t> h((size_t)NPIC * stride * rows);
  uint32_t x = 12345;
  for (auto &b : h) { x = x * 1664525u + 1013904223u; b = (uint8_t)(x >> 24); }
  CK(cudaMemcpy(pics, h.data(), h.size(), cudaMemcpyHostToDevice));
  std::vector<Plane> planes(NPIC);
  std::vector<CUtensorMap> maps(NPIC);
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &q));
  if (!encode) { fprintf(stderr, "cuTensorMapEncodeTiled not available\n"); return 1; }
  for (int z = 0; z < NPIC; z++) {
    uint8_t *base = pics + (size_t)z * stride * rows;
    planes[z] = Plane{base + (size_t)PAD * stride + PAD, stride, W, H};
    const cuuint64_t dims[2] = {(cuuint64_t)stride, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)stride};
    const cuuint32_t box[2] = {BOX_W, BOX_H}, estr[2] = {1, 1};
    const CUresult r = encode(&maps[z], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fprintf(stderr, "cuTensorMapEncodeTiled failed: %d\n", (int)r); return 1; }
  }
  Plane *d_planes;
  CUtensorMap *d_maps;
  uint32_t *d_sums[2];
  CK(cudaMalloc(&d_planes, NPIC * sizeof(Plane)));
  CK(cudaMalloc(&d_maps, NPIC * sizeof(CUtensorMap)));
  CK(cudaMemcpy(d_planes, planes.data(), NPIC * sizeof(Plane), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_maps, maps.data(), NPIC * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  for (int k = 0; k < 2; k++) { CK(cudaMalloc(&d_sums[k], ntiles * 4)); CK(cudaMemset(d_sums[k], 0, ntiles * 4)); }
  const int smem = 2 * BOX_BYTES, grid = prop.multiProcessorCount * 2;
  CK(cudaFuncSetAttribute(k_stage_cpasync, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_stage_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float ms[2] = {0, 0};
  for (int v = 0; v < 2; v++) {
    for (int it = 0; it < 3 + REPS; it++) {
      if (it == 3) CK(cudaEventRecord(e0));
      if (v == 0) k_stage_cpasync<<<grid, THREADS, smem>>>(d_planes, gx, gy, ntiles, d_sums[0]);
      else k_stage_tma<<<grid, THREADS, smem>>>(d_maps, gx, gy, ntiles, d_sums[1]);
    }
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    CK(cudaEventElapsedTime(&ms[v], e0, e1));
    ms[v] /= REPS;
  }
  std::vector<uint32_t> s0(ntiles), s1(ntiles);
  CK(cudaMemcpy(s0.data(), d_sums[0], ntiles * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(s1.data(), d_sums[1], ntiles * 4, cudaMemcpyDeviceToHost));
  int diff = 0;
  for (int i = 0; i < ntiles; i++) diff += s0[i] != s1[i];
  const double gb = (double)ntiles * BOX_BYTES * 1e-9;
  printf("{\"pictures\": %d, \"tiles\": %d, \"box\": \"%dx%d\", \"staged_GB_per_launch\": %.4f, \"cp_async_ms\": %.4f, \"cp_async_GBps\": %.1f, "
         "\"tma_ms\": %.4f, \"tma_GBps\": %.1f, \"tiles_that_differ\": %d, \"grid\": %d, \"threads\": %d}\n",
         NPIC, ntiles, BOX_W, BOX_H, gb, ms[0], gb / (ms[0] * 1e-3), ms[1], gb / (ms[1] * 1e-3), diff, grid, THREADS);
  return diff != 0;
}

// Frame encoder of libthorcuda -- common definitions of the superblock encoder (enc_core.h).
//
// The superblock encoder is written ONCE as "uniform control flow + cooperative pixel loops":
//   * every thread of the CTA executes the same scalar decision code on the same values (decisions live in registers,
//     replicated; there is nothing to broadcast),
//   * pixel / coefficient arrays live in shared memory and are only written inside TFOR loops, which spread the
//     iterations over the CTA and are followed by a barrier before anybody reads what another thread wrote.
// Compiled by nvcc for sm_100a this is the product.  Compiled by g++ with THOR_HOSTEMU (one "thread", barriers are
// no-ops) the very same source runs on the host: tests/hostemu uses that build to check the decision logic against the
// reference encoder byte for byte in a container without a GPU.  The host build is test infrastructure only --
// libthorcuda never contains it and has no CPU path.
#pragma once
#include <stdint.h>
#include "../../../include/thorcuda_enc.h"

#ifdef THOR_HOSTEMU
#include <math.h>
#include <stdlib.h>
#include <string.h>
#define TD static inline
#define TDN static __attribute__((noinline))
#define TCONST static const
#else
#include <cuda_runtime.h>
#define TD __device__ __forceinline__
#define TDN __device__ __noinline__
#define TCONST __device__ const
#endif

// every array the superblock encoder works on lives in shared memory; the device functions are not inlined, so each
// one tells the compiler (LDS / STS with immediate offsets instead of generic loads through a converted address)
#ifdef THOR_HOSTEMU
#define THOR_SHARED(p) ((void)0)
#define THOR_SHS(p) ((void)0)
#define THOR_SHF(p) ((void)0)
#define THOR_SHR(p) ((void)0)
#define THOR_LOCAL(p) ((void)0)
#else
#define THOR_ASSUME_SHARED(p) __builtin_assume(__isShared(p))
// Measured on the B200 (round 2, one build variant per group, .bit compared with the reference's): the hint on the SbMem
// reference (S) and on the reduction scratch (R) kept the streams byte-identical in the 255-register build and took a
// 416x240 frame from 115 to 95 ms (1080p: 1214 -> 1141 ms).  But the same source compiled for 128 registers
// (__launch_bounds__(256, 2) or 512 threads) dies with an illegal memory access in the first bi-predicted frame WITH the S / R
// hints and runs byte-identically WITHOUT them (profiles/r2/crash_matrix.py), the hint on pointer PARAMETERS into SbMem (P)
// changed decisions although every such pointer does point into shared memory, and on the frame description (F) it was
// slower.  Whatever the cause, a 6 % gain does not pay for code whose correctness depends on the register allocation:
// ALL hints are off unless the build asks for them (-DTHOR_ENC_HINTS, plus -DTHOR_HINT_P / -DTHOR_HINT_F).
#ifdef THOR_ENC_HINTS
#define THOR_SHS(p) THOR_ASSUME_SHARED(p)
#define THOR_SHR(p) THOR_ASSUME_SHARED(p)
#else
#define THOR_SHS(p) ((void)0)
#define THOR_SHR(p) ((void)0)
#endif
#ifdef THOR_HINT_P
#define THOR_SHARED(p) THOR_ASSUME_SHARED(p)
#else
#define THOR_SHARED(p) ((void)0)
#endif
#ifdef THOR_HINT_F
#define THOR_SHF(p) THOR_ASSUME_SHARED(p)
#else
#define THOR_SHF(p) ((void)0)
#endif
// The matching hint for per-thread state handed down by reference (__isLocal) is NOT used: with it the kernel hung on the
// B200 in the bi-prediction path (round 2, build variants v1..v3 under gpurun); cause not understood.  THOR_ENC_LOCAL_HINT
// re-enables it for experiments.
#ifdef THOR_ENC_LOCAL_HINT
#define THOR_LOCAL(p) __builtin_assume(__isLocal(p))
#else
#define THOR_LOCAL(p) ((void)0)
#endif
#endif

namespace thorenc {

#ifndef THOR_ENC_NT
#define THOR_ENC_NT 256
#endif
constexpr int NT = THOR_ENC_NT;      // threads per CTA (one CTA walks one superblock row)
constexpr int NW = NT / 32;

// ---- the cooperating thread group ------------------------------------------------------------------------------
struct Grp {
#ifdef THOR_HOSTEMU
  static constexpr int tid = 0;
  static constexpr int n = 1;
  void sync() const {}
  int sum(int v) { return v; }
  template <int K> void sumk(int (&v)[K]) {}
  int any(int p) { return p != 0; }
  int vmax(int v) { return v; }
  void vmax3(int (&v)[3]) {}
  int orall(int m) { return m; }
#else
  int tid;
  static constexpr int n = NT;
  int *red;                          // shared: 2 * NW * 8 ints
  int flip;
  __device__ void sync() const { __syncthreads(); }
  // Reductions over the group: one redux.sync per value inside a warp, the warps' partial results through `red` (double
  // buffered: the next reduction writes the other half, so nobody waits for readers), and a second redux over the partials --
  // no thread walks all warps' slots.
  __device__ int sum(int v) {
    v = __reduce_add_sync(0xffffffffu, v);
    THOR_SHR(red);
    int *r = red + (flip ^= 1) * NW * 8;
    if ((tid & 31) == 0) r[tid >> 5] = v;
    __syncthreads();
    const int lane = tid & 31;
    int t = 0;
#pragma unroll
    for (int i = 0; i < (NW + 31) / 32; i++) t += lane + 32 * i < NW ? r[lane + 32 * i] : 0;
    return __reduce_add_sync(0xffffffffu, t);
  }
  template <int K> __device__ void sumk(int (&v)[K]) {
    static_assert(K <= 8, "reduction scratch holds 8 values per warp");
#pragma unroll
    for (int k = 0; k < K; k++) v[k] = __reduce_add_sync(0xffffffffu, v[k]);
    THOR_SHR(red);
    int *r = red + (flip ^= 1) * NW * 8;
    if ((tid & 31) == 0)
#pragma unroll
      for (int k = 0; k < 8; k++) r[(tid >> 5) * 8 + k] = k < K ? v[k] : 0;
    __syncthreads();
    // slot w * 8 + k: lane l adds the slots l, l + 32, .. (all of value l % 8), then the four lanes of a value add up
    const int lane = tid & 31;
    int x = 0;
#pragma unroll
    for (int i = lane; i < NW * 8; i += 32) x += r[i];
    x += __shfl_xor_sync(0xffffffffu, x, 8);
    x += __shfl_xor_sync(0xffffffffu, x, 16);
#pragma unroll
    for (int k = 0; k < K; k++) v[k] = __shfl_sync(0xffffffffu, x, k);
  }
  __device__ int any(int p) { return __syncthreads_or(p); }
  __device__ int orall(int m) {                                        // bitwise OR of m over the group
    m = (int)__reduce_or_sync(0xffffffffu, (unsigned)m);
    THOR_SHR(red);
    int *r = red + (flip ^= 1) * NW * 8;
    if ((tid & 31) == 0) r[tid >> 5] = m;
    __syncthreads();
    const int lane = tid & 31;
    unsigned t = 0;
#pragma unroll
    for (int i = 0; i < (NW + 31) / 32; i++) t |= lane + 32 * i < NW ? (unsigned)r[lane + 32 * i] : 0u;
    return (int)__reduce_or_sync(0xffffffffu, t);
  }
  __device__ void vmax3(int (&v)[3]) {                                 // three maxima with one barrier
#pragma unroll
    for (int k = 0; k < 3; k++) v[k] = __reduce_max_sync(0xffffffffu, v[k]);
    THOR_SHR(red);
    int *r = red + (flip ^= 1) * NW * 8;
    if ((tid & 31) == 0)
#pragma unroll
      for (int k = 0; k < 3; k++) r[(tid >> 5) * 8 + k] = v[k];
    __syncthreads();
    const int lane = tid & 31;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      int t = r[k];                                                    // warp 0's partial: a valid value for idle lanes
#pragma unroll
      for (int i = 0; i < (NW + 31) / 32; i++)
        if (lane + 32 * i < NW) t = max(t, r[(lane + 32 * i) * 8 + k]);
      v[k] = __reduce_max_sync(0xffffffffu, t);
    }
  }
  __device__ int vmax(int v) {
    v = __reduce_max_sync(0xffffffffu, v);
    THOR_SHR(red);
    int *r = red + (flip ^= 1) * NW * 8;
    if ((tid & 31) == 0) r[tid >> 5] = v;
    __syncthreads();
    const int lane = tid & 31;
    int t = r[0];
#pragma unroll
    for (int i = 0; i < (NW + 31) / 32; i++)
      if (lane + 32 * i < NW) t = max(t, r[lane + 32 * i]);
    return __reduce_max_sync(0xffffffffu, t);
  }
#endif
};


// ---- optional cycle accounting (-DTHOR_ENC_PROF): thread 0 of a CTA adds the clocks spent inside a scope to a category
enum { PF_SB = 0, PF_WAIT, PF_LOAD_ORG, PF_EARLY, PF_PRED_INTER, PF_PRED_INTRA, PF_TB, PF_BITS, PF_SSD, PF_KEEP, PF_SEARCH_INTRA, PF_SAD, PF_HALF,
       PF_QUARTER, PF_ME, PF_FINAL, PF_NEIGH, PF_BIPRED, PF_MODE, PF_COUNT };
#if defined(THOR_ENC_PROF) && !defined(THOR_HOSTEMU)
__device__ unsigned long long g_enc_prof[32];
__device__ __forceinline__ long long prof_clock() {
  long long t = 0;
#ifdef __CUDA_ARCH__
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));
#endif
  return t;
}
struct Prof {
  int c; long long t0;
  __device__ Prof(int cat) : c(cat), t0(prof_clock()) {}
  __device__ ~Prof() { if (threadIdx.x == 0) atomicAdd(&g_enc_prof[c], (unsigned long long)(prof_clock() - t0)); }
};
#define PROF(cat) Prof prof_scope_##cat(cat)
#else
#define PROF(cat) ((void)0)
#endif

// iterations i = 0 .. cnt-1 spread over the group; NOT followed by a barrier (call g.sync() where a reader follows)
#if defined(THOR_ENC_PARANOID) && !defined(THOR_HOSTEMU)
// diagnostic build: a barrier before and after every cooperative loop (finds a missing g.sync() by making it unnecessary)
struct TforGuard {
  const Grp &g;
  bool once;
  __device__ TforGuard(const Grp &g_) : g(g_), once(true) { g.sync(); }
  __device__ ~TforGuard() { g.sync(); }
};
#define TFOR(g, i, cnt) for (TforGuard tfor_guard_(g); tfor_guard_.once; tfor_guard_.once = false) for (int i = (g).tid; i < (cnt); i += (g).n)
#else
// not unrolled: these loops run one to four times per thread, and the kernel is bound by instruction fetch (536 KB of code, SM
// instruction-cache hit rate 67 %, profiles/r2_ncu_encode_frame_1080p_v3.txt) -- four copies of every loop body cost more than
// the branches they save
#ifdef THOR_HOSTEMU
#define TFOR(g, i, cnt) for (int i = (g).tid; i < (cnt); i += (g).n)
#else
#ifdef THOR_TFOR_UNROLL2
#define TFOR(g, i, cnt) _Pragma("unroll 2") for (int i = (g).tid; i < (cnt); i += (g).n)
#else
#define TFOR(g, i, cnt) _Pragma("unroll 1") for (int i = (g).tid; i < (cnt); i += (g).n)
#endif
#endif
#endif

// ---- small helpers ---------------------------------------------------------------------------------------------
TD int imin(int a, int b) { return a < b ? a : b; }
TD int imax(int a, int b) { return a > b ? a : b; }
TD int iabs(int a) { return a < 0 ? -a : a; }
TD int iclip(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }
TD int clip255(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }
TD int ilog2(uint32_t v) {
#ifdef THOR_HOSTEMU
  return 31 - __builtin_clz(v);
#else
  return 31 - __clz((int)v);
#endif
}

// IEEE double product / sum without contraction: the reference's cost terms are (int)(lambda * bits + 0.5) evaluated
// in double by the host compiler (SSE2, no FMA): enc/encode_block.c:863,1225.
TD double dmul(double a, double b) {
#ifdef THOR_HOSTEMU
  volatile double r = a * b;
  return r;
#else
  return __dmul_rn(a, b);
#endif
}
TD double dadd(double a, double b) {
#ifdef THOR_HOSTEMU
  volatile double r = a + b;
  return r;
#else
  return __dadd_rn(a, b);
#endif
}
TD double ddiv(double a, double b) {
#ifdef THOR_HOSTEMU
  volatile double r = a / b;
  return r;
#else
  return __ddiv_rn(a, b);
#endif
}
TD double dsqrt(double a) {
#ifdef THOR_HOSTEMU
  return sqrt(a);
#else
  return __dsqrt_rn(a);
#endif
}

// ---- memory access by kind -------------------------------------------------------------------------------------
// Frame state written during this frame by other superblocks (reconstruction, neighbour units) is read through L2.
TD uint8_t ld_frame8(const uint8_t *p) {
#ifdef THOR_HOSTEMU
  return *p;
#else
  return __ldcg(p);
#endif
}
TD thor_enc_unit ld_unit(const thor_enc_unit *p) {
#ifdef THOR_HOSTEMU
  return *p;
#else
  const int4 v = __ldcg(reinterpret_cast<const int4 *>(p));
  thor_enc_unit u;
  *reinterpret_cast<int4 *>(&u) = v;
  return u;
#endif
}
TD void st_unit(thor_enc_unit *p, const thor_enc_unit &u) {
#ifdef THOR_HOSTEMU
  *p = u;
#else
  *reinterpret_cast<int4 *>(p) = *reinterpret_cast<const int4 *>(&u);
#endif
}
// Read-only pictures (original, references): cached loads; 4 bytes at any alignment.
TD uint8_t ld_ro8(const uint8_t *p) {
#ifdef THOR_HOSTEMU
  return *p;
#else
  return __ldg(p);
#endif
}
TD uint32_t ld_ro32u(const uint8_t *p) {
#ifdef THOR_HOSTEMU
  uint32_t v; memcpy(&v, p, 4); return v;
#else
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
  const uint32_t lo = __ldg(q);
  const unsigned sh = (unsigned)(a & 3) * 8;
  if (sh == 0) return lo;
  return __funnelshift_r(lo, __ldg(q + 1), sh);
#endif
}

// byte-SIMD on packed pixels
TD uint32_t sad4(uint32_t a, uint32_t b) {
#ifdef THOR_HOSTEMU
  uint32_t s = 0;
  for (int k = 0; k < 4; k++) s += (uint32_t)iabs((int)((a >> (8 * k)) & 255) - (int)((b >> (8 * k)) & 255));
  return s;
#else
  return __vsadu4(a, b);
#endif
}
TD uint32_t havg4(uint32_t a, uint32_t b) {          // (a + b) >> 1 per byte
#ifdef THOR_HOSTEMU
  uint32_t r = 0;
  for (int k = 0; k < 4; k++) r |= ((((a >> (8 * k)) & 255) + ((b >> (8 * k)) & 255)) >> 1) << (8 * k);
  return r;
#else
  return __vhaddu4(a, b);
#endif
}

// Index arithmetic of the cooperative loops.  The GPU has no integer divide: `i / w` by a run-time w is a 20-instruction
// sequence, and every loop body starts with one.  Block and transform sides are powers of two (p2div / p2mod); widths that
// come from a picture edge (B.bwidth) can be 24, 40, 56: Div tests once, outside the loop.
TD int p2div(int i, int d) { return i >> ilog2((uint32_t)d); }
TD int p2mod(int i, int d) { return i & (d - 1); }
struct Div { int d, sh; };
TD Div mkdiv(int d) { Div v; v.d = d; v.sh = (d & (d - 1)) ? -1 : ilog2((uint32_t)d); return v; }
TD int dq(const Div &v, int i) { return v.sh >= 0 ? i >> v.sh : i / v.d; }
TD int dr(const Div &v, int i) { return v.sh >= 0 ? i & (v.d - 1) : i % v.d; }

struct Mv { int16_t x, y; };
TD Mv mkmv(int x, int y) { Mv m; m.x = (int16_t)x; m.y = (int16_t)y; return m; }
TD bool mv_eq(Mv a, Mv b) { return a.x == b.x && a.y == b.y; }

// inter_pred_t (common/types.h:111-118)
struct InterPred { Mv mv0, mv1; int ref_idx0, ref_idx1, dir; };

// ---- constant tables -------------------------------------------------------------------------------------------
// reference common/common_block.c:78-83, :97-98
TCONST uint8_t kChromaQp[52] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25,
                                26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41, 42, 43, 44, 45};
TCONST uint16_t kQuant[6] = {26214, 23302, 20560, 18396, 16384, 14564};
TCONST uint16_t kDequant[6] = {40, 45, 51, 57, 64, 72};
// enc/encode_block.c:2975-2979 (iq_8x8)
TCONST uint16_t kIq8x8[52] = {6, 7, 8, 8, 10, 11, 12, 13, 15, 17, 19, 21, 24, 27, 30, 34, 38, 43, 48, 54, 60, 68, 76, 86, 96, 108,
                              121, 136, 152, 171, 192, 216, 242, 272, 305, 342, 384, 431, 484, 543, 610, 684, 768, 862, 968, 1086,
                              1219, 1368, 1536, 1724, 1935, 2172};
// the 32-point cosine ladder the HEVC-style bases of common/transform.c:41-245 are built from (see build_tables)
TCONST int8_t kCos32[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64, 61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0};
// common/inter_prediction.c:47-70
TCONST int8_t kLumaBi[4][6] = {{0, 0, 64, 0, 0, 0}, {2, -10, 59, 17, -5, 1}, {1, -8, 39, 39, -8, 1}, {1, -5, 17, 59, -10, 2}};
TCONST int8_t kLumaUni[4][6] = {{0, 0, 64, 0, 0, 0}, {1, -7, 55, 19, -5, 1}, {1, -7, 38, 38, -7, 1}, {1, -5, 19, 55, -7, 1}};
TCONST int8_t kChroma[8][4] = {{0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-4, 44, 28, -4}, {-4, 36, 36, -4}, {-4, 28, 44, -4}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

}  // namespace thorenc

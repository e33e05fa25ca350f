// Decode-regime (M <= 16) INT4 weight-only GEMV, second generation: weights stream from HBM straight
// into registers in MMA-fragment order.
//
// Replaces, for small token counts, the reference's AWQLinearMethod.apply = awq_dequantize + matmul
// (python/sglang/srt/layers/quantization/awq.py:197-215; kernel sgl-kernel/csrc/gemm/
// awq_kernel.cu:126-219) and vllm's gptq_gemm (quantization/gptq.py:17-31).
//
// Why a second kernel: w4a16_gemv.cu stages [K, N/8] rows through a shared-memory ring with a CTA
// barrier per 64 k rows and regroups nibbles with PRMT; it measured 9-23 % of the HBM peak
// (profiles/r01_bench_w4a16_crossover.json) — latency-bound on 12 warps per SM.  Here:
//   * one-time repack (process_weights_after_loading) into a STREAM layout: tiles of 16 output
//     columns x 64 k, 512 contiguous bytes, ordered [n / 16][k / 64]; lane l = 4 g + t owns one
//     uint4 = { W(g,0), W(g+8,0), W(g,1), W(g+8,1) } where W(c, s) packs the 8 weights of column
//     n0 + c at k = k0 + 32 s + 8 t + {0,2,4,6,1,3,5,7} (nibble order).  A warp's LDG.128 is one
//     512-byte burst, a warp walks K through contiguous memory, and no shared memory, barrier or
//     shuffle sits between the load and the tensor core.
//   * the WEIGHTS are the A operand of mma.sync.m16n8k16 (16 output columns x 16 k), the tokens are
//     the N = 8 side: `(w >> 4i) & 0x000f000f | MAGIC` is an A register as is (k-adjacent pair of one
//     column, exact integers MAGIC + q).  The contraction index is permuted inside each 32-k step
//     (lane t takes k = 8t..8t+7) and the token fragment is loaded with the same permutation: one
//     LDS.128 of x[token g][8t..8t+7].
//   * f16: the integer (alu) pipe is what bounds nibble extraction at B200's bytes per SM clock, so
//     the odd pairs are taken WITHOUT a shift, `w & 0x00f000f0 | MAGIC` = 1024 + 16 q, and the
//     matching activations (k % 4 in {2, 3}) are staged pre-multiplied by 1/16 (exact): 5 alu
//     instructions per 8 weights instead of 7.  bf16 (7 mantissa bits) keeps the shifts.
//   * zero points and scales are folded per group on fp32 partial sums,
//       out[m,n] = sum_g s[g,n] * ( acc_g[m,n] - MAGIC * sum_{k in g} x'[m,k] - z[g,n] * sum_{k in g} x[m,k] ),
//     acc_g = tensor-core sum of x' times the biased integers; both activation sums are computed
//     once per CTA in the prologue (MMA against constant tiles) and kept in shared memory.
//   * CTA = 8 warps = 128 output columns x one K slice; the slice of x (M rows) and of the scales /
//     biased zeros is staged once in shared memory; split-K over a thread-block cluster, partial
//     tiles reduced in fixed rank order through distributed shared memory (deterministic).
#include <cstdlib>

#include "common.cuh"

namespace b200 {

constexpr int kG2Threads = 256;
constexpr int kG2Depth = 6;  // LDG.128 in flight per lane (3 KiB per warp)

struct Gemv2Params {
  const void* x;
  int64_t ldx;
  const uint4* qw;      // stream layout
  const void* zeros;    // [K/g, N] MAGIC + z
  const void* scales;   // [K/g, N]
  const void* bias;
  void* out;
  int64_t ldc;
  int M, N, K, group;
  int kb_per_rank;      // 64-k blocks per cluster rank (a whole number of groups)
  int xs_stride;        // bytes per staged token row
  int sz_off, sum_off, part_off;  // shared-memory offsets: scale/zero rows, x sums, partial tile
};

template <typename T>
struct G2Magic;
template <>
struct G2Magic<__half> {
  static constexpr uint32_t kBits = 0x64006400u;     // (1024, 1024)
  static constexpr uint32_t kOnes = 0x3c003c00u;     // (1, 1)
  static constexpr uint32_t kSixteen = 0x4c004c00u;  // (16, 16)
  static constexpr uint32_t kInv16 = 0x2c002c00u;    // (1/16, 1/16)
  static constexpr float kMagic = 1024.f;
  static constexpr bool kNoShift = true;
};
template <>
struct G2Magic<__nv_bfloat16> {
  static constexpr uint32_t kBits = 0x43004300u;  // (128, 128)
  static constexpr uint32_t kOnes = 0x3f803f80u;
  static constexpr uint32_t kSixteen = 0x3f803f80u;  // unused (no pre-scaling)
  static constexpr uint32_t kInv16 = 0x3f803f80u;
  static constexpr float kMagic = 128.f;
  static constexpr bool kNoShift = false;
};

template <typename T>
__device__ __forceinline__ void g2_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                       uint32_t a3, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void g2_mma<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                               uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void g2_mma<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1,
                                                      uint32_t a2, uint32_t a3, uint32_t b0,
                                                      uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// (w & mask) | magic in ONE LOP3: both constants live in registers (an immediate form can carry
// only one of them and the compiler then emits two instructions)
__device__ __forceinline__ uint32_t g2_and_or(uint32_t w, uint32_t mask, uint32_t magic) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;\n" : "=r"(r) : "r"(w), "r"(mask), "r"(magic));
  return r;
}
__device__ __forceinline__ uint32_t g2_opaque(uint32_t v) {  // keeps a constant in a register
  uint32_t r;
  asm volatile("mov.u32 %0, %1;\n" : "=r"(r) : "r"(v));
  return r;
}
__device__ __forceinline__ uint4 g2_ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t g2_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t g2_cluster_size() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void g2_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n"
               "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t g2_hmul2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.f16x2 %0, %1, %2;\n" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

// MT = number of 8-token tiles (1: M <= 8, 2: M <= 16).  SPG = 32-k steps per quantisation group
// known at compile time (1, 2, 4 = groups of 32, 64, 128: the fold sits at fixed places of the
// unrolled round, no counters) or 0 = any other group size (run-time step counter).
template <typename T, int MT, int SPG>
__global__ void __launch_bounds__(kG2Threads, MT == 1 ? 3 : 2)
w4a16_gemv2_kernel(const Gemv2Params p) {
  using MG = G2Magic<T>;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const uint32_t rank = g2_cluster_rank(), S = g2_cluster_size();
  const int n_grp = blockIdx.x / S;
  const int KB = p.K >> 6;
  const int kb0 = min(KB, static_cast<int>(rank) * p.kb_per_rank);
  const int kb1 = min(KB, kb0 + p.kb_per_rank);
  const int nkb = kb1 - kb0;
  const int nt = n_grp * 8 + warp;
  const bool warp_ok = nt * 16 < p.N;

  // ---- 1. weights do not depend on the previous kernel: put the first loads in flight now ------
  const uint4* wp = p.qw + (static_cast<int64_t>(nt) * KB + kb0) * 32 + lane;
  uint4 buf[kG2Depth];
#pragma unroll
  for (int i = 0; i < kG2Depth; ++i)
    buf[i] = (warp_ok && i < nkb) ? g2_ldg_stream(wp + i * 32) : make_uint4(0, 0, 0, 0);

  // ---- 2. scales and zero points of this slice's groups as fp32: [group][0 = s, 1 = z][128 columns]
  const int grp_lo = (kb0 * 64) / p.group;
  const int grp_hi = nkb > 0 ? ceil_div(kb1 * 64, p.group) : grp_lo;
  const int ng = grp_hi - grp_lo;
  {
    const T* sc = static_cast<const T*>(p.scales);
    const T* zr = static_cast<const T*>(p.zeros);
    float* szf = reinterpret_cast<float*>(smem + p.sz_off);
    const int n0 = n_grp * 128;
    for (int id = threadIdx.x; id < ng * 32; id += kG2Threads) {  // 8 columns per item
      const int gi = id >> 5, which = (id >> 4) & 1, c8 = (id & 15) * 8;
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = 0.f;
      if (n0 + c8 < p.N) {  // N % 8 == 0: the 16-byte load is all-in or all-out
        const T* src = (which ? zr : sc) + static_cast<int64_t>(grp_lo + gi) * p.N + n0 + c8;
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(src));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack2<T>(w[e]);
          v[2 * e] = which ? f.x - MG::kMagic : f.x;
          v[2 * e + 1] = which ? f.y - MG::kMagic : f.y;
        }
      }
      float4* dst = reinterpret_cast<float4*>(szf + (gi * 2 + which) * 128 + c8);
      dst[0] = make_float4(v[0], v[1], v[2], v[3]);
      dst[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
  griddep_wait();  // x is written by the previous kernel
  // ---- 3. x slice: the M real token rows x nkb * 64 k, then ONE zero row that stands in for every
  //         token >= M of the 8-token MMA tiles.  f16: elements with k % 4 in {2, 3} are stored
  //         times 1/16 (see the header comment).
  const int rows = min(p.M, MT * 8);
  {
    const T* xg = static_cast<const T*>(p.x);
    const int cpr = nkb * 8;  // 16-byte chunks per row
    for (int r = warp; r <= rows; r += 8) {
      const bool real = r < rows;
      const uint4* src = reinterpret_cast<const uint4*>(xg + static_cast<int64_t>(real ? r : 0) * p.ldx +
                                                        kb0 * 64);
      uint8_t* dst = smem + r * p.xs_stride;
      for (int c = lane; c < cpr; c += 32) {
        uint4 v = real ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
        if (MG::kNoShift) {
          v.y = g2_hmul2(v.y, MG::kInv16);
          v.w = g2_hmul2(v.w, MG::kInv16);
        }
        *reinterpret_cast<uint4*>(dst + c * 16) = v;
      }
    }
  }
  __syncthreads();

  // per-thread activation row addresses (token g of tile m, or the zero row) + this lane's 16 bytes
  uint32_t xrow[MT];
#pragma unroll
  for (int m = 0; m < MT; ++m)
    xrow[m] = sbase + ((m * 8 + g) < rows ? (m * 8 + g) : rows) * p.xs_stride + t * 16;

  // ---- 4. per (group, token) activation sums: sums[gl][0][tok] = MAGIC * sum x', [gl][1][tok] = sum x
  const int steps_per_group = SPG ? SPG : (p.group >> 5);
  float* sums = reinterpret_cast<float*>(smem + p.sum_off);
  {
    const uint32_t ones = MG::kOnes, sixteen = MG::kSixteen;
    for (int gl = warp; gl < ng; gl += 8) {
      const int st0 = ((grp_lo + gl) * p.group - kb0 * 64) >> 5;
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        float sp[4] = {0.f, 0.f, 0.f, 0.f}, stt[4] = {0.f, 0.f, 0.f, 0.f};
        for (int s = 0; s < steps_per_group; ++s) {
          const uint4 xb = lds128(xrow[m] + (st0 + s) * 64);
          g2_mma<T>(sp, ones, ones, ones, ones, xb.x, xb.y);
          g2_mma<T>(sp, ones, ones, ones, ones, xb.z, xb.w);
          if (MG::kNoShift) {
            g2_mma<T>(stt, ones, ones, sixteen, sixteen, xb.x, xb.y);
            g2_mma<T>(stt, ones, ones, sixteen, sixteen, xb.z, xb.w);
          }
        }
        if (g == 0) {
          float* dst = sums + (gl * 2) * (MT * 8) + m * 8 + 2 * t;
          dst[0] = MG::kMagic * sp[0];
          dst[1] = MG::kMagic * sp[1];
          dst[MT * 8] = MG::kNoShift ? stt[0] : sp[0];
          dst[MT * 8 + 1] = MG::kNoShift ? stt[1] : sp[1];
        }
      }
    }
  }
  __syncthreads();

  float out[MT][4], acc[MT][4];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int e = 0; e < 4; ++e) out[m][e] = acc[m][e] = 0.f;

  const uint32_t kMagic = g2_opaque(MG::kBits);
  const float* szrow = reinterpret_cast<const float*>(smem + p.sz_off) + warp * 16 + g;  // +256 / group
  const float* sumrow = sums + 2 * t;                                                   // +2*MT*8 / group
  int in_group = 0;  // SPG == 0 only

  // fold of a finished group: out += s * (acc - MAGIC * sum x' - z * sum x)
  auto fold = [&]() {
    const float s_lo = szrow[0], s_hi = szrow[8], z_lo = szrow[128], z_hi = szrow[136];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      const float2 mp = *reinterpret_cast<const float2*>(sumrow + m * 8);
      const float2 sx = *reinterpret_cast<const float2*>(sumrow + MT * 8 + m * 8);
      out[m][0] = fmaf(s_lo, fmaf(-z_lo, sx.x, acc[m][0] - mp.x), out[m][0]);
      out[m][1] = fmaf(s_lo, fmaf(-z_lo, sx.y, acc[m][1] - mp.y), out[m][1]);
      out[m][2] = fmaf(s_hi, fmaf(-z_hi, sx.x, acc[m][2] - mp.x), out[m][2]);
      out[m][3] = fmaf(s_hi, fmaf(-z_hi, sx.y, acc[m][3] - mp.y), out[m][3]);
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[m][e] = 0.f;
    }
    szrow += 256;
    sumrow += 2 * MT * 8;
  };
  // one 32-k step: wa / wb = the words of columns g / g + 8, xoff = byte offset of the step's x
  auto step = [&](uint32_t wa, uint32_t wb, uint32_t xoff, bool fold_here) {
    uint4 xb[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) xb[m] = lds128(xrow[m] + xoff);
    uint32_t a0, a1, a2, a3, c0, c1, c2, c3;
    a0 = g2_and_or(wa, 0x000f000fu, kMagic);
    a1 = g2_and_or(wb, 0x000f000fu, kMagic);
    if (MG::kNoShift) {
      a2 = g2_and_or(wa, 0x00f000f0u, kMagic);
      a3 = g2_and_or(wb, 0x00f000f0u, kMagic);
      const uint32_t wa8 = wa >> 8, wb8 = wb >> 8;
      c0 = g2_and_or(wa8, 0x000f000fu, kMagic);
      c1 = g2_and_or(wb8, 0x000f000fu, kMagic);
      c2 = g2_and_or(wa8, 0x00f000f0u, kMagic);
      c3 = g2_and_or(wb8, 0x00f000f0u, kMagic);
    } else {
      a2 = g2_and_or(wa >> 4, 0x000f000fu, kMagic);
      a3 = g2_and_or(wb >> 4, 0x000f000fu, kMagic);
      c0 = g2_and_or(wa >> 8, 0x000f000fu, kMagic);
      c1 = g2_and_or(wb >> 8, 0x000f000fu, kMagic);
      c2 = g2_and_or(wa >> 12, 0x000f000fu, kMagic);
      c3 = g2_and_or(wb >> 12, 0x000f000fu, kMagic);
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      g2_mma<T>(acc[m], a0, a1, a2, a3, xb[m].x, xb[m].y);
      g2_mma<T>(acc[m], c0, c1, c2, c3, xb[m].z, xb[m].w);
    }
    if (SPG) {
      if (fold_here) fold();
    } else if (++in_group == steps_per_group) {
      in_group = 0;
      fold();
    }
  };
  // tile j of a round (64 k = two steps); with a compile-time SPG a round of 6 tiles = 12 steps starts
  // on a group boundary (12 % SPG == 0), so the fold positions are compile-time too
  auto tile = [&](const uint4& w, int j, uint32_t xround) {
    step(w.x, w.y, xround + j * 128, SPG && ((2 * j + 1) % SPG == 0));
    step(w.z, w.w, xround + j * 128 + 64, SPG && ((2 * j + 2) % SPG == 0));
  };
  static_assert(SPG == 0 || (2 * kG2Depth) % SPG == 0, "a round must hold whole groups");

  if (warp_ok) {
    int base = 0;
    const uint4* wq = wp + kG2Depth * 32;  // the tile kG2Depth ahead of the round's first
    uint32_t xround = 0;
    // steady state: every tile of this round has a successor kG2Depth tiles ahead
#pragma unroll 1
    for (; base + 2 * kG2Depth <= nkb; base += kG2Depth, wq += kG2Depth * 32, xround += kG2Depth * 128) {
#pragma unroll
      for (int j = 0; j < kG2Depth; ++j) {
        const uint4 w = buf[j];
        buf[j] = g2_ldg_stream(wq + j * 32);
        tile(w, j, xround);
      }
    }
    // tail: at most two rounds, loads predicated
#pragma unroll 1
    for (; base < nkb; base += kG2Depth, wq += kG2Depth * 32, xround += kG2Depth * 128) {
#pragma unroll
      for (int j = 0; j < kG2Depth; ++j) {
        if (base + j < nkb) {
          const uint4 w = buf[j];
          if (base + j + kG2Depth < nkb) buf[j] = g2_ldg_stream(wq + j * 32);
          tile(w, j, xround);
        }
      }
    }
  }
  griddep_launch_dependents();

  // ---- split-K reduction in fixed rank order through distributed shared memory ------------------
  float* part = reinterpret_cast<float*>(smem + p.part_off);  // [MT * 8][128]
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int e = 0; e < 4; ++e)
      part[(m * 8 + 2 * t + (e & 1)) * 128 + warp * 16 + g + 8 * (e >> 1)] = out[m][e];
  if (S > 1) g2_cluster_sync();
  else __syncthreads();
  {
    const int cpr = 128 / static_cast<int>(S);  // columns reduced by this rank (S is a power of two)
    const int cshift = 31 - __clz(cpr);
    T* o = static_cast<T*>(p.out);
    const T* bias = static_cast<const T*>(p.bias);
    // kG2Threads is a multiple of cpr: a thread keeps the same column for all of its items, so its
    // bias is loaded once (inside the loop the load would wait behind the previous output store)
    const int c = static_cast<int>(rank) * cpr + (threadIdx.x & (cpr - 1));
    const int n = n_grp * 128 + c;
    const float bias_n = (bias && n < p.N) ? to_float<T>(bias[n]) : 0.f;
    for (int item = threadIdx.x; item < p.M * cpr; item += kG2Threads) {
      const int m = item >> cshift;
      const uint32_t local = sbase + p.part_off + (m * 128 + c) * 4;
      float sum = 0.f;
      if (S == 1) {
        sum = part[m * 128 + c];
      } else {
        for (uint32_t q = 0; q < S; ++q) {
          uint32_t remote;
          float v;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"(q));
          asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(remote) : "memory");
          sum += v;
        }
      }
      if (n < p.N) o[static_cast<int64_t>(m) * p.ldc + n] = from_float<T>(sum + bias_n);
    }
  }
  if (S > 1) g2_cluster_sync();  // peers may still be reading this CTA's partial tile
}

// One thread per output word of the stream layout.
__global__ void w4a16_gemv2_repack_kernel(const uint32_t* __restrict__ qweight,
                                          uint32_t* __restrict__ out, int N, int K) {
  const int64_t o = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int KB = K >> 6;
  const int64_t total = static_cast<int64_t>(N / 16) * KB * 128;
  if (o >= total) return;
  const int j = static_cast<int>(o & 3);
  const int lane = static_cast<int>((o >> 2) & 31);
  const int64_t tile = o >> 7;
  const int kb = static_cast<int>(tile % KB);
  const int nt = static_cast<int>(tile / KB);
  const int g = lane >> 2, t = lane & 3;
  const int col = nt * 16 + g + 8 * (j & 1);
  const int kbase = kb * 64 + (j >> 1) * 32 + 8 * t;
  const int cc = col & 7;
  const int shift = 4 * ((cc & 1) ? 4 + (cc >> 1) : (cc >> 1));  // AWQ nibble position of column cc
  const int NC = N / 8;
  uint32_t w = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = kbase + (i < 4 ? 2 * i : 2 * (i - 4) + 1);
    const uint32_t src = qweight[static_cast<int64_t>(k) * NC + (col >> 3)];
    w |= ((src >> shift) & 0xfu) << (4 * i);
  }
  out[o] = w;
}

static std::atomic<int> g_g2_splits{-1};

static bool gemv2_shape_ok(int M, int N, int K, int group) {
  return M >= 1 && M <= 16 && N % 16 == 0 && K % 64 == 0 && group >= 32 && K % group == 0 &&
         (group == 32 || group % 64 == 0);
}

struct Gemv2Plan {
  int S, kb_per_rank, xs_stride, sz_off, sum_off, part_off, smem;
};

static bool gemv2_plan(int M, int N, int K, int group, Gemv2Plan* pl) {
  if (!gemv2_shape_ok(M, N, K, group)) return false;
  const int MT = M <= 8 ? 1 : 2;
  const int n_groups = ceil_div(N, 128);
  const int upg = group >= 64 ? group / 64 : 1;  // 64-k blocks per slice unit (whole groups)
  const int units = (K / 64) / upg;
  int forced = g_g2_splits.load(std::memory_order_relaxed);
  if (forced < 0) {
    const char* e = std::getenv("B200_W4A16_GEMV2_SPLITS");
    forced = (e && *e) ? std::atoi(e) : 0;
    g_g2_splits.store(forced, std::memory_order_relaxed);
  }
  int S = 1;
  if (forced > 0) {
    S = forced > 8 ? 8 : forced;
  } else {
    // at least two CTAs per SM, at least two slice units per rank
    const int sms = device_sm_count();
    while (n_groups * S < 2 * sms && S < 8 && units >= 4 * S) S *= 2;
  }
  if (S > units) S = units;
  if (S < 1) S = 1;
  while (S & (S - 1)) S &= S - 1;  // the reduce splits 128 columns over the ranks: power of two
  for (;; S *= 2) {
    const int kb_per_rank = ceil_div(units, S) * upg;
    const int xs_stride = kb_per_rank * 128 + 64;  // +64: conflict-free LDS.128 across token rows
    const int ng = ceil_div(kb_per_rank * 64, group) + 1;
    pl->S = S;
    pl->kb_per_rank = kb_per_rank;
    pl->xs_stride = xs_stride;
    pl->sz_off = (M + 1) * xs_stride;  // M token rows + the shared zero row
    pl->sum_off = pl->sz_off + ng * 1024;
    pl->part_off = pl->sum_off + ng * 2 * MT * 8 * 4;
    pl->smem = pl->part_off + MT * 8 * 128 * 4;
    if (pl->smem <= 72 * 1024 || S >= 8) break;  // keep >= 3 CTAs per SM when K allows
  }
  return pl->smem <= 200 * 1024;
}

}  // namespace b200

extern "C" {

/* 1 when b200_w4a16_gemv_v2 accepts the shape (else the caller uses b200_w4a16_gemm). */
int b200_w4a16_gemv_v2_supported(int M, int N, int K, int group_size) {
  b200::Gemv2Plan pl;
  return b200::gemv2_plan(M, N, K, group_size, &pl) ? 1 : 0;
}

/* Test/sweep hook: splits > 0 fixes the split-K cluster size, <= 0 restores the default rule. */
int b200_w4a16_gemv_v2_config(int splits) {
  b200::g_g2_splits.store(splits > 0 ? splits : 0, std::memory_order_relaxed);
  return 0;
}

int b200_w4a16_gemv_v2_repack(const void* qweight, void* out, int K, int N, void* stream) {
  using namespace b200;
  B200_REQUIRE(qweight && out, "b200_w4a16_gemv_v2_repack: null pointer");
  B200_REQUIRE(K > 0 && N > 0 && K % 64 == 0 && N % 16 == 0,
               "b200_w4a16_gemv_v2_repack: needs K %% 64 == 0 and N %% 16 == 0 (K=%d N=%d)", K, N);
  const int64_t total = static_cast<int64_t>(N / 16) * (K / 64) * 128;
  const int threads = 256;
  const int64_t blocks = (total + threads - 1) / threads;
  w4a16_gemv2_repack_kernel<<<static_cast<unsigned>(blocks), threads, 0,
                              static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint32_t*>(qweight), static_cast<uint32_t*>(out), N, K);
  B200_CHECK_LAUNCH("w4a16_gemv2_repack_kernel");
  return 0;
}

int b200_w4a16_gemv_v2(const void* x, int64_t ldx, const void* qweight_stream, const void* zeros,
                       const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                       int K, int group_size, int dtype, void* stream) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(x && qweight_stream && zeros && scales && out, "b200_w4a16_gemv_v2: null pointer");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_w4a16_gemv_v2: dtype %d", dtype);
  Gemv2Plan pl;
  B200_REQUIRE(gemv2_plan(M, N, K, group_size, &pl),
               "b200_w4a16_gemv_v2: needs 1 <= M <= 16, N %% 16 == 0, K %% 64 == 0, group 32 or a "
               "multiple of 64 dividing K, and a K slice that fits shared memory "
               "(M=%d N=%d K=%d group=%d)", M, N, K, group_size);
  B200_REQUIRE(aligned16(qweight_stream) && aligned16(zeros) && aligned16(scales) && aligned16(x) &&
                   ldx % 8 == 0 && N % 8 == 0,
               "b200_w4a16_gemv_v2: operands must be 16-byte aligned, ldx %% 8 == 0");
  Gemv2Params p;
  p.x = x;
  p.ldx = ldx;
  p.qw = static_cast<const uint4*>(qweight_stream);
  p.zeros = zeros;
  p.scales = scales;
  p.bias = bias;
  p.out = out;
  p.ldc = ldc;
  p.M = M;
  p.N = N;
  p.K = K;
  p.group = group_size;
  p.kb_per_rank = pl.kb_per_rank;
  p.xs_stride = pl.xs_stride;
  p.sz_off = pl.sz_off;
  p.sum_off = pl.sum_off;
  p.part_off = pl.part_off;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n_groups = ceil_div(N, 128);
  void (*kern)(const Gemv2Params) = nullptr;
  const int spg = group_size == 128 ? 4 : group_size == 64 ? 2 : group_size == 32 ? 1 : 0;
  const int mt = M <= 8 ? 1 : 2;
#define B200_G2_PICK(T_, MT_)                                                             \
  (spg == 4   ? w4a16_gemv2_kernel<T_, MT_, 4>                                            \
   : spg == 2 ? w4a16_gemv2_kernel<T_, MT_, 2>                                            \
   : spg == 1 ? w4a16_gemv2_kernel<T_, MT_, 1>                                            \
              : w4a16_gemv2_kernel<T_, MT_, 0>)
  if (dtype == B200_F16)
    kern = mt == 1 ? B200_G2_PICK(__half, 1) : B200_G2_PICK(__half, 2);
  else
    kern = mt == 1 ? B200_G2_PICK(__nv_bfloat16, 1) : B200_G2_PICK(__nv_bfloat16, 2);
#undef B200_G2_PICK
  const int ai = (dtype == B200_F16 ? 0 : 8) + (mt - 1) * 4 + (spg == 4 ? 3 : spg);
  static int attr_smem[16] = {};
  if (pl.smem > attr_smem[ai]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.smem));
    attr_smem[ai] = pl.smem;
  }
  note_variant("w4a16_gemv2 mt=%d spg=%d splits=%d tiles=%d", mt, spg, pl.S, n_groups);
  {
    ScopedKernelTimer timer("w4a16_gemv2", st);
    B200_CUDA_OK(launch_kernel(kern, dim3(n_groups * pl.S), dim3(kG2Threads), pl.smem, st, pl.S,
                               pdl_family(M, kPdlGemmSmall), p));
  }
  B200_CHECK_LAUNCH("w4a16_gemv2_kernel");
  return 0;
}

}  // extern "C"

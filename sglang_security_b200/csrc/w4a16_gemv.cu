// Decode-regime (M <= 16) INT4 weight-only GEMV for sm_100a: the HBM-bound half of the W4A16 path.
//
// Replaces, for small token counts, the reference's AWQLinearMethod.apply = awq_dequantize + matmul
// (python/sglang/srt/layers/quantization/awq.py:197-215; kernel sgl-kernel/csrc/gemm/
// awq_kernel.cu:126-219) and vllm's gptq_gemm (quantization/gptq.py:17-31): the weight is streamed
// once at 0.5 byte / element instead of being expanded to a [K, N] 16-bit tensor and re-read.
//
//   out[m, n] = sum_g s[g, n] * ( sum_{k in g} x[m, k] * w[k, n]  -  z[g, n] * sum_{k in g} x[m, k] )
//
// which equals x @ ((w - z) * s) without the reference's per-weight 16-bit rounding of the
// dequantised value (|difference| <= 2^-11 relative per term; inside the GEMM tolerance).  The
// INTEGER nibbles are used exactly: B operand = MAGIC + w as a 16-bit float (exact), the stored
// zeros already hold MAGIC + z, and sum_k x comes from one extra MMA against a tile of ones.
// Per weight element that leaves ~1 ALU instruction (PRMT/LOP3 nibble regrouping) instead of the
// ~2 of a (w - z) * s dequant, which is what lets CUDA-core unpacking keep up with 6.5 TB/s.
//
// Layout (unchanged B200 internal format of w4a16.cu): qweight [K, N/8] uint32, nibble p of a word
// = output column 8c + (p < 4 ? 2p : 2(p-4)+1); zeros / scales [K/g, N] 16-bit.
//
// Work decomposition
//   CTA = 4 warps x 64 columns = 256 output columns, one K slice (split-K over a thread-block
//   cluster, reduced in fixed rank order through distributed shared memory like
//   fp8_gemm_smallm.cuh).  Raw packed rows are staged by 16-byte cp.async into a 6-stage ring of
//   64 k-rows x 128 B (XOR-swizzled: conflict-free word reads).  Per k16 step a lane (t = lane%4,
//   g = lane/4) reads the 4 words of rows {2t, 2t+1, 2t+8, 2t+9} at word column g: nibble position
//   p of those words IS the (b0, b1) fragment of mma.m16n8k16 for "n-tile p" (the 8 columns
//   {8(c+g') + col(p)}), so 8 MMAs consume the 4 words with no shuffles.
#include <cstdlib>

#include "common.cuh"

namespace b200 {

struct GemvParams {
  const void* x;
  int64_t ldx;
  const uint32_t* qweight;  // [K, N/8]
  const void* zeros;        // [K/g, N] MAGIC + z
  const void* scales;       // [K/g, N]
  const void* bias;
  void* out;
  int64_t ldc;
  int M, N, K, group;
  int groups_per_split;
};

template <typename T>
struct GemvMagic;
template <>
struct GemvMagic<__half> {
  static constexpr uint32_t kBits = 0x64006400u;   // (1024, 1024)
  static constexpr uint32_t kOnes = 0x3c003c00u;   // (1, 1)
};
template <>
struct GemvMagic<__nv_bfloat16> {
  static constexpr uint32_t kBits = 0x43004300u;   // (128, 128)
  static constexpr uint32_t kOnes = 0x3f803f80u;
};

template <typename T>
__device__ __forceinline__ void gemv_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1);
template <>
__device__ __forceinline__ void gemv_mma<__half>(float (&c)[4], const uint32_t (&a)[4],
                                                 uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void gemv_mma<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4],
                                                        uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(n));
}
__device__ __forceinline__ uint32_t gemv_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t gemv_cluster_size() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void gemv_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n"
               "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

constexpr int kGvWarps = 4;
constexpr int kGvThreads = kGvWarps * 32;
constexpr int kGvCols = kGvWarps * 64;         // 256 output columns per CTA
constexpr int kGvStageRows = 64;               // k rows per pipeline stage
constexpr int kGvStageBytes = kGvStageRows * (kGvCols / 2);  // 64 rows x 128 B = 8 KiB
constexpr int kGvStages = 6;
constexpr int kGvXStageBytes = 16 * kGvStageRows * 2;        // 16 token rows x 64 k x 16 bit = 2 KiB
constexpr int kGvWRing = kGvStages * kGvStageBytes;          // 48 KiB, also holds the partial tile
constexpr int kGvSmem = kGvWRing + kGvStages * kGvXStageBytes;  // + 12 KiB activation ring
static_assert((16 + 8) * kGvCols * 4 <= kGvWRing, "partial tile must fit the ring");

// column (0..7) of nibble position p inside a word: p < 4 -> 2p, else 2(p-4)+1
__device__ __forceinline__ int col_of_nibble(int p) { return p < 4 ? 2 * p : 2 * (p - 4) + 1; }

template <typename T>
__device__ __forceinline__ float elem16(const uint4& v, int i) {  // i-th 16-bit float of a uint4
  const uint32_t w = i < 2 ? v.x : i < 4 ? v.y : i < 6 ? v.z : v.w;
  const float2 f = unpack2<T>(w);
  return (i & 1) ? f.y : f.x;
}

template <typename T>
__global__ void __launch_bounds__(kGvThreads, 3)
w4a16_gemv_kernel(const GemvParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int t = lane & 3, g = lane >> 2;
  const uint32_t rank = gemv_cluster_rank(), S = gemv_cluster_size();
  const int n_tile = blockIdx.x / S;
  const int NC = p.N / 8;                       // packed words per k row
  const int wc0 = n_tile * (kGvCols / 8);       // first word column of this CTA
  const int groups = p.K / p.group;
  const int g0 = min(groups, static_cast<int>(rank) * p.groups_per_split);
  const int g1 = min(groups, g0 + p.groups_per_split);
  const int k_begin = g0 * p.group, k_end = g1 * p.group;
  const int nstage = ceil_div(k_end - k_begin, kGvStageRows);

  // ---- weight pipeline: 16-byte cp.async of raw packed rows into an XOR-swizzled ring ----------
  // activations: 16 token rows x 64 k of this stage, rows >= M zero-filled, 16-byte chunk c of row r
  // stored at chunk c ^ (r & 7) so that ldmatrix reads are conflict-free
  const T* xg = static_cast<const T*>(p.x);
  auto issue_x = [&](int s) {
    if (s < nstage) {
      const int r = threadIdx.x >> 3, c = threadIdx.x & 7;
      const int k = k_begin + s * kGvStageRows + c * 8;
      const bool ok = r < p.M && k < k_end;
      const T* src = xg + static_cast<int64_t>(ok ? r : 0) * p.ldx + (ok ? k : 0);
      cp_async16_zfill(sbase + kGvWRing + (s % kGvStages) * kGvXStageBytes + r * 128 +
                           ((c ^ (r & 7)) << 4), src, ok);
    }
  };
  auto issue_w = [&](int s) {
    if (s < nstage) {
      const uint32_t dst0 = sbase + (s % kGvStages) * kGvStageBytes;
#pragma unroll
      for (int i = 0; i < (kGvStageRows * 8) / kGvThreads; ++i) {
        const int id = threadIdx.x + i * kGvThreads;
        const int r = id >> 3, c = id & 7;
        const int k = k_begin + s * kGvStageRows + r;
        const bool ok = k < k_end && (wc0 + 4 * c) < NC;
        const uint32_t* src = p.qweight + static_cast<int64_t>(ok ? k : 0) * NC + (ok ? wc0 + 4 * c : 0);
        cp_async16_zfill(dst0 + r * 128 + ((c ^ (((r >> 1) & 3) << 1)) << 4), src, ok);
      }
    }
  };
  // Weights never change during a step: start streaming them before the dependency wait.  The
  // activations of the same stages follow after it, as ONE extra commit group (the first compute
  // step therefore waits for the whole prologue; afterwards a group = W + X of one stage).
#pragma unroll
  for (int s = 0; s < kGvStages - 1; ++s) {
    issue_w(s);
    cp_async_commit();
  }
  griddep_wait();  // x is written by the previous kernel
#pragma unroll
  for (int s = 0; s < kGvStages - 1; ++s) issue_x(s);
  cp_async_commit();
  const T* zeros = static_cast<const T*>(p.zeros);
  const T* scales = static_cast<const T*>(p.scales);
  // the 8-column blocks whose columns this lane's accumulators belong to (C fragment n = 2t, 2t+1)
  const int cb0 = n_tile * kGvCols + 8 * (warp * 8 + 2 * t);
  const bool cb_ok0 = cb0 < p.N, cb_ok1 = cb0 + 8 < p.N;

  float out[8][4], acc[8][4], sx[4];
#pragma unroll
  for (int q = 0; q < 8; ++q)
#pragma unroll
    for (int e = 0; e < 4; ++e) out[q][e] = acc[q][e] = 0.f;
  sx[0] = sx[1] = sx[2] = sx[3] = 0.f;
  uint4 sc0, sc1, zb0, zb1;  // scales / biased zeros of the current group for the two blocks
  auto load_group = [&](int gi) {
    const int64_t base = static_cast<int64_t>(gi) * p.N;
    const uint4 zero4 = make_uint4(0, 0, 0, 0);
    sc0 = cb_ok0 ? __ldg(reinterpret_cast<const uint4*>(scales + base + cb0)) : zero4;
    zb0 = cb_ok0 ? __ldg(reinterpret_cast<const uint4*>(zeros + base + cb0)) : zero4;
    sc1 = cb_ok1 ? __ldg(reinterpret_cast<const uint4*>(scales + base + cb0 + 8)) : zero4;
    zb1 = cb_ok1 ? __ldg(reinterpret_cast<const uint4*>(zeros + base + cb0 + 8)) : zero4;
  };
  if (nstage > 0) load_group(g0);

  constexpr uint32_t kMagic = GemvMagic<T>::kBits;
  const int W = warp * 8 + g;                        // word column inside the CTA tile
  const uint32_t word_off = ((W & 3) << 2);
  const int chunk = W >> 2;
  int k_in_group = 0, gi = g0;

  // ldmatrix.x4 lane address inside an activation stage: matrix j = lane / 8 covers token rows
  // (j & 1) * 8 .. +7 and k chunk 2 kk + (j >> 1)
  const int lm_row = (lane & 7) + ((lane >> 3) & 1) * 8;
  const int lm_chunk = lane >> 4;

  for (int s = 0; s < nstage; ++s) {
    // groups outstanding here: (prologue X, if s == 0) + stages s .. s + STAGES - 2
    if (s == 0) cp_async_wait<0>();
    else cp_async_wait<kGvStages - 2>();
    __syncthreads();                  // stage s landed for everyone; stage s-1 is no longer read
    issue_w(s + kGvStages - 1);       // refills the slot of stage s-1
    issue_x(s + kGvStages - 1);
    cp_async_commit();
    const uint32_t st = sbase + (s % kGvStages) * kGvStageBytes;
    const uint32_t sxs = sbase + kGvWRing + (s % kGvStages) * kGvXStageBytes;
#pragma unroll
    for (int kk = 0; kk < kGvStageRows / 16; ++kk) {
      const int k0 = k_begin + s * kGvStageRows + kk * 16;
      if (k0 >= k_end) break;
      // A fragment (x rows g / g+8, k columns 2t,2t+1 and 2t+8,2t+9) with one ldmatrix.x4
      uint32_t a[4];
      {
        const uint32_t addr = sxs + lm_row * 128 + (((2 * kk + lm_chunk) ^ (lm_row & 7)) << 4);
        asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                     : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3])
                     : "r"(addr));
      }
      // the 4 packed words of rows 2t, 2t+1, 2t+8, 2t+9 at this lane's word column
      uint32_t qw[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = kk * 16 + 2 * t + (i & 1) + (i >> 1) * 8;
        const uint32_t addr = st + r * 128 + ((chunk ^ (((r >> 1) & 3) << 1)) << 4) + word_off;
        asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(qw[i]) : "r"(addr));
      }
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        // byte b of (row 2t | row 2t+1) and of (row 2t+8 | row 2t+9): nibbles 2b and 2b+1
        const uint32_t sel = b | (b << 4) | ((4 + b) << 8) | ((4 + b) << 12);
        const uint32_t t01 = __byte_perm(qw[0], qw[1], sel);
        const uint32_t t23 = __byte_perm(qw[2], qw[3], sel);
        gemv_mma<T>(acc[2 * b], a, (t01 & 0x000f000fu) | kMagic, (t23 & 0x000f000fu) | kMagic);
        gemv_mma<T>(acc[2 * b + 1], a, ((t01 >> 4) & 0x000f000fu) | kMagic,
                    ((t23 >> 4) & 0x000f000fu) | kMagic);
      }
      gemv_mma<T>(sx, a, GemvMagic<T>::kOnes, GemvMagic<T>::kOnes);  // sum_k x[m, k]
      k_in_group += 16;
      if (k_in_group == p.group) {
        // fold the finished group: out += s * (acc - (MAGIC + z) * sum_x)
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int ci = col_of_nibble(q);
          const float s0 = elem16<T>(sc0, ci), s1 = elem16<T>(sc1, ci);
          const float z0 = elem16<T>(zb0, ci), z1 = elem16<T>(zb1, ci);
          out[q][0] = fmaf(s0, fmaf(-z0, sx[0], acc[q][0]), out[q][0]);
          out[q][1] = fmaf(s1, fmaf(-z1, sx[0], acc[q][1]), out[q][1]);
          out[q][2] = fmaf(s0, fmaf(-z0, sx[2], acc[q][2]), out[q][2]);
          out[q][3] = fmaf(s1, fmaf(-z1, sx[2], acc[q][3]), out[q][3]);
          acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.f;
        }
        sx[0] = sx[1] = sx[2] = sx[3] = 0.f;
        k_in_group = 0;
        if (++gi < g1) load_group(gi);
      }
    }
  }
  cp_async_wait<0>();
  griddep_launch_dependents();
  __syncthreads();  // the ring is idle: it becomes the partial-tile exchange buffer

  // ---- split-K reduction in fixed rank order through distributed shared memory ----------------
  // row m of every rank's partial goes to rank m % S, slot [src rank][m / S][256 columns]
  const int tpr = ceil_div(p.M, static_cast<int>(S));
  gemv_cluster_sync();  // barrier 1: every peer's ring is idle
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int m = g + 8 * h;
    if (m < p.M) {
      const uint32_t dst_rank = m % S, slot = m / S;
      const uint32_t local = sbase + ((rank * tpr + slot) * kGvCols) * 4;
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"(dst_rank));
#pragma unroll
      for (int q = 0; q < 8; ++q)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int cl = 8 * (warp * 8 + 2 * t + e) + col_of_nibble(q);
          asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(remote + cl * 4),
                       "f"(out[q][2 * h + e]) : "memory");
        }
    }
  }
  gemv_cluster_sync();  // barrier 2: every push into my buffer has landed
  const float* part = reinterpret_cast<const float*>(smem);
  T* o = static_cast<T*>(p.out);
  const T* bias = static_cast<const T*>(p.bias);
  const int c2 = 2 * threadIdx.x;                 // two adjacent columns per thread
  const int n = n_tile * kGvCols + c2;
  if (n < p.N) {
    float b0 = 0.f, b1 = 0.f;
    if (bias) {
      b0 = to_float<T>(bias[n]);
      b1 = to_float<T>(bias[n + 1]);
    }
    int m = rank;
    for (int j = 0; j < tpr; ++j, m += S) {
      if (m >= p.M) break;
      float v0 = 0.f, v1 = 0.f;
      for (uint32_t q = 0; q < S; ++q) {
        const float2 v = *reinterpret_cast<const float2*>(part + (q * tpr + j) * kGvCols + c2);
        v0 += v.x;
        v1 += v.y;
      }
      T* dst = o + static_cast<int64_t>(m) * p.ldc + n;
      dst[0] = from_float<T>(v0 + b0);
      dst[1] = from_float<T>(v1 + b1);
    }
  }
}

static std::atomic<int> g_gemv_max_m{-1};
static std::atomic<int> g_gemv_splits{-1};

// Largest M the dispatch in b200_w4a16_gemm hands to the GEMV for an N-column weight.  Explicit
// settings (b200_w4a16_gemv_config / B200_W4A16_GEMV_MAX_M) win.  Default, from the cfg-4 sweep
// (profiles/r01_bench_w4a16_crossover.json): the swap-AB split-K tcgen05 kernel is as fast or faster
// from M = 1 on every shape whose tiles can be split over a cluster (N / 128 <= 64); the GEMV keeps
// M <= 2 on the wide gate_up weights, which fill the machine without split-K.
int w4a16_gemv_max_m(int N) {
  int v = g_gemv_max_m.load(std::memory_order_relaxed);
  if (v < 0) {
    static const int env = [] {
      const char* e = std::getenv("B200_W4A16_GEMV_MAX_M");
      return (e && *e) ? std::atoi(e) : -1;
    }();
    v = env >= 0 ? env : (N / 128 > 64 ? 2 : 0);
  }
  return v > 16 ? 16 : v;
}

bool w4a16_gemv_supported(int M, int N, int K, int group) {
  return M >= 1 && M <= 16 && N % 32 == 0 && group >= 16 && group % 16 == 0 && K % group == 0 &&
         K % 8 == 0;
}

}  // namespace b200

extern "C" {

/* Test/sweep hook: max_m = 0 disables the GEMV dispatch inside b200_w4a16_gemm, splits > 0 fixes
 * the split-K cluster size, negative values restore the defaults. */
int b200_w4a16_gemv_config(int max_m, int splits) {
  b200::g_gemv_max_m.store(max_m > 16 ? 16 : max_m, std::memory_order_relaxed);
  b200::g_gemv_splits.store(splits, std::memory_order_relaxed);
  return 0;
}

int b200_w4a16_gemv(const void* x, int64_t ldx, const void* qweight, const void* zeros,
                    const void* scales, const void* bias, void* out, int64_t ldc, int M, int N,
                    int K, int group_size, int dtype, void* stream) {
  using namespace b200;
  if (M == 0 || N == 0) return 0;
  B200_REQUIRE(x && qweight && zeros && scales && out, "b200_w4a16_gemv: null pointer");
  B200_REQUIRE(w4a16_gemv_supported(M, N, K, group_size),
               "b200_w4a16_gemv: needs 1 <= M <= 16, N %% 32 == 0, group %% 16 == 0, K %% group == 0 "
               "(M=%d N=%d K=%d group=%d)", M, N, K, group_size);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_w4a16_gemv: dtype %d", dtype);
  B200_REQUIRE(aligned16(qweight) && aligned16(zeros) && aligned16(scales) &&
                   aligned16(x) && ldx % 8 == 0,
               "b200_w4a16_gemv: operands must be 16-byte aligned, ldx %% 8 == 0");
  GemvParams p;
  p.x = x;
  p.ldx = ldx;
  p.qweight = static_cast<const uint32_t*>(qweight);
  p.zeros = zeros;
  p.scales = scales;
  p.bias = bias;
  p.out = out;
  p.ldc = ldc;
  p.M = M;
  p.N = N;
  p.K = K;
  p.group = group_size;
  const int n_tiles = ceil_div(N, kGvCols);
  const int groups = K / group_size;
  int S = g_gemv_splits.load(std::memory_order_relaxed);
  if (S < 0) {
    const char* e = std::getenv("B200_W4A16_GEMV_SPLITS");
    S = (e && *e) ? std::atoi(e) : 0;
    g_gemv_splits.store(S, std::memory_order_relaxed);
  }
  if (S <= 0) {
    // four 48-KiB CTAs fit an SM: aim for >= 2 CTAs per SM, >= 2 groups per CTA
    const int want = ceil_div(2 * device_sm_count(), n_tiles);
    S = 1;
    while (S < want && S * 2 <= groups / 2 && S < 8) S *= 2;
  }
  if (S > 8) S = 8;
  if (S > groups) S = groups;
  if (S < 1) S = 1;
  p.groups_per_split = ceil_div(groups, S);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto kern = dtype == B200_BF16 ? w4a16_gemv_kernel<__nv_bfloat16> : w4a16_gemv_kernel<__half>;
  static bool attr_done[2] = {false, false};
  const int ai = dtype == B200_BF16 ? 0 : 1;
  if (!attr_done[ai]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGvSmem));
    attr_done[ai] = true;
  }
  note_variant("w4a16_gemv splits=%d tiles=%d", S, n_tiles);
  {
    ScopedKernelTimer timer("w4a16_gemv", st);
    B200_CUDA_OK(launch_kernel(kern, dim3(n_tiles * S), dim3(kGvThreads), kGvSmem, st, S,
                               pdl_family(M, kPdlGemmSmall), p));
  }
  B200_CHECK_LAUNCH("w4a16_gemv_kernel");
  return 0;
}

}  // extern "C"

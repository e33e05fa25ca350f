// Paged split-KV decode attention for sm_100a.
//
// Path replaced: AttentionBackend.forward_decode (numerical oracle
// torch_native_backend.py:112-180; GPU reference triton_ops/decode_attention.py:50-557).
//
// Work decomposition
//   grid = (num_splits, Hkv * head_tiles, bs).  One CTA owns one (request, kv head, KV split) and
//   ALL query heads of that kv head's GQA group (up to 16 per head tile), so every K/V row is
//   fetched from HBM exactly once per group.  The KV pool is token-granular (page_size 1): row
//   `slot = req_to_token[req, t]` of k_cache/v_cache, D contiguous elements per kv head.
//   The kernel reads req_to_token directly — no flat kv_indices round trip.
//
// Data movement (the HBM-bound part)
//   Each warp runs its own STAGES-deep cp.async ring (16-byte LDGSTS, L2-only) over NT-token
//   tiles of gathered K and V rows; slot ids are fetched one tile ahead with 16-byte loads.
//   No CTA-wide barrier in the main loop.
//
// Math
//   Q.K^T and P.V of the whole GQA group on mma.sync m16n8k16 (rows = query heads), fp32
//   accumulators and fp32 online softmax in registers (exp2 domain, scale*log2e folded in).
//   Warps are merged once through shared memory; partials (normalised O, log2-domain LSE) go to
//   the workspace and are merged by attn_merge_kernel.
#include <cstdlib>

#include "common.cuh"

namespace b200 {

struct DecodeParams {
  const void* q;
  int64_t q_stride;
  const void* k_cache;
  const void* v_cache;
  int64_t kv_row_stride;
  void* o;
  int64_t o_stride;
  const int32_t* req_to_token;
  int64_t rtt_stride;
  const int64_t* req_pool_indices;
  const int64_t* seq_lens;
  float* part_o;
  float* part_lse;
  int bs, Hq, Hkv, group, head_tiles;
  int num_splits;
  float qk_scale;   // scale*log2e, or plain scale when logit_cap > 0
  float logit_cap;  // <= 0: disabled
  float v_scale;    // FP8 KV pool only: output multiplier (k_scale is folded into qk_scale)
  int pdl_late;
  // Fused tail (b200_attn_decode*_fused, see decode_fused_tail): split partials are merged and,
  // optionally, the output row is quantised to e4m3 per token inside this kernel — the merge kernel
  // and the o_proj input quantiser disappear from the step.
  unsigned* tickets;     // [bs * (1 + gridDim.y)], zero before the first launch, self-resetting
  uint8_t* q_out;        // [bs, q_out_stride] e4m3 or null
  int64_t q_out_stride;
  float* q_scales;       // [bs]
};

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kNegBig = -1.0e30f;

// Tokens per split for a request of `len` tokens: equal parts rounded up to `align` tokens so
// that decode and merge agree on which splits are non-empty.
__host__ __device__ inline int split_chunk(int len, int num_splits, int align) {
  int c = ceil_div(len, num_splits);
  c = ceil_div(c, align) * align;
  return c < align ? align : c;
}

// ------------------------------------------------------------------------------------------------
// The dot products of a GQA group run on mma.sync m16n8k16 (rows = query heads of the group,
// padded to 16), so the per-token instruction count does not scale with the group size and the
// SM front end stays far below the HBM-bound gather stream (a CUDA-core FFMA version of this
// kernel was issue-bound at 78 % of the HBM peak; profiles/r01_decode_tuning.md).
//
//   S[16 heads, NT tokens] = Q[16, D] . K^T   (K rows from swizzled smem via ldmatrix)
//   P = exp2(S*scale - m)  (fp32 registers; the accumulator layout of S is the A layout of P)
//   O[16, D] += P[16, NT] . V[NT, D]          (V via ldmatrix.trans)
//
// Each warp still owns its private STAGES-deep cp.async ring of NT-token tiles.  Every lane copies
// column (lane % CPR) of RPG consecutive rows, so its slot ids are RPG consecutive int32 of
// req_to_token: fetched one tile ahead with 16-byte loads, no shuffles in the issue loop.
// 16-byte chunk c of row r is stored at chunk (c ^ (r & 7)) => ldmatrix is bank-conflict free.
template <int D, int NT, int STAGES, int NW>
struct MmaCfg {
  static constexpr int CPR = D / 8;            // 16-byte chunks per row
  static constexpr int RG = 32 / CPR;          // row groups copied concurrently by a warp
  static constexpr int RPG = NT / RG;          // consecutive rows per lane
  static constexpr int ROW_BYTES = D * 2;
  static constexpr int STAGE_BYTES = NT * ROW_BYTES * 2;  // K rows then V rows
  static constexpr int THREADS = NW * 32;
  static constexpr int ALIGN = NW * NT;
  static constexpr int RING_BYTES = NW * STAGES * STAGE_BYTES;
  static constexpr int FINAL_BYTES = NW * 16 * (D + 2) * 4;
  static constexpr int SMEM_BYTES = RING_BYTES > FINAL_BYTES ? RING_BYTES : FINAL_BYTES;
  static_assert(RPG % 4 == 0, "slot ids are fetched as int4");
  static_assert(NT % 16 == 0, "PV k-step is 16 tokens");
};

template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4],
                                                        uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__half>(float (&c)[4], const uint32_t (&a)[4],
                                                 uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}


// ---- fused tail: merge of the split partials (+ per-token FP8 quantisation) inside the kernel ----
// Two levels of ticket counters (threadFenceReduction pattern; bar.sync + one gpu-scope acq_rel
// atomic by thread 0 orders the CTA's stores before the draw and the winner's loads after it):
//   level 1, one counter per (request, kv-head tile): the LAST of the nvalid split CTAs of that tile
//            merges the partials of its own query heads — same arithmetic as attn_merge_kernel
//            (num * 1/den, one rounding to T) — and writes them to the 16-bit output row;
//   level 2, one counter per request (only when q_out is requested): the last tile to finish
//            quantises the whole row exactly as fp8_quant_per_token_kernel would (scale = amax / 448,
//            q = sat(x * (1/scale)) on the rounded 16-bit values).
// The merge work is spread over all (request, tile) pairs instead of one CTA per request; counters
// re-arm themselves, so CUDA-graph replays need no memset.
// Ticket words: [0, bs) level 2, bs + b * gridDim.y + blockIdx.y level 1.
struct TailDev {
  unsigned* tickets;
  uint8_t* q_out;
  int64_t q_out_stride;
  float* q_scales;
  void* o;
  int64_t o_stride;
  const float* part_o;
  const float* part_lse;
  int Hq, num_splits, bs;
};
__device__ __forceinline__ TailDev tail_of(const DecodeParams& p) {
  TailDev t;
  t.tickets = p.tickets;
  t.q_out = p.q_out;
  t.q_out_stride = p.q_out_stride;
  t.q_scales = p.q_scales;
  t.o = p.o;
  t.o_stride = p.o_stride;
  t.part_o = p.part_o;
  t.part_lse = p.part_lse;
  t.Hq = p.Hq;
  t.num_splits = p.num_splits;
  t.bs = p.bs;
  return t;
}

// true on every thread of the CTA that drew the last of `total` tickets (and re-arms the counter)
__device__ __forceinline__ bool draw_last_ticket(unsigned* ticket, unsigned total, unsigned* s_flag) {
  __syncthreads();  // every thread's stores are ordered before thread 0's release
  if (threadIdx.x == 0) {
    unsigned prev;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;\n"
                 : "=r"(prev) : "l"(ticket) : "memory");
    const bool last = prev + 1 == total;
    if (last) *ticket = 0u;  // every other participant has drawn: safe to re-arm
    *s_flag = last ? 1u : 0u;
  }
  __syncthreads();
  return *s_flag != 0u;
}

template <typename T, int THREADS>
__device__ __noinline__ void decode_fused_tail(const TailDev p, int b, int hq0, int heads_here,
                                               int nvalid, int D) {
  __shared__ unsigned s_flag;
  __shared__ float s_red[THREADS / 32];
  const int NS = p.num_splits;
  T* orow = static_cast<T*>(p.o) + static_cast<int64_t>(b) * p.o_stride;
  if (NS > 1) {
    if (!draw_last_ticket(p.tickets + p.bs + b * gridDim.y + blockIdx.y,
                          static_cast<unsigned>(nvalid), &s_flag))
      return;
    for (int it = threadIdx.x; it < heads_here * D / 4; it += THREADS) {
      const int e0 = it * 4, h = e0 / D, d = e0 - h * D;
      const int hq = hq0 + h;
      const int64_t base = (static_cast<int64_t>(b) * p.Hq + hq) * NS;
      float M = -INFINITY;
      for (int s = 0; s < nvalid; ++s) M = fmaxf(M, __ldcg(p.part_lse + base + s));
      float den = 0.f, num[4] = {0.f, 0.f, 0.f, 0.f};
      for (int s = 0; s < nvalid; ++s) {
        const float w = fast_exp2(__ldcg(p.part_lse + base + s) - M);
        const float4 x = __ldcg(reinterpret_cast<const float4*>(p.part_o + (base + s) * D + d));
        den += w;
        num[0] += w * x.x;
        num[1] += w * x.y;
        num[2] += w * x.z;
        num[3] += w * x.w;
      }
      const float inv = 1.f / den;
      uint2 pk;
      pk.x = pack2<T>(num[0] * inv, num[1] * inv);
      pk.y = pack2<T>(num[2] * inv, num[3] * inv);
      *reinterpret_cast<uint2*>(orow + hq * D + d) = pk;
    }
  }
  if (!p.q_out) return;
  if (!draw_last_ticket(p.tickets + b, gridDim.y, &s_flag)) return;
  // the whole 16-bit row of request b is final: per-token quantisation (two passes over 8 KiB of L2)
  const int n8 = p.Hq * D / 8;
  float amax = 0.f;
  for (int it = threadIdx.x; it < n8; it += THREADS) {
    const uint4 pk = __ldcg(reinterpret_cast<const uint4*>(orow) + it);
    const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 v = unpack2<T>(w[e]);
      amax = fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y)));
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = s_red[0];
#pragma unroll
  for (int w = 1; w < THREADS / 32; ++w) amax = fmaxf(amax, s_red[w]);
  const float scale = amax / 448.0f;
  if (threadIdx.x == 0) p.q_scales[b] = scale;
  const float inv = scale > 0.f ? 1.0f / scale : 0.f;
  uint8_t* qrow = p.q_out + static_cast<int64_t>(b) * p.q_out_stride;
  for (int it = threadIdx.x; it < n8; it += THREADS) {
    const uint4 pk = __ldcg(reinterpret_cast<const uint4*>(orow) + it);
    const uint32_t w[4] = {pk.x, pk.y, pk.z, pk.w};
    uint32_t q[2] = {0u, 0u};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 v = unpack2<T>(w[e]);
      const float x0 = fmaxf(fminf(v.x * inv, 448.0f), -448.0f);
      const float x1 = fmaxf(fminf(v.y * inv, 448.0f), -448.0f);
      uint16_t r;
      asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(x1), "f"(x0));
      q[e >> 1] |= static_cast<uint32_t>(r) << (16 * (e & 1));
    }
    *reinterpret_cast<uint2*>(qrow + it * 8) = make_uint2(q[0], q[1]);
  }
}

}  // namespace b200
#include "attn_decode_fp8.cuh"
namespace b200 {

// HI: the tile holds more than 8 heads (rows 8..15 of the MMA are live).
template <typename T, int D, int NT, int STAGES, int NW, bool HI, bool TAIL>
__global__ void __launch_bounds__(NW * 32, TAIL ? 4 : 0)
attn_decode_mma_kernel(const DecodeParams p) {
  using C = MmaCfg<D, NT, STAGES, NW>;
  constexpr int CPR = C::CPR, RPG = C::RPG, ROW_BYTES = C::ROW_BYTES;
  constexpr int STAGE_BYTES = C::STAGE_BYTES;
  constexpr int KS = D / 16;   // k-steps of Q.K^T
  constexpr int NS = NT / 8;   // n-tiles (8 tokens) of S
  constexpr int NO = D / 8;    // n-tiles (8 dims) of O
  constexpr int GT = HI ? 16 : 8;

  extern __shared__ __align__(128) uint8_t smem[];
  griddep_wait();  // q, the new KV rows and seq_lens are written by the kernels before this one
  if (!p.pdl_late) griddep_launch_dependents();

  const int split = blockIdx.x;
  const int kvh = blockIdx.y / p.head_tiles;
  const int ht = blockIdx.y - kvh * p.head_tiles;
  const int b = blockIdx.z;
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const int len = static_cast<int>(p.seq_lens[b]);
  const int chunk = split_chunk(len, p.num_splits, C::ALIGN);
  const int start = split * chunk;
  if (start >= len) return;
  const int end = min(len, start + chunk);
  const int ntiles = ceil_div(end - start, NT);
  const int n_my = warp < ntiles ? (ntiles - warp + NW - 1) / NW : 0;

  const int32_t* __restrict__ rtt = p.req_to_token + p.req_pool_indices[b] * p.rtt_stride;
  const int64_t row_stride_bytes = p.kv_row_stride * 2;
  const int head0 = ht * GT;
  const int heads_here = min(GT, p.group - head0);

  // ---- Q fragments (A operand, rows = heads) -------------------------------------------------
  uint32_t qa[KS][4];
  {
    const int r_lo = lane >> 2, r_hi = r_lo + 8;
    const T* qb = static_cast<const T*>(p.q) + static_cast<int64_t>(b) * p.q_stride +
                  static_cast<int64_t>(kvh * p.group + head0) * D + (lane & 3) * 2;
    const bool v_lo = r_lo < heads_here, v_hi = HI && r_hi < heads_here;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qa[ks][0] = v_lo ? *reinterpret_cast<const uint32_t*>(qb + r_lo * D + ks * 16) : 0u;
      qa[ks][2] = v_lo ? *reinterpret_cast<const uint32_t*>(qb + r_lo * D + ks * 16 + 8) : 0u;
      qa[ks][1] = v_hi ? *reinterpret_cast<const uint32_t*>(qb + r_hi * D + ks * 16) : 0u;
      qa[ks][3] = v_hi ? *reinterpret_cast<const uint32_t*>(qb + r_hi * D + ks * 16 + 8) : 0u;
    }
  }

  float o[NO][4];
#pragma unroll
  for (int i = 0; i < NO; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_lo = kNegBig, m_hi = kNegBig, l_lo = 0.f, l_hi = 0.f;

  // ---- copy-issue geometry ---------------------------------------------------------------------
  const int ccol = lane % CPR;         // 16-byte column this lane copies
  const int crow0 = (lane / CPR) * RPG;  // first of its RPG consecutive rows
  const char* kcol = static_cast<const char*>(p.k_cache) + static_cast<int64_t>(kvh) * ROW_BYTES +
                     ccol * 16;
  const char* vcol = static_cast<const char*>(p.v_cache) + static_cast<int64_t>(kvh) * ROW_BYTES +
                     ccol * 16;
  const uint32_t ring = smem_u32(smem) + warp * (STAGES * STAGE_BYTES);
  const bool rtt_vec = (p.rtt_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(p.req_to_token) & 15) == 0;

  int slot[RPG];
  auto load_slots = [&](int k) {
    if (k >= n_my) return;
    const int pos = start + (warp + k * NW) * NT + crow0;
    if (rtt_vec && pos + RPG <= end) {
#pragma unroll
      for (int i = 0; i < RPG / 4; ++i) {
        const int4 v = __ldg(reinterpret_cast<const int4*>(rtt + pos) + i);
        slot[4 * i + 0] = v.x; slot[4 * i + 1] = v.y; slot[4 * i + 2] = v.z; slot[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < RPG; ++i) slot[i] = pos + i < end ? __ldg(rtt + pos + i) : 0;
    }
  };
  auto issue_tile = [&](int k) {
    if (k < n_my) {
      const uint32_t sk = ring + (k % STAGES) * STAGE_BYTES;
#pragma unroll
      for (int i = 0; i < RPG; ++i) {
        const int row = crow0 + i;
        const int64_t off = static_cast<int64_t>(slot[i]) * row_stride_bytes;
        const uint32_t dst = sk + row * ROW_BYTES + ((ccol ^ (row & 7)) << 4);
        cp_async16(dst, kcol + off);
        cp_async16(dst + NT * ROW_BYTES, vcol + off);
      }
    }
    cp_async_commit();
  };

  load_slots(0);
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    issue_tile(s);
    load_slots(s + 1);
  }

  const bool capped = p.logit_cap > 0.f;
  const float inv_cap = capped ? p.qk_scale / p.logit_cap : 0.f;  // qk_scale = plain scale if capped
  const float cap_log2 = p.logit_cap * kLog2e;
  const float sc_mul = capped ? 1.f : p.qk_scale;

  // ldmatrix lane geometry
  const int lx = lane & 7;
  const uint32_t k_row_off = lx * ROW_BYTES;                 // + nt * 8 * ROW_BYTES
  const int k_y = (lane >> 3) ^ lx;                          // chunk xor term (QK)
  const uint32_t v_row_off = (((lane >> 3) & 1) * 8 + lx) * ROW_BYTES;  // + kt * 16 * ROW_BYTES
  const int v_y = (lane >> 4) ^ lx;

  for (int it = 0; it < n_my; ++it) {
    issue_tile(it + STAGES - 1);
    load_slots(it + STAGES);
    cp_async_wait<STAGES - 1>();
    __syncwarp();

    const uint32_t sk = ring + (it % STAGES) * STAGE_BYTES;
    const uint32_t sv = sk + NT * ROW_BYTES;
    const int tok0 = start + (warp + it * NW) * NT;

    // ---- S = Q . K^T ---------------------------------------------------------------------------
    float s[NS][4];
#pragma unroll
    for (int nt = 0; nt < NS; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ks += 2) {
        uint32_t kb[4];
        ldsm_x4(kb, sk + nt * 8 * ROW_BYTES + k_row_off + (((2 * ks) ^ k_y) << 4));
        mma16816<T>(s[nt], qa[ks], kb[0], kb[1]);
        mma16816<T>(s[nt], qa[ks + 1], kb[2], kb[3]);
      }
    }
    if (capped) {
#pragma unroll
      for (int nt = 0; nt < NS; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[nt][e] = cap_log2 * fast_tanh(s[nt][e] * inv_cap);
    }
    if (tok0 + NT > end) {  // tail tile: mask tokens past the end (warp-uniform branch)
#pragma unroll
      for (int nt = 0; nt < NS; ++nt) {
        const int t = tok0 + nt * 8 + (lane & 3) * 2;
        if (t >= end) s[nt][0] = s[nt][2] = -INFINITY;
        if (t + 1 >= end) s[nt][1] = s[nt][3] = -INFINITY;
      }
    }

    // ---- online softmax ----------------------------------------------------------------------
    float mt_lo = fmaxf(s[0][0], s[0][1]), mt_hi = fmaxf(s[0][2], s[0][3]);
#pragma unroll
    for (int nt = 1; nt < NS; ++nt) {
      mt_lo = fmaxf(mt_lo, fmaxf(s[nt][0], s[nt][1]));
      if (HI) mt_hi = fmaxf(mt_hi, fmaxf(s[nt][2], s[nt][3]));
    }
    mt_lo = fmaxf(mt_lo, __shfl_xor_sync(0xffffffffu, mt_lo, 1));
    mt_lo = fmaxf(mt_lo, __shfl_xor_sync(0xffffffffu, mt_lo, 2));
    const float mn_lo = fmaxf(m_lo, mt_lo * sc_mul);
    float mn_hi = m_hi;
    if (HI) {
      mt_hi = fmaxf(mt_hi, __shfl_xor_sync(0xffffffffu, mt_hi, 1));
      mt_hi = fmaxf(mt_hi, __shfl_xor_sync(0xffffffffu, mt_hi, 2));
      mn_hi = fmaxf(m_hi, mt_hi * sc_mul);
    }
    if (__any_sync(0xffffffffu, mn_lo > m_lo || (HI && mn_hi > m_hi))) {
      const float a_lo = fast_exp2(m_lo - mn_lo);
      const float a_hi = HI ? fast_exp2(m_hi - mn_hi) : 1.f;
      l_lo *= a_lo;
      l_hi *= a_hi;
#pragma unroll
      for (int i = 0; i < NO; ++i) {
        o[i][0] *= a_lo;
        o[i][1] *= a_lo;
        if (HI) {
          o[i][2] *= a_hi;
          o[i][3] *= a_hi;
        }
      }
      m_lo = mn_lo;
      m_hi = mn_hi;
    }
    uint32_t pa[NT / 16][4];
#pragma unroll
    for (int nt = 0; nt < NS; ++nt) {
      const float p0 = fast_exp2(fmaf(s[nt][0], sc_mul, -m_lo));
      const float p1 = fast_exp2(fmaf(s[nt][1], sc_mul, -m_lo));
      l_lo += p0 + p1;
      pa[nt >> 1][(nt & 1) * 2] = pack2<T>(p0, p1);
      if (HI) {
        const float p2 = fast_exp2(fmaf(s[nt][2], sc_mul, -m_hi));
        const float p3 = fast_exp2(fmaf(s[nt][3], sc_mul, -m_hi));
        l_hi += p2 + p3;
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack2<T>(p2, p3);
      } else {
        pa[nt >> 1][(nt & 1) * 2 + 1] = 0u;
      }
    }

    // ---- O += P . V --------------------------------------------------------------------------
#pragma unroll
    for (int kt = 0; kt < NT / 16; ++kt) {
#pragma unroll
      for (int np = 0; np < NO / 2; ++np) {
        uint32_t vb[4];
        ldsm_x4_t(vb, sv + kt * 16 * ROW_BYTES + v_row_off + (((2 * np) ^ v_y) << 4));
        mma16816<T>(o[2 * np], pa[kt], vb[0], vb[1]);
        mma16816<T>(o[2 * np + 1], pa[kt], vb[2], vb[3]);
      }
    }
    __syncwarp();  // stage may be refilled by the next iteration's issue
  }
  cp_async_wait<0>();
  if (p.pdl_late) griddep_launch_dependents();

  // ---- quad-reduce the row sums, merge warps through shared memory ---------------------------
  l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 1);
  l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 2);
  if (HI) {
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 1);
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 2);
  }
  __syncthreads();
  float* fin = reinterpret_cast<float*>(smem);   // [NW][16][D]
  float* fin_m = fin + NW * 16 * D;               // [NW][16]
  float* fin_l = fin_m + NW * 16;                 // [NW][16]
  {
    const int r_lo = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
    for (int i = 0; i < NO; ++i) {
      *reinterpret_cast<float2*>(fin + (warp * 16 + r_lo) * D + i * 8 + c0) =
          make_float2(o[i][0], o[i][1]);
      if (HI)
        *reinterpret_cast<float2*>(fin + (warp * 16 + r_lo + 8) * D + i * 8 + c0) =
            make_float2(o[i][2], o[i][3]);
    }
    if ((lane & 3) == 0) {
      fin_m[warp * 16 + r_lo] = m_lo;
      fin_l[warp * 16 + r_lo] = l_lo;
      if (HI) {
        fin_m[warp * 16 + r_lo + 8] = m_hi;
        fin_l[warp * 16 + r_lo + 8] = l_hi;
      }
    }
  }
  __syncthreads();

  for (int idx = threadIdx.x; idx < heads_here * D; idx += NW * 32) {
    const int h = idx / D, d = idx - h * D;
    float M = fin_m[h];
#pragma unroll
    for (int w = 1; w < NW; ++w) M = fmaxf(M, fin_m[w * 16 + h]);
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float f = fast_exp2(fin_m[w * 16 + h] - M);  // idle warps: m = -1e30, l = 0 -> f = 0
      num += fin[(w * 16 + h) * D + d] * f;
      den += fin_l[w * 16 + h] * f;
    }
    const float out = num / den;
    const int hq = kvh * p.group + head0 + h;
    if (p.num_splits == 1) {
      static_cast<T*>(p.o)[static_cast<int64_t>(b) * p.o_stride + hq * D + d] = from_float<T>(out);
    } else {
      const int64_t pidx = (static_cast<int64_t>(b) * p.Hq + hq) * p.num_splits + split;
      p.part_o[pidx * D + d] = out;
      if (d == 0) p.part_lse[pidx] = M + log2f(den);
    }
  }
  if constexpr (TAIL)
    decode_fused_tail<T, NW * 32>(tail_of(p), b, kvh * p.group + head0, heads_here, ceil_div(len, chunk), D);
}

// Merge of split partials: one CTA per (query head, request).
template <typename T>
__global__ void attn_merge_kernel(const float* __restrict__ part_o,
                                  const float* __restrict__ part_lse, T* __restrict__ o,
                                  int64_t o_stride, const int64_t* __restrict__ seq_lens, int Hq,
                                  int Dv, int num_splits, int align) {
  griddep_wait();
  griddep_launch_dependents();
  const int hq = blockIdx.x, b = blockIdx.y;
  const int len = static_cast<int>(seq_lens[b]);
  if (len <= 0) return;
  const int chunk = split_chunk(len, num_splits, align);
  const int nvalid = ceil_div(len, chunk);
  const int64_t base = (static_cast<int64_t>(b) * Hq + hq) * num_splits;
  float M = -INFINITY;
  for (int s = 0; s < nvalid; ++s) M = fmaxf(M, __ldg(part_lse + base + s));
  float den = 0.f;
  for (int s = 0; s < nvalid; ++s) den += fast_exp2(__ldg(part_lse + base + s) - M);
  const float inv = 1.f / den;
  for (int d = threadIdx.x; d < Dv; d += blockDim.x) {
    float num = 0.f;
    for (int s = 0; s < nvalid; ++s)
      num += fast_exp2(__ldg(part_lse + base + s) - M) * __ldg(part_o + (base + s) * Dv + d);
    o[static_cast<int64_t>(b) * o_stride + hq * Dv + d] = from_float<T>(num * inv);
  }
}

// Merge of split partials for ALL heads of a request + per-token FP8 quantisation of the merged
// row in one launch.  Replaces attn_merge_kernel followed by fp8_quant_per_token_kernel (same
// arithmetic in the same order, so the bytes are identical); with num_splits == 1 it only quantises
// the row the split kernel wrote.  A request is spread over a CLUSTER of S CTAs (128 threads, one
// 4-element item per thread and pass), the per-token amax is combined through distributed shared
// memory (a maximum, so the result does not depend on the split), every value stays in registers
// between the amax and the quantising pass.
constexpr int kMqThreads = 128;
constexpr int kMqMaxItems = 4;  // 4-element items per thread: covers Hq * Dv <= 8 * 128 * 4 * 4

template <typename T>
__global__ void __launch_bounds__(kMqThreads)
attn_merge_quant_kernel(const float* __restrict__ part_o, const float* __restrict__ part_lse,
                        T* __restrict__ o, int64_t o_stride, uint8_t* __restrict__ q_out,
                        int64_t q_stride, float* __restrict__ q_scales,
                        const int64_t* __restrict__ seq_lens, int Hq, int Dv, int num_splits,
                        int align, int per_cta) {
  __shared__ float red[kMqThreads / 32];
  __shared__ float cl_slots[8];
  uint32_t S, rank;
  asm volatile("mov.u32 %0, %%cluster_nctarank;\n" : "=r"(S));
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(rank));
  if (S > 1)  // every CTA of the cluster runs before any remote store is issued
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
  griddep_wait();
  griddep_launch_dependents();
  const int b = blockIdx.x / S;
  const int len = static_cast<int>(seq_lens[b]);
  T* orow = o + static_cast<int64_t>(b) * o_stride;
  const int n4 = Hq * Dv / 4;
  const int i0 = static_cast<int>(rank) * per_cta, i1 = min(n4, i0 + per_cta);
  const int chunk = split_chunk(max(len, 1), num_splits, align);
  const int nvalid = len > 0 ? ceil_div(len, chunk) : 0;
  uint2 vals[kMqMaxItems];
  float amax = 0.f;
#pragma unroll
  for (int k = 0; k < kMqMaxItems; ++k) {
    const int it = i0 + threadIdx.x + k * kMqThreads;
    vals[k] = make_uint2(0u, 0u);
    if (it < i1 && len > 0) {
      const int e0 = it * 4, h = e0 / Dv, d = e0 - h * Dv;
      if (num_splits > 1) {
        const int64_t base = (static_cast<int64_t>(b) * Hq + h) * num_splits;
        float M = -INFINITY;
        for (int s = 0; s < nvalid; ++s) M = fmaxf(M, __ldg(part_lse + base + s));
        float den = 0.f, num[4] = {0.f, 0.f, 0.f, 0.f};
        for (int s = 0; s < nvalid; ++s) {
          const float w = fast_exp2(__ldg(part_lse + base + s) - M);
          const float4 x = __ldg(reinterpret_cast<const float4*>(part_o + (base + s) * Dv + d));
          den += w;
          num[0] += w * x.x;
          num[1] += w * x.y;
          num[2] += w * x.z;
          num[3] += w * x.w;
        }
        const float inv = 1.f / den;
        vals[k].x = pack2<T>(num[0] * inv, num[1] * inv);
        vals[k].y = pack2<T>(num[2] * inv, num[3] * inv);
        *reinterpret_cast<uint2*>(orow + e0) = vals[k];
      } else {
        vals[k] = *reinterpret_cast<const uint2*>(orow + e0);
      }
      const float2 a = unpack2<T>(vals[k].x), c = unpack2<T>(vals[k].y);
      amax = fmaxf(fmaxf(amax, fmaxf(fabsf(a.x), fabsf(a.y))), fmaxf(fabsf(c.x), fabsf(c.y)));
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = red[0];
#pragma unroll
  for (int w = 1; w < kMqThreads / 32; ++w) amax = fmaxf(amax, red[w]);
  if (S > 1) {
    if (threadIdx.x < S) {  // thread j publishes this CTA's amax into CTA j's slot [rank]
      uint32_t local = smem_u32(cl_slots + rank), remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"(threadIdx.x));
      asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(remote), "f"(amax) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    amax = cl_slots[0];
    for (uint32_t j = 1; j < S; ++j) amax = fmaxf(amax, cl_slots[j]);
  }
  if (len <= 0) return;
  const float scale = amax / 448.0f;
  if (threadIdx.x == 0 && rank == 0) q_scales[b] = scale;
  const float inv = scale > 0.f ? 1.0f / scale : 0.f;
  uint8_t* qrow = q_out + static_cast<int64_t>(b) * q_stride;
#pragma unroll
  for (int k = 0; k < kMqMaxItems; ++k) {
    const int it = i0 + threadIdx.x + k * kMqThreads;
    if (it < i1) {
      const float2 a = unpack2<T>(vals[k].x), c = unpack2<T>(vals[k].y);
      const float x0 = fmaxf(fminf(a.x * inv, 448.0f), -448.0f), x1 = fmaxf(fminf(a.y * inv, 448.0f), -448.0f);
      const float x2 = fmaxf(fminf(c.x * inv, 448.0f), -448.0f), x3 = fmaxf(fminf(c.y * inv, 448.0f), -448.0f);
      uint16_t lo, hi;
      asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(lo) : "f"(x1), "f"(x0));
      asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(hi) : "f"(x3), "f"(x2));
      *reinterpret_cast<uint32_t*>(qrow + it * 4) = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
    }
  }
}

template <typename T>
static int launch_merge_quant(const DecodeParams& p, int Dv, int align, cudaStream_t st) {
  const int n4 = p.Hq * Dv / 4;
  int S = ceil_div(n4, kMqThreads);  // one item per thread where the cluster limit allows it
  if (S > 8) S = 8;
  const int per_cta = ceil_div(n4, S);
  B200_REQUIRE(per_cta <= kMqThreads * kMqMaxItems,
               "attn_merge_quant: row of %d x %d is too long for one cluster", p.Hq, Dv);
  auto kern = attn_merge_quant_kernel<T>;
  {
    ScopedKernelTimer timer("attn_merge_quant", st);
    B200_CUDA_OK(launch_kernel(kern, dim3(p.bs * S), dim3(kMqThreads), 0, st, S,
                               pdl_family(p.bs, kPdlMerge), static_cast<const float*>(p.part_o),
                               static_cast<const float*>(p.part_lse), static_cast<T*>(p.o), p.o_stride,
                               p.q_out, p.q_out_stride, p.q_scales, p.seq_lens, p.Hq, Dv, p.num_splits,
                               align, per_cta));
  }
  B200_CHECK_LAUNCH("attn_merge_quant_kernel");
  return 0;
}

// ---- host-side dispatch -----------------------------------------------------------------------
// One instance family, tuned on the B200 for the HBM-bound regime (profiles/r01_decode_tuning.md):
// 16-token tiles, 2 stages, 3 warps -> 48 KiB per CTA at D=128, 4 CTAs (12 warps) per SM.  More
// resident warps beat deeper per-warp rings; 12 warps x 8 KiB keeps ~100 KiB of gathers in flight.
constexpr int kNT = 16, kStages = 2, kNW = 3;

static int tile_heads(int group) { return group <= 8 ? 8 : 16; }
static int decode_align() { return kNW * kNT; }
static int decode_ctas_per_sm(int D) {
  const int smem = MmaCfg<128, kNT, kStages, kNW>::SMEM_BYTES * D / 128 + 1024;
  int n = (227 * 1024) / smem;
  const int by_regs = 65536 / (kNW * 32 * 168);
  n = n < by_regs ? n : by_regs;
  return n < 1 ? 1 : n;
}

// launcher of the fused-tail instantiations (own attribute flag per kernel through the template)
template <typename C>
static int launch_decode_any(void (*kern)(const DecodeParams), const DecodeParams& p, cudaStream_t st,
                             const char* timer_name, const char* kernel_name) {
  static void (*configured[64])(const DecodeParams) = {};
  static int n_configured = 0;
  bool done = false;
  for (int i = 0; i < n_configured; ++i) done |= configured[i] == kern;
  if (!done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    if (n_configured < 64) configured[n_configured++] = kern;
  }
  dim3 grid(p.num_splits, p.Hkv * p.head_tiles, p.bs);
  {
    ScopedKernelTimer timer(timer_name, st);
    B200_CUDA_OK(launch_kernel(kern, grid, dim3(C::THREADS), C::SMEM_BYTES, st, 1,
                               pdl_family(p.bs, kPdlAttn), p));
  }
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", kernel_name, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return 2;
  }
  return 0;
}

template <typename T, int D, bool HI>
static int launch_decode(const DecodeParams& p, cudaStream_t st) {
  using C = MmaCfg<D, kNT, kStages, kNW>;
  if (p.tickets) return launch_decode_any<C>(attn_decode_mma_kernel<T, D, kNT, kStages, kNW, HI, true>, p, st,
                                             "attn_decode_split", "attn_decode_mma_kernel");
  auto kern = attn_decode_mma_kernel<T, D, kNT, kStages, kNW, HI, false>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(p.num_splits, p.Hkv * p.head_tiles, p.bs);
  {
    ScopedKernelTimer timer("attn_decode_split", st);
    B200_CUDA_OK(launch_kernel(kern, grid, dim3(C::THREADS), C::SMEM_BYTES, st, 1, pdl_family(p.bs, kPdlAttn), p));
  }
  B200_CHECK_LAUNCH("attn_decode_mma_kernel");
  return 0;
}

template <typename T>
static int dispatch_decode(const DecodeParams& p, int D, int G, cudaStream_t st) {
  if (D == 128) return G > 8 ? launch_decode<T, 128, true>(p, st) : launch_decode<T, 128, false>(p, st);
  if (D == 64) return G > 8 ? launch_decode<T, 64, true>(p, st) : launch_decode<T, 64, false>(p, st);
  set_error("b200_attn_decode: unsupported head_dim %d (supported: 64, 128)", D);
  return 1;
}

// ---- FP8 KV pool: 32-token tiles (same 8 KiB per stage as the bf16 kernel's 16-token tiles) ------
constexpr int kNT8 = 32;
static int decode_align_fp8kv() { return kNW * kNT8; }

static int fp8kv_stages() {
  static int v = [] {
    const char* e = std::getenv("B200_FP8KV_STAGES");
    const int s = (e && *e) ? std::atoi(e) : 2;
    return s == 3 ? 3 : 2;
  }();
  return v;
}

template <typename T, int D, bool HI, int ST>
static int launch_decode_fp8kv_st(const DecodeParams& p, cudaStream_t st) {
  using C = Fp8KvCfg<D, kNT8, ST, kNW>;
  if (p.tickets) return launch_decode_any<C>(attn_decode_fp8kv_kernel<T, D, kNT8, ST, kNW, HI, true>, p, st,
                                             "attn_decode_fp8kv", "attn_decode_fp8kv_kernel");
  auto kern = attn_decode_fp8kv_kernel<T, D, kNT8, ST, kNW, HI, false>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(p.num_splits, p.Hkv * p.head_tiles, p.bs);
  {
    ScopedKernelTimer timer("attn_decode_fp8kv", st);
    B200_CUDA_OK(launch_kernel(kern, grid, dim3(C::THREADS), C::SMEM_BYTES, st, 1,
                               pdl_family(p.bs, kPdlAttn), p));
  }
  B200_CHECK_LAUNCH("attn_decode_fp8kv_kernel");
  return 0;
}

template <typename T, int D, bool HI>
static int launch_decode_fp8kv(const DecodeParams& p, cudaStream_t st) {
  return fp8kv_stages() == 3 ? launch_decode_fp8kv_st<T, D, HI, 3>(p, st)
                             : launch_decode_fp8kv_st<T, D, HI, 2>(p, st);
}

// e5m2 pool (`--kv-cache-dtype fp8_e5m2`): the same kernel with the byte -> f16 conversion swapped
// (one PRMT instead of a cvt); 2-stage ring, no in-kernel tail.
template <typename T, int D, bool HI>
static int launch_decode_e5m2kv(const DecodeParams& p, cudaStream_t st) {
  using C = Fp8KvCfg<D, kNT8, 2, kNW>;
  auto kern = attn_decode_fp8kv_kernel<T, D, kNT8, 2, kNW, HI, false, true>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(p.num_splits, p.Hkv * p.head_tiles, p.bs);
  {
    ScopedKernelTimer timer("attn_decode_fp8kv", st);
    B200_CUDA_OK(launch_kernel(kern, grid, dim3(C::THREADS), C::SMEM_BYTES, st, 1,
                               pdl_family(p.bs, kPdlAttn), p));
  }
  B200_CHECK_LAUNCH("attn_decode_fp8kv_kernel<e5m2>");
  return 0;
}

template <typename T>
static int dispatch_decode_fp8kv(const DecodeParams& p, int D, int G, bool e5m2, cudaStream_t st) {
  if (e5m2 && D == 128)
    return G > 8 ? launch_decode_e5m2kv<T, 128, true>(p, st) : launch_decode_e5m2kv<T, 128, false>(p, st);
  if (e5m2 && D == 64)
    return G > 8 ? launch_decode_e5m2kv<T, 64, true>(p, st) : launch_decode_e5m2kv<T, 64, false>(p, st);
  if (D == 128)
    return G > 8 ? launch_decode_fp8kv<T, 128, true>(p, st) : launch_decode_fp8kv<T, 128, false>(p, st);
  if (D == 64)
    return G > 8 ? launch_decode_fp8kv<T, 64, true>(p, st) : launch_decode_fp8kv<T, 64, false>(p, st);
  set_error("b200_attn_decode_fp8kv: unsupported head_dim %d (supported: 64, 128)", D);
  return 1;
}

// Host side of the fused tail (see decode_fused_tail).
struct DecodeTail {
  void* tickets;
  size_t ticket_bytes;
  void* q_out;
  int64_t q_out_stride;
  float* q_scales;
};

static int apply_tail(DecodeParams& p, const DecodeTail* tail, int D, const char* who) {
  p.tickets = nullptr;
  p.q_out = nullptr;
  p.q_out_stride = 0;
  p.q_scales = nullptr;
  if (!tail) return 0;
  const size_t need = static_cast<size_t>(p.bs) * (1 + p.Hkv * p.head_tiles) * sizeof(unsigned);
  B200_REQUIRE(!tail->tickets || tail->ticket_bytes >= need,
               "%s: ticket buffer too small (%zu < %zu bytes)", who, tail->ticket_bytes, need);
  B200_REQUIRE((tail->q_out == nullptr) == (tail->q_scales == nullptr),
               "%s: q_out and q_scales go together", who);
  B200_REQUIRE(tail->q_out || p.num_splits > 1, "%s: nothing to fuse (one split, no q_out)", who);
  B200_REQUIRE(!tail->q_out || (tail->q_out_stride % 8 == 0 &&
                                (reinterpret_cast<uintptr_t>(tail->q_out) & 7u) == 0),
               "%s: q_out must be 8-byte aligned with q_out_stride %% 8 == 0", who);
  B200_REQUIRE(p.o_stride % 8 == 0 && aligned16(p.o) && (p.Hq * D) % 8 == 0,
               "%s: o must be 16-byte aligned with o_stride %% 8 == 0", who);
  p.tickets = static_cast<unsigned*>(tail->tickets);  // null: the tail is one merge(+quant) launch
  p.q_out = static_cast<uint8_t*>(tail->q_out);
  p.q_out_stride = tail->q_out_stride;
  p.q_scales = tail->q_scales;
  return 0;
}

}  // namespace b200

extern "C" {

size_t b200_attn_decode_workspace_bytes(int bs, int Hq, int Dv, int num_splits) {
  if (num_splits <= 1) return 0;
  const size_t rows = static_cast<size_t>(bs) * Hq * num_splits;
  return rows * Dv * sizeof(float) + rows * sizeof(float);
}

int b200_attn_decode_num_splits(int bs, int Hkv, int group, int D, int64_t avg_len, int num_sms) {
  using namespace b200;
  if (bs <= 0 || Hkv <= 0 || group <= 0) return 1;
  if (num_sms <= 0) num_sms = 148;
  const int G = tile_heads(group);
  const int head_tiles = ceil_div(group, G);
  const int64_t base = static_cast<int64_t>(bs) * Hkv * head_tiles;
  const int ctas_per_sm = decode_ctas_per_sm(D);
  const int64_t slots = static_cast<int64_t>(num_sms) * ctas_per_sm;
  const int align = decode_align();
  // HBM-bound: what matters is that enough CTAs stay resident until the end, not whole waves —
  // late CTAs run faster because they share the bandwidth with fewer neighbours.  Take the
  // smallest split count that oversubscribes the resident slots by `kOver` (measured on cfg 2:
  // s = 2 -> 1.7x beats s = 1 -> 0.86x by 20 % and s = 8 by 4 %), with at least 4 CTA tiles
  // (192 tokens) per split.
  constexpr double kOver = 1.5;
  int64_t max_s = avg_len / (4 * align);
  if (max_s > 64) max_s = 64;
  if (max_s < 1) max_s = 1;
  // Few kv heads per rank (tensor parallel): equal-length CTAs that all fit the resident slots at
  // once finish together, while any overflow runs as a thin second wave.  Measured on the per-rank
  // shapes of Llama-3.1-8B at TP 2/4/8, batch 64 x 4k (profiles/r02_sweep_decode_tp_shapes.json):
  // the best split count is the LARGEST one whose CTAs fit one wave — 2 / 4 / 9 splits = 512 / 512 /
  // 576 CTAs on 592 slots (90.8 / 50.7 / 29.9 us) against 93.9 / 56.4 / 35.3 us for the
  // oversubscription rule below.
  {
    const int64_t s1 = slots / base;
    // (s1 == 1, the 512 CTAs of the whole-model bench shape: 170.4 us and no merge launch against
    //  174.4 us + merge with two splits, profiles/r02_bench_n1.json)
    if (s1 >= 1 && s1 <= max_s && static_cast<double>(base * s1) >= 0.85 * static_cast<double>(slots))
      return static_cast<int>(s1);
  }
  int best = static_cast<int>(max_s);
  for (int s = 1; s <= max_s; ++s) {
    if (static_cast<double>(base * s) >= kOver * static_cast<double>(slots)) {
      best = s;
      break;
    }
  }
  return best;
}

static int attn_decode_impl(const void* q, int64_t q_stride, const void* k_cache,
                            const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                            const int32_t* req_to_token, int64_t req_to_token_stride,
                            const int64_t* req_pool_indices, const int64_t* seq_lens, int bs,
                            int Hq, int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                            void* workspace, size_t workspace_bytes, int num_splits, void* stream,
                            const b200::DecodeTail* tail) {
  using namespace b200;
  if (bs == 0) return 0;
  B200_REQUIRE(q && k_cache && v_cache && o && req_to_token && req_pool_indices && seq_lens,
               "b200_attn_decode: null pointer");
  B200_REQUIRE(bs > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0,
               "b200_attn_decode: bad head config bs=%d Hq=%d Hkv=%d", bs, Hq, Hkv);
  B200_REQUIRE(D == Dv, "b200_attn_decode: qk_head_dim (%d) != v_head_dim (%d) is not supported",
               D, Dv);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16,
               "b200_attn_decode: dtype %d not supported (bf16/f16 KV only)", dtype);
  B200_REQUIRE(num_splits >= 1 && num_splits <= 1024, "b200_attn_decode: bad num_splits %d",
               num_splits);
  B200_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && q_stride % 8 == 0 &&
                   kv_row_stride % 8 == 0,
               "b200_attn_decode: q/k_cache/v_cache must be 16-byte aligned with strides %% 8 == 0");
  B200_REQUIRE(bs <= 65535 && Hkv * 8 <= 65535, "b200_attn_decode: grid too large");
  const size_t need = b200_attn_decode_workspace_bytes(bs, Hq, Dv, num_splits);
  B200_REQUIRE(num_splits == 1 || (workspace && workspace_bytes >= need),
               "b200_attn_decode: workspace too small (%zu < %zu)", workspace_bytes, need);

  DecodeParams p;
  p.q = q;
  p.q_stride = q_stride;
  p.k_cache = k_cache;
  p.v_cache = v_cache;
  p.kv_row_stride = kv_row_stride;
  p.o = o;
  p.o_stride = o_stride;
  p.req_to_token = req_to_token;
  p.rtt_stride = req_to_token_stride;
  p.req_pool_indices = req_pool_indices;
  p.seq_lens = seq_lens;
  p.bs = bs;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.group = Hq / Hkv;
  const int G = tile_heads(p.group);
  p.head_tiles = ceil_div(p.group, G);
  p.num_splits = num_splits;
  p.logit_cap = logit_cap;
  p.v_scale = 1.f;
  p.qk_scale = logit_cap > 0.f ? scale : scale * kLog2e;
  { static const char* e = std::getenv("B200_PDL_LATE"); p.pdl_late = e ? std::atoi(e) : 1; }
  p.part_o = static_cast<float*>(workspace);
  p.part_lse = p.part_o ? p.part_o + static_cast<size_t>(bs) * Hq * num_splits * Dv : nullptr;

  if (int rc = apply_tail(p, tail, D, "b200_attn_decode_fused")) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = dtype == B200_BF16 ? dispatch_decode<__nv_bfloat16>(p, D, G, st)
                              : dispatch_decode<__half>(p, D, G, st);
  if (rc) return rc;
  if (tail && !tail->tickets && tail->q_out)
    return dtype == B200_BF16 ? launch_merge_quant<__nv_bfloat16>(p, Dv, decode_align(), st)
                              : launch_merge_quant<__half>(p, Dv, decode_align(), st);
  if (num_splits > 1 && !(tail && tail->tickets))
    return b200_attn_merge(p.part_o, p.part_lse, o, o_stride, seq_lens, bs, Hq, Dv, num_splits,
                           decode_align(), dtype, stream);
  return 0;
}

int b200_attn_decode(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                     int64_t kv_row_stride, void* o, int64_t o_stride,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const int64_t* req_pool_indices, const int64_t* seq_lens, int bs, int Hq,
                     int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                     void* workspace, size_t workspace_bytes, int num_splits, void* stream) {
  return attn_decode_impl(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride, req_to_token,
                          req_to_token_stride, req_pool_indices, seq_lens, bs, Hq, Hkv, D, Dv, scale,
                          logit_cap, dtype, workspace, workspace_bytes, num_splits, stream, nullptr);
}

int b200_attn_decode_fused(const void* q, int64_t q_stride, const void* k_cache,
                           const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens, int bs, int Hq,
                           int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                           void* workspace, size_t workspace_bytes, int num_splits, void* tickets,
                           size_t ticket_bytes, void* q_out, int64_t q_out_stride, float* q_scales,
                           void* stream) {
  b200::DecodeTail tail = {tickets, ticket_bytes, q_out, q_out_stride, q_scales};
  return attn_decode_impl(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride, req_to_token,
                          req_to_token_stride, req_pool_indices, seq_lens, bs, Hq, Hkv, D, Dv, scale,
                          logit_cap, dtype, workspace, workspace_bytes, num_splits, stream, &tail);
}

size_t b200_attn_decode_fused_ticket_bytes(int bs, int Hq, int Hkv) {
  using namespace b200;
  if (bs <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return 0;
  const int group = Hq / Hkv;
  return static_cast<size_t>(bs) * (1 + Hkv * ceil_div(group, tile_heads(group))) * sizeof(unsigned);
}

int b200_attn_decode_fp8kv_split_align(void) { return b200::decode_align_fp8kv(); }

static int attn_decode_fp8kv_impl(const void* q, int64_t q_stride, const void* k_cache,
                                  const void* v_cache, int64_t kv_row_stride, void* o,
                                  int64_t o_stride, const int32_t* req_to_token,
                                  int64_t req_to_token_stride, const int64_t* req_pool_indices,
                                  const int64_t* seq_lens, int bs, int Hq, int Hkv, int D, int Dv,
                                  float scale, float logit_cap, float k_scale, float v_scale,
                                  int dtype, void* workspace, size_t workspace_bytes, int num_splits,
                                  void* stream, const b200::DecodeTail* tail) {
  using namespace b200;
  if (bs == 0) return 0;
  const bool e5m2 = (dtype & B200_KV_FP8_E5M2) != 0;
  dtype &= ~B200_KV_FP8_E5M2;
  DecodeTail tail_e5m2;
  if (e5m2 && tail && tail->tickets) {  // no in-kernel tail instance for e5m2: one merge launch
    tail_e5m2 = *tail;
    tail_e5m2.tickets = nullptr;
    tail_e5m2.ticket_bytes = 0;
    tail = &tail_e5m2;
  }
  B200_REQUIRE(q && k_cache && v_cache && o && req_to_token && req_pool_indices && seq_lens,
               "b200_attn_decode_fp8kv: null pointer");
  B200_REQUIRE(bs > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0,
               "b200_attn_decode_fp8kv: bad head config bs=%d Hq=%d Hkv=%d", bs, Hq, Hkv);
  B200_REQUIRE(D == Dv, "b200_attn_decode_fp8kv: qk_head_dim (%d) != v_head_dim (%d)", D, Dv);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16,
               "b200_attn_decode_fp8kv: q/o dtype %d not supported (bf16/f16)", dtype);
  B200_REQUIRE(num_splits >= 1 && num_splits <= 1024, "b200_attn_decode_fp8kv: bad num_splits %d",
               num_splits);
  B200_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && q_stride % 8 == 0 &&
                   kv_row_stride % 16 == 0,
               "b200_attn_decode_fp8kv: q/k_cache/v_cache must be 16-byte aligned, q_stride %% 8 == 0 "
               "and kv_row_stride %% 16 == 0");
  B200_REQUIRE(bs <= 65535 && Hkv * 8 <= 65535, "b200_attn_decode_fp8kv: grid too large");
  B200_REQUIRE(k_scale > 0.f && v_scale > 0.f, "b200_attn_decode_fp8kv: scales must be positive");
  const size_t need = b200_attn_decode_workspace_bytes(bs, Hq, Dv, num_splits);
  B200_REQUIRE(num_splits == 1 || (workspace && workspace_bytes >= need),
               "b200_attn_decode_fp8kv: workspace too small (%zu < %zu)", workspace_bytes, need);
  DecodeParams p;
  p.q = q;
  p.q_stride = q_stride;
  p.k_cache = k_cache;
  p.v_cache = v_cache;
  p.kv_row_stride = kv_row_stride;
  p.o = o;
  p.o_stride = o_stride;
  p.req_to_token = req_to_token;
  p.rtt_stride = req_to_token_stride;
  p.req_pool_indices = req_pool_indices;
  p.seq_lens = seq_lens;
  p.bs = bs;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.group = Hq / Hkv;
  const int G = tile_heads(p.group);
  p.head_tiles = ceil_div(p.group, G);
  p.num_splits = num_splits;
  p.logit_cap = logit_cap;
  p.v_scale = v_scale;
  p.qk_scale = (logit_cap > 0.f ? scale : scale * kLog2e) * k_scale;
  { static const char* e = std::getenv("B200_PDL_LATE"); p.pdl_late = e ? std::atoi(e) : 1; }
  p.part_o = static_cast<float*>(workspace);
  p.part_lse = p.part_o ? p.part_o + static_cast<size_t>(bs) * Hq * num_splits * Dv : nullptr;
  if (int rc = apply_tail(p, tail, D, "b200_attn_decode_fp8kv_fused")) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rc = dtype == B200_BF16 ? dispatch_decode_fp8kv<__nv_bfloat16>(p, D, G, e5m2, st)
                                    : dispatch_decode_fp8kv<__half>(p, D, G, e5m2, st);
  if (rc) return rc;
  if (tail && !tail->tickets && tail->q_out)
    return dtype == B200_BF16 ? launch_merge_quant<__nv_bfloat16>(p, Dv, decode_align_fp8kv(), st)
                              : launch_merge_quant<__half>(p, Dv, decode_align_fp8kv(), st);
  if (num_splits > 1 && !(tail && tail->tickets))
    return b200_attn_merge(p.part_o, p.part_lse, o, o_stride, seq_lens, bs, Hq, Dv, num_splits,
                           decode_align_fp8kv(), dtype, stream);
  return 0;
}

int b200_attn_decode_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                           const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                           const int32_t* req_to_token, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* seq_lens, int bs,
                           int Hq, int Hkv, int D, int Dv, float scale, float logit_cap,
                           float k_scale, float v_scale, int dtype, void* workspace,
                           size_t workspace_bytes, int num_splits, void* stream) {
  return attn_decode_fp8kv_impl(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride,
                                req_to_token, req_to_token_stride, req_pool_indices, seq_lens, bs, Hq,
                                Hkv, D, Dv, scale, logit_cap, k_scale, v_scale, dtype, workspace,
                                workspace_bytes, num_splits, stream, nullptr);
}

int b200_attn_decode_fp8kv_fused(const void* q, int64_t q_stride, const void* k_cache,
                                 const void* v_cache, int64_t kv_row_stride, void* o,
                                 int64_t o_stride, const int32_t* req_to_token,
                                 int64_t req_to_token_stride, const int64_t* req_pool_indices,
                                 const int64_t* seq_lens, int bs, int Hq, int Hkv, int D, int Dv,
                                 float scale, float logit_cap, float k_scale, float v_scale,
                                 int dtype, void* workspace, size_t workspace_bytes, int num_splits,
                                 void* tickets, size_t ticket_bytes, void* q_out,
                                 int64_t q_out_stride, float* q_scales, void* stream) {
  b200::DecodeTail tail = {tickets, ticket_bytes, q_out, q_out_stride, q_scales};
  return attn_decode_fp8kv_impl(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride,
                                req_to_token, req_to_token_stride, req_pool_indices, seq_lens, bs, Hq,
                                Hkv, D, Dv, scale, logit_cap, k_scale, v_scale, dtype, workspace,
                                workspace_bytes, num_splits, stream, &tail);
}

int b200_attn_decode_split_align(int D, int group) {
  (void)D;
  (void)group;
  return b200::decode_align();
}

int b200_attn_merge(const float* part_o, const float* part_lse, void* o, int64_t o_stride,
                    const int64_t* seq_lens, int bs, int Hq, int Dv, int num_splits,
                    int split_align, int dtype, void* stream) {
  using namespace b200;
  if (bs == 0) return 0;
  const int align = split_align;
  B200_REQUIRE(num_splits >= 1 && split_align >= 1, "b200_attn_merge: bad split config %d/%d",
               num_splits, split_align);
  B200_REQUIRE(part_o && part_lse && o && seq_lens, "b200_attn_merge: null pointer");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_attn_merge: dtype %d", dtype);
  dim3 grid(Hq, bs);
  const int threads = Dv >= 128 ? 128 : 64;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ScopedKernelTimer timer("attn_merge", st);
  if (dtype == B200_BF16)
    B200_CUDA_OK(launch_kernel(attn_merge_kernel<__nv_bfloat16>, grid, dim3(threads), 0, st, 1, pdl_family(bs, kPdlMerge),
                               part_o, part_lse, static_cast<__nv_bfloat16*>(o), o_stride, seq_lens,
                               Hq, Dv, num_splits, align));
  else
    B200_CUDA_OK(launch_kernel(attn_merge_kernel<__half>, grid, dim3(threads), 0, st, 1, pdl_family(bs, kPdlMerge), part_o,
                               part_lse, static_cast<__half*>(o), o_stride, seq_lens, Hq, Dv,
                               num_splits, align));
  B200_CHECK_LAUNCH("attn_merge_kernel");
  return 0;
}

}  // extern "C"

// Paged split-KV decode attention over an FP8 (e4m3, or e5m2 with E5M2) KV pool — included by
// attn_decode.cu.
//
// Path replaced: AttentionBackend.forward_decode with `--kv-cache-dtype fp8_e4m3`
// (model_runner.py:845-858; pool stores e4m3 bytes, memory_pool.py:119-123,390-400; FlashInfer
// semantics flashinfer_backend.py:427-438,499-509: logits *= k_scale, output *= v_scale; the Triton
// backend reads the bytes as values, i.e. k_scale = v_scale = 1, triton_backend.py:808-833).
// Halves the HBM bytes of the bf16 kernel; everything else (grid, split-KV, per-warp cp.async
// rings, online softmax in registers, merge kernel) is the same design as attn_decode_mma_kernel.
//
// Math: the e4m3 bytes are converted to f16 IN REGISTERS (exact) and fed to mma.sync m16n8k16 f16;
// q is converted bf16 -> f16 (exact for |q| < 65504), P is rounded to f16.
//   * Q.K^T: a non-transposed ldmatrix of an 8-token x 16-byte K block hands lane (g, c) the four
//     bytes d = 4c..4c+3 of token g; the contraction index may be permuted freely, so MMA k-index
//     {2c, 2c+1, 2c+8, 2c+9} is mapped to d = {4c, 4c+1, 4c+2, 4c+3} and the Q fragment is loaded
//     with the same permutation (8 contiguous bytes of q per lane and k-step).
//   * P.V: a transposed ldmatrix of an 8-token x 16-byte V block hands lane (g, c) the bytes
//     V[2c][2g], V[2c][2g+1], V[2c+1][2g], V[2c+1][2g+1]; after conversion two byte-permutes give the
//     B fragments of TWO output tiles (even / odd d), so output tile (chunk j, parity e) column n
//     is d = 16j + 2n + e; the epilogue un-permutes when it writes the tile to shared memory.
#pragma once

namespace b200 {

template <int D, int NT, int STAGES, int NW>
struct Fp8KvCfg {
  static constexpr int RB = D;                 // bytes per (token, kv head) row
  static constexpr int CPR = D / 16;           // 16-byte chunks per row
  static constexpr int RG = 32 / CPR;          // row groups copied concurrently by a warp
  static constexpr int RPG = NT / RG;          // consecutive rows per lane
  static constexpr int STAGE_BYTES = NT * RB * 2;  // K rows then V rows
  static constexpr int THREADS = NW * 32;
  static constexpr int ALIGN = NW * NT;
  static constexpr int RING_BYTES = NW * STAGES * STAGE_BYTES;
  static constexpr int FINAL_BYTES = NW * 16 * (D + 2) * 4;
  static constexpr int SMEM_BYTES = RING_BYTES > FINAL_BYTES ? RING_BYTES : FINAL_BYTES;
  static_assert(D == 64 || D == 128, "head_dim");
  static_assert(RPG % 4 == 0, "slot ids are fetched as int4");
  static_assert(NT % 16 == 0, "PV k-step is 16 tokens");
};

// 16-byte chunk c of row r lives at chunk (c ^ swz(r)): 8 consecutive rows of one logical chunk
// hit 8 different 16-byte bank groups (128-byte rows: r & 7; 64-byte rows: two rows per 128 B).
template <int RB>
__device__ __forceinline__ int fp8kv_swz(int row) {
  return RB == 128 ? (row & 7) : ((row >> 1) & 3);
}

// fp8 bytes -> f16x2 registers: cvt_fp8x4_f16x4<E5M2> (common.cuh)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;\n" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
// four consecutive q elements (8 bytes) -> two f16x2 registers
template <typename T>
__device__ __forceinline__ void q4_to_f16(const T* q, uint32_t& lo, uint32_t& hi);
template <>
__device__ __forceinline__ void q4_to_f16<__half>(const __half* q, uint32_t& lo, uint32_t& hi) {
  const uint2 v = *reinterpret_cast<const uint2*>(q);
  lo = v.x;
  hi = v.y;
}
template <>
__device__ __forceinline__ void q4_to_f16<__nv_bfloat16>(const __nv_bfloat16* q, uint32_t& lo,
                                                          uint32_t& hi) {
  const uint2 v = *reinterpret_cast<const uint2*>(q);
  const float2 a = unpack2<__nv_bfloat16>(v.x), b = unpack2<__nv_bfloat16>(v.y);
  const __half2 ha = __floats2half2_rn(a.x, a.y), hb = __floats2half2_rn(b.x, b.y);
  lo = *reinterpret_cast<const uint32_t*>(&ha);
  hi = *reinterpret_cast<const uint32_t*>(&hb);
}

template <typename T, int D, int NT, int STAGES, int NW, bool HI, bool TAIL, bool E5M2 = false>
__global__ void __launch_bounds__(NW * 32, TAIL ? 4 : 0)
attn_decode_fp8kv_kernel(const DecodeParams p) {
  using C = Fp8KvCfg<D, NT, STAGES, NW>;
  constexpr int CPR = C::CPR, RPG = C::RPG, RB = C::RB;
  constexpr int STAGE_BYTES = C::STAGE_BYTES;
  constexpr int KS = D / 16;   // k-steps of Q.K^T (one 16-byte chunk each)
  constexpr int NS = NT / 8;   // n-tiles (8 tokens) of S
  constexpr int NO = D / 8;    // n-tiles (8 dims) of O
  constexpr int GT = HI ? 16 : 8;

  extern __shared__ __align__(128) uint8_t smem[];
  griddep_wait();
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
  const int64_t row_stride_bytes = p.kv_row_stride;  // one byte per element
  const int head0 = ht * GT;
  const int heads_here = min(GT, p.group - head0);

  // ---- Q fragments (A operand, rows = heads), k-index permuted as described above -------------
  uint32_t qa[KS][4];
  {
    const int r_lo = lane >> 2, r_hi = r_lo + 8;
    const T* qb = static_cast<const T*>(p.q) + static_cast<int64_t>(b) * p.q_stride +
                  static_cast<int64_t>(kvh * p.group + head0) * D + (lane & 3) * 4;
    const bool v_lo = r_lo < heads_here, v_hi = HI && r_hi < heads_here;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qa[ks][0] = qa[ks][1] = qa[ks][2] = qa[ks][3] = 0u;
      if (v_lo) q4_to_f16<T>(qb + r_lo * D + ks * 16, qa[ks][0], qa[ks][2]);
      if (v_hi) q4_to_f16<T>(qb + r_hi * D + ks * 16, qa[ks][1], qa[ks][3]);
    }
  }

  float o[NO][4];
#pragma unroll
  for (int i = 0; i < NO; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_lo = kNegBig, m_hi = kNegBig, l_lo = 0.f, l_hi = 0.f;

  // ---- copy-issue geometry ---------------------------------------------------------------------
  const int ccol = lane % CPR;
  const int crow0 = (lane / CPR) * RPG;
  const char* kcol = static_cast<const char*>(p.k_cache) + static_cast<int64_t>(kvh) * RB + ccol * 16;
  const char* vcol = static_cast<const char*>(p.v_cache) + static_cast<int64_t>(kvh) * RB + ccol * 16;
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
        const uint32_t dst = sk + row * RB + ((ccol ^ fp8kv_swz<RB>(row)) << 4);
        cp_async16(dst, kcol + off);
        cp_async16(dst + NT * RB, vcol + off);
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
  const float inv_cap = capped ? p.qk_scale / p.logit_cap : 0.f;
  const float cap_log2 = p.logit_cap * kLog2e;
  const float sc_mul = capped ? 1.f : p.qk_scale;

  // ldmatrix lane geometry.  QK: matrix i of an x4 = chunk (4*jj + i) of the 8 tokens of n-tile nt.
  const int lx = lane & 7;
  const int k_mat = lane >> 3;
  // PV: matrices = (tokens 0-7, chunk j), (tokens 8-15, chunk j), (0-7, j+1), (8-15, j+1)
  const int v_row = ((lane >> 3) & 1) * 8 + lx;
  const int v_mat = lane >> 4;

  for (int it = 0; it < n_my; ++it) {
    issue_tile(it + STAGES - 1);
    load_slots(it + STAGES);
    cp_async_wait<STAGES - 1>();
    __syncwarp();

    const uint32_t sk = ring + (it % STAGES) * STAGE_BYTES;
    const uint32_t sv = sk + NT * RB;
    const int tok0 = start + (warp + it * NW) * NT;

    // ---- S = Q . K^T ---------------------------------------------------------------------------
    float s[NS][4];
#pragma unroll
    for (int nt = 0; nt < NS; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      const int row = nt * 8 + lx;
      const uint32_t rbase = sk + row * RB;
      const int sw = fp8kv_swz<RB>(row);
#pragma unroll
      for (int jj = 0; jj < KS / 4; ++jj) {
        uint32_t kb[4];
        ldsm_x4(kb, rbase + (((4 * jj + k_mat) ^ sw) << 4));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t b0, b1;
          cvt_fp8x4_f16x4<E5M2>(kb[i], b0, b1);
          mma16816<__half>(s[nt], qa[4 * jj + i], b0, b1);
        }
      }
    }
    if (capped) {
#pragma unroll
      for (int nt = 0; nt < NS; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[nt][e] = cap_log2 * fast_tanh(s[nt][e] * inv_cap);
    }
    if (tok0 + NT > end) {
#pragma unroll
      for (int nt = 0; nt < NS; ++nt) {
        const int t = tok0 + nt * 8 + (lane & 3) * 2;
        if (t >= end) s[nt][0] = s[nt][2] = -INFINITY;
        if (t + 1 >= end) s[nt][1] = s[nt][3] = -INFINITY;
      }
    }

    // ---- online softmax (identical to the bf16 kernel) ----------------------------------------
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
      pa[nt >> 1][(nt & 1) * 2] = pack2<__half>(p0, p1);
      if (HI) {
        const float p2 = fast_exp2(fmaf(s[nt][2], sc_mul, -m_hi));
        const float p3 = fast_exp2(fmaf(s[nt][3], sc_mul, -m_hi));
        l_hi += p2 + p3;
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack2<__half>(p2, p3);
      } else {
        pa[nt >> 1][(nt & 1) * 2 + 1] = 0u;
      }
    }

    // ---- O += P . V --------------------------------------------------------------------------
#pragma unroll
    for (int kt = 0; kt < NT / 16; ++kt) {
      const int row = kt * 16 + v_row;
      const uint32_t rbase = sv + row * RB;
      const int sw = fp8kv_swz<RB>(row);
#pragma unroll
      for (int j = 0; j < CPR; j += 2) {
        uint32_t vb[4];
        ldsm_x4_t(vb, rbase + (((j + v_mat) ^ sw) << 4));
#pragma unroll
        for (int h = 0; h < 2; ++h) {  // chunk j + h: vb[2h] = tokens 0-7, vb[2h+1] = tokens 8-15
          uint32_t l0, h0, l1, h1;
          cvt_fp8x4_f16x4<E5M2>(vb[2 * h], l0, h0);
          cvt_fp8x4_f16x4<E5M2>(vb[2 * h + 1], l1, h1);
          const int t = 2 * (j + h);
          mma16816<__half>(o[t], pa[kt], prmt(l0, h0, 0x5410u), prmt(l1, h1, 0x5410u));
          mma16816<__half>(o[t + 1], pa[kt], prmt(l0, h0, 0x7632u), prmt(l1, h1, 0x7632u));
        }
      }
    }
    __syncwarp();
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
    // output tile (chunk j, parity e) column n is d = 16j + 2n + e; this lane holds n = 2c, 2c+1
    const int r_lo = lane >> 2, c4 = (lane & 3) * 4;
#pragma unroll
    for (int j = 0; j < CPR; ++j) {
      *reinterpret_cast<float4*>(fin + (warp * 16 + r_lo) * D + j * 16 + c4) =
          make_float4(o[2 * j][0], o[2 * j + 1][0], o[2 * j][1], o[2 * j + 1][1]);
      if (HI)
        *reinterpret_cast<float4*>(fin + (warp * 16 + r_lo + 8) * D + j * 16 + c4) =
            make_float4(o[2 * j][2], o[2 * j + 1][2], o[2 * j][3], o[2 * j + 1][3]);
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
      const float f = fast_exp2(fin_m[w * 16 + h] - M);
      num += fin[(w * 16 + h) * D + d] * f;
      den += fin_l[w * 16 + h] * f;
    }
    const float out = num / den * p.v_scale;
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

}  // namespace b200

// Paged split-KV decode attention for sm_100a.
//
// Path replaced: AttentionBackend.forward_decode (numerical oracle
// torch_native_backend.py:112-180; GPU reference triton_ops/decode_attention.py:50-557).
//
// Work decomposition
//   grid = (num_splits, Hkv * head_tiles, bs).  One CTA owns one (request, kv head, KV split) and
//   ALL query heads of that kv head's GQA group (G per head tile), so every K/V row is fetched
//   from HBM exactly once per group.  The KV pool is token-granular (page_size 1): row
//   `slot = req_to_token[req, t]` of k_cache/v_cache, D contiguous elements per kv head.
//   The kernel reads req_to_token directly — no flat kv_indices round trip.
//
// Data movement (HBM-bound part)
//   Each warp runs its own STAGES-deep cp.async ring (16-byte LDGSTS, L2-only) over
//   TILE-token tiles of gathered K and V rows; slot ids are prefetched one tile ahead and
//   broadcast with warp shuffles.  No CTA-wide barrier in the main loop.
//
// Math (CUDA cores, fp32)
//   LPT lanes share one token: lane j owns 16-byte chunks {j, j+LPT, ..} of the row, q lives
//   pre-scaled in registers (scale*log2e folded in), dot products use packed FFMA2
//   (fma.rn.f32x2) and a warp-shuffle butterfly.  Every lane group keeps an independent
//   online-softmax state (m, l, acc); groups -> warp -> CTA are merged once at the end.
//   Partials (normalised O, log2-domain LSE) go to the workspace, merged by attn_merge_kernel.
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

template <typename T, int D, int G, int LPT, int R, int STAGES, int NW>
struct DecodeCfg {
  static constexpr int TPW = 32 / LPT;       // tokens handled concurrently by one warp
  static constexpr int TILE = TPW * R;       // tokens per warp tile (one pipeline stage)
  static constexpr int CPR = D / 8;          // 16-byte chunks per row
  static constexpr int NV = CPR / LPT;       // chunks per lane per row
  static constexpr int ROW_BYTES = D * 2;
  static constexpr int STAGE_BYTES = TILE * ROW_BYTES * 2;  // K rows then V rows
  static constexpr int LOADS = TILE * CPR / 32;             // cp.async per lane per tile (K)
  static constexpr int THREADS = NW * 32;
  static constexpr int ALIGN = NW * TILE;    // split length granularity
  static constexpr int RING_BYTES = NW * STAGES * STAGE_BYTES;
  static constexpr int FINAL_BYTES = NW * G * (D + 2) * 4;
  static constexpr int SMEM_BYTES = RING_BYTES > FINAL_BYTES ? RING_BYTES : FINAL_BYTES;
  static_assert(CPR % LPT == 0, "row chunks must divide across lanes");
  static_assert((TILE * CPR) % 32 == 0, "tile must be a whole number of warp-wide copies");
  static_assert(TILE <= 32, "slot ids of a tile are held one per lane");
};

template <typename T, int D, int G, int LPT, int R, int STAGES, int NW>
__global__ void __launch_bounds__(NW * 32)
attn_decode_split_kernel(const DecodeParams p) {
  using C = DecodeCfg<T, D, G, LPT, R, STAGES, NW>;
  constexpr int TPW = C::TPW, TILE = C::TILE, CPR = C::CPR, NV = C::NV;
  constexpr int ROW_BYTES = C::ROW_BYTES, STAGE_BYTES = C::STAGE_BYTES, LOADS = C::LOADS;

  extern __shared__ __align__(16) uint8_t smem[];

  const int split = blockIdx.x;
  const int kvh = blockIdx.y / p.head_tiles;
  const int ht = blockIdx.y - kvh * p.head_tiles;
  const int b = blockIdx.z;
  // Broadcast from lane 0 so the compiler knows the warp index (and everything derived from
  // it, e.g. the tile-loop trip count) is warp-uniform: keeps the shuffles free of re-convergence
  // wrappers.
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int g = lane / LPT, j = lane % LPT;

  const int len = static_cast<int>(p.seq_lens[b]);
  const int chunk = split_chunk(len, p.num_splits, C::ALIGN);
  const int start = split * chunk;
  if (start >= len) return;  // empty split (uniform for the CTA)
  const int end = min(len, start + chunk);
  const int ntiles = ceil_div(end - start, TILE);
  const int n_my = warp < ntiles ? (ntiles - warp + NW - 1) / NW : 0;

  const int32_t* __restrict__ rtt = p.req_to_token + p.req_pool_indices[b] * p.rtt_stride;
  const char* kbase = static_cast<const char*>(p.k_cache) + static_cast<int64_t>(kvh) * ROW_BYTES;
  const char* vbase = static_cast<const char*>(p.v_cache) + static_cast<int64_t>(kvh) * ROW_BYTES;
  const int64_t row_stride_bytes = p.kv_row_stride * 2;
  const int head0 = ht * G;  // first head of this tile inside the GQA group

  // ---- q -> registers, pre-scaled -----------------------------------------------------------
  float2 qf[G][NV][4];
#pragma unroll
  for (int h = 0; h < G; ++h) {
    const bool hv = head0 + h < p.group;
    const T* qrow = static_cast<const T*>(p.q) + static_cast<int64_t>(b) * p.q_stride +
                    static_cast<int64_t>(kvh * p.group + (hv ? head0 + h : 0)) * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      uint4 raw = *reinterpret_cast<const uint4*>(qrow + (j + i * LPT) * 8);
      if (!hv) raw = make_uint4(0, 0, 0, 0);
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float2 f = unpack2<T>(w[e]);
        qf[h][i][e] = make_float2(f.x * p.qk_scale, f.y * p.qk_scale);
      }
    }
  }

  float m_run[G], l_run[G];
  float2 acc[G][NV][4];
#pragma unroll
  for (int h = 0; h < G; ++h) {
    m_run[h] = kNegBig;
    l_run[h] = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[h][i][e] = make_float2(0.f, 0.f);
  }

  const uint32_t ring = smem_u32(smem) + warp * (STAGES * STAGE_BYTES);

  // Slot id of token `lane` of this warp's k-th tile (0 for out-of-range -> dummy row 0).
  auto load_slot = [&](int k) -> int {
    int slot = 0;
    if (k < n_my) {
      const int pos = start + (warp + k * NW) * TILE + lane;
      if (lane < TILE && pos < end) slot = __ldg(rtt + pos);
    }
    return slot;
  };
  // Issue the gathered K/V row copies of this warp's k-th tile into ring stage k % STAGES.
  auto issue_tile = [&](int k, int slot_reg) {
    if (k < n_my) {
      const uint32_t sk = ring + (k % STAGES) * STAGE_BYTES;
      const uint32_t sv = sk + TILE * ROW_BYTES;
#pragma unroll
      for (int i = 0; i < LOADS; ++i) {
        const int c = i * 32 + lane;
        const int row = c / CPR, col = c % CPR;
        const int64_t off = static_cast<int64_t>(__shfl_sync(0xffffffffu, slot_reg, row)) *
                                row_stride_bytes + col * 16;
        cp_async16(sk + row * ROW_BYTES + col * 16, kbase + off);
        cp_async16(sv + row * ROW_BYTES + col * 16, vbase + off);
      }
    }
    cp_async_commit();
  };

  int slot_reg = load_slot(0);
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    const int nxt = load_slot(s + 1);
    issue_tile(s, slot_reg);
    slot_reg = nxt;
  }

  const bool capped = p.logit_cap > 0.f;
  const float inv_cap = capped ? 1.f / p.logit_cap : 0.f;
  const float cap_log2 = p.logit_cap * kLog2e;

  for (int it = 0; it < n_my; ++it) {
    {
      const int nxt = load_slot(it + STAGES);
      issue_tile(it + STAGES - 1, slot_reg);
      slot_reg = nxt;
    }
    cp_async_wait<STAGES - 1>();
    __syncwarp();

    const uint32_t sk = ring + (it % STAGES) * STAGE_BYTES;
    const uint32_t sv = sk + TILE * ROW_BYTES;
    const int tok0 = start + (warp + it * NW) * TILE;

    // ---- scores -----------------------------------------------------------------------------
    float sc[R][G];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int tl = r * TPW + g;
      float2 a2[G];
#pragma unroll
      for (int h = 0; h < G; ++h) a2[h] = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const uint4 kv = lds128(sk + tl * ROW_BYTES + (j + i * LPT) * 16);
        const float2 k0 = unpack2<T>(kv.x), k1 = unpack2<T>(kv.y), k2 = unpack2<T>(kv.z),
                     k3 = unpack2<T>(kv.w);
#pragma unroll
        for (int h = 0; h < G; ++h) {
          a2[h] = ffma2(qf[h][i][0], k0, a2[h]);
          a2[h] = ffma2(qf[h][i][1], k1, a2[h]);
          a2[h] = ffma2(qf[h][i][2], k2, a2[h]);
          a2[h] = ffma2(qf[h][i][3], k3, a2[h]);
        }
      }
      const bool valid = tok0 + tl < end;
#pragma unroll
      for (int h = 0; h < G; ++h) {
        float s = a2[h].x + a2[h].y;
#pragma unroll
        for (int off = LPT / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (capped) s = cap_log2 * fast_tanh(s * inv_cap);
        sc[r][h] = valid ? s : -INFINITY;
      }
    }

    // ---- online softmax (per lane group) ----------------------------------------------------
#pragma unroll
    for (int h = 0; h < G; ++h) {
      float mt = sc[0][h];
#pragma unroll
      for (int r = 1; r < R; ++r) mt = fmaxf(mt, sc[r][h]);
      const float m_new = fmaxf(m_run[h], mt);
      if (m_new > m_run[h]) {
        const float alpha = fast_exp2(m_run[h] - m_new);
        const float2 a2v = make_float2(alpha, alpha);
        l_run[h] *= alpha;
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[h][i][e] = fmul2(acc[h][i][e], a2v);
        m_run[h] = m_new;
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float pr = fast_exp2(sc[r][h] - m_new);
        sc[r][h] = pr;
        l_run[h] += pr;
      }
    }

    // ---- P @ V ------------------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int tl = r * TPW + g;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const uint4 vv = lds128(sv + tl * ROW_BYTES + (j + i * LPT) * 16);
        const float2 v0 = unpack2<T>(vv.x), v1 = unpack2<T>(vv.y), v2 = unpack2<T>(vv.z),
                     v3 = unpack2<T>(vv.w);
#pragma unroll
        for (int h = 0; h < G; ++h) {
          const float2 p2 = make_float2(sc[r][h], sc[r][h]);
          acc[h][i][0] = ffma2(p2, v0, acc[h][i][0]);
          acc[h][i][1] = ffma2(p2, v1, acc[h][i][1]);
          acc[h][i][2] = ffma2(p2, v2, acc[h][i][2]);
          acc[h][i][3] = ffma2(p2, v3, acc[h][i][3]);
        }
      }
    }
    __syncwarp();  // all lanes done with this stage before it is refilled next iteration
  }
  cp_async_wait<0>();

  // ---- merge lane groups inside the warp ------------------------------------------------------
#pragma unroll
  for (int h = 0; h < G; ++h) {
    float mw = m_run[h];
#pragma unroll
    for (int off = LPT; off < 32; off <<= 1) mw = fmaxf(mw, __shfl_xor_sync(0xffffffffu, mw, off));
    const float f = fast_exp2(m_run[h] - mw);
    float lw = l_run[h] * f;
#pragma unroll
    for (int off = LPT; off < 32; off <<= 1) lw += __shfl_xor_sync(0xffffffffu, lw, off);
    m_run[h] = mw;
    l_run[h] = lw;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float x = acc[h][i][e].x * f, y = acc[h][i][e].y * f;
#pragma unroll
        for (int off = LPT; off < 32; off <<= 1) {
          x += __shfl_xor_sync(0xffffffffu, x, off);
          y += __shfl_xor_sync(0xffffffffu, y, off);
        }
        acc[h][i][e] = make_float2(x, y);
      }
  }

  // ---- merge warps through shared memory (aliases the copy ring) ------------------------------
  __syncthreads();
  float* fin = reinterpret_cast<float*>(smem);         // [NW][G][D]
  float* fin_m = fin + NW * G * D;                      // [NW][G]
  float* fin_l = fin_m + NW * G;                        // [NW][G]
  if (g == 0) {
#pragma unroll
    for (int h = 0; h < G; ++h) {
#pragma unroll
      for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          *reinterpret_cast<float2*>(fin + (warp * G + h) * D + (j + i * LPT) * 8 + e * 2) =
              acc[h][i][e];
      if (j == 0) {
        fin_m[warp * G + h] = m_run[h];
        fin_l[warp * G + h] = l_run[h];
      }
    }
  }
  __syncthreads();

  for (int idx = threadIdx.x; idx < G * D; idx += NW * 32) {
    const int h = idx / D, d = idx - h * D;
    if (head0 + h >= p.group) continue;
    float M = fin_m[h];
#pragma unroll
    for (int w = 1; w < NW; ++w) M = fmaxf(M, fin_m[w * G + h]);
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float f = fast_exp2(fin_m[w * G + h] - M);
      num += fin[(w * G + h) * D + d] * f;
      den += fin_l[w * G + h] * f;
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
}

// Merge of split partials: one CTA per (query head, request).
template <typename T>
__global__ void attn_merge_kernel(const float* __restrict__ part_o,
                                  const float* __restrict__ part_lse, T* __restrict__ o,
                                  int64_t o_stride, const int64_t* __restrict__ seq_lens, int Hq,
                                  int Dv, int num_splits, int align) {
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

// ---- host-side dispatch -----------------------------------------------------------------------
struct DecodePlan {
  int G;          // heads per tile
  int align;      // split granularity of the chosen instance
  int smem;       // dynamic shared memory
  int ctas_per_sm;
};

template <typename T, int D, int G, int LPT, int R, int STAGES, int NW>
static int launch_decode(const DecodeParams& p, cudaStream_t st) {
  using C = DecodeCfg<T, D, G, LPT, R, STAGES, NW>;
  auto kern = attn_decode_split_kernel<T, D, G, LPT, R, STAGES, NW>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(p.num_splits, p.Hkv * p.head_tiles, p.bs);
  {
    ScopedKernelTimer timer("attn_decode_split", st);
    kern<<<grid, C::THREADS, C::SMEM_BYTES, st>>>(p);
  }
  B200_CHECK_LAUNCH("attn_decode_split_kernel");
  return 0;
}

// Instance table.  R/STAGES/NW are tuned for the HBM-bound regime: NW=4 warps, 16-token tiles,
// 3 stages -> 96 KiB per CTA for D=128 (2 CTAs/SM, 128 KiB of gathers in flight per SM).
constexpr int kR = 4, kStages = 3, kNW = 4;

static int tile_heads(int group) {
  if (group <= 1) return 1;
  if (group <= 2) return 2;
  if (group <= 4) return 4;
  return 8;
}
static int lanes_per_token(int D, int G) { return (D == 128 && G == 8) ? 16 : 8; }
static int decode_align(int D, int G) { return kNW * (32 / lanes_per_token(D, G)) * kR; }

template <typename T>
static int dispatch_decode(const DecodeParams& p, int D, int G, cudaStream_t st) {
#define B200_DECODE_CASE(DD, GG, LL)                                        \
  if (D == DD && G == GG) return launch_decode<T, DD, GG, LL, kR, kStages, kNW>(p, st);
  B200_DECODE_CASE(128, 1, 8)
  B200_DECODE_CASE(128, 2, 8)
  B200_DECODE_CASE(128, 4, 8)
  B200_DECODE_CASE(128, 8, 16)
  B200_DECODE_CASE(64, 1, 8)
  B200_DECODE_CASE(64, 2, 8)
  B200_DECODE_CASE(64, 4, 8)
  B200_DECODE_CASE(64, 8, 8)
#undef B200_DECODE_CASE
  set_error("b200_attn_decode: unsupported head_dim %d (supported: 64, 128)", D);
  return 1;
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
  const int ctas_per_sm = D <= 64 ? 4 : 2;
  const int64_t slots = static_cast<int64_t>(num_sms) * ctas_per_sm;
  const int align = decode_align(D, G);
  int64_t max_s = avg_len / (4 * align);  // at least 4 CTA-tiles (256 tokens) per split
  if (max_s > 64) max_s = 64;
  if (max_s < 1) max_s = 1;
  int best = 1;
  double best_eff = -1.0;
  for (int s = 1; s <= max_s; ++s) {
    const int64_t ctas = base * s;
    const int64_t waves = (ctas + slots - 1) / slots;
    const double eff = static_cast<double>(ctas) / static_cast<double>(waves * slots);
    if (eff > best_eff + 0.02) {  // prefer fewer splits unless clearly better
      best_eff = eff;
      best = s;
    }
  }
  return best;
}

int b200_attn_decode(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                     int64_t kv_row_stride, void* o, int64_t o_stride,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const int64_t* req_pool_indices, const int64_t* seq_lens, int bs, int Hq,
                     int Hkv, int D, int Dv, float scale, float logit_cap, int dtype,
                     void* workspace, size_t workspace_bytes, int num_splits, void* stream) {
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
  p.qk_scale = logit_cap > 0.f ? scale : scale * kLog2e;
  p.part_o = static_cast<float*>(workspace);
  p.part_lse = p.part_o ? p.part_o + static_cast<size_t>(bs) * Hq * num_splits * Dv : nullptr;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = dtype == B200_BF16 ? dispatch_decode<__nv_bfloat16>(p, D, G, st)
                              : dispatch_decode<__half>(p, D, G, st);
  if (rc) return rc;
  if (num_splits > 1)
    return b200_attn_merge(p.part_o, p.part_lse, o, o_stride, seq_lens, bs, Hq, Dv, num_splits,
                           decode_align(D, G), dtype, stream);
  return 0;
}

int b200_attn_decode_split_align(int D, int group) {
  return b200::decode_align(D, b200::tile_heads(group));
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
    attn_merge_kernel<__nv_bfloat16><<<grid, threads, 0, st>>>(
        part_o, part_lse, static_cast<__nv_bfloat16*>(o), o_stride, seq_lens, Hq, Dv, num_splits,
        align);
  else
    attn_merge_kernel<__half><<<grid, threads, 0, st>>>(part_o, part_lse, static_cast<__half*>(o),
                                                        o_stride, seq_lens, Hq, Dv, num_splits,
                                                        align);
  B200_CHECK_LAUNCH("attn_merge_kernel");
  return 0;
}

}  // extern "C"

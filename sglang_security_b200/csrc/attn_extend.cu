// Extend (prefill with a cached radix prefix) attention on tcgen05 / TMEM for sm_100a.
//
// Path replaced: AttentionBackend.forward_extend — numerical oracle
// torch_native_backend.py:27-110 (every new token attends causally to ALL keys of its request
// gathered from the KV pool: cached prefix + the new tokens, whose K/V were written to the pool
// just before); GPU reference triton_ops/extend_attention.py:42-438.
//
// One CTA = (request, kv head or q head, tile of 128 query rows).  Query rows are (token, head)
// pairs: PACK heads of the same GQA group are packed into the M dimension so one gathered K/V tile
// feeds PACK query heads.  KV tiles of 64 tokens are gathered row by row through req_to_token
// (page_size 1) with 16-byte cp.async into 128-byte-swizzled shared memory — exactly the layout a
// SWIZZLE_128B TMA box would produce, so tcgen05.mma consumes it through ordinary UMMA descriptors.
//
//   warps 0-3 : softmax.  Thread = one query row.  The S tile (128 x 64 fp32) is read from TMEM once
//               into registers and the accumulator released immediately (QK of the next tile runs
//               under this tile's max / exp2 / pack); bf16 P goes to swizzled smem.  O stays in TMEM and is only
//               rescaled when a row max grows by more than 2^8 (lazy rescale), so the common tile
//               costs no TMEM round trip.
//   warp 4    : MMA issuer (one thread): S = Q K^T (M128 N64 K16 x D/16), O += P V
//               (M128 N=D K16 x 4, V consumed as an MN-major B operand); owns the TMEM allocation.
//   warps 5-7 : gather loaders (cp.async double buffer, slot ids from req_to_token).
// Two CTAs are co-resident per SM so one CTA's softmax overlaps the other's MMAs.
#include <cstdlib>
#include <type_traits>

#include "tc_common.cuh"

namespace b200 {

struct ExtendParams {
  const void* q;
  int64_t q_stride;      // elements between consecutive tokens of q
  const void* k_cache;
  const void* v_cache;
  int64_t kv_row_stride; // elements between consecutive slots
  void* o;
  int64_t o_stride;
  const int32_t* req_to_token;
  int64_t rtt_stride;
  const int64_t* req_pool_indices;
  const int64_t* seq_lens;       // [bs] total kv length (prefix + extend)
  const int32_t* prefix_lens;    // [bs]
  const int32_t* extend_lens;    // [bs]
  const int32_t* extend_start;   // [bs] first row of the request in q / o
  int Hq, Hkv, group, pack, head_tiles;
  float qk_scale;                // scale*log2e, or scale when capped
  float logit_cap;
  float out_scale;               // output multiplier (v_scale of an FP8 KV pool; 1 otherwise)
  int causal;
};

constexpr int kExtThreads = 256;
constexpr int kKvTile = 64;
constexpr int kLoaders = 96;
constexpr float kRescaleThreshold = 8.0f;  // log2 units

template <int D>
struct ExtendCfg {
  static constexpr int PANELS = D / 64;                 // 128-byte-wide panels per row
  static constexpr int Q_BYTES = 128 * D * 2;
  static constexpr int KV_TILE_BYTES = kKvTile * D * 2; // one K (or V) tile
  static constexpr int P_BYTES = 128 * kKvTile * 2;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = Q_BYTES;                  // 2 stages
  static constexpr int OFF_V = OFF_K + 2 * KV_TILE_BYTES;
  static constexpr int OFF_P = OFF_V + 2 * KV_TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + P_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 128;
  static constexpr int TMEM_COLS = D == 128 ? 256 : 128;  // S: 64 cols, O: D cols
  static constexpr int O_COL = 64;
};

// Byte offset of 16-byte chunk `col` of row `row` inside a [rows x D] tile stored as D/64 panels of
// 128-byte rows with the 128-byte swizzle (chunk index XOR row%8).
template <int ROWS>
__device__ __forceinline__ uint32_t swz_off(int row, int col) {
  const int panel = col >> 3, c = col & 7;
  return panel * (ROWS * 128) + row * 128 + ((c ^ (row & 7)) << 4);
}

// three-input max (PTX ISA 8.6, sm_100+): one FMNMX3 per two new elements
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;\n" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// FAST (round 2): the softmax warps are the limiter (ncu: issue 56 %, ALU 43 %, tensor 36 %), so the
// per-element instruction count of the common tile is cut from ~7.5 to ~3: the causal / length mask
// is applied only on tiles that cross the row's last visible key, the row max is taken on the RAW
// scores with three-input max (scale > 0 commutes with max), scale and max-subtraction are ONE
// packed FFMA2 per two elements (exp2(s * scale - m)), and the row sum accumulates with FFMA2.
// Same values up to fp32 rounding of the fused multiply-add.  Not used with logit capping or a
// non-positive scale (dispatch falls back to the FAST = false instance, unchanged from round 1).
template <typename T, int D, bool FAST = false>
__global__ void __launch_bounds__(kExtThreads, 2)
attn_extend_kernel(const ExtendParams p) {
  using C = ExtendCfg<D>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();  // swizzled operands need a 1024-byte aligned base

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  // ---- work assignment ---------------------------------------------------------------------
  const int b = blockIdx.z;
  const int ext = p.extend_lens[b];
  const int prefix = p.prefix_lens[b];
  const int tok_per_tile = 128 / p.pack;
  const int tok0 = blockIdx.x * tok_per_tile;
  if (tok0 >= ext) return;  // uniform for the CTA
  const int kvh = blockIdx.y / p.head_tiles;
  const int head_base = kvh * p.group + (blockIdx.y - kvh * p.head_tiles) * p.pack;
  const int kv_len = prefix + ext;
  const int tok_last = min(tok0 + tok_per_tile, ext) - 1;
  const int n_kv = p.causal ? ceil_div(prefix + tok_last + 1, kKvTile) : ceil_div(kv_len, kKvTile);
  const int64_t q_row0 = p.extend_start[b];

  // ---- barriers ----------------------------------------------------------------------------
  const uint32_t bars = sbase + C::OFF_BAR;
  const uint32_t bar_q = bars;
  auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
  auto bar_kv_empty = [&](int s) { return bars + 8u * (3 + s); };
  const uint32_t bar_s_full = bars + 8u * 5, bar_s_free = bars + 8u * 6;
  const uint32_t bar_p_full = bars + 8u * 7, bar_pv_full = bars + 8u * 8;
  const uint32_t tmem_slot = bars + 8u * 9;

  if (threadIdx.x == 0) {
    tc::mbar_init(bar_q, kLoaders);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar_kv_full(s), kLoaders);
      tc::mbar_init(bar_kv_empty(s), 1);
    }
    tc::mbar_init(bar_s_full, 1);
    tc::mbar_init(bar_s_free, 128);
    tc::mbar_init(bar_p_full, 128);
    tc::mbar_init(bar_pv_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 4) {
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + 8 * 9);

  if (warp >= 5) {
    // ===================== gather loaders (96 threads) =====================
    const int lt = threadIdx.x - 5 * 32;
    constexpr int CPR = D / 8;  // 16-byte chunks per row
    const int32_t* rtt = p.req_to_token + p.req_pool_indices[b] * p.rtt_stride;
    const int64_t row_bytes = p.kv_row_stride * 2;
    const char* kbase = static_cast<const char*>(p.k_cache) + static_cast<int64_t>(kvh) * D * 2;
    const char* vbase = static_cast<const char*>(p.v_cache) + static_cast<int64_t>(kvh) * D * 2;
    // Q tile: row r = (token r / pack, head r % pack)
    for (int c = lt; c < 128 * CPR; c += kLoaders) {
      const int r = c / CPR, col = c - r * CPR;
      int tok = tok0 + r / p.pack;
      if (tok >= ext) tok = ext - 1;  // padding rows replicate the last token (never stored)
      const char* src = static_cast<const char*>(p.q) +
                        ((q_row0 + tok) * p.q_stride + (head_base + r % p.pack) * D) * 2 + col * 16;
      cp_async16(sbase + C::OFF_Q + swz_off<128>(r, col), src);
    }
    cp_async_commit();
    // Tile j is issued only after tile j-1 has landed and been published: the MMA warp waits for
    // tile j+1 before it issues PV(j), so publishing must never sit behind a kv_empty wait.
    // A thread owns one 16-byte column of rows lrow + RPP*i; the slot ids of tile j+1 are fetched
    // while tile j is in flight, so the gather itself never waits on a req_to_token load.
    constexpr int RPP = kLoaders / CPR;                 // rows per pass (6 for D = 128)
    constexpr int PASSES = (kKvTile + RPP - 1) / RPP;
    const int lrow = lt / CPR, lcol = lt - lrow * CPR;
    int32_t slots[PASSES];
    auto load_slots = [&](int j) {
#pragma unroll
      for (int i = 0; i < PASSES; ++i) {
        const int pos = j * kKvTile + lrow + RPP * i;
        slots[i] = (lrow + RPP * i < kKvTile && pos < kv_len) ? __ldg(rtt + pos) : 0;  // 0 = dummy row
      }
    };
    auto issue_tile = [&](int j) {
      const int stage = j & 1;
      const uint32_t sk = sbase + C::OFF_K + stage * C::KV_TILE_BYTES;
      const uint32_t sv = sbase + C::OFF_V + stage * C::KV_TILE_BYTES;
#pragma unroll
      for (int i = 0; i < PASSES; ++i) {
        const int r = lrow + RPP * i;
        if (r < kKvTile) {
          const int64_t off = static_cast<int64_t>(slots[i]) * row_bytes + lcol * 16;
          const uint32_t so = swz_off<kKvTile>(r, lcol);
          cp_async16(sk + so, kbase + off);
          cp_async16(sv + so, vbase + off);
        }
      }
      cp_async_commit();
    };
    load_slots(0);
    issue_tile(0);
    if (n_kv > 1) load_slots(1);
    cp_async_wait<1>();  // Q landed
    tc::fence_proxy_async_smem();
    tc::mbar_arrive(bar_q);
    for (int j = 1; j < n_kv; ++j) {
      cp_async_wait<0>();  // tile j-1 landed
      tc::fence_proxy_async_smem();
      tc::mbar_arrive(bar_kv_full((j - 1) & 1));
      if (j >= 2) tc::mbar_wait(bar_kv_empty(j & 1), ((j >> 1) - 1) & 1u);  // PV(j-2) done
      issue_tile(j);
      if (j + 1 < n_kv) load_slots(j + 1);
    }
    cp_async_wait<0>();
    tc::fence_proxy_async_smem();
    tc::mbar_arrive(bar_kv_full((n_kv - 1) & 1));
  } else if (warp == 4) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t fmt = sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      constexpr uint32_t idesc_qk = tc::make_idesc(fmt, fmt, 128, kKvTile, 0, 0);
      constexpr uint32_t idesc_pv = tc::make_idesc(fmt, fmt, 128, D, 0, 1);  // V: MN-major B
      const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + C::O_COL;
      auto issue_qk = [&](int j) {
        const uint32_t sk = sbase + C::OFF_K + (j & 1) * C::KV_TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t koff = (kk >> 2) * (128 * 128) + (kk & 3) * 32;   // Q panel / k-step
          const uint32_t boff = (kk >> 2) * (kKvTile * 128) + (kk & 3) * 32;
          tc::mma_f16(tmem_s, tc::make_smem_desc_sw128(sbase + C::OFF_Q + koff, 1024, 16),
                      tc::make_smem_desc_sw128(sk + boff, 1024, 16), idesc_qk, kk != 0 ? 1u : 0u);
        }
        tc::mma_commit(bar_s_full);
      };
      tc::mbar_wait(bar_q, 0);
      tc::mbar_wait(bar_kv_full(0), 0);
      tc::tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) {
          tc::mbar_wait(bar_kv_full((j + 1) & 1), ((j + 1) >> 1) & 1u);
          tc::mbar_wait(bar_s_free, j & 1u);  // softmax has pulled S(j) out of TMEM
          tc::tc_fence_after();
          issue_qk(j + 1);
        }
        tc::mbar_wait(bar_p_full, j & 1u);
        tc::tc_fence_after();
        const uint32_t sv = sbase + C::OFF_V + (j & 1) * C::KV_TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < kKvTile / 16; ++kk) {
          // A = P [128 x 64 tokens], K-major; B = V rows kk*16.., MN-major: LBO = panel stride
          tc::mma_f16(tmem_o, tc::make_smem_desc_sw128(sbase + C::OFF_P + kk * 32, 1024, 16),
                      tc::make_smem_desc_sw128(sv + kk * 16 * 128, 1024, kKvTile * 128), idesc_pv,
                      (j | kk) != 0 ? 1u : 0u);
        }
        tc::mma_commit(bar_pv_full);
        tc::mma_commit(bar_kv_empty(j & 1));
      }
    }
  } else {
    // ===================== softmax / epilogue (warps 0-3, thread = query row) =====================
    const int r = warp * 32 + lane;
    const int tok = tok0 + r / p.pack;
    const bool row_ok = tok < ext;
    const int limit = p.causal ? prefix + min(tok, ext - 1) : kv_len - 1;  // last visible key
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t tmem_s = tmem_base + lane_addr, tmem_o = tmem_base + lane_addr + C::O_COL;
    const bool capped = p.logit_cap > 0.f;
    const float inv_cap = capped ? 1.f / p.logit_cap : 0.f;
    const float cap_log2 = p.logit_cap * 1.4426950408889634f;
    float m = -INFINITY, l = 0.f;

    auto score = [&](uint32_t raw, int kpos) -> float {
      float s = __uint_as_float(raw) * p.qk_scale;
      if (capped) s = cap_log2 * fast_tanh(s * inv_cap);
      return kpos <= limit ? s : -INFINITY;
    };

    for (int j = 0; j < n_kv; ++j) {
      tc::mbar_wait(bar_s_full, j & 1u);
      tc::tc_fence_after();
      // ---- S(j) -> registers in ONE TMEM read; the accumulator is handed back at once so the MMA
      //      warp issues QK(j+1) while this tile's max / exp2 / pack run from registers
      uint32_t v[2][32];
      tc::tmem_ld32(tmem_s, v[0]);
      tc::tmem_ld32(tmem_s + 32, v[1]);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(bar_s_free);
      float mt = -INFINITY;
      if constexpr (FAST) {
        const int k0 = j * kKvTile;
        if (k0 + kKvTile - 1 > limit) {  // the tile crosses this row's last visible key
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (k0 + h * 32 + c > limit) v[h][c] = 0xff800000u;  // -inf
        }
        float a0 = -INFINITY, a1 = -INFINITY, a2 = -INFINITY, a3 = -INFINITY;
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int c = 0; c < 32; c += 8) {
            a0 = fmax3(a0, __uint_as_float(v[h][c]), __uint_as_float(v[h][c + 1]));
            a1 = fmax3(a1, __uint_as_float(v[h][c + 2]), __uint_as_float(v[h][c + 3]));
            a2 = fmax3(a2, __uint_as_float(v[h][c + 4]), __uint_as_float(v[h][c + 5]));
            a3 = fmax3(a3, __uint_as_float(v[h][c + 6]), __uint_as_float(v[h][c + 7]));
          }
        mt = fmaxf(fmaxf(a0, a1), fmaxf(a2, a3)) * p.qk_scale;  // qk_scale > 0: -inf stays -inf
      } else {
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const float sc = score(v[h][c], j * kKvTile + h * 32 + c);
            v[h][c] = __float_as_uint(sc);
            mt = fmaxf(mt, sc);
          }
      }
      const float m_new = fmaxf(m, mt);
      const bool need = m_new > m + kRescaleThreshold;  // also true for the first finite max
      bool pv_waited = false;
      if (__any_sync(0xffffffffu, need)) {
        if (j > 0) {
          tc::mbar_wait(bar_pv_full, (j - 1) & 1u);  // O holds tiles 0..j-1
          tc::tc_fence_after();
          pv_waited = true;
          const float alpha = (m == -INFINITY) ? 0.f : fast_exp2(m - m_new);
#pragma unroll 1
          for (int cc = 0; cc < D / 32; ++cc) {
            uint32_t o[32];
            tc::tmem_ld32(tmem_o + cc * 32, o);
            tc::tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
            tc::tmem_st32(tmem_o + cc * 32, o);
          }
          tc::tmem_st_wait();
          l *= alpha;
        }
        m = m_new;
      }
      const float m_use = (m == -INFINITY) ? 0.f : m;
      // ---- P = exp2(S - m), 16-bit, packed in registers first
      uint32_t packed[2][16];
      const float2 sc2 = make_float2(p.qk_scale, p.qk_scale), nm2 = make_float2(-m_use, -m_use);
      const float2 one2 = make_float2(1.f, 1.f);
      float2 lsum = make_float2(0.f, 0.f);
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          float p0, p1;
          if constexpr (FAST) {
            const float2 x = ffma2(make_float2(__uint_as_float(v[h][2 * c]),
                                               __uint_as_float(v[h][2 * c + 1])), sc2, nm2);
            p0 = fast_exp2(x.x);
            p1 = fast_exp2(x.y);
            lsum = ffma2(make_float2(p0, p1), one2, lsum);
          } else {
            p0 = fast_exp2(__uint_as_float(v[h][2 * c]) - m_use);
            p1 = fast_exp2(__uint_as_float(v[h][2 * c + 1]) - m_use);
            l += p0 + p1;
          }
          if (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) {
            __nv_bfloat162 t2 = __floats2bfloat162_rn(p0, p1);
            packed[h][c] = *reinterpret_cast<uint32_t*>(&t2);
          } else {
            __half2 t2 = __floats2half2_rn(p0, p1);
            packed[h][c] = *reinterpret_cast<uint32_t*>(&t2);
          }
        }
      if constexpr (FAST) l += lsum.x + lsum.y;
      if (j > 0 && !pv_waited) {
        tc::mbar_wait(bar_pv_full, (j - 1) & 1u);  // PV(j-1) has finished reading P(j-1)
        tc::tc_fence_after();
      }
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int chunk = h * 4 + i;
          const uint32_t addr = sbase + C::OFF_P + r * 128 + ((chunk ^ (r & 7)) << 4);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(addr), "r"(packed[h][4 * i]),
                       "r"(packed[h][4 * i + 1]), "r"(packed[h][4 * i + 2]), "r"(packed[h][4 * i + 3])
                       : "memory");
        }
      tc::fence_proxy_async_smem();
      tc::tc_fence_before();
      tc::mbar_arrive(bar_p_full);
    }

    // ---- epilogue: O / l -> out
    tc::mbar_wait(bar_pv_full, (n_kv - 1) & 1u);
    tc::tc_fence_after();
    const float inv_l = l > 0.f ? p.out_scale / l : 0.f;
    T* orow = static_cast<T*>(p.o) + (q_row0 + (row_ok ? tok : 0)) * p.o_stride +
              static_cast<int64_t>(head_base + r % p.pack) * D;
#pragma unroll 1
    for (int cc = 0; cc < D / 32; ++cc) {
      uint32_t v[32];
      tc::tmem_ld32(tmem_o + cc * 32, v);
      tc::tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float a = __uint_as_float(v[i * 8 + 2 * e]) * inv_l;
            const float c2 = __uint_as_float(v[i * 8 + 2 * e + 1]) * inv_l;
            if (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) {
              __nv_bfloat162 t2 = __floats2bfloat162_rn(a, c2);
              w[e] = *reinterpret_cast<uint32_t*>(&t2);
            } else {
              __half2 t2 = __floats2half2_rn(a, c2);
              w[e] = *reinterpret_cast<uint32_t*>(&t2);
            }
          }
          *reinterpret_cast<uint4*>(orow + cc * 32 + i * 8) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 4) tc::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

// B200_EXTEND_FAST_SOFTMAX=0 selects the round-1 softmax for every call (read per launch: A/B tools flip it)
static bool extend_fast_softmax_enabled() {
  const char* e = std::getenv("B200_EXTEND_FAST_SOFTMAX");
  return !(e && e[0] == '0');
}

template <typename T, int D, bool FAST>
static int launch_extend_inst(const ExtendParams& p, int bs, int max_extend_len, cudaStream_t st) {
  using C = ExtendCfg<D>;
  auto kern = attn_extend_kernel<T, D, FAST>;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      C::SMEM_BYTES));
    attr_done = true;
  }
  const int tok_per_tile = 128 / p.pack;
  dim3 grid(ceil_div(max_extend_len, tok_per_tile), p.Hkv * p.head_tiles, bs);
  {
    ScopedKernelTimer timer("attn_extend", st);
    kern<<<grid, kExtThreads, C::SMEM_BYTES, st>>>(p);
  }
  B200_CHECK_LAUNCH("attn_extend_kernel");
  return 0;
}

template <typename T, int D>
static int launch_extend(const ExtendParams& p, int bs, int max_extend_len, cudaStream_t st) {
  const bool fast = p.logit_cap <= 0.f && p.qk_scale > 0.f && extend_fast_softmax_enabled();
  note_variant("attn_extend d=%d fast_softmax=%d", D, fast ? 1 : 0);
  return fast ? launch_extend_inst<T, D, true>(p, bs, max_extend_len, st)
              : launch_extend_inst<T, D, false>(p, bs, max_extend_len, st);
}

}  // namespace b200

extern "C" {

int b200_attn_extend(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                     int64_t kv_row_stride, void* o, int64_t o_stride,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const int64_t* req_pool_indices, const int64_t* seq_lens,
                     const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens,
                     const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq, int Hkv,
                     int D, int Dv, float scale, float logit_cap, int causal, int dtype,
                     void* stream) {
  return b200_attn_extend_scaled(q, q_stride, k_cache, v_cache, kv_row_stride, o, o_stride,
                                 req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                                 extend_prefix_lens, extend_seq_lens, extend_start_loc, bs,
                                 max_extend_len, Hq, Hkv, D, Dv, scale, logit_cap, 1.0f, causal,
                                 dtype, stream);
}

int b200_attn_extend_scaled(const void* q, int64_t q_stride, const void* k_cache,
                            const void* v_cache, int64_t kv_row_stride, void* o, int64_t o_stride,
                            const int32_t* req_to_token, int64_t req_to_token_stride,
                            const int64_t* req_pool_indices, const int64_t* seq_lens,
                            const int32_t* extend_prefix_lens, const int32_t* extend_seq_lens,
                            const int32_t* extend_start_loc, int bs, int max_extend_len, int Hq,
                            int Hkv, int D, int Dv, float scale, float logit_cap, float out_scale,
                            int causal, int dtype, void* stream) {
  using namespace b200;
  if (bs == 0 || max_extend_len == 0) return 0;
  B200_REQUIRE(q && k_cache && v_cache && o && req_to_token && req_pool_indices && seq_lens &&
                   extend_prefix_lens && extend_seq_lens && extend_start_loc,
               "b200_attn_extend: null pointer");
  B200_REQUIRE(bs > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && max_extend_len > 0,
               "b200_attn_extend: bad config bs=%d Hq=%d Hkv=%d max_extend_len=%d", bs, Hq, Hkv,
               max_extend_len);
  B200_REQUIRE(D == Dv && (D == 64 || D == 128),
               "b200_attn_extend: head dims %d/%d unsupported (qk == v in {64, 128})", D, Dv);
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, "b200_attn_extend: dtype %d", dtype);
  B200_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && aligned16(o) &&
                   q_stride % 8 == 0 && o_stride % 8 == 0 && kv_row_stride % 8 == 0,
               "b200_attn_extend: q/o/k_cache/v_cache must be 16-byte aligned, strides %% 8 == 0");
  ExtendParams p;
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
  p.prefix_lens = extend_prefix_lens;
  p.extend_lens = extend_seq_lens;
  p.extend_start = extend_start_loc;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.group = Hq / Hkv;
  // heads of one GQA group packed into the 128-row tile: largest power of two dividing the group
  int pack = 1;
  while (pack < 8 && p.group % (pack * 2) == 0) pack *= 2;
  p.pack = pack;
  p.head_tiles = p.group / pack;
  p.logit_cap = logit_cap;
  p.out_scale = out_scale;
  p.qk_scale = logit_cap > 0.f ? scale : scale * 1.4426950408889634f;
  p.causal = causal;
  B200_REQUIRE(bs <= 65535 && Hkv * p.head_tiles <= 65535, "b200_attn_extend: grid too large");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16)
    return D == 128 ? launch_extend<__nv_bfloat16, 128>(p, bs, max_extend_len, st)
                    : launch_extend<__nv_bfloat16, 64>(p, bs, max_extend_len, st);
  return D == 128 ? launch_extend<__half, 128>(p, bs, max_extend_len, st)
                  : launch_extend<__half, 64>(p, bs, max_extend_len, st);
}

}  // extern "C"

// tcgen05 / TMEM / TMA GEMM for the DeepFM MLP (forward, dgrad, wgrad) on sm_100a.
//
//   D[M,N] = sum_k A(m,k) * B(n,k)        bf16 operands, fp32 accumulation in tensor memory
//
// One kernel template covers the three passes of `tf.contrib.layers.fully_connected`
// (reference PS:212-214 / HVD:204-206 forward; its autograd for backward) by choosing operand majors:
//   forward  Y  = act(X  W^T + b)   A = X  [B,in]  K-major          B = W [out,in] K-major
//   dgrad    dX = dZ W              A = dZ [B,out] K-major          B = W [out,in] stored [K,N] -> N-major
//   wgrad    dW = dZ^T X            A = dZ [B,out] stored [K,M] -> M-major   B = X [B,in] stored [K,N] -> N-major
// so no transposed copies of activations or weights are ever materialised.
//
// Structure (one 128 x BN output tile per CTA, optional split-K over grid.z):
//   warp 0      TMA producer: cp.async.bulk.tensor 2-D boxes, 128B swizzle, OOB zero fill handles
//               K = 1248 (19.5 k-blocks) and N = 400 without padding anything in global memory
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (UMMA 128 x BN x 16),
//               frees smem stages with tcgen05.commit -> mbarrier
//   warps 2-9   epilogue: tcgen05.ld 32x32b from TMEM, then fused
//                 EPI_BIAS_ACT : + bias, ReLU, inverted dropout (Philox, keep-prob, PS:218), bf16 store
//                 EPI_MASK     : * (relu'/dropout mask of the previous layer, recovered as X_prev != 0) * 1/keep
//                 EPI_ACCUM_F32: red.global.add.v4.f32 into the flat fp32 gradient buffer (split-K safe)
#include "common.cuh"
#include "philox.cuh"

#include <mutex>

namespace dfm {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kEpiWarps = 8;  // 2 warps per TMEM lane quarter
constexpr int kGemmThreads = 64 + 32 * kEpiWarps;

enum { EPI_BIAS_ACT = 0, EPI_MASK = 1, EPI_ACCUM_F32 = 2, EPI_MASK_BIG = 3 };  // MASK_BIG: same epilogue, whole-SM stage budget

struct GemmParams {
  int M, N, K;
  int num_k_blocks;
  int k_blocks_per_split;
  void* out;
  int ldo;
  const float* bias;
  int relu;
  DropoutCfg drop;
  const unsigned long long* step_ptr;  // device-resident step counter (CUDA-graph friendly)
  const __nv_bfloat16* mask;
  int ldmask;
  float mask_scale;
  int out_f32;  // EPI_BIAS_ACT only: store fp32 (pre-batch-norm activations must not be rounded to bf16)
  int trace_id; // timeline id (trace build only; -1 = none)
  int overwrite;  // EPI_ACCUM_F32 only: plain stores instead of red.add (needs a single K split)
  float* bias_out;  // EPI_ACCUM_F32 only: output column N (one past the weight gradient) goes here instead: with a column
                    // of ones appended to the B operand it is sum_k A[k, m] = the bias gradient, on the tensor core for free
  float* colsum;  // EPI_MASK only: colsum[n] += sum over rows of the (masked, scaled) output = the bias gradient of the
                  // layer that produced the mask (reference: BiasAddGrad); null = off
};

// Shared-memory budget per pass (the rest of the step decides it, not the GEMM itself — scripts/step_trace.py):
//   forward  runs alone on the chain: the whole SM (~200 KB, up to 7 stages). With 20 k-blocks of 26 KB behind a ~1.2 us
//            TMA round trip, the bytes in flight are the throughput: 4 stages made mlp_fwd0 1.3 us slower than 7.
//   dgrad    of the upper layers runs alone too (the wgrads are forked after dgrad_1): whole SM (EPI_MASK_BIG);
//   dgrad_0  ~124 KB   and   wgrad ~98 KB: one CTA of each fits on an SM TOGETHER. With ~200 KB each (the first version)
//            the wgrad branch queued for whole SMs behind the dgrad chain, whose programmatic dependent launches keep the
//            next kernel's CTAs resident while they wait: wgrad2 took 23 us beside the chain, 2.3 us alone.
template <int EPI>
struct GemmBudget { static constexpr int BYTES = (EPI == 0 || EPI == 3) ? 200 * 1024 : (EPI == 1 ? 124 * 1024 : 98 * 1024); };

template <int BN, int EPI>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES_RAW = GemmBudget<EPI>::BYTES / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : (STAGES_RAW < 2 ? 2 : STAGES_RAW);
  static constexpr int TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + BN * 4 /*bias*/;
};

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmParams p) {
  using Cfg = GemmCfg<BN, EPI>;
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N for M=128 must be a multiple of 16 in [16,256]");
  static_assert(!B_MN || BN % 64 == 0, "MN-major B tiles are built from 64-wide 128B-swizzle atoms");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tmem_full_bar = empty_bar + Cfg::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + 256);  // [BN]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int kb_begin = blockIdx.z * p.k_blocks_per_split;
  const int kb_end = min(p.num_k_blocks, kb_begin + p.k_blocks_per_split);
  if (kb_begin >= kb_end) return;  // empty split (host avoids this; uniform across the CTA)

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      for (int s = 0; s < Cfg::STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // PDL: everything above (barrier init, descriptor prefetch, TMEM allocation) overlapped the previous kernel's tail.
  // The NEXT kernel is released late, when this CTA's accumulator is complete (see the epilogue): released at entry, its
  // CTAs sat resident for this kernel's whole run and kept registers and thread slots from the wgrad branch.
  pdl_wait(p.trace_id);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sA = smem + stage * Cfg::STAGE_BYTES;
        uint8_t* sB = sA + Cfg::A_BYTES;
        mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
        if constexpr (!A_MN) {
          tma_load_2d(sA, &tmA, &full_bar[stage], kb * BK, m0);
        } else {
#pragma unroll
          for (int g = 0; g < BM / 64; ++g) tma_load_2d(sA + g * (64 * BK * 2), &tmA, &full_bar[stage], m0 + g * 64, kb * BK);
        }
        if constexpr (!B_MN) {
          tma_load_2d(sB, &tmB, &full_bar[stage], kb * BK, n0);
        } else {
#pragma unroll
          for (int g = 0; g < BN / 64; ++g) tma_load_2d(sB + g * (64 * BK * 2), &tmB, &full_bar[stage], n0 + g * 64, kb * BK);
        }
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (single thread)
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + stage * Cfg::STAGE_BYTES);
        const uint32_t b_base = a_base + Cfg::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // K-major, 128B swizzle: 8-row groups 1024 B apart; a K step of 16 bf16 moves 32 B inside the swizzled row.
          // MN-major, 128B swizzle: 64-wide MN groups 8192 B apart (LBO), 8-row K groups 1024 B apart (SBO);
          //                         a K step of 16 rows moves 2048 B.
          const uint64_t adesc = A_MN ? make_smem_desc(a_base + k * 2048, 64 * BK * 2, 1024)
                                      : make_smem_desc(a_base + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc(b_base + k * 2048, 64 * BK * 2, 1024)
                                      : make_smem_desc(b_base + k * 32, 16, 1024);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);  // stage reusable once these MMAs have read it
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    // TMEM lane quarter = warp % 4 (hardware rule); kEpiWarps/4 warps share a quarter and interleave 16-column
    // chunks, so several independent dependency chains per scheduler hide the ALU/LDG latency of the epilogue.
    const int q = warp & 3;
    const int sub = (warp - 2) >> 2;
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < p.M;
    constexpr int NCHUNK = BN / 16;
    constexpr int WPQ = kEpiWarps / 4;
    constexpr int MYCH = (NCHUNK + WPQ - 1) / WPQ;
    const bool n_mult4 = (p.N & 3) == 0;
    DropoutCfg drop = p.drop;

    // ---- work that does not need the accumulator runs NOW, while TMA/MMA are still busy: stage the bias tile in shared
    // memory, draw the dropout keep-bits (Philox) / read the ReLU-dropout mask of the previous layer, 16 bits per chunk
    uint32_t bits[MYCH];
    bool fastc[MYCH];
    if constexpr (EPI == EPI_BIAS_ACT) {
      for (int j = threadIdx.x - 64; j < BN; j += 32 * kEpiWarps)
        s_bias[j] = (p.bias && n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
      if (drop.enabled && p.step_ptr) {
        const unsigned long long step = *p.step_ptr;
        drop.step_lo = (uint32_t)step;
        drop.step_hi = (uint32_t)(step >> 32);
      }
    }
#pragma unroll
    for (int k = 0; k < MYCH; ++k) {
      const int c = (sub + k * WPQ) * 16;
      const int n = n0 + c;
      bits[k] = 0xFFFFu;
      fastc[k] = (c < BN) && row_ok && (n + 16 <= p.N);
      if constexpr (EPI == EPI_BIAS_ACT) {
        fastc[k] = fastc[k] && n_mult4;
        if (fastc[k] && drop.enabled) {
          const uint64_t quad0 = ((uint64_t)(row + drop.row_offset) * (uint64_t)p.N + (uint64_t)n) >> 2;
          uint32_t b = 0;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const uint4 r = dropout_words(drop, quad0 + t);
            b |= (r.x < drop.threshold ? 1u : 0u) << (4 * t) | (r.y < drop.threshold ? 1u : 0u) << (4 * t + 1) |
                 (r.z < drop.threshold ? 1u : 0u) << (4 * t + 2) | (r.w < drop.threshold ? 1u : 0u) << (4 * t + 3);
          }
          bits[k] = b;
        }
      } else if constexpr (EPI == EPI_MASK || EPI == EPI_MASK_BIG) {
        fastc[k] = fastc[k] && (p.ldmask % 8 == 0);
        if (fastc[k] && p.mask) {
          const uint4* mp = reinterpret_cast<const uint4*>(p.mask + (size_t)row * p.ldmask + n);
          const uint4 m0v = __ldg(mp), m1v = __ldg(mp + 1);
          const uint32_t mw[8] = {m0v.x, m0v.y, m0v.z, m0v.w, m1v.x, m1v.y, m1v.z, m1v.w};
          uint32_t b = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            b |= ((mw[j] & 0x7FFFu) ? 1u : 0u) << (2 * j) | ((mw[j] & 0x7FFF0000u) ? 1u : 0u) << (2 * j + 1);
          bits[k] = b;
        }
      }
    }
    if constexpr (EPI == EPI_BIAS_ACT) asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");  // s_bias ready

    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    pdl_launch_dependents();  // one arrival per CTA is what counts; the dependents still wait for this grid to finish

#pragma unroll
    for (int k = 0; k < MYCH; ++k) {
      const int c = (sub + k * WPQ) * 16;
      if (c >= BN) continue;  // warp-uniform
      float v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
      const int n = n0 + c;
      int n_lim = p.N;
      if constexpr (EPI == EPI_ACCUM_F32) n_lim += p.bias_out ? 1 : 0;  // the extra "ones" column
      if (n >= n_lim) continue;  // warp-uniform
      const bool full16 = (n + 16 <= p.N);
      if constexpr (EPI == EPI_MASK || EPI == EPI_MASK_BIG) {
        if (p.colsum) {  // warp-uniform: every lane takes part in the shuffles below, rows past M contribute zeros
          float m[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            bool keep = row_ok && (n + j < p.N);
            if (p.mask && keep) {
              if (fastc[k]) keep = (bits[k] >> j) & 1u;
              else keep = __bfloat162float(p.mask[(size_t)row * p.ldmask + n + j]) != 0.f;
            }
            m[j] = keep ? v[j] * (p.mask ? p.mask_scale : 1.f) : 0.f;
          }
          // transpose-reduce over the warp's 32 rows: each step halves the columns a lane still carries
#pragma unroll
          for (int h = 8, o = 16; h >= 1; h >>= 1, o >>= 1) {
            const bool upper = (lane & o) != 0;
#pragma unroll
            for (int j = 0; j < h; ++j) {
              const float send = upper ? m[j] : m[j + h];
              const float recv = __shfl_xor_sync(0xffffffffu, send, o);
              m[j] = (upper ? m[j + h] : m[j]) + recv;
            }
          }
          m[0] += __shfl_xor_sync(0xffffffffu, m[0], 1);
          // lane bits 4..1 select the column: bit4 -> +8, bit3 -> +4, bit2 -> +2, bit1 -> +1
          const int col = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
          if ((lane & 1) == 0 && n + col < p.N) atomicAdd(p.colsum + n + col, m[0]);
        }
      }
      if (!row_ok) continue;

      if constexpr (EPI == EPI_BIAS_ACT) {
        if (fastc[k]) {
          const float scale = drop.enabled ? drop.scale : 1.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float x = v[j] + s_bias[c + j];
            if (p.relu) x = fmaxf(x, 0.f);
            v[j] = ((bits[k] >> j) & 1u) ? x * scale : 0.f;
          }
        } else {
          uint64_t last_quad = ~0ull;
          uint4 rnd = make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int nn = n + j;
            float x = v[j];
            if (nn < p.N) {
              x += s_bias[c + j];
              if (p.relu) x = fmaxf(x, 0.f);
              if (drop.enabled) {
                const uint64_t elem = (uint64_t)(row + drop.row_offset) * (uint64_t)p.N + (uint64_t)nn;
                const uint64_t quad = elem >> 2;
                if (quad != last_quad) { rnd = dropout_words(drop, quad); last_quad = quad; }
                const uint32_t sel = (uint32_t)(elem & 3);
                const uint32_t u = sel == 0 ? rnd.x : sel == 1 ? rnd.y : sel == 2 ? rnd.z : rnd.w;
                x = (u < drop.threshold) ? x * drop.scale : 0.f;
              }
            }
            v[j] = x;
          }
        }
      } else if constexpr (EPI == EPI_MASK || EPI == EPI_MASK_BIG) {
        if (p.mask) {
          if (fastc[k]) {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = ((bits[k] >> j) & 1u) ? v[j] * p.mask_scale : 0.f;
          } else {
            const __nv_bfloat16* mrow = p.mask + (size_t)row * p.ldmask + n;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (n + j < p.N) {
                const float mv = __bfloat162float(mrow[j]);
                v[j] = (mv != 0.f) ? v[j] * p.mask_scale : 0.f;
              }
            }
          }
        }
      }

      if constexpr (EPI == EPI_ACCUM_F32) {
        float* orow = reinterpret_cast<float*>(p.out) + (size_t)row * p.ldo + n;
        if (p.overwrite) {
          if (full16 && (p.ldo % 4 == 0)) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<float4*>(orow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (n + j < p.N) orow[j] = v[j];
              else if (p.bias_out && n + j == p.N) p.bias_out[row] = v[j];
            }
          }
        } else if (full16 && (p.ldo % 4 == 0)) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) red_add_f32x4(orow + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (n + j < p.N) atomicAdd(orow + j, v[j]);
            else if (p.bias_out && n + j == p.N) atomicAdd(p.bias_out + row, v[j]);
          }
        }
      } else if (EPI == EPI_BIAS_ACT && p.out_f32) {
        float* orow = reinterpret_cast<float*>(p.out) + (size_t)row * p.ldo + n;
        if (full16 && (p.ldo % 4 == 0)) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *reinterpret_cast<float4*>(orow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n + j < p.N) orow[j] = v[j];
        }
      } else {
        __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + n;
        if (full16 && (p.ldo % 8 == 0)) {
          uint4 o0, o1;
          o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
          o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
          o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
          o1.z = pack_bf16x2(v[12], v[13]); o1.w = pack_bf16x2(v[14], v[15]);
          reinterpret_cast<uint4*>(orow)[0] = o0;
          reinterpret_cast<uint4*>(orow)[1] = o1;
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n + j < p.N) orow[j] = __float2bfloat16(v[j]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  trace_end(p.trace_id);
  if (warp == 0) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

// bf16 row-major matrix [rows, cols] with leading dimension ld (elements); box = {64 inner, box_rows}.
int make_tmap(CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows) {  // declared in common.cuh
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    dfm_set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return DFM_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    dfm_set_error("cuTensorMapEncodeTiled failed (%d): base=%p rows=%d cols=%d ld=%d box_rows=%d", (int)r, base, rows,
                  cols, ld, box_rows);
    return DFM_ERR_CUDA;
  }
  return DFM_OK;
}

struct GemmPlan {
  CUtensorMap tmA, tmB;
  GemmParams p;
  int exclusive = 0;  // dgrad: runs with nothing beside it -> whole-SM stage budget
  int mode;  // 0 fwd (KK), 1 dgrad (K,MN), 2 wgrad (MN,MN)
  int bn;
  dim3 grid;
};

template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch_gemm(const GemmPlan& pl, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, EPI>;
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, EPI>;
  static bool attr_set[64] = {false};
  int dev = 0;
  DFM_CHECK(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DFM_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    // largest carve-out, so a dgrad CTA and a wgrad CTA (two kernels, 124 + 98 KB) can share an SM
    DFM_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    attr_set[dev] = true;
  }
  DFM_CHECK(launch_k(kern, pl.grid, dim3(kGemmThreads), (size_t)Cfg::SMEM_BYTES, stream, pl.tmA, pl.tmB, pl.p));
  return DFM_OK;
}

static int pick_bn(int N, const int* cands, int ncand, int m_tiles, bool prefer_wide = false) {
  int best = cands[0];
  long best_cost = -1;
  for (int i = 0; i < ncand; ++i) {
    const int bn = cands[i];
    const int tiles = ceil_div(N, bn);
    // cost ~ padded columns (+ a small per-tile overhead); when few CTAs would result prefer more, smaller tiles
    // cost ~ padded columns; a per-tile overhead only once the grid exceeds one wave — below that, more (smaller) tiles
    // mean more SMs working and a shorter epilogue per CTA
    long cost = (long)tiles * bn * 8;
    if (prefer_wide || (long)tiles * m_tiles > 148) cost += (long)tiles * 64;  // wgrad fills the SMs with split-K instead
    else cost += (long)bn / 8;
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// mode 0: forward   out[M,N] bf16 = act(A[M,K] * B[N,K]^T + bias) with dropout
// mode 1: dgrad     out[M,N] bf16 = (A[M,K] * B[K,N]) * (mask != 0) * mask_scale
// mode 2: wgrad     out[M,N] f32 += A[K,M]^T * B[K,N]          (atomic accumulate; split-K). `bias` non-null in this mode:
//                   B has a column of ones at index N (ldb > N) and bias[m] receives sum_k A[k,m] (the bias gradient)
// lda/ldb/ldo/ldmask are leading dimensions in elements; all bf16 leading dims must be multiples of 8.
int dfm_gemm_plan_create(void** plan_out, int mode, const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                         void* out, int ldo, const float* bias, int relu, float keep_prob, unsigned long long seed,
                         int layer, int row_offset, const unsigned long long* step_ptr, const void* mask, int ldmask,
                         float mask_scale, int split_k) {
  DFM_REQUIRE(mode >= 0 && mode <= 2, "gemm mode must be 0 (fwd), 1 (dgrad) or 2 (wgrad)");
  DFM_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "bf16 leading dimensions must be multiples of 8 (TMA 16-byte stride rule)");
  DFM_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
  GemmPlan* pl = new GemmPlan();
  pl->mode = mode;
  const int m_tiles = ceil_div(M, BM);
  static const int kk_cands[] = {32, 64, 80, 96, 128, 256};
  static const int mn_cands[] = {64, 128, 256};
  pl->bn = mode == 0 ? pick_bn(N, kk_cands, 6, m_tiles) : pick_bn(N, mn_cands, 3, m_tiles, mode == 2);
  if (mode == 2 && split_k == 1) pl->bn = 64;  // one K split: narrow tiles are the only source of CTAs
  int rc;
  if (mode == 0) {  // A [M,K] K-major, B [N,K] K-major
    rc = make_tmap(&pl->tmA, A, M, K, lda, BM);
    if (!rc) rc = make_tmap(&pl->tmB, B, N, K, ldb, pl->bn);
  } else if (mode == 1) {  // A [M,K] K-major, B stored [K,N]
    rc = make_tmap(&pl->tmA, A, M, K, lda, BM);
    if (!rc) rc = make_tmap(&pl->tmB, B, K, N, ldb, BK);
  } else {  // A stored [K,M], B stored [K,N]; with a bias output, B carries a column of ones at index N (ldb > N)
    DFM_REQUIRE(bias == nullptr || ldb > N, "wgrad with a bias output needs the ones column: ldb must exceed N");
    rc = make_tmap(&pl->tmA, A, K, M, lda, BK);
    if (!rc) rc = make_tmap(&pl->tmB, B, K, bias ? N + 1 : N, ldb, BK);
  }
  if (rc) { delete pl; return rc; }
  GemmParams& p = pl->p;
  p.M = M; p.N = N; p.K = K;
  p.num_k_blocks = ceil_div(K, BK);
  int splits = 1;
  if (mode == 2) {
    splits = split_k > 0 ? split_k : 1;
    if (split_k <= 0) {  // auto: fill ~148 SMs
      const int tiles = m_tiles * ceil_div(N, pl->bn);
      splits = tiles >= 148 ? 1 : 148 / tiles;
      if (splits < 1) splits = 1;
    }
    if (splits > p.num_k_blocks) splits = p.num_k_blocks;
  }
  p.k_blocks_per_split = ceil_div(p.num_k_blocks, splits);
  splits = ceil_div(p.num_k_blocks, p.k_blocks_per_split);  // no empty splits
  p.out = out; p.ldo = ldo; p.bias = mode == 2 ? nullptr : bias; p.relu = relu;
  p.bias_out = mode == 2 ? const_cast<float*>(bias) : nullptr;
  p.drop.enabled = (mode == 0 && keep_prob < 1.0f) ? 1u : 0u;
  p.drop.threshold = keep_prob >= 1.0f ? 0xFFFFFFFFu : (uint32_t)floor((double)keep_prob * 4294967296.0);
  p.drop.scale = keep_prob >= 1.0f ? 1.0f : 1.0f / keep_prob;
  p.drop.seed_lo = (uint32_t)seed; p.drop.seed_hi = (uint32_t)(seed >> 32);
  p.drop.step_lo = p.drop.step_hi = 0; p.drop.layer = (uint32_t)layer; p.drop.row_offset = (uint32_t)row_offset;
  p.step_ptr = step_ptr;
  p.mask = reinterpret_cast<const __nv_bfloat16*>(mask); p.ldmask = ldmask; p.mask_scale = mask_scale;
  p.out_f32 = 0;
  p.trace_id = -1;
  p.colsum = nullptr;
  p.overwrite = 0;
  pl->grid = dim3(m_tiles, ceil_div(N + (p.bias_out ? 1 : 0), pl->bn), splits);
  *plan_out = pl;
  return DFM_OK;
}

int dfm_gemm_plan_info(void* plan, int* bn, int* grid_x, int* grid_y, int* grid_z) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  *bn = pl->bn; *grid_x = pl->grid.x; *grid_y = pl->grid.y; *grid_z = pl->grid.z;
  return DFM_OK;
}

// Forward plans only: write the epilogue result as fp32 (ldo then counts fp32 elements). Used by the batch-norm path.
int dfm_gemm_plan_set_out_f32(void* plan, int on) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  DFM_REQUIRE(pl->mode == 0, "out_f32 applies to forward plans only");
  pl->p.out_f32 = on ? 1 : 0;
  return DFM_OK;
}

// wgrad plans created with split_k = 1: out = A^T B with plain stores instead of out += with red.add. In the training
// step the split-K atomics of three wgrads (~1.2 M vector reds) queued in L2 beside emb_bwd's and delayed every kernel
// boundary after them (scripts/step_trace.py); one split per tile needs none and leaves most SMs to the dgrad chain.
int dfm_gemm_plan_set_overwrite(void* plan, int on) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  DFM_REQUIRE(pl->mode == 2, "overwrite applies to wgrad plans only");
  DFM_REQUIRE(!on || pl->grid.z == 1, "overwrite needs a single K split (create the plan with split_k = 1)");
  pl->p.overwrite = on ? 1 : 0;
  return DFM_OK;
}

// dgrad plans: 1 = this GEMM runs with no other GEMM beside it (whole-SM pipeline: all k-blocks of K <= 512 in flight).
int dfm_gemm_plan_set_exclusive(void* plan, int on) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  DFM_REQUIRE(pl->mode == 1, "exclusive applies to dgrad plans only");
  pl->exclusive = on ? 1 : 0;
  return DFM_OK;
}

// dgrad plans: also accumulate the column sums of the output (= bias gradient of the layer below) into colsum[N] (fp32).
int dfm_gemm_plan_set_colsum(void* plan, float* colsum) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  DFM_REQUIRE(pl->mode == 1, "colsum applies to dgrad plans only");
  pl->p.colsum = colsum;
  return DFM_OK;
}

// Timeline id of this plan's launches (trace build, scripts/step_trace.py); ignored by the normal build.
int dfm_gemm_plan_set_trace_id(void* plan, int id) {
  reinterpret_cast<GemmPlan*>(plan)->p.trace_id = id;
  return DFM_OK;
}

int dfm_gemm_plan_set_row_offset(void* plan, int row_offset) {
  reinterpret_cast<GemmPlan*>(plan)->p.drop.row_offset = (uint32_t)row_offset;
  return DFM_OK;
}

void dfm_gemm_plan_destroy(void* plan) { delete reinterpret_cast<GemmPlan*>(plan); }

#define DFM_GEMM_CASE(BN_, AMN, BMN, EPI_) \
  case BN_: return launch_gemm<BN_, AMN, BMN, EPI_>(*pl, stream);

int dfm_gemm_run(void* plan, void* stream_) {
  GemmPlan* pl = reinterpret_cast<GemmPlan*>(plan);
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (pl->mode == 0) {
    switch (pl->bn) {
      DFM_GEMM_CASE(32, false, false, EPI_BIAS_ACT)
      DFM_GEMM_CASE(64, false, false, EPI_BIAS_ACT)
      DFM_GEMM_CASE(80, false, false, EPI_BIAS_ACT)
      DFM_GEMM_CASE(96, false, false, EPI_BIAS_ACT)
      DFM_GEMM_CASE(128, false, false, EPI_BIAS_ACT)
      DFM_GEMM_CASE(256, false, false, EPI_BIAS_ACT)
    }
  } else if (pl->mode == 1 && pl->exclusive) {
    switch (pl->bn) {
      DFM_GEMM_CASE(64, false, true, EPI_MASK_BIG)
      DFM_GEMM_CASE(128, false, true, EPI_MASK_BIG)
      DFM_GEMM_CASE(256, false, true, EPI_MASK_BIG)
    }
  } else if (pl->mode == 1) {
    switch (pl->bn) {
      DFM_GEMM_CASE(64, false, true, EPI_MASK)
      DFM_GEMM_CASE(128, false, true, EPI_MASK)
      DFM_GEMM_CASE(256, false, true, EPI_MASK)
    }
  } else {
    switch (pl->bn) {
      DFM_GEMM_CASE(64, true, true, EPI_ACCUM_F32)
      DFM_GEMM_CASE(128, true, true, EPI_ACCUM_F32)
      DFM_GEMM_CASE(256, true, true, EPI_ACCUM_F32)
    }
  }
  dfm_set_error("no GEMM instantiation for mode %d BN %d", pl->mode, pl->bn);
  return DFM_ERR_UNSUPPORTED;
}

}  // extern "C"

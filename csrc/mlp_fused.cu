// The whole MLP forward (reference: the `fully_connected` stack of PS:212-226 / HVD:204-218) in ONE kernel.
//
//   X_{l+1} = dropout(relu(X_l W_l^T + b_l)),  l = 0 .. L-1        bf16 operands, fp32 accumulation in tensor memory
//
// An MLP acts row by row, so a block of 128 batch rows can run through every layer without talking to any other block.
// A thread-block CLUSTER of C CTAs owns one 128-row block; CTA c computes output columns [c*BN, (c+1)*BN) of every layer
// (C*BN >= the widest layer). After layer l each CTA applies bias / ReLU / dropout to its slice, stores it to global
// memory (the backward pass needs it) and writes it — as bf16, in the 128B-swizzled K-major layout tcgen05.mma reads —
// into the shared memory of EVERY CTA of its cluster (st.shared::cluster, distributed shared memory). After a cluster
// barrier each CTA holds the full [128, H_l] activation block as the resident A operand of layer l+1 and only has to
// fetch its slice of W_{l+1} (prefetched by TMA while the epilogue of layer l runs).
//
// What it would replace: L kernels of csrc/gemm_tcgen05.cu (mode 0), i.e. per layer boundary a grid-completion flush, the
// next kernel's griddepcontrol.wait, a first TMA round trip for A and the pipeline fill (~3 us each at B = 1024,
// scripts/step_trace.py). Layer 0 (K = F*K, streamed from global memory through a 7-stage TMA ring) is unchanged.
//
// STATUS: correct (bit-identical to the per-layer chain, tests/test_kernels_gpu.py) and NOT used by the model. Measured on
// a B200 at the headline shape (scripts/fused_mlp_bench.py, in-kernel stamps): the DSMEM broadcast of one 128 x 400 bf16
// block to 5 CTAs takes 4-7 us per layer — distributed shared memory moves ~17-21 B/cycle/SM — against the ~3 us of the
// kernel boundary it removes: 32 us fused vs 16 us for the three per-layer kernels. Kept as the measured experiment behind
// that statement, and as the starting point if a later part has a faster SM-to-SM path.
//
//   warp 0      TMA producer (layer 0: A and B tiles through the stage ring; layers >= 1: the whole W_l slice at once)
//   warp 1      MMA issuer (one elected thread, tcgen05.mma.cta_group::1.kind::f16, UMMA 128 x BN x 16)
//   warps 2-9   epilogue (tcgen05.ld, bias + ReLU + Philox dropout, global store, DSMEM broadcast)
//
// Shared memory is ONE region used twice: layer 0 cuts it into 7 stages of (16 KB A + BN*128 B) bytes; from layer 1 on
// its first 7*16 KB hold the resident activation block (7 k-blocks of 64 columns: H <= 448) and the rest the W slice.
#include "common.cuh"
#include "philox.cuh"

#include <math.h>
#include <string.h>

namespace dfm {

constexpr int FBM = 128;
constexpr int FBK = 64;
constexpr int kFMaxLayers = 4;
constexpr int kFMaxKB = 7;  // k-blocks of the resident activation block: layer widths up to 448
constexpr int kFEpiWarps = 8;
constexpr int kFThreads = 64 + 32 * kFEpiWarps;

struct MlpFwdParams {
  int M, L, K0;
  int H[kFMaxLayers];                 // layer widths (multiples of 16)
  __nv_bfloat16* out[kFMaxLayers];    // activations of every layer in global memory
  int ldo[kFMaxLayers];               // their row strides in elements (multiples of 8)
  const float* bias[kFMaxLayers];
  DropoutCfg drop[kFMaxLayers];
  int relu;
  const unsigned long long* step_ptr;
  int trace_id;
  unsigned long long* dbg;  // optional: [cta][32] globaltimer stamps (development aid; null in production)
};

template <int BN>
struct FusedCfg {
  static constexpr int A_BYTES = FBM * FBK * 2;  // 16 KB
  static constexpr int B_BYTES = BN * FBK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = kFMaxKB;  // the ring and the resident layout cover the same bytes
  static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  static constexpr int TMEM_COLS = BN <= 64 ? 64 : 128;
  static constexpr int NBAR = 2 * STAGES + 2;  // full[], empty[], tmem_full, b_full
  static constexpr int SMEM_BYTES = RING_BYTES + 1024 /*align slack*/ + 256 /*barriers + tmem slot*/ + kFMaxLayers * BN * 4;
};

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}

template <int BN>
__global__ void __launch_bounds__(kFThreads, 1)
mlp_fwd_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW0,
                     const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
                     const __grid_constant__ CUtensorMap tmW3, const MlpFwdParams p) {
  using Cfg = FusedCfg<BN>;
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 128, "UMMA N for M=128: multiple of 16");
  static_assert(Cfg::B_BYTES % 1024 == 0, "W k-blocks must keep the 1024-byte swizzle-atom alignment");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_res = smem;                                  // [kFMaxKB][128 rows][64 cols] bf16, swizzled, layers >= 1
  uint8_t* b_res = smem + kFMaxKB * Cfg::A_BYTES;         // [kFMaxKB][BN rows][64 cols] bf16, swizzled, layers >= 1
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tmem_full_bar = empty_bar + Cfg::STAGES;
  uint64_t* b_full_bar = tmem_full_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_full_bar + 1);
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::RING_BYTES + 256);  // [kFMaxLayers][BN]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int crank = (int)cluster_ctarank();    // column slice of this CTA
  const int csize = (int)cluster_nctarank();
  const int m0 = blockIdx.y * FBM;
  const int n0 = crank * BN;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW0);
      tma_prefetch_desc(&tmW1);
      tma_prefetch_desc(&tmW2);
      tma_prefetch_desc(&tmW3);
      for (int s = 0; s < Cfg::STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
      mbar_init(b_full_bar, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // every CTA of the cluster is running (its shared memory exists) before anybody writes into a peer
  cluster_sync_all();
  pdl_wait(p.trace_id);

  // a CTA whose slice lies beyond a layer's width idles through that layer (it still takes part in the barriers)
  auto active = [&](int l) -> bool { return n0 < p.H[l]; };
  auto nkb_of = [&](int l) -> int { return ((l == 0 ? p.K0 : p.H[l - 1]) + FBK - 1) / FBK; };
  const int cta_lin = blockIdx.y * gridDim.x + blockIdx.x;
  auto stamp = [&](int slot) { if (p.dbg) p.dbg[cta_lin * 32 + slot] = global_timer_ns(); };
  auto tmw = [&](int l) -> const CUtensorMap* { return l == 0 ? &tmW0 : l == 1 ? &tmW1 : l == 2 ? &tmW2 : &tmW3; };

  if (warp == 0) {
    // ================================================================== TMA producer
    uint32_t full_phase = 0;  // parity of tmem_full_bar for the next ACTIVE layer's completion
    if (active(0)) {
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        const int nkb = nkb_of(0);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sA = smem + stage * Cfg::STAGE_BYTES;
          mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
          tma_load_2d(sA, &tmX, &full_bar[stage], kb * FBK, m0);
          tma_load_2d(sA + Cfg::A_BYTES, &tmW0, &full_bar[stage], kb * FBK, n0);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
      __syncwarp();
    }
    uint32_t b_loads = 0;
#pragma unroll
    for (int l = 1; l < kFMaxLayers; ++l) {
      if (l < p.L) {
        if (active(l - 1)) {  // this CTA's own MMAs of layer l-1 are what still reads its ring / W slice
          if (elect_one()) mbar_wait(tmem_full_bar, full_phase);
          __syncwarp();
          full_phase ^= 1;
        }
        if (active(l)) {
          if (elect_one()) {
            const int nkb = nkb_of(l);
            mbar_expect_tx(b_full_bar, (uint32_t)nkb * Cfg::B_BYTES);
            for (int kb = 0; kb < nkb; ++kb) tma_load_2d(b_res + kb * Cfg::B_BYTES, tmw(l), b_full_bar, kb * FBK, n0);
          }
          __syncwarp();
          ++b_loads;
        }
        cluster_sync_all();  // B1: every CTA's MMAs of layer l-1 are complete
        cluster_sync_all();  // B2: every CTA's slice of X_l has landed in everybody's shared memory
      }
    }
    (void)b_loads;
  } else if (warp == 1) {
    // ================================================================== MMA issuer
    constexpr uint32_t idesc = make_idesc_bf16(FBM, BN, 0, 0);
    if (active(0)) {
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        const int nkb = nkb_of(0);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t b_base = a_base + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < FBK / 16; ++k)
            umma_bf16(tmem_base, make_smem_desc(a_base + k * 32, 16, 1024), make_smem_desc(b_base + k * 32, 16, 1024),
                      idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(tmem_full_bar);
      }
      __syncwarp();
    }
    uint32_t b_phase = 0;
#pragma unroll
    for (int l = 1; l < kFMaxLayers; ++l) {
      if (l < p.L) {
        cluster_sync_all();  // B1
        cluster_sync_all();  // B2: X_l is resident here, and the epilogue warps have drained the accumulator
        if (active(l)) {
          if (elect_one()) {
            mbar_wait(b_full_bar, b_phase);
            fence_proxy_async();  // the activation block was written with ordinary stores; the tensor core reads it
            tc_fence_after();     // through the async proxy
            const int nkb = nkb_of(l);
            const uint32_t a0 = smem_u32(a_res), b0 = smem_u32(b_res);
            for (int kb = 0; kb < nkb; ++kb) {
#pragma unroll
              for (int k = 0; k < FBK / 16; ++k)
                umma_bf16(tmem_base, make_smem_desc(a0 + kb * Cfg::A_BYTES + k * 32, 16, 1024),
                          make_smem_desc(b0 + kb * Cfg::B_BYTES + k * 32, 16, 1024), idesc, (kb > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit(tmem_full_bar);
          }
          __syncwarp();
          b_phase ^= 1;
        }
      }
    }
  } else {
    // ================================================================== epilogue warps
    const int q = warp & 3;           // TMEM lane quarter (hardware rule: warp % 4)
    const int sub = (warp - 2) >> 2;  // two warps share a quarter and interleave 16-column chunks
    const int r_in = q * 32 + lane;   // row inside the block
    const int row = m0 + r_in;
    const bool row_ok = row < p.M;
    constexpr int NCHUNK = BN / 16;
    constexpr int WPQ = kFEpiWarps / 4;
    constexpr int MYCH = (NCHUNK + WPQ - 1) / WPQ;
    for (int j = threadIdx.x - 64; j < kFMaxLayers * BN; j += 32 * kFEpiWarps) {
      const int l = j / BN, c = j - l * BN;
      float b = 0.f;
      if (l < p.L) {
        const float* bp = l == 0 ? p.bias[0] : l == 1 ? p.bias[1] : l == 2 ? p.bias[2] : p.bias[3];
        const int hl = l == 0 ? p.H[0] : l == 1 ? p.H[1] : l == 2 ? p.H[2] : p.H[3];
        if (bp && n0 + c < hl) b = __ldg(bp + n0 + c);
      }
      s_bias[j] = b;
    }
    unsigned long long step = 0ull;
    if (p.step_ptr) step = *p.step_ptr;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * kFEpiWarps) : "memory");  // s_bias ready

    uint32_t full_phase = 0;
#pragma unroll
    for (int l = 0; l < kFMaxLayers; ++l) {
      if (l < p.L) {
        const bool act = active(l);
        const bool last = (l == p.L - 1);
        uint4 pk[MYCH][2];
        if (act) {
          DropoutCfg drop = p.drop[l];
          drop.step_lo = (uint32_t)step;
          drop.step_hi = (uint32_t)(step >> 32);
          const int Hl = p.H[l];
          // keep-bits first: Philox runs while TMA / MMA are still busy
          uint32_t bits[MYCH];
#pragma unroll
          for (int k = 0; k < MYCH; ++k) {
            const int c = (sub + k * WPQ) * 16;
            const int n = n0 + c;
            bits[k] = 0xFFFFu;
            if (c < BN && n < Hl && drop.enabled) {
              const uint64_t quad0 = ((uint64_t)(row + drop.row_offset) * (uint64_t)Hl + (uint64_t)n) >> 2;
              uint32_t b = 0;
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const uint4 r = dropout_words(drop, quad0 + t);
                b |= (r.x < drop.threshold ? 1u : 0u) << (4 * t) | (r.y < drop.threshold ? 1u : 0u) << (4 * t + 1) |
                     (r.z < drop.threshold ? 1u : 0u) << (4 * t + 2) | (r.w < drop.threshold ? 1u : 0u) << (4 * t + 3);
              }
              bits[k] = b;
            }
          }
          if (threadIdx.x == 64 && l == 0) stamp(0);
          mbar_wait(tmem_full_bar, full_phase);
          full_phase ^= 1;
          tc_fence_after();
          if (threadIdx.x == 64) stamp(1 + l * 6);
          if (last) pdl_launch_dependents();  // late, as in the single-layer GEMM: the accumulator of the last layer is done
          const float scale = drop.enabled ? drop.scale : 1.f;
          __nv_bfloat16* orow = p.out[l] + (size_t)row * p.ldo[l];
#pragma unroll
          for (int k = 0; k < MYCH; ++k) {
            const int c = (sub + k * WPQ) * 16;
            const int n = n0 + c;
            pk[k][0] = make_uint4(0u, 0u, 0u, 0u);
            pk[k][1] = make_uint4(0u, 0u, 0u, 0u);
            if (c < BN && n < Hl) {  // warp-uniform; widths are multiples of 16, so a chunk is inside or outside
              float v[16];
              tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                float x = v[j] + s_bias[l * BN + c + j];
                if (p.relu) x = fmaxf(x, 0.f);
                v[j] = ((bits[k] >> j) & 1u) ? x * scale : 0.f;
              }
              pk[k][0].x = pack_bf16x2(v[0], v[1]);   pk[k][0].y = pack_bf16x2(v[2], v[3]);
              pk[k][0].z = pack_bf16x2(v[4], v[5]);   pk[k][0].w = pack_bf16x2(v[6], v[7]);
              pk[k][1].x = pack_bf16x2(v[8], v[9]);   pk[k][1].y = pack_bf16x2(v[10], v[11]);
              pk[k][1].z = pack_bf16x2(v[12], v[13]); pk[k][1].w = pack_bf16x2(v[14], v[15]);
              if (row_ok) {
                reinterpret_cast<uint4*>(orow + n)[0] = pk[k][0];
                reinterpret_cast<uint4*>(orow + n)[1] = pk[k][1];
              }
            }
          }
        }
        if (!last) {
          if (threadIdx.x == 64) stamp(2 + l * 6);
          tc_fence_before();   // the accumulator has been read: the next layer's MMAs may overwrite it after B2
          cluster_sync_all();  // B1: nobody's tensor core reads the old activation block / ring any more
          if (threadIdx.x == 64) stamp(3 + l * 6);
          if (act) {
            const int Hl = p.H[l];
            const uint32_t a_local = smem_u32(a_res);
            for (int peer = 0; peer < csize; ++peer) {
              const uint32_t a_peer = mapa_u32(a_local, (uint32_t)peer);
#pragma unroll
              for (int k = 0; k < MYCH; ++k) {
                const int c = (sub + k * WPQ) * 16;
                const int n = n0 + c;
                if (c < BN && n < Hl) {
                  // element (r_in, n) of the K-major 128B-swizzled block: k-block n/64, 8-row groups 1024 B apart, rows
                  // 128 B apart, 16-byte unit index XOR (row % 8)
                  const uint32_t kb = (uint32_t)n >> 6, u = ((uint32_t)n & 63u) >> 3;
                  const uint32_t base = a_peer + kb * Cfg::A_BYTES + ((uint32_t)r_in >> 3) * 1024u + ((uint32_t)r_in & 7u) * 128u;
                  st_cluster_v4(base + ((u ^ ((uint32_t)r_in & 7u)) << 4), pk[k][0]);
                  st_cluster_v4(base + (((u + 1u) ^ ((uint32_t)r_in & 7u)) << 4), pk[k][1]);
                }
              }
            }
          }
          // columns [H_l, 64 * ceil(H_l / 64)) of the resident block are multiplied by zero-filled W columns, but 0 * NaN
          // is NaN: clear whatever the layer-0 ring left there (one warp of each row quarter, own shared memory)
          if (sub == 0) {
            const int Hl = p.H[l];
            const uint32_t u_end = (uint32_t)((Hl + FBK - 1) / FBK) * 8u;
            for (uint32_t u = (uint32_t)Hl >> 3; u < u_end; ++u) {
              const uint32_t kb = u >> 3;
              uint8_t* dst = a_res + kb * Cfg::A_BYTES + (r_in >> 3) * 1024 + (r_in & 7) * 128 + (((u & 7u) ^ (r_in & 7)) << 4);
              *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
            }
          }
          if (threadIdx.x == 64) stamp(4 + l * 6);
          fence_proxy_async();  // generic-proxy stores -> visible to the async proxy (tcgen05.mma operand reads)
          if (threadIdx.x == 64) stamp(5 + l * 6);
          cluster_sync_all();   // B2
          if (threadIdx.x == 64) stamp(6 + l * 6);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  trace_end(p.trace_id);
  if (warp == 0) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  cluster_sync_all();  // nobody leaves (and frees its shared memory) while a peer could still be writing into it
}

}  // namespace dfm

using namespace dfm;

namespace {
struct MlpFwdPlan {
  CUtensorMap tmX, tmW[kFMaxLayers];
  MlpFwdParams p;
  int bn, cluster, m_tiles;
};

template <int BN>
int launch_fused(const MlpFwdPlan& pl, cudaStream_t stream) {
  using Cfg = FusedCfg<BN>;
  auto kern = mlp_fwd_fused_kernel<BN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  DFM_CHECK(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DFM_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    DFM_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)pl.cluster, (unsigned)pl.m_tiles, 1);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  int na = 0;
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = (unsigned)pl.cluster;
  attr[na].val.clusterDim.y = 1;
  attr[na].val.clusterDim.z = 1;
  ++na;
  if (dfm_pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  const cudaAccessPolicyWindow* w = dfm_launch_no_window() ? nullptr : dfm_l2_window();
  if (w) {
    attr[na].id = cudaLaunchAttributeAccessPolicyWindow;
    attr[na].val.accessPolicyWindow = *w;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  DFM_CHECK(cudaLaunchKernelEx(&cfg, kern, pl.tmX, pl.tmW[0], pl.tmW[1], pl.tmW[2], pl.tmW[3], pl.p));
  return DFM_OK;
}
}  // namespace

extern "C" {

// Fused MLP forward over L <= 4 layers: X_{l+1}[M, H_l] = dropout(relu(X_l W_l^T + b_l)), X_0 = X [M, K0].
// W[l]: bf16 [H_l, K_l] row-major (K_l = K0 or H_{l-1}) with leading dimension ldw[l]; out[l]: bf16 [M, H_l] with leading
// dimension ldo[l]; keep[l]: dropout keep probability (1 = none); layer_ids[l]: Philox stream of the layer (the same ids
// the per-layer GEMMs use, so both paths draw identical masks). Requirements (else DFM_ERR_ARG: use the per-layer GEMMs):
// every H_l a multiple of 16 and at most 448, leading dimensions multiples of 8.
int dfm_mlp_fwd_plan_create(void** plan_out, const void* X, int ldx, int M, int K0, int L, const void* const* W,
                            const int* ldw, const int* H, void* const* out, const int* ldo, const float* const* bias,
                            int relu, const float* keep, unsigned long long seed, const int* layer_ids, int row_offset,
                            const unsigned long long* step_ptr) {
  DFM_REQUIRE(L >= 1 && L <= kFMaxLayers, "fused mlp: 1..%d layers", kFMaxLayers);
  DFM_REQUIRE(M > 0 && K0 > 0 && ldx % 8 == 0, "fused mlp: empty problem or ldx not a multiple of 8");
  int maxh = 0;
  for (int l = 0; l < L; ++l) {
    DFM_REQUIRE(H[l] > 0 && H[l] % 16 == 0 && H[l] <= kFMaxKB * FBK, "fused mlp: layer widths must be multiples of 16, <= %d",
                kFMaxKB * FBK);
    DFM_REQUIRE(ldw[l] % 8 == 0 && ldo[l] % 8 == 0, "fused mlp: leading dimensions must be multiples of 8");
    maxh = H[l] > maxh ? H[l] : maxh;
  }
  MlpFwdPlan* pl = new MlpFwdPlan();
  const int pad80 = ceil_div(maxh, 80) * 80 - maxh, pad64 = ceil_div(maxh, 64) * 64 - maxh;
  pl->bn = pad80 < pad64 ? 80 : 64;
  pl->cluster = ceil_div(maxh, pl->bn);
  if (pl->cluster > 8) { delete pl; DFM_REQUIRE(false, "fused mlp: layer too wide for one cluster"); }
  pl->m_tiles = ceil_div(M, FBM);
  int rc = make_tmap(&pl->tmX, X, M, K0, ldx, FBM);
  for (int l = 0; l < kFMaxLayers && !rc; ++l) {
    const int s = l < L ? l : L - 1;  // unused maps repeat the last layer's
    rc = make_tmap(&pl->tmW[l], W[s], H[s], s == 0 ? K0 : H[s - 1], ldw[s], pl->bn);
  }
  if (rc) { delete pl; return rc; }
  MlpFwdParams& p = pl->p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.L = L; p.K0 = K0; p.relu = relu; p.step_ptr = step_ptr; p.trace_id = -1;
  for (int l = 0; l < L; ++l) {
    p.H[l] = H[l];
    p.out[l] = reinterpret_cast<__nv_bfloat16*>(out[l]);
    p.ldo[l] = ldo[l];
    p.bias[l] = bias ? bias[l] : nullptr;
    const float kp = keep ? keep[l] : 1.0f;
    DropoutCfg& d = p.drop[l];
    d.enabled = kp < 1.0f ? 1u : 0u;
    d.threshold = kp >= 1.0f ? 0xFFFFFFFFu : (uint32_t)floor((double)kp * 4294967296.0);
    d.scale = kp >= 1.0f ? 1.0f : 1.0f / kp;
    d.seed_lo = (uint32_t)seed; d.seed_hi = (uint32_t)(seed >> 32);
    d.step_lo = d.step_hi = 0; d.layer = (uint32_t)(layer_ids ? layer_ids[l] : l); d.row_offset = (uint32_t)row_offset;
  }
  *plan_out = pl;
  return DFM_OK;
}

int dfm_mlp_fwd_plan_info(void* plan, int* bn, int* cluster, int* m_tiles) {
  DFM_REQUIRE(plan, "null plan");
  const MlpFwdPlan* pl = (const MlpFwdPlan*)plan;
  if (bn) *bn = pl->bn;
  if (cluster) *cluster = pl->cluster;
  if (m_tiles) *m_tiles = pl->m_tiles;
  return DFM_OK;
}

int dfm_mlp_fwd_plan_set_trace_id(void* plan, int id) {
  DFM_REQUIRE(plan, "null plan");
  ((MlpFwdPlan*)plan)->p.trace_id = id;
  return DFM_OK;
}

int dfm_mlp_fwd_plan_set_debug(void* plan, unsigned long long* dbg) {
  DFM_REQUIRE(plan, "null plan");
  ((MlpFwdPlan*)plan)->p.dbg = dbg;
  return DFM_OK;
}

int dfm_mlp_fwd_run(void* plan, void* stream) {
  DFM_REQUIRE(plan, "null plan");
  const MlpFwdPlan& pl = *(const MlpFwdPlan*)plan;
  return pl.bn == 80 ? launch_fused<80>(pl, (cudaStream_t)stream) : launch_fused<64>(pl, (cudaStream_t)stream);
}

void dfm_mlp_fwd_plan_destroy(void* plan) { delete (MlpFwdPlan*)plan; }

}  // extern "C"

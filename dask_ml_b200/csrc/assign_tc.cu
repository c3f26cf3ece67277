// assign_tc.cu -- K1 on the 5th-generation tensor cores: 3xTF32 tcgen05.mma with TMEM accumulators.
//
// Same contract as assign_tiled_f32_kernel (labels, winner distance, near-tie flags), for d <= 128.
// Replaces dask_ml.metrics.pairwise_distances_argmin_min (reference dask_ml/metrics/pairwise.py:19-47)
// on the Lloyd path (dask_ml/cluster/k_means.py:576-579) when k*d makes the FFMA pipe the bound.
//
// Score matrix  S = [x_hi | x_lo | 1 1 1] . [ (-2c)_hi | (-2c)_lo | cn_h cn_m cn_l ]^T  computed as
//   D  = Xhi.Bhi^T + Xhi.Blo^T + Xlo.Bhi^T   (error-compensated TF32 split, fp32 accumulate in TMEM)
//      + Xaug.Baug^T                         (one extra K-step that adds ||c||^2 exactly)
// so the epilogue only has to track (best, second, argmin) per row -- no per-element norm add.
//
// CTA = 320 threads, persistent, one per SM, 128 rows (one M=128 row tile) per step:
//   warp 0      : producer -- one 1-D bulk async copy (TMA, mbarrier complete_tx) per K-step block of a
//                 256-centroid column tile (16 KB: hi + lo parts of 8 features)
//   warp 1      : TMEM allocator + tcgen05.mma issue (M=128, N=256, K=8, kind::tf32, one elected lane)
//   warps 2..9  : X staging (LDG.128 -> cvt.rna.tf32 split -> STS in the UMMA canonical no-swizzle
//                 K-major layout, double-buffered so the next step is staged under the current one) and
//                 the epilogue (tcgen05.ld 32x32b.x32, double-buffered against the argmin tracking)
// N = 256 per MMA keeps the tensor core's shared-memory operand fetch (A 4 KB + B 8 KB per K-step) below
// its 128-cycle math time; with N = 64 the fetch dominated and the pipe ran at 36 % (profiles/).
// Pipelines: B ring (5 smem stages), accumulator ring (2 TMEM stages of 256 columns), 2 X buffers.
#include "tc_common.cuh"

namespace {

struct TcBars {
    uint64_t b_full[TC_BSTAGES_MAX];
    uint64_t b_empty[TC_BSTAGES_MAX];
    uint64_t a_full[TC_ASTAGES];
    uint64_t a_empty[TC_ASTAGES];
    uint64_t x_ready[2];
    uint64_t x_free[2];
    uint64_t m_ready[2];           // both column halves of a step have published their (best, second, argmin)
    uint64_t s_ready[2];           // the upper half has published its partial winner distance
    uint32_t tmem_base;
    uint32_t pad;
    // per step buffer [2], per column half [2], per row [128]
    float m_best[2][2][128];
    float m_second[2][2][128];
    int m_idx[2][2][128];
    float xn_part[2][2][128];
    float s_part[2][128];
    float xn_pub[2][128];
    alignas(16) float mu[KM_TC_MAX_D];   // the pack's centring vector (zero padded), read as float4
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// TC image of the centres (runs after pack_centers_kernel on the same stream)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) pack_tc_kernel(int k, int d, unsigned char *__restrict__ pack, int pair) {
    // pair != 0: each 256-column tile is stored as two 128-column halves (one per CTA of the pair), each
    // half = [aug | K-step blocks]; otherwise as one 256-row operand.  BR = operand rows per block.
    const KmPackLayout L = km_pack_layout(k, d);
    const float *c32 = reinterpret_cast<const float *>(pack + L.off_c32);
    const float *cn32 = reinterpret_cast<const float *>(pack + L.off_cn32);
    const int j = blockIdx.x;
    const int tile = j / KM_TC_BN, jj = j % KM_TC_BN;
    const int kpad = km_tc_kpad(d);
    const int BR = pair ? KM_TC_BN / 2 : KM_TC_BN;
    const int half = pair ? jj / BR : 0, jr = jj - half * BR;
    const size_t aug_bytes = (size_t)2 * BR * 16, ks_bytes = (size_t)4 * BR * 16;
    unsigned char *img = pack + L.off_tc + (size_t)tile * km_tc_tile_bytes(d) +
                         (size_t)half * (aug_bytes + (size_t)(kpad / 8) * ks_bytes);
    for (int f = threadIdx.x; f < kpad; f += blockDim.x) {
        float v = 0.f;
        if (j < k && f < d) v = -2.f * c32[(size_t)j * d + f];
        const float h = tf32_rna(v);
        const float l = v - h;            // exact; read by the tensor core as TF32 (low 13 bits ignored)
        const int ks = f >> 3, fc = f & 7;
        unsigned char *blk = img + aug_bytes + (size_t)ks * ks_bytes;
        const size_t o = ((size_t)(fc >> 2) * BR + jr) * 4 + (fc & 3);
        reinterpret_cast<float *>(blk)[o] = h;
        reinterpret_cast<float *>(blk + ks_bytes / 2)[o] = l;
    }
    if (threadIdx.x < 8) {
        // augmented K-step: chunk 0 = [cn_h, cn_m, cn_l, 1], chunk 1 = [1, 1, 0, 0].  K1 multiplies it with
        // [1, 1, 1, 0] [0, 0, 0, 0] (adds ||c||^2), K7 with [1, 1, 1, xn_h] [xn_m, xn_l, 0, 0] (adds ||x||^2 as well)
        float v = 0.f;
        if (threadIdx.x < 3) {
            if (j < k) {
                const float cn = cn32[j];
                const float h = tf32_rna(cn);
                const float m = tf32_rna(cn - h);
                const float l = tf32_rna(cn - h - m);
                v = threadIdx.x == 0 ? h : (threadIdx.x == 1 ? m : l);
            } else {
                v = threadIdx.x == 0 ? 1e30f : 0.f;  // padded columns never win
            }
        } else if (threadIdx.x < 6 && j < k) {
            v = 1.f;
        }
        const int f = threadIdx.x;
        reinterpret_cast<float *>(img)[((size_t)(f >> 2) * BR + jr) * 4 + (f & 3)] = v;
    }
}

bool km_tc_pair_mode();  // assign.cu: true when K1 runs on CTA pairs (cta_group::2)
bool km_tc_small_enabled();   // assign.cu: false when the small-k kernel is switched off (variant 2 forces the general one)

int km_launch_pack_tc(int k, int d, void *pack, cudaStream_t st) {
    if (d > KM_TC_WIDE_MAX_D) return 0;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    pack_tc_kernel<<<ntc * KM_TC_BN, 64, 0, st>>>(k, d, static_cast<unsigned char *>(pack), km_tc_pair_mode() ? 1 : 0);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
// DBUF: two X buffers (d <= 64) so the next step's rows are staged under the current step; for
// 64 < d <= 128 one buffer fills shared memory and staging happens at the step boundary instead.
// PAIR: the kernel runs on clusters of two CTAs (one SM pair); one tcgen05.mma.cta_group::2 of M = 256 covers
// 128 rows of each CTA, and each CTA streams only ITS 128-column half of every centroid tile -- half the
// TMA writes and half the B operand reads per SM, which is what keeps shared memory from capping the pipe.
template <bool VEC, bool DBUF, bool PAIR>
__global__ void __launch_bounds__(TC_THREADS, 1)
assign_tc_f32_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx,
                     const unsigned char *__restrict__ tc_img, const float *__restrict__ c32,
                     const float *__restrict__ mu, const unsigned int *__restrict__ cmax_bits, int k,
                     int32_t *__restrict__ labels, float *__restrict__ mind2,
                     int32_t *__restrict__ flag_rows, int32_t *__restrict__ flag_count, int sparse_d2) {
    // sparse_d2 != 0 (b200km_assign_lloyd_f32): the winner distance is not rebuilt here (K2 sums ||x - c||^2 itself);
    // a flagged row gets ||x'||^2 + best score, which is what the pruned re-decision needs.  That drops the gather of
    // the winner's centre row (256 B per row through L2) and the re-read of the staged row from every step.
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;                 // 16-byte K chunks of the real features
    const int nks = kpad / 8;                 // K-steps
    const uint32_t xhi_bytes = (uint32_t)(nch + 2) * 128 * 16;
    const uint32_t xlo_bytes = (uint32_t)nch * 128 * 16;
    const uint32_t xbuf_bytes = xhi_bytes + xlo_bytes;     // one X buffer: hi part then lo part
    const uint32_t tile_bytes = (uint32_t)km_tc_tile_bytes(d);
    constexpr int TC_BSTAGES = PAIR ? 8 : 5;
    constexpr uint32_t BROWS = PAIR ? KM_TC_BN / 2 : KM_TC_BN;   // B operand rows held by this CTA
    constexpr uint32_t TC_STAGE_BYTES = 4 * BROWS * 16;           // one K-step block: hi + lo, 2 chunks each
    constexpr uint32_t AUG_BYTES = 2 * BROWS * 16;
    unsigned char *xb = tc_smem;                           // [DBUF ? 2 : 1][xbuf_bytes]
    unsigned char *bst = xb + (DBUF ? 2 : 1) * xbuf_bytes; // [TC_BSTAGES][TC_STAGE_BYTES]
    TcBars *bars = reinterpret_cast<TcBars *>(bst + TC_BSTAGES * TC_STAGE_BYTES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    const uint32_t crank = PAIR ? cluster_ctarank() : 0u;          // 0 = leader of the pair
    const int64_t step_rows = PAIR ? 2 * TC_ROWS : TC_ROWS;        // rows per step of the CTA (pair)
    const int64_t nsteps = (n + step_rows - 1) / step_rows;
    const int64_t step0 = PAIR ? blockIdx.x / 2 : blockIdx.x;
    const int64_t step_inc = PAIR ? gridDim.x / 2 : gridDim.x;

    if (tid == 0) {
        // pair mode: the leader's b_full / a_empty / x_ready also collect the peer CTA's arrivals
        const uint32_t np = PAIR ? 2u : 1u;
        for (int i = 0; i < TC_BSTAGES; ++i) {
            km_mbar_init(&bars->b_full[i], (PAIR && crank == 0) ? 2 : 1);
            km_mbar_init(&bars->b_empty[i], 1);
        }
        for (int i = 0; i < TC_ASTAGES; ++i) { km_mbar_init(&bars->a_full[i], 1); km_mbar_init(&bars->a_empty[i], np * TC_EPI); }
        for (int i = 0; i < 2; ++i) {
            km_mbar_init(&bars->x_ready[i], np * TC_EPI);
            km_mbar_init(&bars->x_free[i], 1);
            km_mbar_init(&bars->m_ready[i], TC_EPI);
            km_mbar_init(&bars->s_ready[i], TC_EPI / 2);
        }
        km_fence_mbar_init();
    }
    if (warp == 1) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             km_smem_u32(&bars->tmem_base)),
                         "r"(512u)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             km_smem_u32(&bars->tmem_base)),
                         "r"(512u)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    for (int f = tid; f < kpad; f += TC_THREADS) bars->mu[f] = __ldg(mu + f);
    // constant augmented chunks of Xhi: chunk nch = [1,1,1,0], chunk nch+1 = 0 (both X buffers)
    for (int e = tid; e < (DBUF ? 2 : 1) * 2 * 128; e += TC_THREADS) {
        const int bf = e / 256, c = (e / 128) & 1, r = e & 127;
        float4 v = c == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(xb + (size_t)bf * xbuf_bytes + ((size_t)(nch + c) * 128 + r) * 16) = v;
    }
    km_fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t step = step0; step < nsteps; step += step_inc) {
                for (int t = 0; t < ntc; ++t) {
                    // pair mode: this CTA streams only its 128-column half of the tile
                    const unsigned char *src = tc_img + (size_t)t * tile_bytes + (size_t)crank * (tile_bytes / 2);
                    for (int b = 0; b <= nks; ++b, ++it) {
                        const int s = it % TC_BSTAGES;
                        const uint32_t ph = (it / TC_BSTAGES) & 1;
                        const uint32_t bytes = b == 0 ? AUG_BYTES : TC_STAGE_BYTES;
                        mbar_wait_bounded(&bars->b_empty[s], ph ^ 1);
                        km_mbar_expect_tx(&bars->b_full[s], bytes);
                        km_bulk_g2s(bst + (size_t)s * TC_STAGE_BYTES, src, bytes, &bars->b_full[s]);
                        src += bytes;
                    }
                }
            }
        }
    } else if (warp == 1 && PAIR && crank != 0) {
        // ================= peer CTA: relay "my half of block b has landed" to the leader =============
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t step = step0; step < nsteps; step += step_inc)
                for (int t = 0; t < ntc; ++t)
                    for (int b = 0; b <= nks; ++b, ++it) {
                        const int s = it % TC_BSTAGES;
                        mbar_wait_bounded(&bars->b_full[s], (it / TC_BSTAGES) & 1);
                        mbar_arrive_remote_relaxed(&bars->b_full[s], 0);
                    }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (warp-uniform control flow, one elected lane issues) =======
        // instruction descriptor: D=F32, A=B=TF32, both K-major, N=256, M=128 (256 across a CTA pair)
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KM_TC_BN >> 3) << 17) |
                               ((uint32_t)((PAIR ? 256 : 128) >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);           // SBO = 128 B, version = 1
        const uint32_t a_lbo = (2048u >> 4) << 16;                   // A: 128 rows x 16 B per K chunk
        const uint32_t b_lbo = ((BROWS * 16) >> 4) << 16;            // B: BROWS rows x 16 B per K chunk
        const uint32_t xb_a = km_smem_u32(xb), bst_a = km_smem_u32(bst);
        const bool leader = elect_one();
        uint32_t it = 0, ita = 0, li = 0;
        for (int64_t step = step0; step < nsteps; step += step_inc, ++li) {
            const uint32_t xbuf = DBUF ? (li & 1) : 0u;
            mbar_wait_bounded(&bars->x_ready[xbuf], (DBUF ? (li >> 1) : li) & 1);
            tc_fence_after();
            const uint32_t ah = (((xb_a + xbuf * xbuf_bytes) & 0x3FFFFu) >> 4) | a_lbo;
            const uint32_t al = ah + (xhi_bytes >> 4);
            for (int t = 0; t < ntc; ++t, ++ita) {
                const int a = ita % TC_ASTAGES;
                const uint32_t aph = (ita / TC_ASTAGES) & 1;
                mbar_wait_bounded(&bars->a_empty[a], aph ^ 1);
                const uint32_t dcol = tmem_base + (uint32_t)(a * KM_TC_BN);
                for (int b = 0; b <= nks; ++b, ++it) {
                    const int s = it % TC_BSTAGES;
                    const uint32_t bph = (it / TC_BSTAGES) & 1;
                    mbar_wait_bounded(&bars->b_full[s], bph);
                    tc_fence_after();
                    const uint32_t bh = (((bst_a + s * TC_STAGE_BYTES) & 0x3FFFFu) >> 4) | b_lbo;
                    if (leader) {
                        const uint32_t ao = (uint32_t)(b - 1) * 256u;
                        const uint32_t bl = bh + ((TC_STAGE_BYTES / 2) >> 4);
                        if (PAIR) {
                            if (b == 0) {
                                mma_tf32_pair(dcol, ah + (uint32_t)nch * 128u, bh, desc_hi, idesc, 0u);
                            } else {
                                mma_tf32_pair(dcol, al + ao, bh, desc_hi, idesc, 1u);
                                mma_tf32_pair(dcol, ah + ao, bl, desc_hi, idesc, 1u);
                                mma_tf32_pair(dcol, ah + ao, bh, desc_hi, idesc, 1u);
                            }
                            mma_commit_pair(&bars->b_empty[s]);
                            if (b == nks) mma_commit_pair(&bars->a_full[a]);
                        } else {
                            if (b == 0) {
                                // augmented step first (overwrites the accumulator): + ||c||^2
                                mma_tf32(dcol, ah + (uint32_t)nch * 128u, bh, desc_hi, idesc, 0u);
                            } else {
                                mma_tf32(dcol, al + ao, bh, desc_hi, idesc, 1u);
                                mma_tf32(dcol, ah + ao, bl, desc_hi, idesc, 1u);
                                mma_tf32(dcol, ah + ao, bh, desc_hi, idesc, 1u);
                            }
                            mma_commit(&bars->b_empty[s]);  // smem stage reusable once these MMAs retire
                            if (b == nks) mma_commit(&bars->a_full[a]);  // accumulator ready for the epilogue
                        }
                    }
                    __syncwarp();
                }
            }
            if (leader) {   // X buffer no longer read by the tensor core
                if (PAIR) mma_commit_pair(&bars->x_free[xbuf]); else mma_commit(&bars->x_free[xbuf]);
            }
            __syncwarp();
        }
    } else {
        // ================= X staging + epilogue (warps 2..9, 256 threads) =================
        // Two threads per row: half g = 0 (warps 2..5) owns columns [0,128) of every 256-column tile and the
        // lower K chunks of the row, half g = 1 (warps 6..9) the rest.  Two warps per scheduler keep the ALU
        // pipe busy through the dependent min/max chains and the TMEM-load waits of one another.
        const int g = (warp - 2) >> 2;           // column / K half
        const int q4 = warp & 3;                 // TMEM lane quarter this warp may access
        const int rl = q4 * 32 + lane;           // row within the tile == TMEM lane
        const float cmax = __uint_as_float(*cmax_bits);
        const float bound_scale = 2.0f * (float)(d + 14) * 1.1920929e-07f;
        constexpr int NCH_H = DBUF ? KM_TC_MAX_D / 16 : KM_TC_MAX_D / 8;   // K chunks per half (max; d <= 64 when DBUF)
        const int nch_h = nch / 2;               // nch is even (kpad multiple of 8)
        const int c_lo = g * nch_h;              // first K chunk of this half

        // ---- software-pipelined helpers (loads are issued one epilogue tile before they are used) ----
        using LdRegs = float4[NCH_H];             // in-flight global loads (half an X row or half a centre row)
        LdRegs ld;
        auto issue_row_loads = [&](LdRegs &ld, const float *src, bool live) {
#pragma unroll
            for (int c = 0; c < NCH_H; ++c) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                const int cc = c_lo + c;
                if (c < nch_h && live) {
                    if (VEC) {
                        if (cc * 4 < d) v = __ldg(reinterpret_cast<const float4 *>(src) + cc);
                    } else {
                        const int f = cc * 4;
                        v.x = f + 0 < d ? __ldg(src + f + 0) : 0.f;
                        v.y = f + 1 < d ? __ldg(src + f + 1) : 0.f;
                        v.z = f + 2 < d ? __ldg(src + f + 2) : 0.f;
                        v.w = f + 3 < d ? __ldg(src + f + 3) : 0.f;
                    }
                }
                ld[c] = v;
            }
        };
        // split ld[] (half an X row) into TF32 hi/lo parts in X buffer `bf`; `use` = earlier stagings of it
        auto stage_store = [&](const LdRegs &ld, uint32_t bf, uint32_t use) {
            if (use > 0) mbar_wait_bounded(&bars->x_free[bf], (use - 1) & 1);  // tensor core done with it
            unsigned char *ph = xb + (size_t)bf * xbuf_bytes + (size_t)rl * 16;
            unsigned char *pl = ph + xhi_bytes;
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < NCH_H; ++c) {
                if (c < nch_h) {
                    float4 v = ld[c];
                    const float4 mq = *reinterpret_cast<const float4 *>(&bars->mu[(c_lo + c) * 4]);   // x - mu
                    v.x -= mq.x; v.y -= mq.y; v.z -= mq.z; v.w -= mq.w;
                    s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                    float4 h, l;
                    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
                    // low parts are stored exactly (fp32): the tensor core ignores the 13 low mantissa bits of a
                    // TF32 operand, and h + l == x lets the epilogue rebuild the row without error
                    l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;
                    *reinterpret_cast<float4 *>(ph + (size_t)(c_lo + c) * 2048) = h;
                    *reinterpret_cast<float4 *>(pl + (size_t)(c_lo + c) * 2048) = l;
                }
            }
            bars->xn_part[bf][g][rl] = s;
            km_fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
            if (PAIR) mbar_arrive_remote(&bars->x_ready[bf], 0); else mbar_arrive(&bars->x_ready[bf]);
        };
        // pending result of the previous step (its outputs are written one tile into the next step)
        bool pend = false;
        int64_t p_row = 0;
        int p_bi = 0;
        uint32_t p_buf = 0, p_use = 0;
        float p_best = 0.f, p_second = 0.f;
        // merge the two column halves of the pending step (both threads of a row compute the same result)
        auto merge_pending = [&]() {
            mbar_wait_bounded(&bars->m_ready[p_buf], p_use & 1);
            const float b0 = bars->m_best[p_buf][0][rl], b1 = bars->m_best[p_buf][1][rl];
            const float s0 = bars->m_second[p_buf][0][rl], s1 = bars->m_second[p_buf][1][rl];
            const int i0 = bars->m_idx[p_buf][0][rl], i1 = bars->m_idx[p_buf][1][rl];
            const bool take1 = b1 < b0 || (b1 == b0 && i1 < i0);
            p_best = take1 ? b1 : b0;
            const int bi = take1 ? i1 : i0;
            p_second = fminf(fminf(s0, s1), take1 ? b0 : b1);
            p_bi = bi >= k ? k - 1 : bi;
        };
        // partial winner distance (direct form, fp32) over this half's K chunks from ld[] (= centre row half);
        // the upper half publishes its part, the lower half adds it and writes labels / mind2 / near-tie flag
        auto output_pending = [&](const LdRegs &ld) {
            const unsigned char *ph = xb + (size_t)p_buf * xbuf_bytes + (size_t)rl * 16;
            const unsigned char *pl = ph + xhi_bytes;
            float s = 0.f;
            if (!sparse_d2) {
#pragma unroll
              for (int c = 0; c < NCH_H; ++c) {
                if (c < nch_h) {
                    const float4 h = *reinterpret_cast<const float4 *>(ph + (size_t)(c_lo + c) * 2048);
                    const float4 l = *reinterpret_cast<const float4 *>(pl + (size_t)(c_lo + c) * 2048);
                    const float4 cc = ld[c];
                    float df;
                    df = (h.x + l.x) - cc.x; s = fmaf(df, df, s);
                    df = (h.y + l.y) - cc.y; s = fmaf(df, df, s);
                    df = (h.z + l.z) - cc.z; s = fmaf(df, df, s);
                    df = (h.w + l.w) - cc.w; s = fmaf(df, df, s);
                }
              }
            }
            if (g == 1) {
                bars->s_part[p_buf][rl] = s;
                bars->xn_pub[p_buf][rl] = bars->xn_part[p_buf][1][rl];   // before this thread restages the buffer
                mbar_arrive(&bars->s_ready[p_buf]);
            } else {
                mbar_wait_bounded(&bars->s_ready[p_buf], p_use & 1);
                s += bars->s_part[p_buf][rl];
                if (p_row < n) {
                    labels[p_row] = p_bi;
                    const float xn = bars->xn_part[p_buf][0][rl] + bars->xn_pub[p_buf][rl];
                    const bool near_tie = flag_rows && (p_second - p_best <= bound_scale * (xn + cmax));
                    if (sparse_d2) {
                        if (mind2 && near_tie) mind2[p_row] = fmaxf(xn + p_best, 0.f);
                    } else if (mind2) {
                        mind2[p_row] = s;
                    }
                    if (near_tie) {
                        const int pos = atomicAdd(flag_count, 1);
                        flag_rows[pos] = (int32_t)p_row;
                    }
                }
            }
        };

        // two independent trackers (even / odd columns) double the ILP of the dependent min/max chains
        float best2[2], second2[2];
        int bidx2[2];
        // this thread's 128-column half of column tile t (accumulator stage of ring position `it`):
        // TMEM -> registers -> trackers, 4 chunks of 32 columns
        auto track_tile = [&](int t, uint32_t it) {
            constexpr int NC = 4;
            const int col0 = g * 128;
            const int a = it % TC_ASTAGES;
            const uint32_t aph = (it / TC_ASTAGES) & 1;
            mbar_wait_bounded(&bars->a_full[a], aph);
            tc_fence_after();
            const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(a * KM_TC_BN + col0);
            uint32_t ra[32], rb[32];
            tmem_ld32(tbase, ra);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                uint32_t (&cur)[32] = (c & 1) ? rb : ra;
                uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
                if (c + 1 < NC) tmem_ld32(tbase + (uint32_t)((c + 1) * 32), nxt);  // flies under this chunk
                const int cb = t * KM_TC_BN + col0 + c * 32;
                if (cb < k) {               // chunks made of padding columns only cannot win (warp-uniform)
                    const float before0 = best2[0], before1 = best2[1];
                    int jloc0 = 0, jloc1 = 0;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        track(__uint_as_float(cur[j]), j, best2[0], second2[0], jloc0);
                        track(__uint_as_float(cur[j + 1]), j + 1, best2[1], second2[1], jloc1);
                    }
                    bidx2[0] = best2[0] < before0 ? cb + jloc0 : bidx2[0];
                    bidx2[1] = best2[1] < before1 ? cb + jloc1 : bidx2[1];
                }
                if (c + 1 < NC) tmem_ld_wait();
            }
            tc_fence_before();
            if (PAIR) mbar_arrive_remote_relaxed(&bars->a_empty[a], 0); else mbar_arrive(&bars->a_empty[a]);
        };
        // publish this half's result of a step (an exact tie keeps the lower column; it is flagged anyway)
        auto publish = [&](uint32_t xbuf, int64_t row, uint32_t use) {
            const bool takeb = best2[1] < best2[0] || (best2[1] == best2[0] && bidx2[1] < bidx2[0]);
            bars->m_best[xbuf][g][rl] = takeb ? best2[1] : best2[0];
            bars->m_idx[xbuf][g][rl] = takeb ? bidx2[1] : bidx2[0];
            bars->m_second[xbuf][g][rl] = fminf(fminf(second2[0], second2[1]), takeb ? best2[0] : best2[1]);
            mbar_arrive(&bars->m_ready[xbuf]);
            pend = true;
            p_row = row;
            p_buf = xbuf;
            p_use = use;
        };

        uint32_t it = 0, li = 0;
        if (step0 < nsteps) {
            const int64_t row = step0 * step_rows + crank * TC_ROWS + rl;
            issue_row_loads(ld, X + (row < n ? row : 0) * ldx, row < n);
            stage_store(ld, 0, 0);
        }
        if (DBUF && ntc == 1) {
            // ---- one column tile per step (k <= 256): nothing inside a step to hide the step boundary under, so
            // the pipeline runs one step deeper -- step s+1 is staged BEFORE the scores of step s are tracked
            // (its MMAs then run under that epilogue) and the rows of step s+2 are already in flight.
            LdRegs ldc;                       // winner's centre row of the pending step
            if (step0 + step_inc < nsteps) {
                const int64_t row = (step0 + step_inc) * step_rows + crank * TC_ROWS + rl;
                issue_row_loads(ld, X + (row < n ? row : 0) * ldx, row < n);
            }
            for (int64_t step = step0; step < nsteps; step += step_inc, ++li, ++it) {
                const uint32_t xbuf = li & 1;
                const bool have_next = step + step_inc < nsteps;
                const bool have_next2 = step + 2 * step_inc < nsteps;
                if (pend) {                   // step s-1 leaves X buffer xbuf^1 ...
                    output_pending(ldc);
                    pend = false;
                }
                if (have_next) stage_store(ld, xbuf ^ 1, (li + 1) >> 1);   // ... and step s+1 moves in
                if (have_next2) {
                    const int64_t row = (step + 2 * step_inc) * step_rows + crank * TC_ROWS + rl;
                    issue_row_loads(ld, X + (row < n ? row : 0) * ldx, row < n);
                }
                best2[0] = best2[1] = second2[0] = second2[1] = INFINITY;
                bidx2[0] = bidx2[1] = 0;
                track_tile(0, it);
                publish(xbuf, step * step_rows + crank * TC_ROWS + rl, li >> 1);
                merge_pending();
                if (!sparse_d2) issue_row_loads(ldc, c32 + (size_t)p_bi * d, true);
            }
            if (pend) output_pending(ldc);
        } else {
            // tile after which the NEXT step's rows are staged (never inside the tile loop with one X buffer)
            const int t_stage = DBUF ? (ntc > 1 ? 1 : 0) : -1;
            for (int64_t step = step0; step < nsteps; step += step_inc, ++li) {
                const int64_t row0 = step * step_rows + crank * TC_ROWS;
                const uint32_t xbuf = DBUF ? (li & 1) : 0u;
                const bool have_next = step + step_inc < nsteps;
                const int64_t nrow = (step + step_inc) * step_rows + crank * TC_ROWS + rl;
                best2[0] = best2[1] = second2[0] = second2[1] = INFINITY;
                bidx2[0] = bidx2[1] = 0;
                for (int t = 0; t < ntc; ++t, ++it) {
                    // loads issued here land while this tile's scores are being tracked
                    if (t == 0 && pend) {
                        merge_pending();
                        if (!sparse_d2) issue_row_loads(ld, c32 + (size_t)p_bi * d, true);
                    }
                    if (t == t_stage && have_next && !(t == 0 && pend))
                        issue_row_loads(ld, X + (nrow < n ? nrow : 0) * ldx, nrow < n);
                    track_tile(t, it);
                    if (t == 0 && pend) {          // previous step's outputs (its X buffer is still intact)
                        output_pending(ld);
                        pend = false;
                        if (t_stage == 0 && have_next) issue_row_loads(ld, X + (nrow < n ? nrow : 0) * ldx, nrow < n);
                    }
                    // stage the NEXT step's rows into the other X buffer (free: the previous step is written
                    // out), so the tensor core never waits for X at a step boundary
                    if (t == t_stage && have_next) stage_store(ld, xbuf ^ 1, (li + 1) >> 1);
                }
                publish(xbuf, row0 + rl, DBUF ? (li >> 1) : li);
                if (!DBUF) {
                    // one X buffer: write this step out now, then restage once the tensor core has let go of it
                    merge_pending();
                    if (!sparse_d2) issue_row_loads(ld, c32 + (size_t)p_bi * d, true);
                    output_pending(ld);
                    pend = false;
                    if (have_next) {
                        issue_row_loads(ld, X + (nrow < n ? nrow : 0) * ldx, nrow < n);
                        stage_store(ld, 0, li + 1);
                    }
                }
            }
            if (pend) {
                merge_pending();
                if (!sparse_d2) issue_row_loads(ld, c32 + (size_t)p_bi * d, true);
                output_pending(ld);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();   // neither CTA may leave (or free TMEM) while the pair's MMAs / arrivals are in flight
    if (warp == 1) {
        tc_fence_after();
        if (PAIR)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// small-k variant: k <= 256 (one column tile per step) and d <= 32
// ---------------------------------------------------------------------------------------------
// With a single column tile there is nothing inside a step to overlap the step boundary with, and the whole
// centre image (<= 72 KB) fits in shared memory next to four X buffers.  So: the image is loaded ONCE per CTA
// (no per-step re-streaming from L2), and the eight staging/epilogue warps form two groups that own alternate
// 128-row steps end to end -- one thread per row: it stages the row, tracks all 256 columns of the step's
// accumulator (TMEM stage = group), and writes the row's outputs with no cross-thread merge.  The groups run
// out of phase, so one group's waits (row loads, centre-row gather, barriers) sit under the other's tracking.
#ifdef KM_TC_TRACE
// debug build only (make EXTRA=-DKM_TC_TRACE): CTA 0 stamps clock64 into a fixed slot per (role, step, event) --
// plain stores, no atomics, so the stamps barely disturb the schedule.  roles: 0/1 = row groups, 2/3 = MMAs of group 0/1
__device__ long long g_tc_trace[4 * 256 * 8];
__device__ __forceinline__ void tc_trace(int role, int step, int ev) {
    if (blockIdx.x != 0 || step >= 256) return;
    g_tc_trace[(role * 256 + step) * 8 + ev] = clock64();
}
extern "C" int b200km_debug_read_trace(long long *host, int max_events) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(host, g_tc_trace, sizeof(long long) * 4 * 256 * 8);
    return 4 * 256 * 8;
}
#define TC_TRACE(role, step, ev) do { if (lane == 0) tc_trace(role, (int)(step), ev); } while (0)
#else
#define TC_TRACE(role, step, ev) do { } while (0)
#endif

#ifndef TC_SMALL_TOKEN_AT
#define TC_SMALL_TOKEN_AT 6
#endif
struct TcSmallBars {
    uint64_t b_ready;
    uint64_t a_full[2];
    uint64_t a_empty[2];
    uint64_t x_ready[2][2];        // [group][buffer]
    uint64_t x_free[2][2];
    uint64_t turn[2];              // tracking token: turn[g] completes when the OTHER group has finished a step
    uint32_t tmem_base;
    uint32_t pad;
    float xn[2][2][128];           // ||x||^2 of the staged rows
};

template <bool VEC>
__global__ void __launch_bounds__(TC_THREADS, 1)
assign_tc_small_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx,
                       const unsigned char *__restrict__ tc_img, const float *__restrict__ c32,
                       const float *__restrict__ mu, const unsigned int *__restrict__ cmax_bits, int k,
                       int32_t *__restrict__ labels, float *__restrict__ mind2,
                       int32_t *__restrict__ flag_rows, int32_t *__restrict__ flag_count, int sparse_d2) {
    // sparse_d2 != 0 (the Lloyd iteration, b200km_assign_lloyd_f32): the winner distance is NOT rebuilt here -- K2
    // sums ||x - c||^2 in float64 from the rows it reads anyway -- and mind2 gets, for the flagged rows only, the
    // expanded-form value ||x'||^2 + best score, which is all the pruned re-decision needs.  That takes the
    // winner's centre-row gather and the re-read of the staged row (2/5 of the kernel's shared-memory traffic, the
    // resource it was bound by) out of the step.
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    constexpr int NCH = 8;                                 // 16-byte K chunks of a row at d = 32
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;
    const int nks = kpad / 8;
    const uint32_t xhi_bytes = (uint32_t)(nch + 2) * 128 * 16;
    const uint32_t xlo_bytes = (uint32_t)nch * 128 * 16;
    const uint32_t xbuf_bytes = xhi_bytes + xlo_bytes;
    const uint32_t img_bytes = (uint32_t)km_tc_tile_bytes(d);
    unsigned char *xb = tc_smem;                           // [group][buffer][xbuf_bytes]
    unsigned char *bimg = xb + 4 * xbuf_bytes;             // [aug | K-step blocks], resident
    TcSmallBars *bars = reinterpret_cast<TcSmallBars *>(bimg + img_bytes);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t ntiles = (n + TC_ROWS - 1) / TC_ROWS;
    const int64_t my_tiles =
        (int64_t)blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;   // tile = bid + li * grid

    if (tid == 0) {
        km_mbar_init(&bars->b_ready, 1);
        for (int g = 0; g < 2; ++g) {
            km_mbar_init(&bars->a_full[g], 1);
            km_mbar_init(&bars->a_empty[g], TC_EPI / 2);
            km_mbar_init(&bars->turn[g], TC_EPI / 2);
            for (int b = 0; b < 2; ++b) {
                km_mbar_init(&bars->x_ready[g][b], TC_EPI / 2);
                km_mbar_init(&bars->x_free[g][b], 1);
            }
        }
        km_fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         km_smem_u32(&bars->tmem_base)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // constant augmented chunks of Xhi in all four buffers: chunk nch = [1,1,1,0], chunk nch+1 = 0
    for (int e = tid; e < 4 * 2 * 128; e += TC_THREADS) {
        const int bf = e / 256, c = (e / 128) & 1, r = e & 127;
        float4 v = c == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(xb + (size_t)bf * xbuf_bytes + ((size_t)(nch + c) * 128 + r) * 16) = v;
    }
    km_fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================= the centre image, once =================
        if (lane == 0 && my_tiles > 0) {
            km_mbar_expect_tx(&bars->b_ready, img_bytes);
            km_bulk_g2s(bimg, tc_img, KM_TC_AUG_BYTES, &bars->b_ready);
            for (int b = 0; b < nks; ++b)
                km_bulk_g2s(bimg + KM_TC_AUG_BYTES + (size_t)b * KM_TC_KS_BYTES,
                            tc_img + KM_TC_AUG_BYTES + (size_t)b * KM_TC_KS_BYTES, KM_TC_KS_BYTES, &bars->b_ready);
        }
    } else if (warp == 1) {
        // ================= MMA issuer for both groups =================
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KM_TC_BN >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);           // SBO = 128 B, version = 1
        const uint32_t a_lbo = (2048u >> 4) << 16;
        const uint32_t b_lbo = ((KM_TC_BN * 16) >> 4) << 16;
        const uint32_t xb_a = km_smem_u32(xb), bimg_a = km_smem_u32(bimg);
        const bool leader = elect_one();
        if (my_tiles > 0) {
            mbar_wait_bounded(&bars->b_ready, 0);
            tc_fence_after();
        }
        // Serve whichever group is ready (its next rows staged, its accumulator drained): strict alternation
        // would hold one group's MMAs behind the other group's tracking.
        const uint32_t nj0 = (uint32_t)((my_tiles + 1) / 2), nj1 = (uint32_t)(my_tiles / 2);
        uint32_t j0 = 0, j1 = 0, g = 0, spins = 0;
        unsigned long long wait_t0 = 0;
        while (j0 < nj0 || j1 < nj1) {
            const uint32_t j = g ? j1 : j0, buf = j & 1;
            const bool more = g ? (j1 < nj1) : (j0 < nj0);
            const bool ready = more && mbar_test(&bars->x_ready[g][buf], (j >> 1) & 1) &&
                               mbar_test(&bars->a_empty[g], (j & 1) ^ 1);
            if (!ready) {
                g ^= 1;
                if (((++spins) & 0xFFFFu) == 0 && km_wait_expired(wait_t0)) __trap();
                continue;
            }
            spins = 0;
            wait_t0 = 0;
            tc_fence_after();
            TC_TRACE(2 + (int)g, j, 0);
            const uint32_t ah = (((xb_a + (g * 2 + buf) * xbuf_bytes) & 0x3FFFFu) >> 4) | a_lbo;
            const uint32_t al = ah + (xhi_bytes >> 4);
            const uint32_t dcol = tmem_base + g * (uint32_t)KM_TC_BN;
            if (leader) {
                const uint32_t baug = ((bimg_a & 0x3FFFFu) >> 4) | b_lbo;
                mma_tf32(dcol, ah + (uint32_t)nch * 128u, baug, desc_hi, idesc, 0u);   // + ||c||^2, overwrites
                for (int b = 0; b < nks; ++b) {
                    const uint32_t bh =
                        (((bimg_a + KM_TC_AUG_BYTES + (uint32_t)b * KM_TC_KS_BYTES) & 0x3FFFFu) >> 4) | b_lbo;
                    const uint32_t bl = bh + ((KM_TC_KS_BYTES / 2) >> 4);
                    const uint32_t ao = (uint32_t)b * 256u;
                    mma_tf32(dcol, al + ao, bh, desc_hi, idesc, 1u);
                    mma_tf32(dcol, ah + ao, bl, desc_hi, idesc, 1u);
                    mma_tf32(dcol, ah + ao, bh, desc_hi, idesc, 1u);
                }
                mma_commit(&bars->a_full[g]);
                mma_commit(&bars->x_free[g][buf]);
            }
            __syncwarp();
            if (g) ++j1; else ++j0;
            g ^= 1;
        }
    } else {
        // ================= the two row groups (warps 2..5 and 6..9) =================
        // Tracking and outputs: one thread per row (row rl of the tile == TMEM lane).  Staging: the warp's 32 rows
        // are loaded cooperatively -- lane (r8, cq) takes rows r8, r8+8, r8+16, r8+24 and the 16-byte chunks cq and
        // cq+4 -- so a load instruction touches 8 rows x 64 contiguous bytes instead of 32 separate lines, and a
        // quarter-warp's shared-memory stores are 128 contiguous bytes.
        const int g = (warp - 2) >> 2;
        const int q4 = warp & 3;                 // TMEM lane quarter this warp may access
        const int rl = q4 * 32 + lane;           // row within the tile == TMEM lane
        const int r8 = lane & 7, cq = lane >> 3;
        const float cmax = __uint_as_float(*cmax_bits);
        const float bound_scale = 2.0f * (float)(d + 14) * 1.1920929e-07f;
        // threshold tracker: BIG = 2^(30 - exponent of cmax) resolves 2^-30 of the data's scale (far below a score's
        // own ulp), pad_abs = 2^-28 of that scale keeps every score within the true bound a FULL hit (u == 1)
        int big_e = 284 - (int)((*cmax_bits >> 23) & 0xFFu);
        big_e = big_e < 1 ? 1 : (big_e > 227 ? 227 : big_e);
        const float big = __uint_as_float((uint32_t)big_e << 23);
        const float nbig = -big;
        const float pad_abs = cmax * 3.7252903e-09f;
        // the centring vector's two chunks this lane stages (mu is zero padded to kpad)
        const float4 mu0 = cq < nch ? __ldg(reinterpret_cast<const float4 *>(mu) + cq) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 mu1 =
            4 + cq < nch ? __ldg(reinterpret_cast<const float4 *>(mu) + 4 + cq) : make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t nj = my_tiles > g ? (my_tiles - g + 1) / 2 : 0;      // this group's steps: li = 2j + g
        auto tile_row0 = [&](int64_t j) { return ((int64_t)blockIdx.x + (2 * j + g) * (int64_t)gridDim.x) * TC_ROWS; };
        using LdRegs = float4[NCH];
        auto issue_x = [&](LdRegs &ld, int64_t j) {
            const int64_t base = tile_row0(j) + q4 * 32 + r8;
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                const int64_t row = base + 8 * (i & 3);
                const int c = 4 * (i >> 2) + cq;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c < nch && row < n) {
                    const float *src = X + row * ldx;
                    if (VEC) {
                        if (c * 4 < d) v = __ldg(reinterpret_cast<const float4 *>(src) + c);
                    } else {
                        const int f = c * 4;
                        v.x = f + 0 < d ? __ldg(src + f + 0) : 0.f;
                        v.y = f + 1 < d ? __ldg(src + f + 1) : 0.f;
                        v.z = f + 2 < d ? __ldg(src + f + 2) : 0.f;
                        v.w = f + 3 < d ? __ldg(src + f + 3) : 0.f;
                    }
                }
                ld[i] = v;
            }
        };
        // split the pieces in ld[] into TF32 hi / exact lo parts in this group's buffer j & 1
        auto stage = [&](const LdRegs &ld, int64_t j) {
            const uint32_t bf = (uint32_t)(j & 1), use = (uint32_t)(j >> 1);
            if (use > 0) mbar_wait_bounded(&bars->x_free[g][bf], (use - 1) & 1);   // tensor core done with it
            __syncwarp();             // every lane is done with this buffer's previous rows (output of step j-2)
            unsigned char *ph = xb + (size_t)(g * 2 + bf) * xbuf_bytes + (size_t)(q4 * 32 + r8) * 16;
            unsigned char *pl = ph + xhi_bytes;
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                const int c = 4 * (i >> 2) + cq;
                if (c < nch) {
                    float4 v = ld[i];
                    const float4 mq = (i >> 2) ? mu1 : mu0;                      // x - mu
                    v.x -= mq.x; v.y -= mq.y; v.z -= mq.z; v.w -= mq.w;
                    float s = s4[i & 3];
                    s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                    s4[i & 3] = s;
                    float4 h, l;
                    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
                    l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;   // exact: h + l == x
                    const size_t o = (size_t)c * 2048 + (size_t)(8 * (i & 3)) * 16;
                    *reinterpret_cast<float4 *>(ph + o) = h;
                    *reinterpret_cast<float4 *>(pl + o) = l;
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {         // row norms: the row's chunks sit in the 4 lanes with the same r8
                float s = s4[q];
                s += __shfl_xor_sync(0xffffffffu, s, 8);
                s += __shfl_xor_sync(0xffffffffu, s, 16);
                if (cq == 0) bars->xn[g][bf][q4 * 32 + 8 * q + r8] = s;
            }
            km_fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
            mbar_arrive(&bars->x_ready[g][bf]);
            __syncwarp();             // xn[] written by other lanes of this warp is read in output()
        };
        // labels / winner distance / near-tie flag of row rl of step j.  The winner's centre row comes from the
        // resident image: -(hi + lo) / 2 is the float32 centre exactly (lo is stored unrounded).
        auto output = [&](int64_t j, bool near_tie, int bi, float best_score) {
            const uint32_t bf = (uint32_t)(j & 1);
            const unsigned char *ph = xb + (size_t)(g * 2 + bf) * xbuf_bytes + (size_t)rl * 16;
            const unsigned char *pl = ph + xhi_bytes;
            const unsigned char *pc = bimg + KM_TC_AUG_BYTES + (size_t)bi * 16;
            float s = 0.f;
            if (sparse_d2) {
                s = fmaxf(bars->xn[g][bf][rl] + best_score, 0.f);
            } else if (mind2) {
#pragma unroll
              for (int c = 0; c < NCH; ++c) {
                if (c < nch) {
                    const float4 h = *reinterpret_cast<const float4 *>(ph + (size_t)c * 2048);
                    const float4 l = *reinterpret_cast<const float4 *>(pl + (size_t)c * 2048);
                    const unsigned char *pk = pc + (size_t)(c >> 1) * KM_TC_KS_BYTES + (size_t)(c & 1) * (KM_TC_BN * 16);
                    const float4 ch = *reinterpret_cast<const float4 *>(pk);
                    const float4 cl = *reinterpret_cast<const float4 *>(pk + KM_TC_KS_BYTES / 2);
                    float df;
                    df = fmaf(0.5f, ch.x + cl.x, h.x + l.x); s = fmaf(df, df, s);
                    df = fmaf(0.5f, ch.y + cl.y, h.y + l.y); s = fmaf(df, df, s);
                    df = fmaf(0.5f, ch.z + cl.z, h.z + l.z); s = fmaf(df, df, s);
                    df = fmaf(0.5f, ch.w + cl.w, h.w + l.w); s = fmaf(df, df, s);
                }
              }
            }
            const int64_t row = tile_row0(j) + rl;
            if (row < n) {
                labels[row] = bi;
                if (mind2 && (!sparse_d2 || near_tie)) mind2[row] = s;
                if (flag_rows && near_tie) {
                    const int pos = atomicAdd(flag_count, 1);
                    flag_rows[pos] = (int32_t)row;
                }
            }
        };

        LdRegs ld;
        if (nj > 0) {
            issue_x(ld, 0);
            stage(ld, 0);
            if (nj > 1) issue_x(ld, 1);
        }
        const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(g * KM_TC_BN);
        for (int64_t j = 0; j < nj; ++j) {
            if (q4 == 2) TC_TRACE(g, j, 1);
            if (j + 1 < nj) stage(ld, j + 1);       // its MMAs run under the tracking below
            if (j + 2 < nj) issue_x(ld, j + 2);
            if (q4 == 2) TC_TRACE(g, j, 3);
            // ---- scores of step j: all 256 columns of this group's accumulator stage (threshold tracker) ----
            const float thr = bound_scale * (bars->xn[g][j & 1][rl] + cmax);
            const float thrp = fmaf(thr, 1.0001f, pad_abs);     // padded upward: can only flag more rows
            float m = INFINITY, TB = 0.f;
            float cnt0 = 0.f, cnt1 = 0.f, ix0 = 0.f, ix1 = 0.f;
            mbar_wait_bounded(&bars->a_full[g], (uint32_t)(j & 1));
            if (q4 == 2) TC_TRACE(g, j, 2);
            // The groups track strictly in turn (0, 1, 0, 1, ...).  Left alone they fall into lockstep -- both
            // tracking at once on shared pipes, then both idle while their MMAs queue on the tensor core.
            if (g == 1) mbar_wait_bounded(&bars->turn[1], (uint32_t)(j & 1));
            else if (j > 0) mbar_wait_bounded(&bars->turn[0], (uint32_t)((j - 1) & 1));
            tc_fence_after();
            if (q4 == 2) TC_TRACE(g, j, 4);
            uint32_t ra[32], rb[32];
            tmem_ld32(tbase, ra);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t (&cur)[32] = (c & 1) ? rb : ra;
                uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
                if (c + 1 < 8) tmem_ld32(tbase + (uint32_t)((c + 1) * 32), nxt);  // flies under this chunk
                // The step is paced by the TMEM reads (a warp's 32 x 32 floats take ~256 cycles), so the token goes
                // to the other group as soon as this group's LAST read is queued: its first read then queues right
                // behind instead of waiting for this group's arithmetic on the last chunks.
                if (c == TC_SMALL_TOKEN_AT) mbar_arrive(&bars->turn[g ^ 1]);
                const int cb = c * 32;
                if (cb < k) {               // chunks made of padding columns only cannot win (warp-uniform)
                    // pass 1: the chunk's minimum
                    float cm0 = __uint_as_float(cur[0]), cm1 = __uint_as_float(cur[1]);
#pragma unroll
                    for (int jj = 2; jj < 32; jj += 4) {
                        cm0 = min3f(cm0, __uint_as_float(cur[jj]), __uint_as_float(cur[jj + 1]));
                        if (jj + 3 < 32) cm1 = min3f(cm1, __uint_as_float(cur[jj + 2]), __uint_as_float(cur[jj + 3]));
                    }
                    const float m_old = m;
                    m = min3f(m, cm0, cm1);
                    float T = m + thrp;
                    T = fmaf(fabsf(T), 4.8e-7f, T);          // covers the rounding of the sum above
                    // Earlier hits all lie at or above the old minimum.  If that is now outside the threshold they
                    // are void; if it is inside, the row has two scores within the threshold and the stale count
                    // (>= 2 with the new minimum's own hit) says so.
                    const bool reset = m_old > T;
                    TB = T * big;
                    cnt0 = reset ? 0.f : cnt0;
                    cnt1 = reset ? 0.f : cnt1;
                    ix0 = reset ? 0.f : ix0;
                    ix1 = reset ? 0.f : ix1;
                    // pass 2: hits of this chunk (FMA pipe only)
#pragma unroll
                    for (int jj = 0; jj < 32; jj += 2) {
                        const float u0 = fma_sat(__uint_as_float(cur[jj]), nbig, TB);
                        const float u1 = fma_sat(__uint_as_float(cur[jj + 1]), nbig, TB);
                        cnt0 += u0;
                        cnt1 += u1;
                        ix0 = fmaf(u0, (float)(cb + jj), ix0);
                        ix1 = fmaf(u1, (float)(cb + jj + 1), ix1);
                    }
                }
                if (c + 1 < 8) tmem_ld_wait();
            }
            const float cnt = cnt0 + cnt1;
            const bool near_tie = !(cnt == 1.0f);        // also catches a NaN score
            int bi = (int)(ix0 + ix1);
            // (sparse_d2: the re-decision that follows looks at scores only, never at a flagged row's provisional label)
            if (!sparse_d2 && __any_sync(0xffffffffu, near_tie)) {
                // exact provisional label for the rows that need one: lowest column whose score equals the minimum
                // (ArgKmin's strict `<`); the accumulator stage is still ours, so it is simply read again
                int bx = 0;
                tmem_ld32(tbase + 7u * 32u, ra);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    uint32_t (&cur)[32] = (c & 1) ? rb : ra;
                    uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
                    if (c + 1 < 8) tmem_ld32(tbase + (uint32_t)((7 - (c + 1)) * 32), nxt);
                    const int cb = (7 - c) * 32;            // chunks from the top down: the last match is the lowest column
                    if (cb < k) {
#pragma unroll
                        for (int jj = 31; jj >= 0; --jj) bx = (__uint_as_float(cur[jj]) == m) ? cb + jj : bx;
                    }
                    if (c + 1 < 8) tmem_ld_wait();
                }
                if (near_tie) bi = bx;
            }
            tc_fence_before();
            mbar_arrive(&bars->a_empty[g]);
            if (q4 == 2) TC_TRACE(g, j, 5);
            bi = bi < 0 ? 0 : bi;
            output(j, near_tie, bi >= k ? k - 1 : bi, m);
            if (q4 == 2) TC_TRACE(g, j, 6);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// candidate kernel: which centres can still win a flagged row?
// ---------------------------------------------------------------------------------------------
// Second, small tensor-core pass over the rows K1 flagged (0.2-0.5 % of the rows).  A flagged row's float64
// winner has a float32 score within the rounding bound of the row's best score, so instead of re-scanning all k
// centres in float64 the row's scores are recomputed here (same 3xTF32 MMAs, rows gathered through flag_rows) and
// every column with  score <= (mind2 - ||x'||^2) + 2 * bound  is listed: typically 2-3 columns.  The float64
// re-decision then looks at those only (refine_cand_kernel).  (mind2 - ||x'||^2 is the winner's score up to the
// bound, so the threshold needs nothing stored by K1.)
// Scratch: the unused tail of flag_rows (n entries, *flag_count used), laid out by capacity M = (n-1) / (CAP + 3):
//   [0, M) row indices | [M, M + M*(CAP+1)) per flagged row: count, CAP columns | then the overflow list | [n-1] its size
// Runs only when *flag_count <= M; otherwise the full re-decision kernel does the work.
// (capacity of a row's list: km_cand_cap in common.cuh)

struct TcCandBars {
    uint64_t b_full[5];
    uint64_t b_empty[5];
    uint64_t a_full[TC_ASTAGES];
    uint64_t a_empty[TC_ASTAGES];
    uint64_t x_ready;
    uint64_t x_free;
    uint64_t l_done;               // both column halves of a step have finished listing
    uint32_t tmem_base;
    uint32_t pad;
    alignas(16) float mu[KM_TC_MAX_D];
    float xn_part[2][128];
    int ccnt[128];
};

template <bool VEC>
__global__ void __launch_bounds__(TC_THREADS, 1)
cand_tc_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx, const unsigned char *__restrict__ tc_img,
               const float *__restrict__ mu, const unsigned int *__restrict__ cmax_bits, int k,
               int32_t *__restrict__ F, int64_t flen, const int32_t *__restrict__ flag_count,
               const float *__restrict__ mind2) {
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    const int64_t M = km_cand_rows_max(flen);
    const int64_t cnt = (int64_t)(*flag_count);
    if (cnt > M) return;                                   // uniform: the full re-decision handles this call
    if (blockIdx.x == 0 && threadIdx.x == 0) F[flen - 1] = 0; // size of the overflow list
    if (cnt <= 0) return;
    const int cap = km_cand_cap(flen, cnt);
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;
    const int nks = kpad / 8;
    const uint32_t xhi_bytes = (uint32_t)(nch + 2) * 128 * 16;
    const uint32_t xlo_bytes = (uint32_t)nch * 128 * 16;
    const uint32_t xbuf_bytes = xhi_bytes + xlo_bytes;
    const uint32_t tile_bytes = (uint32_t)km_tc_tile_bytes(d);
    constexpr int NST = 5;
    constexpr uint32_t STAGE = KM_TC_KS_BYTES, AUG = KM_TC_AUG_BYTES;
    unsigned char *xb = tc_smem;
    unsigned char *bst = xb + xbuf_bytes;
    TcCandBars *bars = reinterpret_cast<TcCandBars *>(bst + NST * STAGE);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    const int64_t nsteps = (cnt + TC_ROWS - 1) / TC_ROWS;
    int32_t *cand = F + M;

    if (tid == 0) {
        for (int i = 0; i < NST; ++i) { km_mbar_init(&bars->b_full[i], 1); km_mbar_init(&bars->b_empty[i], 1); }
        for (int i = 0; i < TC_ASTAGES; ++i) { km_mbar_init(&bars->a_full[i], 1); km_mbar_init(&bars->a_empty[i], TC_EPI); }
        km_mbar_init(&bars->x_ready, TC_EPI);
        km_mbar_init(&bars->x_free, 1);
        km_mbar_init(&bars->l_done, TC_EPI);
        km_fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         km_smem_u32(&bars->tmem_base)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int f = tid; f < kpad; f += TC_THREADS) bars->mu[f] = __ldg(mu + f);
    for (int e = tid; e < 2 * 128; e += TC_THREADS) {
        const int c = e / 128, r = e & 127;
        float4 v = c == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(xb + ((size_t)(nch + c) * 128 + r) * 16) = v;
    }
    km_fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x)
                for (int t = 0; t < ntc; ++t) {
                    const unsigned char *src = tc_img + (size_t)t * tile_bytes;
                    for (int b = 0; b <= nks; ++b, ++it) {
                        const int s = it % NST;
                        const uint32_t bytes = b == 0 ? AUG : STAGE;
                        mbar_wait_bounded(&bars->b_empty[s], ((it / NST) & 1) ^ 1);
                        km_mbar_expect_tx(&bars->b_full[s], bytes);
                        km_bulk_g2s(bst + (size_t)s * STAGE, src, bytes, &bars->b_full[s]);
                        src += bytes;
                    }
                }
        }
    } else if (warp == 1) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KM_TC_BN >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);
        const uint32_t a_lbo = (2048u >> 4) << 16;
        const uint32_t b_lbo = ((KM_TC_BN * 16) >> 4) << 16;
        const uint32_t xb_a = km_smem_u32(xb), bst_a = km_smem_u32(bst);
        const bool leader = elect_one();
        uint32_t it = 0, ita = 0, li = 0;
        for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x, ++li) {
            mbar_wait_bounded(&bars->x_ready, li & 1);
            tc_fence_after();
            const uint32_t ah = ((xb_a & 0x3FFFFu) >> 4) | a_lbo;
            const uint32_t al = ah + (xhi_bytes >> 4);
            for (int t = 0; t < ntc; ++t, ++ita) {
                const int a = ita % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_empty[a], ((ita / TC_ASTAGES) & 1) ^ 1);
                const uint32_t dcol = tmem_base + (uint32_t)(a * KM_TC_BN);
                for (int b = 0; b <= nks; ++b, ++it) {
                    const int s = it % NST;
                    mbar_wait_bounded(&bars->b_full[s], (it / NST) & 1);
                    tc_fence_after();
                    const uint32_t bh = (((bst_a + s * STAGE) & 0x3FFFFu) >> 4) | b_lbo;
                    if (leader) {
                        const uint32_t ao = (uint32_t)(b - 1) * 256u;
                        const uint32_t bl = bh + ((STAGE / 2) >> 4);
                        if (b == 0) {
                            mma_tf32(dcol, ah + (uint32_t)nch * 128u, bh, desc_hi, idesc, 0u);
                        } else {
                            mma_tf32(dcol, al + ao, bh, desc_hi, idesc, 1u);
                            mma_tf32(dcol, ah + ao, bl, desc_hi, idesc, 1u);
                            mma_tf32(dcol, ah + ao, bh, desc_hi, idesc, 1u);
                        }
                        mma_commit(&bars->b_empty[s]);
                        if (b == nks) mma_commit(&bars->a_full[a]);
                    }
                    __syncwarp();
                }
            }
            if (leader) mma_commit(&bars->x_free);
            __syncwarp();
        }
    } else {
        const int g = (warp - 2) >> 2;
        const int q4 = warp & 3;
        const int rl = q4 * 32 + lane;
        const float cmax = __uint_as_float(*cmax_bits);
        const float bound_scale = 2.0f * (float)(d + 14) * 1.1920929e-07f;
        const int nch_h = nch / 2, c_lo = g * nch_h;
        uint32_t it = 0, li = 0;
        for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x, ++li) {
            const int64_t pos = step * TC_ROWS + rl;
            const bool live = pos < cnt;
            const int64_t row = live ? (int64_t)__ldg(F + pos) : 0;
            const float *src = X + row * ldx;
            if (li > 0) mbar_wait_bounded(&bars->x_free, (li - 1) & 1);    // tensor core done with the previous rows
            if (g == 0) bars->ccnt[rl] = 0;
            {
                unsigned char *ph = xb + (size_t)rl * 16;
                unsigned char *pl = ph + xhi_bytes;
                float s = 0.f;
                for (int c = 0; c < nch_h; ++c) {
                    const int cc = c_lo + c;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (live) {
                        if (VEC) {
                            if (cc * 4 < d) v = __ldg(reinterpret_cast<const float4 *>(src) + cc);
                        } else {
                            const int f = cc * 4;
                            v.x = f + 0 < d ? __ldg(src + f + 0) : 0.f;
                            v.y = f + 1 < d ? __ldg(src + f + 1) : 0.f;
                            v.z = f + 2 < d ? __ldg(src + f + 2) : 0.f;
                            v.w = f + 3 < d ? __ldg(src + f + 3) : 0.f;
                        }
                    }
                    const float4 mq = *reinterpret_cast<const float4 *>(&bars->mu[cc * 4]);
                    v.x -= mq.x; v.y -= mq.y; v.z -= mq.z; v.w -= mq.w;
                    s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                    float4 h, l;
                    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
                    l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;
                    *reinterpret_cast<float4 *>(ph + (size_t)cc * 2048) = h;
                    *reinterpret_cast<float4 *>(pl + (size_t)cc * 2048) = l;
                }
                bars->xn_part[g][rl] = s;
                km_fence_proxy_async();
                mbar_arrive(&bars->x_ready);
            }
            // both halves of the row are staged once x_ready completes: the row's threshold
            mbar_wait_bounded(&bars->x_ready, li & 1);
            const float xn = bars->xn_part[0][rl] + bars->xn_part[1][rl];
            const float T = live ? (__ldg(mind2 + row) - xn) + 2.0f * bound_scale * (xn + cmax) : -INFINITY;
            for (int t = 0; t < ntc; ++t, ++it) {
                const int a = it % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_full[a], (it / TC_ASTAGES) & 1);
                tc_fence_after();
                const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(a * KM_TC_BN + g * 128);
                for (int c = 0; c < 4; ++c) {
                    uint32_t r[32];
                    tmem_ld32(tbase + (uint32_t)(c * 32), r);
                    tmem_ld_wait();
                    const int cb = t * KM_TC_BN + g * 128 + c * 32;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (__uint_as_float(r[j]) <= T && cb + j < k) {     // (T = -inf for a row that is not live)
                            const int q = atomicAdd(&bars->ccnt[rl], 1);
                            if (q < cap) cand[pos * (cap + 1) + 1 + q] = cb + j;
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&bars->a_empty[a]);
            }
            mbar_arrive(&bars->l_done);
            mbar_wait_bounded(&bars->l_done, li & 1);
            if (g == 0 && live) cand[pos * (cap + 1)] = bars->ccnt[rl];
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

int km_launch_cand_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, int32_t *flag_rows,
                      int64_t flen, const int32_t *flag_count, const float *mind2, cudaStream_t st, bool precentred) {
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    const int nch = km_tc_kpad(d) / 4;
    const size_t smem = (size_t)(nch + 2) * 128 * 16 + (size_t)nch * 128 * 16 + 5 * (size_t)KM_TC_KS_BYTES +
                        sizeof(TcCandBars) + 64;
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    const unsigned char *img = pk + L.off_tc;
    const float *mu = reinterpret_cast<const float *>(pk + (precentred ? L.off_zero : L.off_mu));
    const unsigned int *cmax = reinterpret_cast<const unsigned int *>(pk + L.off_cmax);
    static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(cand_tc_kernel<true>, 227 * 1024, attr_v)) return rc;
    if (int rc = km_set_smem_once(cand_tc_kernel<false>, 227 * 1024, attr_s)) return rc;
    if (vec)
        cand_tc_kernel<true><<<km_num_sms(), TC_THREADS, smem, st>>>(X, n, d, ldx, img, mu, cmax, k, flag_rows, flen, flag_count, mind2);
    else
        cand_tc_kernel<false><<<km_num_sms(), TC_THREADS, smem, st>>>(X, n, d, ldx, img, mu, cmax, k, flag_rows, flen, flag_count, mind2);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// wide rows: 128 < d <= B200KM_MAX_D
// ---------------------------------------------------------------------------------------------
// A 128-row tile of X no longer fits in shared memory next to the centre ring (hi + lo parts: 8 bytes per element),
// so the K dimension is cut into PANELS of 64 features.  The two X buffers of the d <= 64 layout hold alternate
// panels; for every column tile the staging warps re-stage the step's rows panel by panel (re-read from L2, ntc
// times per step) while the MMA warp accumulates the tile over all panels in TMEM.  Tracking of tile t-1 is slotted
// in after the second panel of tile t has been staged -- by then tile t-1's accumulator is complete and the tensor
// core has two panels queued.  The winner's distance is rebuilt from global memory at the end of the step.
constexpr int TCW_PF = 64;              // features per panel
constexpr int TCW_PCH = TCW_PF / 4;     // 16-byte chunks per panel
constexpr int TCW_PKS = TCW_PF / 8;     // K-steps per panel

// CAND: the candidate pass of the pruned re-decision for wide rows -- the rows are flag_rows[0 .. *flag_count)
// (gathered), the epilogue lists the columns with score <= threshold instead of tracking, nothing else is written
// (see cand_tc_kernel for the threshold and the scratch layout in the tail of flag_rows).
template <bool VEC, bool CAND>
__global__ void __launch_bounds__(TC_THREADS, 1)
assign_tc_wide_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx,
                      const unsigned char *__restrict__ tc_img, const float *__restrict__ c32,
                      const float *__restrict__ mu, const unsigned int *__restrict__ cmax_bits, int k,
                      int32_t *__restrict__ labels, float *__restrict__ mind2,
                      int32_t *__restrict__ flag_rows, int32_t *__restrict__ flag_count, int64_t flen) {
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    int64_t nrows = n;                                          // rows (or flagged rows) to process
    int64_t cand_M = 0;
    int cand_cap = 0;
    if (CAND) {
        cand_M = km_cand_rows_max(flen);
        const int64_t cnt = (int64_t)(*flag_count);
        if (cnt > cand_M) return;                               // uniform: the full re-decision handles this call
        if (blockIdx.x == 0 && threadIdx.x == 0) flag_rows[flen - 1] = 0;   // size of the overflow list
        if (cnt <= 0) return;
        nrows = cnt;
        cand_cap = km_cand_cap(flen, cnt);
    }
    const int kpad = km_tc_kpad(d);
    const int nks = kpad / 8;                                   // K-steps of a whole row
    const int npan = (kpad + TCW_PF - 1) / TCW_PF;              // panels per row (>= 3)
    constexpr uint32_t xhi_bytes = (uint32_t)(TCW_PCH + 2) * 128 * 16;
    constexpr uint32_t xlo_bytes = (uint32_t)TCW_PCH * 128 * 16;
    constexpr uint32_t xbuf_bytes = xhi_bytes + xlo_bytes;
    const uint32_t tile_bytes = (uint32_t)km_tc_tile_bytes(d);
    constexpr int NST = 5;
    constexpr uint32_t STAGE = KM_TC_KS_BYTES, AUG = KM_TC_AUG_BYTES;
    unsigned char *xb = tc_smem;                                // [2][xbuf_bytes]
    unsigned char *bst = xb + 2 * xbuf_bytes;                   // [NST][STAGE]
    TcBars *bars = reinterpret_cast<TcBars *>(bst + NST * STAGE);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    const int64_t nsteps = (nrows + TC_ROWS - 1) / TC_ROWS;

    if (tid == 0) {
        for (int i = 0; i < NST; ++i) { km_mbar_init(&bars->b_full[i], 1); km_mbar_init(&bars->b_empty[i], 1); }
        for (int i = 0; i < TC_ASTAGES; ++i) { km_mbar_init(&bars->a_full[i], 1); km_mbar_init(&bars->a_empty[i], TC_EPI); }
        for (int i = 0; i < 2; ++i) {
            km_mbar_init(&bars->x_ready[i], TC_EPI);
            km_mbar_init(&bars->x_free[i], 1);
            km_mbar_init(&bars->m_ready[i], TC_EPI);
            km_mbar_init(&bars->s_ready[i], TC_EPI / 2);
        }
        km_fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         km_smem_u32(&bars->tmem_base)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // constant augmented chunks of Xhi in both panel buffers: chunk PCH = [1,1,1,0], chunk PCH+1 = 0
    for (int e = tid; e < 2 * 2 * 128; e += TC_THREADS) {
        const int bf = e / 256, c = (e / 128) & 1, r = e & 127;
        float4 v = c == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(xb + (size_t)bf * xbuf_bytes + ((size_t)(TCW_PCH + c) * 128 + r) * 16) = v;
    }
    km_fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================= producer: the centre image, K-step block by K-step block =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x)
                for (int t = 0; t < ntc; ++t) {
                    const unsigned char *src = tc_img + (size_t)t * tile_bytes;
                    for (int b = 0; b <= nks; ++b, ++it) {
                        const int s = it % NST;
                        const uint32_t bytes = b == 0 ? AUG : STAGE;
                        mbar_wait_bounded(&bars->b_empty[s], ((it / NST) & 1) ^ 1);
                        km_mbar_expect_tx(&bars->b_full[s], bytes);
                        km_bulk_g2s(bst + (size_t)s * STAGE, src, bytes, &bars->b_full[s]);
                        src += bytes;
                    }
                }
        }
    } else if (warp == 1) {
        // ================= MMA issuer: a tile accumulates over all panels =================
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KM_TC_BN >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);
        const uint32_t a_lbo = (2048u >> 4) << 16;
        const uint32_t b_lbo = ((KM_TC_BN * 16) >> 4) << 16;
        const uint32_t xb_a = km_smem_u32(xb), bst_a = km_smem_u32(bst);
        const bool leader = elect_one();
        uint32_t it = 0, ita = 0, sq = 0;
        for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x) {
            for (int t = 0; t < ntc; ++t, ++ita) {
                const int a = ita % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_empty[a], ((ita / TC_ASTAGES) & 1) ^ 1);
                const uint32_t dcol = tmem_base + (uint32_t)(a * KM_TC_BN);
                for (int p = 0; p < npan; ++p, ++sq) {
                    const uint32_t buf = sq & 1;
                    mbar_wait_bounded(&bars->x_ready[buf], (sq >> 1) & 1);
                    tc_fence_after();
                    const uint32_t ah = (((xb_a + buf * xbuf_bytes) & 0x3FFFFu) >> 4) | a_lbo;
                    const uint32_t al = ah + (xhi_bytes >> 4);
                    const int ks_p = (nks - p * TCW_PKS) < TCW_PKS ? (nks - p * TCW_PKS) : TCW_PKS;
                    // panel 0 starts with the augmented K-step (block 0 of the tile image)
                    for (int b = (p == 0 ? -1 : 0); b < ks_p; ++b, ++it) {
                        const int s = it % NST;
                        mbar_wait_bounded(&bars->b_full[s], (it / NST) & 1);
                        tc_fence_after();
                        const uint32_t bh = (((bst_a + s * STAGE) & 0x3FFFFu) >> 4) | b_lbo;
                        if (leader) {
                            if (b < 0) {
                                mma_tf32(dcol, ah + (uint32_t)TCW_PCH * 128u, bh, desc_hi, idesc, 0u);
                            } else {
                                const uint32_t ao = (uint32_t)b * 256u;
                                const uint32_t bl = bh + ((STAGE / 2) >> 4);
                                mma_tf32(dcol, al + ao, bh, desc_hi, idesc, 1u);
                                mma_tf32(dcol, ah + ao, bl, desc_hi, idesc, 1u);
                                mma_tf32(dcol, ah + ao, bh, desc_hi, idesc, 1u);
                            }
                            mma_commit(&bars->b_empty[s]);
                        }
                        __syncwarp();
                    }
                    if (leader) {
                        mma_commit(&bars->x_free[buf]);                       // panel buffer reusable
                        if (p == npan - 1) mma_commit(&bars->a_full[a]);      // tile complete
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ================= staging + tracking + outputs (warps 2..9) =================
        // Tracking: one thread per (row, column half): row rl == TMEM lane, half g.  Global loads (staging, and the
        // winner's distance at the end of the step) are cooperative instead: warp (g, q4) covers its 32 rows and the
        // 32 features [64p + 32g, +32) of each panel with lane (r8, cq) taking rows r8, r8+8, r8+16, r8+24 and the
        // 16-byte chunks cq, cq+4 -- a load instruction touches 8 rows x 64 contiguous bytes, not 32 lines.
        const int g = (warp - 2) >> 2;
        const int q4 = warp & 3;
        const int rl = q4 * 32 + lane;
        const int r8 = lane & 7, cq = lane >> 3;
        const float cmax = __uint_as_float(*cmax_bits);
        const float bound_scale = 2.0f * (float)(d + 14) * 1.1920929e-07f;
        constexpr int HCH = TCW_PCH / 2;         // chunks per half panel (8): two per lane and row
        uint32_t it = 0, sq = 0, li = 0;
        for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x, ++li) {
            // the row this thread TRACKS and the 4 rows this lane LOADS (CAND: through the flagged-row list)
            const int64_t pos_t = step * TC_ROWS + rl;
            const bool live_t = pos_t < nrows;
            const int64_t row_t = CAND ? (live_t ? (int64_t)__ldg(flag_rows + pos_t) : 0) : pos_t;
            int64_t row4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t pos = step * TC_ROWS + q4 * 32 + 8 * q + r8;
                row4[q] = pos < nrows ? (CAND ? (int64_t)__ldg(flag_rows + pos) : pos) : -1;
            }
            float cand_T = -INFINITY;
            int *ccnt = reinterpret_cast<int *>(&bars->m_best[0][0][0]);      // CAND: per-row list length
            float best2[2] = {INFINITY, INFINITY}, second2[2] = {INFINITY, INFINITY};
            int bidx2[2] = {0, 0};
            float xn4[4] = {0.f, 0.f, 0.f, 0.f};
            auto track_tile = [&](int t, uint32_t itile) {
                const int a = itile % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_full[a], (itile / TC_ASTAGES) & 1);
                tc_fence_after();
                const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(a * KM_TC_BN + g * 128);
                uint32_t ra[32], rb[32];
                tmem_ld32(tbase, ra);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t (&cur)[32] = (c & 1) ? rb : ra;
                    uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
                    if (c + 1 < 4) tmem_ld32(tbase + (uint32_t)((c + 1) * 32), nxt);
                    const int cb = t * KM_TC_BN + g * 128 + c * 32;
                    if (CAND) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (__uint_as_float(cur[j]) <= cand_T && cb + j < k) {   // (cand_T = -inf: row not live)
                                const int q = atomicAdd(&ccnt[rl], 1);
                                if (q < cand_cap) flag_rows[cand_M + pos_t * (cand_cap + 1) + 1 + q] = cb + j;
                            }
                        }
                    } else if (cb < k) {
                        const float before0 = best2[0], before1 = best2[1];
                        int jloc0 = 0, jloc1 = 0;
#pragma unroll
                        for (int j = 0; j < 32; j += 2) {
                            track(__uint_as_float(cur[j]), j, best2[0], second2[0], jloc0);
                            track(__uint_as_float(cur[j + 1]), j + 1, best2[1], second2[1], jloc1);
                        }
                        bidx2[0] = best2[0] < before0 ? cb + jloc0 : bidx2[0];
                        bidx2[1] = best2[1] < before1 ? cb + jloc1 : bidx2[1];
                    }
                    if (c + 1 < 4) tmem_ld_wait();
                }
                tc_fence_before();
                mbar_arrive(&bars->a_empty[a]);
            };
            // piece i of a half panel: row 8*(i&3) + r8 of the warp's 32, chunk 4*(i>>2) + cq of the half's 8
            auto load_piece = [&](const float *base, int64_t row, int f) -> float4 {
                float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
                if (row >= 0 && f < d) {
                    const float *src = base + f;
                    if (VEC) {
                        x = __ldg(reinterpret_cast<const float4 *>(src));
                    } else {
                        x.x = __ldg(src);
                        x.y = f + 1 < d ? __ldg(src + 1) : 0.f;
                        x.z = f + 2 < d ? __ldg(src + 2) : 0.f;
                        x.w = f + 3 < d ? __ldg(src + 3) : 0.f;
                    }
                }
                return x;
            };
            for (int t = 0; t < ntc; ++t) {
                for (int p = 0; p < npan; ++p, ++sq) {
                    const uint32_t buf = sq & 1, use = sq >> 1;
                    float4 v[HCH];
#pragma unroll
                    for (int i = 0; i < HCH; ++i) {
                        const int64_t row = row4[i & 3];
                        const int f = p * TCW_PF + (g * HCH + 4 * (i >> 2) + cq) * 4;
                        v[i] = load_piece(X + (row >= 0 ? row : 0) * ldx, row, f);
                    }
                    float4 mq[2];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int f = p * TCW_PF + (g * HCH + 4 * h + cq) * 4;
                        mq[h] = f < kpad ? __ldg(reinterpret_cast<const float4 *>(mu + f)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    if (use > 0) mbar_wait_bounded(&bars->x_free[buf], (use - 1) & 1);
                    unsigned char *ph = xb + (size_t)buf * xbuf_bytes + (size_t)(q4 * 32 + r8) * 16;
                    unsigned char *pl = ph + xhi_bytes;
#pragma unroll
                    for (int i = 0; i < HCH; ++i) {
                        const int64_t row = row4[i & 3];
                        const int f = p * TCW_PF + (g * HCH + 4 * (i >> 2) + cq) * 4;
                        float4 x = v[i];
                        const float4 m4 = mq[i >> 2];
                        x.x -= m4.x; x.y -= m4.y; x.z -= m4.z; x.w -= m4.w;
                        if (row < 0 || f >= d) x = make_float4(0.f, 0.f, 0.f, 0.f);
                        else if (f + 3 >= d) {          // padding features of the row's last chunk stay exactly zero
                            if (f + 1 >= d) x.y = 0.f;
                            if (f + 2 >= d) x.z = 0.f;
                            x.w = 0.f;
                        }
                        if (t == 0) {
                            float sx = xn4[i & 3];
                            sx = fmaf(x.x, x.x, sx); sx = fmaf(x.y, x.y, sx); sx = fmaf(x.z, x.z, sx); sx = fmaf(x.w, x.w, sx);
                            xn4[i & 3] = sx;
                        }
                        float4 h, l;
                        h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
                        l.x = x.x - h.x; l.y = x.y - h.y; l.z = x.z - h.z; l.w = x.w - h.w;
                        const size_t o = (size_t)(g * HCH + 4 * (i >> 2) + cq) * 2048 + (size_t)(8 * (i & 3)) * 16;
                        *reinterpret_cast<float4 *>(ph + o) = h;
                        *reinterpret_cast<float4 *>(pl + o) = l;
                    }
                    km_fence_proxy_async();
                    mbar_arrive(&bars->x_ready[buf]);
                    // tile t-1 is complete once this panel's buffer was free again: track it under tile t's MMAs
                    if (p == 1 && t > 0) { track_tile(t - 1, it); ++it; }
                }
                if (CAND && t == 0) {
                    // every panel of the rows has been staged once: the rows' norms, hence their thresholds
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        xn4[q] += __shfl_xor_sync(0xffffffffu, xn4[q], 8);
                        xn4[q] += __shfl_xor_sync(0xffffffffu, xn4[q], 16);
                    }
                    bars->xn_part[0][g][rl] = cq == 0 ? xn4[0] : (cq == 1 ? xn4[1] : (cq == 2 ? xn4[2] : xn4[3]));
                    if (g == 0) ccnt[rl] = 0;
                    mbar_arrive(&bars->m_ready[1]);
                    mbar_wait_bounded(&bars->m_ready[1], li & 1);
                    const float xn_c = bars->xn_part[0][0][rl] + bars->xn_part[0][1][rl];
                    cand_T = live_t ? (__ldg(mind2 + row_t) - xn_c) + 2.0f * bound_scale * (xn_c + cmax) : -INFINITY;
                }
            }
            track_tile(ntc - 1, it); ++it;
            if (CAND) {
                mbar_arrive(&bars->m_ready[0]);                 // both column halves have finished listing
                mbar_wait_bounded(&bars->m_ready[0], li & 1);
                if (g == 0 && live_t) flag_rows[cand_M + pos_t * (cand_cap + 1)] = ccnt[rl];
            } else {
            // ---- merge the two column halves; every thread ends up with its row's winner -------------------
            float xn_half;
            {
#pragma unroll
                for (int q = 0; q < 4; ++q) {       // the row's pieces sit in the 4 lanes with the same r8
                    xn4[q] += __shfl_xor_sync(0xffffffffu, xn4[q], 8);
                    xn4[q] += __shfl_xor_sync(0xffffffffu, xn4[q], 16);
                }
                xn_half = cq == 0 ? xn4[0] : (cq == 1 ? xn4[1] : (cq == 2 ? xn4[2] : xn4[3]));   // row 8*cq + r8 == lane
            }
            const bool takeb = best2[1] < best2[0] || (best2[1] == best2[0] && bidx2[1] < bidx2[0]);
            bars->m_best[0][g][rl] = takeb ? best2[1] : best2[0];
            bars->m_idx[0][g][rl] = takeb ? bidx2[1] : bidx2[0];
            bars->m_second[0][g][rl] = fminf(fminf(second2[0], second2[1]), takeb ? best2[0] : best2[1]);
            bars->xn_part[0][g][rl] = xn_half;
            mbar_arrive(&bars->m_ready[0]);
            mbar_wait_bounded(&bars->m_ready[0], li & 1);
            const float b0 = bars->m_best[0][0][rl], b1 = bars->m_best[0][1][rl];
            const float s0 = bars->m_second[0][0][rl], s1 = bars->m_second[0][1][rl];
            const int i0 = bars->m_idx[0][0][rl], i1 = bars->m_idx[0][1][rl];
            const bool take1 = b1 < b0 || (b1 == b0 && i1 < i0);
            const float p_best = take1 ? b1 : b0;
            int bi = take1 ? i1 : i0;
            const float p_second = fminf(fminf(s0, s1), take1 ? b0 : b1);
            bi = bi >= k ? k - 1 : bi;
            const float xn = bars->xn_part[0][0][rl] + bars->xn_part[0][1][rl];
            // ---- the winner's distance in direct form, from global memory, with the cooperative load pattern ---
            float d4[4] = {0.f, 0.f, 0.f, 0.f};
            int bi4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) bi4[q] = __shfl_sync(0xffffffffu, bi, 8 * q + r8);
            for (int p = 0; p < npan; ++p) {
                float4 xv[HCH], cv[HCH];
#pragma unroll
                for (int i = 0; i < HCH; ++i) {
                    const int64_t row = row4[i & 3];
                    const int f = p * TCW_PF + (g * HCH + 4 * (i >> 2) + cq) * 4;
                    xv[i] = load_piece(X + (row >= 0 ? row : 0) * ldx, row, f);
                    cv[i] = load_piece(c32 + (size_t)bi4[i & 3] * d, row, f);
                }
#pragma unroll
                for (int i = 0; i < HCH; ++i) {
                    const int f = p * TCW_PF + (g * HCH + 4 * (i >> 2) + cq) * 4;
                    if (f < d) {
                        const float4 m4 = __ldg(reinterpret_cast<const float4 *>(mu + f));
                        float sx = d4[i & 3];
                        float df;
                        df = (xv[i].x - m4.x) - cv[i].x; sx = fmaf(df, df, sx);
                        if (f + 1 < d) { df = (xv[i].y - m4.y) - cv[i].y; sx = fmaf(df, df, sx); }
                        if (f + 2 < d) { df = (xv[i].z - m4.z) - cv[i].z; sx = fmaf(df, df, sx); }
                        if (f + 3 < d) { df = (xv[i].w - m4.w) - cv[i].w; sx = fmaf(df, df, sx); }
                        d4[i & 3] = sx;
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                d4[q] += __shfl_xor_sync(0xffffffffu, d4[q], 8);
                d4[q] += __shfl_xor_sync(0xffffffffu, d4[q], 16);
            }
            float s = cq == 0 ? d4[0] : (cq == 1 ? d4[1] : (cq == 2 ? d4[2] : d4[3]));   // this thread's row, this half
            if (g == 1) {
                bars->s_part[0][rl] = s;
                mbar_arrive(&bars->s_ready[0]);
            } else {
                mbar_wait_bounded(&bars->s_ready[0], li & 1);
                s += bars->s_part[0][rl];
                if (live_t) {
                    labels[row_t] = bi;
                    if (mind2) mind2[row_t] = s;
                    if (flag_rows) {
                        const float thr = bound_scale * (xn + cmax);
                        if (p_second - p_best <= thr) {
                            const int pos = atomicAdd(flag_count, 1);
                            flag_rows[pos] = (int32_t)row_t;
                        }
                    }
                }
            }
            }   // !CAND
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

int km_launch_cand_tc_wide(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                           int32_t *flag_rows, int64_t flen, const int32_t *flag_count, float *mind2, cudaStream_t st,
                           bool precentred) {
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    const size_t xbuf_w = (size_t)(TCW_PCH + 2) * 128 * 16 + (size_t)TCW_PCH * 128 * 16;
    const size_t smem_w = 2 * xbuf_w + 5 * (size_t)KM_TC_KS_BYTES + sizeof(TcBars) + 64;
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    const unsigned char *img = pk + L.off_tc;
    const float *c32 = reinterpret_cast<const float *>(pk + L.off_c32);
    const float *mu = reinterpret_cast<const float *>(pk + (precentred ? L.off_zero : L.off_mu));
    const unsigned int *cmax = reinterpret_cast<const unsigned int *>(pk + L.off_cmax);
    static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(assign_tc_wide_kernel<true, true>, 227 * 1024, attr_v)) return rc;
    if (int rc = km_set_smem_once(assign_tc_wide_kernel<false, true>, 227 * 1024, attr_s)) return rc;
    if (vec)
        assign_tc_wide_kernel<true, true><<<km_num_sms(), TC_THREADS, smem_w, st>>>(
            X, n, d, ldx, img, c32, mu, cmax, k, nullptr, mind2, flag_rows, const_cast<int32_t *>(flag_count), flen);
    else
        assign_tc_wide_kernel<false, true><<<km_num_sms(), TC_THREADS, smem_w, st>>>(
            X, n, d, ldx, img, c32, mu, cmax, k, nullptr, mind2, flag_rows, const_cast<int32_t *>(flag_count), flen);
    return (int)cudaGetLastError();
}

// host-side launcher, called from b200km_assign_f32
int km_launch_assign_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, int32_t *labels,
                        float *mind2, int32_t *flag_rows, int32_t *flag_count, cudaStream_t st, bool precentred,
                        bool sparse_d2) {
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;
    const bool dbuf = d <= 64;
    const bool pair = km_tc_pair_mode();
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    const unsigned char *img = pk + L.off_tc;
    const float *c32 = reinterpret_cast<const float *>(pk + L.off_c32);
    const unsigned int *cmax = reinterpret_cast<const unsigned int *>(pk + L.off_cmax);
    const float *mu = reinterpret_cast<const float *>(pk + (precentred ? L.off_zero : L.off_mu));
    if (d > KM_TC_MAX_D) {
        // wide rows: K cut into 64-feature panels (assign_tc_wide_kernel); same shared-memory footprint as d = 64
        const size_t xbuf_w = (size_t)(TCW_PCH + 2) * 128 * 16 + (size_t)TCW_PCH * 128 * 16;
        const size_t smem_w = 2 * xbuf_w + 5 * (size_t)KM_TC_KS_BYTES + sizeof(TcBars) + 64;
        const int64_t steps_w = (n + TC_ROWS - 1) / TC_ROWS;
        const unsigned grid = (unsigned)(steps_w < km_num_sms() ? steps_w : km_num_sms());
        static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(assign_tc_wide_kernel<true, false>, 227 * 1024, attr_v)) return rc;
        if (int rc = km_set_smem_once(assign_tc_wide_kernel<false, false>, 227 * 1024, attr_s)) return rc;
        if (vec)
            assign_tc_wide_kernel<true, false><<<grid, TC_THREADS, smem_w, st>>>(X, n, d, ldx, img, c32, mu, cmax, k, labels, mind2,
                                                                         flag_rows, flag_count, 0);
        else
            assign_tc_wide_kernel<false, false><<<grid, TC_THREADS, smem_w, st>>>(X, n, d, ldx, img, c32, mu, cmax, k, labels, mind2,
                                                                          flag_rows, flag_count, 0);
        return (int)cudaGetLastError();
    }
    if (!pair && k <= KM_TC_BN && d <= 32 && km_tc_small_enabled()) {
        // one column tile, resident centre image, two row groups (assign_tc_small_kernel)
        const size_t xbuf = (size_t)(nch + 2) * 128 * 16 + (size_t)nch * 128 * 16;
        const size_t smem_s = 4 * xbuf + km_tc_tile_bytes(d) + sizeof(TcSmallBars) + 64;
        const int64_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
        const unsigned grid = (unsigned)(tiles < km_num_sms() ? tiles : km_num_sms());
        static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(assign_tc_small_kernel<true>, 227 * 1024, attr_v)) return rc;
        if (int rc = km_set_smem_once(assign_tc_small_kernel<false>, 227 * 1024, attr_s)) return rc;
        if (vec)
            assign_tc_small_kernel<true><<<grid, TC_THREADS, smem_s, st>>>(X, n, d, ldx, img, c32, mu, cmax, k, labels, mind2,
                                                                          flag_rows, flag_count, sparse_d2 ? 1 : 0);
        else
            assign_tc_small_kernel<false><<<grid, TC_THREADS, smem_s, st>>>(X, n, d, ldx, img, c32, mu, cmax, k, labels, mind2,
                                                                           flag_rows, flag_count, sparse_d2 ? 1 : 0);
        return (int)cudaGetLastError();
    }
    const size_t stage_bytes = pair ? 8192 : 16384;
    const int nstages = pair ? 8 : 5;
    const size_t smem = (dbuf ? 2 : 1) * ((size_t)(nch + 2) * 128 * 16 + (size_t)nch * 128 * 16) +
                        (size_t)nstages * stage_bytes + sizeof(TcBars) + 64;
    const int64_t step_rows = pair ? 2 * TC_ROWS : TC_ROWS;
    int64_t steps = (n + step_rows - 1) / step_rows;
    const int max_units = pair ? km_num_sms() / 2 : km_num_sms();
    const int units = (int)(steps < max_units ? steps : max_units);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pair ? 2 * units : units);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr1[1];
    attr1[0].id = cudaLaunchAttributeClusterDimension;
    attr1[0].val.clusterDim.x = 2;
    attr1[0].val.clusterDim.y = 1;
    attr1[0].val.clusterDim.z = 1;
    cfg.attrs = attr1;
    cfg.numAttrs = pair ? 1 : 0;
    const int sparse_i = (sparse_d2 && flag_rows) ? 1 : 0;
#define TC_LAUNCH(VEC_, DBUF_, PAIR_)                                                                      \
    {                                                                                                       \
        auto kern = assign_tc_f32_kernel<VEC_, DBUF_, PAIR_>;                                               \
        static bool attr[KM_MAX_DEVICES] = {};                                                              \
        if (int rc = km_set_smem_once(kern, 227 * 1024, attr)) return rc;                                   \
        cudaError_t le = cudaLaunchKernelEx(&cfg, kern, X, n, d, ldx, img, c32, mu, cmax, k, labels, mind2,     \
                                            flag_rows, flag_count, sparse_i);                               \
        if (le != cudaSuccess) return (int)le;                                                              \
    }
#define TC_LAUNCH_VD(PAIR_)                     \
    if (vec && dbuf) TC_LAUNCH(true, true, PAIR_) \
    else if (vec) TC_LAUNCH(true, false, PAIR_)   \
    else if (dbuf) TC_LAUNCH(false, true, PAIR_)  \
    else TC_LAUNCH(false, false, PAIR_)
    if (pair) { TC_LAUNCH_VD(true) } else { TC_LAUNCH_VD(false) }
#undef TC_LAUNCH_VD
#undef TC_LAUNCH
    return (int)cudaGetLastError();
}

// assign_tc.cu -- K1 on the 5th-generation tensor cores: 3xTF32 tcgen05.mma with TMEM accumulators.
//
// Same contract as assign_tiled_f32_kernel (labels, winner distance, near-tie flags), for d <= 64.
// Replaces dask_ml.metrics.pairwise_distances_argmin_min (reference dask_ml/metrics/pairwise.py:19-47)
// on the Lloyd path (dask_ml/cluster/k_means.py:576-579) when k*d makes the FFMA pipe the bound.
//
// Score matrix  S = [x_hi | x_lo | 1 1 1] . [ (-2c)_hi | (-2c)_lo | cn_h cn_m cn_l ]^T  computed as
//   D  = Xhi.Bhi^T + Xhi.Blo^T + Xlo.Bhi^T   (error-compensated TF32 split, fp32 accumulate in TMEM)
//      + Xaug.Baug^T                         (one extra K-step that adds ||c||^2 exactly)
// so the epilogue only has to track (best, second, argmin) per row -- no per-element norm add.
//
// CTA = 192 threads, persistent, one per SM, 256 rows (two M=128 row tiles) per step:
//   warp 0      : producer -- one 1-D bulk async copy (TMA, mbarrier complete_tx) per 64-centroid stage
//   warp 1      : TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=64, K=8, kind::tf32)
//   warps 2..5  : X staging (LDG.128 -> cvt.rna.tf32 split -> STS in the UMMA canonical no-swizzle
//                 K-major layout) and the epilogue (tcgen05.ld 32x32b.x32 -> argmin tracking)
// Pipelines: B ring (2 smem stages), accumulator ring (4 TMEM stages of 2x64 columns).
// Both row tiles share every B stage, which halves the L2 -> SMEM centroid traffic per row.
#include "common.cuh"

namespace {

constexpr int TC_THREADS = 192;
constexpr int TC_ROWS = 256;     // rows per CTA step (2 row tiles of 128)
constexpr int TC_BSTAGES = 2;
constexpr int TC_ASTAGES = 4;    // accumulator stages: 4 x (2 x 64) = 512 TMEM columns
constexpr uint32_t TC_SPIN_LIMIT = 1u << 27;

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(km_smem_u32(bar)) : "memory");
}
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_bounded(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = km_smem_u32(bar);
    uint32_t done;
    uint32_t spins = 0;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (!done && ++spins > TC_SPIN_LIMIT) __trap();
    } while (!done);
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    // no-swizzle ("interleave") K-major canonical layout: 8-row x 16-byte core matrices;
    // LBO = byte distance between the two 16-byte K chunks of one MMA, SBO = between 8-row groups.
    const uint32_t lo = ((saddr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
    const uint32_t hi = ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14);  // version = 1 (Blackwell)
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u), "r"(0u), "r"(0u), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(km_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    return __uint_as_float(u);
}

__device__ __forceinline__ void track(float v, int col, float &best, float &second, int &bidx) {
    const bool lt = v < best;
    second = lt ? best : fminf(second, v);
    bidx = lt ? col : bidx;
    best = fminf(best, v);
}

struct TcBars {
    uint64_t b_full[TC_BSTAGES];
    uint64_t b_empty[TC_BSTAGES];
    uint64_t a_full[TC_ASTAGES];
    uint64_t a_empty[TC_ASTAGES];
    uint64_t x_ready;
    uint64_t x_free;
    uint32_t tmem_base;
    uint32_t pad;
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// TC image of the centres (runs after pack_centers_kernel on the same stream)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) pack_tc_kernel(int k, int d, unsigned char *__restrict__ pack) {
    const KmPackLayout L = km_pack_layout(k, d);
    const float *c32 = reinterpret_cast<const float *>(pack + L.off_c32);
    const float *cn32 = reinterpret_cast<const float *>(pack + L.off_cn32);
    const int j = blockIdx.x;
    const int tile = j / KM_TC_BN, jj = j % KM_TC_BN;
    const int kpad = km_tc_kpad(d);
    unsigned char *img = pack + L.off_tc + (size_t)tile * km_tc_tile_bytes(d);
    float *hi = reinterpret_cast<float *>(img);
    float *lo = reinterpret_cast<float *>(img + km_tc_hi_bytes(d));
    for (int f = threadIdx.x; f < kpad; f += blockDim.x) {
        float v = 0.f;
        if (j < k && f < d) v = -2.f * c32[(size_t)j * d + f];
        const float h = tf32_rna(v);
        const float l = tf32_rna(v - h);
        const size_t o = ((size_t)(f >> 2) * KM_TC_BN + jj) * 4 + (f & 3);
        hi[o] = h;
        lo[o] = l;
    }
    if (threadIdx.x < 8) {
        // augmented K-step: chunk kpad/4 = [cn_h, cn_m, cn_l, 0], chunk kpad/4+1 = 0
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
        }
        const int f = kpad + threadIdx.x;
        hi[((size_t)(f >> 2) * KM_TC_BN + jj) * 4 + (f & 3)] = v;
    }
}

int km_launch_pack_tc(int k, int d, void *pack, cudaStream_t st) {
    if (d > KM_TC_MAX_D) return 0;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    pack_tc_kernel<<<ntc * KM_TC_BN, 64, 0, st>>>(k, d, static_cast<unsigned char *>(pack));
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(TC_THREADS, 1)
assign_tc_f32_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx,
                     const unsigned char *__restrict__ tc_img, const float *__restrict__ c32,
                     const unsigned int *__restrict__ cmax_bits, int k,
                     int32_t *__restrict__ labels, float *__restrict__ mind2,
                     int32_t *__restrict__ flag_rows, int32_t *__restrict__ flag_count) {
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;                 // 16-byte K chunks of the real features
    const uint32_t xhi_bytes = (uint32_t)(nch + 2) * 128 * 16;  // per row tile
    const uint32_t xlo_bytes = (uint32_t)nch * 128 * 16;
    const uint32_t bhi_bytes = (uint32_t)km_tc_hi_bytes(d);
    const uint32_t blo_bytes = (uint32_t)km_tc_lo_bytes(d);
    const uint32_t stage_bytes = bhi_bytes + blo_bytes;
    unsigned char *xhi = tc_smem;                          // [2][nch+2][128][16B]
    unsigned char *xlo = xhi + 2 * xhi_bytes;              // [2][nch][128][16B]
    unsigned char *bst = xlo + 2 * xlo_bytes;              // [TC_BSTAGES][stage_bytes]
    TcBars *bars = reinterpret_cast<TcBars *>(bst + TC_BSTAGES * stage_bytes);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntc = (k + KM_TC_BN - 1) / KM_TC_BN;
    const int64_t nsteps = (n + TC_ROWS - 1) / TC_ROWS;

    if (tid == 0) {
        for (int i = 0; i < TC_BSTAGES; ++i) { km_mbar_init(&bars->b_full[i], 1); km_mbar_init(&bars->b_empty[i], 1); }
        for (int i = 0; i < TC_ASTAGES; ++i) { km_mbar_init(&bars->a_full[i], 1); km_mbar_init(&bars->a_empty[i], 128); }
        km_mbar_init(&bars->x_ready, 128);
        km_mbar_init(&bars->x_free, 1);
        km_fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         km_smem_u32(&bars->tmem_base)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // constant augmented chunks of Xhi: chunk nch = [1,1,1,0], chunk nch+1 = 0 (both row tiles)
    for (int e = tid; e < 2 * 2 * 128; e += TC_THREADS) {
        const int rt = e / 256, c = (e / 128) & 1, r = e & 127;
        float4 v = c == 0 ? make_float4(1.f, 1.f, 1.f, 0.f) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(xhi + (size_t)rt * xhi_bytes + ((size_t)(nch + c) * 128 + r) * 16) = v;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x) {
                for (int t = 0; t < ntc; ++t, ++it) {
                    const int s = it % TC_BSTAGES;
                    const uint32_t ph = (it / TC_BSTAGES) & 1;
                    mbar_wait_bounded(&bars->b_empty[s], ph ^ 1);
                    km_mbar_expect_tx(&bars->b_full[s], stage_bytes);
                    km_bulk_g2s(bst + (size_t)s * stage_bytes, tc_img + (size_t)t * stage_bytes, stage_bytes,
                                &bars->b_full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            // instruction descriptor: D=F32, A=B=TF32, both K-major, N=64, M=128
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KM_TC_BN >> 3) << 17) |
                                   ((uint32_t)(128 >> 4) << 24);
            const uint32_t xhi_a = km_smem_u32(xhi), xlo_a = km_smem_u32(xlo), bst_a = km_smem_u32(bst);
            uint32_t it = 0, xphase = 0;
            for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x, xphase ^= 1) {
                mbar_wait_bounded(&bars->x_ready, xphase);
                tc_fence_after();
                for (int t = 0; t < ntc; ++t, ++it) {
                    const int s = it % TC_BSTAGES;
                    const uint32_t bph = (it / TC_BSTAGES) & 1;
                    const int a = it % TC_ASTAGES;
                    const uint32_t aph = (it / TC_ASTAGES) & 1;
                    mbar_wait_bounded(&bars->a_empty[a], aph ^ 1);
                    mbar_wait_bounded(&bars->b_full[s], bph);
                    tc_fence_after();
                    const uint32_t bhi_a = bst_a + s * stage_bytes, blo_a = bhi_a + bhi_bytes;
#pragma unroll 1
                    for (int rt = 0; rt < 2; ++rt) {
                        const uint32_t dcol = tmem_base + (uint32_t)(a * 128 + rt * 64);
                        const uint32_t ah = xhi_a + rt * xhi_bytes, al = xlo_a + rt * xlo_bytes;
                        // augmented step first (overwrites the accumulator): + ||c||^2
                        mma_tf32(dcol, make_smem_desc(ah + (uint32_t)nch * 2048u, 2048u, 128u),
                                 make_smem_desc(bhi_a + (uint32_t)nch * 1024u, 1024u, 128u), idesc, 0u);
                        for (int ks = 0; ks < nch / 2; ++ks) {
                            const uint64_t da_h = make_smem_desc(ah + (uint32_t)ks * 4096u, 2048u, 128u);
                            const uint64_t da_l = make_smem_desc(al + (uint32_t)ks * 4096u, 2048u, 128u);
                            const uint64_t db_h = make_smem_desc(bhi_a + (uint32_t)ks * 2048u, 1024u, 128u);
                            const uint64_t db_l = make_smem_desc(blo_a + (uint32_t)ks * 2048u, 1024u, 128u);
                            mma_tf32(dcol, da_l, db_h, idesc, 1u);
                            mma_tf32(dcol, da_h, db_l, idesc, 1u);
                            mma_tf32(dcol, da_h, db_h, idesc, 1u);
                        }
                    }
                    mma_commit(&bars->b_empty[s]);  // smem stage reusable once these MMAs retire
                    mma_commit(&bars->a_full[a]);   // accumulator stage ready for the epilogue
                }
                mma_commit(&bars->x_free);  // X tile no longer read by the tensor core
            }
        }
    } else {
        // ================= X staging + epilogue (warps 2..5, 128 threads) =================
        const int q4 = warp & 3;                 // TMEM lane quarter this warp may access
        const int rl = q4 * 32 + lane;           // row within a row tile == TMEM lane (also staged by this thread)
        const float cmax = __uint_as_float(*cmax_bits);
        const float bound_scale = 2.0f * (float)(d + 8) * 1.1920929e-07f;
        uint32_t it = 0, fphase = 0;
        bool first = true;
        for (int64_t step = blockIdx.x; step < nsteps; step += gridDim.x) {
            const int64_t row0 = step * TC_ROWS;
            if (!first) {
                mbar_wait_bounded(&bars->x_free, fphase);
                fphase ^= 1;
            }
            first = false;
            // ---- stage both row tiles: this thread owns rows rl and 128+rl of this step ---------------
            float xn[2];
#pragma unroll
            for (int rt = 0; rt < 2; ++rt) {
                const int64_t row = row0 + rt * 128 + rl;
                const bool live = row < n;
                const float *xr = X + (live ? row : 0) * ldx;
                unsigned char *ph = xhi + (size_t)rt * xhi_bytes + (size_t)rl * 16;
                unsigned char *pl = xlo + (size_t)rt * xlo_bytes + (size_t)rl * 16;
                float s = 0.f;
                for (int c = 0; c < nch; ++c) {
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (live) {
                        if (VEC) {
                            v = __ldg(reinterpret_cast<const float4 *>(xr) + c);
                        } else {
                            const int f = c * 4;
                            v.x = f + 0 < d ? __ldg(xr + f + 0) : 0.f;
                            v.y = f + 1 < d ? __ldg(xr + f + 1) : 0.f;
                            v.z = f + 2 < d ? __ldg(xr + f + 2) : 0.f;
                            v.w = f + 3 < d ? __ldg(xr + f + 3) : 0.f;
                        }
                    }
                    s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                    float4 h, l;
                    h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
                    l.x = tf32_rna(v.x - h.x); l.y = tf32_rna(v.y - h.y);
                    l.z = tf32_rna(v.z - h.z); l.w = tf32_rna(v.w - h.w);
                    *reinterpret_cast<float4 *>(ph + (size_t)c * 2048) = h;
                    *reinterpret_cast<float4 *>(pl + (size_t)c * 2048) = l;
                }
                xn[rt] = s;
            }
            km_fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
            mbar_arrive(&bars->x_ready);

            // ---- epilogue over the column tiles ----------------------------------------------------
            float best[2] = {INFINITY, INFINITY}, second[2] = {INFINITY, INFINITY};
            int bidx[2] = {0, 0};
            for (int t = 0; t < ntc; ++t, ++it) {
                const int a = it % TC_ASTAGES;
                const uint32_t aph = (it / TC_ASTAGES) & 1;
                mbar_wait_bounded(&bars->a_full[a], aph);
                tc_fence_after();
                const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(a * 128);
#pragma unroll
                for (int rt = 0; rt < 2; ++rt) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t r[32];
                        tmem_ld32(tbase + (uint32_t)(rt * 64 + h * 32), r);
                        tmem_ld_wait();
                        const int colbase = t * KM_TC_BN + h * 32;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            track(__uint_as_float(r[j]), colbase + j, best[rt], second[rt], bidx[rt]);
                    }
                }
                tc_fence_before();
                mbar_arrive(&bars->a_empty[a]);
            }

            // ---- winner distance (direct form, fp32), outputs, near-tie flags ---------------------
#pragma unroll
            for (int rt = 0; rt < 2; ++rt) {
                const int64_t row = row0 + rt * 128 + rl;
                if (row >= n) continue;
                int bi = bidx[rt];
                if (bi >= k) bi = k - 1;
                const unsigned char *ph = xhi + (size_t)rt * xhi_bytes + (size_t)rl * 16;
                const unsigned char *pl = xlo + (size_t)rt * xlo_bytes + (size_t)rl * 16;
                const float *crow = c32 + (size_t)bi * d;
                float s = 0.f;
                for (int c = 0; c < nch; ++c) {
                    const float4 h = *reinterpret_cast<const float4 *>(ph + (size_t)c * 2048);
                    const float4 l = *reinterpret_cast<const float4 *>(pl + (size_t)c * 2048);
                    const float xv[4] = {h.x + l.x, h.y + l.y, h.z + l.z, h.w + l.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int f = c * 4 + i;
                        if (f < d) {
                            const float df = xv[i] - __ldg(crow + f);
                            s = fmaf(df, df, s);
                        }
                    }
                }
                labels[row] = bi;
                if (mind2) mind2[row] = s;
                if (flag_rows) {
                    const float thr = bound_scale * (xn[rt] + cmax);
                    if (second[rt] - best[rt] <= thr) {
                        const int pos = atomicAdd(flag_count, 1);
                        flag_rows[pos] = (int32_t)row;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// host-side launcher, called from b200km_assign_f32
int km_launch_assign_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, int32_t *labels,
                        float *mind2, int32_t *flag_rows, int32_t *flag_count, cudaStream_t st) {
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;
    const size_t smem = 2 * (size_t)(nch + 2) * 128 * 16 + 2 * (size_t)nch * 128 * 16 +
                        TC_BSTAGES * km_tc_tile_bytes(d) + sizeof(TcBars) + 64;
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(assign_tc_f32_kernel<true>, 227 * 1024, attr_v)) return rc;
    if (int rc = km_set_smem_once(assign_tc_f32_kernel<false>, 227 * 1024, attr_s)) return rc;
    int64_t steps = (n + TC_ROWS - 1) / TC_ROWS;
    const int grid = (int)(steps < KM_NUM_SMS ? steps : KM_NUM_SMS);
    const unsigned char *img = pk + L.off_tc;
    const float *c32 = reinterpret_cast<const float *>(pk + L.off_c32);
    const unsigned int *cmax = reinterpret_cast<const unsigned int *>(pk + L.off_cmax);
    if (vec)
        assign_tc_f32_kernel<true><<<grid, TC_THREADS, smem, st>>>(X, n, d, ldx, img, c32, cmax, k, labels, mind2,
                                                                    flag_rows, flag_count);
    else
        assign_tc_f32_kernel<false><<<grid, TC_THREADS, smem, st>>>(X, n, d, ldx, img, c32, cmax, k, labels, mind2,
                                                                     flag_rows, flag_count);
    return (int)cudaGetLastError();
}

// cdist_tc.cu -- K7 on the tensor cores: the (n x k) matrix of distances between rows and centres, streamed out.
//
// Replaces dask_ml.metrics.euclidean_distances (reference dask_ml/metrics/pairwise.py:70-105, the body of
// KMeans.transform, dask_ml/cluster/k_means.py:225-228), pairwise_distances (:50-67) and rbf_kernel (:141-151) for
// float32 rows of up to 128 features; everything else stays on the FFMA / DFMA tile kernel in misc.cu.
//
// The op is bound by its OUTPUT (k floats written per row against 2*k*d flops), so the kernel is built around the
// store: the roles of the operands are SWAPPED with respect to K1 -- the centres are the M side of the MMA (128 per
// accumulator, TMEM lane = centre) and 256 rows of X the N side (TMEM column = row).  A thread of the epilogue then
// holds, register by register, one row after the other, and the 32 lanes of a warp 32 neighbouring centres: every
// store instruction writes 128 contiguous bytes of one output row.  (With K1's orientation a lane owns a row and a
// warp store scatters 16-byte pieces over 32 output rows.)
//
//   D = [(-2c)_hi | (-2c)_lo | cn_h cn_m cn_l 1 1 1] . [x_hi | x_lo | 1 1 1 xn_h xn_m xn_l]^T   (3xTF32, fp32 in TMEM)
//     = ||c||^2 + ||x||^2 - 2 x.c     -- both norms ride in the augmented K-step, the epilogue adds nothing.
// Rows and centres are centred on the pack's mu (distances are translation invariant; the norms that the
// cancellation error scales with become those of the centred data).
//
// CTA = 320 threads, persistent, one per SM, 256 rows per block (128 for 64 < d <= 128):
//   warp 0     : producer -- the 256-centre tiles of the pack's tcgen05 image, K-step block by K-step block (1-D bulk
//                async copies through a 5-stage ring), once per 128-centre half
//   warp 1     : tcgen05.mma M = 128 (centres), N = 256 (rows), K = 8, kind::tf32; two accumulator stages of 256 columns
//   warps 2..9 : stage the block's rows (TF32 hi / exact lo parts + the augmented chunks) and run the epilogue:
//                tcgen05.ld 32x32b.x32 -> max(0) -> sqrt / exp -> coalesced stores
#include "tc_common.cuh"

namespace {

// rows of X per block (N of the MMA): 256 while a 256-row block (hi + lo parts) fits next to the ring (d <= 64),
// 128 beyond that (d <= 128; two staging threads per row then)
constexpr int CD_NST = 5;         // ring stages of 16 KB
constexpr int CD_MAX_D = 128;

__device__ __forceinline__ float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

struct CdBars {
    uint64_t b_full[CD_NST];
    uint64_t b_empty[CD_NST];
    uint64_t a_full[TC_ASTAGES];
    uint64_t a_empty[TC_ASTAGES];
    uint64_t x_ready;
    uint64_t x_free;
    uint32_t tmem_base;
    uint32_t pad;
    alignas(16) float mu[CD_MAX_D];
    float xn_part[128];            // CD_ROWS = 128: the second staging thread's half of ||x||^2
    float mins[2][4][4][32];       // MODE 3: [row half][32-row chunk][lane quarter][row]: the quarters' minima
};

// MODE 0: squared distance, 1: distance, 2: exp(-gamma * squared distance)
// MODE 3 (K4 against many new candidates, b200km_mindist_tc_f32): nothing of the matrix is stored -- `out` is the
//   running per-row minimum: out[i] = min(out[i], min_j ||x_i - c_j||^2) (init != 0: the old value is ignored) and
//   *cost += sum_i out[i].  The k <= 128 candidates are the 128 TMEM lanes of ONE accumulator stage; the warps that
//   own lanes with real candidates read their 32 x 32 pieces and a transposing min-reduction over the 32 lanes leaves
//   row i's minimum in lane i -- coalesced access to the minima, one exchange per value instead of five.
template <bool VEC, int MODE, int CD_ROWS>
__global__ void __launch_bounds__(TC_THREADS, 1)
cdist_tc_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx, const unsigned char *__restrict__ tc_img,
                const float *__restrict__ mu, int k, float *__restrict__ out, int64_t ldo, float gamma,
                double *__restrict__ cost, int init) {
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    const int kpad = km_tc_kpad(d);
    const int nch = kpad / 4;                 // 16-byte K chunks of the real features
    const int nks = kpad / 8;                 // K-steps
    const uint32_t xhi_bytes = (uint32_t)(nch + 2) * CD_ROWS * 16;
    const uint32_t xlo_bytes = (uint32_t)nch * CD_ROWS * 16;
    const uint32_t tile_bytes = (uint32_t)km_tc_tile_bytes(d);
    unsigned char *xb = tc_smem;                                   // [hi: nch + 2 chunks | lo: nch chunks] x 256 rows x 16 B
    unsigned char *bst = xb + xhi_bytes + xlo_bytes;               // [CD_NST][16 KB]
    CdBars *bars = reinterpret_cast<CdBars *>(bst + CD_NST * KM_TC_KS_BYTES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nst = (k + 127) / 128;                               // accumulator stages per block: 128 centres each
    const int64_t nblocks = (n + CD_ROWS - 1) / CD_ROWS;

    if (tid == 0) {
        for (int i = 0; i < CD_NST; ++i) { km_mbar_init(&bars->b_full[i], 1); km_mbar_init(&bars->b_empty[i], 1); }
        for (int i = 0; i < TC_ASTAGES; ++i) { km_mbar_init(&bars->a_full[i], 1); km_mbar_init(&bars->a_empty[i], TC_EPI); }
        km_mbar_init(&bars->x_ready, TC_EPI);
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
    for (int f = tid; f < CD_MAX_D; f += TC_THREADS) bars->mu[f] = f < kpad ? __ldg(mu + f) : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
                for (int s = 0; s < nst; ++s) {
                    const unsigned char *src = tc_img + (size_t)(s >> 1) * tile_bytes;   // both halves read the same tile
                    for (int b = 0; b <= nks; ++b, ++it) {
                        const int st = it % CD_NST;
                        const uint32_t ph = (it / CD_NST) & 1;
                        const uint32_t bytes = b == 0 ? KM_TC_AUG_BYTES : KM_TC_KS_BYTES;
                        mbar_wait_bounded(&bars->b_empty[st], ph ^ 1);
                        km_mbar_expect_tx(&bars->b_full[st], bytes);
                        km_bulk_g2s(bst + (size_t)st * KM_TC_KS_BYTES, src, bytes, &bars->b_full[st]);
                        src += bytes;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // D = F32, A = B = TF32, both K-major, M = 128 (centres), N = 256 (rows)
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(CD_ROWS >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);           // SBO = 128 B, version = 1
        const uint32_t lbo = ((256u * 16u) >> 4) << 16;              // centres: the image's 256 rows x 16 B per K chunk
        const uint32_t xlbo = (((uint32_t)CD_ROWS * 16u) >> 4) << 16;   // rows: CD_ROWS x 16 B per K chunk
        const uint32_t xb_a = km_smem_u32(xb), bst_a = km_smem_u32(bst);
        const bool leader = elect_one();
        uint32_t it = 0, ita = 0, li = 0;
        const uint32_t xh = ((xb_a & 0x3FFFFu) >> 4) | xlbo;         // X (the B operand), hi parts
        const uint32_t xl = xh + (xhi_bytes >> 4);
        for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x, ++li) {
            mbar_wait_bounded(&bars->x_ready, li & 1);
            tc_fence_after();
            for (int s = 0; s < nst; ++s, ++ita) {
                const int a = ita % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_empty[a], ((ita / TC_ASTAGES) & 1) ^ 1);
                const uint32_t dcol = tmem_base + (uint32_t)(a * CD_ROWS);
                const uint32_t half = (uint32_t)(s & 1) * (128u * 16u >> 4);   // rows 128.. of every chunk of the tile
                for (int b = 0; b <= nks; ++b, ++it) {
                    const int st = it % CD_NST;
                    mbar_wait_bounded(&bars->b_full[st], (it / CD_NST) & 1);
                    tc_fence_after();
                    const uint32_t ch = ((((bst_a + st * KM_TC_KS_BYTES) & 0x3FFFFu) >> 4) | lbo) + half;   // centres, hi
                    if (leader) {
                        if (b == 0) {
                            // augmented step first (overwrites the accumulator): ||c||^2 + ||x||^2
                            mma_tf32(dcol, ch, xh + (uint32_t)nch * (CD_ROWS * 16u >> 4), desc_hi, idesc, 0u);
                        } else {
                            const uint32_t cl = ch + ((KM_TC_KS_BYTES / 2) >> 4);
                            const uint32_t xo = (uint32_t)(b - 1) * (2u * CD_ROWS * 16u >> 4);
                            mma_tf32(dcol, cl, xh + xo, desc_hi, idesc, 1u);
                            mma_tf32(dcol, ch, xl + xo, desc_hi, idesc, 1u);
                            mma_tf32(dcol, ch, xh + xo, desc_hi, idesc, 1u);
                        }
                        mma_commit(&bars->b_empty[st]);
                        if (b == nks) mma_commit(&bars->a_full[a]);
                    }
                    __syncwarp();
                }
            }
            if (leader) mma_commit(&bars->x_free);       // the row block is no longer read by the tensor core
            __syncwarp();
        }
    } else {
        // ================= row staging + epilogue (warps 2..9) =================
        // staging: CD_ROWS = 256 -> thread et stages row et; CD_ROWS = 128 -> two threads per row, alternate chunks
        const int et = tid - 64;                 // 0..255
        constexpr int TPR = 256 / CD_ROWS;       // staging threads per row
        const int srow = et % CD_ROWS, shalf = et / CD_ROWS;
        const float ngamma = -gamma;
        const int g = (warp - 2) >> 2;           // epilogue: column half of the accumulator (rows of the block)
        const int q4 = warp & 3;                 // epilogue: TMEM lane quarter (32 centres)
        constexpr int NCHM = 16;                 // chunks per staging thread: 64 features (x 2 threads at CD_ROWS = 128)
        constexpr int EC = CD_ROWS / 64;         // 32-column chunks per epilogue warp
        float4 ld[NCHM];
        double cost_acc = 0.0;                   // MODE 3
        auto issue_loads = [&](int64_t blk) {
            const int64_t row = blk * CD_ROWS + srow;
            const float *src = X + row * ldx;
#pragma unroll
            for (int i = 0; i < NCHM; ++i) {
                const int c = i * TPR + shalf;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c < nch && row < n) {
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
        if ((int64_t)blockIdx.x < nblocks) issue_loads(blockIdx.x);
        uint32_t ita = 0, li = 0;
        for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x, ++li) {
            // ---- stage this block's rows: x - mu split into TF32 hi / exact lo parts, norms into the augmented chunks
            if (li > 0) mbar_wait_bounded(&bars->x_free, (li - 1) & 1);
            {
                const bool live = blk * CD_ROWS + srow < n;
                unsigned char *ph = xb + (size_t)srow * 16;
                unsigned char *pl = ph + xhi_bytes;
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < NCHM; ++i) {
                    const int c = i * TPR + shalf;
                    if (c < nch) {
                        float4 v = ld[i];
                        const float4 mq = *reinterpret_cast<const float4 *>(&bars->mu[c * 4]);
                        if (live) { v.x -= mq.x; v.y -= mq.y; v.z -= mq.z; v.w -= mq.w; }
                        // (features beyond d: the loads gave 0 and mu is zero padded, so they stay 0)
                        s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                        float4 h, l;
                        h.x = tf32_rna(v.x); h.y = tf32_rna(v.y); h.z = tf32_rna(v.z); h.w = tf32_rna(v.w);
                        l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;
                        *reinterpret_cast<float4 *>(ph + (size_t)c * (CD_ROWS * 16)) = h;
                        *reinterpret_cast<float4 *>(pl + (size_t)c * (CD_ROWS * 16)) = l;
                    }
                }
                if (TPR == 2) {      // the row's two halves meet: the second thread publishes, the first adds
                    if (shalf == 1) bars->xn_part[srow] = s;
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (shalf == 0) s += bars->xn_part[srow];
                }
                if (shalf == 0) {
                    // augmented chunks: [1, 1, 1, xn_h] [xn_m, xn_l, 0, 0] against the image's [cn_h, cn_m, cn_l, 1] [1, 1, 0, 0]
                    const float sh = tf32_rna(s);
                    const float sm = tf32_rna(s - sh);
                    const float sl = (s - sh) - sm;
                    *reinterpret_cast<float4 *>(ph + (size_t)nch * (CD_ROWS * 16)) = make_float4(1.f, 1.f, 1.f, sh);
                    *reinterpret_cast<float4 *>(ph + (size_t)(nch + 1) * (CD_ROWS * 16)) = make_float4(sm, sl, 0.f, 0.f);
                }
                km_fence_proxy_async();
                mbar_arrive(&bars->x_ready);
            }
            const int64_t nblk = blk + gridDim.x;
            if (nblk < nblocks) issue_loads(nblk);       // lands while this block's stages are computed and stored
            // ---- epilogue: one accumulator stage = 128 centres x CD_ROWS rows
            const int64_t row0 = blk * CD_ROWS + g * (CD_ROWS / 2);
            for (int s = 0; s < nst; ++s, ++ita) {
                const int a = ita % TC_ASTAGES;
                mbar_wait_bounded(&bars->a_full[a], (ita / TC_ASTAGES) & 1);
                tc_fence_after();
                const int j = s * 128 + q4 * 32 + lane;                 // this lane's centre
                const bool jok = j < k;
                const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(a * CD_ROWS + g * (CD_ROWS / 2));
                if (MODE == 3) {
                    // lane quarters 1..3 have work only when k > 32 * q4; padded candidates carry ||c||^2 = 1e30 and
                    // never win.  mmin[c]: after the reduction, lane i holds the minimum of row r0 + i over this
                    // warp's 32 candidates; the quarters are merged through shared memory (mins[g][c][q4][lane]).
                    const bool busy = 32 * q4 < k;
#pragma unroll
                    for (int c = 0; c < EC; ++c) {
                        float v[32];
                        if (busy) {
                            uint32_t r[32];
                            tmem_ld32(tbase + (uint32_t)(c * 32), r);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                            // transposing min-reduction: at every step a lane hands on half of what it holds
#pragma unroll
                            for (int O = 16, H = 16; O >= 1; O >>= 1, H >>= 1) {
                                const bool up = (lane & O) != 0;
#pragma unroll
                                for (int i = 0; i < 16; ++i) {
                                    if (i < H) {
                                        const float send = up ? v[i] : v[i + H];
                                        const float keep = up ? v[i + H] : v[i];
                                        v[i] = fminf(keep, __shfl_xor_sync(0xffffffffu, send, O));
                                    }
                                }
                            }
                        } else {
                            v[0] = INFINITY;
                        }
                        bars->mins[g][c][q4][lane] = v[0];
                    }
                    tc_fence_before();
                    mbar_arrive(&bars->a_empty[a]);
                    // the four quarters of this row half meet (named barrier per half: 128 threads), quarter 0 finishes
                    asm volatile("bar.sync %0, 128;" ::"r"(2 + g) : "memory");
                    if (q4 == 0) {
#pragma unroll
                        for (int c = 0; c < EC; ++c) {
                            const int64_t r = row0 + c * 32 + lane;
                            if (r < n) {
                                float mn = fminf(fminf(bars->mins[g][c][0][lane], bars->mins[g][c][1][lane]),
                                                 fminf(bars->mins[g][c][2][lane], bars->mins[g][c][3][lane]));
                                mn = fmaxf(mn, 0.f);
                                if (!init) mn = fminf(mn, out[r]);
                                out[r] = mn;
                                cost_acc += (double)mn;
                            }
                        }
                    }
                    asm volatile("bar.sync %0, 128;" ::"r"(2 + g) : "memory");   // mins is rewritten by the next block
                    continue;
                }
                uint32_t ra[32], rb[32];
                tmem_ld32(tbase, ra);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < EC; ++c) {
                    uint32_t (&cur)[32] = (c & 1) ? rb : ra;
                    uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
                    if (c + 1 < EC) tmem_ld32(tbase + (uint32_t)((c + 1) * 32), nxt);   // flies under this chunk's stores
                    const int64_t r0 = row0 + c * 32;
                    if (r0 < n) {                                        // warp-uniform
                        float *po = out + r0 * ldo + j;
                        const int lim = (int)((n - r0) < 32 ? (n - r0) : 32);
                        // one MUFU per element (sqrt.approx / ex2.approx: ~1 ulp, far inside the float32 expanded
                        // form's own error); the IEEE sqrtf's fix-up path made the epilogue the bound
                        float v[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            v[i] = fmaxf(__uint_as_float(cur[i]), 0.f);
                            if (MODE == 1) v[i] = sqrt_approx(v[i]);
                            if (MODE == 2) v[i] = __expf(ngamma * v[i]);
                        }
                        if (lim == 32) {                                 // warp-uniform: no per-element predicate
                            if (jok) {
#pragma unroll
                                for (int i = 0; i < 32; ++i) po[(int64_t)i * ldo] = v[i];
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (jok && i < lim) po[(int64_t)i * ldo] = v[i];
                        }
                    }
                    if (c + 1 < EC) tmem_ld_wait();
                }
                tc_fence_before();
                mbar_arrive(&bars->a_empty[a]);
            }
        }
        if (MODE == 3) {
            cost_acc = km_warp_sum(cost_acc);
            if (lane == 0 && cost_acc != 0.0) atomicAdd(cost, cost_acc);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace

bool km_tc_pair_mode();   // assign.cu

// host-side launcher, called from b200km_cdist_pack_f32 (misc.cu)
int km_launch_cdist_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, float *out, int64_t ldo,
                       int mode, float gamma, cudaStream_t st, double *cost, int init) {
    if (d > CD_MAX_D || km_tc_pair_mode()) return B200KM_ETOOBIG;
    if (mode == 3 && k > 128) return B200KM_ETOOBIG;
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    const int nch = km_tc_kpad(d) / 4;
    const int rows = d <= 64 ? 256 : 128;
    const size_t smem = (size_t)(nch + 2) * rows * 16 + (size_t)nch * rows * 16 + (size_t)CD_NST * KM_TC_KS_BYTES +
                        sizeof(CdBars) + 64;
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    const unsigned char *img = pk + L.off_tc;
    const float *mu = reinterpret_cast<const float *>(pk + L.off_mu);
    const int64_t nblocks = (n + rows - 1) / rows;
    const unsigned grid = (unsigned)(nblocks < km_num_sms() ? nblocks : km_num_sms());
#define CD_LAUNCH(VEC_, MODE_, ROWS_)                                                                              \
    {                                                                                                              \
        static bool attr[KM_MAX_DEVICES] = {};                                                                     \
        if (int rc = km_set_smem_once(cdist_tc_kernel<VEC_, MODE_, ROWS_>, 227 * 1024, attr)) return rc;           \
        cdist_tc_kernel<VEC_, MODE_, ROWS_><<<grid, TC_THREADS, smem, st>>>(X, n, d, ldx, img, mu, k, out, ldo, gamma,  \
                                                                      cost, init);                             \
    }
#define CD_LAUNCH_M(VEC_, ROWS_)                                                                                   \
    if (mode == 0) CD_LAUNCH(VEC_, 0, ROWS_) else if (mode == 1) CD_LAUNCH(VEC_, 1, ROWS_)                          \
    else if (mode == 2) CD_LAUNCH(VEC_, 2, ROWS_) else CD_LAUNCH(VEC_, 3, ROWS_)
    if (rows == 256) {
        if (vec) { CD_LAUNCH_M(true, 256) } else { CD_LAUNCH_M(false, 256) }
    } else {
        if (vec) { CD_LAUNCH_M(true, 128) } else { CD_LAUNCH_M(false, 128) }
    }
#undef CD_LAUNCH_M
#undef CD_LAUNCH
    return (int)cudaGetLastError();
}

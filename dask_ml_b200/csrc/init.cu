// init.cu -- k-means|| kernels: K4 running min-distance / cost, K5 Bernoulli oversampling,
// K6 row gather.
//
// Replaces evaluate_cost (reference dask_ml/cluster/k_means.py:516-519), _sample_points
// (:522-541) and the fancy-index gathers X[sorted(c_idx)] (:479), X[locs] (:500), X[idx]
// (:431, :452).  The reference recomputes the distance to EVERY candidate each round and
// materialises the n x |C| block matrices; here a running minimum per row is kept, so a round
// costs one read of X plus 2 words per row.
#include "common.cuh"
#include <type_traits>

int km_k4_variant();  // assign.cu (b200km_set_option)

constexpr int MD_THREADS = 256;
constexpr int MD_MAXM = B200KM_MINDIST_MAX_M;

// LPR lanes cooperate on one row; lane l owns features l, l+LPR, ...  (coalesced, no alignment
// requirement).  32/LPR rows per warp step.
template <typename T, int LPR>
__global__ void __launch_bounds__(MD_THREADS)
mindist_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const T *__restrict__ cand, int m,
               T *__restrict__ mind2, int init, double *__restrict__ cost_out) {
    extern __shared__ __align__(16) unsigned char md_smem[];
    T *cs = reinterpret_cast<T *>(md_smem);  // [m][d]
    __shared__ double red[32];
    for (int e = threadIdx.x; e < m * d; e += MD_THREADS) cs[e] = cand[e];
    __syncthreads();

    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int sub = lane / LPR, l = lane % LPR;
    const int64_t warp_global = ((int64_t)blockIdx.x * MD_THREADS + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * MD_THREADS) >> 5;
    double cost = 0.0;
    for (int64_t r0 = warp_global * RPW; r0 < n; r0 += nwarps * RPW) {
        const int64_t row = r0 + sub;
        const bool live = row < n;
        T acc[MD_MAXM];
#pragma unroll
        for (int j = 0; j < MD_MAXM; ++j) acc[j] = (T)0;
        const T *xr = X + (live ? row : 0) * ldx;
#pragma unroll 4
        for (int f = l; f < d; f += LPR) {
            const T xv = live ? __ldg(xr + f) : (T)0;
#pragma unroll
            for (int j = 0; j < MD_MAXM; ++j) {
                if (j < m) {
                    const T df = xv - cs[j * d + f];
                    acc[j] = fma(df, df, acc[j]);
                }
            }
        }
        T best = (T)INFINITY;
#pragma unroll
        for (int j = 0; j < MD_MAXM; ++j) {
            if (j < m) {
                T v = acc[j];
#pragma unroll
                for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                best = v < best ? v : best;
            }
        }
        if (live && l == 0) {
            if (!init) {
                const T old = mind2[row];
                best = old < best ? old : best;
            }
            mind2[row] = best;
            cost += (double)best;
        }
    }
    km_block_atomic_add(cost, cost_out, red);
}

// Sum NV per-lane values over the O*2 lanes of a group with a TRANSPOSING reduction: at every step a lane passes on
// half of what it holds and adds what it receives to the other half, so the group needs NV - NV/LPR shuffles in
// all instead of NV * log2(LPR) (32 values over 16 lanes: 30 instead of 128 -- the butterfly made K4 shuffle-bound
// once a round brought more than two new candidates).  Afterwards lane l of the group holds, in v[0 .. NV/LPR),
// the full sums of the original entries l * NV/LPR ... .
template <typename T, int NV, int CNT, int O>
struct KmTrReduce {
    static __device__ __forceinline__ void run(T (&v)[NV], int l) {
        constexpr int H = CNT / 2;
        const bool up = (l & O) != 0;
#pragma unroll
        for (int i = 0; i < H; ++i) {
            const T send = up ? v[i] : v[i + H];
            const T keep = up ? v[i + H] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, O);
        }
        if constexpr (O > 1) KmTrReduce<T, NV, H, O / 2>::run(v, l);
    }
};

// Vector path (rows 16-byte aligned, d % VEC == 0): LPR lanes cooperate on one row, each lane loads whole
// 16-byte vectors, and U row groups are in flight per warp so enough bytes are outstanding to cover the
// HBM latency (the scalar kernel above kept ~32 KB per SM in flight and reached 25 % of the HBM peak).
template <typename T, int LPR, int U, int MM>
__global__ void __launch_bounds__(MD_THREADS)
mindist_vec_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const T *__restrict__ cand, int m,
                   T *__restrict__ mind2, int init, double *__restrict__ cost_out) {
    constexpr int VEC = 16 / sizeof(T);
    using V = typename std::conditional<sizeof(T) == 4, float4, double2>::type;
    extern __shared__ __align__(16) unsigned char md_smem[];
    T *cs = reinterpret_cast<T *>(md_smem);  // [m][d]
    __shared__ double red[32];
    for (int e = threadIdx.x; e < m * d; e += MD_THREADS) cs[e] = cand[e];
    __syncthreads();

    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int sub = lane / LPR, l = lane % LPR;
    const int dv = d / VEC;
    const int64_t warp_global = ((int64_t)blockIdx.x * MD_THREADS + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * MD_THREADS) >> 5;
    double cost = 0.0;
    for (int64_t r0 = warp_global * (RPW * U); r0 < n; r0 += nwarps * (RPW * U)) {
        // One lane per row owns its running minimum: loaded up front (not on the critical path after the
        // reduction) and written back coalesced.  TR: the transposing reduction leaves row u's sums with the G lanes
        // u*G .. u*G+G-1 and lane u*G owns the row; otherwise (more lanes than values) lane u does.
        constexpr bool TR = (U * MM >= LPR) && (LPR > 1);
        constexpr int G = TR ? LPR / U : 1;
        const int own_u = TR ? l / G : l;
        const int64_t myrow = r0 + (int64_t)own_u * RPW + sub;
        const bool mine = (TR ? (l % G) == 0 : l < U) && myrow < n;
        T oldv = (T)INFINITY;
        if (mine && !init) oldv = mind2[myrow];
        T mybest = (T)INFINITY;
        T acc[U][MM];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < MM; ++j) acc[u][j] = (T)0;
        for (int q0 = 0; q0 < dv; q0 += LPR) {
            const int q = q0 + l;
            V xv[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t row = r0 + u * RPW + sub;
                V v;
                if (sizeof(T) == 4) { float4 z = make_float4(0.f, 0.f, 0.f, 0.f); v = *reinterpret_cast<V *>(&z); }
                else { double2 z = make_double2(0.0, 0.0); v = *reinterpret_cast<V *>(&z); }
                if (row < n && q < dv) v = __ldg(reinterpret_cast<const V *>(X + row * ldx) + q);
                xv[u] = v;
            }
            if (q < dv) {
#pragma unroll
                for (int j = 0; j < MM; ++j) {
                    if (j < m) {
                        const V cv = *reinterpret_cast<const V *>(cs + (size_t)j * d + (size_t)q * VEC);
                        const T *cp = reinterpret_cast<const T *>(&cv);
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const T *xp = reinterpret_cast<const T *>(&xv[u]);
#pragma unroll
                            for (int e = 0; e < VEC; ++e) {
                                const T df = xp[e] - cp[e];
                                acc[u][j] = fma(df, df, acc[u][j]);
                            }
                        }
                    }
                }
            }
        }
        if constexpr (TR) {
            constexpr int NV = U * MM, R = NV / LPR;
            T v[NV];
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int j = 0; j < MM; ++j) v[u * MM + j] = acc[u][j];
            KmTrReduce<T, NV, NV, LPR / 2>::run(v, l);
            // this lane now holds candidates j = (l % G) * R + i, i < R, of row l / G
            const int j0 = (l % G) * R;
            T best = (T)INFINITY;
#pragma unroll
            for (int i = 0; i < R; ++i) best = (j0 + i < m && v[i] < best) ? v[i] : best;
#pragma unroll
            for (int o = 1; o < G; o <<= 1) {
                const T other = __shfl_xor_sync(0xffffffffu, best, o);
                best = other < best ? other : best;
            }
            mybest = best;
        } else {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                T best = (T)INFINITY;
#pragma unroll
                for (int j = 0; j < MM; ++j) {
                    if (j < m) {
                        T v = acc[u][j];
#pragma unroll
                        for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                        best = v < best ? v : best;
                    }
                }
                if (l == u) mybest = best;
            }
        }
        if (mine) {
            const T r = oldv < mybest ? oldv : mybest;
            mind2[myrow] = r;
            cost += (double)r;
        }
    }
    km_block_atomic_add(cost, cost_out, red);
}

// ---------------------------------------------------------------------------------------------
// K4 on a TMA ring (float32, contiguous 16-byte aligned rows): mindist_vec_kernel's arithmetic, fed from shared memory.
// The register path above keeps U row groups of loads in flight per warp and reaches 0.99 of the HBM peak with one new
// candidate but 0.87 / 0.76 with two / four -- more accumulators per row leave fewer registers for loads in flight.
// Here the loads are not the threads' business: one persistent CTA per SM, warp 0 streams 256-row pieces of X into a
// ring of MT_NST shared-memory stages with 1-D bulk async copies (mbarrier full / empty per stage; a copy stream of this
// kind alone reads 7.4 TB/s), and the eight compute warps take 32 rows of a stage each -- same per-lane accumulation
// order and same cross-lane reduction tree as above, so every row's result is bit-identical to the register path's.
// A warp's 32 running minima are one coalesced load issued before the wait for the stage, and one coalesced store.
// ---------------------------------------------------------------------------------------------
#ifndef MT_NW_DEF
#define MT_NW_DEF 16
#endif
constexpr int MT_NW = MT_NW_DEF;               // compute warps
constexpr int MT_THREADS = 32 * (MT_NW + 1);   // + the producer warp
constexpr int MT_NST = 3;
constexpr int MT_STAGE_BYTES = 64 * 1024;

struct MtShared {
    uint64_t full[MT_NST];
    uint64_t empty[MT_NST];
    float rowmin[MT_NW][32];
    double red[32];
};

template <int LPR, int U, int MM>
__global__ void __launch_bounds__(MT_THREADS, 1)
mindist_tma_kernel(const float *__restrict__ X, int64_t n, int d, const float *__restrict__ cand, int m,
                   float *__restrict__ mind2, int init, double *__restrict__ cost_out, int RW) {
    extern __shared__ __align__(128) unsigned char mt_smem[];
    unsigned char *ring = mt_smem;                                              // [MT_NST][MT_STAGE_BYTES]
    float *cs = reinterpret_cast<float *>(mt_smem + (size_t)MT_NST * MT_STAGE_BYTES);   // [m][d]
    __shared__ MtShared sh;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int e = tid; e < m * d; e += MT_THREADS) cs[e] = cand[e];
    if (tid == 0) {
        for (int i = 0; i < MT_NST; ++i) { km_mbar_init(&sh.full[i], 1); km_mbar_init(&sh.empty[i], MT_NW); }
        km_fence_mbar_init();
    }
    __syncthreads();
    const int R = MT_NW * RW;                                                    // rows per stage
    const int64_t nsub = (n + R - 1) / R;
    const uint32_t rb = (uint32_t)d * 4u;
    double cost = 0.0;

    if (warp == 0) {
        // ================= producer =================
        if (lane == 0) {
            uint32_t i = 0;
            for (int64_t sub = blockIdx.x; sub < nsub; sub += gridDim.x, ++i) {
                const uint32_t st = i % MT_NST;
                const int64_t base = sub * R;
                const uint32_t rows = (uint32_t)((n - base) < R ? (n - base) : R);
                if (i >= MT_NST) km_mbar_wait(&sh.empty[st], ((i / MT_NST) - 1) & 1);
                km_mbar_expect_tx(&sh.full[st], rows * rb);
                km_bulk_g2s(ring + (size_t)st * MT_STAGE_BYTES, X + base * d, rows * rb, &sh.full[st]);
            }
        }
    } else {
        // ================= compute warps: warp w takes rows [w RW, (w + 1) RW) of every stage =================
        constexpr int RPW = 32 / LPR;
        constexpr bool TR = (U * MM >= LPR) && (LPR > 1);
        constexpr int G = TR ? LPR / U : 1;
        const int w = warp - 1;
        const int sub_l = lane / LPR, l = lane % LPR;
        const int dv = d / 4;
        const int own_u = TR ? l / G : l;
        const bool owner = TR ? (l % G) == 0 : l < U;
        uint32_t i = 0;
        for (int64_t sub = blockIdx.x; sub < nsub; sub += gridDim.x, ++i) {
            const uint32_t st = i % MT_NST;
            const int64_t base = sub * R + (int64_t)w * RW;              // this warp's first row
            const int64_t left = n - base;                                // rows of the warp's slice that exist
            float oldv = INFINITY;
            if (!init && lane < RW && lane < left) oldv = mind2[base + lane];
            km_mbar_wait(&sh.full[st], (i / MT_NST) & 1);
            const float *xs = reinterpret_cast<const float *>(ring + (size_t)st * MT_STAGE_BYTES) + (size_t)w * RW * d;
            for (int g0 = 0; g0 < RW; g0 += RPW * U) {
                float acc[U][MM];
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int j = 0; j < MM; ++j) acc[u][j] = 0.f;
                for (int q0 = 0; q0 < dv; q0 += LPR) {
                    const int q = q0 + l;
                    float4 xv[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int r = g0 + u * RPW + sub_l;
                        xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (r < left && q < dv) xv[u] = *reinterpret_cast<const float4 *>(xs + (size_t)r * d + q * 4);
                    }
                    if (q < dv) {
#pragma unroll
                        for (int j = 0; j < MM; ++j) {
                            if (j < m) {
                                const float4 cv = *reinterpret_cast<const float4 *>(cs + (size_t)j * d + (size_t)q * 4);
                                const float *cp = reinterpret_cast<const float *>(&cv);
#pragma unroll
                                for (int u = 0; u < U; ++u) {
                                    const float *xp = reinterpret_cast<const float *>(&xv[u]);
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        const float df = xp[e] - cp[e];
                                        acc[u][j] = fmaf(df, df, acc[u][j]);
                                    }
                                }
                            }
                        }
                    }
                }
                float mybest = INFINITY;
                if constexpr (TR) {
                    constexpr int NV = U * MM, RR = NV / LPR;
                    float v[NV];
#pragma unroll
                    for (int u = 0; u < U; ++u)
#pragma unroll
                        for (int j = 0; j < MM; ++j) v[u * MM + j] = acc[u][j];
                    KmTrReduce<float, NV, NV, LPR / 2>::run(v, l);
                    const int j0 = (l % G) * RR;
                    float best = INFINITY;
#pragma unroll
                    for (int ii = 0; ii < RR; ++ii) best = (j0 + ii < m && v[ii] < best) ? v[ii] : best;
#pragma unroll
                    for (int o = 1; o < G; o <<= 1) {
                        const float other = __shfl_xor_sync(0xffffffffu, best, o);
                        best = other < best ? other : best;
                    }
                    mybest = best;
                } else {
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        float best = INFINITY;
#pragma unroll
                        for (int j = 0; j < MM; ++j) {
                            if (j < m) {
                                float v = acc[u][j];
#pragma unroll
                                for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                                best = v < best ? v : best;
                            }
                        }
                        if (l == u) mybest = best;
                    }
                }
                if (owner) {
                    const int r = g0 + own_u * RPW + sub_l;
                    if (r < RW) sh.rowmin[w][r] = mybest;
                }
            }
            __syncwarp();                                     // rowmin complete; every lane is done with the stage
            if (lane == 0)
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(km_smem_u32(&sh.empty[st])) : "memory");
            if (lane < RW && lane < left) {
                const float nb = sh.rowmin[w][lane];
                const float r = oldv < nb ? oldv : nb;
                mind2[base + lane] = r;
                cost += (double)r;
            }
            __syncwarp();                                     // rowmin is rewritten by the next stage's groups
        }
    }
    __syncthreads();
    km_block_atomic_add(cost, cost_out, sh.red);
}

// ---------------------------------------------------------------------------------------------
// K4 for 3 .. 8 new candidates (float32 rows of 128 / 256 / 512 bytes): one THREAD per row.
// With LPR lanes per row (the kernels above) every candidate's partial sums must be reduced across the lanes: from three
// candidates on that reduction and its bookkeeping are most of the instruction stream (88 instructions per row at 8
// candidates, 36 of them arithmetic) and the pass is bound by its issue slots, not by HBM.  Here a warp stages ITS 32
// rows with cp.async into one of two private shared-memory buffers (16-byte chunks XOR-swizzled by the row, so that
// the 32 lanes reading chunk c of 32 different rows hit different banks), each lane then walks its own row chunk by
// chunk against the candidates (broadcast loads) with one accumulator per candidate -- no cross-lane step at all, and
// lane = row makes the minima one coalesced load and store.  No block barrier in the loop.
// The sum of a row runs over its elements in order, whatever the shard or the launch shape: every row's result
// depends on the row and the candidates alone.
// ---------------------------------------------------------------------------------------------
// warps per CTA (one CTA per SM): 24, 12 for 512-byte rows -- ONE 32-row buffer per warp (192 KB in all at d = 64 and
// 128): a warp's copy is not overlapped with its own arithmetic but with the other warps', and six warps per scheduler
// cover the latencies of the walk better than three with two buffers each did (5.4 -> ms per 6e7 rows at 8 candidates)
__host__ __device__ constexpr int mr_nw(int dv) { return dv <= 16 ? 24 : 12; }

template <int MM, int DV>
__global__ void __launch_bounds__(32 * mr_nw(DV), 1)
mindist_rows_kernel(const float *__restrict__ X, int64_t n, int64_t ldx, const float *__restrict__ cand, int m,
                    float *__restrict__ mind2, int init, double *__restrict__ cost_out) {
    constexpr int D = DV * 4;
    constexpr int RB = DV * 16;                // row bytes
    constexpr int MR_NW = mr_nw(DV);
    extern __shared__ __align__(128) unsigned char mr_smem[];
    float *cs = reinterpret_cast<float *>(mr_smem);                              // [m][D]
    unsigned char *bufs = mr_smem + (((size_t)MM * D * 4 + 127) & ~(size_t)127); // [MR_NW][32][RB]
    __shared__ double red[32];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int e = tid; e < m * D; e += 32 * MR_NW) cs[e] = cand[e];
    __syncthreads();
    unsigned char *mine = bufs + (size_t)warp * 32 * RB;
    const uint32_t mine32 = km_smem_u32(mine);
    // staging: lane l moves chunks l, l + 32, ... of the slice (consecutive lanes -> consecutive 16-byte pieces of a row)
    auto issue = [&](int64_t base) {
        const int64_t left = n - base;
#pragma unroll
        for (int i = 0; i < DV; ++i) {
            const int e = lane + 32 * i, r = e / DV, c = e % DV;
            if (r < left)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(mine32 + (uint32_t)(r * RB + ((c ^ (r & 7)) << 4))),
                             "l"(X + (base + r) * ldx + c * 4) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const int64_t nslices = (n + 31) / 32;
    const int64_t stride = (int64_t)gridDim.x * MR_NW;
    double cost = 0.0;
    const unsigned char *xr = mine + (size_t)lane * RB;
    const uint32_t my_sw = (uint32_t)(lane & 7);
    for (int64_t sl = (int64_t)blockIdx.x * MR_NW + warp; sl < nslices; sl += stride) {
        const int64_t row = sl * 32 + lane;
        issue(sl * 32);
        float oldv = INFINITY;
        if (!init && row < n) oldv = mind2[row];
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();                       // the slice has landed for every lane
        float acc[MM];
#pragma unroll
        for (int j = 0; j < MM; ++j) acc[j] = 0.f;
#pragma unroll 4
        for (int c = 0; c < DV; ++c) {
            const float4 xv = *reinterpret_cast<const float4 *>(xr + ((c ^ my_sw) << 4));
#pragma unroll
            for (int j = 0; j < MM; ++j) {
                if (j < m) {
                    const float4 cv = *reinterpret_cast<const float4 *>(cs + j * D + c * 4);
                    float df;
                    df = xv.x - cv.x; acc[j] = fmaf(df, df, acc[j]);
                    df = xv.y - cv.y; acc[j] = fmaf(df, df, acc[j]);
                    df = xv.z - cv.z; acc[j] = fmaf(df, df, acc[j]);
                    df = xv.w - cv.w; acc[j] = fmaf(df, df, acc[j]);
                }
            }
        }
        __syncwarp();                       // every lane has read its row: the buffer can be refilled
        if (row < n) {
            float best = oldv;
#pragma unroll
            for (int j = 0; j < MM; ++j) best = (j < m && acc[j] < best) ? acc[j] : best;
            mind2[row] = best;
            cost += (double)best;
        }
    }
    km_block_atomic_add(cost, cost_out, red);
}

// (or -1: not for this shape).  Taken for EVERY n when it fits, so that a row's minimum does not depend on which
// kernel the size of its shard would otherwise select.
static int mindist_rows(const float *X, int64_t n, int d, int64_t ldx, const float *cand, int m, float *mind2, int init,
                        double *cost_out, cudaStream_t st) {
    if (m < 3 || m > MD_MAXM || d % 4 != 0 || ldx % 4 != 0 || (reinterpret_cast<uintptr_t>(X) & 15) != 0) return -1;
    if (km_k4_variant() == 1) return -1;
    const int dv = d / 4;
    if (dv != 8 && dv != 16 && dv != 32) return -1;
#define MR_LAUNCH(MM_, DV_)                                                                                      \
    {                                                                                                            \
        constexpr int NW_ = mr_nw(DV_);                                                                          \
        int64_t blocks = ((n + 31) / 32 + NW_ - 1) / NW_;                                                        \
        if (blocks > km_num_sms()) blocks = km_num_sms();                                                        \
        auto kern = mindist_rows_kernel<MM_, DV_>;                                                               \
        static bool attr_set[KM_MAX_DEVICES] = {};                                                               \
        if (int rc = km_set_smem_once(kern, 216 * 1024, attr_set)) return rc;                                    \
        const size_t smem = (((size_t)MM_ * DV_ * 16 + 127) & ~(size_t)127) + (size_t)NW_ * 32 * DV_ * 16;       \
        kern<<<(unsigned)blocks, 32 * NW_, smem, st>>>(X, n, ldx, cand, m, mind2, init, cost_out);               \
    }
#define MR_LAUNCH_D(MM_)                                                                                         \
    if (dv == 8) MR_LAUNCH(MM_, 8) else if (dv == 16) MR_LAUNCH(MM_, 16) else MR_LAUNCH(MM_, 32)
    if (m <= 4) { MR_LAUNCH_D(4) } else { MR_LAUNCH_D(8) }
#undef MR_LAUNCH_D
#undef MR_LAUNCH
    return (int)cudaGetLastError();
}

template <typename T>
static int mindist_impl(const T *X, int64_t n, int d, int64_t ldx, const T *cand, int m, T *mind2, int init,
                        double *cost_out, void *stream) {
    KM_CHECK_ARG(X && cand && mind2 && cost_out, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && m > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(m <= MD_MAXM && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    const size_t smem = sizeof(T) * (size_t)m * d;
    KM_CHECK_ARG(smem <= 128 * 1024, B200KM_ETOOBIG);
    cudaStream_t st = km_stream(stream);
    constexpr int VECW = 16 / sizeof(T);
    if (d % VECW == 0 && ldx % VECW == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0) {
        const int dv = d / VECW;
        int64_t cap = (int64_t)km_num_sms() * 8;
        if constexpr (sizeof(T) == 4) {
            const int rc_rows = mindist_rows(reinterpret_cast<const float *>(X), n, d, ldx, reinterpret_cast<const float *>(cand), m,
                                             reinterpret_cast<float *>(mind2), init, cost_out, st);
            if (rc_rows >= 0) return rc_rows;
        }
        if constexpr (sizeof(T) == 4) {
            // the TMA ring (B200KM_OPT_K4_VARIANT = 1 keeps the register path): contiguous rows, a row group size that
            // divides a warp's slice of a stage, and enough rows to fill the ring on every SM
            // (measured, tools/k4_bench.py: the ring wins whenever the pass is HBM-bound -- one or two new candidates, any
            //  row width -- and for 132..256-byte rows with any number; with 3+ candidates on narrow or wide rows the
            //  pass is bound by its arithmetic and the register path's 8 resident CTAs per SM do as well or better)
            const bool ring_pays = m <= 2 || (dv > 8 && dv <= 16) || (dv > 16 && m > 4);
            if (km_k4_variant() != 1 && (ring_pays || km_k4_variant() == 2) && ldx == d && n >= (int64_t)km_num_sms() * 1024) {
                int rc_done = -1;
#define MTV_TRY(LPR_, U_, MM_)                                                                                   \
    {                                                                                                            \
        constexpr int GR = (32 / LPR_) * U_;              /* rows of one accumulation group */                   \
        int RW = 32;                                                                                             \
        while (RW >= GR && (size_t)MT_NW * RW * d * 4 > (size_t)MT_STAGE_BYTES) RW /= 2;                         \
        if (GR <= 32 && RW >= GR && RW % GR == 0) {                                                              \
            auto kern = mindist_tma_kernel<LPR_, U_, MM_>;                                                       \
            static bool attr_set[KM_MAX_DEVICES] = {};                                                           \
            if (int rc = km_set_smem_once(kern, 216 * 1024, attr_set)) return rc;                                \
            const size_t tsm = (size_t)MT_NST * MT_STAGE_BYTES + smem;                                           \
            int64_t blocks = (n + (int64_t)MT_NW * RW - 1) / ((int64_t)MT_NW * RW);                              \
            if (blocks > km_num_sms()) blocks = km_num_sms();                                                    \
            kern<<<(unsigned)blocks, MT_THREADS, tsm, st>>>(reinterpret_cast<const float *>(X), n, d,            \
                                                            reinterpret_cast<const float *>(cand), m,            \
                                                            reinterpret_cast<float *>(mind2), init, cost_out, RW); \
            rc_done = (int)cudaGetLastError();                                                                   \
        }                                                                                                        \
    }
                /* rows in flight per lane group: U * MM == LPR values per lane (one transposing reduction, one value
                   per lane afterwards), at most 8 -- the loads come from shared memory, not from HBM */
#define MTV_LAUNCH(LPR_)                                                                                         \
    if (m <= 2) MTV_TRY(LPR_, (LPR_ / 2 < 8 ? LPR_ / 2 : 8), 2)                                                  \
    else if (m <= 4) MTV_TRY(LPR_, (LPR_ / 4 < 8 ? LPR_ / 4 : 8), 4)                                             \
    else MTV_TRY(LPR_, (LPR_ / 8 < 8 ? LPR_ / 8 : 8), MD_MAXM)
                if (dv <= 4) { }                      /* rows of <= 64 bytes: copies too small to pay */
                else if (dv <= 8) MTV_LAUNCH(8)
                else if (dv <= 16) MTV_LAUNCH(16)
                else if (dv <= 128) MTV_LAUNCH(32)
#undef MTV_LAUNCH
#undef MTV_TRY
                if (rc_done >= 0) return rc_done;
            }
        }
#define MDV_LAUNCH(LPR_, U_)  /* U_ <= LPR_: every row of a warp step needs an owner lane */                   \
    if (m <= 2) MDV_LAUNCH2(LPR_, (2 * U_ <= LPR_ ? 2 * U_ : U_), 2)                                             \
    else if (m <= 4) MDV_LAUNCH2(LPR_, (2 * U_ <= LPR_ ? 2 * U_ : U_), 4)                                       \
    else MDV_LAUNCH2(LPR_, U_, MD_MAXM)
#define MDV_LAUNCH2(LPR_, U_, MM_)                                                                               \
    {                                                                                                            \
        auto kern = mindist_vec_kernel<T, LPR_, U_, MM_>;                                                        \
        static bool attr_set[KM_MAX_DEVICES] = {};                                                               \
        if (int rc = km_set_smem_once(kern, 128 * 1024, attr_set)) return rc;                                    \
        const int64_t rows_per_block = (int64_t)(32 / LPR_) * U_ * (MD_THREADS / 32);                            \
        int64_t blocks = (n + rows_per_block - 1) / rows_per_block;                                              \
        if (blocks > cap) blocks = cap;                                                                          \
        kern<<<(unsigned)blocks, MD_THREADS, smem, st>>>(X, n, d, ldx, cand, m, mind2, init, cost_out);          \
    }
        if (dv <= 1) MDV_LAUNCH(1, 1)
        else if (dv <= 2) MDV_LAUNCH(2, 2)
        else if (dv <= 4) MDV_LAUNCH(4, 4)
        else if (dv <= 8) MDV_LAUNCH(8, 4)
        else if (dv <= 16) MDV_LAUNCH(16, 4)
        else MDV_LAUNCH(32, 4)
#undef MDV_LAUNCH
#undef MDV_LAUNCH2
        KM_LAUNCH_RET();
    }
    int lpr = 32;
    if (d <= 4) lpr = 1; else if (d <= 8) lpr = 2; else if (d <= 16) lpr = 4; else if (d <= 32) lpr = 8;
    else if (d <= 64) lpr = 16;
    const int rpw = 32 / lpr;
    int64_t blocks = (n + (int64_t)rpw * (MD_THREADS / 32) - 1) / ((int64_t)rpw * (MD_THREADS / 32));
    const int64_t cap = (int64_t)km_num_sms() * 8;
    if (blocks > cap) blocks = cap;
#define MD_LAUNCH(LPR_)                                                                                          \
    {                                                                                                            \
        auto kern = mindist_kernel<T, LPR_>;                                                                     \
        static bool attr_set[KM_MAX_DEVICES] = {};                                                               \
        if (int rc = km_set_smem_once(kern, 128 * 1024, attr_set)) return rc;                                    \
        kern<<<(unsigned)blocks, MD_THREADS, smem, st>>>(X, n, d, ldx, cand, m, mind2, init, cost_out);          \
    }
    switch (lpr) {
        case 1: MD_LAUNCH(1) break;
        case 2: MD_LAUNCH(2) break;
        case 4: MD_LAUNCH(4) break;
        case 8: MD_LAUNCH(8) break;
        case 16: MD_LAUNCH(16) break;
        default: MD_LAUNCH(32) break;
    }
#undef MD_LAUNCH
    KM_LAUNCH_RET();
}

extern "C" int b200km_mindist_update_f32(const float *X, int64_t n, int d, int64_t ldx, const float *cand, int m,
                                         float *mind2, int init, double *cost_out, void *stream) {
    return mindist_impl<float>(X, n, d, ldx, cand, m, mind2, init, cost_out, stream);
}
extern "C" int b200km_mindist_update_f64(const double *X, int64_t n, int d, int64_t ldx, const double *cand, int m,
                                         double *mind2, int init, double *cost_out, void *stream) {
    return mindist_impl<double>(X, n, d, ldx, cand, m, mind2, init, cost_out, stream);
}

// Many new candidates at once (a large oversampling factor: the paper's l = 2k brings ~2k picks per round): one K4
// launch handles 8, so a round would cost m / 8 passes over X.  The host then takes K1 against the m candidates (one
// tensor-core pass, the distance matrix never exists) and this kernel merges its winner distances into the running
// minimum and sums the cost: mind2[i] = min(mind2[i], other[i]).
template <typename T>
__global__ void __launch_bounds__(256)
mindist_merge_kernel(const T *__restrict__ other, int64_t n, T *__restrict__ mind2, int init, double *__restrict__ cost) {
    __shared__ double red[32];
    double c = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        T v = __ldg(other + i);
        if (!init) {
            const T o = mind2[i];
            v = o < v ? o : v;
        }
        mind2[i] = v;
        c += (double)v;
    }
    km_block_atomic_add(c, cost, red);
}
template <typename T>
static int mindist_merge_impl(const T *other, int64_t n, T *mind2, int init, double *cost, void *stream) {
    KM_CHECK_ARG(other && mind2 && cost && n > 0, B200KM_EINVAL);
    int64_t blocks = (n + 255) / 256;
    if (blocks > km_num_sms() * 16) blocks = km_num_sms() * 16;
    mindist_merge_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(other, n, mind2, init, cost);
    KM_LAUNCH_RET();
}
extern "C" int b200km_mindist_merge_f32(const float *other, int64_t n, float *mind2, int init, double *cost_out,
                                        void *stream) {
    return mindist_merge_impl<float>(other, n, mind2, init, cost_out, stream);
}
extern "C" int b200km_mindist_merge_f64(const double *other, int64_t n, double *mind2, int init, double *cost_out,
                                        void *stream) {
    return mindist_merge_impl<double>(other, n, mind2, init, cost_out, stream);
}

// ---------------------------------------------------------------------------------------------
// k-means++ trial potentials: pots[j] = sum_i min(mind2[i], ||x_i - cand_j||^2), mind2 left untouched.
// Replaces, per k-means++ step, scikit-learn's `distance_to_candidates = euclidean_distances(X[candidate_ids], X)`
// + `np.minimum(closest_dist_sq, ...)` + `@ sample_weight` (the (trials x n) matrix that the reference's init_pp,
// dask_ml/cluster/k_means.py:404-423 -> sklearn _kmeans_plusplus, materialises on the host).
// ---------------------------------------------------------------------------------------------
template <typename T, int LPR>
__global__ void __launch_bounds__(MD_THREADS)
mindist_pots_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const T *__restrict__ cand, int m,
                    const T *__restrict__ mind2, double *__restrict__ pots_out) {
    extern __shared__ __align__(16) unsigned char md_smem[];
    T *cs = reinterpret_cast<T *>(md_smem);  // [m][d]
    __shared__ double red[32];
    for (int e = threadIdx.x; e < m * d; e += MD_THREADS) cs[e] = cand[e];
    __syncthreads();
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int sub = lane / LPR, l = lane % LPR;
    const int64_t warp_global = ((int64_t)blockIdx.x * MD_THREADS + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * MD_THREADS) >> 5;
    double pot[MD_MAXM];
#pragma unroll
    for (int j = 0; j < MD_MAXM; ++j) pot[j] = 0.0;
    for (int64_t r0 = warp_global * RPW; r0 < n; r0 += nwarps * RPW) {
        const int64_t row = r0 + sub;
        const bool live = row < n;
        T acc[MD_MAXM];
#pragma unroll
        for (int j = 0; j < MD_MAXM; ++j) acc[j] = (T)0;
        const T *xr = X + (live ? row : 0) * ldx;
#pragma unroll 4
        for (int f = l; f < d; f += LPR) {
            const T xv = live ? __ldg(xr + f) : (T)0;
#pragma unroll
            for (int j = 0; j < MD_MAXM; ++j) {
                if (j < m) {
                    const T df = xv - cs[j * d + f];
                    acc[j] = fma(df, df, acc[j]);
                }
            }
        }
        const T old = (live && l == 0) ? mind2[row] : (T)0;
#pragma unroll
        for (int j = 0; j < MD_MAXM; ++j) {
            if (j < m) {
                T v = acc[j];
#pragma unroll
                for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (live && l == 0) pot[j] += (double)(old < v ? old : v);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < MD_MAXM; ++j) {
        if (j < m) {
            km_block_atomic_add(pot[j], pots_out + j, red);
            __syncthreads();   // red[] is reused by the next candidate
        }
    }
}

template <typename T>
static int mindist_pots_impl(const T *X, int64_t n, int d, int64_t ldx, const T *cand, int m, const T *mind2,
                             double *pots_out, void *stream) {
    KM_CHECK_ARG(X && cand && mind2 && pots_out, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && m > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(m <= MD_MAXM && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    const size_t smem = sizeof(T) * (size_t)m * d;
    KM_CHECK_ARG(smem <= 128 * 1024, B200KM_ETOOBIG);
    cudaStream_t st = km_stream(stream);
    int lpr = 32;
    if (d <= 4) lpr = 1; else if (d <= 8) lpr = 2; else if (d <= 16) lpr = 4; else if (d <= 32) lpr = 8;
    else if (d <= 64) lpr = 16;
    const int rpw = 32 / lpr;
    int64_t blocks = (n + (int64_t)rpw * (MD_THREADS / 32) - 1) / ((int64_t)rpw * (MD_THREADS / 32));
    const int64_t cap = (int64_t)km_num_sms() * 8;
    if (blocks > cap) blocks = cap;
#define MDP_LAUNCH(LPR_)                                                                                         \
    {                                                                                                            \
        auto kern = mindist_pots_kernel<T, LPR_>;                                                                \
        static bool attr_set[KM_MAX_DEVICES] = {};                                                               \
        if (int rc = km_set_smem_once(kern, 128 * 1024, attr_set)) return rc;                                    \
        kern<<<(unsigned)blocks, MD_THREADS, smem, st>>>(X, n, d, ldx, cand, m, mind2, pots_out);                \
    }
    switch (lpr) {
        case 1: MDP_LAUNCH(1) break;
        case 2: MDP_LAUNCH(2) break;
        case 4: MDP_LAUNCH(4) break;
        case 8: MDP_LAUNCH(8) break;
        case 16: MDP_LAUNCH(16) break;
        default: MDP_LAUNCH(32) break;
    }
#undef MDP_LAUNCH
    KM_LAUNCH_RET();
}

extern "C" int b200km_mindist_pots_f32(const float *X, int64_t n, int d, int64_t ldx, const float *cand, int m,
                                       const float *mind2, double *pots_out, void *stream) {
    return mindist_pots_impl<float>(X, n, d, ldx, cand, m, mind2, pots_out, stream);
}
extern "C" int b200km_mindist_pots_f64(const double *X, int64_t n, int d, int64_t ldx, const double *cand, int m,
                                       const double *mind2, double *pots_out, void *stream) {
    return mindist_pots_impl<double>(X, n, d, ldx, cand, m, mind2, pots_out, stream);
}

// ---------------------------------------------------------------------------------------------
// K5 Bernoulli oversampling
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
bernoulli_kernel(const T *__restrict__ mind2, int64_t n, double scale, uint32_t k0, uint32_t k1, uint32_t round,
                 int64_t row_offset, const double *__restrict__ draws, int64_t *__restrict__ picked,
                 int32_t *__restrict__ count, int capacity) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double p = scale * (double)__ldg(mind2 + i);  // k_means.py:535
        double u;
        if (draws) {
            u = __ldg(draws + i);
        } else {
            const uint64_t g = (uint64_t)(row_offset + i);
            const KmPhilox4 w = km_philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), round, 0x6B6D7C7Cu, k0, k1);
            u = km_u53(w.x, w.y);
        }
        if (p > u) {  // :538
            const int pos = atomicAdd(count, 1);
            if (pos < capacity) picked[pos] = row_offset + i;
        }
    }
}

template <typename T>
static int bernoulli_impl(const T *mind2, int64_t n, double scale, uint64_t seed, uint32_t round, int64_t row_offset,
                          const double *draws, int64_t *picked, int32_t *count, int capacity, void *stream) {
    KM_CHECK_ARG(mind2 && picked && count && n > 0 && capacity > 0, B200KM_EINVAL);
    int64_t blocks = (n + 255) / 256;
    if (blocks > km_num_sms() * 16) blocks = km_num_sms() * 16;
    bernoulli_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(
        mind2, n, scale, (uint32_t)seed, (uint32_t)(seed >> 32), round, row_offset, draws, picked, count, capacity);
    KM_LAUNCH_RET();
}

extern "C" int b200km_bernoulli_pick_f32(const float *mind2, int64_t n, double scale, uint64_t seed, uint32_t round,
                                         int64_t row_offset, const double *draws, int64_t *picked, int32_t *count,
                                         int capacity, void *stream) {
    return bernoulli_impl<float>(mind2, n, scale, seed, round, row_offset, draws, picked, count, capacity, stream);
}
extern "C" int b200km_bernoulli_pick_f64(const double *mind2, int64_t n, double scale, uint64_t seed, uint32_t round,
                                         int64_t row_offset, const double *draws, int64_t *picked, int32_t *count,
                                         int capacity, void *stream) {
    return bernoulli_impl<double>(mind2, n, scale, seed, round, row_offset, draws, picked, count, capacity, stream);
}

// ---------------------------------------------------------------------------------------------
// K6 gather rows
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128)
gather_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const int64_t *__restrict__ idx, int m,
              T *__restrict__ out) {
    for (int j = blockIdx.x; j < m; j += gridDim.x) {
        const int64_t r = idx[j];
        const bool ok = r >= 0 && r < n;
        for (int f = threadIdx.x; f < d; f += blockDim.x) out[(size_t)j * d + f] = ok ? X[r * ldx + f] : (T)NAN;
    }
}

template <typename T>
static int gather_impl(const T *X, int64_t n, int d, int64_t ldx, const int64_t *idx, int m, T *out, void *stream) {
    KM_CHECK_ARG(X && idx && out && n > 0 && d > 0 && m > 0 && ldx >= d, B200KM_EINVAL);
    int blocks = m < km_num_sms() * 16 ? m : km_num_sms() * 16;
    gather_kernel<T><<<blocks, 128, 0, km_stream(stream)>>>(X, n, d, ldx, idx, m, out);
    KM_LAUNCH_RET();
}

extern "C" int b200km_gather_rows_f32(const float *X, int64_t n, int d, int64_t ldx, const int64_t *idx, int m,
                                      float *out, void *stream) {
    return gather_impl<float>(X, n, d, ldx, idx, m, out, stream);
}
extern "C" int b200km_gather_rows_f64(const double *X, int64_t n, int d, int64_t ldx, const int64_t *idx, int m,
                                      double *out, void *stream) {
    return gather_impl<double>(X, n, d, ldx, idx, m, out, stream);
}

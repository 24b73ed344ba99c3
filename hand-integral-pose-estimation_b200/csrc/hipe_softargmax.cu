// Soft-argmax over per-joint 3-D heatmap volumes: forward, backward, volume softmax and the
// fused JointLocationLoss forward+backward, for sm_100a.
//
// Replaces common/nets/loss.py:23-91 of the reference (and what autograd derives from it).
//
// One CLUSTER of S CTAs owns one volume (V = D*H*W contiguous elements, x fastest).  The
// volume is seen as NR = D*H rows of W elements; CTA c of the cluster owns the row range
// [NR*c/S, NR*(c+1)/S).  Inside a CTA 256 consumer threads tile rows x 16-byte column
// groups: thread t owns column group t % G (G = W*sizeof(T)/16) and visits rows
// t/G, t/G + R, t/G + 2R, ... (R = 256/G), so that
//   * every visit of the CTA is one fully coalesced run of R*G*16 contiguous bytes,
//   * a thread's x indices never change: the x marginal is a per-thread register
//     accumulator per column and Sum(e*x) costs nothing per element,
//   * y and z change once per visit: one FMA each per 16-byte vector.
// Data reaches the consumers either through a ring of 16 KB shared-memory stages filled by
// a producer warp with 1-D TMA bulk copies completing on mbarriers (HIPE_VARIANT_TMA), or by
// 128-bit ld.global.nc straight into registers (HIPE_VARIANT_LDG).
//
// Softmax is "online": each thread keeps a private running max m and rescales its four
// accumulators when a stage raises it; (m, S0, Sx, Sy, Sz) partials are merged thread ->
// warp (shuffles) -> CTA (shared memory) -> cluster (DSMEM).  The backward recomputes
// p_i = exp(s_i - M)/L from the logits and the saved (M, L): probabilities never exist in
// HBM.
//
// Precision.  The gradient is p_i * sum_a g_a (idx_a - E_a)/dim_a, so an absolute error in
// an expectation E_a is amplified by 1/|idx_a - E_a| at exactly the voxels that carry the
// largest gradient (a sharp peak sits within a fraction of a voxel of E).  fp32 sums of
// e*idx and e drift apart by ~1e-6 relative = 3e-5 voxels, which alone breaks a 1e-5 gate.
// So: exponentials are fp32 (MUFU.EX2), but every 16-byte vector's (sum e, sum j*e) is
// widened once and the four moment sums are carried in fp64 -- all of them built from the
// SAME rounded row sums, so rounding errors only re-weight voxels (an error ~ eps*sigma, not
// eps*E) -- and the expectations travel to the backward as fp64, where they are split into
// hi+lo floats so that (idx - E) keeps its absolute accuracy in fp32 arithmetic.
#include "hipe_common.cuh"
#include <stdlib.h>

namespace hipe {

constexpr int kConsumers = 256;                   // consumer threads per CTA
constexpr int kPasses = 4;                        // 16-byte vectors per thread per stage
constexpr int kStages = 4;                        // TMA ring depth
constexpr int kStageBytes = kPasses * kConsumers * 16;  // 16 KB
constexpr int kMaxCluster = 8;

enum Op : int {
    OP_REDUCE_SOFTMAX = 0,  // forward
    OP_REDUCE_PLAIN = 1,    // forward on already-normalised heatmaps (no exp)
    OP_WRITE_GRAD = 2,      // backward
    OP_WRITE_PROB = 3,      // softmax probabilities
    OP_FUSED = 4            // forward + L1 gradient + backward in one launch
};

struct Params {
    const unsigned char* in;
    unsigned char* out;
    long long n_vol;
    long long vol_bytes;
    int W, H, D;
    int G;   // 16-byte column groups per row
    int R;   // rows per visit of the CTA
    int NR;  // rows per volume
    int reverse;
    int pol1, pol2;  // L2 policy of the TMA loads of pass 1 / pass 2: 0 normal, 1 evict_first, 2 evict_last
    // reduce outputs
    float* coords;
    double* expect;
    float* stats;
    // write inputs
    const float* stats_in;
    const double* expect_in;
    const float* gcoords;
    const float* gscale;
    // fused
    const void* gt;
    const void* vis;
    int tgt_f64;
    double norm;
    double* loss_terms;
};

struct SmemLayout {
    // dynamic shared memory: [stages][kStageBytes] then this struct
    unsigned long long full[kStages];
    unsigned long long empty[kStages];
    double wred[kConsumers / 32][4];
    float wmax[kConsumers / 32];
    double part[4];                // this CTA's partial (S0, Sx, Sy, Sz), read by peers through DSMEM
    float part_m;                  // ... and the maximum it is relative to
    double all[kMaxCluster][4];
    float all_m[kMaxCluster];
    double finE[3];                // Ex, Ey, Ez (voxel units)
    float fin[5];                  // M, L, gx, gy, gz
};

__device__ __forceinline__ void consumer_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
}

template <bool kSoftmax>
struct Acc {
    float m;    // running max (natural units)
    float mL;   // fl(m * log2e): the exponent offset every e of this thread is relative to
    double s0;  // sum e
    double sx;  // sum e * (x - x0)   (x0 = the thread's first column)
    double sy;  // sum e * y
    double sz;  // sum e * z
    __device__ __forceinline__ void init() {
        m = kSoftmax ? kNegInf : 0.f;
        mL = kSoftmax ? kNegInf : 0.f;
        s0 = sx = sy = sz = 0.0;
    }
    __device__ __forceinline__ void scale(float f) {
        const double d = (double)f;
        s0 *= d; sx *= d; sy *= d; sz *= d;
    }
};

// Weight that re-bases a partial computed against maximum m_from onto maximum m_to >= m_from.
// Every exponential is 2^(s*log2e - fl(m*log2e)), so the weight is 2^(fl(m_from*log2e) -
// fl(m_to*log2e)) with the SAME rounded products (no contraction): a partial is re-based by
// exactly the offset its exponentials were computed against.
__device__ __forceinline__ float rebase_weight(float m_from, float m_to) {
    return (m_from == kNegInf) ? 0.f
                               : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

template <typename T, bool kTma, int kOp>
__global__ void __launch_bounds__(kTma ? kConsumers + 32 : kConsumers)
    k_stream(const Params p) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    constexpr bool kHasReduce = (kOp == OP_REDUCE_SOFTMAX || kOp == OP_REDUCE_PLAIN || kOp == OP_FUSED);
    constexpr bool kHasWrite = (kOp == OP_WRITE_GRAD || kOp == OP_WRITE_PROB || kOp == OP_FUSED);
    constexpr bool kSoftmax = (kOp != OP_REDUCE_PLAIN);
    constexpr int kNumPassesOverVolume = (kOp == OP_FUSED) ? 2 : 1;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    SmemLayout& sm = *reinterpret_cast<SmemLayout*>(smem_raw + (kTma ? kStages * kStageBytes : 0));

    const int tid = threadIdx.x;
    const uint32_t S = cluster_nctarank();
    const uint32_t crank = cluster_ctarank();
    long long vol = blockIdx.x / S;
    if (p.reverse) vol = p.n_vol - 1 - vol;

    // this CTA's rows
    const int r0 = (int)(((long long)p.NR * crank) / S);
    const int r1 = (int)(((long long)p.NR * (crank + 1)) / S);
    const int nrows = r1 - r0;
    const int row_bytes = p.G * 16;
    const int pass_bytes = p.R * row_bytes;
    const long long cta_bytes = (long long)nrows * row_bytes;
    const int stage_bytes = kPasses * pass_bytes;
    const int nst = (int)((cta_bytes + stage_bytes - 1) / stage_bytes);
    const unsigned char* src = p.in + vol * p.vol_bytes + (long long)r0 * row_bytes;

    // ---------------- barrier setup (TMA variant) ----------------
    if constexpr (kTma) {
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < kStages; ++s) {
                mbar_init(smem_u32(&sm.full[s]), 1);
                mbar_init(smem_u32(&sm.empty[s]), kConsumers / 32);
            }
            mbar_fence_init();
            fence_proxy_async();
        }
        __syncthreads();

        if (tid >= kConsumers) {
            // ---------------- producer warp ----------------
            // It has nothing to contribute to the cluster reduction: arrive right away so the
            // consumers' cluster barrier can complete while this warp is still streaming.
            if (S > 1 && kHasReduce) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
            if (tid == kConsumers) {
                const uint64_t pn = policy_evict_normal(), pf = policy_evict_first(), pl = policy_evict_last();
                const uint64_t pol_a = p.pol1 == 0 ? pn : (p.pol1 == 1 ? pf : pl);
                const uint64_t pol_b = p.pol2 == 0 ? pn : (p.pol2 == 1 ? pf : pl);
                const int total = nst * kNumPassesOverVolume;
                for (int it = 0; it < total; ++it) {
                    const int slot = it % kStages;
                    const int i = (it >= nst) ? it - nst : it;
                    if (it >= kStages) mbar_wait(smem_u32(&sm.empty[slot]), ((it / kStages) - 1) & 1);
                    const long long off = (long long)i * stage_bytes;
                    const long long left = cta_bytes - off;
                    const uint32_t bytes = (uint32_t)(left < stage_bytes ? left : stage_bytes);
                    const uint32_t bar = smem_u32(&sm.full[slot]);
                    mbar_arrive_expect_tx(bar, bytes);
                    tma_load_1d(smem_u32(smem_raw + slot * kStageBytes), src + off, bytes, bar,
                                (it < nst) ? pol_a : pol_b);
                }
            }
            __syncwarp();
            if (S > 1 && kHasReduce) {
                asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
                // second cluster phase: peers must not exit while their `part` is being read
                asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
                asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
            }
            return;
        }
    }

    // ---------------- consumers ----------------
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const bool active = tid < p.G * p.R;
    const int cg = active ? tid % p.G : 0;
    const int rr = active ? tid / p.G : 0;
    // number of row visits of this thread inside the CTA's range
    const int nvis = (active && nrows > rr) ? (nrows - rr + p.R - 1) / p.R : 0;
    const int x0i = cg * VEC;
    const int row_first = r0 + rr;
    const int y_first = row_first % p.H;
    const int z_first = row_first / p.H;
    const uint32_t lds_base = kTma ? smem_u32(smem_raw) + 16u * tid : 0u;
    const unsigned char* gsrc = src + 16ll * tid;
    int it = 0;  // stage counter across both passes (ring position)

    float M = 0.f, invL = 1.f, gx = 0.f, gy = 0.f, gz = 0.f;
    double Ex = 0.0, Ey = 0.0, Ez = 0.0;

    if constexpr (kHasReduce) {
        Acc<kSoftmax> a;
        a.init();
        double yd = (double)y_first, zd = (double)z_first;
        const double Hd = (double)p.H, Rd = (double)p.R;
        for (int i = 0; i < nst; ++i, ++it) {
            float v[kPasses][VEC];
            // ---- fetch the stage into registers ----
            if constexpr (kTma) {
                const int slot = it % kStages;
                mbar_wait(smem_u32(&sm.full[slot]), (it / kStages) & 1);
#pragma unroll
                for (int q = 0; q < kPasses; ++q) {
                    if (i * kPasses + q < nvis) {
                        V16::unpack(lds128(lds_base + slot * kStageBytes + q * pass_bytes), v[q]);
                    } else {
#pragma unroll
                        for (int j = 0; j < VEC; ++j) v[q][j] = kSoftmax ? kNegInf : 0.f;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&sm.empty[slot]));
            } else {
#pragma unroll
                for (int q = 0; q < kPasses; ++q) {
                    if (i * kPasses + q < nvis) {
                        V16::unpack(ldg_stream(gsrc + (long long)(i * kPasses + q) * pass_bytes), v[q]);
                    } else {
#pragma unroll
                        for (int j = 0; j < VEC; ++j) v[q][j] = kSoftmax ? kNegInf : 0.f;
                    }
                }
            }
            // ---- online max ----
            if constexpr (kSoftmax) {
                float smax = kNegInf;
#pragma unroll
                for (int q = 0; q < kPasses; ++q)
#pragma unroll
                    for (int j = 0; j < VEC; ++j) smax = fmaxf(smax, v[q][j]);
                if (smax > a.m) {
                    a.scale(rebase_weight(a.m, smax));
                    a.m = smax;
                    a.mL = __fmul_rn(smax, kLog2e);
                }
            }
            // ---- accumulate: fp32 inside one 16-byte vector, fp64 across vectors ----
            const bool live = kSoftmax ? (a.m != kNegInf) : true;
#pragma unroll
            for (int q = 0; q < kPasses; ++q) {
                if (live) {
                    float rs = 0.f, xs = 0.f;
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const float e = kSoftmax ? ex2(fmaf(v[q][j], kLog2e, -a.mL)) : v[q][j];
                        rs += e;
                        if (j) xs = fmaf((float)j, e, xs);
                    }
                    const double drs = (double)rs;
                    a.s0 += drs;
                    a.sx += (double)xs;
                    a.sy = fma(yd, drs, a.sy);
                    a.sz = fma(zd, drs, a.sz);
                }
                yd += Rd;
                while (yd >= Hd) {
                    yd -= Hd;
                    zd += 1.0;
                }
            }
        }

        // ---- thread -> warp -> CTA ----
        double s0 = a.s0, sy = a.sy, sz = a.sz;
        double sx = fma((double)x0i, a.s0, a.sx);
        float m = a.m;
        if constexpr (kSoftmax) {
            const float wm = warp_max(m);
            const double f = (double)rebase_weight(m, wm);
            s0 *= f; sx *= f; sy *= f; sz *= f;
            m = wm;
        }
        s0 = warp_sum(s0); sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
        if (lane == 0) {
            sm.wred[warp][0] = s0; sm.wred[warp][1] = sx; sm.wred[warp][2] = sy; sm.wred[warp][3] = sz;
            sm.wmax[warp] = m;
        }
        consumer_barrier();
        if (warp == 0) {
            constexpr int NW = kConsumers / 32;
            float pm = (lane < NW) ? sm.wmax[lane] : (kSoftmax ? kNegInf : 0.f);
            double q0 = (lane < NW) ? sm.wred[lane][0] : 0.0;
            double q1 = (lane < NW) ? sm.wred[lane][1] : 0.0;
            double q2 = (lane < NW) ? sm.wred[lane][2] : 0.0;
            double q3 = (lane < NW) ? sm.wred[lane][3] : 0.0;
            if constexpr (kSoftmax) {
                const float wm = warp_max(pm);
                const double f = (double)rebase_weight(pm, wm);
                q0 *= f; q1 *= f; q2 *= f; q3 *= f;
                pm = wm;
            }
            q0 = warp_sum(q0); q1 = warp_sum(q1); q2 = warp_sum(q2); q3 = warp_sum(q3);
            if (lane == 0) {
                sm.part[0] = q0; sm.part[1] = q1; sm.part[2] = q2; sm.part[3] = q3;
                sm.part_m = pm;
            }
        }
        // ---- CTA -> cluster (DSMEM) ----
        if (S > 1) {
            cluster_sync_all();  // every CTA's `part` is written and visible cluster-wide
            if (tid < 4 * (int)S) sm.all[tid / 4][tid % 4] = dsmem_ld_f64(smem_u32(&sm.part[tid % 4]), tid / 4);
            if (tid >= 32 && tid < 32 + (int)S) sm.all_m[tid - 32] = dsmem_ld_f32(smem_u32(&sm.part_m), tid - 32);
            // done reading peers: arrive now, wait before exiting
            consumer_barrier();
            asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        } else {
            consumer_barrier();
            if (tid < 4) sm.all[0][tid] = sm.part[tid];
            if (tid == 4) sm.all_m[0] = sm.part_m;
            consumer_barrier();
        }
        if (tid == 0) {
            float fm = kSoftmax ? kNegInf : 0.f;
            if constexpr (kSoftmax)
                for (uint32_t r = 0; r < S; ++r) fm = fmaxf(fm, sm.all_m[r]);
            double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
            for (uint32_t r = 0; r < S; ++r) {
                const double f = kSoftmax ? (double)rebase_weight(sm.all_m[r], fm) : 1.0;
                t0 = fma(f, sm.all[r][0], t0);
                t1 = fma(f, sm.all[r][1], t1);
                t2 = fma(f, sm.all[r][2], t2);
                t3 = fma(f, sm.all[r][3], t3);
            }
            // loss.py:39-41 (expectations) and :54-56 (normalisation), in fp64, rounded once
            const double ex_ = kSoftmax ? t1 / t0 : t1;
            const double ey_ = kSoftmax ? t2 / t0 : t2;
            const double ez_ = kSoftmax ? t3 / t0 : t3;
            const float cxn = (float)(ex_ / (double)p.W - 0.5);
            const float cyn = (float)(ey_ / (double)p.H - 0.5);
            const float czn = (float)(ez_ / (double)p.D - 0.5);
            float g0 = 0.f, g1 = 0.f, g2 = 0.f;
            if constexpr (kOp == OP_FUSED) {
                // loss.py:86-89 on this volume's three coordinates, in the targets' precision
                double t[3], w[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    if (p.tgt_f64) {
                        t[k] = reinterpret_cast<const double*>(p.gt)[vol * 3 + k];
                        w[k] = reinterpret_cast<const double*>(p.vis)[vol * 3 + k];
                    } else {
                        t[k] = reinterpret_cast<const float*>(p.gt)[vol * 3 + k];
                        w[k] = reinterpret_cast<const float*>(p.vis)[vol * 3 + k];
                    }
                }
                const double c[3] = {(double)cxn, (double)cyn, (double)czn};
                float g[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const double dlt = c[k] - t[k];
                    const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
                    g[k] = (float)(sg * w[k] * p.norm);
                    if (crank == 0 && p.loss_terms) p.loss_terms[vol * 3 + k] = fabs(dlt) * w[k];
                }
                g0 = g[0]; g1 = g[1]; g2 = g[2];
            }
            if (crank == 0) {
                if (p.coords) { p.coords[vol * 3 + 0] = cxn; p.coords[vol * 3 + 1] = cyn; p.coords[vol * 3 + 2] = czn; }
                if (p.expect) { p.expect[vol * 3 + 0] = ex_; p.expect[vol * 3 + 1] = ey_; p.expect[vol * 3 + 2] = ez_; }
                if (p.stats) { p.stats[vol * 2 + 0] = fm; p.stats[vol * 2 + 1] = (float)t0; }
            }
            sm.finE[0] = ex_; sm.finE[1] = ey_; sm.finE[2] = ez_;
            sm.fin[0] = fm; sm.fin[1] = (float)t0; sm.fin[2] = g0; sm.fin[3] = g1; sm.fin[4] = g2;
        }
        if constexpr (kOp == OP_FUSED) {
            consumer_barrier();
            M = sm.fin[0]; invL = 1.f / sm.fin[1];
            Ex = sm.finE[0]; Ey = sm.finE[1]; Ez = sm.finE[2];
            gx = sm.fin[2]; gy = sm.fin[3]; gz = sm.fin[4];
        }
    } else {
        // write-only ops read their per-volume scalars from the forward's outputs
        M = p.stats_in[vol * 2 + 0];
        invL = 1.f / p.stats_in[vol * 2 + 1];
        if constexpr (kOp == OP_WRITE_GRAD) {
            const float gs = p.gscale ? *p.gscale : 1.f;
            Ex = p.expect_in[vol * 3 + 0]; Ey = p.expect_in[vol * 3 + 1]; Ez = p.expect_in[vol * 3 + 2];
            gx = p.gcoords[vol * 3 + 0] * gs; gy = p.gcoords[vol * 3 + 1] * gs; gz = p.gcoords[vol * 3 + 2] * gs;
        }
    }

    if constexpr (kHasWrite) {
        // dlogits_i = e_i/L * ( gx (x-Ex)/W + gy (y-Ey)/H + gz (z-Ez)/D )
        // (idx - E) must keep its ABSOLUTE accuracy (see the header): x once per thread in fp64,
        // y and z per row against E split into hi + lo floats.
        const float ML = __fmul_rn(M, kLog2e);
        float ax[VEC];
        float ky = 0.f, kz = 0.f;
        const float Eyh = (float)Ey, Eyl = (float)(Ey - (double)Eyh);
        const float Ezh = (float)Ez, Ezl = (float)(Ez - (double)Ezh);
        if constexpr (kOp == OP_WRITE_PROB) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) ax[j] = invL;
        } else {
            const float kx = gx / (float)p.W * invL;
            ky = gy / (float)p.H * invL;
            kz = gz / (float)p.D * invL;
#pragma unroll
            for (int j = 0; j < VEC; ++j) ax[j] = kx * (float)((double)(x0i + j) - Ex);
        }
        unsigned char* gdst = p.out + vol * p.vol_bytes + (long long)r0 * row_bytes + 16ll * tid;
        const float Hf = (float)p.H, Rf = (float)p.R;
        float yf = (float)y_first, zf = (float)z_first;
        for (int i = 0; i < nst; ++i, ++it) {
            float v[kPasses][VEC];
            if constexpr (kTma) {
                const int slot = it % kStages;
                mbar_wait(smem_u32(&sm.full[slot]), (it / kStages) & 1);
#pragma unroll
                for (int q = 0; q < kPasses; ++q)
                    if (i * kPasses + q < nvis)
                        V16::unpack(lds128(lds_base + slot * kStageBytes + q * pass_bytes), v[q]);
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&sm.empty[slot]));
            } else {
#pragma unroll
                for (int q = 0; q < kPasses; ++q)
                    if (i * kPasses + q < nvis)
                        V16::unpack(ldg_stream(gsrc + (long long)(i * kPasses + q) * pass_bytes), v[q]);
            }
#pragma unroll
            for (int q = 0; q < kPasses; ++q) {
                if (i * kPasses + q < nvis) {
                    float rowterm = 0.f;
                    if constexpr (kOp != OP_WRITE_PROB)
                        rowterm = fmaf(ky, (yf - Eyh) - Eyl, kz * ((zf - Ezh) - Ezl));
                    float o[VEC];
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const float e = ex2(fmaf(v[q][j], kLog2e, -ML));
                        o[j] = e * (ax[j] + rowterm);
                    }
                    stg_stream(gdst + (long long)(i * kPasses + q) * pass_bytes, V16::pack(o));
                }
                yf += Rf;
                while (yf >= Hf) {
                    yf -= Hf;
                    zf += 1.f;
                }
            }
        }
    }

    if constexpr (kHasReduce) {
        if (S > 1) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
}

// ---------------------------------------------------------------------------------------
// Generic scalar kernels for shapes the vector path cannot take (row bytes not a multiple
// of 16, misaligned base, or more than 256 column groups): one CTA per volume, per-element
// index arithmetic.  Correctness path for odd geometries, not tuned.
// ---------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float ld_elem(const unsigned char* base, long long i) {
    return Vec16<T>::to_f32(reinterpret_cast<const T*>(base)[i]);
}

template <typename T, int kOp>
__global__ void __launch_bounds__(256) k_generic(const Params p) {
    constexpr bool kSoftmax = (kOp != OP_REDUCE_PLAIN);
    constexpr bool kHasReduce = (kOp == OP_REDUCE_SOFTMAX || kOp == OP_REDUCE_PLAIN || kOp == OP_FUSED);
    constexpr bool kHasWrite = (kOp == OP_WRITE_GRAD || kOp == OP_WRITE_PROB || kOp == OP_FUSED);
    __shared__ float redm[8];
    __shared__ double red[8][4];
    __shared__ double finE[3];
    __shared__ float fin[5];
    const long long vol = blockIdx.x;
    const long long V = (long long)p.D * p.H * p.W;
    const unsigned char* src = p.in + vol * p.vol_bytes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int HW = p.H * p.W;
    float M = 0.f, L = 1.f, gx = 0.f, gy = 0.f, gz = 0.f;
    double Ex = 0.0, Ey = 0.0, Ez = 0.0;

    if constexpr (kHasReduce) {
        float m = kSoftmax ? kNegInf : 0.f;
        if constexpr (kSoftmax) {
            for (long long i = tid; i < V; i += 256) m = fmaxf(m, ld_elem<T>(src, i));
            m = warp_max(m);
            if (lane == 0) redm[warp] = m;
            __syncthreads();
            m = redm[0];
            for (int w = 1; w < 8; ++w) m = fmaxf(m, redm[w]);
            __syncthreads();
        }
        const float mL = __fmul_rn(m, kLog2e);
        double l = 0.0, sx = 0.0, sy = 0.0, sz = 0.0;
        for (long long i = tid; i < V; i += 256) {
            const float s = ld_elem<T>(src, i);
            const double e = (double)(kSoftmax ? ex2(fmaf(s, kLog2e, -mL)) : s);
            const int z = (int)(i / HW);
            const int rem = (int)(i - (long long)z * HW);
            const int y = rem / p.W;
            const int x = rem - y * p.W;
            l += e;
            sx = fma((double)x, e, sx);
            sy = fma((double)y, e, sy);
            sz = fma((double)z, e, sz);
        }
        l = warp_sum(l); sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
        if (lane == 0) { red[warp][0] = l; red[warp][1] = sx; red[warp][2] = sy; red[warp][3] = sz; }
        __syncthreads();
        if (tid == 0) {
            double fl = 0.0, fx = 0.0, fy = 0.0, fz = 0.0;
            for (int w = 0; w < 8; ++w) { fl += red[w][0]; fx += red[w][1]; fy += red[w][2]; fz += red[w][3]; }
            const double ex_ = kSoftmax ? fx / fl : fx;
            const double ey_ = kSoftmax ? fy / fl : fy;
            const double ez_ = kSoftmax ? fz / fl : fz;
            const float cxn = (float)(ex_ / (double)p.W - 0.5);
            const float cyn = (float)(ey_ / (double)p.H - 0.5);
            const float czn = (float)(ez_ / (double)p.D - 0.5);
            float g[3] = {0.f, 0.f, 0.f};
            if constexpr (kOp == OP_FUSED) {
                const double c[3] = {(double)cxn, (double)cyn, (double)czn};
                for (int k = 0; k < 3; ++k) {
                    const double t = p.tgt_f64 ? reinterpret_cast<const double*>(p.gt)[vol * 3 + k]
                                               : (double)reinterpret_cast<const float*>(p.gt)[vol * 3 + k];
                    const double w = p.tgt_f64 ? reinterpret_cast<const double*>(p.vis)[vol * 3 + k]
                                               : (double)reinterpret_cast<const float*>(p.vis)[vol * 3 + k];
                    const double dlt = c[k] - t;
                    const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
                    g[k] = (float)(sg * w * p.norm);
                    if (p.loss_terms) p.loss_terms[vol * 3 + k] = fabs(dlt) * w;
                }
            }
            if (p.coords) { p.coords[vol * 3 + 0] = cxn; p.coords[vol * 3 + 1] = cyn; p.coords[vol * 3 + 2] = czn; }
            if (p.expect) { p.expect[vol * 3 + 0] = ex_; p.expect[vol * 3 + 1] = ey_; p.expect[vol * 3 + 2] = ez_; }
            if (p.stats) { p.stats[vol * 2 + 0] = m; p.stats[vol * 2 + 1] = (float)fl; }
            finE[0] = ex_; finE[1] = ey_; finE[2] = ez_;
            fin[0] = m; fin[1] = (float)fl; fin[2] = g[0]; fin[3] = g[1]; fin[4] = g[2];
        }
        __syncthreads();
        M = fin[0]; L = fin[1]; Ex = finE[0]; Ey = finE[1]; Ez = finE[2];
        gx = fin[2]; gy = fin[3]; gz = fin[4];
    } else {
        M = p.stats_in[vol * 2 + 0];
        L = p.stats_in[vol * 2 + 1];
        if constexpr (kOp == OP_WRITE_GRAD) {
            const float gs = p.gscale ? *p.gscale : 1.f;
            Ex = p.expect_in[vol * 3 + 0]; Ey = p.expect_in[vol * 3 + 1]; Ez = p.expect_in[vol * 3 + 2];
            gx = p.gcoords[vol * 3 + 0] * gs; gy = p.gcoords[vol * 3 + 1] * gs; gz = p.gcoords[vol * 3 + 2] * gs;
        }
    }

    if constexpr (kHasWrite) {
        const float ML = __fmul_rn(M, kLog2e);
        const float invL = 1.f / L;
        const float kx = gx / (float)p.W * invL, ky = gy / (float)p.H * invL, kz = gz / (float)p.D * invL;
        T* dst = reinterpret_cast<T*>(p.out + vol * p.vol_bytes);
        for (long long i = tid; i < V; i += 256) {
            const float e = ex2(fmaf(ld_elem<T>(src, i), kLog2e, -ML));
            float o;
            if constexpr (kOp == OP_WRITE_PROB) {
                o = e * invL;
            } else {
                const int z = (int)(i / HW);
                const int rem = (int)(i - (long long)z * HW);
                const int y = rem / p.W;
                const int x = rem - y * p.W;
                const float dx = (float)((double)x - Ex), dy = (float)((double)y - Ey), dz = (float)((double)z - Ez);
                o = e * (kx * dx + fmaf(ky, dy, kz * dz));
            }
            dst[i] = Vec16<T>::from_f32(o);
        }
    }
}

// ---------------------------------------------------------------------------------------
// Host-side launch
// ---------------------------------------------------------------------------------------
static int check_device() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return (int)e;
    return major == 10 ? 0 : HIPE_E_ARCH;
}

static int pick_cluster(long long n_vol, long long vol_bytes, int NR) {
    // Aim for >= ~6 resident-CTA waves worth of work items over 148 SMs while keeping every
    // CTA's share >= 64 KB so the per-CTA prologue/reduction stays a small fraction.
    int S = 1;
    while (S < kMaxCluster && n_vol * S < 148ll * 24 && vol_bytes / (2 * S) >= 65536 && NR / (2 * S) >= 1) S *= 2;
    return S;
}

template <typename T, bool kTma, int kOp>
static int launch_stream(const Params& p, int S, cudaStream_t st) {
    auto kern = k_stream<T, kTma, kOp>;
    const size_t smem = sizeof(SmemLayout) + (kTma ? (size_t)kStages * kStageBytes : 0);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(p.n_vol * S), 1, 1);
    cfg.blockDim = dim3(kTma ? kConsumers + 32 : kConsumers, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, p);
    return (int)e;
}

// Default L2 policies: a forward is followed by a backward that walks the volumes in reverse,
// so its lines are left at normal priority; write-only passes never need their input again;
// the fused launch re-reads pass 1 from L2.  HIPE_L2_POLICY="a,b" overrides (tuning only).
static void pick_policy(int op, int* p1, int* p2) {
    *p1 = 0; *p2 = 1;
    if (op == OP_WRITE_GRAD || op == OP_WRITE_PROB) *p1 = 1;
    if (op == OP_FUSED) { *p1 = 2; *p2 = 1; }
    const char* e = getenv("HIPE_L2_POLICY");
    if (e && e[0] >= '0' && e[0] <= '2' && e[1] == ',' && e[2] >= '0' && e[2] <= '2') {
        *p1 = e[0] - '0';
        *p2 = e[2] - '0';
    }
}

template <typename T, int kOp>
static int dispatch(Params p, int cluster, int variant, cudaStream_t st) {
    const int esz = (int)sizeof(T);
    pick_policy(kOp, &p.pol1, &p.pol2);
    const long long row_b = (long long)p.W * esz;
    const bool vec_ok = (row_b % 16 == 0) && (row_b / 16 <= kConsumers) &&
                        ((reinterpret_cast<uintptr_t>(p.in) & 15) == 0) &&
                        (!p.out || (reinterpret_cast<uintptr_t>(p.out) & 15) == 0) &&
                        ((long long)p.D * p.H < (1ll << 30));
    p.vol_bytes = (long long)p.D * p.H * p.W * esz;
    p.NR = p.D * p.H;
    if (!vec_ok) {
        k_generic<T, kOp><<<(unsigned)p.n_vol, 256, 0, st>>>(p);
        return (int)cudaGetLastError();
    }
    p.G = (int)(row_b / 16);
    p.R = kConsumers / p.G;
    int S = cluster ? cluster : pick_cluster(p.n_vol, p.vol_bytes, p.NR);
    while (S > 1 && p.NR < S) S /= 2;
    const bool tma = (variant == HIPE_VARIANT_AUTO) ? true : (variant == HIPE_VARIANT_TMA);
    return tma ? launch_stream<T, true, kOp>(p, S, st) : launch_stream<T, false, kOp>(p, S, st);
}

static int check_geom(const void* in, int dtype, long long n_vol, int D, int H, int W, int cluster, int variant) {
    if (!in && n_vol != 0) return HIPE_E_NULL;
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (n_vol < 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    if ((long long)D * H * W > (1ll << 31)) return HIPE_E_SHAPE;
    if (n_vol * kMaxCluster > 0x7fffffffll) return HIPE_E_SHAPE;
    if (cluster != 0 && cluster != 1 && cluster != 2 && cluster != 4 && cluster != 8) return HIPE_E_CLUSTER;
    if (variant < 0 || variant > 2) return HIPE_E_SHAPE;
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    if (reinterpret_cast<uintptr_t>(in) % esz) return HIPE_E_ALIGN;
    return check_device();
}

template <int kOp>
static int dispatch_dtype(const Params& p, int dtype, int cluster, int variant, cudaStream_t st) {
    if (p.n_vol == 0) return 0;
    return dtype == HIPE_F32 ? dispatch<float, kOp>(p, cluster, variant, st)
                             : dispatch<__nv_bfloat16, kOp>(p, cluster, variant, st);
}

}  // namespace hipe

using namespace hipe;

extern "C" int hipe_softargmax_fwd(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                   int apply_softmax, float* coords, double* expect, float* stats,
                                   int cluster, int variant, void* stream) {
    int rc = check_geom(logits, dtype, n_vol, D, H, W, cluster, variant);
    if (rc) return rc;
    if (n_vol == 0) return 0;
    Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.n_vol = n_vol; p.D = D; p.H = H; p.W = W;
    p.coords = coords; p.expect = expect; p.stats = stats;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return apply_softmax ? dispatch_dtype<OP_REDUCE_SOFTMAX>(p, dtype, cluster, variant, st)
                         : dispatch_dtype<OP_REDUCE_PLAIN>(p, dtype, cluster, variant, st);
}

extern "C" int hipe_softargmax_bwd(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                   const float* stats, const double* expect, const float* gcoords,
                                   const float* gscale, void* dlogits, int cluster, int variant,
                                   void* stream) {
    int rc = check_geom(logits, dtype, n_vol, D, H, W, cluster, variant);
    if (rc) return rc;
    if (n_vol == 0) return 0;
    if (!stats || !expect || !gcoords || !dlogits) return HIPE_E_NULL;
    Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.out = static_cast<unsigned char*>(dlogits);
    p.n_vol = n_vol; p.D = D; p.H = H; p.W = W;
    p.stats_in = stats; p.expect_in = expect; p.gcoords = gcoords; p.gscale = gscale;
    p.reverse = 1;  // the forward's last volumes are the ones still in L2
    return dispatch_dtype<OP_WRITE_GRAD>(p, dtype, cluster, variant, static_cast<cudaStream_t>(stream));
}

extern "C" int hipe_volume_softmax(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                   const float* stats, void* probs, int cluster, int variant,
                                   void* stream) {
    int rc = check_geom(logits, dtype, n_vol, D, H, W, cluster, variant);
    if (rc) return rc;
    if (n_vol == 0) return 0;
    if (!stats || !probs) return HIPE_E_NULL;
    Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.out = static_cast<unsigned char*>(probs);
    p.n_vol = n_vol; p.D = D; p.H = H; p.W = W;
    p.stats_in = stats;
    p.reverse = 1;
    return dispatch_dtype<OP_WRITE_PROB>(p, dtype, cluster, variant, static_cast<cudaStream_t>(stream));
}

extern "C" int hipe_joint_loss_fused(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                     const void* gt, const void* vis, int tgt_dtype, double norm,
                                     float* coords, double* expect, float* stats, double* loss_terms,
                                     void* dlogits, int cluster, int variant, void* stream) {
    int rc = check_geom(logits, dtype, n_vol, D, H, W, cluster, variant);
    if (rc) return rc;
    if (n_vol == 0) return 0;
    if (!gt || !vis || !dlogits) return HIPE_E_NULL;
    if (tgt_dtype != HIPE_F32 && tgt_dtype != HIPE_F64) return HIPE_E_DTYPE;
    Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.out = static_cast<unsigned char*>(dlogits);
    p.n_vol = n_vol; p.D = D; p.H = H; p.W = W;
    p.coords = coords; p.expect = expect; p.stats = stats;
    p.gt = gt; p.vis = vis; p.tgt_f64 = (tgt_dtype == HIPE_F64); p.norm = norm; p.loss_terms = loss_terms;
    return dispatch_dtype<OP_FUSED>(p, dtype, cluster, variant, static_cast<cudaStream_t>(stream));
}

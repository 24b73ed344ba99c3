// JointLocationLoss forward + backward in ONE persistent launch that reads every heatmap volume
// from HBM exactly once and writes its gradient exactly once (sm_100a).
//
// Replaces common/nets/loss.py:74-91 followed by loss.backward() (main/train.py:113) of the
// reference.  The two-launch path (hipe_softargmax_fwd + hipe_softargmax_bwd) moves
// 2 V s + V g bytes per volume; this launch moves V s + V g: the volume stays ON CHIP between
// the reduction and the gradient write.
//
// A 64^3 fp32 volume is 1 MiB, more than one SM's shared memory, so a volume is cut into P
// ITEMS of <= 4 stages (a stage = SR whole rows <= 16 KB) and the items of one volume are
// taken by P different CTAs of a persistent grid (one CTA per SM, cooperative launch so that
// all of them are co-resident).  Per CTA:
//
//   producer warp   streams the CTA's items, stage by stage, into a 13-slot x 16 KB ring with
//                   1-D TMA bulk copies (evict_first: HBM lines are never needed again);
//   reduce team     (256 threads) REDUCE a stage: online-max softmax; the exponentials are
//                   written back over the logits in the ring slot ("stash", fp32 only) together with the offset
//                   they are relative to, so the gradient pass needs no second MUFU.EX2;
//                   per-thread fp32 accumulators per owned (row-in-stage, column) position;
//                   PUBLISH the item's partial sums to a global workspace and bump the
//                   volume's arrival counter (release);
//   write team      (256 threads)
//                   WAIT until the counter shows all P partials (acquire), merge them in a
//                   fixed order (every CTA of the volume computes bit-identical scalars), form
//                   the L1 gradient sign(c-gt)*vis*norm (loss.py:86-89);
//                   WRITE: dlogits = e*2^(off-M)/L * (gx(x-Ex)/W + gy(y-Ey)/H + gz(z-Ez)/D)
//                   from the stashed ring slot, then release the slot to the producer.
//
// REDUCE/PUBLISH and WAIT/WRITE are done by two different 256-thread teams of the CTA working
// on consecutive items through the same ring: the reduce team runs up to (13 - n) stages (2-3 items) ahead
// of the write team, so the cross-CTA latency (global flag round trip, ~1 us) and each team's
// per-item serial work are covered by the other team's streaming and the ring always has loads
// in flight.  A reduce team never waits on another CTA, a write team only waits on volumes
// whose items were all published by reduce teams that needed strictly earlier volumes to
// drain: with all CTAs co-resident (cooperative launch) the schedule cannot deadlock.
//
// Precision (same contract as hipe_stream_kernel.cuh): exponentials are fp32; each expectation
// is a ratio of fp64 sums built from the SAME rounded fp32 atoms, so rounding only re-weights
// voxels.  Two families of atoms are used, each with its own denominator:
//   c[q][j]  per-thread accumulators over the item's stages for a fixed (row-in-stage, column)
//            -> S0c, Sx, and the row-offset parts Syo, Szo of the y/z moments;
//   t_s      per-thread per-stage sums -> S0t and the stage-base parts Syb, Szb.
//   Ex = Sx/S0c,  Ey = Syo/S0c + Syb/S0t,  Ez = Szo/S0c + Szb/S0t.
#include "hipe_common.cuh"
#include <stdlib.h>

namespace hipe {
namespace res {

constexpr int kC = 256;               // consumer threads
constexpr int kNW = kC / 32;          // consumer warps
constexpr int kSlots = 13;            // ring depth
constexpr int kSlotBytes = 16384;     // ring slot
constexpr int kKP = 4;                // 16-byte vectors per thread per stage
constexpr int kMaxItemStages = 4;     // an item never exceeds 4 stages (64 KB): >= 2 whole items of look-ahead in the ring
constexpr int kRD = 16;               // reduce->write hand-off barriers (the ring never holds more than 13 items)
constexpr int kMaxRtRows = 1024;      // rows of one item (<= 4 stages x 256 rows)
constexpr int kNSums = 7;             // S0c Sx Syo Szo S0t Syb Szb
constexpr unsigned kSpinLimit = 1u << 22;

struct Params {
    const unsigned char* in;
    unsigned char* out;
    long long n_vol;
    long long vol_bytes;
    long long n_items;
    int W, H, D;
    int G;            // 16-byte column groups per row
    int R;            // rows per visit of the CTA
    int SR;           // rows per stage (= kKP * R); SR | H or H | SR
    int stage_bytes;  // SR * row bytes (<= 16 KB)
    int pass_bytes;   // R * row bytes
    int nact;         // active consumer threads (= R * G)
    int NSv;          // stages per volume
    int P;            // items per volume
    // outputs (nullable except out)
    float* coords;
    double* expect;
    float* stats;
    double* loss_terms;
    // targets
    const void* gt;
    const void* vis;
    int tgt_f64;
    double norm;
    // workspace
    double* partials;     // n_vol * P * 8 doubles
    unsigned* counters;   // n_vol
    unsigned* err;        // 1 word: set when a spin limit was hit
};

struct Smem {
    float mls[kSlots][kC];        // offset (in log2 units) the stashed exponentials of (slot, thread) are relative to
    float rt[kMaxRtRows];         // per-row gradient term of the item being written
    unsigned long long full[kSlots];
    unsigned long long empty[kSlots];
    unsigned long long red_done[kRD];  // reduce team -> write team: every stage of the item is stashed
    double wred[kNW][8];          // reduce team scratch
    float wmax[kNW];
    double finE[3];               // write team: Ex, Ey, Ez
    double finK[3];               //             g_a / dim_a / L per axis
    float finM;                   //             volume maximum
};

struct Item {
    long long vol;
    int i;      // item index inside the volume
    int sv0;    // first stage (volume-relative)
    int n;      // stages
};

__device__ __forceinline__ Item decode(const Params& p, long long w) {
    Item it;
    it.vol = w / p.P;
    it.i = (int)(w - it.vol * p.P);
    it.sv0 = (int)(((long long)p.NSv * it.i) / p.P);
    it.n = (int)(((long long)p.NSv * (it.i + 1)) / p.P) - it.sv0;
    return it;
}

template <int kId>
__device__ __forceinline__ void team_barrier() {
    asm volatile("bar.sync %0, %1;" ::"n"(kId), "n"(kC) : "memory");
}

// mbarrier wait that traps instead of hanging the GPU if the pipeline logic is ever wrong
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (unsigned spin = 0; spin < (1u << 26); ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}

__device__ __forceinline__ void sts128(uint32_t saddr, const uint4& v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float rebase_w(float m_from, float m_to) {
    // 2^(fl(m_from*log2e) - fl(m_to*log2e)): the same rounded products the exponentials used
    return (m_from == kNegInf) ? 0.f : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

// Thread roles inside the CTA (kThreads = 2*kC + 32):
//   [0, kC)        REDUCE team: online softmax + stash + publish of item k+1
//   [kC, 2kC)      WRITE team : flag wait + scalars + gradient write of item k, releases slots
//   [2kC, 2kC+32)  producer warp (one elected lane)
// The reduce team runs at most 13 - n stages ahead of the write team (ring capacity), which is
// what hides the cross-CTA flag latency and the per-item serial work of either team.
template <typename T, bool kStash>
__global__ void __launch_bounds__(2 * kC + 32, 1) k_resident(const Params p) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw + kSlots * kSlotBytes);
    const uint32_t ring = smem_u32(smem_raw);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kSlots; ++s) {
            mbar_init(smem_u32(&sm.full[s]), 1);
            mbar_init(smem_u32(&sm.empty[s]), kNW);
        }
#pragma unroll
        for (int s = 0; s < kRD; ++s) mbar_init(smem_u32(&sm.red_done[s]), kNW);
        mbar_fence_init();
        fence_proxy_async();
    }
    __syncthreads();

    if (threadIdx.x >= 2 * kC) {
        // ------------------------------ producer warp ------------------------------
        if (threadIdx.x == 2 * kC) {
            const uint64_t pol = policy_evict_first();
            unsigned g = 0;
            for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x) {
                const Item it = decode(p, w);
                const unsigned char* src = p.in + it.vol * p.vol_bytes + (long long)it.sv0 * p.stage_bytes;
                for (int s = 0; s < it.n; ++s, ++g) {
                    const unsigned slot = g % kSlots;
                    if (g >= (unsigned)kSlots) mbar_wait_bounded(smem_u32(&sm.empty[slot]), ((g / kSlots) - 1) & 1);
                    const uint32_t bar = smem_u32(&sm.full[slot]);
                    mbar_arrive_expect_tx(bar, (uint32_t)p.stage_bytes);
                    tma_load_1d(ring + slot * kSlotBytes, src + (long long)s * p.stage_bytes, (uint32_t)p.stage_bytes,
                                bar, pol);
                }
            }
        }
        return;
    }

    const bool is_writer = threadIdx.x >= kC;
    const int tid = is_writer ? threadIdx.x - kC : threadIdx.x;   // team-local
    const int lane = tid & 31, warp = tid >> 5;
    const bool active = tid < p.nact;
    const int cg = active ? tid % p.G : 0;
    const int rr = active ? tid / p.G : 0;
    const int x0 = cg * VEC;
    const uint32_t lds_base = ring + 16u * tid;

    if (!is_writer) {
        // =============================== REDUCE team ===============================
        unsigned g = 0, item_no = 0;
        // row offsets of this thread's kKP visits inside a stage
        int yo[kKP], zo[kKP];
#pragma unroll
        for (int q = 0; q < kKP; ++q) {
            const int rho = rr + q * p.R;
            zo[q] = rho / p.H;
            yo[q] = rho - zo[q] * p.H;
        }
        for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
            const Item it = decode(p, w);
            float m = kNegInf, mL = kNegInf;
            float c[kKP][VEC];
#pragma unroll
            for (int q = 0; q < kKP; ++q)
#pragma unroll
                for (int j = 0; j < VEC; ++j) c[q][j] = 0.f;
            double s0t = 0.0, syb = 0.0, szb = 0.0;
            int rowbase = it.sv0 * p.SR;
            int zb = rowbase / p.H;
            int yb = rowbase - zb * p.H;
            for (int s = 0; s < it.n; ++s, ++g) {
                const unsigned slot = g % kSlots;
                mbar_wait_bounded(smem_u32(&sm.full[slot]), (g / kSlots) & 1);
                if (active) {
                    const uint32_t sbase = lds_base + slot * kSlotBytes;
                    float v[kKP][VEC];
#pragma unroll
                    for (int q = 0; q < kKP; ++q) V16::unpack(lds128(sbase + q * p.pass_bytes), v[q]);
                    float smax = kNegInf;
#pragma unroll
                    for (int q = 0; q < kKP; ++q)
#pragma unroll
                        for (int j = 0; j < VEC; ++j) smax = fmaxf(smax, v[q][j]);
                    if (smax > m) {
                        const float f = rebase_w(m, smax);
#pragma unroll
                        for (int q = 0; q < kKP; ++q)
#pragma unroll
                            for (int j = 0; j < VEC; ++j) c[q][j] *= f;
                        const double fd = (double)f;
                        s0t *= fd; syb *= fd; szb *= fd;
                        m = smax;
                        mL = __fmul_rn(smax, kLog2e);
                    }
                    // all -inf so far: exponentiate against 0 so that e = 2^-inf = 0 (no NaN)
                    const float off = (m == kNegInf) ? 0.f : mL;
                    float rowsum[kKP];
#pragma unroll
                    for (int q = 0; q < kKP; ++q) {
                        float e[VEC];
#pragma unroll
                        for (int j = 0; j < VEC; ++j) {
                            e[j] = ex2(fmaf(v[q][j], kLog2e, -off));
                            c[q][j] += e[j];
                        }
                        if constexpr (kStash)
                            sts128(sbase + q * p.pass_bytes, make_uint4(__float_as_uint(e[0]), __float_as_uint(e[1]),
                                                                        __float_as_uint(e[2]), __float_as_uint(e[3])));
                        if constexpr (VEC == 4)
                            rowsum[q] = (e[0] + e[1]) + (e[2] + e[3]);
                        else
                            rowsum[q] = ((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[VEC - 1]));
                    }
                    if constexpr (kStash) sm.mls[slot][tid] = mL;
                    const double dt = (double)((rowsum[0] + rowsum[1]) + (rowsum[2] + rowsum[3]));
                    s0t += dt;
                    syb = fma((double)yb, dt, syb);
                    szb = fma((double)zb, dt, szb);
                }
                // next stage's base row (uniform): SR | H or H | SR, so one of the two moves
                yb += p.SR;
                while (yb >= p.H) { yb -= p.H; ++zb; }
            }
            // the item is stashed: hand it to the write team (release of this warp's smem writes)
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.red_done[item_no % kRD]));
            // ---- fold the thread's accumulators, merge over the team, publish the item's partial ----
            double q7[kNSums];
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q7[k] = 0.0;
            float pmx = kNegInf;
            if (active) {
                pmx = m;
                double col[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) col[j] = 0.0;
#pragma unroll
                for (int q = 0; q < kKP; ++q) {
                    double row = 0.0;
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const double cd = (double)c[q][j];
                        row += cd;
                        col[j] += cd;
                    }
                    q7[2] = fma((double)yo[q], row, q7[2]);
                    q7[3] = fma((double)zo[q], row, q7[3]);
                }
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    q7[0] += col[j];
                    q7[1] = fma((double)(x0 + j), col[j], q7[1]);
                }
                q7[4] = s0t; q7[5] = syb; q7[6] = szb;
            }
            const float wm = warp_max(pmx);
            const double f = (double)rebase_w(pmx, wm);
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q7[k] = warp_sum(q7[k] * f);
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < kNSums; ++k) sm.wred[warp][k] = q7[k];
                sm.wmax[warp] = wm;
            }
            team_barrier<1>();
            if (warp == 0) {
                const float pm = (lane < kNW) ? sm.wmax[lane] : kNegInf;
                double a[kNSums];
#pragma unroll
                for (int k = 0; k < kNSums; ++k) a[k] = (lane < kNW) ? sm.wred[lane][k] : 0.0;
                float cm = pm;
#pragma unroll
                for (int o = kNW / 2; o > 0; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
                const double f2 = (double)rebase_w(pm, cm);
#pragma unroll
                for (int k = 0; k < kNSums; ++k) {
                    a[k] *= f2;
#pragma unroll
                    for (int o = kNW / 2; o > 0; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
                }
                if (lane == 0) {
                    double* dst = p.partials + (it.vol * p.P + it.i) * 8;
#pragma unroll
                    for (int k = 0; k < kNSums; ++k) __stcg(dst + k, a[k]);
                    __stcg(dst + 7, (double)cm);
                    __threadfence();
                    atomicAdd(p.counters + it.vol, 1u);
                }
            }
            team_barrier<1>();   // wred / wmax are free again
        }
        return;
    }

    // ================================ WRITE team ================================
    unsigned g = 0;
    unsigned item_no = 0;
    for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
        const Item it = decode(p, w);
        if (warp == 0) {
            // targets first: their latency overlaps the flag wait
            double tk = 0.0, wk = 0.0;
            if (lane < 3) {
                tk = p.tgt_f64 ? reinterpret_cast<const double*>(p.gt)[it.vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.gt)[it.vol * 3 + lane];
                wk = p.tgt_f64 ? reinterpret_cast<const double*>(p.vis)[it.vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.vis)[it.vol * 3 + lane];
            }
            unsigned spins = 0;
            while (ld_acquire_u32(p.counters + it.vol) < (unsigned)p.P) {
                __nanosleep(20);
                if (++spins > kSpinLimit) {
                    if (lane == 0) atomicExch(p.err, 1u);
                    break;
                }
            }
            float mi = kNegInf;
            double a[kNSums];
#pragma unroll
            for (int k = 0; k < kNSums; ++k) a[k] = 0.0;
            for (int i = lane; i < p.P; i += 32) {
                const double* src = p.partials + (it.vol * p.P + i) * 8;
                double b[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) b[k] = __ldcg(src + k);
                const float pm = (float)b[7];
                if (pm > mi) {
                    const double fr = (double)rebase_w(mi, pm);
#pragma unroll
                    for (int k = 0; k < kNSums; ++k) a[k] *= fr;
                    mi = pm;
                }
                const double fr = (double)rebase_w(pm, mi);
#pragma unroll
                for (int k = 0; k < kNSums; ++k) a[k] = fma(fr, b[k], a[k]);
            }
            const float M = warp_max(mi);
            const double fr = (double)rebase_w(mi, M);
#pragma unroll
            for (int k = 0; k < kNSums; ++k) a[k] = warp_sum(a[k] * fr);
            // loss.py:39-41 (expectations) and :54-56 (normalisation), in fp64, rounded once.
            // Lane k (< 3) finishes axis k: x, y, z.
            const double ic = 1.0 / a[0], itt = 1.0 / a[4];
            const double num_c = lane == 0 ? a[1] : (lane == 1 ? a[2] : a[3]);
            const double num_t = lane == 0 ? 0.0 : (lane == 1 ? a[5] : a[6]);
            const double dim = lane == 0 ? (double)p.W : (lane == 1 ? (double)p.H : (double)p.D);
            const double ex_ = num_c * ic + (lane == 0 ? 0.0 : num_t * itt);
            const float cn = (float)(ex_ / dim - 0.5);
            // loss.py:86-89 on this volume's three coordinates, in the targets' precision
            const double dlt = (double)cn - tk;
            const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
            const float g1 = (float)(sg * wk * p.norm);
            if (lane < 3) {
                if (it.i == 0) {
                    if (p.loss_terms) p.loss_terms[it.vol * 3 + lane] = fabs(dlt) * wk;
                    if (p.coords) p.coords[it.vol * 3 + lane] = cn;
                    if (p.expect) p.expect[it.vol * 3 + lane] = ex_;
                }
                sm.finE[lane] = ex_;
                sm.finK[lane] = (double)g1 / dim * ic;     // g_a / dim_a / L
            }
            if (lane == 0) {
                sm.finM = M;
                if (it.i == 0 && p.stats) { p.stats[it.vol * 2 + 0] = M; p.stats[it.vol * 2 + 1] = (float)a[0]; }
            }
        }
        team_barrier<2>();
        const double kxd = sm.finK[0], kyd = sm.finK[1], kzd = sm.finK[2];
        const double Ex = sm.finE[0], Ey = sm.finE[1], Ez = sm.finE[2];
        const float ML = __fmul_rn(sm.finM, kLog2e);
        float ax[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) ax[j] = (float)(kxd * ((double)(x0 + j) - Ex));
        {
            // per-row term: (idx - E) keeps its absolute accuracy, formed in fp64, rounded once
            const int nrows = it.n * p.SR;
            const int row0 = it.sv0 * p.SR;
            for (int lr = tid; lr < nrows; lr += kC) {
                const int row = row0 + lr;
                const int z = row / p.H;
                const int y = row - z * p.H;
                sm.rt[lr] = (float)(kyd * ((double)y - Ey) + kzd * ((double)z - Ez));
            }
        }
        team_barrier<2>();   // rt visible; fin* may be overwritten after the NEXT barrier pair only
        // the reduce team has stashed every stage of this item
        mbar_wait_bounded(smem_u32(&sm.red_done[item_no % kRD]), (item_no / kRD) & 1);

        unsigned char* gdst = p.out + it.vol * p.vol_bytes + (long long)it.sv0 * p.stage_bytes + 16ll * tid;
        for (int s = 0; s < it.n; ++s, ++g) {
            const unsigned slot = g % kSlots;
            if (active) {
                const uint32_t sbase = lds_base + slot * kSlotBytes;
                float f = 1.f;
                if constexpr (kStash) f = ex2(__fsub_rn(sm.mls[slot][tid], ML));
                float fax[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) fax[j] = f * ax[j];
                uint4 raw[kKP];
#pragma unroll
                for (int q = 0; q < kKP; ++q) raw[q] = lds128(sbase + q * p.pass_bytes);
#pragma unroll
                for (int q = 0; q < kKP; ++q) {
                    const float frt = f * sm.rt[s * p.SR + rr + q * p.R];
                    float v[VEC], o[VEC];
                    V16::unpack(raw[q], v);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const float e = kStash ? v[j] : ex2(fmaf(v[j], kLog2e, -ML));
                        o[j] = e * (fax[j] + frt);
                    }
                    stg_stream(gdst + (long long)s * p.stage_bytes + q * p.pass_bytes, V16::pack(o));
                }
            }
            __syncwarp();
            if (lane == 0) {
                // the slot was written through the generic proxy (stash): order that before the
                // producer's next bulk copy (async proxy) into it
                fence_proxy_async();
                mbar_arrive(smem_u32(&sm.empty[slot]));
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct Geom {
    int G, R, SR, stage_bytes, pass_bytes, nact, NSv, P;
};

static bool pick_geom(int dtype, int D, int H, int W, Geom* g) {
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    const long long row_b = (long long)W * esz;
    if (row_b % 16 != 0 || row_b / 16 > kC) return false;
    const int G = (int)(row_b / 16);
    const long long NR = (long long)D * H;
    if (NR >= (1ll << 30)) return false;
    for (int R = kC / G; R >= 1; --R) {
        const int SR = kKP * R;
        if (SR > kMaxRtRows / kMaxItemStages) continue;
        if (2 * R * G < kC) break;   // fewer than half the consumers would work
        const bool ok = (SR % H == 0 && NR % SR == 0) || (H % SR == 0);
        if (!ok) continue;
        g->G = G; g->R = R; g->SR = SR;
        g->pass_bytes = (int)(R * row_b);
        g->stage_bytes = (int)(SR * row_b);
        g->nact = R * G;
        g->NSv = (int)(NR / SR);
        g->P = (g->NSv + kMaxItemStages - 1) / kMaxItemStages;
        return true;
    }
    return false;
}

static size_t ws_counters_bytes(long long n_vol) {
    return (size_t)(((n_vol + 1) * sizeof(unsigned) + 255) / 256 * 256);   // counters + the error word
}

}  // namespace res
}  // namespace hipe

using namespace hipe;

extern "C" long long hipe_resident_workspace_bytes(int dtype, long long n_vol, int D, int H, int W) {
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return 0;
    if (n_vol <= 0 || D <= 0 || H <= 0 || W <= 0) return 0;
    res::Geom g;
    if (!res::pick_geom(dtype, D, H, W, &g)) return 0;
    return (long long)(res::ws_counters_bytes(n_vol) + (size_t)n_vol * g.P * 8 * sizeof(double));
}

extern "C" int hipe_joint_loss_resident(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                        const void* gt, const void* vis, int tgt_dtype, double norm,
                                        float* coords, double* expect, float* stats, double* loss_terms,
                                        void* dlogits, void* workspace, long long workspace_bytes, void* stream) {
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (tgt_dtype != HIPE_F32 && tgt_dtype != HIPE_F64) return HIPE_E_DTYPE;
    if (n_vol < 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    if (n_vol == 0) return 0;
    if (!logits || !gt || !vis || !dlogits || !workspace) return HIPE_E_NULL;
    if ((reinterpret_cast<uintptr_t>(logits) & 15) || (reinterpret_cast<uintptr_t>(dlogits) & 15) ||
        (reinterpret_cast<uintptr_t>(workspace) & 255))
        return HIPE_E_ALIGN;
    res::Geom g;
    if (!res::pick_geom(dtype, D, H, W, &g)) return HIPE_E_UNSUPPORTED;
    const long long need = hipe_resident_workspace_bytes(dtype, n_vol, D, H, W);
    if (workspace_bytes < need) return HIPE_E_SHAPE;

    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    int major = 0, sms = 0, coop = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (major != 10) return HIPE_E_ARCH;
    if (!coop) return HIPE_E_UNSUPPORTED;

    const bool stash = dtype == HIPE_F32;
    const void* kern = stash ? (const void*)res::k_resident<float, true>
                             : (const void*)res::k_resident<__nv_bfloat16, false>;
    const size_t smem = (size_t)res::kSlots * res::kSlotBytes + sizeof(res::Smem);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 2 * res::kC + 32, smem);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return HIPE_E_UNSUPPORTED;
    long long grid = (long long)sms;   // one CTA per SM: the ring takes the whole shared memory
    if (const char* env = getenv("HIPE_RESIDENT_GRID")) {
        const long long v = atoll(env);
        if (v > 0 && v < grid) grid = v;
    }
    if (grid < g.P) return HIPE_E_UNSUPPORTED;   // the items of one volume must sit on distinct CTAs

    res::Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.out = static_cast<unsigned char*>(dlogits);
    p.n_vol = n_vol;
    p.vol_bytes = (long long)D * H * W * (dtype == HIPE_F32 ? 4 : 2);
    p.n_items = n_vol * g.P;
    p.W = W; p.H = H; p.D = D;
    p.G = g.G; p.R = g.R; p.SR = g.SR; p.stage_bytes = g.stage_bytes; p.pass_bytes = g.pass_bytes;
    p.nact = g.nact; p.NSv = g.NSv; p.P = g.P;
    p.coords = coords; p.expect = expect; p.stats = stats; p.loss_terms = loss_terms;
    p.gt = gt; p.vis = vis; p.tgt_f64 = (tgt_dtype == HIPE_F64); p.norm = norm;
    const size_t cb = res::ws_counters_bytes(n_vol);
    p.counters = static_cast<unsigned*>(workspace);
    p.err = p.counters + n_vol;
    p.partials = reinterpret_cast<double*>(static_cast<unsigned char*>(workspace) + cb);
    if (grid > p.n_items) grid = p.n_items;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    e = cudaMemsetAsync(workspace, 0, cb, st);
    if (e != cudaSuccess) return (int)e;
    void* args[] = {(void*)&p};
    e = cudaLaunchCooperativeKernel(kern, dim3((unsigned)grid), dim3(2 * res::kC + 32), args, smem, st);
    if (e == cudaErrorCooperativeLaunchTooLarge) {
        (void)cudaGetLastError();
        return HIPE_E_UNSUPPORTED;
    }
    return (int)e;
}

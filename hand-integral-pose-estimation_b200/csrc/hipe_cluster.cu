// JointLocationLoss forward + backward in ONE persistent launch of thread-block CLUSTERS that keep
// each heatmap volume in (distributed) shared memory between the reduction and the gradient
// write: HBM sees every volume once on the way in and its gradient once on the way out, and the
// L2 is not asked to hold anything (sm_100a).
//
// Replaces common/nets/loss.py:74-91 followed by loss.backward() (main/train.py:113) of the
// reference, like hipe_joint_loss_resident / hipe_joint_loss_fused, for every geometry whose
// volume fits the shared memory of one cluster (fp32 64^3 = 1 MiB on 8 CTAs, bf16 64^3 = 512 KiB
// on 4, fp32 56^3 = 686 KiB on 4 or 8, ...).
//
// A cluster of C CTAs owns one volume at a time; CTA c owns the stage range
// [NSv*c/C, NSv*(c+1)/C) of the volume's NSv stages (a stage = SR whole rows <= 16 KB, the same
// tiling as hipe_resident.cu).  The cluster walks the batch: volume = cluster + k * n_clusters.
// Every CTA runs four roles on one ring of `nslots` shared-memory slots:
//
//   producer warp  1-D TMA bulk copies (cp.async.bulk, L2 evict_first) of the CTA's stages into
//                  the ring, as far ahead as free slots allow -- the next volumes' stages stream
//                  in while the current one is still being written;
//   REDUCE team    256 threads: online-max softmax over each stage as it lands; does NOT free the
//                  slot.  Per volume: fold to fp64, merge over the team, hand to the merge warp;
//   merge warp     stores the CTA's partial into every peer's shared memory (st.shared::cluster)
//                  and arrives on the peer's mbarrier (release.cluster); waits for the C partials
//                  in its own shared memory (acquire.cluster), merges them in rank order (every
//                  CTA of the cluster gets bit-identical scalars), finishes the coordinates and
//                  the L1 gradient sign(c-gt)*vis*norm (loss.py:86-89);
//   WRITE team     256 threads: walks the SAME slots again (still resident: no second read from
//                  HBM or L2), writes dlogits = e/L * (gx(x-Ex)/W + gy(y-Ey)/H + gz(z-Ez)/D),
//                  and only then hands the slot back to the producer.
//
// The reduce team runs up to (nslots - stages per share) stages ahead of the write team, so the
// cross-CTA merge of volume n+1 happens under the gradient write of volume n; no role ever waits
// on global memory, and clusters do not depend on each other (no cooperative launch, no
// co-residency requirement).
//
// Exchange-buffer safety (two buffers, xch[k & 1]): the merge warp is sequential, so before it
// publishes volume k it has completed the merge of k-1, i.e. received every peer's partial of
// k-1, which each peer sent only after it had read the partials of k-2 -- the buffer being
// overwritten.  The same chain shows that a peer's arrive for volume k+2 cannot reach an
// mbarrier whose phase for volume k is still open.
//
// Precision: identical contract to hipe_resident.cu (fp32 exponentials; every expectation a
// ratio of fp64 sums built from the same rounded fp32 atoms; two atom families with their own
// denominators: Ex = Sx/S0c, Ey = Syo/S0c + Syb/S0t, Ez = Szo/S0c + Szb/S0t).
#include "hipe_common.cuh"
#include <stdio.h>
#include <stdlib.h>

namespace hipe {
namespace clu {

constexpr int kC = 256;               // threads per team
constexpr int kNW = kC / 32;          // warps per team
constexpr int kKP = 4;                // 16-byte vectors per thread per stage
constexpr int kMaxSlots = 24;
constexpr int kMaxCluster = 8;        // portable cluster size
constexpr int kNSums = 7;             // S0c Sx Syo Szo S0t Syb Szb
constexpr int kNB = 4;                // scalar records in flight (merge warp -> write team)
constexpr int kPB = 2;                // team partials in flight (reduce team -> merge warp)
constexpr int kThreads = 2 * kC + 64; // two teams + producer warp + merge warp
constexpr float kLazy = 16.f;         // lazy re-basing threshold of the online softmax
constexpr int kSmemMax = 232448;      // 227 KB: the opt-in dynamic shared memory of one CTA
constexpr int kMinSpare = 3;          // slots beyond one share the ring must have (look-ahead)

struct Params {
    const unsigned char* in;
    unsigned char* out;
    long long n_vol;
    long long vol_bytes;
    int W, H, D;
    int G;            // 16-byte column groups per row
    int R;            // rows per visit of a team
    int SR;           // rows per stage (= kKP * R); SR | H or H | SR
    int stage_bytes;  // SR * row bytes (<= 16 KB)
    int pass_bytes;   // R * row bytes
    int nact;         // active threads per team (= R * G)
    int NSv;          // stages per volume
    int nslots;       // ring slots per CTA
    int slot_bytes;   // stage_bytes rounded up to 128
    int n_clusters;
    int st_hint;      // gradient stores carry an L2 evict_first hint
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
    unsigned* ticket;     // 1 word, zeroed before the launch: clusters that have finished
    double* loss_out;     // nullable: norm * sum of loss_terms, written by the last cluster
};

struct Scal {
    double E[3];              // Ex, Ey, Ez
    double K[3];              // g_a / dim_a / L per axis
    float M;                  // volume maximum
    float pad;
};

struct Smem {
    unsigned long long full[kMaxSlots], empty[kMaxSlots];
    unsigned long long red_ready[kPB], red_free[kPB];      // reduce team <-> merge warp
    unsigned long long scal_ready[kNB], scal_free[kNB];    // merge warp <-> write team
    unsigned long long xfull[2];                           // C remote arrivals per volume
    double wred[kPB][kNW][8];
    float wmax[kPB][kNW];
    double xch[2][kMaxCluster][8];                         // [buffer][source rank][7 sums + max]
    Scal scal[kNB];
};

template <int kId>
__device__ __forceinline__ void team_barrier() {
    asm volatile("bar.sync %0, %1;" ::"n"(kId), "n"(kC) : "memory");
}

// mbarrier waits that trap instead of hanging the GPU if the pipeline logic is ever wrong
__device__ __forceinline__ void wait_cta(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
#pragma unroll 1
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
__device__ __forceinline__ void wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
#pragma unroll 1
    for (unsigned spin = 0; spin < (1u << 26); ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}
__device__ __forceinline__ uint32_t mapa(uint32_t local_saddr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_saddr), "r"(rank));
    return remote;
}
__device__ __forceinline__ void st_cluster_f64(uint32_t remote_saddr, double v) {
    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(remote_saddr), "d"(v) : "memory");
}
// arrive on the mbarrier at the same offset in CTA `rank`; releases this thread's earlier
// st.shared::cluster stores at cluster scope
__device__ __forceinline__ void arrive_remote(uint32_t remote_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar) : "memory");
}

__device__ __forceinline__ float rebase_w(float m_from, float m_to) {
    // 2^(fl(m_from*log2e) - fl(m_to*log2e)): the same rounded products the exponentials used
    return (m_from == kNegInf) ? 0.f : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

// ring position of a role: slot index and the parity of the slot's current use
struct Ring {
    int slot;
    uint32_t lap;
    __device__ __forceinline__ void init() { slot = 0; lap = 0; }
    __device__ __forceinline__ void next(int nslots) {
        if (++slot == nslots) { slot = 0; ++lap; }
    }
};

// Thread roles inside the CTA (kThreads = 2*kC + 64):
//   [0, kC)          REDUCE team
//   [kC, 2kC)        WRITE team
//   warp 2kC/32      producer (one elected lane)
//   warp 2kC/32 + 1  merge warp
template <typename T>
__global__ void __maxnreg__(112) k_cluster(const Params p) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw + (size_t)p.nslots * p.slot_bytes);
    const uint32_t ring = smem_u32(smem_raw);

    const uint32_t C = cluster_nctarank();
    const uint32_t crank = cluster_ctarank();
    const long long cluster_id = blockIdx.x / C;
    const int sv0 = (int)(((long long)p.NSv * crank) / C);                   // first stage of this CTA's share
    const int nst = (int)(((long long)p.NSv * (crank + 1)) / C) - sv0;       // stages of the share
    // volumes this cluster takes: cluster_id, cluster_id + n_clusters, ...
    const long long nv = (p.n_vol > cluster_id) ? (p.n_vol - cluster_id + p.n_clusters - 1) / p.n_clusters : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nslots; ++s) {
            mbar_init(smem_u32(&sm.full[s]), 1);
            mbar_init(smem_u32(&sm.empty[s]), kNW);
        }
#pragma unroll
        for (int s = 0; s < kPB; ++s) {
            mbar_init(smem_u32(&sm.red_ready[s]), kNW);
            mbar_init(smem_u32(&sm.red_free[s]), 1);
        }
#pragma unroll
        for (int s = 0; s < kNB; ++s) {
            mbar_init(smem_u32(&sm.scal_ready[s]), 1);
            mbar_init(smem_u32(&sm.scal_free[s]), kNW);
        }
        mbar_init(smem_u32(&sm.xfull[0]), C);
        mbar_init(smem_u32(&sm.xfull[1]), C);
        mbar_fence_init();
        fence_proxy_async();
    }
    __syncthreads();
    // every CTA's barriers exist before any peer arrives on them
    cluster_sync_all();

    if (threadIdx.x >= 2 * kC + 32) {
        // ------------------------------- merge warp -------------------------------
        const int lane = threadIdx.x & 31;
        for (long long k = 0; k < nv; ++k) {
            const long long vol = cluster_id + k * p.n_clusters;
            const unsigned sb = (unsigned)(k % kPB), xb = (unsigned)(k & 1), rb = (unsigned)(k % kNB);
            // targets first: their latency overlaps the waits below
            double tk = 0.0, wk = 0.0;
            if (lane < 3) {
                tk = p.tgt_f64 ? reinterpret_cast<const double*>(p.gt)[vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.gt)[vol * 3 + lane];
                wk = p.tgt_f64 ? reinterpret_cast<const double*>(p.vis)[vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.vis)[vol * 3 + lane];
            }
            // ---- this CTA's partial: merge the reduce team's warp partials ----
            wait_cta(smem_u32(&sm.red_ready[sb]), (uint32_t)(k / kPB) & 1);
            const float pm = (lane < kNW) ? sm.wmax[sb][lane] : kNegInf;
            double a[kNSums];
#pragma unroll
            for (int i = 0; i < kNSums; ++i) a[i] = (lane < kNW) ? sm.wred[sb][lane][i] : 0.0;
            {
                // hand the buffer back only once the loads above have returned (see
                // mbar_arrive_after_loads; a warp's scoreboard covers all its lanes)
                uint32_t dep = __float_as_uint(pm);
#pragma unroll
                for (int i = 0; i < kNSums; ++i) dep ^= (uint32_t)__double2loint(a[i]);
                __syncwarp();
                if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.red_free[sb]), dep);
            }
            float cm = pm;
#pragma unroll
            for (int o = kNW / 2; o > 0; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
            const double f2 = (double)rebase_w(pm, cm);
#pragma unroll
            for (int i = 0; i < kNSums; ++i) {
                a[i] *= f2;
#pragma unroll
                for (int o = kNW / 2; o > 0; o >>= 1) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
            }
            // lanes 0..7 now all hold the CTA's sums; broadcast lane 0's copy to the whole warp
            cm = __shfl_sync(0xffffffffu, cm, 0);
#pragma unroll
            for (int i = 0; i < kNSums; ++i) a[i] = __shfl_sync(0xffffffffu, a[i], 0);
            // ---- publish to every CTA of the cluster (self included), lane r -> rank r ----
            if (lane < (int)C) {
                const uint32_t dst = mapa(smem_u32(&sm.xch[xb][crank][0]), (uint32_t)lane);
#pragma unroll
                for (int i = 0; i < kNSums; ++i) st_cluster_f64(dst + 8u * i, a[i]);
                st_cluster_f64(dst + 56u, (double)cm);
                arrive_remote(mapa(smem_u32(&sm.xfull[xb]), (uint32_t)lane));
            }
            // ---- all C partials of this volume, merged in rank order ----
            wait_cluster(smem_u32(&sm.xfull[xb]), (uint32_t)(k >> 1) & 1);
            double b[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) b[i] = (lane < (int)C) ? sm.xch[xb][lane][i] : 0.0;
            const float mi = (lane < (int)C) ? (float)b[7] : kNegInf;
            const float M = warp_max(mi);
            const double fr = (double)rebase_w(mi, M);
#pragma unroll
            for (int i = 0; i < kNSums; ++i) a[i] = warp_sum(b[i] * fr);
            // loss.py:39-41 (expectations) and :54-56 (normalisation), in fp64, rounded once.
            // Lane a (< 3) finishes axis a: x, y, z.
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
            if (k >= kNB) wait_cta(smem_u32(&sm.scal_free[rb]), (uint32_t)((k / kNB) - 1) & 1);
            Scal& rec = sm.scal[rb];
            if (lane < 3) {
                if (crank == 0) {
                    if (p.loss_terms) p.loss_terms[vol * 3 + lane] = fabs(dlt) * wk;
                    if (p.coords) p.coords[vol * 3 + lane] = cn;
                    if (p.expect) p.expect[vol * 3 + lane] = ex_;
                }
                rec.E[lane] = ex_;
                rec.K[lane] = (double)g1 / dim * ic;     // g_a / dim_a / L
            }
            if (lane == 0) {
                rec.M = M;
                if (crank == 0 && p.stats) { p.stats[vol * 2 + 0] = M; p.stats[vol * 2 + 1] = (float)a[0]; }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.scal_ready[rb]));
        }
        // The loss scalar without a second launch: rank 0's merge warps are the writers of
        // loss_terms, so the last of them to finish sums the terms (fixed order: lane-strided
        // partials, then the shuffle tree) -- loss.py:88-89.
        if (crank == 0 && p.loss_out && p.loss_terms) {
            __threadfence();
            __syncwarp();
            unsigned t = 0;
            if (lane == 0) t = atomicAdd(p.ticket, 1u);
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t == (unsigned)p.n_clusters - 1) {
                __threadfence();
                double acc = 0.0;
                for (long long i = lane; i < p.n_vol * 3; i += 32) acc += __ldcg(p.loss_terms + i);
                acc = warp_sum(acc);
                if (lane == 0) p.loss_out[0] = acc * p.norm;
            }
        }
        return;
    }

    if (threadIdx.x >= 2 * kC) {
        // ------------------------------- producer warp -------------------------------
        if ((threadIdx.x & 31) == 0) {
            const uint64_t pol = policy_evict_first();
            Ring r;
            r.init();
            for (long long k = 0; k < nv; ++k) {
                const long long vol = cluster_id + k * p.n_clusters;
                const unsigned char* src = p.in + vol * p.vol_bytes + (long long)sv0 * p.stage_bytes;
                for (int s = 0; s < nst; ++s) {
                    if (r.lap > 0) wait_cta(smem_u32(&sm.empty[r.slot]), (r.lap - 1) & 1);
                    const uint32_t bar = smem_u32(&sm.full[r.slot]);
                    mbar_arrive_expect_tx(bar, (uint32_t)p.stage_bytes);
                    tma_load_1d(ring + (uint32_t)r.slot * (uint32_t)p.slot_bytes, src + (long long)s * p.stage_bytes,
                                (uint32_t)p.stage_bytes, bar, pol);
                    r.next(p.nslots);
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
    // row offsets of this thread's kKP visits inside a stage
    int yo[kKP], zo[kKP];
#pragma unroll
    for (int q = 0; q < kKP; ++q) {
        const int rho = rr + q * p.R;
        zo[q] = rho / p.H;
        yo[q] = rho - zo[q] * p.H;
    }
    const int rowbase0 = sv0 * p.SR;              // first row of the share
    const int zb0 = rowbase0 / p.H, yb0 = rowbase0 - zb0 * p.H;
    const uint32_t lds_base = ring + 16u * tid;

    if (!is_writer) {
        // =============================== REDUCE team ===============================
        Ring r;
        r.init();
        for (long long k = 0; k < nv; ++k) {
            const unsigned sb = (unsigned)(k % kPB);
            float m = kNegInf, mL = kNegInf, tm = kNegInf;   // offset in use, fl(m*log2e), true maximum
            float2 c2[kKP][VEC / 2];
#pragma unroll
            for (int q = 0; q < kKP; ++q)
#pragma unroll
                for (int j = 0; j < VEC / 2; ++j) c2[q][j] = make_float2(0.f, 0.f);
            double s0t = 0.0, syb = 0.0, szb = 0.0;
            int zb = zb0, yb = yb0;
            auto fetch = [&](uint4 (&dst)[kKP]) {
                wait_cta(smem_u32(&sm.full[r.slot]), r.lap & 1);
                if (active) {
                    const uint32_t sbase = lds_base + (uint32_t)r.slot * (uint32_t)p.slot_bytes;
#pragma unroll
                    for (int q = 0; q < kKP; ++q) dst[q] = lds128(sbase + q * p.pass_bytes);
                }
                r.next(p.nslots);
            };
            auto compute = [&](const uint4 (&raw)[kKP]) {
                if (active) {
                    float v[kKP][VEC];
#pragma unroll
                    for (int q = 0; q < kKP; ++q) V16::unpack(raw[q], v[q]);
                    float smax = kNegInf;
#pragma unroll
                    for (int q = 0; q < kKP; ++q)
#pragma unroll
                        for (int j = 0; j < VEC; ++j) smax = fmaxf(smax, v[q][j]);
                    tm = fmaxf(tm, smax);
                    // LAZY re-basing: the offset m only moves when a value exceeds it by more than
                    // kLazy (e <= exp(kLazy) ~ 9e6 stays far inside fp32); the fold re-bases once,
                    // onto the thread's true maximum tm.
                    if (smax > m + kLazy) {
                        const float f = rebase_w(m, smax);
                        const float2 f2 = make_float2(f, f);
#pragma unroll
                        for (int q = 0; q < kKP; ++q)
#pragma unroll
                            for (int j = 0; j < VEC / 2; ++j) c2[q][j] = __fmul2_rn(c2[q][j], f2);
                        const double fd = (double)f;
                        s0t *= fd; syb *= fd; szb *= fd;
                        m = smax;
                        mL = __fmul_rn(smax, kLog2e);
                    }
                    // all -inf so far: exponentiate against 0 so that e = 2^-inf = 0 (no NaN)
                    const float noff = (m == kNegInf) ? 0.f : -mL;
                    const float2 noff2 = make_float2(noff, noff), l2e2 = make_float2(kLog2e, kLog2e);
                    float rowsum[kKP];
#pragma unroll
                    for (int q = 0; q < kKP; ++q) {
                        float2 e2[VEC / 2];
#pragma unroll
                        for (int j = 0; j < VEC / 2; ++j) {
                            const float2 a2 = __ffma2_rn(make_float2(v[q][2 * j], v[q][2 * j + 1]), l2e2, noff2);
                            e2[j] = make_float2(ex2(a2.x), ex2(a2.y));
                            c2[q][j] = __fadd2_rn(c2[q][j], e2[j]);
                        }
                        float2 r2 = __fadd2_rn(e2[0], e2[1]);
                        if constexpr (VEC == 8) r2 = __fadd2_rn(r2, __fadd2_rn(e2[2], e2[VEC / 2 - 1]));
                        rowsum[q] = r2.x + r2.y;
                    }
                    const float2 t2 = __fadd2_rn(make_float2(rowsum[0], rowsum[2]), make_float2(rowsum[1], rowsum[3]));
                    const double dt = (double)(t2.x + t2.y);
                    s0t += dt;
                    syb = fma((double)yb, dt, syb);
                    szb = fma((double)zb, dt, szb);
                }
                // next stage's base row (uniform): SR | H or H | SR, so one of the two moves
                yb += p.SR;
                while (yb >= p.H) { yb -= p.H; ++zb; }
            };
            // Stage s+1 is fetched into registers before the arithmetic of stage s: two stages'
            // dependency chains overlap in every warp.
            // (4-byte logits only: with 8 elements per vector the second buffer spills.)
            if constexpr (VEC == 4) {
                uint4 rawA[kKP], rawB[kKP];
                fetch(rawA);
                for (int s = 0; s < nst; s += 2) {
                    if (s + 1 < nst) fetch(rawB);
                    compute(rawA);
                    if (s + 1 < nst) {
                        if (s + 2 < nst) fetch(rawA);
                        compute(rawB);
                    }
                }
            } else {
                uint4 rawA[kKP];
                for (int s = 0; s < nst; ++s) {
                    fetch(rawA);
                    compute(rawA);
                }
            }
            // ---- fold the thread's accumulators, merge over the warp, hand to the merge warp ----
            double q7[kNSums];
#pragma unroll
            for (int i = 0; i < kNSums; ++i) q7[i] = 0.0;
            float pmx = kNegInf;
            if (active) {
                pmx = tm;
                double col[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) col[j] = 0.0;
#pragma unroll
                for (int q = 0; q < kKP; ++q) {
                    double row = 0.0;
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const double cd = (double)((j & 1) ? c2[q][j / 2].y : c2[q][j / 2].x);
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
                // from the lazy offset onto the thread's true maximum (weight <= 1)
                const double fl_ = (double)rebase_w(m, tm);
#pragma unroll
                for (int i = 0; i < kNSums; ++i) q7[i] *= fl_;
            }
            const float wm = warp_max(pmx);
            const double f = (double)rebase_w(pmx, wm);
#pragma unroll
            for (int i = 0; i < kNSums; ++i) q7[i] = warp_sum(q7[i] * f);
            if (k >= kPB) wait_cta(smem_u32(&sm.red_free[sb]), (uint32_t)((k / kPB) - 1) & 1);
            if (lane == 0) {
#pragma unroll
                for (int i = 0; i < kNSums; ++i) sm.wred[sb][warp][i] = q7[i];
                sm.wmax[sb][warp] = wm;
                mbar_arrive(smem_u32(&sm.red_ready[sb]));
            }
        }
        return;
    }

    // ================================ WRITE team ================================
    float yof[kKP], zof[kKP];
#pragma unroll
    for (int q = 0; q < kKP; ++q) { yof[q] = (float)yo[q]; zof[q] = (float)zo[q]; }
    const float Hf = (float)p.H, SRf = (float)p.SR;
    const uint64_t st_pol = policy_evict_first();
    Ring r;
    r.init();
    for (long long k = 0; k < nv; ++k) {
        const long long vol = cluster_id + k * p.n_clusters;
        const unsigned rb = (unsigned)(k % kNB);
        wait_cta(smem_u32(&sm.scal_ready[rb]), (uint32_t)(k / kNB) & 1);
        const Scal& rec = sm.scal[rb];
        const double kxd = rec.K[0], Ex = rec.E[0], Ey = rec.E[1], Ez = rec.E[2];
        const float ky = (float)rec.K[1], kz = (float)rec.K[2];
        const float ML = __fmul_rn(rec.M, kLog2e);
        {
            // the record slot goes back to the merge warp only once these loads have returned
            const uint32_t dep = (uint32_t)__double2loint(kxd) ^ (uint32_t)__double2loint(Ex) ^
                                 (uint32_t)__double2loint(Ey) ^ (uint32_t)__double2loint(Ez) ^
                                 __float_as_uint(ky) ^ __float_as_uint(kz) ^ __float_as_uint(ML);
            __syncwarp();
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.scal_free[rb]), dep);
        }
        const float2 nML2 = make_float2(-ML, -ML), l2e2 = make_float2(kLog2e, kLog2e);
        float ax[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) ax[j] = (float)(kxd * ((double)(x0 + j) - Ex));
        // (idx - E) must keep its ABSOLUTE accuracy: E split into hi + lo floats, the row index an
        // exact small float, so (y - Eh) is exact near the peak and - El restores the rest
        const float Eyh = (float)Ey, Eyl = (float)(Ey - (double)Eyh);
        const float Ezh = (float)Ez, Ezl = (float)(Ez - (double)Ezh);
        float ybf = (float)yb0, zbf = (float)zb0;
        unsigned char* gdst = p.out + vol * p.vol_bytes + (long long)sv0 * p.stage_bytes + 16ll * tid;

        auto fetch = [&](uint4 (&dst)[kKP]) {
            // already complete (the reduce team consumed this stage); makes the TMA's bytes
            // visible to this thread as well
            wait_cta(smem_u32(&sm.full[r.slot]), r.lap & 1);
            if (active) {
                const uint32_t sbase = lds_base + (uint32_t)r.slot * (uint32_t)p.slot_bytes;
#pragma unroll
                for (int q = 0; q < kKP; ++q) dst[q] = lds128(sbase + q * p.pass_bytes);
            }
            // the slot goes back to the producer only once the loads have RETURNED
            uint32_t dep = 0;
#pragma unroll
            for (int q = 0; q < kKP; ++q) dep ^= dst[q].x;
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.empty[r.slot]), active ? dep : 0u);
            r.next(p.nslots);
        };
        auto compute = [&](const uint4 (&raw)[kKP], int s) {
            if (active) {
#pragma unroll
                for (int q = 0; q < kKP; ++q) {
                    const float y = ybf + yof[q], z = zbf + zof[q];
                    const float rtv = fmaf(ky, (y - Eyh) - Eyl, kz * ((z - Ezh) - Ezl));
                    float v[VEC], o[VEC];
                    V16::unpack(raw[q], v);
                    const float2 rt2 = make_float2(rtv, rtv);
#pragma unroll
                    for (int j = 0; j < VEC / 2; ++j) {
                        const float2 a2 = __ffma2_rn(make_float2(v[2 * j], v[2 * j + 1]), l2e2, nML2);
                        const float2 w2 = __fadd2_rn(make_float2(ax[2 * j], ax[2 * j + 1]), rt2);
                        const float2 o2 = __fmul2_rn(make_float2(ex2(a2.x), ex2(a2.y)), w2);
                        o[2 * j] = o2.x;
                        o[2 * j + 1] = o2.y;
                    }
                    unsigned char* dst = gdst + (long long)s * p.stage_bytes + q * p.pass_bytes;
                    if (p.st_hint) stg_stream_hint(dst, V16::pack(o), st_pol);
                    else stg_stream(dst, V16::pack(o));
                }
            }
            ybf += SRf;
            while (ybf >= Hf) { ybf -= Hf; zbf += 1.f; }
        };
        uint4 rawA[kKP], rawB[kKP];
        fetch(rawA);
        for (int s = 0; s < nst; s += 2) {
            if (s + 1 < nst) fetch(rawB);
            compute(rawA, s);
            if (s + 1 < nst) {
                if (s + 2 < nst) fetch(rawA);
                compute(rawB, s + 1);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct Geom {
    int G, R, SR, stage_bytes, pass_bytes, nact, NSv, slot_bytes, nslots, C, max_share;
};

static bool pick_geom(int dtype, long long n_vol, int D, int H, int W, Geom* g) {
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    const long long row_b = (long long)W * esz;
    if (row_b % 16 != 0 || row_b / 16 > kC) return false;
    const int G = (int)(row_b / 16);
    const long long NR = (long long)D * H;
    if (NR >= (1ll << 30)) return false;
    bool found = false;
    for (int R = kC / G; R >= 1; --R) {
        const int SR = kKP * R;
        if (2 * R * G < kC) break;   // fewer than half the team would work
        const bool ok = (SR % H == 0 && NR % SR == 0) || (H % SR == 0);
        if (!ok) continue;
        g->G = G; g->R = R; g->SR = SR;
        g->pass_bytes = (int)(R * row_b);
        g->stage_bytes = (int)(SR * row_b);
        g->nact = R * G;
        g->NSv = (int)(NR / SR);
        found = true;
        break;
    }
    if (!found) return false;
    g->slot_bytes = (g->stage_bytes + 127) / 128 * 128;
    int nslots = (int)((kSmemMax - (int)sizeof(Smem) - 128) / g->slot_bytes);
    if (nslots > kMaxSlots) nslots = kMaxSlots;
    if (const char* env = getenv("HIPE_CLUSTER_SLOTS")) {
        const int v = atoi(env);
        if (v >= 2 && v < nslots) nslots = v;
    }
    g->nslots = nslots;
    // smallest cluster whose largest share leaves kMinSpare slots of look-ahead
    int want = 0;
    if (const char* env = getenv("HIPE_CLUSTER_C")) want = atoi(env);
    g->C = 0;
    for (int C = 1; C <= kMaxCluster; ++C) {
        if (want ? (C != want) : ((C & (C - 1)) != 0)) continue;   // auto: powers of two only
        if (g->NSv < C) break;
        const int share = (g->NSv + C - 1) / C;
        if (share + (want ? 1 : kMinSpare) > nslots) continue;
        g->C = C;
        g->max_share = share;
        break;
    }
    (void)n_vol;
    return g->C > 0;
}

template <typename T>
static int launch(const Params& p0, const Geom& g, cudaStream_t st) {
    auto kern = k_cluster<T>;
    const size_t smem = (size_t)g.nslots * g.slot_bytes + sizeof(Smem);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)g.C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // how many clusters of this shape the device can hold at once: the persistent grid
    cfg.gridDim = dim3((unsigned)g.C, 1, 1);
    int ncl = 0;
    e = cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg);
    if (getenv("HIPE_DEBUG"))
        fprintf(stderr, "hipe_cluster: C=%d smem=%zu threads=%d -> cudaOccupancyMaxActiveClusters: %s, %d clusters\n",
                g.C, smem, kThreads, cudaGetErrorString(e), ncl);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return (int)e; }
    if (ncl < 1) return HIPE_E_UNSUPPORTED;
    if (const char* env = getenv("HIPE_CLUSTER_GRID")) {
        const int v = atoi(env);
        if (v > 0 && v < ncl) ncl = v;
    }
    Params p = p0;
    p.n_clusters = (int)((long long)ncl < p.n_vol ? ncl : p.n_vol);
    cfg.gridDim = dim3((unsigned)(p.n_clusters * g.C), 1, 1);
    e = cudaLaunchKernelEx(&cfg, kern, p);
    return (int)e;
}

}  // namespace clu
}  // namespace hipe

using namespace hipe;

extern "C" long long hipe_cluster_workspace_bytes(int dtype, long long n_vol, int D, int H, int W) {
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return 0;
    if (n_vol <= 0 || D <= 0 || H <= 0 || W <= 0) return 0;
    clu::Geom g;
    if (!clu::pick_geom(dtype, n_vol, D, H, W, &g)) return 0;
    return 256;
}

extern "C" int hipe_cluster_plan(int dtype, long long n_vol, int D, int H, int W, int* cluster, int* nslots,
                                 int* share_stages, int* stage_bytes) {
    clu::Geom g;
    if ((dtype != HIPE_F32 && dtype != HIPE_BF16) || n_vol <= 0 || D <= 0 || H <= 0 || W <= 0 ||
        !clu::pick_geom(dtype, n_vol, D, H, W, &g))
        return HIPE_E_UNSUPPORTED;
    if (cluster) *cluster = g.C;
    if (nslots) *nslots = g.nslots;
    if (share_stages) *share_stages = g.max_share;
    if (stage_bytes) *stage_bytes = g.stage_bytes;
    return 0;
}

extern "C" int hipe_joint_loss_cluster(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                       const void* gt, const void* vis, int tgt_dtype, double norm,
                                       float* coords, double* expect, float* stats, double* loss_terms,
                                       double* loss, void* dlogits, void* workspace, long long workspace_bytes,
                                       void* stream) {
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (tgt_dtype != HIPE_F32 && tgt_dtype != HIPE_F64) return HIPE_E_DTYPE;
    if (n_vol < 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    if (n_vol == 0) return 0;
    if (!logits || !gt || !vis || !dlogits || !workspace) return HIPE_E_NULL;
    if (loss && !loss_terms) return HIPE_E_NULL;
    if ((reinterpret_cast<uintptr_t>(logits) & 15) || (reinterpret_cast<uintptr_t>(dlogits) & 15) ||
        (reinterpret_cast<uintptr_t>(workspace) & 255))
        return HIPE_E_ALIGN;
    if (n_vol > 0x7fffffffll) return HIPE_E_SHAPE;
    clu::Geom g;
    if (!clu::pick_geom(dtype, n_vol, D, H, W, &g)) return HIPE_E_UNSUPPORTED;
    if (workspace_bytes < 256) return HIPE_E_SHAPE;

    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    int major = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (major != 10) return HIPE_E_ARCH;

    clu::Params p = {};
    p.in = static_cast<const unsigned char*>(logits);
    p.out = static_cast<unsigned char*>(dlogits);
    p.n_vol = n_vol;
    p.vol_bytes = (long long)D * H * W * (dtype == HIPE_F32 ? 4 : 2);
    p.W = W; p.H = H; p.D = D;
    p.G = g.G; p.R = g.R; p.SR = g.SR; p.stage_bytes = g.stage_bytes; p.pass_bytes = g.pass_bytes;
    p.nact = g.nact; p.NSv = g.NSv; p.nslots = g.nslots; p.slot_bytes = g.slot_bytes;
    p.st_hint = 0;
    if (const char* env = getenv("HIPE_ST_HINT")) p.st_hint = (env[0] != '0') ? 1 : 0;
    p.coords = coords; p.expect = expect; p.stats = stats; p.loss_terms = loss_terms;
    p.gt = gt; p.vis = vis; p.tgt_f64 = (tgt_dtype == HIPE_F64); p.norm = norm;
    p.ticket = static_cast<unsigned*>(workspace);
    p.loss_out = loss;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    e = cudaMemsetAsync(workspace, 0, 256, st);
    if (e != cudaSuccess) return (int)e;
    return dtype == HIPE_F32 ? clu::launch<float>(p, g, st) : clu::launch<__nv_bfloat16>(p, g, st);
}

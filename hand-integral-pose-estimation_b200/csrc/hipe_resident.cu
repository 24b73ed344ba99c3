// JointLocationLoss forward + backward in ONE persistent launch in which HBM sees every heatmap
// volume once on the way in and its gradient once on the way out (sm_100a).
//
// Replaces common/nets/loss.py:74-91 followed by loss.backward() (main/train.py:113) of the
// reference.  The two-launch path (hipe_softargmax_fwd + hipe_softargmax_bwd) moves
// 2 V s + V g bytes per volume through HBM; this launch moves V s + V g: the second read of
// the logits, which the gradient needs, is served ON CHIP by the 126 MB L2, because the
// gradient pass trails the reduction pass by a bounded window of a few hundred KB per SM.
//
// A persistent grid (one CTA per SM, cooperative launch: all CTAs co-resident) walks the batch
// in ITEMS: a volume is cut into P items of <= 8 stages, a stage is SR whole rows (<= 16 KB),
// and the P items of one volume go to P different CTAs.  Every CTA runs two decoupled
// 256-thread teams, each fed by its own TMA producer warp and its own shared-memory ring:
//
//   REDUCE team   streams its items (1-D TMA bulk copies, L2 evict_last): online-max softmax,
//                 per-thread fp32 accumulators per owned (row-in-stage, column) position, folded
//                 to fp64 per item; PUBLISHES the item's partial sums to a global workspace and
//                 bumps the volume's arrival counter (release).  Never waits on another CTA.
//   WRITE team    for the same items, up to `lag` items behind: WAITS until the volume's counter
//                 shows all P partials (acquire), merges them in a fixed order (every CTA of the
//                 volume computes bit-identical scalars), forms the L1 gradient
//                 sign(c-gt)*vis*norm (loss.py:86-89), streams the item AGAIN (TMA, L2 hit,
//                 evict_first) and writes
//                 dlogits = e/L * (gx(x-Ex)/W + gy(y-Ey)/H + gz(z-Ez)/D).
//
// The reduce producer is throttled to stay at most `lag` items ahead of the write team, which
// bounds the L2 window to grid * lag * 128 KB (57 MB at lag 3) and gives the write team enough
// slack that the skew between the CTAs sharing a volume never stalls it.  A first version kept
// the exponentials in a 13-slot shared-memory ring instead of re-reading through L2; with so
// little look-ahead every item behaved like a barrier across the P CTAs (write team polling the
// flag ~17 times per item, 0.59 ms per 1344 volumes of 64^3 fp32 against 0.43 ms at HBM speed).
//
// Deadlock freedom: reduce teams depend only on their own write team's progress on EARLIER
// items; write teams wait on volumes whose partials come from reduce teams: no cycle, provided
// all CTAs are co-resident.
//
// Tried and dropped: CLAIMING items from a global counter instead of dealing them round-robin
// (so that an SM with less bandwidth takes fewer).  The claim has to come after the lag
// throttle to stay deadlock-free, which puts an atomic round trip on the one path the launch
// is sensitive to: 0.69 ms against 0.57 ms for the round-robin in the same build.  The L2
// policies of the two rings and of the gradient stores were swept too (evict_last / normal /
// evict_first x lag 3-5): nothing beats evict_last reads + evict_first re-reads + plain stores.
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

constexpr int kC = 256;               // threads per team
constexpr int kNW = kC / 32;          // warps per team
// When a consumer warp hands a ring slot back relative to the arithmetic of the previous stage:
// 0 = right after issuing the loads (the arrive then waits for them: no overlap of two stages'
// dependency chains), 1 = half-way through the previous stage's arithmetic, 2 = after it.
// Measured (scripts/ab_resident_release.sh, fp32 64^3 x 1344): 0 -> 0.519 ms, 1 -> 0.568, 2 -> 0.572:
// the launch is far more sensitive to WHEN a slot comes back than to how many there are.
#ifndef HIPE_RES_RELEASE
#define HIPE_RES_RELEASE 0
#endif
// How a volume's P partials travel between CTAs.  0: data, __threadfence, arrival counter, then
// the scalar warps read the data.  1: every 8-byte word of a partial validates itself (the
// workspace is preset to all-ones, which no stored word can equal), so the publish warp needs no
// fence + counter and the scalar warps poll the data itself: one global round trip instead of
// three.  Parity-green, but NOT faster (scripts/ab_resident_partials.sh, same box: 0.543 ms vs
// 0.549-0.551, the difference being the larger workspace preset), which says the publish -> merge
// latency is not what paces the launch.  Kept as a build option.
#ifndef HIPE_RES_SELFVALID
#define HIPE_RES_SELFVALID 0
#endif
#ifndef HIPE_RES_POLL_NS
#define HIPE_RES_POLL_NS 0
#endif
#ifndef HIPE_RES_RSLOTS
#define HIPE_RES_RSLOTS 7
#endif
#ifndef HIPE_RES_WSLOTS
#define HIPE_RES_WSLOTS 6
#endif
constexpr int kRSlots = HIPE_RES_RSLOTS;   // reduce ring depth (HBM latency to cover)
constexpr int kWSlots = HIPE_RES_WSLOTS;   // write ring depth (L2 latency to cover)
constexpr int kSlotBytes = 16384;     // ring slot
constexpr int kKP = 4;                // 16-byte vectors per thread per stage
constexpr int kMaxItemStages = 16;    // unroll bound: an item never exceeds 16 stages (256 KB)
constexpr int kDefItemStages = 8;     // default item size in stages (128 KB)
constexpr int kMaxRtRows = 1024;      // rows of one item (<= 8 stages x 128 rows)
constexpr int kNSums = 7;             // S0c Sx Syo Szo S0t Syb Szb
constexpr int kRD = 32;               // write-team progress barriers (> any lag)
constexpr int kMaxLag = 24;
constexpr int kNB = 8;                // scalar records in flight (scalar warp -> write team)
constexpr int kPB = 4;                // per-item warp partials in flight (reduce team -> publish warp)
constexpr int kThreads = 2 * kC + 128; // two teams + two producer warps + the scalar warp + the publish warp
constexpr unsigned kSpinLimit = 1u << 22;
constexpr int kProfN = 20;                // wait classes of the diagnostic profile (HIPE_RES_PROF)
constexpr float kLazy = 16.f;         // lazy re-basing threshold of the online softmax (natural-log units)

struct Params {
    const unsigned char* in;
    unsigned char* out;
    long long n_vol;
    long long vol_bytes;
    long long n_items;
    int W, H, D;
    int G;            // 16-byte column groups per row
    int R;            // rows per visit of a team
    int SR;           // rows per stage (= kKP * R); SR | H or H | SR
    int stage_bytes;  // SR * row bytes (<= 16 KB)
    int pass_bytes;   // R * row bytes
    int nact;         // active threads per team (= R * G)
    int NSv;          // stages per volume
    int P;            // items per volume
    int lag;          // items the reduce side may run ahead of the write side
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
    unsigned* ticket;     // 1 word: scalar warps that have finished (the last one sums the loss)
    double* loss_out;     // nullable: norm * sum of loss_terms, written by the last scalar warp
};

struct Smem {
    float rt[2][kMaxRtRows];      // per-row gradient term of the item being written (double-buffered)
    unsigned long long rfull[kRSlots], rempty[kRSlots];
    unsigned long long wfull[kWSlots], wempty[kWSlots];
    unsigned long long w_done[kRD];   // write team finished item (index item_no % kRD)
    unsigned long long scal_ready[kNB];   // scalar warp -> write team
    unsigned long long red_ready[kPB];    // reduce team (kNW arrivals) -> publish warp
    unsigned long long red_free[kPB];     // publish warp -> reduce team
    double wred[kPB][kNW][8];     // per-warp partials of an item
    float wmax[kPB][kNW];
    struct Scal {
        double E[3];              // Ex, Ey, Ez
        double K[3];              // g_a / dim_a / L per axis
        float M;                  // volume maximum
        float pad;
    } scal[kNB];
#ifdef HIPE_RES_PROF
    unsigned long long prof[kProfN];
    unsigned prof_done;
#endif
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

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Wait-time profile (diagnostic builds only, -DHIPE_RES_PROF; scripts/diag_resident_waits.sh):
// lane 0 of every warp adds the clock cycles it spent in each class of wait to a shared-memory
// table; the last warp of the CTA to finish adds the table to g_prof.
#ifdef HIPE_RES_PROF
__device__ unsigned long long g_prof[kProfN];
#define RES_PROF_T0() const long long prof_t0_ = clock64()
#define RES_PROF_ADD(cls)                                                                              \
    do {                                                                                               \
        if ((threadIdx.x & 31) == 0) atomicAdd(&sm.prof[cls], (unsigned long long)(clock64() - prof_t0_)); \
    } while (0)
#define RES_WAIT(cls, bar, par)          \
    do {                                 \
        RES_PROF_T0();                   \
        mbar_wait_bounded(bar, par);     \
        RES_PROF_ADD(cls);               \
    } while (0)
#define RES_FINISH(cls)                                                                      \
    do {                                                                                     \
        __syncwarp();                                                                        \
        if ((threadIdx.x & 31) == 0) {                                                       \
            atomicAdd(&sm.prof[cls], (unsigned long long)(clock64() - prof_start_));         \
            __threadfence_block();                                                           \
            if (atomicAdd(&sm.prof_done, 1u) == kThreads / 32 - 1) {                         \
                __threadfence_block();                                                       \
                for (int k_ = 0; k_ < kProfN; ++k_) atomicAdd(&g_prof[k_], sm.prof[k_]);     \
            }                                                                                \
        }                                                                                    \
    } while (0)
#else
#define RES_WAIT(cls, bar, par) mbar_wait_bounded(bar, par)
#define RES_FINISH(cls) do { } while (0)
#endif

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
constexpr unsigned long long kUnset = ~0ull;   // what cudaMemsetAsync(0xFF) leaves in a word

__device__ __forceinline__ float rebase_w(float m_from, float m_to) {
    // 2^(fl(m_from*log2e) - fl(m_to*log2e)): the same rounded products the exponentials used
    return (m_from == kNegInf) ? 0.f : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

// Thread roles inside the CTA (kThreads = 2*kC + 64):
//   [0, kC)          REDUCE team
//   [kC, 2kC)        WRITE team
//   warp 2kC/32      reduce-ring producer (one elected lane), throttled by the write team
//   warp 2kC/32 + 1  write-ring producer (one elected lane)
//   warp 2kC/32 + 2  scalar warp: flag wait, merge of the P partials, L1 gradient
//   warp 2kC/32 + 3  publish warp: merges the reduce team's warp partials, writes + flags them
template <typename T>
__global__ void __launch_bounds__(kThreads, 1) k_resident(const Params p) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw + (kRSlots + kWSlots) * kSlotBytes);
    const uint32_t rring = smem_u32(smem_raw);
    const uint32_t wring = rring + kRSlots * kSlotBytes;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kRSlots; ++s) {
            mbar_init(smem_u32(&sm.rfull[s]), 1);
            mbar_init(smem_u32(&sm.rempty[s]), kNW);
        }
#pragma unroll
        for (int s = 0; s < kWSlots; ++s) {
            mbar_init(smem_u32(&sm.wfull[s]), 1);
            mbar_init(smem_u32(&sm.wempty[s]), kNW);
        }
#pragma unroll
        for (int s = 0; s < kRD; ++s) mbar_init(smem_u32(&sm.w_done[s]), kNW);
#pragma unroll
        for (int s = 0; s < kNB; ++s) mbar_init(smem_u32(&sm.scal_ready[s]), 1);
#pragma unroll
        for (int s = 0; s < kPB; ++s) {
            mbar_init(smem_u32(&sm.red_ready[s]), kNW);
            mbar_init(smem_u32(&sm.red_free[s]), 1);
        }
        mbar_fence_init();
        fence_proxy_async();
#ifdef HIPE_RES_PROF
        for (int k = 0; k < kProfN; ++k) sm.prof[k] = 0;
        sm.prof_done = 0;
#endif
    }
    __syncthreads();
#ifdef HIPE_RES_PROF
    const long long prof_start_ = clock64();
#endif

    if (threadIdx.x >= 2 * kC + 96) {
        // ------------------------------- publish warp -------------------------------
        // Merges the reduce team's kNW warp partials of an item, writes the item's partial to the
        // workspace and bumps the volume's arrival counter (release).
        const int lane = threadIdx.x & 31;
        unsigned item_no = 0;
        for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
            const Item it = decode(p, w);
            const unsigned sb = item_no % kPB;
            RES_WAIT(15, smem_u32(&sm.red_ready[sb]), (item_no / kPB) & 1);
            const float pm = (lane < kNW) ? sm.wmax[sb][lane] : kNegInf;
            double a[kNSums];
#pragma unroll
            for (int k = 0; k < kNSums; ++k) a[k] = (lane < kNW) ? sm.wred[sb][lane][k] : 0.0;
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.red_free[sb]));
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
#if HIPE_RES_SELFVALID
            if (lane < 8) {   // the xor tree leaves the sums in every lane of the first group
                double mine = (double)cm;
#pragma unroll
                for (int k = 0; k < kNSums; ++k) mine = (lane == k) ? a[k] : mine;
                unsigned long long bits = (unsigned long long)__double_as_longlong(mine);
                if (bits == kUnset) bits = 0x7ff8000000000000ull;   // a NaN stays a NaN
                __stcg(reinterpret_cast<unsigned long long*>(p.partials + (it.vol * p.P + it.i) * 8) + lane, bits);
            }
#else
            if (lane == 0) {
                double* dst = p.partials + (it.vol * p.P + it.i) * 8;
#pragma unroll
                for (int k = 0; k < kNSums; ++k) __stcg(dst + k, a[k]);
                __stcg(dst + 7, (double)cm);
                __threadfence();
                atomicAdd(p.counters + it.vol, 1u);
            }
#endif
        }
        RES_FINISH(16);
        return;
    }

    if (threadIdx.x >= 2 * kC + 64) {
        // ------------------------------- scalar warp -------------------------------
        // Per item: wait for the volume's P partials, merge them, finish the coordinates and the
        // L1 gradient, and leave the record the write team needs.  Runs ahead of the write team
        // by up to kNB items, so none of these global round trips is on a team's critical path.
        const int lane = threadIdx.x & 31;
        unsigned item_no = 0;
        for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
            const Item it = decode(p, w);
            if (item_no >= (unsigned)kNB) {
                const unsigned d = item_no - (unsigned)kNB;   // its record slot must have been consumed
                RES_WAIT(9, smem_u32(&sm.w_done[d % kRD]), (d / kRD) & 1);
            }
            Smem::Scal& rec = sm.scal[item_no % kNB];
            // targets first: their latency overlaps the flag wait
            double tk = 0.0, wk = 0.0;
            if (lane < 3) {
                tk = p.tgt_f64 ? reinterpret_cast<const double*>(p.gt)[it.vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.gt)[it.vol * 3 + lane];
                wk = p.tgt_f64 ? reinterpret_cast<const double*>(p.vis)[it.vol * 3 + lane]
                               : (double)reinterpret_cast<const float*>(p.vis)[it.vol * 3 + lane];
            }
            unsigned spins = 0;
            bool gave_up = false;       // an inter-CTA wait hit its spin limit: the outputs are poisoned below
#ifdef HIPE_RES_PROF
            RES_PROF_T0();
#endif
            float mi = kNegInf;
            double a[kNSums];
#pragma unroll
            for (int k = 0; k < kNSums; ++k) a[k] = 0.0;
#if HIPE_RES_SELFVALID
            // Poll the partials themselves: lane i keeps partial i's eight words once all of them
            // are set; partials beyond 32 (P > 32: 96^3) are only checked here and read below.
            unsigned long long bw[8];
            for (;;) {
                bool ok = true;
                if (lane < p.P) {
                    const double* src = p.partials + (it.vol * p.P + lane) * 8;
#pragma unroll
                    for (int k = 0; k < 8; ++k) bw[k] = ld_relaxed_u64(src + k);
#pragma unroll
                    for (int k = 0; k < 8; ++k) ok = ok && (bw[k] != kUnset);
                }
                for (int i = lane + 32; i < p.P; i += 32) {
                    const double* src = p.partials + (it.vol * p.P + i) * 8;
#pragma unroll
                    for (int k = 0; k < 8; ++k) ok = ok && (ld_relaxed_u64(src + k) != kUnset);
                }
                if (__all_sync(0xffffffffu, ok)) break;
                if (HIPE_RES_POLL_NS > 0) __nanosleep(HIPE_RES_POLL_NS);
                if (++spins > kSpinLimit) {
                    if (lane == 0) atomicExch(p.err, 1u);
                    gave_up = true;
                    break;
                }
            }
#ifdef HIPE_RES_PROF
            RES_PROF_ADD(6);
#endif
            for (int i = lane; i < p.P; i += 32) {
                double b[8];
                if (i < 32) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) b[k] = __longlong_as_double((long long)bw[k]);
                } else {
                    const double* src = p.partials + (it.vol * p.P + i) * 8;
#pragma unroll
                    for (int k = 0; k < 8; ++k) b[k] = __longlong_as_double((long long)ld_relaxed_u64(src + k));
                }
#else
            while (ld_acquire_u32(p.counters + it.vol) < (unsigned)p.P) {
                __nanosleep(64);
                if (++spins > kSpinLimit) {
                    if (lane == 0) atomicExch(p.err, 1u);
                    gave_up = true;
                    break;
                }
            }
#ifdef HIPE_RES_PROF
            RES_PROF_ADD(6);
#endif
            for (int i = lane; i < p.P; i += 32) {
                const double* src = p.partials + (it.vol * p.P + i) * 8;
                double b[8];
    #pragma unroll
                for (int k = 0; k < 8; ++k) b[k] = __ldcg(src + k);
#endif
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
            // A wait that gave up merged whatever was in the workspace: nothing computed from it may
            // look valid.  NaN here poisons the coordinates, the loss term, the loss scalar and (through
            // rec.K) every gradient element of this CTA's item of the volume.
            const double ex_ = gave_up ? __longlong_as_double(0x7ff8000000000000ll)
                                       : num_c * ic + (lane == 0 ? 0.0 : num_t * itt);
            const float cn = (float)(ex_ / dim - 0.5);
            // loss.py:86-89 on this volume's three coordinates, in the targets' precision
            const double dlt = (double)cn - tk;
            const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
            const float g1 = gave_up ? __int_as_float(0x7fc00000) : (float)(sg * wk * p.norm);
            if (lane < 3) {
                if (it.i == 0 || gave_up) {
                    if (p.loss_terms) p.loss_terms[it.vol * 3 + lane] = gave_up ? ex_ : fabs(dlt) * wk;
                    if (p.coords) p.coords[it.vol * 3 + lane] = cn;
                    if (p.expect) p.expect[it.vol * 3 + lane] = ex_;
                }
                rec.E[lane] = ex_;
                rec.K[lane] = (double)g1 / dim * ic;     // g_a / dim_a / L
            }
            if (lane == 0) {
                rec.M = M;
                if (it.i == 0 && p.stats) { p.stats[it.vol * 2 + 0] = M; p.stats[it.vol * 2 + 1] = (float)a[0]; }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.scal_ready[item_no % kNB]));
        }
        // The loss scalar without a second launch: the scalar warps are the writers of loss_terms,
        // so the last of them to finish sums the terms (fixed order: lane-strided partials, then
        // the shuffle tree) -- loss.py:88-89.
        if (p.loss_out && p.loss_terms) {
            __threadfence();
            unsigned t = 0;
            if (lane == 0) t = atomicAdd(p.ticket, 1u) + (HIPE_RES_SELFVALID ? 1u : 0u);   // preset to ~0 / 0
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t == gridDim.x - 1) {
                __threadfence();
                double acc = 0.0;
                for (long long i = lane; i < p.n_vol * 3; i += 32) acc += __ldcg(p.loss_terms + i);
                acc = warp_sum(acc);
                if (lane == 0) p.loss_out[0] = acc * p.norm;
            }
        }
        RES_FINISH(14);
        return;
    }

    if (threadIdx.x >= 2 * kC) {
        // ------------------------------ producer warps ------------------------------
        const bool for_write = threadIdx.x >= 2 * kC + 32;
        if ((threadIdx.x & 31) == 0) {
            const uint64_t pol = for_write ? policy_evict_first() : policy_evict_last();
            const int nslots = for_write ? kWSlots : kRSlots;
            const uint32_t ring = for_write ? wring : rring;
            unsigned long long* full = for_write ? sm.wfull : sm.rfull;
            unsigned long long* empty = for_write ? sm.wempty : sm.rempty;
            unsigned g = 0, item_no = 0;
            for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
                const Item it = decode(p, w);
                // L2 window: the reduce side stays at most `lag` items ahead of the write side
                if (!for_write && item_no >= (unsigned)p.lag) {
                    const unsigned d = item_no - (unsigned)p.lag;
                    RES_WAIT(4, smem_u32(&sm.w_done[d % kRD]), (d / kRD) & 1);
                }
                const unsigned char* src = p.in + it.vol * p.vol_bytes + (long long)it.sv0 * p.stage_bytes;
                for (int s = 0; s < it.n; ++s, ++g) {
                    const unsigned slot = g % nslots;
                    if (g >= (unsigned)nslots) {
                        if (for_write) RES_WAIT(3, smem_u32(&empty[slot]), ((g / nslots) - 1) & 1);
                        else RES_WAIT(2, smem_u32(&empty[slot]), ((g / nslots) - 1) & 1);
                    }
                    const uint32_t bar = smem_u32(&full[slot]);
                    mbar_arrive_expect_tx(bar, (uint32_t)p.stage_bytes);
                    tma_load_1d(ring + slot * kSlotBytes, src + (long long)s * p.stage_bytes, (uint32_t)p.stage_bytes,
                                bar, pol);
                }
            }
        }
        RES_FINISH(threadIdx.x >= 2 * kC + 32 ? 13 : 12);
        return;
    }

    const bool is_writer = threadIdx.x >= kC;
    const int tid = is_writer ? threadIdx.x - kC : threadIdx.x;   // team-local
    const int lane = tid & 31, warp = tid >> 5;
    const bool active = tid < p.nact;
    const int cg = active ? tid % p.G : 0;
    const int rr = active ? tid / p.G : 0;
    const int x0 = cg * VEC;

    if (!is_writer) {
        // =============================== REDUCE team ===============================
        const uint32_t lds_base = rring + 16u * tid;
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
            const unsigned sb = item_no % kPB;
            float m = kNegInf, mL = kNegInf, tm = kNegInf;   // offset in use, fl(m*log2e), true maximum
            float2 c2[kKP][VEC / 2];
#pragma unroll
            for (int q = 0; q < kKP; ++q)
#pragma unroll
                for (int j = 0; j < VEC / 2; ++j) c2[q][j] = make_float2(0.f, 0.f);
            double s0t = 0.0, syb = 0.0, szb = 0.0;
            int rowbase = it.sv0 * p.SR;
            int zb = rowbase / p.H;
            int yb = rowbase - zb * p.H;
            // Stage s+1 is fetched into registers (and its slot handed back) before the arithmetic
            // of stage s: two stages' dependency chains overlap in every warp.
            uint4 raw[2][kKP];
            auto fetch = [&](unsigned gg, uint4 (&dst)[kKP]) {
                const unsigned slot = gg % kRSlots;
                RES_WAIT(0, smem_u32(&sm.rfull[slot]), (gg / kRSlots) & 1);
                if (active) {
                    const uint32_t sbase = lds_base + slot * kSlotBytes;
#pragma unroll
                    for (int q = 0; q < kKP; ++q) dst[q] = lds128(sbase + q * p.pass_bytes);
                }
            };
            // The slot goes back to the producer only once the loads have RETURNED (the arrive
            // consumes a word of every load instruction; a warp's scoreboard covers all its
            // lanes), and one compute phase later, so that waiting for them stalls nothing.
            auto release = [&](unsigned gg, const uint4 (&src)[kKP]) {
                uint32_t dep = 0;
#pragma unroll
                for (int q = 0; q < kKP; ++q) dep ^= src[q].x;
                if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.rempty[gg % kRSlots]), active ? dep : 0u);
            };
            fetch(g, raw[0]);
            release(g, raw[0]);
#pragma unroll
            for (int s = 0; s < kMaxItemStages; ++s) {
                if (s >= it.n) break;
                if (s + 1 < it.n) {
                    fetch(g + s + 1, raw[(s + 1) & 1]);
                    if (HIPE_RES_RELEASE == 0) release(g + s + 1, raw[(s + 1) & 1]);
                }
                if (active) {
                    float v[kKP][VEC];
#pragma unroll
                    for (int q = 0; q < kKP; ++q) V16::unpack(raw[s & 1][q], v[q]);
                    float smax = kNegInf;
#pragma unroll
                    for (int q = 0; q < kKP; ++q)
#pragma unroll
                        for (int j = 0; j < VEC; ++j) smax = fmaxf(smax, v[q][j]);
                    tm = fmaxf(tm, smax);
                    // LAZY re-basing: the offset m only moves when a value exceeds it by more than
                    // kLazy (e <= exp(kLazy) ~ 9e6 stays far inside fp32); the fold re-bases once,
                    // onto the thread's true maximum tm.  Saves ~30 instructions per stage.
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
                        if (HIPE_RES_RELEASE == 1 && q == kKP / 2 && s + 1 < it.n) release(g + s + 1, raw[(s + 1) & 1]);
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
                if (HIPE_RES_RELEASE == 2 && s + 1 < it.n) release(g + s + 1, raw[(s + 1) & 1]);
                // next stage's base row (uniform): SR | H or H | SR, so one of the two moves
                yb += p.SR;
                while (yb >= p.H) { yb -= p.H; ++zb; }
            }
            g += it.n;
            // ---- fold the thread's accumulators, merge over the team, publish the item's partial ----
            double q7[kNSums];
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q7[k] = 0.0;
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
                for (int k = 0; k < kNSums; ++k) q7[k] *= fl_;
            }
            const float wm = warp_max(pmx);
            const double f = (double)rebase_w(pmx, wm);
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q7[k] = warp_sum(q7[k] * f);
            // hand the warp's partial to the publish warp: no team barrier, nobody waits for the
            // global round trip
            if (item_no >= (unsigned)kPB) RES_WAIT(7, smem_u32(&sm.red_free[sb]), ((item_no / kPB) - 1) & 1);
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < kNSums; ++k) sm.wred[sb][warp][k] = q7[k];
                sm.wmax[sb][warp] = wm;
                mbar_arrive(smem_u32(&sm.red_ready[sb]));
            }
        }
        RES_FINISH(10);
        return;
    }

    // ================================ WRITE team ================================
    const uint32_t lds_base = wring + 16u * tid;
    unsigned g = 0;
    unsigned item_no = 0;
    for (long long w = blockIdx.x; w < p.n_items; w += gridDim.x, ++item_no) {
        const Item it = decode(p, w);
        RES_WAIT(5, smem_u32(&sm.scal_ready[item_no % kNB]), (item_no / kNB) & 1);
        const Smem::Scal& rec = sm.scal[item_no % kNB];
        float* rt = sm.rt[item_no & 1];
        const double kxd = rec.K[0], kyd = rec.K[1], kzd = rec.K[2];
        const double Ex = rec.E[0], Ey = rec.E[1], Ez = rec.E[2];
        const float ML = __fmul_rn(rec.M, kLog2e);
        const float2 nML2 = make_float2(-ML, -ML), l2e2 = make_float2(kLog2e, kLog2e);
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
                rt[lr] = (float)(kyd * ((double)y - Ey) + kzd * ((double)z - Ez));
            }
        }
        // rt[b] visible; rt[1-b] is refilled only after every warp has passed here
#ifdef HIPE_RES_PROF
        {
            RES_PROF_T0();
            team_barrier<2>();
            RES_PROF_ADD(8);
        }
#else
        team_barrier<2>();
#endif

        unsigned char* gdst = p.out + it.vol * p.vol_bytes + (long long)it.sv0 * p.stage_bytes + 16ll * tid;
        uint4 raw[2][kKP];
        auto fetch = [&](unsigned gg, uint4 (&dst)[kKP]) {
            const unsigned slot = gg % kWSlots;
            RES_WAIT(1, smem_u32(&sm.wfull[slot]), (gg / kWSlots) & 1);
            if (active) {
                const uint32_t sbase = lds_base + slot * kSlotBytes;
#pragma unroll
                for (int q = 0; q < kKP; ++q) dst[q] = lds128(sbase + q * p.pass_bytes);
            }
        };
        auto release = [&](unsigned gg, const uint4 (&src)[kKP]) {   // see the reduce team
            uint32_t dep = 0;
#pragma unroll
            for (int q = 0; q < kKP; ++q) dep ^= src[q].x;
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.wempty[gg % kWSlots]), active ? dep : 0u);
        };
        fetch(g, raw[0]);
        release(g, raw[0]);
#pragma unroll
        for (int s = 0; s < kMaxItemStages; ++s) {
            if (s >= it.n) break;
            if (s + 1 < it.n) {
                    fetch(g + s + 1, raw[(s + 1) & 1]);
                    if (HIPE_RES_RELEASE == 0) release(g + s + 1, raw[(s + 1) & 1]);
                }
            if (active) {
#pragma unroll
                for (int q = 0; q < kKP; ++q) {
                    if (HIPE_RES_RELEASE == 1 && q == kKP / 2 && s + 1 < it.n) release(g + s + 1, raw[(s + 1) & 1]);
                    const float rtv = rt[s * p.SR + rr + q * p.R];
                    float v[VEC], o[VEC];
                    V16::unpack(raw[s & 1][q], v);
                    const float2 rt2 = make_float2(rtv, rtv);
#pragma unroll
                    for (int j = 0; j < VEC / 2; ++j) {
                        const float2 a2 = __ffma2_rn(make_float2(v[2 * j], v[2 * j + 1]), l2e2, nML2);
                        const float2 w2 = __fadd2_rn(make_float2(ax[2 * j], ax[2 * j + 1]), rt2);
                        const float2 o2 = __fmul2_rn(make_float2(ex2(a2.x), ex2(a2.y)), w2);
                        o[2 * j] = o2.x;
                        o[2 * j + 1] = o2.y;
                    }
                    stg_stream(gdst + (long long)s * p.stage_bytes + q * p.pass_bytes, V16::pack(o));
                }
            }
            if (HIPE_RES_RELEASE == 2 && s + 1 < it.n) release(g + s + 1, raw[(s + 1) & 1]);
        }
        g += it.n;
        // progress report for the reduce producer's L2-window throttle
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&sm.w_done[item_no % kRD]));
    }
    RES_FINISH(11);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct Geom {
    int G, R, SR, stage_bytes, pass_bytes, nact, NSv, P;
};

static int item_stages() {
    int n = kDefItemStages;
    if (const char* env = getenv("HIPE_RESIDENT_ITEM_STAGES")) {
        const int v = atoi(env);
        if (v >= 1 && v <= kMaxItemStages) n = v;
    }
    return n;
}

static bool pick_geom(int dtype, int D, int H, int W, Geom* g) {
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    const long long row_b = (long long)W * esz;
    if (row_b % 16 != 0 || row_b / 16 > kC) return false;
    const int G = (int)(row_b / 16);
    const long long NR = (long long)D * H;
    if (NR >= (1ll << 30)) return false;
    for (int R = kC / G; R >= 1; --R) {
        const int SR = kKP * R;
        if (SR > kMaxRtRows / kDefItemStages) continue;
        if (2 * R * G < kC) break;   // fewer than half the consumers would work
        const bool ok = (SR % H == 0 && NR % SR == 0) || (H % SR == 0);
        if (!ok) continue;
        g->G = G; g->R = R; g->SR = SR;
        g->pass_bytes = (int)(R * row_b);
        g->stage_bytes = (int)(SR * row_b);
        g->nact = R * G;
        g->NSv = (int)(NR / SR);
        int ist = item_stages();
        if (ist * SR > kMaxRtRows) ist = kMaxRtRows / SR;
        g->P = (g->NSv + ist - 1) / ist;
        return true;
    }
    return false;
}

static size_t ws_counters_bytes(long long n_vol) {
    return (size_t)(((n_vol + 2) * sizeof(unsigned) + 255) / 256 * 256);   // counters + error word + ticket
}

}  // namespace res
}  // namespace hipe

using namespace hipe;

#ifdef HIPE_RES_PROF
// diagnostic builds only: read (and optionally clear) the wait-time table
extern "C" __attribute__((visibility("default"))) int hipe_resident_prof(unsigned long long* out, int reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out, res::g_prof, sizeof(unsigned long long) * res::kProfN);
    if (e != cudaSuccess) return (int)e;
    if (reset) {
        unsigned long long z[res::kProfN] = {};
        e = cudaMemcpyToSymbol(res::g_prof, z, sizeof(z));
    }
    return (int)e;
}
#endif

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

    const void* kern = dtype == HIPE_F32 ? (const void*)res::k_resident<float>
                                         : (const void*)res::k_resident<__nv_bfloat16>;
    const size_t smem = (size_t)(res::kRSlots + res::kWSlots) * res::kSlotBytes + sizeof(res::Smem);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, res::kThreads, smem);
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
    // L2 window = grid * lag * (<= 128 KB): 57 MB at the default, inside the 126 MB L2
    p.lag = 3;
    if (const char* env = getenv("HIPE_RESIDENT_LAG")) {
        const int v = atoi(env);
        if (v >= 1 && v <= res::kMaxLag) p.lag = v;
    }
    p.coords = coords; p.expect = expect; p.stats = stats; p.loss_terms = loss_terms;
    p.gt = gt; p.vis = vis; p.tgt_f64 = (tgt_dtype == HIPE_F64); p.norm = norm;
    const size_t cb = res::ws_counters_bytes(n_vol);
    p.counters = static_cast<unsigned*>(workspace);
    p.err = p.counters + n_vol;
    p.ticket = p.counters + n_vol + 1;
    p.loss_out = loss;
    p.partials = reinterpret_cast<double*>(static_cast<unsigned char*>(workspace) + cb);
    if (grid > p.n_items) grid = p.n_items;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
#if HIPE_RES_SELFVALID
    e = cudaMemsetAsync(workspace, 0xFF, (size_t)need, st);   // every word "unset" (err, ticket: ~0)
#else
    e = cudaMemsetAsync(workspace, 0, cb, st);
#endif
    if (e != cudaSuccess) return (int)e;
    void* args[] = {(void*)&p};
    e = cudaLaunchCooperativeKernel(kern, dim3((unsigned)grid), dim3(res::kThreads), args, smem, st);
    if (e == cudaErrorCooperativeLaunchTooLarge) {
        (void)cudaGetLastError();
        return HIPE_E_UNSUPPORTED;
    }
    return (int)e;
}

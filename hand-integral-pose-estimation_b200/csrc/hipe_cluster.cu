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
// STREAMED shares (kMode 3): a volume that does not fit the shared memory of any cluster (fp32 96^3 =
// 3.4 MiB) keeps the same roles, but nothing stays resident.  The ring is split in two: slots
// [0, nslots_r) are fed from HBM (L2 evict_last) by the producer warp, reduced and handed straight
// back by the REDUCE team; the other slots are fed by a SECOND producer warp that re-reads the same
// stages (L2 evict_first) once the CTA's share of the volume has been reduced, and are consumed by
// the WRITE team.  The first producer starts volume k+2 only when the write team has finished volume
// k, so the reduction of volume k+1 runs under the gradient write of volume k and at most two
// volumes per cluster -- (SMs / C) * 2 * vol_bytes, 106 MB for fp32 96^3 at C = 8, of which the
// already re-read half is marked evict_first -- are live in the 126 MB L2: the second read is served
// on chip, where the one-volume-per-cluster streaming launch (hipe_joint_loss_fused: 2 CTAs per SM,
// 37 volumes = 131 MB in flight) re-fetched 46 % of it from HBM.
//
// WRAPPED visits (kWrap): when a row is not a power-of-two number of 16-byte groups (W = 96 fp32:
// 24), whole-row visits leave a quarter of each team idle.  A stage is then KP = 3 visits of all 256
// threads over 12 288 contiguous bytes = 32 whole rows: thread t's visit q covers group
// (t + 256 q) mod 24 of row (t + 256 q) / 24, so its column differs per visit (kept per visit) and a
// stage starts at y = 0, 32 or 64.
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
constexpr int kMaxClusterStr = 16;    // streamed shares: the largest (non-portable) cluster, one per GPC
constexpr int kNSums = 7;             // S0c Sx Syo Szo S0t Syb Szb
constexpr int kNB = 4;                // scalar records in flight (merge warp -> write team)
constexpr int kPB = 2;                // team partials in flight (reduce team -> merge warp)
constexpr int kThreads = 2 * kC + 64; // two teams + producer warp + merge warp
// 18 warps = 5 on the fullest scheduler partition, each partition has 16384 registers: 5 * 32 * regs
// must stay <= 16384, i.e. <= 102 registers per thread (112 makes the launch impossible: the
// occupancy query answers 0 clusters)
constexpr int kMaxRegs = 96;
// Two write teams (kMode 2, 2-byte logits): 26 warps = 7 on the fullest partition -> <= 73 registers.
constexpr int kMaxRegs2 = 72;
constexpr int kThreads2 = 3 * kC + 64;
// Streamed shares (kMode 3): a second producer warp for the write team's own ring.  19 warps are
// still 5 on the fullest partition, so the 96-register cap stands.
constexpr int kThreads3 = 2 * kC + 96;
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
    int stasync;      // partial exchange by st.async + transaction bytes (1) or plain stores + remote arrives (0)
    double inv_W, inv_H, inv_D;
    int wteams2;      // two write teams taking alternate stages (832 threads, 72 registers)
    int sym;          // both teams reduce AND write, alternating volumes (0: team 0 reduces, team 1 writes)
    int hoist;        // H | SR: every stage starts at row y = 0 (row term = per-volume part + per-stage part)
    int nslots_r;     // streamed shares (kMode 3): slots [0, nslots_r) feed the reduce team, the rest the write team
    int lead;         // streamed shares: stages the first producer may run ahead of the write team (> share)
    int pol1;         // streamed shares: L2 policy of the first read (0 normal, 1 evict_first, 2 evict_last)
    int pf;           // resident form: stages the producer prefetches into L2 when it blocks on a held slot (0: none)
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
    double term[3];           // |c_a - gt_a| * vis_a
    float cn[3];              // normalised coordinates
    float M;                  // volume maximum
    float L;                  // sum of exponentials relative to M
    float pad;
};

template <int MC>
struct SmemT {
    unsigned long long full[kMaxSlots], empty[kMaxSlots];
    unsigned long long red_ready[kPB], red_free[kPB];      // reduce team <-> merge warp
    unsigned long long scal_ready[kNB], scal_free[kNB];    // merge warp <-> write team
    unsigned long long xfull[2];                           // C remote arrivals per volume
    // Streamed shares, pacing only (no data is ordered by them, so plain counters instead of mbarriers:
    // a one-directional mbarrier would lose a phase if its producer ever ran two volumes ahead of its
    // consumer): warps of the reduce team that have finished their volume (kNW per volume), and
    // stages the write team has taken (warp 0).
    unsigned rcount, wprog;
    double wred[kPB][kNW][8];
    float wmax[kPB][kNW];
    double xch[2][MC][8];                                  // [buffer][source rank][7 sums + max]
    Scal scal[kNB];
};
using Smem = SmemT<kMaxCluster>;
using SmemStr = SmemT<kMaxClusterStr>;
static_assert(sizeof(Smem) <= 3072, "14 slots of 16 KB + Smem must fit the 227 KB of one CTA");

// Wait-time profile (diagnostic builds only, -DHIPE_CLU_PROF; scripts/diag_cluster_waits.py): lane 0
// of every warp adds the clock cycles it spent in each class of wait, and its total run time, to a
// global table.
//   0 producer: empty slot      1 reduce: stage landed (full)   2 reduce: red_free
//   3 merge: red_ready          4 merge: peers' partials        5 merge: scal_free
//   6 write: scal_ready         7 write: full (already landed)
//   8 producer total   9 reduce total   10 merge total   11 write total   (warp-cycles)
#ifdef HIPE_CLU_PROF
constexpr int kProfN = 12;
__device__ unsigned long long g_prof[kProfN];
// per-warp register accumulators (three wait classes per role at most), flushed once at role exit:
// an atomic per wait would itself become the bottleneck
// event trace of CTA 0 for volumes kTraceK0 .. kTraceK0+7 (clock64 of one SM):
//   0 first stage issued   1 last stage issued   2 last stage landed (reduce warp 0)   3 fold handed over
//   4 merge: team partial seen   5 merge: published   6 merge: peers seen   7 merge: record ready
//   8 write warp 0: record seen (write starts)   9 write warp 0: last stage stored
constexpr int kTraceK0 = 24, kTraceN = 8, kTraceE = 10;
__device__ long long g_trace[kTraceN][kTraceE];
#define CLU_TRACE(k, ev)                                                                               \
    do {                                                                                               \
        if (blockIdx.x == 0 && (k) >= kTraceK0 && (k) < kTraceK0 + kTraceN && (threadIdx.x & 31) == 0)  \
            g_trace[(k) - kTraceK0][ev] = clock64();                                                   \
    } while (0)
#define CLU_SLOT(cls) ((cls) == 0 || (cls) == 1 || (cls) == 3 || (cls) == 6 ? 0 : ((cls) == 5 ? 2 : 1))
#define CLU_WAIT(cls, call)                                          \
    do {                                                             \
        const long long t0_ = clock64();                             \
        call;                                                        \
        prof_acc_[CLU_SLOT(cls)] += (unsigned long long)(clock64() - t0_); \
    } while (0)
#define CLU_TOTAL(role, c0, c1, c2)                                                             \
    do {                                                                                        \
        if ((threadIdx.x & 31) == 0) {                                                          \
            atomicAdd(&g_prof[role], (unsigned long long)(clock64() - prof_start_));            \
            if ((c0) >= 0) atomicAdd(&g_prof[(c0) < 0 ? 0 : (c0)], prof_acc_[0]);               \
            if ((c1) >= 0) atomicAdd(&g_prof[(c1) < 0 ? 0 : (c1)], prof_acc_[1]);               \
            if ((c2) >= 0) atomicAdd(&g_prof[(c2) < 0 ? 0 : (c2)], prof_acc_[2]);               \
        }                                                                                       \
    } while (0)
#else
#define CLU_WAIT(cls, call) call
#define CLU_TOTAL(role, c0, c1, c2) do { } while (0)
#define CLU_TRACE(k, ev) do { } while (0)
#endif

// How long one try_wait may keep the warp suspended before it returns false (ns): a waiting warp
// then re-issues its poll loop that rarely instead of competing for issue slots with the working
// teams (ncu on the first version: 40 % of all executed instructions were poll loops).  The wake-up
// on completion is immediate either way.
#ifndef HIPE_CLU_SUSPEND_NS
#define HIPE_CLU_SUSPEND_NS 20000
#endif
constexpr uint32_t kSuspendNs = HIPE_CLU_SUSPEND_NS;

// mbarrier waits that trap instead of hanging the GPU if the pipeline logic is ever wrong.  A failed
// try_wait costs up to kSuspendNs (20 us), so kSpinLimit of them are ~10 s: four orders of magnitude
// above any legitimate wait inside a launch that lasts milliseconds, and well inside a test's time-out.
constexpr unsigned kSpinLimit = 1u << 19;
__device__ __forceinline__ void wait_cta(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
#pragma unroll 1
    for (unsigned spin = 0; spin < kSpinLimit; ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity), "r"(kSuspendNs)
            : "memory");
        if (done) return;
    }
    __trap();
}
__device__ __forceinline__ void wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
#pragma unroll 1
    for (unsigned spin = 0; spin < kSpinLimit; ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity), "r"(kSuspendNs)
            : "memory");
        if (done) return;
    }
    __trap();
}
// Build option (-DHIPE_CLU_PEEK=1, off): streamed shares ask one stage early, without blocking, whether
// the next stage has landed, so that a landed stage costs no try_wait round trip in the warp's chain.
// Measured: no difference (scripts/ab_cluster_peek.sh, two builds alternating on one box: fp32 96^3
// 0.833-0.851 ms off, 0.809-0.856 on; fp32 64^3 streamed 0.473-0.495 off, 0.488-0.506 on).
__device__ __forceinline__ uint32_t test_cta(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    return done;
}
#ifndef HIPE_CLU_PEEK
#define HIPE_CLU_PEEK 0
#endif
__device__ __forceinline__ uint32_t mapa(uint32_t local_saddr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_saddr), "r"(rank));
    return remote;
}
__device__ __forceinline__ void st_cluster_f64(uint32_t remote_saddr, double v) {
    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(remote_saddr), "d"(v) : "memory");
}
// 8-byte store into a peer's shared memory that completes 8 bytes of the peer's mbarrier transaction
// count: data and signal travel together through the async proxy, no fence on the sender
__device__ __forceinline__ void st_async_f64(uint32_t remote_saddr, double v, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_saddr),
                 "l"(__double_as_longlong(v)), "r"(remote_bar)
                 : "memory");
}
// arrive on the mbarrier at the same offset in CTA `rank`; releases this thread's earlier
// st.shared::cluster stores at cluster scope
__device__ __forceinline__ void arrive_remote(uint32_t remote_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar) : "memory");
}

__device__ __forceinline__ __nv_bfloat162 u32_as_bf162(uint32_t w) {
    return *reinterpret_cast<__nv_bfloat162*>(&w);
}

__device__ __forceinline__ float rebase_w(float m_from, float m_to) {
    // 2^(fl(m_from*log2e) - fl(m_to*log2e)): the same rounded products the exponentials used
    return (m_from == kNegInf) ? 0.f : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

// Build option (-DHIPE_CLU_POLY=1|2, off): evaluate every other / every exponential of the 2-byte gradient
// pass with ex2_poly2 below.  Measured slower (bf16 64^3 x 2688: 0.618 -> 0.630 / 0.74 ms): the launch is
// bound by the latency of each warp's chain, not by XU throughput, and the polynomial is a longer chain.
#ifndef HIPE_CLU_POLY
#define HIPE_CLU_POLY 0
#endif
// 2^a for a <= ~0 on the FMA / ALU pipes instead of the quarter-rate XU (MUFU.EX2): round-to-nearest
// split a = n + f through the 1.5 * 2^23 constant, a degree-4 minimax polynomial for 2^f on
// [-0.5, 0.5] (relative error 2.7e-6, far below the bf16 output's 2^-9 rounding) in packed FFMA2,
// and n added straight into the exponent field.  Used for a share of the bf16 gradient pass, whose
// two exponentials per 2 bytes of traffic otherwise keep the XU pipe ~80 % busy at HBM speed.
__device__ __forceinline__ float2 ex2_poly2(float2 a) {
    a.x = fmaxf(a.x, -125.f);
    a.y = fmaxf(a.y, -125.f);
    const float2 magic = make_float2(12582912.f, 12582912.f), nmagic = make_float2(-12582912.f, -12582912.f);
    const float2 t = __fadd2_rn(a, magic);
    const float2 nf = __fadd2_rn(t, nmagic);
    const float2 f = __ffma2_rn(nf, make_float2(-1.f, -1.f), a);
    float2 q = __ffma2_rn(f, make_float2(0.009570102207362652f, 0.009570102207362652f),
                          make_float2(0.05591785907745361f, 0.05591785907745361f));
    q = __ffma2_rn(q, f, make_float2(0.240247443318367f, 0.240247443318367f));
    q = __ffma2_rn(q, f, make_float2(0.6931217908859253f, 0.6931217908859253f));
    q = __ffma2_rn(q, f, make_float2(0.9999992847442627f, 0.9999992847442627f));
    return make_float2(__int_as_float(__float_as_int(q.x) + (__float_as_int(t.x) << 23)),
                       __int_as_float(__float_as_int(q.y) + (__float_as_int(t.y) << 23)));
}

// bulk prefetch global -> L2 (no shared-memory slot needed): what the producer cannot load yet
__device__ __forceinline__ void prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

// ring position of a role: slot index and the parity of the slot's current use
struct Ring {
    int slot;
    uint32_t lap;
    uint32_t off;      // slot * slot_bytes
    __device__ __forceinline__ void init() { slot = 0; lap = 0; off = 0; }
    __device__ __forceinline__ void next(int nslots, int slot_bytes) {
        off += (uint32_t)slot_bytes;
        if (++slot == nslots) { slot = 0; off = 0; ++lap; }
    }
    __device__ __forceinline__ void advance(int n, int nslots, int slot_bytes) {
        slot += n;
        while (slot >= nslots) { slot -= nslots; ++lap; }
        off = (uint32_t)slot * (uint32_t)slot_bytes;
    }
};

// fp32 >= 0 (normal or zero) -> fp64 without the conversion unit (F2F.F64.F32 shares the quarter-rate
// XU pipe with MUFU.EX2): re-bias the exponent with integer ops.  0 maps to 2^-127 (5.9e-39), far
// below one ulp of any sum that contains the volume maximum's e = 1; denormals cannot occur
// (ex2.ftz flushes them).  Same helper as hipe_stream_kernel.cuh.
__device__ __forceinline__ double widen_nonneg(float f) {
    const uint32_t b = __float_as_uint(f);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
}

// Sum 8 doubles per lane over the 32 lanes of a warp with 9 double-shuffles instead of 40: each
// step a lane hands the half of its values it will not keep to its partner.  On return every lane
// holds the warp total of value number (lane >> 2) & 7.  Fixed order: bit-reproducible.
__device__ __forceinline__ double warp_sum8(const double (&v)[8], int lane) {
    double w[4], u[2];
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double mine = b4 ? v[i + 4] : v[i], other = b4 ? v[i] : v[i + 4];
        w[i] = mine + __shfl_xor_sync(0xffffffffu, other, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const double mine = b3 ? w[i + 2] : w[i], other = b3 ? w[i] : w[i + 2];
        u[i] = mine + __shfl_xor_sync(0xffffffffu, other, 8);
    }
    double t = (b2 ? u[1] : u[0]) + __shfl_xor_sync(0xffffffffu, b2 ? u[0] : u[1], 4);
    t += __shfl_xor_sync(0xffffffffu, t, 2);
    t += __shfl_xor_sync(0xffffffffu, t, 1);
    return t;
}

// Thread roles inside the CTA (kThreads = 2*kC + 64):
//   [0, kC)          REDUCE team
//   [kC, 2kC)        WRITE team
//   warp 2kC/32      producer (one elected lane)
//   warp 2kC/32 + 1  merge warp
// kPass != 0: a geometry known at compile time -- kPass bytes per visit (= 16 bytes x active team
// threads), stages = slots of 4 visits, every stage starting at y = 0 (H | SR): 4096 for 64^3 and
// 32^3 (fp32 and bf16: all 256 team threads active), 3136 for 56^3 (196 threads).  The byte offsets
// become immediates and the row bookkeeping collapses.  The launch is instruction-bound (one CTA per
// SM, every role a dependency chain), so the instructions per stage are what is tuned.  kPass == 0
// takes any geometry pick_geom accepts.
template <typename T, int kPass, int kMode, int KP, bool kWrap>
__global__ void __maxnreg__(kMode == 2 ? kMaxRegs2 : kMaxRegs) k_cluster(const Params p) {
    constexpr bool kSym = kMode == 1;
    constexpr bool kStr = kMode == 3;                   // streamed shares: two rings, nothing resident
    constexpr int kWT = kMode == 2 ? 2 : 1;             // write teams
    constexpr int kTeamThreads = (1 + kWT) * kC;
    static_assert(!kWrap || (kPass == 16 * kC), "wrapped visits: every team thread takes one vector per visit");
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr bool kFix = kPass != 0;
    const int pass_bytes = kFix ? kPass : p.pass_bytes;
    const int stage_bytes = kFix ? KP * kPass : p.stage_bytes;
    const int slot_bytes = kFix ? KP * kPass : p.slot_bytes;    // the specialised sizes are multiples of 128
    const bool hoist = kWrap ? false : (kFix ? true : (p.hoist != 0));   // H | SR: every stage starts at row y = 0
    // streamed shares: the reduce team's ring is slots [0, nR), the write team's [nR, nR + nW)
    const int nR = kStr ? p.nslots_r : p.nslots;
    const int nW = kStr ? p.nslots - p.nslots_r : p.nslots;
    const int wbase = kStr ? p.nslots_r : 0;
    constexpr int kMC = kWrap ? kMaxClusterStr : kMaxCluster;   // (the one geometry whose volumes need 16 CTAs)
    using SmemK = SmemT<kMC>;
    SmemK& sm = *reinterpret_cast<SmemK*>(smem_raw + (size_t)p.nslots * slot_bytes);
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
            mbar_init(smem_u32(&sm.scal_free[s]), kNW * kWT);
        }
        // st.async exchange: one local arrive.expect_tx per volume, the peers' stores complete the bytes;
        // plain-store exchange: one remote arrive per peer
        mbar_init(smem_u32(&sm.xfull[0]), p.stasync ? 1 : C);
        mbar_init(smem_u32(&sm.xfull[1]), p.stasync ? 1 : C);
        sm.rcount = 0;
        sm.wprog = 0;
        mbar_fence_init();
        fence_proxy_async();
    }
    __syncthreads();
    // every CTA's barriers exist before any peer arrives on them
    cluster_sync_all();
#ifdef HIPE_CLU_PROF
    const long long prof_start_ = clock64();
    unsigned long long prof_acc_[3] = {0ull, 0ull, 0ull};
#endif

    if (kStr && threadIdx.x >= kTeamThreads + 64) {
        // -------------------- second producer warp (streamed shares) --------------------
        // Re-reads the share for the write team, from L2: volume k's stages are requested only after
        // this CTA's reduce team has consumed them all (rcount), and as far ahead of the write team as
        // its ring allows -- the prefetch runs under the cross-CTA merge of the volume.
        if ((threadIdx.x & 31) == 0) {
            const uint64_t pol = policy_evict_first();
            Ring r;
            r.init();
            for (long long k = 0; k < nv; ++k) {
                const long long vol = cluster_id + k * p.n_clusters;
                const unsigned char* src = p.in + vol * p.vol_bytes + (long long)sv0 * stage_bytes;
                {
                    // (only so that the re-read finds the lines in L2: the first read of the share is done)
                    const unsigned need = (unsigned)(k + 1) * (unsigned)kNW;
                    unsigned spin = 0;
                    while ((int)(*(volatile unsigned*)&sm.rcount - need) < 0) {
                        __nanosleep(100);
                        if (++spin > (1u << 26)) __trap();
                    }
                }
                for (int s = 0; s < nst; ++s) {
                    if (r.lap > 0) wait_cta(smem_u32(&sm.empty[wbase + r.slot]), (r.lap - 1) & 1);
                    const uint32_t bar = smem_u32(&sm.full[wbase + r.slot]);
                    mbar_arrive_expect_tx(bar, (uint32_t)stage_bytes);
                    tma_load_1d(ring + (uint32_t)wbase * (uint32_t)slot_bytes + r.off, src, (uint32_t)stage_bytes, bar, pol);
                    src += stage_bytes;
                    r.next(nW, slot_bytes);
                }
            }
        }
        return;
    }

    if (threadIdx.x >= kTeamThreads + 32) {
        // ------------------------------- merge warp -------------------------------
        // Lane l (< 7) owns sum number l through both merges, lane 7 the maximum: every merge is a short
        // sequential sum in a fixed order per lane (no shuffle trees on this warp's critical path).
        const int lane = threadIdx.x & 31;
        const int li = lane < kNSums ? lane : kNSums - 1;       // the word this lane reads
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
            if (p.stasync && lane == 0) mbar_arrive_expect_tx(smem_u32(&sm.xfull[xb]), C * 8u * (kNSums + 1));
            // ---- this CTA's partial: the reduce team's kNW warp partials, in warp order ----
            CLU_WAIT(3, wait_cta(smem_u32(&sm.red_ready[sb]), (uint32_t)(k / kPB) & 1));
            CLU_TRACE(k, 4);
            float wmx[kNW];
            double wv[kNW];
#pragma unroll
            for (int w = 0; w < kNW; ++w) {
                wmx[w] = sm.wmax[sb][w];
                wv[w] = sm.wred[sb][w][li];
            }
            {
                // hand the buffer back only once the loads above have returned (see
                // mbar_arrive_after_loads; a warp's scoreboard covers all its lanes)
                uint32_t dep = 0;
#pragma unroll
                for (int w = 0; w < kNW; ++w) dep ^= __float_as_uint(wmx[w]) ^ (uint32_t)__double2loint(wv[w]);
                __syncwarp();
                if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.red_free[sb]), dep);
            }
            float cm = wmx[0];
#pragma unroll
            for (int w = 1; w < kNW; ++w) cm = fmaxf(cm, wmx[w]);
            double mine = 0.0;
#pragma unroll
            for (int w = 0; w < kNW; ++w) mine = fma((double)rebase_w(wmx[w], cm), wv[w], mine);
            if (lane == kNSums) mine = (double)cm;
            // ---- publish to every CTA of the cluster (self included): lane l stores word l ----
            if (p.stasync) {
                if (lane <= kNSums) {
                    const uint32_t src = smem_u32(&sm.xch[xb][crank][lane]), bar = smem_u32(&sm.xfull[xb]);
                    for (uint32_t r = 0; r < C; ++r) st_async_f64(mapa(src, r), mine, mapa(bar, r));
                }
            } else {
                if (lane <= kNSums) {
                    const uint32_t src = smem_u32(&sm.xch[xb][crank][lane]);
                    for (uint32_t r = 0; r < C; ++r) st_cluster_f64(mapa(src, r), mine);
                }
                __syncwarp();       // the arriving lanes' release covers the other lanes' stores
                if (lane < (int)C) arrive_remote(mapa(smem_u32(&sm.xfull[xb]), (uint32_t)lane));
            }
            CLU_TRACE(k, 5);
            // ---- all C partials of this volume, merged in rank order ----
            CLU_WAIT(4, wait_cluster(smem_u32(&sm.xfull[xb]), (uint32_t)(k >> 1) & 1));
            CLU_TRACE(k, 6);
            // all loads first (independent), then the short dependent chains
            float mr[kMC];
            double vr[kMC];
#pragma unroll
            for (int r = 0; r < kMC; ++r) {
                const bool on = (uint32_t)r < C;
                mr[r] = on ? (float)sm.xch[xb][r][kNSums] : kNegInf;
                vr[r] = on ? sm.xch[xb][r][li] : 0.0;
            }
            float M = mr[0];
#pragma unroll
            for (int r = 1; r < kMC; ++r) M = fmaxf(M, mr[r]);
            double a = 0.0;
#pragma unroll
            for (int r = 0; r < kMC; ++r) a = fma((double)rebase_w(mr[r], M), vr[r], a);
            // loss.py:39-41 (expectations) and :54-56 (normalisation), in fp64, rounded once.
            // Lane t (< 3) finishes axis t: x, y, z.  Sums: 0 S0c 1 Sx 2 Syo 3 Szo 4 S0t 5 Syb 6 Szb.
            const double a0 = __shfl_sync(0xffffffffu, a, 0), a4 = __shfl_sync(0xffffffffu, a, 4);
            const double num_c = __shfl_sync(0xffffffffu, a, (lane < 3) ? 1 + lane : lane);
            const double num_t = __shfl_sync(0xffffffffu, a, (lane == 1 || lane == 2) ? 4 + lane : lane);
            const double ic = 1.0 / a0, itt = 1.0 / a4;
            const double inv_dim = lane == 0 ? p.inv_W : (lane == 1 ? p.inv_H : p.inv_D);
            const double ex_ = num_c * ic + (lane == 0 ? 0.0 : num_t * itt);
            const float cn = (float)(ex_ * inv_dim - 0.5);
            // loss.py:86-89 on this volume's three coordinates, in the targets' precision
            const double dlt = (double)cn - tk;
            const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
            const float g1 = (float)(sg * wk * p.norm);
            if (k >= kNB) CLU_WAIT(5, wait_cta(smem_u32(&sm.scal_free[rb]), (uint32_t)((k / kNB) - 1) & 1));
            Scal& rec = sm.scal[rb];
            if (lane < 3) {
                rec.E[lane] = ex_;
                rec.K[lane] = (double)g1 * inv_dim * ic;     // g_a / dim_a / L
                rec.term[lane] = fabs(dlt) * wk;
                rec.cn[lane] = cn;
            }
            if (lane == 0) { rec.M = M; rec.L = (float)a0; }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.scal_ready[rb]));
            CLU_TRACE(k, 7);
            // (the volume's outputs go to global memory from the write team: a global store here would
            // sit in front of this warp's next cluster-scope release and stall it)
        }
        CLU_TOTAL(10, 3, 4, 5);
        return;
    }

    if (threadIdx.x >= kTeamThreads) {
        // ------------------------------- producer warp -------------------------------
        if ((threadIdx.x & 31) == 0) {
            // streamed shares: the lines are wanted again within two volumes' time
            const uint64_t pol = !kStr || p.pol1 == 1 ? policy_evict_first()
                                                      : (p.pol1 == 2 ? policy_evict_last() : policy_evict_normal());
            Ring r;
            r.init();
            // Resident form: a slot is held until its gradient is written, so the last (share - spare)
            // stages of the next volume can only be requested once the write team frees slots, and their
            // HBM latency (2-3.5 us under load) sits in the per-volume loop write(k) -> load -> reduce ->
            // merge -> write(k+1).  When the producer is about to block on a held slot it therefore first
            // asks the L2 for the next p.pf stages (bulk prefetch, no slot needed): the late loads take an
            // L2 hit.
            long long pk = 0;      // prefetch cursor: volume, stage (never behind the load cursor)
            int ps = 0;
            auto prefetch_from = [&](long long k, int s) {     // stages (k, s) .. (k, s) + p.pf - 1, each once
                if (pk < k || (pk == k && ps < s)) { pk = k; ps = s; }
                long long tk = k;
                int ts = s + p.pf;
                while (ts >= nst) { ts -= nst; ++tk; }
                while (pk < nv && (pk < tk || (pk == tk && ps < ts))) {
                    const long long pvol = cluster_id + pk * p.n_clusters;
                    prefetch_l2(p.in + pvol * p.vol_bytes + (long long)(sv0 + ps) * stage_bytes, (uint32_t)stage_bytes);
                    if (++ps == nst) { ps = 0; ++pk; }
                }
            };
            for (long long k = 0; k < nv; ++k) {
                const long long vol = cluster_id + k * p.n_clusters;
                const unsigned char* src = p.in + vol * p.vol_bytes + (long long)sv0 * stage_bytes;
                for (int s = 0; s < nst; ++s) {
                    if constexpr (!kStr) {
                        if (p.pf > 0 && r.lap > 0 && !test_cta(smem_u32(&sm.empty[r.slot]), (r.lap - 1) & 1)) prefetch_from(k, s);
                    }
                    if constexpr (kStr) {
                        // Streamed shares: stay at most `lead` stages ahead of the write team -- what has
                        // been read once and not yet again is what the L2 must hold (lead x stage bytes x
                        // CTAs).  lead > share, so the wait only ever concerns volumes whose reduction
                        // needs nothing more from this warp: no cycle.
                        const unsigned q0 = (unsigned)(k * nst + s);
                        if (q0 >= (unsigned)p.lead) {
                            unsigned spin = 0;
                            while ((int)(*(volatile unsigned*)&sm.wprog + (unsigned)p.lead - q0) <= 0) {
                                __nanosleep(100);
                                if (++spin > (1u << 26)) __trap();
                            }
                        }
                    }
                    if (r.lap > 0) CLU_WAIT(0, wait_cta(smem_u32(&sm.empty[r.slot]), (r.lap - 1) & 1));
                    if (s == 0) CLU_TRACE(k, 0);
                    if (s == nst - 1) CLU_TRACE(k, 1);
                    const uint32_t bar = smem_u32(&sm.full[r.slot]);
                    mbar_arrive_expect_tx(bar, (uint32_t)stage_bytes);
                    tma_load_1d(ring + r.off, src, (uint32_t)stage_bytes, bar, pol);
                    src += stage_bytes;
                    r.next(nR, slot_bytes);
                }
            }
        }
        CLU_TOTAL(8, 0, -1, -1);
        return;
    }

    const bool is_writer = threadIdx.x >= kC;
    const int wt = is_writer ? (int)(threadIdx.x / kC) - 1 : 0;   // which write team (kMode 2: 0 or 1)
    const int tid = threadIdx.x % kC;                             // team-local
    const int lane = tid & 31, warp = tid >> 5;
    const bool active = kFix ? (kPass / 16 >= kC || tid < kPass / 16) : (tid < p.nact);
    const int cg = active ? tid % p.G : 0;
    const int rr = active ? tid / p.G : 0;
    const int x0 = cg * VEC;
    const int rowbase0 = sv0 * p.SR;              // first row of the share
    const int zb0 = rowbase0 / p.H, yb0 = rowbase0 - zb0 * p.H;
    const uint32_t lds_base = ring + 16u * tid;

    // streamed shares: has the next stage of this team's ring landed already?  (asked one stage early)
    constexpr bool kPeek = kStr && HIPE_CLU_PEEK;
    uint32_t peek = 0;
    // =============================== reduction of one volume ===============================
    // (the team's share of volume k: the stages at ring position r, which is advanced past them)
    auto reduce_volume = [&](long long k, Ring& r) {
        {
            const double dzd = (double)(p.SR / p.H);      // hoist: z advance per stage
            const unsigned sb = (unsigned)(k % kPB);
            float m = kNegInf, mL = kNegInf, tm = kNegInf;   // offset in use, fl(m*log2e), true maximum
            float2 c2[KP][VEC / 2];
#pragma unroll
            for (int q = 0; q < KP; ++q)
#pragma unroll
                for (int j = 0; j < VEC / 2; ++j) c2[q][j] = make_float2(0.f, 0.f);
            double s0t = 0.0, syb = 0.0, szb = 0.0;
            int zb = zb0, yb = yb0;                 // general geometry
            double zbd = (double)zb0;               // hoist: the stage's first z as a double, y is 0
            auto fetch = [&](uint4 (&dst)[KP]) {
                if (!(kPeek && peek)) CLU_WAIT(1, wait_cta(smem_u32(&sm.full[r.slot]), r.lap & 1));
                if (active) {
                    const uint32_t sbase = lds_base + r.off;
#pragma unroll
                    for (int q = 0; q < KP; ++q) dst[q] = lds128(sbase + q * pass_bytes);
                }
                if constexpr (kStr) {
                    // streamed shares: nothing stays resident -- the slot goes back to the producer as
                    // soon as the loads have returned
                    uint32_t dep = 0;
#pragma unroll
                    for (int q = 0; q < KP; ++q) dep ^= dst[q].x;
                    if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.empty[r.slot]), active ? dep : 0u);
                }
                r.next(nR, slot_bytes);
                if constexpr (kPeek) peek = test_cta(smem_u32(&sm.full[r.slot]), r.lap & 1);   // the stage after this one
            };
            auto compute = [&](const uint4 (&raw)[KP]) {
                if (active) {
                    // the stage's maximum straight from the raw words (2-byte logits: packed bf16x2
                    // maxima, half the instructions and nothing unpacked yet)
                    float smax;
                    if constexpr (VEC == 8) {
                        __nv_bfloat162 mx = u32_as_bf162(raw[0].x);
#pragma unroll
                        for (int q = 0; q < KP; ++q) {
                            mx = __hmax2(mx, __hmax2(__hmax2(u32_as_bf162(raw[q].x), u32_as_bf162(raw[q].y)),
                                                     __hmax2(u32_as_bf162(raw[q].z), u32_as_bf162(raw[q].w))));
                        }
                        smax = fmaxf(__low2float(mx), __high2float(mx));
                    } else {
                        smax = kNegInf;
#pragma unroll
                        for (int q = 0; q < KP; ++q)
                            smax = fmaxf(smax, fmaxf(fmaxf(__uint_as_float(raw[q].x), __uint_as_float(raw[q].y)),
                                                     fmaxf(__uint_as_float(raw[q].z), __uint_as_float(raw[q].w))));
                    }
                    tm = fmaxf(tm, smax);
                    // LAZY re-basing: the offset m only moves when a value exceeds it by more than
                    // kLazy (e <= exp(kLazy) ~ 9e6 stays far inside fp32); the fold re-bases once.
                    if (smax > m + kLazy) {
                        const float f = rebase_w(m, smax);
                        const float2 f2 = make_float2(f, f);
#pragma unroll
                        for (int q = 0; q < KP; ++q)
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
                    float2 rs2 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int q = 0; q < KP; ++q) {
                        float v[VEC];
                        V16::unpack(raw[q], v);
                        float2 e2[VEC / 2];
#pragma unroll
                        for (int j = 0; j < VEC / 2; ++j) {
                            const float2 a2 = __ffma2_rn(make_float2(v[2 * j], v[2 * j + 1]), l2e2, noff2);
                            e2[j] = make_float2(ex2(a2.x), ex2(a2.y));
                            c2[q][j] = __fadd2_rn(c2[q][j], e2[j]);
                        }
                        float2 r2 = __fadd2_rn(e2[0], e2[1]);
                        if constexpr (VEC == 8) r2 = __fadd2_rn(r2, __fadd2_rn(e2[2], e2[VEC / 2 - 1]));
                        rs2 = (q == 0) ? r2 : __fadd2_rn(rs2, r2);
                    }
                    const double dt = widen_nonneg(rs2.x + rs2.y);
                    s0t += dt;
                    if (hoist) {
                        szb = fma(zbd, dt, szb);
                    } else {
                        syb = fma((double)yb, dt, syb);
                        szb = fma((double)zb, dt, szb);
                    }
                }
                // next stage's base row (uniform): SR | H or H | SR, so one of the two moves
                if (hoist) {
                    zbd += dzd;
                } else {
                    yb += p.SR;
                    while (yb >= p.H) { yb -= p.H; ++zb; }
                }
            };
            // Stage s+1 is fetched into registers before the arithmetic of stage s: two stages'
            // dependency chains overlap in every warp.
            // (4-byte logits only: with 8 elements per vector the second buffer spills.)
            if constexpr (VEC == 4) {
                uint4 rawA[KP], rawB[KP];
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
                uint4 rawA[KP];
                for (int s = 0; s < nst; ++s) {
                    fetch(rawA);
                    compute(rawA);
                }
            }
            if constexpr (kStr) {
                // the whole share has been requested and consumed: the second producer may re-read it
                if (lane == 0) atomicAdd(&sm.rcount, 1u);
            }
            if (warp == 0) CLU_TRACE(k, 2);
            // ---- fold the thread's accumulators, sum over the warp, hand to the merge warp ----
            double q8[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) q8[i] = 0.0;
            if (active) {
                double col[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) col[j] = 0.0;
#pragma unroll
                for (int q = 0; q < KP; ++q) {
                    // row offsets of this thread's visit q inside a stage
                    const int rho = kWrap ? (tid + q * kC) / p.G : rr + q * p.R;
                    const int zo = rho / p.H, yo = rho - zo * p.H;
                    const int xq = kWrap ? ((tid + q * kC) % p.G) * VEC : x0;   // wrapped visits: the column moves too
                    double row = 0.0;
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        const double cd = widen_nonneg((j & 1) ? c2[q][j / 2].y : c2[q][j / 2].x);
                        row += cd;
                        if constexpr (kWrap) q8[1] = fma((double)(xq + j), cd, q8[1]);
                        else col[j] += cd;
                    }
                    if constexpr (kWrap) q8[0] += row;
                    q8[2] = fma((double)yo, row, q8[2]);
                    q8[3] = fma((double)zo, row, q8[3]);
                }
                if constexpr (!kWrap) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) {
                        q8[0] += col[j];
                        q8[1] = fma((double)(x0 + j), col[j], q8[1]);
                    }
                }
                q8[4] = s0t; q8[5] = syb; q8[6] = szb;
            }
            // from the thread's lazy offset straight onto the warp's true maximum (weight <= 1)
            const float wm = warp_max(active ? tm : kNegInf);
            const double f = active ? (double)rebase_w(m, wm) : 0.0;
#pragma unroll
            for (int i = 0; i < kNSums; ++i) q8[i] *= f;
            const double tot = warp_sum8(q8, lane);
            if (k >= kPB) CLU_WAIT(2, wait_cta(smem_u32(&sm.red_free[sb]), (uint32_t)((k / kPB) - 1) & 1));
            if ((lane & 3) == 0) sm.wred[sb][warp][lane >> 2] = tot;
            if (lane == 0) sm.wmax[sb][warp] = wm;
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&sm.red_ready[sb]));
            if (warp == 0) CLU_TRACE(k, 3);
        }
    };

    // ============================== gradient write of one volume ==============================
    unsigned wcount = 0;
    auto write_volume = [&](long long k, Ring& r) {
        // row offsets of this thread's kKP visits inside a stage, as exact small floats
        float yof[KP], zof[KP];
#pragma unroll
        for (int q = 0; q < KP; ++q) {
            const int rho = kWrap ? (tid + q * kC) / p.G : rr + q * p.R;
            const int zo = rho / p.H;
            zof[q] = (float)zo;
            yof[q] = (float)(rho - zo * p.H);
        }
        const float Hf = (float)p.H, SRf = (float)p.SR, dzf = (float)(p.SR / p.H);
        const float2 l2e2 = make_float2(kLog2e, kLog2e);
        const long long vol = cluster_id + k * p.n_clusters;
        const unsigned rb = (unsigned)(k % kNB);
        CLU_WAIT(6, wait_cta(smem_u32(&sm.scal_ready[rb]), (uint32_t)(k / kNB) & 1));
        if (warp == 0) CLU_TRACE(k, 8);
        const Scal& rec = sm.scal[rb];
        const double kxd = rec.K[0], Ex = rec.E[0], Ey = rec.E[1], Ez = rec.E[2];
        const float ky = (float)rec.K[1], kz = (float)rec.K[2];
        const float ML = __fmul_rn(rec.M, kLog2e);
        uint32_t dep_out = 0;
        if (crank == 0 && warp == 0 && wt == 0) {
            // the volume's outputs (rank 0's first write warp: lanes 0-2 one axis each)
            if (lane < 3) {
                const double term = rec.term[lane], ea = rec.E[lane];
                const float cn = rec.cn[lane];
                dep_out = (uint32_t)__double2loint(term) ^ __float_as_uint(cn);
                if (p.loss_terms) p.loss_terms[vol * 3 + lane] = term;
                if (p.coords) p.coords[vol * 3 + lane] = cn;
                if (p.expect) p.expect[vol * 3 + lane] = ea;
            }
            if (lane == 0 && p.stats) {
                const float L = rec.L;
                dep_out ^= __float_as_uint(L);
                p.stats[vol * 2 + 0] = rec.M;
                p.stats[vol * 2 + 1] = L;
            }
            dep_out = __shfl_sync(0xffffffffu, dep_out, 0) ^ __shfl_sync(0xffffffffu, dep_out, 1);
        }
        {
            // the record slot goes back to the merge warp only once these loads have returned
            const uint32_t dep = (uint32_t)__double2loint(kxd) ^ (uint32_t)__double2loint(Ex) ^
                                 (uint32_t)__double2loint(Ey) ^ (uint32_t)__double2loint(Ez) ^
                                 __float_as_uint(ky) ^ __float_as_uint(kz) ^ __float_as_uint(ML) ^ dep_out;
            __syncwarp();
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.scal_free[rb]), dep);
        }
        const float2 nML2 = make_float2(-ML, -ML);
        // column terms kx (x - Ex): one set per thread, or one per visit when the visits wrap
        constexpr int kAX = kWrap ? KP : 1;
        float2 ax2[kAX][VEC / 2];
#pragma unroll
        for (int q = 0; q < kAX; ++q) {
            const int xq = kWrap ? ((tid + q * kC) % p.G) * VEC : x0;
#pragma unroll
            for (int j = 0; j < VEC / 2; ++j)
                ax2[q][j] = make_float2((float)(kxd * ((double)(xq + 2 * j) - Ex)), (float)(kxd * ((double)(xq + 2 * j + 1) - Ex)));
        }
        // (idx - E) must keep its ABSOLUTE accuracy: E split into hi + lo floats, the row index an
        // exact small float, so (y - Eh) is exact near the peak and - El restores the rest
        const float Eyh = (float)Ey, Eyl = (float)(Ey - (double)Eyh);
        const float Ezh = (float)Ez, Ezl = (float)(Ez - (double)Ezh);
        // hoist (every stage starts at y = 0): the visit's row term is A[q] + B(stage),
        //   A[q] = ky (yo_q - Ey) + kz zo_q      per volume,     B = kz (z_stage - Ez)   per stage
        float rowA[KP];
#pragma unroll
        for (int q = 0; q < KP; ++q) rowA[q] = fmaf(kz, zof[q], ky * ((yof[q] - Eyh) - Eyl));
        float ybf = (float)yb0, zbf = (float)zb0;
        unsigned char* gdst = p.out + vol * p.vol_bytes + (long long)sv0 * stage_bytes + 16ll * tid;

        auto fetch = [&](uint4 (&dst)[KP]) {
            // The stage landed long ago: the reduce team observed its `full` barrier, and the chain
            // reduce team -> red_ready -> merge warp -> scal_ready -> this thread (each an mbarrier
            // release / acquire pair) orders the TMA's bytes before these loads.  Re-checking the
            // barrier here cost 14 % of the write team's time.
            // (Streamed shares: the slot is this team's own re-read, so its barrier is the only order.)
            if constexpr (kStr) {
                if (!(kPeek && peek)) CLU_WAIT(7, wait_cta(smem_u32(&sm.full[wbase + r.slot]), r.lap & 1));
            }
            if (active) {
                const uint32_t sbase = lds_base + (uint32_t)wbase * (uint32_t)slot_bytes + r.off;
#pragma unroll
                for (int q = 0; q < KP; ++q) dst[q] = lds128(sbase + q * pass_bytes);
            }
            // the slot goes back to the producer only once the loads have RETURNED
            uint32_t dep = 0;
#pragma unroll
            for (int q = 0; q < KP; ++q) dep ^= dst[q].x;
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.empty[wbase + r.slot]), active ? dep : 0u);
            if constexpr (kStr) {
                // progress of the write team, for the first producer's throttle (a pace, not an order)
                if (warp == 0 && lane == 0) *(volatile unsigned*)&sm.wprog = ++wcount;
            }
            r.next(nW, slot_bytes);
            if constexpr (kPeek) peek = test_cta(smem_u32(&sm.full[wbase + r.slot]), r.lap & 1);
        };
        auto compute = [&](const uint4 (&raw)[KP]) {
            if (active) {
                const float dzs = (zbf - Ezh) - Ezl;             // z_stage - Ez, absolute accuracy kept
                const float kzt = kz * dzs;                      // wrapped visits: a stage lies in one z
#pragma unroll
                for (int q = 0; q < KP; ++q) {
                    float rtv;
                    if (kWrap) {
                        rtv = fmaf(ky, ((ybf + yof[q]) - Eyh) - Eyl, kzt);
                    } else if (hoist) {
                        rtv = fmaf(kz, dzs, rowA[q]);
                    } else {
                        const float y = ybf + yof[q], z = zbf + zof[q];
                        rtv = fmaf(ky, (y - Eyh) - Eyl, kz * ((z - Ezh) - Ezl));
                    }
                    float v[VEC], o[VEC];
                    V16::unpack(raw[q], v);
                    const float2 rt2 = make_float2(rtv, rtv);
#pragma unroll
                    for (int j = 0; j < VEC / 2; ++j) {
                        const float2 a2 = __ffma2_rn(make_float2(v[2 * j], v[2 * j + 1]), l2e2, nML2);
                        const float2 w2 = __fadd2_rn(ax2[kWrap ? q : 0][j], rt2);
                        float2 e2;
#if HIPE_CLU_POLY
                        if (VEC == 8 && (HIPE_CLU_POLY == 2 || (q & 1))) e2 = ex2_poly2(a2);
                        else
#endif
                        e2 = make_float2(ex2(a2.x), ex2(a2.y));
                        const float2 o2 = __fmul2_rn(e2, w2);
                        o[2 * j] = o2.x;
                        o[2 * j + 1] = o2.y;
                    }
                    stg_stream(gdst + q * pass_bytes, V16::pack(o));
                }
            }
            gdst += stage_bytes;
            if (hoist) {
                zbf += dzf;
            } else {
                ybf += SRf;
                while (ybf >= Hf) { ybf -= Hf; zbf += 1.f; }
            }
        };
        if constexpr (kWT == 1) {
            uint4 rawA[KP], rawB[KP];
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
            // two write teams: alternate stages of the same volume (each frees the slots it wrote)
            uint4 rawA[KP];
            for (int s = 0; s < nst; ++s) {
                if ((s & 1) == wt) {
                    fetch(rawA);
                    compute(rawA);
                } else {
                    r.next(p.nslots, slot_bytes);
                    gdst += stage_bytes;
                    if (hoist) {
                        zbf += dzf;
                    } else {
                        ybf += SRf;
                        while (ybf >= Hf) { ybf -= Hf; zbf += 1.f; }
                    }
                }
            }
        }
        if (warp == 0) CLU_TRACE(k, 9);
    };

    // (wcount: stages write warp 0 has taken so far, all volumes)
    // Also built and dropped (bf16 64^3, resident form): the reduce team leaving its exponentials IN PLACE
    // of the logits (bf16, relative to a warp-uniform offset kept per slot) so that the write team
    // scales them instead of exponentiating again -- one MUFU.EX2 per element instead of two.  Parity-
    // green (one more bf16 rounding, inside the 1-ulp gate) and the gradient write of a volume did drop
    // from 5.0 to 4.1 us, but the launch did not move (0.531 vs 0.531 ms x 2688; 2.18 vs 2.19-2.20 x
    // 10752): the period per volume (6.6 us) is the loop write(k) frees 2 slots -> last TMA of volume k+1
    // -> 2.2-3.4 us until it lands -> last stage + fold 0.8 -> merge incl. the slowest peer 1.8 ->
    // write(k+1), not the write team's arithmetic.
    // Team roles.  kMode 0 (what ships): team 0 reduces every volume, team 1 writes every volume.
    // Two other layouts were built to relieve the 2-byte gradient pass (the write team is > 90 % busy
    // there) and both lost on every shape; they compile only with -DHIPE_CLU_EXPERIMENTS and are
    // selected by HIPE_CLUSTER_SYM=1 / HIPE_CLUSTER_WT2=1:
    //   kMode 1, symmetric: each team takes every other volume and does both passes on it, so that two
    //     gradient writes could overlap.  bf16 64^3 x 2688: 0.641 vs 0.562 ms; fp32 64^3: 0.631 vs 0.487:
    //     a team idles through its own volume's merge, and the ring never holds enough of the next
    //     volume for two writes to overlap much.
    //   kMode 2, one reduce team + two write teams on alternate stages (832 threads, hence 72
    //     registers): bf16 64^3 x 2688: 0.663 vs 0.530 ms; fp32 64^3: 0.796 vs 0.482: the spills and the
    //     extra warps on the XU pipe (already 72 % busy for bf16, ncu r2) cost more than they hide.
    if constexpr (!kSym) {
        Ring r;
        r.init();
        if (!is_writer) {
            for (long long k = 0; k < nv; ++k) reduce_volume(k, r);
            CLU_TOTAL(9, 1, 2, -1);
            return;
        }
        for (long long k = 0; k < nv; ++k) write_volume(k, r);
    } else {
        const int team = is_writer ? 1 : 0;
        Ring r;
        r.init();
        for (long long k = 0; k < nv; ++k) {
            if ((int)(k & 1) == team) {
                Ring rw = r;
                reduce_volume(k, r);
                write_volume(k, rw);
            } else {
                r.advance(nst, p.nslots, slot_bytes);
            }
        }
    }
    // The loss scalar without a second launch: rank 0's first warps of the writing teams are the
    // writers of loss_terms, so the last of them to finish sums the terms (fixed order: lane-strided
    // partials, then the shuffle tree) -- loss.py:88-89.
    if (crank == 0 && warp == 0 && wt == 0 && p.loss_out && p.loss_terms) {
        __threadfence();
        __syncwarp();
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(p.ticket, 1u);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t == (unsigned)p.n_clusters * (kSym ? 2u : 1u) - 1u) {
            __threadfence();
            double acc = 0.0;
            for (long long i = lane; i < p.n_vol * 3; i += 32) acc += __ldcg(p.loss_terms + i);
            acc = warp_sum(acc);
            if (lane == 0) p.loss_out[0] = acc * p.norm;
        }
    }
    CLU_TOTAL(11, 6, 7, -1);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct Geom {
    int G, R, SR, stage_bytes, pass_bytes, nact, NSv, slot_bytes, nslots, C, max_share;
    int wrap;        // wrapped visits: stages of 3 x 4096 bytes = whole rows (W = 96 fp32)
    int streamed;    // the share does not fit the ring: two rings, second read through L2
    int nslots_r;    // streamed: slots of the reduce team's ring
    int lead;        // streamed: stages the first read may run ahead of the second (0: resident form)
};
constexpr int kWrapKP = 3;
constexpr int kStrMargin = 13;   // streamed shares: default lead = share + this many stages
constexpr int kStrMinMargin = 9;
// what the 126 MB L2 serves of a read-once / reuse-once stream while the gradient stores pass through it
// (ncu DRAM read bytes against the lead, profiles/r2_configs.md section 5)
constexpr double kStrL2Bytes = 42e6;

static bool pick_geom(int dtype, long long n_vol, int D, int H, int W, Geom* g) {
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    const long long row_b = (long long)W * esz;
    if (row_b % 16 != 0 || row_b / 16 > kC) return false;
    const int G = (int)(row_b / 16);
    const long long NR = (long long)D * H;
    if (NR >= (1ll << 30)) return false;
    bool found = false;
    g->wrap = 0; g->streamed = 0; g->nslots_r = 0; g->lead = 0;
    // Wrapped visits: whole-row visits would idle part of each team (R * G < 256) but 3 visits of all
    // 256 threads are a whole number of rows that divides H (a stage never straddles two z).
    {
        const long long sb = (long long)kWrapKP * kC * 16;
        const long long nsv = (sb % row_b == 0) ? NR / (sb / row_b) : 0;
        const long long ring = (kSmemMax - (long long)sizeof(Smem)) / sb;
        // (built for 4-byte logits and streamed shares only: the one shape that needs it is fp32 96^3)
        if (dtype == HIPE_F32 && kC % G != 0 && sb % row_b == 0 && H % (sb / row_b) == 0 &&
            (nsv + kMaxCluster - 1) / kMaxCluster + kMinSpare > ring && !getenv("HIPE_CLUSTER_GENERIC")) {
            g->G = G; g->R = 0; g->SR = (int)(sb / row_b);
            g->pass_bytes = kC * 16;
            g->stage_bytes = (int)sb;
            g->nact = kC;
            g->NSv = (int)(NR / g->SR);
            g->wrap = 1;
            found = true;
        }
    }
    for (int R = kC / G; R >= 1 && !found; --R) {
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
    int nslots = (int)((kSmemMax - (int)sizeof(Smem)) / g->slot_bytes);   // the ring is a multiple of 128 bytes: Smem follows it directly
    if (nslots > kMaxSlots) nslots = kMaxSlots;
    if (const char* env = getenv("HIPE_CLUSTER_SLOTS")) {
        const int v = atoi(env);
        if (v >= 2 && v < nslots) nslots = v;
    }
    g->nslots = nslots;
    // Cluster size.  What a B200 can hold at once (cudaOccupancyMaxActiveClusters with this kernel's
    // footprint, profiles/r2_cluster_occupancy.md): clusters of 1 / 2 CTAs reach all 148 SMs, 3 -> 135,
    // 4 -> 132, 5 -> 130, 6 -> 132, 7 -> 105, 8 -> 120.  Among the sizes whose largest share leaves
    // kMinSpare slots of look-ahead, take the best score = SMs in use, less 1 % per slot of look-ahead
    // short of 6 (ties: the smaller cluster -- fewer partials to wait for).  Measured on 64^3 (ms, 1344 fp32 /
    // 2688 bf16 volumes): fp32 C=6 0.487 vs C=8 0.518, C=5 0.511, C=7 0.59; bf16 C=4 0.562 vs C=3
    // 0.588, C=5 0.66; fp32 56^3 x 672: C=6 0.177 vs C=8 0.183, C=5 0.185, C=4 0.187.
    static const int kSMs[kMaxCluster + 1] = {0, 148, 148, 135, 132, 130, 132, 105, 120};
    int want = 0;
    if (const char* env = getenv("HIPE_CLUSTER_C")) want = atoi(env);
    g->C = 0;
    double best = -1.0;
    for (int C = 1; C <= kMaxCluster; ++C) {
        if (want && C != want) continue;
        if (g->NSv < C) break;
        const int share = (g->NSv + C - 1) / C;
        if (share + (want ? 1 : kMinSpare) > nslots) continue;
        const int spare = nslots - share;
        const double score = kSMs[C] * (1.0 - 0.01 * (spare < 6 ? 6 - spare : 0));
        if (score > best) {
            best = score;
            g->C = C;
            g->max_share = share;
        }
    }
    (void)n_vol;
    // Streamed shares (two rings, second read through L2).  Any cluster size works, nothing is resident;
    // what matters is what the L2 must hold between the two reads: `lead` stages per CTA = one volume
    // per cluster + a margin per CTA.  Measured with ncu (fp32, DRAM read bytes / algorithmic): 64^3 C=8
    // lead 21 (41 MB) 1.03, C=6 lead 24 (52 MB) 1.21, C=4 (63 MB) 1.42; 96^3 C=8 lead 37 (55 MB) 1.40, the
    // same on 12 / 9 clusters (47 / 35 MB) 1.14 / 1.01, C=16 lead 31 (43 MB) 1.03: a little over 40 MB of
    // the 126 MB L2 serve this pattern.
    // Taken when no cluster can hold the volume, and -- because the two rings decouple the teams -- for
    // 4-byte volumes of the 4096-byte geometry with >= 64 stages for which some cluster size fits that
    // bound (fp32 64^3 x 1344, same box, alternating: 0.473-0.495 ms streamed at C=8 vs 0.500-0.518
    // resident; x 672: 0.257 vs 0.269).  It loses where the share is short or the gradient pass is
    // XU-bound (bf16 64^3 x 2688: 0.60 vs 0.54; fp32 56^3 x 672: 0.190-0.205 vs 0.187-0.192; fp32 32^3 x
    // 5376: 0.264 vs 0.242).  HIPE_CLUSTER_STREAM=1 / 0 forces / forbids it.
    // The cluster size of the streamed form: what has been read once and not yet twice is
    // lead x stage bytes x CTAs, and it must stay inside kStrL2Bytes.  Among the sizes that leave a
    // margin (lead - share) of >= kStrMinMargin stages for the cross-CTA merge and the HBM latency, the
    // one that keeps the most SMs busy (ties: the larger margin); lead = share + kStrMargin, capped by
    // the L2 bound.  fp32 64^3: C=8, lead 21 (120 SMs, 41 MB); fp32 96^3: C=16, lead 30 (112 SMs, 41 MB).
    // The bound is sharp and it pays to stay clear of it: C=6 at lead 20 (132 SMs, 43.3 MB) wins a
    // back-to-back loop of launches on a warm box (scripts/ab_cluster.py lead lead2: 0.468-0.472 ms vs
    // 0.50-0.53) but loses the training step, where other kernels run between the launches (bench.py,
    // alternating processes: 0.507-0.509 vs 0.468-0.473 ms per step), and at lead 22 loses both.
    // No size fits (fp32 128^3): C = 8 with the plain lead; the second read then misses L2 in part, like
    // the streaming launch's.
    const int maxC = g->wrap ? kMaxClusterStr : kMaxCluster;     // (16-CTA exchange buffers: wrapped kernel only)
    static const int kSMsStr[kMaxClusterStr + 1] = {0, 148, 148, 135, 132, 130, 132, 105, 120, 0, 0, 0, 0, 0, 0, 0, 112};
    int bestC = 0, bestLead = 0, bestSMs = 0, bestMargin = 0;
    for (int C = 1; C <= maxC; ++C) {
        if (kSMsStr[C] == 0 || g->NSv < C) continue;
        const int share = (g->NSv + C - 1) / C;
        const int lead_max = (int)(kStrL2Bytes / ((double)kSMsStr[C] * g->stage_bytes));
        const int margin = lead_max - share;
        if (margin < kStrMinMargin) continue;
        if (kSMsStr[C] > bestSMs || (kSMsStr[C] == bestSMs && margin > bestMargin)) {
            bestC = C; bestSMs = kSMsStr[C]; bestMargin = margin;
            bestLead = share + kStrMargin < lead_max ? share + kStrMargin : lead_max;
        }
    }
    {
        int force = -1;
        if (const char* env = getenv("HIPE_CLUSTER_STREAM")) force = atoi(env);
        const bool prefer = esz == 4 && g->stage_bytes == kKP * kC * 16 && g->nact == kC && g->SR % H == 0 &&
                            g->NSv >= 8 * kMaxCluster && bestC > 0 && !want;
        if (g->C > 0 && (force == 0 || (force < 0 && !prefer))) return true;
        if (g->C == 0 && force == 0) return false;
    }
    nslots = (int)((kSmemMax - (int)(g->wrap ? sizeof(SmemStr) : sizeof(Smem))) / g->slot_bytes);
    if (nslots > kMaxSlots) nslots = kMaxSlots;
    if (const char* env = getenv("HIPE_CLUSTER_SLOTS")) {
        const int v = atoi(env);
        if (v >= 6 && v < nslots) nslots = v;
    }
    if (nslots < 6) return false;
    g->nslots = nslots;
    int C = bestC > 0 ? bestC : kMaxCluster;
    if (want >= 1 && want <= maxC) C = want;
    while (C > 1 && g->NSv < C) --C;
    g->C = C;
    g->max_share = (g->NSv + C - 1) / C;
    g->streamed = 1;
    g->nslots_r = (nslots * 5 + 4) / 9;          // HBM latency on the reduce ring, L2 latency on the other
    if (const char* env = getenv("HIPE_CLUSTER_RSLOTS")) {
        const int v = atoi(env);
        if (v >= 2 && v <= nslots - 2) g->nslots_r = v;
    }
    g->lead = (C == bestC) ? bestLead : g->max_share + kStrMargin;
    if (const char* env = getenv("HIPE_CLUSTER_LEAD")) {
        const int v = atoi(env);
        if (v > g->max_share) g->lead = v;
    }
    return true;
}

template <typename T, int kPass, int kMode, int KP = kKP, bool kWrap = false>
static int launch(const Params& p0, const Geom& g, cudaStream_t st) {
    auto kern = k_cluster<T, kPass, kMode, KP, kWrap>;
    constexpr int kBlock = kMode == 2 ? kThreads2 : (kMode == 3 ? kThreads3 : kThreads);
    const size_t smem = (size_t)g.nslots * g.slot_bytes + (kWrap ? sizeof(SmemStr) : sizeof(Smem));
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    if (g.C > kMaxCluster) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return (int)e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kBlock, 1, 1);
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
                g.C, smem, kBlock, cudaGetErrorString(e), ncl);
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

#ifdef HIPE_CLU_PROF
// diagnostic builds only: read (and optionally clear) the wait-time table
extern "C" __attribute__((visibility("default"))) int hipe_cluster_prof(unsigned long long* out, int reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out, clu::g_prof, sizeof(unsigned long long) * clu::kProfN);
    if (e != cudaSuccess) return (int)e;
    if (reset) {
        unsigned long long z[clu::kProfN] = {};
        e = cudaMemcpyToSymbol(clu::g_prof, z, sizeof(z));
    }
    return (int)e;
}
extern "C" __attribute__((visibility("default"))) int hipe_cluster_trace(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, clu::g_trace, sizeof(long long) * clu::kTraceN * clu::kTraceE);
}
#endif

extern "C" long long hipe_cluster_workspace_bytes(int dtype, long long n_vol, int D, int H, int W) {
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return 0;
    if (n_vol <= 0 || D <= 0 || H <= 0 || W <= 0) return 0;
    clu::Geom g;
    if (!clu::pick_geom(dtype, n_vol, D, H, W, &g)) return 0;
    return 256;
}

extern "C" int hipe_cluster_plan(int dtype, long long n_vol, int D, int H, int W, int* cluster, int* nslots,
                                 int* share_stages, int* stage_bytes, int* lead_stages) {
    clu::Geom g;
    if ((dtype != HIPE_F32 && dtype != HIPE_BF16) || n_vol <= 0 || D <= 0 || H <= 0 || W <= 0 ||
        !clu::pick_geom(dtype, n_vol, D, H, W, &g))
        return HIPE_E_UNSUPPORTED;
    if (cluster) *cluster = g.C;
    if (nslots) *nslots = g.nslots;
    if (share_stages) *share_stages = g.max_share;
    if (stage_bytes) *stage_bytes = g.stage_bytes;
    if (lead_stages) *lead_stages = g.streamed ? g.lead : 0;
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
    p.nslots_r = g.nslots_r;
    p.lead = g.lead;
    p.pol1 = 2;
    if (const char* env = getenv("HIPE_CLUSTER_POL1")) p.pol1 = atoi(env);
    // L2 prefetch of the stages the producer cannot load yet (resident form), where it was measured to pay:
    // fp32 56^3 x 672 0.179-0.184 vs 0.184-0.195 ms, bf16 32^3 x 5376 0.172-0.175 vs 0.190-0.198; neutral or
    // slightly negative on bf16 64^3 (0.542-0.561 vs 0.530-0.564 on a drifting box), fp32 32^3 (0.254 vs
    // 0.243) and the resident form of fp32 64^3.  HIPE_CLUSTER_PF overrides (stages; 0 = off).
    p.pf = ((dtype == HIPE_F32 && g.pass_bytes == 3136) || (dtype == HIPE_BF16 && g.C == 1 && !g.streamed)) ? 4 : 0;
    if (const char* env = getenv("HIPE_CLUSTER_PF")) p.pf = atoi(env);
    p.coords = coords; p.expect = expect; p.stats = stats; p.loss_terms = loss_terms;
    p.gt = gt; p.vis = vis; p.tgt_f64 = (tgt_dtype == HIPE_F64); p.norm = norm;
    p.ticket = static_cast<unsigned*>(workspace);
    p.loss_out = loss;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    e = cudaMemsetAsync(workspace, 0, 256, st);
    if (e != cudaSuccess) return (int)e;
    p.hoist = (g.SR % H == 0) ? 1 : 0;
    p.inv_W = 1.0 / (double)W; p.inv_H = 1.0 / (double)H; p.inv_D = 1.0 / (double)D;
    p.wteams2 = 0;
    if (const char* env = getenv("HIPE_CLUSTER_WT2")) p.wteams2 = (env[0] != '0') ? 1 : 0;
    p.sym = 0;     // symmetric teams: an experiment that lost on every shape (profiles/r2_configs.md)
    if (const char* env = getenv("HIPE_CLUSTER_SYM")) p.sym = (env[0] != '0') ? 1 : 0;
    p.stasync = 1;
    if (const char* env = getenv("HIPE_CLUSTER_STASYNC")) p.stasync = (env[0] != '0') ? 1 : 0;
    int pass = 0;      // specialised geometry?
    if (p.hoist && g.stage_bytes == 4 * g.pass_bytes && g.slot_bytes == g.stage_bytes && g.nact * 16 == g.pass_bytes &&
        !getenv("HIPE_CLUSTER_GENERIC"))
        pass = (g.pass_bytes == 4096 || g.pass_bytes == 3136) ? g.pass_bytes : 0;
#define HIPE_CLU_LAUNCH(T, MODE)                                                                   \
    (pass == 4096 ? clu::launch<T, 4096, MODE>(p, g, st)                                          \
                  : (pass == 3136 ? clu::launch<T, 3136, MODE>(p, g, st) : clu::launch<T, 0, MODE>(p, g, st)))
#ifdef HIPE_CLU_EXPERIMENTS
    // the two team layouts that lost (see the kernel): built only on request
    const int mode = p.sym ? 1 : p.wteams2 ? 2 : 0;
    if (mode == 1) return dtype == HIPE_F32 ? HIPE_CLU_LAUNCH(float, 1) : HIPE_CLU_LAUNCH(__nv_bfloat16, 1);
    if (mode == 2) return dtype == HIPE_F32 ? HIPE_CLU_LAUNCH(float, 2) : HIPE_CLU_LAUNCH(__nv_bfloat16, 2);
#endif
    if (g.streamed) {
        // streamed shares: the wrapped fp32 geometry (W = 96), the two specialised whole-row ones, or any
        if (g.wrap) {
            if (dtype != HIPE_F32) return HIPE_E_UNSUPPORTED;
            return clu::launch<float, 16 * clu::kC, 3, clu::kWrapKP, true>(p, g, st);
        }
        return dtype == HIPE_F32 ? HIPE_CLU_LAUNCH(float, 3) : HIPE_CLU_LAUNCH(__nv_bfloat16, 3);
    }
    return dtype == HIPE_F32 ? HIPE_CLU_LAUNCH(float, 0) : HIPE_CLU_LAUNCH(__nv_bfloat16, 0);
#undef HIPE_CLU_LAUNCH
}

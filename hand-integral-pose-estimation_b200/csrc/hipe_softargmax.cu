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
constexpr int kNSums = 5;                         // S0 Sx Sy Sz S0c
constexpr int kStages = 4;                        // TMA ring depth
constexpr int kStageBytes = kPasses * kConsumers * 16;  // 16 KB
constexpr int kMaxCluster = 16;                  // 16 is non-portable (opt-in attribute), used by the resident fused launch

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
    int nv2;      // reduce pass reads two adjacent 16-byte vectors per visit (rows are 32-byte multiples)
    int reverse;
    int st_hint;     // gradient stores carry an L2 evict_first hint
    int resident;    // fused launch: every CTA's share fits the ring, pass 2 re-reads shared memory not L2/HBM
    int rev;         // fused launch: the second pass walks the CTA's share backwards (L2 recency)
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
    double wred[kConsumers / 32][kNSums];
    float wmax[kConsumers / 32];
    double part[kNSums];           // this CTA's partial (S0, Sx, Sy, Sz, S0c), read by peers through DSMEM
    float part_m;                  // ... and the maximum it is relative to
    double all[kMaxCluster][kNSums];
    float all_m[kMaxCluster];
    double finE[3];                // Ex, Ey, Ez (voxel units)
    float fin[5];                  // M, L, gx, gy, gz
    double tgt[6];                 // fused: gt[0..2], vis[0..2] of the volume, fetched before the reduce pass
};

__device__ __forceinline__ void consumer_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
}

}  // namespace hipe

#include "hipe_stream_kernel.cuh"

namespace hipe {

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

static int pick_cluster(int op, long long n_vol, long long vol_bytes, int NR, int esz) {
    // Write-only launches: aim for >= ~8 waves of 3 resident CTAs over 148 SMs while
    // keeping every CTA's share >= 64 KB (finer grain = shorter tail; no reduction to pay for).
    // Reduce-only launches pay a cluster barrier + DSMEM merge per CTA, so they split a volume
    // only when there are fewer volumes than two per SM (measured on B200, fp32 64^3: 1344 volumes
    // 0.236 ms at S=1, 0.291 at S=4; 672 volumes 0.119 at S=1, 0.123 at S=2; 336 volumes 0.062 at
    // S=1, 0.070 at S=2, 0.072 at S=4; bf16 672 volumes 0.084 at S=1, 0.095 at S=2).
    // Fused launches (reduce, meet, write) want a share of 256-450 KB per CTA for 4-byte logits and
    // up to ~900 KB for 2-byte logits -- smaller shares pay the serial section between the passes
    // too often, larger ones lose the L2 re-read -- and at least ~8 CTAs per SM in the grid.
    // Measured (fused_ms at S = 1/2/4/8): fp32 64^3 x 1344: -/0.625/0.570/0.692, x 672:
    // 0.340/0.308/0.291/0.359; 56^3 x 2016: 0.594/0.564/0.614/-, x 672: 0.232/0.205/0.216/0.291,
    // x 336: 0.148/0.123/0.115/0.156; 96^3 x 672: -/-/1.035/0.999; bf16 64^3 x 2688:
    // 0.613/0.660/-/-; bf16 96^3 x 672: 0.603/0.588/0.595/-.
    if (op == OP_FUSED) {
        const long long max_share = (esz == 2) ? (896ll << 10) : (448ll << 10);
        int S = 1;
        while (S < 8 && vol_bytes / S > max_share && NR / (2 * S) >= 1) S *= 2;
        while (S < 8 && n_vol * S < 148ll * 8 && vol_bytes / (2 * S) >= (128ll << 10) && NR / (2 * S) >= 1) S *= 2;
        return S;
    }
    const bool reduce_only = (op == OP_REDUCE_SOFTMAX || op == OP_REDUCE_PLAIN);
    const long long want = reduce_only ? 148ll * 2 : 148ll * 24;
    const long long min_share = reduce_only ? 131072 : 65536;
    int S = 1;
    while (S < kMaxCluster && n_vol * S < want && vol_bytes / (2 * S) >= min_share && NR / (2 * S) >= 1) S *= 2;
    return S;
}

template <typename T, bool kTma, int kOp>
static int launch_stream(const Params& p, int S, cudaStream_t st) {
    auto kern = k_stream<T, kTma, kOp>;
    size_t smem = sizeof(SmemLayout) + (kTma ? (size_t)kStages * kStageBytes : 0);
    if (const char* e = getenv("HIPE_PAD_SMEM")) smem += (size_t)atoi(e);   // tuning: lowers CTAs/SM
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    if (S > 8) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return (int)e;
    }
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
    int S = cluster ? cluster : pick_cluster(kOp, p.n_vol, p.vol_bytes, p.NR, esz);
    const bool tma_ = (variant == HIPE_VARIANT_AUTO) ? true : (variant == HIPE_VARIANT_TMA);
    const long long ring_rows = (long long)kStages * kPasses * p.R;   // rows one CTA's ring holds
    if (kOp == OP_FUSED && tma_ && !cluster && getenv("HIPE_RESIDENT")) {
        // smallest cluster whose CTAs keep their whole share in the shared-memory ring: the
        // second pass then re-reads shared memory, and HBM sees exactly one read + one write.
        // Opt-in only: measured on B200 it is SLOWER than re-reading through L2 (64^3 fp32:
        // 0.81 ms at S=16 vs 0.57 ms at S=4) because 64 KB CTAs are too short-lived to hide the
        // cluster launch + barrier latency; the persistent form is the one that pays.
        for (int c = 1; c <= kMaxCluster; c *= 2)
            if ((p.NR + c - 1) / c <= ring_rows && p.NR >= c) { S = c; break; }
    }
    while (S > 1 && p.NR < S) S /= 2;
    p.resident = (kOp == OP_FUSED && tma_ && (p.NR + S - 1) / S <= ring_rows) ? 1 : 0;
    const bool tma = (variant == HIPE_VARIANT_AUTO) ? true : (variant == HIPE_VARIANT_TMA);
    // two adjacent vectors per visit in the reduce pass when rows are 32-byte multiples (shared
    // memory reads only: from global memory it would fetch every 32-byte sector twice)
    p.nv2 = (tma && p.G % 2 == 0) ? 1 : 0;
    if (const char* e = getenv("HIPE_NV2")) p.nv2 = (p.nv2 && e[0] != '0') ? 1 : 0;
    p.rev = (kOp == OP_FUSED) ? 1 : 0;
    if (const char* e = getenv("HIPE_FUSED_REV")) p.rev = (p.rev && e[0] != '0') ? 1 : 0;
    p.st_hint = 0;
    if (const char* e = getenv("HIPE_ST_HINT")) p.st_hint = (e[0] != '0') ? 1 : 0;
    return tma ? launch_stream<T, true, kOp>(p, S, st) : launch_stream<T, false, kOp>(p, S, st);
}

static int check_geom(const void* in, int dtype, long long n_vol, int D, int H, int W, int cluster, int variant) {
    if (!in && n_vol != 0) return HIPE_E_NULL;
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (n_vol < 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    if ((long long)D * H * W > (1ll << 31)) return HIPE_E_SHAPE;
    if (n_vol * kMaxCluster > 0x7fffffffll) return HIPE_E_SHAPE;
    if (cluster != 0 && cluster != 1 && cluster != 2 && cluster != 4 && cluster != 8 && cluster != 16) return HIPE_E_CLUSTER;
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

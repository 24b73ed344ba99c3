// FreiHAND-style keypoint measures on the device: what EvalUtil.feed +
// EvalUtil.get_measures (main/eval.py:25-106) compute from per-sample per-keypoint distances --
// mean and median end-point error per keypoint, the PCK curve over linspace thresholds, its
// trapezoid AUC, and the means over the keypoints that have at least one visible sample.
// The distances are the pa_dist / raw_dist arrays hipe_procrustes_batched already produces, so
// the whole scoring stays on the GPU.  Latency-bound (N*K doubles, 0.67 MB at 3960 x 21): one
// CTA per keypoint, integer histograms only (bit-deterministic), no roofline claimed.  sm_100a.
#include "hipe_common.cuh"
#include <mutex>

namespace hipe {

constexpr int kMT = 1024;          // threads per keypoint CTA
constexpr int kMaxSteps = 1024;    // thresholds held in shared memory

// np.linspace(val_min, val_max, steps) (main/eval.py:65): start + i*step with two roundings,
// last point pinned to the stop value.
__device__ __forceinline__ double linspace_at(double a, double b, int steps, int i) {
    if (steps == 1) return a;
    if (i == steps - 1) return b;
    const double step = __ddiv_rn(__dsub_rn(b, a), (double)(steps - 1));
    return __dadd_rn(__dmul_rn((double)i, step), a);
}

// k-th smallest (0-based) of the visible distances of one keypoint by an 8 x 8-bit radix select
// on the IEEE bit patterns (non-negative doubles order like their unsigned bit patterns).
__device__ unsigned long long radix_select(const double* __restrict__ dist, const unsigned char* __restrict__ vis,
                                           long long N, int K, int kp, long long rank, unsigned* hist /*[256]*/,
                                           unsigned long long* bcast) {
    unsigned long long prefix = 0ull, mask = 0ull;
    for (int shift = 56; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += kMT) hist[i] = 0u;
        __syncthreads();
        for (long long n = threadIdx.x; n < N; n += kMT) {
            if (vis && !vis[n * K + kp]) continue;
            const unsigned long long key = (unsigned long long)__double_as_longlong(dist[n * K + kp]);
            if ((key & mask) == prefix) atomicAdd(&hist[(unsigned)(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            long long cum = 0;
            int d = 0;
            for (; d < 255; ++d) {
                if (cum + hist[d] > rank) break;
                cum += hist[d];
            }
            rank -= cum;
            bcast[0] = prefix | ((unsigned long long)d << shift);
            bcast[1] = (unsigned long long)rank;
        }
        __syncthreads();
        prefix = bcast[0];
        rank = (long long)bcast[1];
        mask |= 255ull << shift;
        __syncthreads();
    }
    return prefix;
}

// per keypoint: count, mean, median, pck[steps], auc
__global__ void __launch_bounds__(kMT)
    k_keypoint_measures(const double* __restrict__ dist, const unsigned char* __restrict__ vis, long long N, int K,
                        double val_min, double val_max, int steps, double* __restrict__ kp_count,
                        double* __restrict__ kp_mean, double* __restrict__ kp_median, double* __restrict__ kp_auc,
                        double* __restrict__ kp_pck /* K*steps */) {
    __shared__ double thr[kMaxSteps];
    __shared__ unsigned cnt[kMaxSteps + 1];
    __shared__ unsigned hist[256];
    __shared__ unsigned long long bcast[2];
    __shared__ double red[32];
    __shared__ unsigned long long nvis_sh;
    const int kp = blockIdx.x;
    for (int i = threadIdx.x; i < steps; i += kMT) thr[i] = linspace_at(val_min, val_max, steps, i);
    for (int i = threadIdx.x; i <= steps; i += kMT) cnt[i] = 0u;
    if (threadIdx.x == 0) nvis_sh = 0ull;
    __syncthreads();
    // one sweep: sum, count, and for every distance the first threshold it passes (d <= thr[t])
    double acc = 0.0;
    unsigned long long mine = 0;
    for (long long n = threadIdx.x; n < N; n += kMT) {
        if (vis && !vis[n * K + kp]) continue;
        const double d = dist[n * K + kp];
        acc += d;
        ++mine;
        int lo = 0, hi = steps;            // first t with d <= thr[t]; steps = none (NaN lands here too)
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (d <= thr[mid]) hi = mid; else lo = mid + 1;
        }
        if (!(d <= thr[steps - 1])) lo = steps;
        atomicAdd(&cnt[lo], 1u);
    }
    atomicAdd(&nvis_sh, mine);
    // fixed-order fp64 sum: thread-strided partials -> warp shuffles -> 32 warp partials
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    const long long nvis = (long long)nvis_sh;
    if (threadIdx.x < 32) {
        double r = red[threadIdx.x];
        r = warp_sum(r);
        if (threadIdx.x == 0) {
            kp_count[kp] = (double)nvis;
            kp_mean[kp] = nvis ? r / (double)nvis : 0.0;
        }
    }
    if (nvis == 0) {       // main/eval.py:52-53,61-62: the keypoint is skipped
        if (threadIdx.x == 0) { kp_median[kp] = 0.0; kp_auc[kp] = 0.0; }
        for (int i = threadIdx.x; i < steps; i += kMT) kp_pck[(long long)kp * steps + i] = 0.0;
        return;
    }
    // pck[t] = #(d <= thr[t]) / n (main/eval.py:50-57), then np.trapz (main/eval.py:66,95-97)
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        double area = 0.0, norm = 0.0, prev = 0.0;
        for (int t = 0; t < steps; ++t) {
            run += cnt[t];
            const double pck = (double)run / (double)nvis;
            kp_pck[(long long)kp * steps + t] = pck;
            if (t > 0) {
                const double dx = thr[t] - thr[t - 1];
                area += dx * (pck + prev) / 2.0;
                norm += dx * (1.0 + 1.0) / 2.0;
            }
            prev = pck;
        }
        kp_auc[kp] = area / norm;
    }
    // np.median: middle element, or the mean of the two middle ones
    const unsigned long long lo_key = radix_select(dist, vis, N, K, kp, (nvis - 1) / 2, hist, bcast);
    unsigned long long hi_key = lo_key;
    if ((nvis & 1) == 0) hi_key = radix_select(dist, vis, N, K, kp, nvis / 2, hist, bcast);
    if (threadIdx.x == 0)
        kp_median[kp] = (nvis & 1) ? __longlong_as_double((long long)lo_key)
                                   : (__longlong_as_double((long long)lo_key) + __longlong_as_double((long long)hi_key)) / 2.0;
}

// means over the keypoints that were fed at least once (main/eval.py:100-104)
__global__ void __launch_bounds__(256)
    k_measures_final(const double* __restrict__ kp_count, const double* __restrict__ kp_mean,
                     const double* __restrict__ kp_median, const double* __restrict__ kp_auc,
                     const double* __restrict__ kp_pck, int K, int steps, double* __restrict__ summary /*[3]*/,
                     double* __restrict__ pck_curve /*[steps]*/) {
    int valid = 0;
    for (int k = 0; k < K; ++k) valid += kp_count[k] > 0.0;
    if (threadIdx.x == 0) {
        double m = 0.0, md = 0.0, a = 0.0;
        for (int k = 0; k < K; ++k)
            if (kp_count[k] > 0.0) { m += kp_mean[k]; md += kp_median[k]; a += kp_auc[k]; }
        summary[0] = valid ? m / valid : 0.0;
        summary[1] = valid ? md / valid : 0.0;
        summary[2] = valid ? a / valid : 0.0;
    }
    for (int t = threadIdx.x; t < steps; t += blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < K; ++k)
            if (kp_count[k] > 0.0) s += kp_pck[(long long)k * steps + t];
        pck_curve[t] = valid ? s / valid : 0.0;
    }
}


// ---------------------------------------------------------------------------------------
// Per-volume softmax entropy: what main/generate_correlation_graph.py:116-141
// gets from softmax_integral_tensor2 (common/nets/loss.py:61-67) -> .cpu() -> scipy.stats.entropy per
// joint, without ever materialising the (B, J, V) probabilities or copying them to the host:
//   H = -sum p ln p = ln L - ln2 * (sum e_i a_i) / L,   a_i = s_i log2e - fl(M log2e),  e_i = 2^a_i
// with (M, L) the statistics the forward launch already produced.  One streaming read of the
// volume (128-bit ld.global.nc, 4 loads in flight per thread), one CTA per volume, fp32 products
// summed per thread in fp64, fixed-order block reduction.  HBM-bound: V*s bytes per volume.
// ---------------------------------------------------------------------------------------
constexpr int kET = 512;

template <typename T>
__global__ void __launch_bounds__(kET)
    k_entropy(const unsigned char* __restrict__ in, long long n_vol, long long V, const float* __restrict__ stats,
              double* __restrict__ entropy, int vec_ok) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    __shared__ double red[kET / 32];
    const long long vol = blockIdx.x;
    const float M = stats[vol * 2 + 0];
    const double Lsum = (double)stats[vol * 2 + 1];
    const float ML = __fmul_rn(M, kLog2e);
    const T* src = reinterpret_cast<const T*>(in) + vol * V;
    double acc = 0.0;
    long long done = 0;
    if (vec_ok) {
        const long long nvec = V / VEC;
        const unsigned char* base = reinterpret_cast<const unsigned char*>(src);
        for (long long i0 = 0; i0 < nvec; i0 += 4ll * kET) {
            uint4 r[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long i = i0 + (long long)u * kET + threadIdx.x;
                ok[u] = i < nvec;
                if (ok[u]) r[u] = ldg_stream(base + i * 16);
            }
            float part = 0.f;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (!ok[u]) continue;
                float v[VEC];
                V16::unpack(r[u], v);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    // clamp: a = -inf (a -inf logit) would give 0 * inf; 2^-200 flushes to 0 all the same
                    const float a = fmaxf(fmaf(v[j], kLog2e, -ML), -200.f);
                    part = fmaf(ex2(a), a, part);     // e * a in (-0.54, 0]
                }
            }
            acc += (double)part;
        }
        done = nvec * VEC;
    }
    for (long long i = done + threadIdx.x; i < V; i += kET) {
        const float a = fmaxf(fmaf(V16::to_f32(src[i]), kLog2e, -ML), -200.f);
        acc += (double)(ex2(a) * a);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double r = (threadIdx.x < kET / 32) ? red[threadIdx.x] : 0.0;
        r = warp_sum(r);
        if (threadIdx.x == 0) entropy[vol] = log(Lsum) - 0.6931471805599453 * r / Lsum;
    }
}

}  // namespace hipe

using namespace hipe;

extern "C" long long hipe_keypoint_measures_workspace_bytes(int K, int steps) {
    if (K <= 0 || steps <= 0) return 0;
    return (long long)sizeof(double) * ((long long)K * 4 + (long long)K * steps);
}

extern "C" int hipe_keypoint_measures(const double* dist, const unsigned char* vis, long long N, int K,
                                      double val_min, double val_max, int steps, double* summary,
                                      double* pck_curve, double* workspace, long long workspace_bytes,
                                      void* stream) {
    if (!dist || !summary || !pck_curve || !workspace) return HIPE_E_NULL;
    if (N < 0 || K <= 0 || steps <= 0 || steps > kMaxSteps) return HIPE_E_SHAPE;
    if (workspace_bytes < hipe_keypoint_measures_workspace_bytes(K, steps)) return HIPE_E_SHAPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    double* cnt = workspace;
    double* mean = cnt + K;
    double* med = mean + K;
    double* auc = med + K;
    double* pck = auc + K;
    k_keypoint_measures<<<K, kMT, 0, st>>>(dist, vis, N, K, val_min, val_max, steps, cnt, mean, med, auc, pck);
    k_measures_final<<<1, 256, 0, st>>>(cnt, mean, med, auc, pck, K, steps, summary, pck_curve);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// hipe_keypoint_measures_noalloc: the same two launches with the per-keypoint scratch in a
// file-scope device array instead of caller-owned memory, for callers that must not allocate
// (CUDA graph capture, serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the array until exit.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) double g_measures_scratch[HIPE_NOALLOC_MEASURES_MAX_DOUBLES];

// The array's name is a host-side stand-in in host code, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
static int measures_scratch(double** out) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static double* ptr[kMaxDev] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDev) return HIPE_E_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_measures_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!p) return HIPE_E_NULL;
        ptr[dev] = static_cast<double*>(p);
    }
    *out = ptr[dev];
    return 0;
}

extern "C" int hipe_keypoint_measures_noalloc(const double* dist, const unsigned char* vis, long long N, int K,
                                              double val_min, double val_max, int steps, double* summary,
                                              double* pck_curve, void* stream) {
    if (!dist || !summary || !pck_curve) return HIPE_E_NULL;
    if (N < 0 || K <= 0 || steps <= 0 || steps > kMaxSteps) return HIPE_E_SHAPE;
    if (K > HIPE_NOALLOC_MEASURES_MAX_K ||
        (long long)K * (4 + (long long)steps) > (long long)HIPE_NOALLOC_MEASURES_MAX_DOUBLES)
        return HIPE_E_LIMIT;
    double* ws = nullptr;
    const int rc = measures_scratch(&ws);
    if (rc != 0) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array keeps the previous call's contents: the part this call uses starts at zero
    const size_t used = sizeof(double) * ((size_t)K * 4 + (size_t)K * steps);
    cudaError_t e = cudaMemsetAsync(ws, 0, used, st);
    if (e != cudaSuccess) return (int)e;
    double* cnt = ws;
    double* mean = cnt + K;
    double* med = mean + K;
    double* auc = med + K;
    double* pck = auc + K;
    k_keypoint_measures<<<K, kMT, 0, st>>>(dist, vis, N, K, val_min, val_max, steps, cnt, mean, med, auc, pck);
    k_measures_final<<<1, 256, 0, st>>>(cnt, mean, med, auc, pck, K, steps, summary, pck_curve);
    return (int)cudaGetLastError();
}

extern "C" int hipe_volume_entropy(const void* logits, int dtype, long long n_vol, int D, int H, int W,
                                   const float* stats, double* entropy, void* stream) {
    if (n_vol < 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    if (n_vol == 0) return 0;
    if (!logits || !stats || !entropy) return HIPE_E_NULL;
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (n_vol > 0x7fffffffll) return HIPE_E_SHAPE;
    const long long V = (long long)D * H * W;
    const int esz = dtype == HIPE_F32 ? 4 : 2;
    if (reinterpret_cast<uintptr_t>(logits) % esz) return HIPE_E_ALIGN;
    const int vec_ok = (reinterpret_cast<uintptr_t>(logits) % 16 == 0) && ((V * esz) % 16 == 0);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == HIPE_F32)
        k_entropy<float><<<(unsigned)n_vol, kET, 0, st>>>(static_cast<const unsigned char*>(logits), n_vol, V, stats, entropy, vec_ok);
    else
        k_entropy<__nv_bfloat16><<<(unsigned)n_vol, kET, 0, st>>>(static_cast<const unsigned char*>(logits), n_vol, V, stats, entropy, vec_ok);
    return (int)cudaGetLastError();
}

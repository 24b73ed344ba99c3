// Small fp64 kernels of the path: the L1 integral loss on coordinates
// (common/nets/loss.py:86-89), a deterministic fp64 sum, and the batched Procrustes
// alignment (common/augment.py:18-77, main/eval.py:142-165) with the per-joint error
// distances of data/FreiHand/FreiHand.py:766-767,804-807.  All latency-bound; sm_100a.
#include "hipe_common.cuh"

namespace hipe {

constexpr int kRedThreads = 1024;

// Fixed-order block reduction: thread-strided partials -> warp shuffles -> 32 warp partials.
__device__ __forceinline__ double block_sum_f64(double v, double* sh /*[32]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double r = 0.0;
    if (warp == 0) {
        r = (lane < (int)(blockDim.x >> 5)) ? sh[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;  // valid in warp 0
}

template <typename TT>
__global__ void __launch_bounds__(kRedThreads)
    k_l1_loss(const float* __restrict__ coords, const TT* __restrict__ gt, const TT* __restrict__ vis,
              long long n, double norm, double* __restrict__ loss, float* __restrict__ gcoords) {
    __shared__ double sh[32];
    double acc = 0.0;
    for (long long i = threadIdx.x; i < n; i += kRedThreads) {
        // type promotion as torch does it: fp32 coords against fp64 targets -> fp64
        const double d = (double)coords[i] - (double)gt[i];
        const double w = (double)vis[i];
        acc += fabs(d) * w;
        if (gcoords) {
            const double sg = (d > 0.0) ? 1.0 : ((d < 0.0) ? -1.0 : 0.0);
            gcoords[i] = (float)(sg * w * norm);
        }
    }
    const double tot = block_sum_f64(acc, sh);
    if (threadIdx.x == 0) loss[0] = tot * norm;
}

__global__ void __launch_bounds__(kRedThreads)
    k_sum_f64(const double* __restrict__ in, long long n, double scale, double* __restrict__ out) {
    __shared__ double sh[32];
    double acc = 0.0;
    for (long long i = threadIdx.x; i < n; i += kRedThreads) acc += in[i];
    const double tot = block_sum_f64(acc, sh);
    if (threadIdx.x == 0) out[0] = tot * scale;
}

// x *= *gscale, skipped entirely (no memory traffic) when the scalar is exactly 1: lets the
// fused loss hand back the gradient it computed in the forward launch without a host sync.
template <typename T>
__global__ void __launch_bounds__(256) k_scale_inplace(unsigned char* __restrict__ x, long long n_vec,
                                                         long long n_tail_start, long long n,
                                                         const float* __restrict__ gscale) {
    const float g = *gscale;
    if (g == 1.0f) return;
    using V16 = Vec16<T>;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec;
         i += (long long)gridDim.x * blockDim.x) {
        float v[V16::N];
        uint4 r = *reinterpret_cast<const uint4*>(x + i * 16);
        V16::unpack(r, v);
#pragma unroll
        for (int j = 0; j < V16::N; ++j) v[j] *= g;
        *reinterpret_cast<uint4*>(x + i * 16) = V16::pack(v);
    }
    if (blockIdx.x == 0) {
        T* xe = reinterpret_cast<T*>(x);
        for (long long i = n_tail_start + threadIdx.x; i < n; i += blockDim.x)
            xe[i] = V16::from_f32(V16::to_f32(xe[i]) * g);
    }
}

// ---------------------------------------------------------------------------------------
// 3x3 SVD in registers by one-sided (Hestenes) Jacobi: rotate column pairs of A until they
// are mutually orthogonal; then A V = U diag(s).  High relative accuracy, no squaring of
// the condition number.  Returns singular values sorted descending, as LAPACK does
// (the reflection repair acts on the LAST = smallest one, augment.py:60-61).
// ---------------------------------------------------------------------------------------
struct Mat3 {
    double a[3][3];
};

__device__ __forceinline__ void rot_cols(Mat3& M, int p, int q, double c, double s) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double mp = M.a[i][p], mq = M.a[i][q];
        M.a[i][p] = c * mp - s * mq;
        M.a[i][q] = s * mp + c * mq;
    }
}

__device__ __forceinline__ void swap_cols(Mat3& M, int p, int q) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double t = M.a[i][p];
        M.a[i][p] = M.a[i][q];
        M.a[i][q] = t;
    }
}

__device__ void svd3(const Mat3& A, Mat3& U, double (&sv)[3], Mat3& V) {
    Mat3 B = A;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) V.a[i][j] = (i == j) ? 1.0 : 0.0;

    for (int sweep = 0; sweep < 40; ++sweep) {
        bool rotated = false;
#pragma unroll
        for (int pair = 0; pair < 3; ++pair) {
            const int p = (pair == 2) ? 1 : 0;
            const int q = (pair == 0) ? 1 : 2;
            double alpha = 0.0, beta = 0.0, gamma = 0.0;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                alpha += B.a[i][p] * B.a[i][p];
                beta += B.a[i][q] * B.a[i][q];
                gamma += B.a[i][p] * B.a[i][q];
            }
            if (fabs(gamma) > 1e-17 * sqrt(alpha * beta) && gamma != 0.0) {
                rotated = true;
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = ((zeta >= 0.0) ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t);
                const double s = c * t;
                rot_cols(B, p, q, c, s);
                rot_cols(V, p, q, c, s);
            }
        }
        if (!rotated) break;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
        sv[k] = sqrt(B.a[0][k] * B.a[0][k] + B.a[1][k] * B.a[1][k] + B.a[2][k] * B.a[2][k]);
    // sort descending (3-element network), permuting the columns of B and V alike
#define HIPE_CSWAP(P, Q)                 \
    if (sv[P] < sv[Q]) {                 \
        const double t_ = sv[P];         \
        sv[P] = sv[Q];                   \
        sv[Q] = t_;                      \
        swap_cols(B, P, Q);              \
        swap_cols(V, P, Q);              \
    }
    HIPE_CSWAP(0, 1)
    HIPE_CSWAP(1, 2)
    HIPE_CSWAP(0, 1)
#undef HIPE_CSWAP
    // left vectors: u0 = b0/s0; u1 = b1/s1 re-orthogonalised against u0; u2 = +-(u0 x u1)
    // with the sign of b2 (an exactly orthonormal U, like LAPACK's, also when s2 ~ 0).
    double u0[3], u1[3], u2[3];
    if (sv[0] > 0.0) {
#pragma unroll
        for (int i = 0; i < 3; ++i) u0[i] = B.a[i][0] / sv[0];
    } else {
        u0[0] = 1.0; u0[1] = 0.0; u0[2] = 0.0;
    }
    if (sv[1] > 1e-300 && sv[1] > 1e-15 * sv[0]) {
        double dp = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i) u1[i] = B.a[i][1] / sv[1];
#pragma unroll
        for (int i = 0; i < 3; ++i) dp += u1[i] * u0[i];
        double nn = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            u1[i] -= dp * u0[i];
            nn += u1[i] * u1[i];
        }
        nn = sqrt(nn);
#pragma unroll
        for (int i = 0; i < 3; ++i) u1[i] /= nn;
    } else {
        // rank <= 1: any unit vector orthogonal to u0
        const int k = (fabs(u0[0]) <= fabs(u0[1]) && fabs(u0[0]) <= fabs(u0[2])) ? 0
                      : (fabs(u0[1]) <= fabs(u0[2]) ? 1 : 2);
        double e[3] = {0.0, 0.0, 0.0};
        e[k] = 1.0;
        const double dp = u0[k];
        double nn = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            u1[i] = e[i] - dp * u0[i];
            nn += u1[i] * u1[i];
        }
        nn = sqrt(nn);
#pragma unroll
        for (int i = 0; i < 3; ++i) u1[i] /= nn;
    }
    u2[0] = u0[1] * u1[2] - u0[2] * u1[1];
    u2[1] = u0[2] * u1[0] - u0[0] * u1[2];
    u2[2] = u0[0] * u1[1] - u0[1] * u1[0];
    const double dir = u2[0] * B.a[0][2] + u2[1] * B.a[1][2] + u2[2] * B.a[2][2];
    const double sg = (dir < 0.0) ? -1.0 : 1.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        U.a[i][0] = u0[i];
        U.a[i][1] = u1[i];
        U.a[i][2] = sg * u2[i];
    }
}

__device__ __forceinline__ double det3(const Mat3& M) {
    return M.a[0][0] * (M.a[1][1] * M.a[2][2] - M.a[1][2] * M.a[2][1]) -
           M.a[0][1] * (M.a[1][0] * M.a[2][2] - M.a[1][2] * M.a[2][0]) +
           M.a[0][2] * (M.a[1][0] * M.a[2][1] - M.a[1][1] * M.a[2][0]);
}

constexpr int kProcWarps = 4;

// One warp per sample; lanes stride over joints.
__global__ void __launch_bounds__(kProcWarps * 32)
    k_procrustes(const double* __restrict__ X, const double* __restrict__ Y, long long N, int J,
                 int mode, int optimal_scale, double* __restrict__ d_out, double* __restrict__ Z_out,
                 double* __restrict__ T_out, double* __restrict__ b_out, double* __restrict__ c_out,
                 double* __restrict__ pa_dist, double* __restrict__ raw_dist) {
    const int lane = threadIdx.x & 31;
    const long long n = (long long)blockIdx.x * kProcWarps + (threadIdx.x >> 5);
    if (n >= N) return;
    const double* x = X + n * J * 3;
    const double* y = Y + n * J * 3;

    // centroids (augment.py:35-36)
    double mx[3] = {0, 0, 0}, my[3] = {0, 0, 0};
    for (int j = lane; j < J; j += 32)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mx[k] += x[j * 3 + k];
            my[k] += y[j * 3 + k];
        }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        mx[k] = warp_sum(mx[k]) / (double)J;
        my[k] = warp_sum(my[k]) / (double)J;
    }
    // squared Frobenius norms of the centred sets (augment.py:38-46)
    double ssx = 0.0, ssy = 0.0;
    for (int j = lane; j < J; j += 32)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double a = x[j * 3 + k] - mx[k], b = y[j * 3 + k] - my[k];
            ssx += a * a;
            ssy += b * b;
        }
    ssx = warp_sum(ssx);
    ssy = warp_sum(ssy);
    double nx = sqrt(ssx), ny = sqrt(ssy);
    if (mode == 1) {  // eval.py:151-154
        nx += 1e-8;
        ny += 1e-8;
    }
    // A = X0^T Y0 on the unit-scaled sets (augment.py:49-53)
    Mat3 A;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) A.a[i][k] = 0.0;
    for (int j = lane; j < J; j += 32) {
        double a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            a[k] = (x[j * 3 + k] - mx[k]) / nx;
            b[k] = (y[j * 3 + k] - my[k]) / ny;
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int k = 0; k < 3; ++k) A.a[i][k] += a[i] * b[k];
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) A.a[i][k] = warp_sum(A.a[i][k]);

    // every lane runs the same 3x3 SVD (augment.py:54-56): T = V U^T
    Mat3 U, V;
    double s[3];
    svd3(A, U, s, V);
    Mat3 T;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            T.a[i][k] = V.a[i][0] * U.a[k][0] + V.a[i][1] * U.a[k][1] + V.a[i][2] * U.a[k][2];
    if (mode == 0) {
        // reflection repair (augment.py:59-62); np.sign(0) = 0 is mirrored
        const double dt = det3(T);
        const double sg = (dt > 0.0) ? 1.0 : ((dt < 0.0) ? -1.0 : 0.0);
        s[2] *= sg;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int k = 0; k < 3; ++k)
                T.a[i][k] = V.a[i][0] * U.a[k][0] + V.a[i][1] * U.a[k][1] + sg * V.a[i][2] * U.a[k][2];
    }
    const double tr = s[0] + s[1] + s[2];  // augment.py:64
    double b, d, zscale;
    if (mode == 1) {  // eval.py:157-161
        b = tr * nx / ny;
        d = 1.0 - tr * tr;
        zscale = tr * nx;
    } else if (optimal_scale) {  // augment.py:66-69
        b = tr * nx / ny;
        d = 1.0 - tr * tr;
        zscale = nx * tr;
    } else {  // augment.py:70-73
        b = 1.0;
        d = 1.0 + ssy / ssx - 2.0 * tr * ny / nx;
        zscale = ny;
    }
    if (lane == 0) {
        if (d_out) d_out[n] = d;
        if (b_out) b_out[n] = b;
        if (T_out) {
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int k = 0; k < 3; ++k) T_out[n * 9 + i * 3 + k] = T.a[i][k];
        }
        if (c_out) {  // augment.py:75
#pragma unroll
            for (int k = 0; k < 3; ++k)
                c_out[n * 3 + k] = mx[k] - b * (my[0] * T.a[0][k] + my[1] * T.a[1][k] + my[2] * T.a[2][k]);
        }
    }
    // Z = zscale * Y0 T + muX, and the per-joint distances
    for (int j = lane; j < J; j += 32) {
        double y0[3], z[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) y0[k] = (y[j * 3 + k] - my[k]) / ny;
        double e1 = 0.0, e2 = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            z[k] = zscale * (y0[0] * T.a[0][k] + y0[1] * T.a[1][k] + y0[2] * T.a[2][k]) + mx[k];
            const double da = z[k] - x[j * 3 + k], dr = y[j * 3 + k] - x[j * 3 + k];
            e1 += da * da;
            e2 += dr * dr;
        }
        if (Z_out) {
#pragma unroll
            for (int k = 0; k < 3; ++k) Z_out[(n * J + j) * 3 + k] = z[k];
        }
        if (pa_dist) pa_dist[n * J + j] = sqrt(e1);
        if (raw_dist) raw_dist[n * J + j] = sqrt(e2);
    }
}

}  // namespace hipe

using namespace hipe;

extern "C" int hipe_l1_loss_fwd(const float* coords, const void* gt, const void* vis, int tgt_dtype,
                                long long B, int K, double norm, double* loss, float* gcoords,
                                void* stream) {
    if (!coords || !gt || !vis || !loss) return HIPE_E_NULL;
    if (B < 0 || K <= 0) return HIPE_E_SHAPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long n = B * K;
    if (tgt_dtype == HIPE_F64)
        k_l1_loss<double><<<1, kRedThreads, 0, st>>>(coords, static_cast<const double*>(gt),
                                                     static_cast<const double*>(vis), n, norm, loss, gcoords);
    else if (tgt_dtype == HIPE_F32)
        k_l1_loss<float><<<1, kRedThreads, 0, st>>>(coords, static_cast<const float*>(gt),
                                                    static_cast<const float*>(vis), n, norm, loss, gcoords);
    else
        return HIPE_E_DTYPE;
    return (int)cudaGetLastError();
}

extern "C" int hipe_sum_f64(const double* in, long long n, double scale, double* out, void* stream) {
    if (!out || (!in && n > 0)) return HIPE_E_NULL;
    if (n < 0) return HIPE_E_SHAPE;
    k_sum_f64<<<1, kRedThreads, 0, static_cast<cudaStream_t>(stream)>>>(in, n, scale, out);
    return (int)cudaGetLastError();
}

extern "C" int hipe_scale_inplace(void* x, int dtype, long long n, const float* gscale, void* stream) {
    if (!x || !gscale) return HIPE_E_NULL;
    if (n < 0) return HIPE_E_SHAPE;
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (n == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int per = dtype == HIPE_F32 ? 4 : 8;
    const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    const long long n_vec = aligned ? n / per : 0;
    long long grid = (n_vec + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    if (grid < 1) grid = 1;
    if (dtype == HIPE_F32)
        k_scale_inplace<float><<<(unsigned)grid, 256, 0, st>>>(static_cast<unsigned char*>(x), n_vec, n_vec * per, n, gscale);
    else
        k_scale_inplace<__nv_bfloat16><<<(unsigned)grid, 256, 0, st>>>(static_cast<unsigned char*>(x), n_vec, n_vec * per, n, gscale);
    return (int)cudaGetLastError();
}

extern "C" int hipe_procrustes_batched(const double* X, const double* Y, long long N, int J, int mode,
                                       int optimal_scale, double* d, double* Z, double* T, double* b,
                                       double* c, double* pa_dist, double* raw_dist, void* stream) {
    if (!X || !Y) return HIPE_E_NULL;
    if (N < 0 || J <= 0) return HIPE_E_SHAPE;
    if (mode != 0 && mode != 1) return HIPE_E_SHAPE;
    if (N == 0) return 0;
    const unsigned grid = (unsigned)((N + kProcWarps - 1) / kProcWarps);
    k_procrustes<<<grid, kProcWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        X, Y, N, J, mode, optimal_scale, d, Z, T, b, c, pa_dist, raw_dist);
    return (int)cudaGetLastError();
}

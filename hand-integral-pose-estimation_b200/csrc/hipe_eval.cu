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
template <typename T, typename S>
__global__ void __launch_bounds__(256) k_scale_inplace(unsigned char* __restrict__ x, long long n_vec,
                                                         long long n_tail_start, long long n,
                                                         const S* __restrict__ gscale) {
    const float g = (float)*gscale;
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

// ---------------------------------------------------------------------------------------
// Patch -> camera coordinate chain between the soft-argmax and the Procrustes alignment
// (SURVEY.md 8f row 1).  Per joint, all fp64 like the reference's numpy:
//   get_joint_location_result rescale         common/augment.py:176-178
//   trans_coords_from_patch_to_org_3d         common/augment.py:186-217, trans_point2d :508-511
//       (u, v) = first two rows of trans (x, y, 1) -- no perspective divide, as there --, z + tprime
//   pixel2cam   K^-1 (u, v, 1) z / 1000       common/augment.py:89-103
//   R^T                                       data/FreiHand/FreiHand.py:722
//   scale_result(method='scale')              data/FreiHand/FreiHand.py:632-653
// `trans` either comes from the caller (gts[n]['trans']) or is formed from the crop box in closed
// form: gen_trans_from_patch_cv(..., inv=True) (common/augment.py:457-506) of two axis-aligned
// rectangles is a per-axis scale and offset; the reference's float32 rounding of the corners
// (:473-476) is reproduced.
// ---------------------------------------------------------------------------------------
struct ChainArgs {
    const void* coords;      // kind 0: float (N, 3J) soft-argmax output; kind 1: double (N, J, stride) patch pixels
    int kind, stride;
    double patch_w, patch_h;
    const double* trans;     // N*9 row-major, or NULL -> bbox
    const double* bbox;      // N*4 (centre x, centre y, width, height)
    double box_scale, dst_w, dst_h;
    const double* tprime;    // N
    const double* K;         // N*9
    const double* R;         // N*9 or NULL
    const double* bone_len;  // N or NULL
    int ja, jb;
    double* img_out;         // N*J*3 or NULL
    double* cam_out;         // N*J*3 or NULL
};

// One warp per sample; lanes stride over the joints and leave the camera-space joints in `ysm`
// (J*3 doubles of this warp's shared memory).
__device__ __forceinline__ void chain_warp(const ChainArgs& c, long long n, int J, int lane, double* ysm) {
    double t[2][3];
    if (c.trans) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int k = 0; k < 3; ++k) t[i][k] = c.trans[n * 9 + i * 3 + k];
    } else {
        const double cx = c.bbox[n * 4 + 0], cy = c.bbox[n * 4 + 1];
        const double sw = c.bbox[n * 4 + 2] * c.box_scale, sh = c.bbox[n * 4 + 3] * c.box_scale;
        const double xl = (double)(float)(cx - sw * 0.5), xr = (double)(float)(cx + sw * 0.5);
        const double yt = (double)(float)(cy - sh * 0.5), yb = (double)(float)(cy + sh * 0.5);
        t[0][0] = (xr - xl) / (double)(float)c.dst_w; t[0][1] = 0.0; t[0][2] = xl;
        t[1][0] = 0.0; t[1][1] = (yb - yt) / (double)(float)c.dst_h; t[1][2] = yt;
    }
    // K^-1 by the adjugate
    Mat3 Km, Ki;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) Km.a[i][k] = c.K[n * 9 + i * 3 + k];
    const double idet = 1.0 / det3(Km);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int r0 = (k + 1) % 3, r1 = (k + 2) % 3, c0 = (i + 1) % 3, c1 = (i + 2) % 3;
            Ki.a[i][k] = (Km.a[r0][c0] * Km.a[r1][c1] - Km.a[r0][c1] * Km.a[r1][c0]) * idet;
        }
    const double tp = c.tprime[n];
    for (int j = lane; j < J; j += 32) {
        double x, y, z;
        if (c.kind == 0) {
            const float* f = static_cast<const float*>(c.coords) + (n * J + j) * 3;
            x = ((double)f[0] + 0.5) * c.patch_w;
            y = ((double)f[1] + 0.5) * c.patch_h;
            z = (double)f[2] * c.patch_w;
        } else {
            const double* d = static_cast<const double*>(c.coords) + (n * J + j) * c.stride;
            x = d[0]; y = d[1]; z = d[2];
        }
        const double u = t[0][0] * x + t[0][1] * y + t[0][2];
        const double v = t[1][0] * x + t[1][1] * y + t[1][2];
        const double zi = z + tp;
        if (c.img_out) {
            c.img_out[(n * J + j) * 3 + 0] = u;
            c.img_out[(n * J + j) * 3 + 1] = v;
            c.img_out[(n * J + j) * 3 + 2] = zi;
        }
        const double zs = zi / 1000.0;
        double q[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) q[i] = (Ki.a[i][0] * u + Ki.a[i][1] * v + Ki.a[i][2]) * zs;
        if (c.R) {
            const double* Rn = c.R + n * 9;
#pragma unroll
            for (int k = 0; k < 3; ++k) ysm[j * 3 + k] = q[0] * Rn[0 * 3 + k] + q[1] * Rn[1 * 3 + k] + q[2] * Rn[2 * 3 + k];
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) ysm[j * 3 + k] = q[k];
        }
    }
    __syncwarp();
    if (c.bone_len) {
        double dd = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double df = ysm[c.ja * 3 + k] - ysm[c.jb * 3 + k];
            dd += df * df;
        }
        const double alpha = c.bone_len[n] / sqrt(dd);
        __syncwarp();
        for (int j = lane; j < J; j += 32)
#pragma unroll
            for (int k = 0; k < 3; ++k) ysm[j * 3 + k] *= alpha;
        __syncwarp();
    }
    if (c.cam_out)
        for (int j = lane; j < J; j += 32)
#pragma unroll
            for (int k = 0; k < 3; ++k) c.cam_out[(n * J + j) * 3 + k] = ysm[j * 3 + k];
}

constexpr int kProcWarps = 4;

// One warp per sample; lanes stride over joints.  kChain: the inputs Y are first produced in
// shared memory by the patch -> camera chain (the eval path then never leaves the device between
// the soft-argmax and the error distances); kAlign = false stops after the chain.
template <bool kChain, bool kAlign>
__global__ void __launch_bounds__(kProcWarps * 32)
    k_procrustes(const ChainArgs ca, const double* __restrict__ X, const double* __restrict__ Y, long long N, int J,
                 int mode, int optimal_scale, double* __restrict__ d_out, double* __restrict__ Z_out,
                 double* __restrict__ T_out, double* __restrict__ b_out, double* __restrict__ c_out,
                 double* __restrict__ pa_dist, double* __restrict__ raw_dist) {
    extern __shared__ double ysm_all[];
    const int lane = threadIdx.x & 31;
    const long long n = (long long)blockIdx.x * kProcWarps + (threadIdx.x >> 5);
    if (n >= N) return;
    const double* y = Y + n * J * 3;
    if constexpr (kChain) {
        double* ysm = ysm_all + (size_t)(threadIdx.x >> 5) * J * 3;
        chain_warp(ca, n, J, lane, ysm);
        y = ysm;
    }
    if constexpr (!kAlign) return;
    const double* x = X + n * J * 3;

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


// ---------------------------------------------------------------------------------------
// Batched 3x3 orthonormalisation with backward: make_orthonormal
// (procrustes_encoding/nrsfm/nrsfm_modules.py:46-67) = U diag(1, 1, det(U V^T)) V^T from the SVD
// of each matrix, which the reference gets from the un-vendored torch_batch_svd extension
// (nrsfm/batch_svd.py:5-35, cuSOLVER gesvdjBatched).  One thread per matrix, the same register
// Jacobi SVD as the Procrustes kernel, fp64 inside whatever the I/O type.
//   forward : R = U D V^T, D = diag(1, 1, d), d = sign(det(U V^T));  aux = U (9) V (9) s' (3) with
//             s'_i = s_i D_ii the signed singular values
//   backward: with P = U^T G V,  dA = U Q V^T,  Q_ij = (P_ij - D_ii D_jj P_ji) / (s'_i + s'_j), Q_ii = 0
//             (the derivative of the polar-type factor; what autograd through the SVD amounts to,
//             the det factor being locally constant).
// ---------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128)
    k_orthonormalize3_fwd(const T* __restrict__ A, long long N, T* __restrict__ R, double* __restrict__ aux) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    Mat3 M, U, V;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) M.a[i][k] = (double)A[n * 9 + i * 3 + k];
    double s[3];
    svd3(M, U, s, V);
    Mat3 O;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            O.a[i][k] = U.a[i][0] * V.a[k][0] + U.a[i][1] * V.a[k][1] + U.a[i][2] * V.a[k][2];
    const double d = (det3(O) < 0.0) ? -1.0 : 1.0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            R[n * 9 + i * 3 + k] = (T)(U.a[i][0] * V.a[k][0] + U.a[i][1] * V.a[k][1] + d * U.a[i][2] * V.a[k][2]);
    if (aux) {
        double* a = aux + n * 21;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                a[i * 3 + k] = U.a[i][k];
                a[9 + i * 3 + k] = V.a[i][k];
            }
        a[18] = s[0]; a[19] = s[1]; a[20] = d * s[2];
    }
}

template <typename T>
__global__ void __launch_bounds__(128)
    k_orthonormalize3_bwd(const T* __restrict__ G, long long N, const double* __restrict__ aux, T* __restrict__ dA) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const double* a = aux + n * 21;
    Mat3 U, V, Gm, P, Q;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            U.a[i][k] = a[i * 3 + k];
            V.a[i][k] = a[9 + i * 3 + k];
            Gm.a[i][k] = (double)G[n * 9 + i * 3 + k];
        }
    const double sp[3] = {a[18], a[19], a[20]};
    const double dd[3] = {1.0, 1.0, (sp[2] < 0.0) ? -1.0 : 1.0};
    // P = U^T G V
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double acc = 0.0;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) acc += U.a[r][i] * Gm.a[r][c] * V.a[c][k];
            P.a[i][k] = acc;
        }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            Q.a[i][k] = (i == k) ? 0.0 : (P.a[i][k] - dd[i] * dd[k] * P.a[k][i]) / (sp[i] + sp[k]);
    // dA = U Q V^T
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double acc = 0.0;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) acc += U.a[i][r] * Q.a[r][c] * V.a[k][c];
            dA[n * 9 + i * 3 + k] = (T)acc;
        }
}

}  // namespace hipe

using namespace hipe;

extern "C" int hipe_orthonormalize3_fwd(const void* A, int dtype, long long N, void* R, double* aux, void* stream) {
    if (N < 0) return HIPE_E_SHAPE;
    if (N == 0) return 0;
    if (!A || !R) return HIPE_E_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)((N + 127) / 128);
    if (dtype == HIPE_F32)
        k_orthonormalize3_fwd<float><<<grid, 128, 0, st>>>(static_cast<const float*>(A), N, static_cast<float*>(R), aux);
    else if (dtype == HIPE_F64)
        k_orthonormalize3_fwd<double><<<grid, 128, 0, st>>>(static_cast<const double*>(A), N, static_cast<double*>(R), aux);
    else
        return HIPE_E_DTYPE;
    return (int)cudaGetLastError();
}

extern "C" int hipe_orthonormalize3_bwd(const void* G, int dtype, long long N, const double* aux, void* dA,
                                        void* stream) {
    if (N < 0) return HIPE_E_SHAPE;
    if (N == 0) return 0;
    if (!G || !aux || !dA) return HIPE_E_NULL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)((N + 127) / 128);
    if (dtype == HIPE_F32)
        k_orthonormalize3_bwd<float><<<grid, 128, 0, st>>>(static_cast<const float*>(G), N, aux, static_cast<float*>(dA));
    else if (dtype == HIPE_F64)
        k_orthonormalize3_bwd<double><<<grid, 128, 0, st>>>(static_cast<const double*>(G), N, aux, static_cast<double*>(dA));
    else
        return HIPE_E_DTYPE;
    return (int)cudaGetLastError();
}

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

template <typename S>
static int scale_inplace_impl(void* x, int dtype, long long n, const S* gscale, void* stream) {
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
        k_scale_inplace<float, S><<<(unsigned)grid, 256, 0, st>>>(static_cast<unsigned char*>(x), n_vec, n_vec * per, n, gscale);
    else
        k_scale_inplace<__nv_bfloat16, S><<<(unsigned)grid, 256, 0, st>>>(static_cast<unsigned char*>(x), n_vec, n_vec * per, n, gscale);
    return (int)cudaGetLastError();
}

extern "C" int hipe_scale_inplace(void* x, int dtype, long long n, const float* gscale, void* stream) {
    return scale_inplace_impl<float>(x, dtype, n, gscale, stream);
}

extern "C" int hipe_scale_inplace_s(void* x, int dtype, long long n, const void* gscale, int gscale_dtype,
                                    void* stream) {
    if (gscale_dtype == HIPE_F32) return scale_inplace_impl<float>(x, dtype, n, static_cast<const float*>(gscale), stream);
    if (gscale_dtype == HIPE_F64) return scale_inplace_impl<double>(x, dtype, n, static_cast<const double*>(gscale), stream);
    return HIPE_E_DTYPE;
}

extern "C" int hipe_procrustes_batched(const double* X, const double* Y, long long N, int J, int mode,
                                       int optimal_scale, double* d, double* Z, double* T, double* b,
                                       double* c, double* pa_dist, double* raw_dist, void* stream) {
    if (!X || !Y) return HIPE_E_NULL;
    if (N < 0 || J <= 0) return HIPE_E_SHAPE;
    if (mode != 0 && mode != 1) return HIPE_E_SHAPE;
    if (N == 0) return 0;
    const unsigned grid = (unsigned)((N + kProcWarps - 1) / kProcWarps);
    k_procrustes<false, true><<<grid, kProcWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        ChainArgs{}, X, Y, N, J, mode, optimal_scale, d, Z, T, b, c, pa_dist, raw_dist);
    return (int)cudaGetLastError();
}

static int chain_args(ChainArgs& c, const void* coords, int coords_kind, int coords_stride, long long N, int J,
                      double patch_w, double patch_h, const double* trans, const double* bbox, double box_scale,
                      double dst_w, double dst_h, const double* tprime, const double* K, const double* R,
                      const double* bone_len, int joint_a, int joint_b, double* img, double* cam) {
    if (!coords || !tprime || !K || (!trans && !bbox)) return HIPE_E_NULL;
    if (N < 0 || J <= 0 || J > 512) return HIPE_E_SHAPE;   // 4 warps x J x 3 doubles of static-limit shared memory
    if (coords_kind != 0 && coords_kind != 1) return HIPE_E_DTYPE;
    if (coords_kind == 1 && coords_stride < 3) return HIPE_E_SHAPE;
    if (bone_len && (joint_a < 0 || joint_b < 0 || joint_a >= J || joint_b >= J)) return HIPE_E_SHAPE;
    c.coords = coords; c.kind = coords_kind; c.stride = coords_stride;
    c.patch_w = patch_w; c.patch_h = patch_h;
    c.trans = trans; c.bbox = bbox; c.box_scale = box_scale; c.dst_w = dst_w; c.dst_h = dst_h;
    c.tprime = tprime; c.K = K; c.R = R; c.bone_len = bone_len; c.ja = joint_a; c.jb = joint_b;
    c.img_out = img; c.cam_out = cam;
    return 0;
}

extern "C" int hipe_patch_to_camera(const void* coords, int coords_kind, int coords_stride, long long N, int J,
                                    double patch_w, double patch_h, const double* trans, const double* bbox,
                                    double box_scale, double dst_w, double dst_h, const double* tprime,
                                    const double* K, const double* R, const double* bone_len, int joint_a,
                                    int joint_b, double* img, double* cam, void* stream) {
    ChainArgs c;
    const int rc = chain_args(c, coords, coords_kind, coords_stride, N, J, patch_w, patch_h, trans, bbox, box_scale,
                              dst_w, dst_h, tprime, K, R, bone_len, joint_a, joint_b, img, cam);
    if (rc) return rc;
    if (N == 0) return 0;
    const unsigned grid = (unsigned)((N + kProcWarps - 1) / kProcWarps);
    const size_t smem = (size_t)kProcWarps * J * 3 * sizeof(double);
    k_procrustes<true, false><<<grid, kProcWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        c, nullptr, nullptr, N, J, 0, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    return (int)cudaGetLastError();
}

extern "C" int hipe_eval_chain_procrustes(const void* coords, int coords_kind, int coords_stride, long long N, int J,
                                          double patch_w, double patch_h, const double* trans, const double* bbox,
                                          double box_scale, double dst_w, double dst_h, const double* tprime,
                                          const double* K, const double* R, const double* bone_len, int joint_a,
                                          int joint_b, const double* X, int mode, int optimal_scale, double* cam,
                                          double* Z, double* pa_dist, double* raw_dist, void* stream) {
    ChainArgs c;
    const int rc = chain_args(c, coords, coords_kind, coords_stride, N, J, patch_w, patch_h, trans, bbox, box_scale,
                              dst_w, dst_h, tprime, K, R, bone_len, joint_a, joint_b, nullptr, cam);
    if (rc) return rc;
    if (!X) return HIPE_E_NULL;
    if (mode != 0 && mode != 1) return HIPE_E_SHAPE;
    if (N == 0) return 0;
    const unsigned grid = (unsigned)((N + kProcWarps - 1) / kProcWarps);
    const size_t smem = (size_t)kProcWarps * J * 3 * sizeof(double);
    k_procrustes<true, true><<<grid, kProcWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        c, X, nullptr, N, J, mode, optimal_scale, nullptr, Z, nullptr, nullptr, nullptr, pa_dist, raw_dist);
    return (int)cudaGetLastError();
}

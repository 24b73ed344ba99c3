// Host-buffer entry points of libhipe_b200.so: chunked, copy/compute-overlapped drivers
// around the device-pointer API, plus version / error strings.  These are what a non-PyTorch
// caller (or the end-to-end measurement) binds; they own only transient stream-ordered
// scratch from the device's default memory pool.
#include "hipe_common.cuh"
#include <stdint.h>

#define HIPE_CK(call)                             \
    do {                                          \
        cudaError_t e_ = (call);                  \
        if (e_ != cudaSuccess) {                  \
            rc = (int)e_;                         \
            goto done;                            \
        }                                         \
    } while (0)
#define HIPE_RC(call)                             \
    do {                                          \
        rc = (call);                              \
        if (rc != 0) goto done;                   \
    } while (0)

extern "C" int hipe_version(void) { return 100; }

extern "C" const char* hipe_strerror(int code) {
    switch (code) {
        case 0: return "ok";
        case HIPE_E_DTYPE: return "unsupported element type";
        case HIPE_E_SHAPE: return "bad shape or size argument";
        case HIPE_E_ALIGN: return "pointer not aligned to its element size";
        case HIPE_E_NULL: return "required pointer is NULL";
        case HIPE_E_CLUSTER: return "cluster size must be 0, 1, 2, 4, 8 or 16";
        case HIPE_E_ARCH: return "current device is not sm_100 (Blackwell B200)";
        case HIPE_E_UNSUPPORTED: return "geometry or launch not supported by this entry point";
        case HIPE_E_LIMIT: return "size beyond the fixed scratch of this no-alloc entry point (see the limits in hipe_b200.h); use the general entry point";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown hipe error";
}

static void keep_pool_memory() {
    // let the default pool retain freed scratch between calls (otherwise every call pays
    // cudaMalloc-class latency)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) != cudaSuccess) return;
    uint64_t thr = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
}

extern "C" int hipe_joint_location_loss_host(const void* logits_host, int dtype, long long B, int J,
                                             int D, int H, int W, const void* gt_host,
                                             const void* vis_host, int tgt_dtype, int size_average,
                                             double* loss_host, float* coords_host,
                                             void* dlogits_host, void* dlogits_dev,
                                             long long chunk_samples) {
    if (!logits_host || !gt_host || !vis_host || !loss_host) return HIPE_E_NULL;
    if (dtype != HIPE_F32 && dtype != HIPE_BF16) return HIPE_E_DTYPE;
    if (tgt_dtype != HIPE_F32 && tgt_dtype != HIPE_F64) return HIPE_E_DTYPE;
    if (B <= 0 || J <= 0 || D <= 0 || H <= 0 || W <= 0) return HIPE_E_SHAPE;
    const size_t esz = dtype == HIPE_F32 ? 4 : 2;
    const size_t tsz = tgt_dtype == HIPE_F64 ? 8 : 4;
    const size_t sample_bytes = (size_t)J * D * H * W * esz;
    const int K = 3 * J;
    if (chunk_samples <= 0) {
        // ~128 MB per chunk: long enough to amortise launches, short enough to pipeline
        chunk_samples = (long long)((128ull << 20) / sample_bytes);
        if (chunk_samples < 1) chunk_samples = 1;
    }
    if (chunk_samples > B) chunk_samples = B;
    const double norm = size_average ? 1.0 / (double)B : 1.0;
    const bool to_host = dlogits_host != nullptr;
    const bool want_grad = to_host || dlogits_dev != nullptr;
    // with a caller-owned device gradient buffer the backward writes straight into it
    const bool stage_grad = to_host && dlogits_dev == nullptr;
    constexpr int NBUF = 3;

    int rc = 0;
    cudaStream_t s_in = nullptr, s_cmp = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[NBUF] = {}, ev_cmp[NBUF] = {}, ev_out[NBUF] = {};
    unsigned char* d_logits[NBUF] = {};
    unsigned char* d_grad[NBUF] = {};
    unsigned char *d_gt = nullptr, *d_vis = nullptr;
    float *d_coords = nullptr, *d_stats = nullptr, *d_g = nullptr;
    double *d_expect = nullptr, *d_loss = nullptr;
    const long long nchunks = (B + chunk_samples - 1) / chunk_samples;

    keep_pool_memory();
    HIPE_CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
    HIPE_CK(cudaStreamCreateWithFlags(&s_cmp, cudaStreamNonBlocking));
    HIPE_CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
    for (int i = 0; i < NBUF; ++i) {
        HIPE_CK(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
        HIPE_CK(cudaEventCreateWithFlags(&ev_cmp[i], cudaEventDisableTiming));
        HIPE_CK(cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming));
        HIPE_CK(cudaMallocAsync((void**)&d_logits[i], chunk_samples * sample_bytes, s_cmp));
        if (stage_grad) HIPE_CK(cudaMallocAsync((void**)&d_grad[i], chunk_samples * sample_bytes, s_cmp));
    }
    HIPE_CK(cudaMallocAsync((void**)&d_gt, B * K * tsz, s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_vis, B * K * tsz, s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_coords, B * K * sizeof(float), s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_expect, B * K * sizeof(double), s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_stats, B * J * 2 * sizeof(float), s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_g, B * K * sizeof(float), s_cmp));
    HIPE_CK(cudaMallocAsync((void**)&d_loss, 2 * sizeof(double), s_cmp));
    HIPE_CK(cudaMemcpyAsync(d_gt, gt_host, B * K * tsz, cudaMemcpyHostToDevice, s_cmp));
    HIPE_CK(cudaMemcpyAsync(d_vis, vis_host, B * K * tsz, cudaMemcpyHostToDevice, s_cmp));
    // allocations are ordered on s_cmp: make the copy streams wait for them
    HIPE_CK(cudaEventRecord(ev_cmp[0], s_cmp));
    HIPE_CK(cudaStreamWaitEvent(s_in, ev_cmp[0], 0));
    HIPE_CK(cudaStreamWaitEvent(s_out, ev_cmp[0], 0));

    for (long long c = 0; c < nchunks; ++c) {
        const int bi = (int)(c % NBUF);
        const long long b0 = c * chunk_samples;
        const long long nb = (B - b0 < chunk_samples) ? (B - b0) : chunk_samples;
        const long long n_vol = nb * J;
        // the buffer pair is free once chunk c-NBUF has been computed (and copied out)
        if (c >= NBUF) {
            HIPE_CK(cudaStreamWaitEvent(s_in, ev_cmp[bi], 0));
            if (stage_grad) HIPE_CK(cudaStreamWaitEvent(s_cmp, ev_out[bi], 0));
        }
        HIPE_CK(cudaMemcpyAsync(d_logits[bi], (const unsigned char*)logits_host + b0 * sample_bytes,
                                nb * sample_bytes, cudaMemcpyHostToDevice, s_in));
        HIPE_CK(cudaEventRecord(ev_in[bi], s_in));
        HIPE_CK(cudaStreamWaitEvent(s_cmp, ev_in[bi], 0));
        HIPE_RC(hipe_softargmax_fwd(d_logits[bi], dtype, n_vol, D, H, W, 1, d_coords + b0 * K,
                                    d_expect + b0 * K, d_stats + b0 * J * 2, 0, 0, s_cmp));
        unsigned char* grad_dst = stage_grad ? d_grad[bi]
                                             : (unsigned char*)dlogits_dev + b0 * sample_bytes;
        if (want_grad) {
            HIPE_RC(hipe_l1_loss_fwd(d_coords + b0 * K, d_gt + b0 * K * tsz, d_vis + b0 * K * tsz,
                                     tgt_dtype, nb, K, norm, d_loss + 1, d_g + b0 * K, s_cmp));
            HIPE_RC(hipe_softargmax_bwd(d_logits[bi], dtype, n_vol, D, H, W, d_stats + b0 * J * 2,
                                        d_expect + b0 * K, d_g + b0 * K, nullptr, grad_dst, 0, 0, s_cmp));
        }
        HIPE_CK(cudaEventRecord(ev_cmp[bi], s_cmp));
        if (to_host) {
            HIPE_CK(cudaStreamWaitEvent(s_out, ev_cmp[bi], 0));
            HIPE_CK(cudaMemcpyAsync((unsigned char*)dlogits_host + b0 * sample_bytes, grad_dst,
                                    nb * sample_bytes, cudaMemcpyDeviceToHost, s_out));
            HIPE_CK(cudaEventRecord(ev_out[bi], s_out));
        }
    }
    // the scalar: one fixed-order fp64 reduction over all B*3J terms (loss.py:86-89)
    HIPE_RC(hipe_l1_loss_fwd(d_coords, d_gt, d_vis, tgt_dtype, B, K, norm, d_loss, nullptr, s_cmp));
    HIPE_CK(cudaMemcpyAsync(loss_host, d_loss, sizeof(double), cudaMemcpyDeviceToHost, s_cmp));
    if (coords_host)
        HIPE_CK(cudaMemcpyAsync(coords_host, d_coords, B * K * sizeof(float), cudaMemcpyDeviceToHost, s_cmp));
    HIPE_CK(cudaStreamSynchronize(s_out));
    HIPE_CK(cudaStreamSynchronize(s_cmp));

done:
    if (s_cmp) {
        cudaStreamSynchronize(s_cmp);
        if (s_in) cudaStreamSynchronize(s_in);
        if (s_out) cudaStreamSynchronize(s_out);
        for (int i = 0; i < NBUF; ++i) {
            if (d_logits[i]) cudaFreeAsync(d_logits[i], s_cmp);
            if (d_grad[i]) cudaFreeAsync(d_grad[i], s_cmp);
        }
        void* small[] = {d_gt, d_vis, d_coords, (void*)d_expect, d_stats, d_g, d_loss};
        for (void* q : small)
            if (q) cudaFreeAsync(q, s_cmp);
        cudaStreamSynchronize(s_cmp);
    }
    for (int i = 0; i < NBUF; ++i) {
        if (ev_in[i]) cudaEventDestroy(ev_in[i]);
        if (ev_cmp[i]) cudaEventDestroy(ev_cmp[i]);
        if (ev_out[i]) cudaEventDestroy(ev_out[i]);
    }
    if (s_in) cudaStreamDestroy(s_in);
    if (s_cmp) cudaStreamDestroy(s_cmp);
    if (s_out) cudaStreamDestroy(s_out);
    return rc;
}

extern "C" int hipe_procrustes_host(const double* X_host, const double* Y_host, long long N, int J,
                                    int mode, int optimal_scale, double* Z_host, double* pa_mpjpe,
                                    double* mpjpe) {
    if (!X_host || !Y_host) return HIPE_E_NULL;
    if (N <= 0 || J <= 0) return HIPE_E_SHAPE;
    int rc = 0;
    cudaStream_t st = nullptr;
    double *dX = nullptr, *dY = nullptr, *dZ = nullptr, *dpa = nullptr, *draw = nullptr, *dout = nullptr;
    const size_t nb = (size_t)N * J * 3 * sizeof(double);
    double out[2] = {0, 0};
    keep_pool_memory();
    HIPE_CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    HIPE_CK(cudaMallocAsync((void**)&dX, nb, st));
    HIPE_CK(cudaMallocAsync((void**)&dY, nb, st));
    HIPE_CK(cudaMallocAsync((void**)&dZ, nb, st));
    HIPE_CK(cudaMallocAsync((void**)&dpa, (size_t)N * J * sizeof(double), st));
    HIPE_CK(cudaMallocAsync((void**)&draw, (size_t)N * J * sizeof(double), st));
    HIPE_CK(cudaMallocAsync((void**)&dout, 2 * sizeof(double), st));
    HIPE_CK(cudaMemcpyAsync(dX, X_host, nb, cudaMemcpyHostToDevice, st));
    HIPE_CK(cudaMemcpyAsync(dY, Y_host, nb, cudaMemcpyHostToDevice, st));
    HIPE_RC(hipe_procrustes_batched(dX, dY, N, J, mode, optimal_scale, nullptr, dZ, nullptr, nullptr,
                                    nullptr, dpa, draw, st));
    // data/FreiHand/FreiHand.py:804-807: mean over all samples and joints
    HIPE_RC(hipe_sum_f64(dpa, N * J, 1.0 / (double)(N * J), dout, st));
    HIPE_RC(hipe_sum_f64(draw, N * J, 1.0 / (double)(N * J), dout + 1, st));
    HIPE_CK(cudaMemcpyAsync(out, dout, sizeof(out), cudaMemcpyDeviceToHost, st));
    if (Z_host) HIPE_CK(cudaMemcpyAsync(Z_host, dZ, nb, cudaMemcpyDeviceToHost, st));
    HIPE_CK(cudaStreamSynchronize(st));
    if (pa_mpjpe) *pa_mpjpe = out[0];
    if (mpjpe) *mpjpe = out[1];
done:
    if (st) {
        void* bufs[] = {dX, dY, dZ, dpa, draw, dout};
        for (void* q : bufs)
            if (q) cudaFreeAsync(q, st);
        cudaStreamSynchronize(st);
        cudaStreamDestroy(st);
    }
    return rc;
}

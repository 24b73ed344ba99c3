/*
 * hipe_b200.h -- C ABI of libhipe_b200.so: the integral-regression hot path of
 * mqadri9/hand-integral-pose-estimation as hand-written sm_100a CUDA kernels.
 *
 * Conventions:
 *   - every pointer named *_dev is DEVICE memory owned by the caller; the library allocates
 *     nothing on those calls, keeps no global mutable state and never synchronises the
 *     device.  Work is enqueued on `stream` (a cudaStream_t passed as void*) of the CURRENT
 *     device of the calling thread.  Entry points are re-entrant and thread-safe (the
 *     reference runs one Python thread per GPU, common/nets/balanced_parallel.py:149-173).
 *   - the *_host entry points take HOST pointers, allocate/free their own device scratch,
 *     copy in, compute, copy out and synchronise; they exist for non-PyTorch callers and
 *     for the end-to-end measurement.
 *   - return value: 0 = ok; < 0 = argument error (HIPE_E_*); > 0 = a cudaError_t.
 *   - a heatmap tensor is `n_vol` back-to-back volumes of D*H*W elements, x fastest, which
 *     is exactly main/model.py:14-20,40-44's contiguous (B, J*D, H, W) output
 *     (n_vol = B*J, channel = j*D + z).
 *
 * Citations are into /root/reference (the replaced Python interface).
 */
#ifndef HIPE_B200_H
#define HIPE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define HIPE_API __attribute__((visibility("default")))
#else
#define HIPE_API
#endif

/* element types */
#define HIPE_F32 0
#define HIPE_BF16 1
#define HIPE_F64 3

/* argument errors */
#define HIPE_E_DTYPE (-1)    /* unsupported element type for this argument            */
#define HIPE_E_SHAPE (-2)    /* non-positive / overflowing dimension                  */
#define HIPE_E_ALIGN (-3)    /* pointer not aligned to its element size               */
#define HIPE_E_NULL (-4)     /* required pointer is NULL                              */
#define HIPE_E_CLUSTER (-5)  /* cluster size not in {0,1,2,4,8,16}                    */
#define HIPE_E_ARCH (-6)     /* current device is not compute capability 10.x         */
#define HIPE_E_UNSUPPORTED (-7) /* this entry point cannot take the geometry / launch: use the general one */
#define HIPE_E_LIMIT (-8)    /* size beyond the fixed scratch of a *_noalloc entry point: use the general one */

/* variant selector for the streaming kernels (0 = library default) */
#define HIPE_VARIANT_AUTO 0
#define HIPE_VARIANT_TMA 1   /* cp.async.bulk + mbarrier ring through shared memory   */
#define HIPE_VARIANT_LDG 2   /* 128-bit ld.global.nc straight to registers            */

HIPE_API int hipe_version(void);

/* Short static description of an error code returned by this library. */
HIPE_API const char* hipe_strerror(int code);

/*
 * Soft-argmax forward: replaces softmax_integral_tensor (common/nets/loss.py:46-59) and,
 * with `apply_softmax == 0`, generate_3d_integral_preds_tensor (common/nets/loss.py:23-43)
 * on already-normalised heatmaps.
 *   logits_dev : n_vol * D*H*W elements of `dtype` (HIPE_F32 | HIPE_BF16), read-only
 *   coords_dev : n_vol*3 floats  (x/W-0.5, y/H-0.5, z/D-0.5)  -- loss.py:54-58   (nullable)
 *   expect_dev : n_vol*3 DOUBLES raw expectations in voxel units -- loss.py:39-41 (nullable);
 *                fp64 because the backward needs (idx - E) to ~1e-7 voxel, below fp32's
 *                resolution at E ~ 32
 *   stats_dev  : n_vol*2 floats  (max logit, sum exp(logit-max)); input of the backward
 *                (nullable; written as (0, sum) when apply_softmax == 0)
 *   cluster    : CTAs cooperating on one volume: 0 = auto, else 1|2|4|8|16
 *   variant    : HIPE_VARIANT_*
 */
HIPE_API int hipe_softargmax_fwd(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                        int apply_softmax, float* coords_dev, double* expect_dev,
                        float* stats_dev, int cluster, int variant, void* stream);

/*
 * Soft-argmax backward: what autograd does for loss.backward() through
 * common/nets/loss.py:46-59 (main/train.py:113), recomputing the softmax from the logits:
 *   dlogits[i] = p_i * ( gx (x_i - Ex)/W + gy (y_i - Ey)/H + gz (z_i - Ez)/D ) * gscale
 *   gcoords_dev : n_vol*3 floats, upstream gradient w.r.t. coords
 *   gscale_dev  : optional device float multiplied into gcoords (NULL = 1)
 *   dlogits_dev : n_vol * D*H*W elements of `dtype`
 */
HIPE_API int hipe_softargmax_bwd(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                        const float* stats_dev, const double* expect_dev,
                        const float* gcoords_dev, const float* gscale_dev, void* dlogits_dev,
                        int cluster, int variant, void* stream);

/*
 * Volume softmax: replaces softmax_integral_tensor2 (common/nets/loss.py:61-67); needs the
 * stats a forward call produced.  probs_dev: n_vol * D*H*W elements of `dtype`.
 */
HIPE_API int hipe_volume_softmax(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                        const float* stats_dev, void* probs_dev, int cluster, int variant,
                        void* stream);

/*
 * Per-volume entropy of the softmax distribution, -sum p ln p: replaces the chain
 * softmax_integral_tensor2 (common/nets/loss.py:61-67) -> .cpu().numpy() -> scipy.stats.entropy per
 * joint of main/generate_correlation_graph.py:116-141.  One streaming read of the logits; needs the
 * stats a forward call produced.  entropy_dev: n_vol doubles (nats).
 */
HIPE_API int hipe_volume_entropy(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                        const float* stats_dev, double* entropy_dev, void* stream);

/*
 * L1 integral loss on coordinates: replaces common/nets/loss.py:86-89.
 *   coords_dev : B*K floats (K = 3*J);  gt_dev, vis_dev : B*K of tgt_dtype (HIPE_F32|HIPE_F64)
 *   norm       : the value the masked L1 sum is multiplied by (1/B for size_average)
 *   loss_dev   : one double, sum |c-gt|*vis * norm, accumulated in fp64 in a fixed order
 *   gcoords_dev: B*K floats, sign(c-gt)*vis*norm  (torch.abs backward is sgn)   (nullable)
 */
HIPE_API int hipe_l1_loss_fwd(const float* coords_dev, const void* gt_dev, const void* vis_dev,
                     int tgt_dtype, long long B, int K, double norm, double* loss_dev,
                     float* gcoords_dev, void* stream);

/*
 * Fused JointLocationLoss forward+backward in ONE launch per batch
 * (common/nets/loss.py:74-91 + main/train.py:113 with upstream gradient 1): every cluster
 * reduces its volume, forms the three local gradients sign(c-gt)*vis*norm, and streams the
 * same volume again (from L2 / shared memory) to write dlogits.  Also writes coords,
 * expect, stats and per-coordinate loss terms; hipe_sum_f64 / hipe_l1_loss_fwd finish the
 * scalar.  gt/vis are indexed [vol*3 + a], i.e. the (B, 3J) target layout.
 */
HIPE_API int hipe_joint_loss_fused(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                          const void* gt_dev, const void* vis_dev, int tgt_dtype, double norm,
                          float* coords_dev, double* expect_dev, float* stats_dev,
                          double* loss_terms_dev /* n_vol*3 */, void* dlogits_dev,
                          int cluster, int variant, void* stream);

/*
 * The same contract as hipe_joint_loss_fused, as ONE persistent cooperative launch that keeps
 * every volume on chip between the reduction and the gradient write: HBM sees each volume once
 * in and its gradient once out (V*s + V*g bytes per volume instead of 2*V*s + V*g).  A volume is
 * cut into items of <= 128 KB taken by different CTAs (one CTA per SM, 13 x 16 KB TMA ring);
 * the items' partial sums are merged through `workspace_dev`.
 *   workspace_dev   : caller-owned device scratch, 256-byte aligned, of at least
 *                     hipe_resident_workspace_bytes(...) bytes; holds nothing between calls.
 *                     Word [n_vol] (uint32) is set non-zero if an inter-CTA wait ever timed out.
 * Returns HIPE_E_UNSUPPORTED when the geometry cannot be tiled for it (rows not a multiple of
 * 16 bytes, stage rows neither dividing nor a multiple of H, ...) or the grid cannot be made
 * co-resident; callers then use hipe_joint_loss_fused.  hipe_resident_workspace_bytes returns 0
 * in exactly the geometry cases.
 */
HIPE_API long long hipe_resident_workspace_bytes(int dtype, long long n_vol, int D, int H, int W);
HIPE_API int hipe_joint_loss_resident(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                             const void* gt_dev, const void* vis_dev, int tgt_dtype, double norm,
                             float* coords_dev, double* expect_dev, float* stats_dev,
                             double* loss_terms_dev /* n_vol*3 */,
                             double* loss_dev /* nullable: norm * sum(loss_terms), by the last CTA */,
                             void* dlogits_dev, void* workspace_dev, long long workspace_bytes,
                             void* stream);

/*
 * The same contract again, as ONE persistent launch of thread-block CLUSTERS (C CTAs x up to 14
 * 16 KB TMA slots; the C partial sums travel CTA to CTA through st.async into distributed shared
 * memory): HBM sees each volume once in and its gradient once out, and no role waits on global
 * memory.  Two forms, chosen per geometry (hipe_cluster_plan reports which):
 *   resident  the volume stays in the cluster's shared memory between the reduction and the gradient
 *             write, the L2 holds nothing (bf16 64^3 on 4 CTAs, fp32 56^3 on 6, 32^3 on 1, ...);
 *   streamed  two rings per CTA: the reduce team's is fed from HBM, the write team's re-reads the same
 *             stages through L2 at most `lead` stages later, the lead sized so that what has been read
 *             once and not yet twice stays inside ~40 MB (fp32 64^3 on 8 CTAs; volumes no cluster can
 *             hold: fp32 96^3 on 16 CTAs, the one non-portable cluster size used).
 *   workspace_dev : >= hipe_cluster_workspace_bytes(...) (256) bytes, 256-byte aligned
 * Returns HIPE_E_UNSUPPORTED when the geometry cannot be tiled (hipe_cluster_workspace_bytes returns 0
 * in exactly those cases); hipe_cluster_plan reports the cluster size, ring slots per CTA, stages
 * per CTA share, bytes per stage, and the lead in stages (0 = resident form).
 */
HIPE_API long long hipe_cluster_workspace_bytes(int dtype, long long n_vol, int D, int H, int W);
HIPE_API int hipe_cluster_plan(int dtype, long long n_vol, int D, int H, int W, int* cluster, int* nslots,
                      int* share_stages, int* stage_bytes, int* lead_stages);
HIPE_API int hipe_joint_loss_cluster(const void* logits_dev, int dtype, long long n_vol, int D, int H, int W,
                            const void* gt_dev, const void* vis_dev, int tgt_dtype, double norm,
                            float* coords_dev, double* expect_dev, float* stats_dev,
                            double* loss_terms_dev /* n_vol*3 */,
                            double* loss_dev /* nullable: norm * sum(loss_terms), by the last cluster */,
                            void* dlogits_dev, void* workspace_dev, long long workspace_bytes,
                            void* stream);

/*
 * x[i] *= *gscale_dev for n elements of `dtype`, in place; returns without touching memory when
 * the scalar is exactly 1.  Used to apply an upstream loss gradient to the dlogits the fused
 * launch already produced, without a host synchronisation.
 */
HIPE_API int hipe_scale_inplace(void* x_dev, int dtype, long long n, const float* gscale_dev, void* stream);
/* the same with the scalar read as HIPE_F32 or HIPE_F64 (the loss, hence its gradient, is fp64 in the
 * reference: no conversion launch in front) */
HIPE_API int hipe_scale_inplace_s(void* x_dev, int dtype, long long n, const void* gscale_dev, int gscale_dtype,
                         void* stream);

/* Deterministic fp64 sum of n doubles into out_dev[0] (times `scale`). */
HIPE_API int hipe_sum_f64(const double* in_dev, long long n, double scale, double* out_dev, void* stream);

/*
 * Batched similarity (Procrustes) alignment: replaces the per-sample loop around
 * augment.compute_similarity_transform (common/augment.py:18-77, called at
 * data/FreiHand/FreiHand.py:742) and, with mode == 1, eval.align_w_scale
 * (main/eval.py:142-165).  All arrays fp64.
 *   X_dev, Y_dev : N*J*3 targets / inputs
 *   mode         : 0 = compute_similarity_transform (reflection repaired),
 *                  1 = align_w_scale (no repair, 1e-8 added to the norms)
 *   optimal_scale: compute_optimal_scale of augment.py:18 (mode 0 only)
 *   d N, Z N*J*3, T N*9 (row-major), b N, c N*3      -- each nullable
 *   pa_dist N*J : ||Z - X||_2 per joint;  raw_dist N*J : ||Y - X||_2 per joint
 *                 (data/FreiHand/FreiHand.py:766-767,804-807)   -- each nullable
 */
HIPE_API int hipe_procrustes_batched(const double* X_dev, const double* Y_dev, long long N, int J,
                            int mode, int optimal_scale, double* d_dev, double* Z_dev,
                            double* T_dev, double* b_dev, double* c_dev, double* pa_dist_dev,
                            double* raw_dist_dev, void* stream);

/*
 * Patch -> camera coordinate chain that sits between the soft-argmax and the alignment in the
 * reference's evaluators (data/FreiHand/FreiHand.py:683-737 and :853-883), per joint in fp64:
 *   get_joint_location_result rescale (common/augment.py:176-178; coords_kind 0 only),
 *   trans_coords_from_patch_to_org_3d (common/augment.py:186-217,508-511),
 *   pixel2cam (common/augment.py:89-103), R^T (FreiHand.py:722; R_dev nullable),
 *   scale_result(method='scale') on the bone joint_a-joint_b (FreiHand.py:632-653; bone_len_dev nullable).
 *   coords_dev   : coords_kind 0: N*3J floats, the soft-argmax output (x, y, z in [-0.5, 0.5));
 *                  coords_kind 1: N*J*coords_stride doubles, patch pixels (stride 4 = with score)
 *   trans_dev    : N*9 row-major patch->image matrices (gts[n]['trans']), or NULL to build
 *                  gen_trans_from_patch_cv(cx, cy, w, h, dst_w, dst_h, box_scale, inv=True)
 *                  (common/augment.py:457-506) from bbox_dev N*4 = (cx, cy, w, h)
 *   tprime_dev N, K_dev N*9, R_dev N*9, bone_len_dev N   (all fp64)
 *   img_dev N*J*3 : image-space (u, v, z_mm) (nullable);  cam_dev N*J*3 : camera space, metres (nullable)
 * J <= 512.
 */
HIPE_API int hipe_patch_to_camera(const void* coords_dev, int coords_kind, int coords_stride, long long N, int J,
                         double patch_w, double patch_h, const double* trans_dev, const double* bbox_dev,
                         double box_scale, double dst_w, double dst_h, const double* tprime_dev,
                         const double* K_dev, const double* R_dev, const double* bone_len_dev, int joint_a,
                         int joint_b, double* img_dev, double* cam_dev, void* stream);

/*
 * The same chain fused in front of hipe_procrustes_batched (one launch, the camera-space joints
 * stay in shared memory): the per-sample body of FreiHand.evaluate (FreiHand.py:703-767).
 *   X_dev N*J*3 : ground-truth joints (gt_3d_kpt);  cam_dev, Z_dev, pa_dist_dev, raw_dist_dev nullable.
 */
HIPE_API int hipe_eval_chain_procrustes(const void* coords_dev, int coords_kind, int coords_stride, long long N,
                               int J, double patch_w, double patch_h, const double* trans_dev,
                               const double* bbox_dev, double box_scale, double dst_w, double dst_h,
                               const double* tprime_dev, const double* K_dev, const double* R_dev,
                               const double* bone_len_dev, int joint_a, int joint_b, const double* X_dev,
                               int mode, int optimal_scale, double* cam_dev, double* Z_dev,
                               double* pa_dist_dev, double* raw_dist_dev, void* stream);

/*
 * FreiHAND keypoint measures: replaces EvalUtil.feed + EvalUtil.get_measures (main/eval.py:25-106)
 * on per-sample per-keypoint distances that are already on the device (pa_dist / raw_dist above).
 *   dist_dev      : N*K doubles, ||gt - pred||_2 per keypoint (main/eval.py:40-41)
 *   vis_dev       : N*K bytes, non-zero = the keypoint was visible and is fed (nullable = all)
 *   thresholds    : np.linspace(val_min, val_max, steps) (main/eval.py:65); steps <= 1024
 *   summary_dev   : 3 doubles: mean EPE, median EPE, AUC, each the mean over fed keypoints
 *   pck_curve_dev : `steps` doubles, PCK per threshold averaged over fed keypoints
 *   workspace_dev : hipe_keypoint_measures_workspace_bytes(K, steps) bytes of caller-owned scratch;
 *                   afterwards it holds, per keypoint, [count K][mean K][median K][auc K][pck K*steps]
 */
HIPE_API long long hipe_keypoint_measures_workspace_bytes(int K, int steps);
HIPE_API int hipe_keypoint_measures(const double* dist_dev, const unsigned char* vis_dev, long long N, int K,
                           double val_min, double val_max, int steps, double* summary_dev,
                           double* pck_curve_dev, double* workspace_dev, long long workspace_bytes,
                           void* stream);

/*
 * hipe_keypoint_measures without caller-owned scratch and without any allocation: NO COUNTERPART IN THE
 * REFERENCE -- added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two launches,
 * same results as hipe_keypoint_measures; the per-keypoint scratch is a file-scope __device__ array of
 * HIPE_NOALLOC_MEASURES_MAX_DOUBLES doubles (514 KiB; every GPU that loads the library holds it until
 * exit).  Calls neither cudaMalloc / cudaFree nor any allocating library routine (the device address of
 * the array is asked once per device, on that device's first call, with cudaGetSymbolAddress).
 *   limits : K <= HIPE_NOALLOC_MEASURES_MAX_K and K * (4 + steps) <= HIPE_NOALLOC_MEASURES_MAX_DOUBLES
 *            (steps <= 1024 as for the general entry point); beyond them returns HIPE_E_LIMIT
 *            ("size beyond the fixed scratch ...") and launches nothing.
 *   All calls on a device share the one array, which keeps the last call's contents (the part a call
 *   uses is cleared in stream order first): calls that OVERLAP on the GPU -- different streams without
 *   an event between them -- race on it.  One call at a time per device; use the general entry point
 *   with separate workspaces for concurrent streams.  The per-keypoint values are not exposed.
 */
#define HIPE_NOALLOC_MEASURES_MAX_K 64
#define HIPE_NOALLOC_MEASURES_MAX_DOUBLES (64 * (4 + 1024))
HIPE_API int hipe_keypoint_measures_noalloc(const double* dist_dev, const unsigned char* vis_dev, long long N, int K,
                                   double val_min, double val_max, int steps, double* summary_dev,
                                   double* pck_curve_dev, void* stream);

/*
 * Batched 3x3 orthonormalisation: replaces make_orthonormal
 * (procrustes_encoding/nrsfm/nrsfm_modules.py:46-67) and the SVD it takes from the external
 * torch_batch_svd extension (procrustes_encoding/nrsfm/batch_svd.py:5-35).
 *   A_dev N*9 (row-major 3x3) of `dtype` (HIPE_F32 | HIPE_F64) -> R_dev N*9 = U diag(1,1,det(UV^T)) V^T
 *   aux_dev N*21 doubles (U, V, signed singular values): input of the backward (nullable)
 *   backward: G_dev = dL/dR -> dA_dev = dL/dA
 */
HIPE_API int hipe_orthonormalize3_fwd(const void* A_dev, int dtype, long long N, void* R_dev, double* aux_dev,
                             void* stream);
HIPE_API int hipe_orthonormalize3_bwd(const void* G_dev, int dtype, long long N, const double* aux_dev,
                             void* dA_dev, void* stream);

/*
 * The exchange step of the batch-sharded path as ONE kernel over NVLink peer memory: replaces
 * Reduce.apply(*outputs) / len(outputs) (common/nets/balanced_parallel.py:46-55,127) and the
 * gather of the predicted joints.  Every rank calls it once per step with the same `seq`
 * (1, 2, 3, ...); rank r stores [in_scale * sum(loss_in[0..n_in)), B, coords (B x K)] (the sum in
 * the fixed order of hipe_sum_f64: n_in = 1 for a finished scalar, or the loss terms of
 * hipe_joint_loss_fused / _resident) into slot r of every
 * peer's buffer, publishes `seq`, waits (at most timeout_s, default 20 s) for all peers, and writes
 *   value_out[0] = value_scale * sum_r loss_sum_r   (fixed rank order, identical on every rank)
 *   sums_out[world] (nullable), coords_all_dev[world * B * K] doubles (nullable)
 *   peers_dev : DEVICE array of `world` pointers, entry p = rank p's symmetric buffer of
 *               hipe_peer_exchange_bytes(world, B, K) bytes (zeroed once before the first step),
 *               mapped into this process (torch.distributed._symmetric_memory / cudaIpc)
 *   err_dev   : set to 1 if a peer did not arrive in time (value_out is then NaN)   (nullable)
 *   scratch_dev : world + 1 doubles of caller-owned device scratch, zeroed once before the first step
 * Equal B on every rank (the training loop's equal shards); other cases use NCCL.
 */
HIPE_API long long hipe_peer_exchange_reclen(long long B, int K);
HIPE_API long long hipe_peer_exchange_bytes(int world, long long B, int K);
HIPE_API int hipe_peer_exchange(const double* loss_in_dev, long long n_in, double in_scale,
                       const float* coords_dev, long long B, int K, void* const* peers_dev, int world, int rank, unsigned long long seq,
                       double value_scale, double* value_out_dev, double* sums_out_dev,
                       double* coords_all_dev, unsigned* err_dev, double* scratch_dev, double timeout_s,
                       void* stream);

/*
 * The same exchange split in two launches, so that the rendezvous with the slowest rank leaves the
 * stream that runs the loss kernels:
 *   hipe_peer_publish  sums the loss terms, stores this rank's record into every peer's buffer and
 *                      publishes `seq`; never waits.
 *   hipe_peer_collect  waits for every peer's `seq` (bounded), copies the records out and combines
 *                      them exactly like hipe_peer_exchange; the caller launches it on a side stream.
 *                      err_dev: `world` words (device or host-mapped memory), word p set to 1 if
 *                      rank p did not arrive in time (value_out is then NaN).
 * Required ordering on each rank: publish(seq) is stream-ordered after collect(seq - 1).
 */
HIPE_API int hipe_peer_publish(const double* loss_in_dev, long long n_in, double in_scale, const float* coords_dev,
                      long long B, int K, void* const* peers_dev, int world, int rank, unsigned long long seq,
                      void* stream);
HIPE_API int hipe_peer_collect(void* const* peers_dev, int world, int rank, unsigned long long seq, long long B, int K,
                      double value_scale, double* value_out_dev, double* sums_out_dev, double* coords_all_dev,
                      unsigned* err_dev, double* scratch_dev, double timeout_s, void* stream);

/*
 * Host-buffer convenience entry points (synchronous).
 * hipe_joint_location_loss_host: JointLocationLoss forward (+ backward when dlogits_host or
 * dlogits_dev is non-NULL) on host logits, streamed to the current device in chunks of
 * `chunk_samples` samples (0 = auto) with copies overlapped against the kernels.
 *   logits_host : B * J * D*H*W elements (pinned memory recommended)
 *   gt/vis_host : B * 3J of tgt_dtype
 *   loss_host   : one double;  coords_host : B*3J floats (nullable)
 *   dlogits_host: gradient copied back to the host (nullable)
 *   dlogits_dev : caller-owned DEVICE buffer of B * J * D*H*W elements that receives the
 *                 gradient where the next consumer (the head's backward) wants it (nullable)
 */
HIPE_API int hipe_joint_location_loss_host(const void* logits_host, int dtype, long long B, int J, int D,
                                  int H, int W, const void* gt_host, const void* vis_host,
                                  int tgt_dtype, int size_average, double* loss_host,
                                  float* coords_host, void* dlogits_host, void* dlogits_dev,
                                  long long chunk_samples);

HIPE_API int hipe_procrustes_host(const double* X_host, const double* Y_host, long long N, int J, int mode,
                         int optimal_scale, double* Z_host, double* pa_mpjpe, double* mpjpe);

#ifdef __cplusplus
}
#endif
#endif /* HIPE_B200_H */

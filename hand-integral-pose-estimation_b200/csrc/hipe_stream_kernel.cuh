// The streaming kernel of the soft-argmax path (included by hipe_softargmax.cu only).
//
// CTA c of an S-CTA cluster owns rows [NR*c/S, NR*(c+1)/S) of one volume.  A producer warp
// streams them through a ring of kStages x 16 KB shared-memory stages with 1-D TMA bulk
// copies (or the consumers load straight from global memory in the LDG variant).  A stage
// holds SR = kPasses * R rows.  Two thread mappings read it:
//   reduce pass  NV = 1 or 2 adjacent 16-byte vectors per thread per visit (2 whenever a row
//                is a multiple of 32 bytes): the per-visit fp64 bookkeeping is paid once per
//                NV*VEC elements;
//   write pass   always one 16-byte vector per thread, so that every store instruction of a
//                warp covers one contiguous 512-byte run.
#pragma once
#include "hipe_common.cuh"

// Resident CTAs per SM the TMA kernels are compiled for (register cap 65536 / (288 * n)).
// Measured on B200 (scripts/ab_minctas.sh): the reduce-only launches like 3 CTAs of <= 72
// registers (bf16 forward 5.24 vs 5.16 TB/s); the launches with a write pass are 1-9 % faster
// with 2 CTAs and no spills (bf16 backward 6.67 vs 6.45 TB/s, fused 56^3 0.566 vs 0.618 ms).
#ifndef HIPE_MIN_CTAS_REDUCE
#define HIPE_MIN_CTAS_REDUCE 3
#endif
#ifndef HIPE_MIN_CTAS_WRITE
#define HIPE_MIN_CTAS_WRITE 2
#endif

namespace hipe {

// geometry of one CTA's share of its volume (uniform over the CTA)
struct Tile {
    const unsigned char* src;   // first byte of this CTA's rows
    long long cta_bytes;
    int r0, nrows;              // first row, number of rows
    int row_bytes;
    int stage_bytes;            // bytes per full stage (= kPasses * R * row_bytes <= 16 KB)
    int nst;                    // stages this CTA streams per pass
};

template <bool kSoftmax>
struct Acc {
    float m;    // running max (natural units)
    float mL;   // fl(m * log2e): the exponent offset every e of this thread is relative to
    double s0;  // sum of the per-visit row sums rs            (denominator of Ey, Ez; L)
    double sx;  // sum C_j * (x_j - xc)   (xc = centre of the thread's columns, C_j column sums)
    double sy;  // sum rs * y
    double sz;  // sum rs * z
    double s0c; // sum C_j                                     (denominator of Ex)
    __device__ __forceinline__ void init() {
        m = kSoftmax ? kNegInf : 0.f;
        mL = kSoftmax ? kNegInf : 0.f;
        s0 = sx = sy = sz = s0c = 0.0;
    }
    __device__ __forceinline__ void scale(float f) {
        const double d = (double)f;
        s0 *= d; sx *= d; sy *= d; sz *= d; s0c *= d;
    }
};

// Weight that re-bases a partial computed against maximum m_from onto maximum m_to >= m_from.
// Every exponential is 2^(s*log2e - fl(m*log2e)), so the weight is 2^(fl(m_from*log2e) -
// fl(m_to*log2e)) with the SAME rounded products (no contraction): a partial is re-based by
// exactly the offset its exponentials were computed against.
__device__ __forceinline__ float rebase_weight(float m_from, float m_to) {
    return (m_from == kNegInf) ? 0.f
                               : ex2(__fsub_rn(__fmul_rn(m_from, kLog2e), __fmul_rn(m_to, kLog2e)));
}

// Pairwise (tree) sum over the N exponentials of one visit.  Every expectation is a ratio of
// two fp64 sums built from the SAME rounded fp32 atoms, so an fp32 rounding only re-weights
// voxels (error ~ eps * spread of the distribution, zero for a one-hot volume):
//   Ey, Ez  atoms = per-visit row sums rs (a visit lies in one row):  Sy/S0, Sz/S0;
//   Ex      atoms = per-thread column sums C_j kept in fp32 over kFlush stages and folded into
//           fp64 with their exact column index:                      Sx/S0c.
// (An earlier version rounded sum (j-c) e_j per visit on its own; that error is eps * N/2 voxels
// whatever the spread, and a near-one-hot volume whose E sits within 0.005 voxel of the peak
// missed the 1e-5 gradient gate by 4x: scripts/diag_resident.py.)
// Blackwell packed fp32: one FFMA2 / FADD2 / FMUL2 issue slot does two lanes' worth of work.  The
// reduce pass is issue-bound for bf16 logits (ncu: 75 % issue, 70 % XU at 60 % of HBM), so the
// scale-and-offset, the column sums and the per-visit tree sum all run on float2 pairs.
template <int NP>
struct PairTree {   // sum of NP float2 values, pairwise
    __device__ __forceinline__ static float2 sum(const float2* a) {
        return __fadd2_rn(PairTree<NP / 2>::sum(a), PairTree<NP / 2>::sum(a + NP / 2));
    }
};
template <>
struct PairTree<1> {
    __device__ __forceinline__ static float2 sum(const float2* a) { return a[0]; }
};

// fp32 -> fp64 of a float without the conversion unit (F2F.F64.F32 shares the
// quarter-rate XU pipe with MUFU.EX2, the kernel's busiest pipe): re-bias the exponent with
// integer ops.  +-0 maps to +-2^-127 (5.9e-39) instead of 0, far below one ulp of any sum that
// contains the volume maximum's e = 1; denormals cannot occur (ex2.ftz flushes them, and sums
// / products of flushed values by weights >= 0.5 stay normal or exactly zero).
__device__ __forceinline__ double widen_nonneg(float f) {
    const uint32_t b = __float_as_uint(f);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
}
__device__ __forceinline__ double widen_signed(float f) {
    const uint32_t b = __float_as_uint(f);
    const uint32_t hi = (((b & 0x7fffffffu) >> 3) + 0x38000000u) | (b & 0x80000000u);
    return __hiloint2double((int)hi, (int)(b << 29));
}

// ---------------------------------------------------------------------------------------
// Reduce pass: online softmax + the four moment sums of this thread's share.
//   NV      16-byte vectors per visit (1 or 2)
//   R       rows covered by one visit of the CTA under this mapping
//   KP      visits per stage
// ---------------------------------------------------------------------------------------
template <typename T, bool kTma, bool kSoftmax, int NV>
__device__ __forceinline__ void reduce_pass(const Params& p, const Tile& t, unsigned char* smem_raw,
                                            SmemLayout& sm, int tid, int& it, Acc<kSoftmax>& a,
                                            double& xc_out) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    constexpr int EPV = NV * VEC;            // elements per visit
    constexpr int KP = kPasses / NV;         // visits per stage
    const int lane = tid & 31;
    const int G = p.G / NV;                  // column groups per row under this mapping
    const int R = p.R * NV;                  // rows per visit of the CTA
    const bool active = tid < G * R;
    const int cg = active ? tid % G : 0;
    const int rr = active ? tid / G : 0;
    const int nvis = (active && t.nrows > rr) ? (t.nrows - rr + R - 1) / R : 0;
    const int full = nvis / KP;              // stages whose KP visits are all inside the range
    const int pass_bytes = R * t.row_bytes;
    const int row_first = t.r0 + rr;
    // row coordinates of the current visit: small integers, exact in fp32 (FADD/FSETP instead of
    // fp64 adds and int->double conversions, which share the XU pipe with MUFU.EX2)
    float yf = (float)(row_first % p.H), zf = (float)(row_first / p.H);
    const float dyf = (float)(R % p.H), dzf = (float)(R / p.H), Hf = (float)p.H;
    const uint32_t lds_base = kTma ? smem_u32(smem_raw) + (uint32_t)(16 * NV) * tid : 0u;
    const unsigned char* gsrc = t.src + (long long)(16 * NV) * tid;
    xc_out = (double)(cg * EPV) + 0.5 * (double)(EPV - 1);
    a.init();
    // column sums of this thread, fp32 for kFlush stages at a time (softmax terms are in [0, 1]:
    // 8 stages x KP visits = 16-32 additions per column between folds)
    constexpr int kFlush = kSoftmax ? 8 : 1;
    constexpr int NP = EPV / 2;
    float2 col[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) col[j] = make_float2(0.f, 0.f);

    for (int i = 0; i < t.nst; ++i, ++it) {
        float v[KP][EPV];
        // ---- fetch the stage into registers ----
        uint32_t sbase = 0;
        if constexpr (kTma) {
            const int slot = it % kStages;
            mbar_wait(smem_u32(&sm.full[slot]), (it / kStages) & 1);
            sbase = lds_base + slot * kStageBytes;
        }
        if (i < full) {
#pragma unroll
            for (int q = 0; q < KP; ++q)
#pragma unroll
                for (int h = 0; h < NV; ++h) {
                    float w[VEC];
                    if constexpr (kTma)
                        V16::unpack(lds128(sbase + q * pass_bytes + 16 * h), w);
                    else
                        V16::unpack(ldg_stream(gsrc + (long long)(i * KP + q) * pass_bytes + 16 * h), w);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) v[q][h * VEC + j] = w[j];
                }
        } else {
#pragma unroll
            for (int q = 0; q < KP; ++q) {
                const bool ok = i * KP + q < nvis;
#pragma unroll
                for (int h = 0; h < NV; ++h) {
                    float w[VEC];
#pragma unroll
                    for (int j = 0; j < VEC; ++j) w[j] = kSoftmax ? kNegInf : 0.f;
                    if (ok) {
                        if constexpr (kTma)
                            V16::unpack(lds128(sbase + q * pass_bytes + 16 * h), w);
                        else
                            V16::unpack(ldg_stream(gsrc + (long long)(i * KP + q) * pass_bytes + 16 * h), w);
                    }
#pragma unroll
                    for (int j = 0; j < VEC; ++j) v[q][h * VEC + j] = w[j];
                }
            }
        }
        if constexpr (kTma) {
            // hand the slot back only once the loads of this stage have returned
            // (mbar_arrive_after_loads: the arrive consumes a word of every load instruction; a
            // warp's scoreboard covers all its lanes)
            uint32_t dep = 0;
#pragma unroll
            for (int q = 0; q < KP; ++q)
#pragma unroll
                for (int h = 0; h < NV; ++h) dep ^= __float_as_uint(v[q][h * VEC]);
            if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.empty[it % kStages]), dep);
        }
        // ---- online max ----
        if constexpr (kSoftmax) {
            float smax = kNegInf;
#pragma unroll
            for (int q = 0; q < KP; ++q)
#pragma unroll
                for (int j = 0; j < EPV; ++j) smax = fmaxf(smax, v[q][j]);
            if (smax > a.m) {
                const float wgt = rebase_weight(a.m, smax);
                a.scale(wgt);
#pragma unroll
                for (int j = 0; j < NP; ++j) col[j] = __fmul2_rn(col[j], make_float2(wgt, wgt));
                a.m = smax;
                a.mL = __fmul_rn(smax, kLog2e);
            }
        }
        // all -inf so far: exponentiate against 0 so that e = 2^-inf = 0 (no NaN, no branch)
        const float off = (kSoftmax && a.m == kNegInf) ? 0.f : a.mL;
        // ---- accumulate: fp32 inside one visit, fp64 across visits ----
        const float2 sc2 = make_float2(kLog2e, kLog2e), off2 = make_float2(-off, -off);
#pragma unroll
        for (int q = 0; q < KP; ++q) {
            float2 e[NP];
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                const float2 x = make_float2(v[q][2 * j], v[q][2 * j + 1]);
                if constexpr (kSoftmax) {
                    const float2 a2 = __ffma2_rn(x, sc2, off2);
                    e[j] = make_float2(ex2(a2.x), ex2(a2.y));
                } else {
                    e[j] = x;
                }
            }
            const float2 r2 = PairTree<NP>::sum(e);
            const float rs = r2.x + r2.y;
#pragma unroll
            for (int j = 0; j < NP; ++j) col[j] = __fadd2_rn(col[j], e[j]);
            const double drs = kSoftmax ? widen_nonneg(rs) : (double)rs;
            a.s0 += drs;
            a.sy = fma(widen_nonneg(yf), drs, a.sy);
            a.sz = fma(widen_nonneg(zf), drs, a.sz);
            // next visited row: R = dz*H + dy rows further, one carry at most
            yf += dyf;
            zf += dzf;
            if (yf >= Hf) {
                yf -= Hf;
                zf += 1.f;
            }
        }
        if ((i + 1) % kFlush == 0 || i + 1 == t.nst) {
#pragma unroll
            for (int j = 0; j < EPV; ++j) {
                const float cj = (j & 1) ? col[j / 2].y : col[j / 2].x;
                const double dc = kSoftmax ? widen_nonneg(cj) : (double)cj;
                a.s0c += dc;
                a.sx = fma((double)j - 0.5 * (double)(EPV - 1), dc, a.sx);
            }
#pragma unroll
            for (int j = 0; j < NP; ++j) col[j] = make_float2(0.f, 0.f);
        }
    }
}

// ---------------------------------------------------------------------------------------
// Merge the per-thread partials: thread -> warp (shuffles) -> CTA (shared memory) ->
// cluster (DSMEM).  On return thread 0 of every CTA of the cluster holds the totals
// (relative to the volume maximum *fm).  Leaves the second cluster-barrier phase ARRIVED
// but not waited (the caller waits before exiting, so that peers' `part` stays readable).
// ---------------------------------------------------------------------------------------
template <bool kSoftmax>
__device__ __forceinline__ void merge_partials(SmemLayout& sm, int tid, uint32_t S, const Acc<kSoftmax>& a,
                                               double xc, double (&tot)[kNSums], float& fm) {
    const int lane = tid & 31, warp = tid >> 5;
    // order: S0 (row-sum atoms), Sx, Sy, Sz, S0c (column-sum atoms)
    double q[kNSums] = {a.s0, fma(xc, a.s0c, a.sx), a.sy, a.sz, a.s0c};
    float m = a.m;
    if constexpr (kSoftmax) {
        const float wm = warp_max(m);
        const double f = (double)rebase_weight(m, wm);
#pragma unroll
        for (int k = 0; k < kNSums; ++k) q[k] *= f;
        m = wm;
    }
#pragma unroll
    for (int k = 0; k < kNSums; ++k) q[k] = warp_sum(q[k]);
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < kNSums; ++k) sm.wred[warp][k] = q[k];
        sm.wmax[warp] = m;
    }
    consumer_barrier();
    if (warp == 0) {
        constexpr int NW = kConsumers / 32;
        float pm = (lane < NW) ? sm.wmax[lane] : (kSoftmax ? kNegInf : 0.f);
#pragma unroll
        for (int k = 0; k < kNSums; ++k) q[k] = (lane < NW) ? sm.wred[lane][k] : 0.0;
        if constexpr (kSoftmax) {
            float wm = pm;
#pragma unroll
            for (int o = NW / 2; o > 0; o >>= 1) wm = fmaxf(wm, __shfl_xor_sync(0xffffffffu, wm, o));
            const double f = (double)rebase_weight(pm, wm);
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q[k] *= f;
            pm = wm;
        }
#pragma unroll
        for (int o = NW / 2; o > 0; o >>= 1)
#pragma unroll
            for (int k = 0; k < kNSums; ++k) q[k] += __shfl_xor_sync(0xffffffffu, q[k], o);
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < kNSums; ++k) sm.part[k] = q[k];
            sm.part_m = pm;
        }
    }
    if (S > 1) {
        cluster_sync_all();  // every CTA's `part` is written and visible cluster-wide
        if (tid < kNSums * (int)S)
            sm.all[tid / kNSums][tid % kNSums] = dsmem_ld_f64(smem_u32(&sm.part[tid % kNSums]), tid / kNSums);
        if (tid >= 64 && tid < 64 + (int)S) sm.all_m[tid - 64] = dsmem_ld_f32(smem_u32(&sm.part_m), tid - 64);
        consumer_barrier();
        // done reading peers: arrive now, the caller waits before exiting
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    } else {
        consumer_barrier();
        if (tid < kNSums) sm.all[0][tid] = sm.part[tid];
        if (tid == 32) sm.all_m[0] = sm.part_m;
        consumer_barrier();
    }
    fm = kSoftmax ? kNegInf : 0.f;
#pragma unroll
    for (int k = 0; k < kNSums; ++k) tot[k] = 0.0;
    if (tid == 0) {
        if constexpr (kSoftmax)
            for (uint32_t r = 0; r < S; ++r) fm = fmaxf(fm, sm.all_m[r]);
        for (uint32_t r = 0; r < S; ++r) {
            const double f = kSoftmax ? (double)rebase_weight(sm.all_m[r], fm) : 1.0;
#pragma unroll
            for (int k = 0; k < kNSums; ++k) tot[k] = fma(f, sm.all[r][k], tot[k]);
        }
    }
}

// ---------------------------------------------------------------------------------------
// Write pass: out_i = e_i * (ax(x_i) + rowterm(y_i, z_i)), one 16-byte vector per thread per
// visit.  kProb: out_i = e_i / L.
//   dlogits_i = e_i/L * ( gx (x-Ex)/W + gy (y-Ey)/H + gz (z-Ez)/D )
// (idx - E) must keep its ABSOLUTE accuracy: x once per thread in fp64, y and z per row
// against E split into hi + lo floats.
// ---------------------------------------------------------------------------------------
template <typename T, bool kTma, bool kProb>
__device__ __forceinline__ void write_pass(const Params& p, const Tile& t, unsigned char* smem_raw,
                                           SmemLayout& sm, int tid, int& it, long long vol, float M,
                                           float invL, double Ex, double Ey, double Ez, float gx, float gy,
                                           float gz) {
    using V16 = Vec16<T>;
    constexpr int VEC = V16::N;
    const int lane = tid & 31;
    const bool resident = kTma && p.resident;
    const bool active = tid < p.G * p.R;
    const int cg = active ? tid % p.G : 0;
    const int rr = active ? tid / p.G : 0;
    const int nvis = (active && t.nrows > rr) ? (t.nrows - rr + p.R - 1) / p.R : 0;
    const int full = nvis / kPasses;
    const int pass_bytes = p.R * t.row_bytes;
    const int row_first = t.r0 + rr;
    const int x0i = cg * VEC;
    const uint32_t lds_base = kTma ? smem_u32(smem_raw) + 16u * tid : 0u;
    const unsigned char* gsrc = t.src + 16ll * tid;
    unsigned char* gdst = p.out + vol * p.vol_bytes + (long long)t.r0 * t.row_bytes + 16ll * tid;

    const float ML = __fmul_rn(M, kLog2e);
    float ax[VEC];
    float ky = 0.f, kz = 0.f;
    const float Eyh = (float)Ey, Eyl = (float)(Ey - (double)Eyh);
    const float Ezh = (float)Ez, Ezl = (float)(Ez - (double)Ezh);
    if constexpr (kProb) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) ax[j] = invL;
    } else {
        const float kx = gx / (float)p.W * invL;
        ky = gy / (float)p.H * invL;
        kz = gz / (float)p.D * invL;
#pragma unroll
        for (int j = 0; j < VEC; ++j) ax[j] = kx * (float)((double)(x0i + j) - Ex);
    }
    const uint64_t st_pol = policy_evict_first();
    const float dyf = (float)(p.R % p.H), dzf = (float)(p.R / p.H), Hf = (float)p.H;
    // Stage order.  The second pass of a fused launch walks the CTA's share BACKWARDS: the bytes
    // read last are the ones most likely still in L2 (the re-read distance of stage i is
    // ~2*(nst-i) stages x the resident CTAs instead of a constant 2*nst), measured on bf16 64^3 as
    // the difference between re-reading everything from HBM and hitting L2 for the tail.
    const bool rev = p.rev && !resident;
    const int srows = kPasses * p.R;                       // rows a stage advances
    const float sdy = (float)(srows % p.H), sdz = (float)(srows / p.H);
    float ys, zs;                                          // first row of the current stage (this thread)
    {
        const int row_s = row_first + (rev ? (t.nst - 1) * srows : 0);
        ys = (float)(row_s % p.H);
        zs = (float)(row_s / p.H);
    }

    for (int ii = 0; ii < t.nst; ++ii, ++it) {
        const int i = rev ? t.nst - 1 - ii : ii;
        float yf = ys, zf = zs;
        if (rev) {
            ys -= sdy; zs -= sdz;
            if (ys < 0.f) { ys += Hf; zs -= 1.f; }
        } else {
            ys += sdy; zs += sdz;
            if (ys >= Hf) { ys -= Hf; zs += 1.f; }
        }
        float v[kPasses][VEC];
        uint32_t sbase = 0;
        if constexpr (kTma) {
            if (resident) {
                sbase = lds_base + i * kStageBytes;   // stage i still sits in slot i: no reload, no wait
            } else {
                const int slot = it % kStages;
                mbar_wait(smem_u32(&sm.full[slot]), (it / kStages) & 1);
                sbase = lds_base + slot * kStageBytes;
            }
        }
        const bool whole = i < full;
#pragma unroll
        for (int q = 0; q < kPasses; ++q) {
            if (whole || i * kPasses + q < nvis) {
                if constexpr (kTma)
                    V16::unpack(lds128(sbase + q * pass_bytes), v[q]);
                else
                    V16::unpack(ldg_stream(gsrc + (long long)(i * kPasses + q) * pass_bytes), v[q]);
            }
        }
        if constexpr (kTma) {
            if (!resident) {
                uint32_t dep = 0;
#pragma unroll
                for (int q = 0; q < kPasses; ++q)
                    if (whole || i * kPasses + q < nvis) dep ^= __float_as_uint(v[q][0]);
                if (lane == 0) mbar_arrive_after_loads(smem_u32(&sm.empty[it % kStages]), dep);
            }
        }
#pragma unroll
        for (int q = 0; q < kPasses; ++q) {
            if (whole || i * kPasses + q < nvis) {
                float rowterm = 0.f;
                if constexpr (!kProb) rowterm = fmaf(ky, (yf - Eyh) - Eyl, kz * ((zf - Ezh) - Ezl));
                float o[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    const float e = ex2(fmaf(v[q][j], kLog2e, -ML));
                    o[j] = e * (ax[j] + rowterm);
                }
                if (p.st_hint)
                    stg_stream_hint(gdst + (long long)(i * kPasses + q) * pass_bytes, V16::pack(o), st_pol);
                else
                    stg_stream(gdst + (long long)(i * kPasses + q) * pass_bytes, V16::pack(o));
            }
            yf += dyf;
            zf += dzf;
            if (yf >= Hf) {
                yf -= Hf;
                zf += 1.f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// The kernel.  kOp selects reduce-only (forward), write-only (backward / probabilities) or
// reduce + L1 gradient + write in one launch (fused).
// ---------------------------------------------------------------------------------------
template <typename T, bool kTma, int kOp>
__global__ void __launch_bounds__(kTma ? kConsumers + 32 : kConsumers, !kTma ? 1 : ((kOp == OP_REDUCE_SOFTMAX || kOp == OP_REDUCE_PLAIN) ? HIPE_MIN_CTAS_REDUCE : HIPE_MIN_CTAS_WRITE))
    k_stream(const Params p) {
    constexpr bool kHasReduce = (kOp == OP_REDUCE_SOFTMAX || kOp == OP_REDUCE_PLAIN || kOp == OP_FUSED);
    constexpr bool kHasWrite = (kOp == OP_WRITE_GRAD || kOp == OP_WRITE_PROB || kOp == OP_FUSED);
    constexpr bool kSoftmax = (kOp != OP_REDUCE_PLAIN);
    constexpr int kNumPassesOverVolume = (kOp == OP_FUSED) ? 2 : 1;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    SmemLayout& sm = *reinterpret_cast<SmemLayout*>(smem_raw + (kTma ? kStages * kStageBytes : 0));

    const int tid = threadIdx.x;
    const uint32_t S = cluster_nctarank();
    const uint32_t crank = cluster_ctarank();
    long long vol = blockIdx.x / S;
    if (p.reverse) vol = p.n_vol - 1 - vol;

    Tile t;
    t.r0 = (int)(((long long)p.NR * crank) / S);
    t.nrows = (int)(((long long)p.NR * (crank + 1)) / S) - t.r0;
    t.row_bytes = p.G * 16;
    t.cta_bytes = (long long)t.nrows * t.row_bytes;
    t.stage_bytes = kPasses * p.R * t.row_bytes;
    t.nst = (int)((t.cta_bytes + t.stage_bytes - 1) / t.stage_bytes);
    t.src = p.in + vol * p.vol_bytes + (long long)t.r0 * t.row_bytes;

    if constexpr (kOp == OP_FUSED) {
        // targets of this volume: their global-memory latency is paid under the reduce pass, not
        // in the serial section between the passes (0.6-0.7 % on fp32 56^3 / 64^3, same-box A/B)
        if (tid < 6) {
            const void* base = tid < 3 ? p.gt : p.vis;
            const long long j = vol * 3 + (tid < 3 ? tid : tid - 3);
            sm.tgt[tid] = p.tgt_f64 ? reinterpret_cast<const double*>(base)[j] : (double)reinterpret_cast<const float*>(base)[j];
        }
    }
    if constexpr (kTma) {
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < kStages; ++s) {
                mbar_init(smem_u32(&sm.full[s]), 1);
                mbar_init(smem_u32(&sm.empty[s]), kConsumers / 32);
            }
            mbar_fence_init();
            fence_proxy_async();
        }
        __syncthreads();

        if (tid >= kConsumers) {
            // ---------------- producer warp ----------------
            // It has nothing to contribute to the cluster reduction: arrive right away so the
            // consumers' cluster barrier can complete while this warp is still streaming.
            if (S > 1 && kHasReduce) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
            if (tid == kConsumers) {
                const uint64_t pn = policy_evict_normal(), pf = policy_evict_first(), pl = policy_evict_last();
                const uint64_t pol_a = p.pol1 == 0 ? pn : (p.pol1 == 1 ? pf : pl);
                const uint64_t pol_b = p.pol2 == 0 ? pn : (p.pol2 == 1 ? pf : pl);
                const int total = t.nst * (p.resident ? 1 : kNumPassesOverVolume);
                for (int it = 0; it < total; ++it) {
                    const int slot = it % kStages;
                    // second pass of a fused launch: LAST stage first (see write_pass)
                    const int i = (it >= t.nst) ? (p.rev ? 2 * t.nst - 1 - it : it - t.nst) : it;
                    if (it >= kStages) mbar_wait(smem_u32(&sm.empty[slot]), ((it / kStages) - 1) & 1);
                    const long long off = (long long)i * t.stage_bytes;
                    const long long left = t.cta_bytes - off;
                    const uint32_t bytes = (uint32_t)(left < t.stage_bytes ? left : t.stage_bytes);
                    const uint32_t bar = smem_u32(&sm.full[slot]);
                    mbar_arrive_expect_tx(bar, bytes);
                    tma_load_1d(smem_u32(smem_raw + slot * kStageBytes), t.src + off, bytes, bar,
                                (it < t.nst) ? pol_a : pol_b);
                }
            }
            __syncwarp();
            if (S > 1 && kHasReduce) {
                asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
                // second cluster phase: peers must not exit while their `part` is being read
                asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
                asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
            }
            return;
        }
    }

    // ---------------- consumers ----------------
    int it = 0;  // stage counter across both passes (ring position)
    float M = 0.f, invL = 1.f, gx = 0.f, gy = 0.f, gz = 0.f;
    double Ex = 0.0, Ey = 0.0, Ez = 0.0;

    if constexpr (kHasReduce) {
        Acc<kSoftmax> a;
        double xc = 0.0;
        if (p.nv2)
            reduce_pass<T, kTma, kSoftmax, 2>(p, t, smem_raw, sm, tid, it, a, xc);
        else
            reduce_pass<T, kTma, kSoftmax, 1>(p, t, smem_raw, sm, tid, it, a, xc);
        double tot[kNSums];
        float fm;
        merge_partials<kSoftmax>(sm, tid, S, a, xc, tot, fm);
        if (tid == 0) {
            // loss.py:39-41 (expectations) and :54-56 (normalisation), in fp64, rounded once
            const double ex_ = kSoftmax ? tot[1] / tot[4] : tot[1];
            const double ey_ = kSoftmax ? tot[2] / tot[0] : tot[2];
            const double ez_ = kSoftmax ? tot[3] / tot[0] : tot[3];
            const float cxn = (float)(ex_ / (double)p.W - 0.5);
            const float cyn = (float)(ey_ / (double)p.H - 0.5);
            const float czn = (float)(ez_ / (double)p.D - 0.5);
            float g[3] = {0.f, 0.f, 0.f};
            if constexpr (kOp == OP_FUSED) {
                // loss.py:86-89 on this volume's three coordinates, in the targets' precision
                const double c[3] = {(double)cxn, (double)cyn, (double)czn};
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const double tk = sm.tgt[k], wk = sm.tgt[3 + k];
                    const double dlt = c[k] - tk;
                    const double sg = (dlt > 0.0) ? 1.0 : ((dlt < 0.0) ? -1.0 : 0.0);
                    g[k] = (float)(sg * wk * p.norm);
                    if (crank == 0 && p.loss_terms) p.loss_terms[vol * 3 + k] = fabs(dlt) * wk;
                }
            }
            if (crank == 0) {
                if (p.coords) { p.coords[vol * 3 + 0] = cxn; p.coords[vol * 3 + 1] = cyn; p.coords[vol * 3 + 2] = czn; }
                if (p.expect) { p.expect[vol * 3 + 0] = ex_; p.expect[vol * 3 + 1] = ey_; p.expect[vol * 3 + 2] = ez_; }
                if (p.stats) { p.stats[vol * 2 + 0] = fm; p.stats[vol * 2 + 1] = (float)tot[0]; }
            }
            sm.finE[0] = ex_; sm.finE[1] = ey_; sm.finE[2] = ez_;
            sm.fin[0] = fm; sm.fin[1] = (float)tot[0]; sm.fin[2] = g[0]; sm.fin[3] = g[1]; sm.fin[4] = g[2];
        }
        if constexpr (kOp == OP_FUSED) {
            consumer_barrier();
            M = sm.fin[0]; invL = 1.f / sm.fin[1];
            Ex = sm.finE[0]; Ey = sm.finE[1]; Ez = sm.finE[2];
            gx = sm.fin[2]; gy = sm.fin[3]; gz = sm.fin[4];
        }
    } else {
        // write-only ops read their per-volume scalars from the forward's outputs
        M = p.stats_in[vol * 2 + 0];
        invL = 1.f / p.stats_in[vol * 2 + 1];
        if constexpr (kOp == OP_WRITE_GRAD) {
            const float gs = p.gscale ? *p.gscale : 1.f;
            Ex = p.expect_in[vol * 3 + 0]; Ey = p.expect_in[vol * 3 + 1]; Ez = p.expect_in[vol * 3 + 2];
            gx = p.gcoords[vol * 3 + 0] * gs; gy = p.gcoords[vol * 3 + 1] * gs; gz = p.gcoords[vol * 3 + 2] * gs;
        }
    }

    if constexpr (kHasWrite)
        write_pass<T, kTma, kOp == OP_WRITE_PROB>(p, t, smem_raw, sm, tid, it, vol, M, invL, Ex, Ey, Ez, gx, gy, gz);

    if constexpr (kHasReduce) {
        if (S > 1) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
}

}  // namespace hipe

// The one exchange step of the batch-sharded path, as ONE kernel over NVLink peer
// memory instead of an NCCL all-gather between small pack / combine launches.
//
// Replaces, per training step, what the reference does with Reduce.apply(*outputs) / len(outputs)
// (common/nets/balanced_parallel.py:46-55,127) plus the gather of predicted joints: every rank
// holds a record [ loss_sum_r , B_r , coords_r (B x 3J) ] and needs all of them.  Each rank's
// kernel (one CTA per peer) stores its record straight into slot [parity][rank] of EVERY peer's
// symmetric buffer (16-byte P2P stores through NVSwitch), publishes a monotonically increasing
// sequence number into the peer's flag word with a system-scope release, acquires that peer's flag
// in its own buffer, copies the record out, and the last CTA combines the world loss sums in fixed
// rank order -- bit-identical on every rank.  (A first version did all peers from one CTA: 29 us
// at 2 GPUs but 108 / 167 us at 4 / 8, i.e. serial in the world size.)  ~32 KB per rank per step: latency-sized, so the point is launch count and the absence
// of stream hand-offs, not bandwidth.
//
// Buffer layout per rank (doubles unless noted), both parities:
//   slot[2][world][reclen]   reclen = 2 + B*K rounded up to even
//   flag[2][world]           (unsigned long long) last sequence number published by that rank
// Parity = seq & 1.  A peer can only overwrite parity p again at step seq+2, which needs this
// rank's flag of step seq+1, which this rank publishes after its step-seq kernel has finished
// reading (stream order): no reuse race.  Flags are never reset.
#include "hipe_common.cuh"

namespace hipe {

constexpr int kXT = 512;       // threads per CTA of the one-kernel form
// The split form's two launches run beside the next step's persistent loss launch (608 threads x 96
// registers per SM): with 128 threads x 32 registers their CTAs fit on an SM that already holds one
// of its CTAs, so a collect that is still polling for a slow rank never keeps a cluster from launching.
constexpr int kXTSplit = 128;
constexpr int kXMaxWorld = 64;

__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// grid = world CTAs: CTA p serves peer p (store + publish), waits for rank p's record, copies it out;
// the last CTA to finish (atomic ticket) combines the `world` loss sums in rank order.  No CTA
// waits on another CTA of this grid, so co-residency is not required for progress.
// kPhase 0: the whole exchange in one launch (publish, wait, combine).  kPhase 1: PUBLISH only (sum,
// stores, flag; never waits).  kPhase 2: COLLECT only (wait for the peers' flags, copy out, combine):
// launched on a side stream one step later, it takes the rendezvous with the slowest rank off the
// stream that runs the loss kernels -- nothing on the GPU needs the global loss value, the gradient
// is final before the exchange.  Flow control of the split form (hipe_b200/parallel.py): a rank's
// publish(t) is stream-ordered after its collect(t-1); so a peer that publishes t+1 into parity t-1
// has completed collect(t), which needed this rank's publish(t), which came after this rank's
// collect(t-1): the slot being overwritten has been read.
template <int kPhase>
__global__ void __launch_bounds__(kXT)
    k_peer_exchange(const double* __restrict__ loss_in, long long n_in, double in_scale, const float* __restrict__ coords,
                    long long B, int K, double* const* __restrict__ peers, int world, int rank,
                    unsigned long long seq, long long reclen, double value_scale, double* __restrict__ value_out,
                    double* __restrict__ sums_out, double* __restrict__ coords_all, unsigned* __restrict__ err,
                    double* __restrict__ scratch /* world sums */, unsigned* __restrict__ ticket,
                    unsigned long long timeout_ns) {
    __shared__ int timed_out;
    __shared__ double red[32];
    __shared__ double my_sum_sh;
    __shared__ unsigned my_ticket;
    const int p = blockIdx.x;                 // the peer this CTA serves / the rank it listens to
    const int nthr = (int)blockDim.x;         // kXT or kXTSplit
    const int par = (int)(seq & 1ull);
    const long long n = B * K;
    // 0. this rank's loss sum, in a fixed order (thread-strided partials -> warp shuffles ->
    //    kXT/32 warp partials); every CTA forms the same bits
    if (threadIdx.x == 0) timed_out = 0;
    if constexpr (kPhase != 2) {
        // kXT virtual lanes whatever the CTA's size (kXT or kXTSplit threads): thread t accumulates lanes
        // t, t + nthr, ...; each group of 32 consecutive lanes is one shuffle tree, as in the kXT-thread
        // launch -- the one-kernel form and the split form give the same bits
        const int per = kXT / nthr;
        double acc[kXT / kXTSplit];
#pragma unroll
        for (int j = 0; j < kXT / kXTSplit; ++j) {
            acc[j] = 0.0;
            if (j < per)
                for (long long i = threadIdx.x + (long long)j * nthr; i < n_in; i += kXT) acc[j] += loss_in[i];
        }
#pragma unroll
        for (int j = 0; j < kXT / kXTSplit; ++j) {
            if (j < per) {
                const double w = warp_sum(acc[j]);
                if ((threadIdx.x & 31) == 0) red[(threadIdx.x + j * nthr) >> 5] = w;
            }
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            double r = warp_sum((threadIdx.x < kXT / 32) ? red[threadIdx.x] : 0.0);
            if (threadIdx.x == 0) my_sum_sh = r * in_scale;
        }
        __syncthreads();
    }
    // 1. my record into slot [par][rank] of peer p (16-byte stores over NVLink; own buffer when p == rank)
    if constexpr (kPhase != 2) {
        double* dst = peers[p] + ((long long)par * world + rank) * reclen;
        if (threadIdx.x == 0) *reinterpret_cast<double2*>(dst) = make_double2(my_sum_sh, (double)B);
        const long long npair = n >> 1;
        for (long long i = threadIdx.x; i < npair; i += nthr)
            *reinterpret_cast<double2*>(dst + 2 + 2 * i) = make_double2((double)coords[2 * i], (double)coords[2 * i + 1]);
        if ((n & 1) && threadIdx.x == 0) dst[2 + n - 1] = (double)coords[n - 1];
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish to peer p, 3. wait for rank p's record in my own buffer (bounded by wall time)
    if (threadIdx.x == 0) {
        if constexpr (kPhase != 2) {
            unsigned long long* fp = reinterpret_cast<unsigned long long*>(peers[p] + 2ll * world * reclen) +
                                     (long long)par * world + rank;
            st_release_sys_u64(fp, seq);
        }
    }
    if constexpr (kPhase == 1) return;
    if (threadIdx.x == 0) {
        const unsigned long long* fm = reinterpret_cast<const unsigned long long*>(peers[rank] + 2ll * world * reclen) +
                                       (long long)par * world + p;
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys_u64(fm) < seq) {
            if (globaltimer_ns() - t0 > timeout_ns) {
                timed_out = 1;
                break;
            }
            __nanosleep(32);
        }
    }
    __syncthreads();
    // 4. rank p's record -> ordinary tensors
    const double* slot = peers[rank] + ((long long)par * world + p) * reclen;
    if (!timed_out) {
        if (coords_all) {
            double* out = coords_all + (long long)p * n;
            for (long long i = threadIdx.x; i < n; i += nthr) out[i] = ld_relaxed_sys_f64(slot + 2 + i);
        }
        if (threadIdx.x == 0) scratch[p] = ld_relaxed_sys_f64(slot);
    } else if (threadIdx.x == 0) {
        // kPhase 2: err is an array of `world` words (may be host-mapped): word p names the missing rank
        if (err) {
            if constexpr (kPhase == 2) *reinterpret_cast<volatile unsigned*>(err + p) = 1u;
            else atomicExch(err, 1u);
        }
        scratch[p] = __longlong_as_double(0x7ff8000000000000ll);
    }
    // 5. last CTA out combines in fixed rank order
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) my_ticket = atomicAdd(ticket, 1u);
    __syncthreads();
    if (my_ticket == (unsigned)world - 1u && threadIdx.x == 0) {
        __threadfence();
        double tot = 0.0;
        for (int r = 0; r < world; ++r) {
            const double v = *reinterpret_cast<volatile double*>(scratch + r);
            if (sums_out) sums_out[r] = v;
            tot += v;
        }
        value_out[0] = tot * value_scale;
        *ticket = 0u;                          // ready for the next launch (stream order)
    }
}

}  // namespace hipe

using namespace hipe;

extern "C" long long hipe_peer_exchange_reclen(long long B, int K) {
    if (B < 0 || K <= 0) return 0;
    const long long r = 2 + B * K;
    return (r + 1) & ~1ll;
}

extern "C" long long hipe_peer_exchange_bytes(int world, long long B, int K) {
    if (world <= 0 || world > kXMaxWorld) return 0;
    const long long reclen = hipe_peer_exchange_reclen(B, K);
    if (reclen == 0) return 0;
    return 8ll * (2ll * world * reclen + 2ll * world);
}

extern "C" int hipe_peer_exchange(const double* loss_in, long long n_in, double in_scale, const float* coords,
                                  long long B, int K,
                                  void* const* peers_dev, int world, int rank, unsigned long long seq,
                                  double value_scale, double* value_out, double* sums_out, double* coords_all,
                                  unsigned* err, double* scratch, double timeout_s, void* stream) {
    if (!loss_in || !coords || !peers_dev || !value_out || !scratch) return HIPE_E_NULL;
    if (world <= 0 || world > kXMaxWorld || rank < 0 || rank >= world || B <= 0 || K <= 0 || seq == 0 || n_in <= 0)
        return HIPE_E_SHAPE;
    const long long reclen = hipe_peer_exchange_reclen(B, K);
    const unsigned long long tns = (unsigned long long)((timeout_s > 0 ? timeout_s : 20.0) * 1e9);
    // scratch: `world` doubles, then one 32-bit ticket (zero before the first launch)
    k_peer_exchange<0><<<world, kXT, 0, static_cast<cudaStream_t>(stream)>>>(
        loss_in, n_in, in_scale, coords, B, K, reinterpret_cast<double* const*>(peers_dev), world, rank, seq, reclen,
        value_scale, value_out, sums_out, coords_all, err, scratch, reinterpret_cast<unsigned*>(scratch + world), tns);
    return (int)cudaGetLastError();
}

extern "C" int hipe_peer_publish(const double* loss_in, long long n_in, double in_scale, const float* coords,
                                 long long B, int K, void* const* peers_dev, int world, int rank,
                                 unsigned long long seq, void* stream) {
    if (!loss_in || !coords || !peers_dev) return HIPE_E_NULL;
    if (world <= 0 || world > kXMaxWorld || rank < 0 || rank >= world || B <= 0 || K <= 0 || seq == 0 || n_in <= 0)
        return HIPE_E_SHAPE;
    const long long reclen = hipe_peer_exchange_reclen(B, K);
    k_peer_exchange<1><<<world, kXTSplit, 0, static_cast<cudaStream_t>(stream)>>>(
        loss_in, n_in, in_scale, coords, B, K, reinterpret_cast<double* const*>(peers_dev), world, rank, seq, reclen,
        0.0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0ull);
    return (int)cudaGetLastError();
}

extern "C" int hipe_peer_collect(void* const* peers_dev, int world, int rank, unsigned long long seq, long long B,
                                 int K, double value_scale, double* value_out, double* sums_out,
                                 double* coords_all, unsigned* err, double* scratch, double timeout_s,
                                 void* stream) {
    if (!peers_dev || !value_out || !scratch) return HIPE_E_NULL;
    if (world <= 0 || world > kXMaxWorld || rank < 0 || rank >= world || B <= 0 || K <= 0 || seq == 0)
        return HIPE_E_SHAPE;
    const long long reclen = hipe_peer_exchange_reclen(B, K);
    const unsigned long long tns = (unsigned long long)((timeout_s > 0 ? timeout_s : 20.0) * 1e9);
    k_peer_exchange<2><<<world, kXTSplit, 0, static_cast<cudaStream_t>(stream)>>>(
        nullptr, 0, 0.0, nullptr, B, K, reinterpret_cast<double* const*>(peers_dev), world, rank, seq, reclen,
        value_scale, value_out, sums_out, coords_all, err, scratch, reinterpret_cast<unsigned*>(scratch + world), tns);
    return (int)cudaGetLastError();
}

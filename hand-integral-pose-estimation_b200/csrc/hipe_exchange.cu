// The one exchange step of the batch-sharded path, as ONE kernel over NVLink peer
// memory instead of an NCCL all-gather between small pack / combine launches.
//
// Replaces, per training step, what the reference does with Reduce.apply(*outputs) / len(outputs)
// (common/nets/balanced_parallel.py:46-55,127) plus the gather of predicted joints: every rank
// holds a record [ loss_sum_r , B_r , coords_r (B x 3J) ] and needs all of them.  Each rank's
// kernel (1 CTA) stores its record straight into slot [parity][rank] of EVERY peer's symmetric
// buffer (P2P stores through NVSwitch), publishes a monotonically increasing sequence number
// into the peer's flag word with a system-scope release, then acquires the flags of all peers in
// its own buffer and combines the world loss sums in fixed rank order -- bit-identical on every
// rank.  ~32 KB per rank per step: latency-sized, so the point is launch count and the absence
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

constexpr int kXT = 1024;
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

__global__ void __launch_bounds__(kXT)
    k_peer_exchange(const double* __restrict__ loss_in, long long n_in, double in_scale, const float* __restrict__ coords, long long B,
                    int K, double* const* __restrict__ peers, int world, int rank, unsigned long long seq,
                    long long reclen, double value_scale, double* __restrict__ value_out,
                    double* __restrict__ sums_out, double* __restrict__ coords_all, unsigned* __restrict__ err,
                    unsigned long long timeout_ns) {
    __shared__ int timed_out;
    __shared__ double red[32];
    __shared__ double my_sum_sh;
    const int par = (int)(seq & 1ull);
    const long long n = B * K;
    // 0. this rank's loss sum, in the fixed order of k_sum_f64 (thread-strided partials -> warp
    //    shuffles -> 32 warp partials)
    {
        double acc = 0.0;
        for (long long i = threadIdx.x; i < n_in; i += kXT) acc += loss_in[i];
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        if (threadIdx.x == 0) timed_out = 0;
        __syncthreads();
        if (threadIdx.x < 32) {
            double r = warp_sum(red[threadIdx.x]);
            if (threadIdx.x == 0) my_sum_sh = r * in_scale;
        }
        __syncthreads();
    }
    const double my_sum = my_sum_sh;
    // 1. my record into slot [par][rank] of every peer (own buffer included)
    for (int p = 0; p < world; ++p) {
        double* dst = peers[p] + ((long long)par * world + rank) * reclen;
        if (threadIdx.x == 0) {
            dst[0] = my_sum;
            dst[1] = (double)B;
        }
        for (long long i = threadIdx.x; i < n; i += kXT) dst[2 + i] = (double)coords[i];
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish: flag[par][rank] of every peer <- seq
    if (threadIdx.x < world) {
        unsigned long long* f = reinterpret_cast<unsigned long long*>(peers[threadIdx.x] + 2ll * world * reclen) +
                                (long long)par * world + rank;
        st_release_sys_u64(f, seq);
    }
    // 3. wait for every peer's record in my own buffer (bounded by wall time, never hangs the GPU)
    if (threadIdx.x < world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(peers[rank] + 2ll * world * reclen) +
                                      (long long)par * world + threadIdx.x;
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys_u64(f) < seq) {
            if (globaltimer_ns() - t0 > timeout_ns) {
                timed_out = 1;
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
    if (timed_out) {
        if (threadIdx.x == 0) {
            if (err) atomicExch(err, 1u);
            value_out[0] = __longlong_as_double(0x7ff8000000000000ll);
        }
        return;
    }
    // 4. combine in fixed rank order; hand the gathered joints to an ordinary tensor
    const double* mine = peers[rank] + (long long)par * world * reclen;
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int r = 0; r < world; ++r) {
            const double s = ld_relaxed_sys_f64(mine + r * reclen);
            if (sums_out) sums_out[r] = s;
            tot += s;
        }
        value_out[0] = tot * value_scale;
    }
    if (coords_all)
        for (long long i = threadIdx.x; i < (long long)world * n; i += kXT) {
            const long long r = i / n, j = i - r * n;
            coords_all[i] = ld_relaxed_sys_f64(mine + r * reclen + 2 + j);
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
                                  unsigned* err, double timeout_s, void* stream) {
    if (!loss_in || !coords || !peers_dev || !value_out) return HIPE_E_NULL;
    if (world <= 0 || world > kXMaxWorld || rank < 0 || rank >= world || B <= 0 || K <= 0 || seq == 0 || n_in <= 0)
        return HIPE_E_SHAPE;
    const long long reclen = hipe_peer_exchange_reclen(B, K);
    const unsigned long long tns = (unsigned long long)((timeout_s > 0 ? timeout_s : 20.0) * 1e9);
    k_peer_exchange<<<1, kXT, 0, static_cast<cudaStream_t>(stream)>>>(
        loss_in, n_in, in_scale, coords, B, K, reinterpret_cast<double* const*>(peers_dev), world, rank, seq, reclen,
        value_scale, value_out, sums_out, coords_all, err, tns);
    return (int)cudaGetLastError();
}

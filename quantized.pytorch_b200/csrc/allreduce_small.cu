// allreduce_small.cu - the cross-rank statistic exchange of batch-sharded training as ONE
// one-shot kernel over NVLink peer memory instead of NCCL calls.
//
// The messages are tiny (16 B ... 12 KB: QuantMeasure double[8], RangeBN double[3C] / double[2C], gradient range
// float[2]) and sit on the critical path of every layer, ~160 of them per ResNet-18 step, so they are latency bound: a
// NCCL all-reduce costs ~20 us each inside a CUDA graph (plus the small torch kernels that pack MIN/MAX into a second
// message), i.e. ~3.5 ms of a 20 ms step.  Here every rank owns a peer-mapped ("symmetric") buffer
//     [ control: seq, error | flags[2][W] | slots[2][W][SLOT_BYTES] ]
// and one CTA per call (a) stores its block into slot[parity][my rank] of EVERY peer over NVLink, (b) releases a flag
// there, (c) waits for the W flags in its own buffer, (d) reduces the W slots in rank order - so all ranks obtain
// bit-identical results - with a per-element operation (SUM / MIN / MAX layouts of the three statistic blocks in ONE
// message).  The sequence number lives in device memory, so the kernel can be captured in a CUDA graph and replayed.
// Slots are double buffered by call parity: a rank can run at most one call ahead of its slowest peer, because call
// s+1 cannot complete without that peer's flag for s+1, which the peer only raises after finishing call s.
// A bounded spin (about two seconds) records an error instead of hanging the GPU if a peer never arrives.
#include <stdlib.h>

#include "common.cuh"

constexpr int AR_MAX_WORLD = 16;
constexpr int AR_CTRL_BYTES = 4096;
constexpr int AR_THREADS = 512;

struct ArCtrl {
    unsigned long long seq;          // calls completed by this rank
    unsigned int error;              // != 0: a wait timed out (results of that call are invalid)
    unsigned int pad[13];
    unsigned long long flags[2][AR_MAX_WORLD];
};

// A slot holds a message in the flagged ("LL") form below: every 4 payload bytes travel with a 4-byte tag, so it is twice
// the message size.
extern "C" size_t qb_allreduce_small_buffer_bytes(int world, size_t slot_bytes) {
    if (world < 1 || world > AR_MAX_WORLD) return 0;
    return (size_t)AR_CTRL_BYTES + 2 * (size_t)world * (2 * slot_bytes);
}

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

constexpr unsigned long long AR_TIMEOUT_NS = 120ull * 1000ull * 1000ull * 1000ull;

struct ArPeers {
    unsigned char* buf[AR_MAX_WORLD];
};

// mode 0: SUM everywhere; mode 1: QuantMeasure block (element 4 MIN, element 5 MAX, others SUM); mode 2: {MIN, MAX} pair
template <typename T>
__device__ __forceinline__ T ar_combine(T a, T b, int i, int mode) {
    if (mode == 1) return i == 4 ? (b < a ? b : a) : (i == 5 ? (b > a ? b : a) : a + b);
    if (mode == 2) return i == 0 ? (b < a ? b : a) : (b > a ? b : a);
    return a + b;
}

template <typename T>
__global__ void __launch_bounds__(AR_THREADS) k_allreduce_oneshot(T* __restrict__ data, int n, int mode, ArPeers peers, int rank, int world,
                                                                  size_t slot_bytes) {
    __shared__ unsigned long long s_seq;
    __shared__ int s_fail;
    ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(peers.buf[rank]);
    if (threadIdx.x == 0) { s_seq = ctrl->seq + 1ull; s_fail = 0; }
    __syncthreads();
    const unsigned long long seq = s_seq;
    const int par = (int)(seq & 1ull);
    slot_bytes *= 2;                                               // slot stride of the buffer layout (flagged form: 2x)
    const size_t slot_off = (size_t)AR_CTRL_BYTES + ((size_t)par * world + rank) * slot_bytes;
    // (a) my block into my slot on every rank (own rank included: plain local stores)
    for (int i = threadIdx.x; i < n; i += AR_THREADS) {
        const T v = data[i];
        for (int p = 0; p < world; ++p) reinterpret_cast<T*>(peers.buf[p] + slot_off)[i] = v;
    }
    __syncthreads();
    // (b) publish: one thread per peer; the barrier above orders the CTA's stores before this thread's system fence
    if (threadIdx.x < world) {
        __threadfence_system();
        st_release_sys(&reinterpret_cast<ArCtrl*>(peers.buf[threadIdx.x])->flags[par][rank], seq);
    }
    // (c) wait for every rank's flag in MY buffer
    if (threadIdx.x < world) {
        const unsigned long long* f = &ctrl->flags[par][threadIdx.x];
        // Rank skew of seconds is normal in training (a rank writing a checkpoint, a data-loader stall, first-step autotune),
        // so the wait is long (AR_TIMEOUT_NS, two minutes) and a timeout is FATAL: reducing stale slots would silently
        // de-synchronise the running statistics of the ranks.  The flag is left for the host (PeerReducer.error()).
        unsigned int spins = 0;
        unsigned long long t0 = 0ull;
        while (ld_acquire_sys(f) < seq) {
            if (++spins > 64) {
                __nanosleep(200);
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0ull) t0 = now;
                else if (now - t0 > AR_TIMEOUT_NS) { s_fail = 1; break; }
            }
        }
    }
    __syncthreads();
    if (s_fail) {
        if (threadIdx.x == 0) { ctrl->error = 1u; __threadfence_system(); }
        __syncthreads();
        __trap();
    }
    // (d) reduce the slots in rank order
    const unsigned char* mine = peers.buf[rank] + (size_t)AR_CTRL_BYTES + (size_t)par * world * slot_bytes;
    for (int i = threadIdx.x; i < n; i += AR_THREADS) {
        T acc = reinterpret_cast<const volatile T*>(mine)[i];
        for (int q = 1; q < world; ++q) acc = ar_combine<T>(acc, reinterpret_cast<const volatile T*>(mine + (size_t)q * slot_bytes)[i], i, mode);
        data[i] = acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ctrl->seq = seq;
        if (s_fail) ctrl->error = 1u;
    }
}

// ---------------------------------------------------------------------------------------------------
// Flagged one-shot exchange (default).  The separate flag + system fence of the kernel above costs a full NVLink round trip
// per call on top of the data: here every 8-byte word that crosses the link carries 4 payload bytes and the 4-byte call tag
// (aligned 8-byte stores are indivisible), so a reader that sees the tag has the data - no fence, no flag, no barrier between
// the write and the read phase (a thread polls exactly the elements it wrote).  Same slots, double buffering and rank-ordered
// reduction as above; a timeout is fatal in the same way.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <typename T>
__global__ void __launch_bounds__(AR_THREADS) k_allreduce_ll(T* __restrict__ data, int n, int mode, ArPeers peers, int rank, int world,
                                                             size_t slot_bytes) {
    constexpr int WORDS = sizeof(T) / 4;                           // 8-byte flagged words per element
    __shared__ unsigned long long s_seq;
    __shared__ int s_fail;
    ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(peers.buf[rank]);
    if (threadIdx.x == 0) { s_seq = ctrl->seq + 1ull; s_fail = 0; }
    __syncthreads();
    const unsigned long long seq = s_seq;
    const unsigned long long tag = (seq & 0xffffffffull) << 32;
    const int par = (int)(seq & 1ull);
    const size_t stride = 2 * slot_bytes;
    const size_t my_off = (size_t)AR_CTRL_BYTES + ((size_t)par * world + rank) * stride;
    for (int i = threadIdx.x; i < n; i += AR_THREADS) {
        const T v = data[i];
        unsigned long long w[WORDS];
        if (WORDS == 2) {
            const unsigned long long b = (unsigned long long)__double_as_longlong((double)v);
            w[0] = (b & 0xffffffffull) | tag;
            w[WORDS - 1] = (b >> 32) | tag;
        } else {
            w[0] = (unsigned long long)__float_as_uint((float)v) | tag;
        }
        for (int p = 0; p < world; ++p) {
            unsigned long long* dst = reinterpret_cast<unsigned long long*>(peers.buf[p] + my_off) + (size_t)i * WORDS;
#pragma unroll
            for (int k = 0; k < WORDS; ++k) st_volatile_u64(dst + k, w[k]);
        }
    }
    const unsigned char* mine = peers.buf[rank] + (size_t)AR_CTRL_BYTES + (size_t)par * world * stride;
    for (int i = threadIdx.x; i < n; i += AR_THREADS) {
        T acc = T(0);
        for (int q = 0; q < world; ++q) {
            const unsigned long long* src = reinterpret_cast<const unsigned long long*>(mine + (size_t)q * stride) + (size_t)i * WORDS;
            unsigned long long w[WORDS];
            unsigned int spins = 0;
            unsigned long long t0 = 0ull;
            bool ok = false;
            while (!ok) {
                ok = true;
#pragma unroll
                for (int k = 0; k < WORDS; ++k) {
                    w[k] = ld_volatile_u64(src + k);
                    ok = ok && ((w[k] & 0xffffffff00000000ull) == tag);
                }
                if (!ok && ++spins > 256) {
                    __nanosleep(100);
                    unsigned long long now;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                    if (t0 == 0ull) t0 = now;
                    else if (now - t0 > AR_TIMEOUT_NS) { s_fail = 1; break; }
                }
            }
            T v;
            if (WORDS == 2) v = (T)__longlong_as_double((long long)((w[0] & 0xffffffffull) | (w[WORDS - 1] << 32)));
            else v = (T)__uint_as_float((unsigned int)(w[0] & 0xffffffffull));
            acc = q == 0 ? v : ar_combine<T>(acc, v, i, mode);
        }
        data[i] = acc;
    }
    __syncthreads();
    if (s_fail) {
        if (threadIdx.x == 0) { ctrl->error = 1u; __threadfence_system(); }
        __syncthreads();
        __trap();
    }
    if (threadIdx.x == 0) ctrl->seq = seq;
}

/* data: n elements of double (dtype 0) or float (dtype 1), reduced IN PLACE across the `world` ranks.
 * peer_bufs[world]: the peer-mapped address of every rank's buffer (qb_allreduce_small_buffer_bytes bytes, zero-filled
 * once before the first call, same slot_bytes everywhere).  Every rank must issue the same sequence of calls. */
extern "C" int qb_allreduce_small(void* data, int64_t n, int dtype, int mode, const uint64_t* peer_bufs, int rank, int world,
                                  size_t slot_bytes, void* stream) {
    QB_CHECK_ARG(data && peer_bufs && n > 0, "qb_allreduce_small: null pointer / empty message");
    QB_CHECK_ARG(world >= 1 && world <= AR_MAX_WORLD && rank >= 0 && rank < world, "qb_allreduce_small: rank %d of %d", rank, world);
    QB_CHECK_ARG(dtype == 0 || dtype == 1, "qb_allreduce_small: dtype %d", dtype);
    QB_CHECK_ARG(mode >= 0 && mode <= 2, "qb_allreduce_small: mode %d", mode);
    const size_t bytes = (size_t)n * (dtype == 0 ? 8 : 4);
    if (bytes > slot_bytes) {
        qb_set_error("qb_allreduce_small: message of %zu bytes exceeds the %zu-byte slot", bytes, slot_bytes);
        return QB_ERR_UNSUPPORTED;
    }
    ArPeers peers;
    for (int p = 0; p < AR_MAX_WORLD; ++p) peers.buf[p] = p < world ? reinterpret_cast<unsigned char*>(peer_bufs[p]) : nullptr;
    cudaStream_t st = qb_stream(stream);
    static const bool flagged = []() { const char* e = getenv("QB_AR_FLAGGED"); return !(e && e[0] == '0'); }();
    if (flagged) {
        if (dtype == 0) k_allreduce_ll<double><<<1, AR_THREADS, 0, st>>>(static_cast<double*>(data), (int)n, mode, peers, rank, world, slot_bytes);
        else k_allreduce_ll<float><<<1, AR_THREADS, 0, st>>>(static_cast<float*>(data), (int)n, mode, peers, rank, world, slot_bytes);
    } else if (dtype == 0) k_allreduce_oneshot<double><<<1, AR_THREADS, 0, st>>>(static_cast<double*>(data), (int)n, mode, peers, rank, world, slot_bytes);
    else k_allreduce_oneshot<float><<<1, AR_THREADS, 0, st>>>(static_cast<float*>(data), (int)n, mode, peers, rank, world, slot_bytes);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

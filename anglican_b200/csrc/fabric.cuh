// fabric.cuh — peer-memory fabric of a sharded particle population.
//
// The reference has no multi-process code; this layer is new.
// A population of N particles is split into `world` contiguous shards, one
// per GPU (global id = rank * n_local + local).  Every rank owns one
// "symmetric heap" (a single cudaMalloc: mailbox + both state buffers + the
// ancestor array) that all peers map — through CUDA IPC between processes, or
// directly inside one process.  Kernels then talk over NVLink/NVSwitch with
// plain loads and stores:
//   * k_ancestors stores the offspring indices straight into the ancestor
//     array of the rank that owns the output slot (the particle migration of
//     SURVEY §8(e) never materialises as a separate all-to-all);
//   * k_propagate reads the parent state through the owner's heap pointer;
//   * the scalars of the per-step exchange (local max / weight sum, each shard's
//     contribution to the sequential fold of smc.cljc:94,118-137 -- a composable
//     transducer where the shard stays in one binade, its exact output where it
//     does not -- and the accurate sums) travel through a small mailbox with
//     release/acquire flags at system scope.
// NCCL is used only to bootstrap (exchange of the IPC handles) and for the
// IPMCMC log-Z allgather.
#pragma once
#include <stdint.h>

namespace apop {

constexpr int kMaxWorld = 8;

struct MailSlot {
    double v[3];
    unsigned long long seq;
};

struct Mailbox {
    MailSlot fin[2][kMaxWorld];   // per source rank: (local max, local sum rel. local max, err bits)
    MailSlot acc[2][kMaxWorld];   // per source rank: accurate (sum w, sum w^2, sum w x)
    MailSlot comp[2][2][kMaxWorld];   // [parity][lap][source rank]: the shard's part of the exact
                                      //   fold (CK_EASY: transducer + binade, CK_EXACT: running
                                      //   sum leaving the shard), see k_chain
    unsigned long long bar[kMaxWorld];
    unsigned long long pad[8];
};

constexpr size_t kMailboxBytes = (sizeof(Mailbox) + 4095) & ~(size_t)4095;

// by-value kernel argument
struct Fabric {
    int rank, world;
    long long n_local;
    long long spin_cycles;        // bound of every wait on a peer (APOP_SPIN_SECONDS, default 3 s)
    Mailbox* mbox[kMaxWorld];     // mbox[q]: rank q's mailbox as mapped here
};
struct PeerF64 {
    const double* p[kMaxWorld];
};
struct PeerI32 {
    int32_t* p[kMaxWorld];
};

// which wait timed out, recorded beside EB_COMM: (what << 16) | (peer << 24)
enum { CW_FIN = 1, CW_ACC = 2, CW_COMP0 = 3, CW_COMP1 = 4, CW_BARRIER = 5 };
#if defined(__CUDACC__)
constexpr long long kSpinLimitCycles = 6000000000LL;     // ~3 s at 2 GHz: a hang becomes an error

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mail_send(MailSlot* s, double a, double b, double c,
                                          unsigned long long seq) {
    volatile double* v = s->v;
    v[0] = a; v[1] = b; v[2] = c;
    __threadfence_system();
    st_release_sys(&s->seq, seq);
}
// false on timeout (the peer never arrived)
__device__ __forceinline__ bool spin_until(const unsigned long long* flag, unsigned long long seq,
                                           long long limit = kSpinLimitCycles) {
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < seq) {
        if (clock64() - t0 > limit) return false;
        __nanosleep(64);
    }
    return true;
}
// raw 64-bit payloads (integer transducers travel bit for bit)
__device__ __forceinline__ void mail_send_bits(MailSlot* s, long long a, long long b, long long c,
                                               unsigned long long seq) {
    volatile long long* v = reinterpret_cast<volatile long long*>(s->v);
    v[0] = a; v[1] = b; v[2] = c;
    __threadfence_system();
    st_release_sys(&s->seq, seq);
}
__device__ __forceinline__ bool mail_wait_bits(MailSlot* s, unsigned long long seq,
                                               long long (&out)[3],
                                               long long limit = kSpinLimitCycles) {
    const bool ok = spin_until(&s->seq, seq, limit);
    volatile long long* v = reinterpret_cast<volatile long long*>(s->v);
    out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    return ok;
}
__device__ __forceinline__ bool mail_wait(MailSlot* s, unsigned long long seq, double (&out)[3],
                                          long long limit = kSpinLimitCycles) {
    const bool ok = spin_until(&s->seq, seq, limit);
    volatile double* v = s->v;
    out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    return ok;
}
#endif

} // namespace apop

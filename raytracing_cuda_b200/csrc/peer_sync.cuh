/*
 * peer_sync.cuh -- device side of the peer-frame hand-shake (peer.cu, fxaa.cu): system-scope flag loads / stores
 * and the two halves a frame's last kernel carries (RtPeerSync, rt_device.h).
 */
#ifndef RTB_PEER_SYNC_CUH
#define RTB_PEER_SYNC_CUH

#include "rt_device.h"

__device__ __forceinline__ unsigned int rt_peer_load(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void rt_peer_store(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long rt_peer_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

/* spin (one thread) until *flag >= seq (signed difference: sequence numbers may wrap); false after timeout_ms, with
 * `code` left in the error word -- the frame is then invalid and the caller must not store into the buffer */
__device__ __forceinline__ bool rt_peer_spin(unsigned int* flags, int flag, unsigned int seq, unsigned int timeout_ms,
                                             unsigned int code) {
    if ((int)(rt_peer_load(flags + flag) - seq) >= 0) return true;
    const unsigned long long t0 = rt_peer_now(), lim = (unsigned long long)timeout_ms * 1000000ull;
    while ((int)(rt_peer_load(flags + flag) - seq) < 0) {
        if (rt_peer_now() - t0 > lim) {
            rt_peer_store(flags + RT_PEER_FLAG_ERROR, code);
            return false;
        }
        __nanosleep(200);
    }
    return true;
}

/* Every thread of every CTA of the frame's last kernel calls this after its last store into the peer buffer (CTAs
 * that stored nothing too): the CTA's stores are fenced at system scope, the CTA takes a ticket, and the CTA that
 * draws the last one re-arms the counter and publishes the rank's frame number.  Causality: stores -> fence.sys ->
 * ticket RMW (every CTA) ... last RMW -> fence.sys -> st.release.sys, so a consumer that acquires the flag sees every
 * CTA's pixels. */
__device__ __forceinline__ void rt_peer_publish(const RtPeerSync& p, unsigned int n_ctas, bool leader) {
    __threadfence_system();
    __syncthreads();
    if (leader) {
        if (atomicAdd(p.ticket, 1u) == n_ctas - 1u) {
            *p.ticket = 0u;
            __threadfence_system();
            rt_peer_store(p.flags + RT_PEER_FLAG_DONE(p.rank), p.seq);
        }
    }
}

#endif

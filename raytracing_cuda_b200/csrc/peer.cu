/*
 * peer.cu -- flags of the peer frame: one frame over N GPUs without a gather.
 *
 * every GPU traces + filters its interleaved screen tiles.  Instead of packing them, gathering
 * them over NCCL and unpacking them on GPU 0, the FXAA (or copy) kernel of every rank stores its pixels
 * straight into GPU 0's frame buffer through a peer mapping (NVLink 5 / NVSwitch; cudaIpc handle opened by the
 * other processes) -- the compute step and the transfer are one kernel.  What is left of the collective is the
 * hand-shake below, four tiny kernels on the ranks' own streams over flag words that live behind the frame
 * in GPU 0's allocation:
 *
 *   producer k, frame t:  rt_peer_acquire  wait until the consumer released frame t - buffers (remote loads)
 *                         ... cull, trace, FXAA -> peer stores ...
 *                         rt_peer_signal   done[k] = t   (after a system-scope fence)
 *   consumer,   frame t:  rt_peer_wait     spin until done[k] >= t for every k
 *                         ... use the frame ...
 *                         rt_peer_release  released = t
 *
 * Stream order puts a rank's signal after its FXAA kernel; a kernel boundary completes that kernel's stores.
 * Every wait gives up after RT_PEER_TIMEOUT_NS and raises the error word instead of hanging the GPU.
 */
#include <cuda_runtime.h>

#include "rt_device.h"

#define RT_PEER_TIMEOUT_NS 2000000000ull   /* 2 s */

/* flag block layout: rt_device.h (RT_PEER_FLAG_*) */

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

__global__ void rt_peer_signal_kernel(unsigned int* flags, int rank, unsigned int seq) {
    __threadfence_system();
    rt_peer_store(flags + RT_PEER_FLAG_DONE(rank), seq);
}

/* sequence numbers are compared as signed differences so that they may wrap */
__global__ void rt_peer_wait_kernel(unsigned int* flags, int n_ranks, unsigned int seq) {
    const int k = threadIdx.x;
    if (k >= n_ranks) return;
    const unsigned long long t0 = rt_peer_now();
    while ((int)(rt_peer_load(flags + RT_PEER_FLAG_DONE(k)) - seq) < 0) {
        if (rt_peer_now() - t0 > RT_PEER_TIMEOUT_NS) {
            rt_peer_store(flags + RT_PEER_FLAG_ERROR, 1u + (unsigned int)k);
            return;
        }
        __nanosleep(200);
    }
}

__global__ void rt_peer_release_kernel(unsigned int* flags, unsigned int seq) {
    rt_peer_store(flags + RT_PEER_FLAG_RELEASED, seq);
}

__global__ void rt_peer_acquire_kernel(unsigned int* flags, unsigned int seq) {
    const unsigned long long t0 = rt_peer_now();
    while ((int)(rt_peer_load(flags + RT_PEER_FLAG_RELEASED) - seq) < 0) {
        if (rt_peer_now() - t0 > RT_PEER_TIMEOUT_NS) {
            rt_peer_store(flags + RT_PEER_FLAG_ERROR, 0x100u);
            return;
        }
        __nanosleep(500);
    }
}

extern "C" cudaError_t rt_launch_peer_signal(unsigned int* flags, int rank, unsigned int seq, cudaStream_t stream) {
    rt_peer_signal_kernel<<<1, 1, 0, stream>>>(flags, rank, seq);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_wait(unsigned int* flags, int n_ranks, unsigned int seq, cudaStream_t stream) {
    rt_peer_wait_kernel<<<1, 32, 0, stream>>>(flags, n_ranks, seq);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_release(unsigned int* flags, unsigned int seq, cudaStream_t stream) {
    rt_peer_release_kernel<<<1, 1, 0, stream>>>(flags, seq);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_acquire(unsigned int* flags, unsigned int seq, cudaStream_t stream) {
    rt_peer_acquire_kernel<<<1, 1, 0, stream>>>(flags, seq);
    return cudaGetLastError();
}

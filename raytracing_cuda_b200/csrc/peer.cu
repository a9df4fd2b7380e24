/*
 * peer.cu -- flags of the peer frame: one frame over N GPUs without a gather.
 *
 * every GPU traces + filters its interleaved screen tiles.  Instead of packing them, gathering
 * them over NCCL and unpacking them on GPU 0, the FXAA (or copy) kernel of every rank stores its pixels
 * straight into GPU 0's frame buffer through a peer mapping (NVLink 5 / NVSwitch; cudaIpc handle opened by the
 * other processes) -- the compute step and the transfer are one kernel.  What is left of the collective is the
 * hand-shake below over flag words that live behind the frame in GPU 0's allocation:
 *
 *   producer k, frame t:  acquire   wait until the consumer released frame t - buffers (remote loads)
 *                         ... cull, trace, FXAA -> peer stores ...
 *                         signal    done[k] = t   (after a system-scope fence)
 *   consumer,   frame t:  rt_peer_wait     spin until done[k] >= t for every k
 *                         ... use the frame ...
 *                         rt_peer_release  released = t
 *
 * rt_render_peer folds the producer's two halves into the frame's last kernel (RtPeerSync, peer_sync.cuh, fxaa.cu:
 * every CTA acquires before its first peer store, the last CTA to finish signals), so a producer's frame is three
 * launches -- pre-pass, tracer, FXAA -- and nothing else; the one-thread acquire / signal kernels below remain for
 * callers that drive the hand-shake themselves (rt_peer_frame_acquire / _signal).
 * Every wait gives up after the frame's timeout (default RT_PEER_TIMEOUT_MS_DEFAULT, rt_peer_frame_set_timeout) and
 * raises the error word instead of hanging the GPU; the frame is then INVALID (a producer that gave up stores
 * nothing) and stays so until rt_peer_frame_reset_error.
 */
#include <cuda_runtime.h>

#include "peer_sync.cuh"

/* flag block layout: rt_device.h (RT_PEER_FLAG_*) */

__global__ void rt_peer_signal_kernel(unsigned int* flags, int rank, unsigned int seq) {
    __threadfence_system();
    rt_peer_store(flags + RT_PEER_FLAG_DONE(rank), seq);
}

/* sequence numbers are compared as signed differences so that they may wrap */
__global__ void rt_peer_wait_kernel(unsigned int* flags, int n_ranks, unsigned int seq, unsigned int timeout_ms) {
    const int k = threadIdx.x;
    if (k >= n_ranks) return;
    rt_peer_spin(flags, RT_PEER_FLAG_DONE(k), seq, timeout_ms, 1u + (unsigned int)k);
}

__global__ void rt_peer_release_kernel(unsigned int* flags, unsigned int seq) {
    rt_peer_store(flags + RT_PEER_FLAG_RELEASED, seq);
}

__global__ void rt_peer_acquire_kernel(unsigned int* flags, unsigned int seq, unsigned int timeout_ms) {
    rt_peer_spin(flags, RT_PEER_FLAG_RELEASED, seq, timeout_ms, 0x100u);
}

__global__ void rt_peer_reset_error_kernel(unsigned int* flags) { rt_peer_store(flags + RT_PEER_FLAG_ERROR, 0u); }

extern "C" cudaError_t rt_launch_peer_signal(unsigned int* flags, int rank, unsigned int seq, cudaStream_t stream) {
    rt_peer_signal_kernel<<<1, 1, 0, stream>>>(flags, rank, seq);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_wait(unsigned int* flags, int n_ranks, unsigned int seq, unsigned int timeout_ms,
                                           cudaStream_t stream) {
    rt_peer_wait_kernel<<<1, 32, 0, stream>>>(flags, n_ranks, seq, timeout_ms);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_release(unsigned int* flags, unsigned int seq, cudaStream_t stream) {
    rt_peer_release_kernel<<<1, 1, 0, stream>>>(flags, seq);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_acquire(unsigned int* flags, unsigned int seq, unsigned int timeout_ms,
                                              cudaStream_t stream) {
    rt_peer_acquire_kernel<<<1, 1, 0, stream>>>(flags, seq, timeout_ms);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}
extern "C" cudaError_t rt_launch_peer_reset_error(unsigned int* flags, cudaStream_t stream) {
    rt_peer_reset_error_kernel<<<1, 1, 0, stream>>>(flags);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}

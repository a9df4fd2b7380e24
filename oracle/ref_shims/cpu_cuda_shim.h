/* Host stand-ins for the CUDA device vocabulary used by the reference's
 * kernel.cu:9-403, so that the reference's OWN text runs on the CPU as a
 * cross-check of oracle/rt_oracle.cpp.  TEST INFRASTRUCTURE ONLY.
 *
 *  - __device__/__global__/__constant__ vanish, __shared__ becomes per-host-thread
 *    static storage;
 *  - threadIdx/blockIdx/blockDim are thread-local variables set by the driver
 *    loop in ref_trace_cpu_wrap.inc;
 *  - __syncthreads() splits a block into two sequential passes: in pass 0 it
 *    unwinds the emulated thread (everything before the barrier only fills the
 *    shared tile), in pass 1 it is a no-op;
 *  - tex2D is a point-sampled, clamp-addressed, normalised-coordinate fetch;
 *  - __powf(x,y) = exp2f(y*log2f(x)); max(double,float) as in CUDA's overload set.
 */
#pragma once
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector_types.h>

#undef __device__
#undef __global__
#undef __constant__
#undef __shared__
#undef __host__
#define __device__
#define __global__
#define __constant__
#define __host__
#define __shared__ static thread_local

enum { cudaReadModeElementType = 0 };
template <class T, int Dim, int Mode>
struct texture {
    const unsigned char* data = nullptr;
    int w = 0, h = 0;
};
template <class T, int Dim, int Mode>
inline uchar4 tex2D(const texture<T, Dim, Mode>& t, float u, float v) {
    int ix = (int)floorf(u * (float)t.w);
    int iy = (int)floorf(v * (float)t.h);
    ix = std::min(std::max(ix, 0), t.w - 1);
    iy = std::min(std::max(iy, 0), t.h - 1);
    const unsigned char* p = t.data + 4 * ((size_t)iy * t.w + ix);
    uchar4 r;
    r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3];
    return r;
}

inline double max(double a, float b) { return a > (double)b ? a : (double)b; }
inline float __powf(float x, float y) { return exp2f(y * log2f(x)); }

static thread_local uint3 threadIdx, blockIdx;
static thread_local dim3 blockDim;
static thread_local int ref_sync_pass = 1;
struct RefSyncReached {};
inline void __syncthreads() {
    if (ref_sync_pass == 0) throw RefSyncReached();
}

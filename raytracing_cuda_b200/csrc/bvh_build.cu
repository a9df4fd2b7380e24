/*
 * bvh_build.cu -- GPU LBVH build for the tracer (no counterpart in the reference,
 * which loops linearly over 133 objects held in __constant__ memory,
 * reference kernel.cu:14, 144-151).
 *
 * Pipeline (all on the context's stream, no host round trip):
 *   rt_setup_kernel    Object AoS (68 B, reference structs.h:27-38) -> leaf records, materials,
 *                      per-primitive AABB, centroid bounds (atomic min/max on ordered ints)
 *   rt_morton_kernel   30-bit Morton code of the AABB centroid in the centroid bounds
 *   sort               stable LSD radix sort of (code, id): 4 passes x 8 bits, per pass
 *                      histogram -> scan -> ranked scatter (warp match/ballot ranking);
 *                      scenes <= 2048 primitives use a single-CTA bitonic sort instead
 *   rt_karras_kernel   Karras 2012 hierarchy: one thread per internal node, common-prefix
 *                      length via clz, index as tie-breaker for equal codes
 *   rt_refit_kernel    bottom-up: leaves write their padded box into the parent's record and
 *                      the second arrival at each node (atomic counter) carries the union up
 *
 * Node / leaf-record layout: see rt_device.h.  Refit alone (same topology, moved
 * primitives) re-runs rt_setup_kernel + rt_refit_kernel.
 *
 * Leaf boxes are padded by 2^-15 of their largest |coordinate| plus 2^-20 of the scene
 * extent, which is >100x the rounding slack of the slab test and of the Moller-Trumbore
 * acceptance region, so the hierarchy never culls a primitive the leaf test would accept.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "rt_device.h"

#define RT_BUILD_THREADS 256

/* order-preserving float <-> uint mapping for atomicMin/Max */
__device__ __forceinline__ unsigned int f2ord(float f) {
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

/* bounds[0..2] = min centroid, [3..5] = max centroid, [6] = max |coordinate| (all ordered-uint) */
extern "C" __global__ void rt_bounds_init_kernel(unsigned int* bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xFFFFFFFFu;
    else if (threadIdx.x < 7) bounds[threadIdx.x] = 0u;
}

/* One object of the reference's AoS (68 B, structs.h:27-38) -> leaf record, material, padded-later AABB.
 * Fills the ordered-uint centroid bounds of this object into bmin/bmax (neutral elements for planes). */
__device__ __forceinline__ void rt_setup_object(int i, const float* o, const int* __restrict__ plane_idx,
                                                int n_planes, float4* __restrict__ mats, float4* __restrict__ recs,
                                                float4* __restrict__ planes, int* __restrict__ bvh_src,
                                                float4* __restrict__ prim_lo, float4* __restrict__ prim_hi,
                                                unsigned int bmin[3], unsigned int bmax[4]) {
    /* o: the 17 words of object i (global, shared or parameter space) */
    const int type = __float_as_int(o[0]);
    const float shine = o[1], specular = o[2], mirror = o[3];
    const float3 color = make_float3(o[4], o[5], o[6]);
    const float3 pos = make_float3(o[7], o[8], o[9]);
    const float3 size = make_float3(o[10], o[11], o[12]);
    const float3 third = make_float3(o[13], o[14], o[15]);
    const bool emissive = (__float_as_uint(o[16]) & 0xFFu) != 0;

    unsigned int tag = (unsigned int)i | (emissive ? RT_TAG_EMISSIVE : 0u);
    mats[2 * i] = make_float4(color.x, color.y, color.z, mirror);
    mats[2 * i + 1] = make_float4(shine, specular, emissive ? 1.0f : 0.0f, (float)type);

    int planes_before = 0, my_plane = -1;
    for (int p = 0; p < n_planes; p++) {
        const int pi = plane_idx[p];
        if (pi < i) planes_before++;
        if (pi == i) my_plane = p;
    }
    if (my_plane >= 0) {
        tag |= RT_TAG_PLANE;
        /* unknown primitive types never hit in the reference (kernel.cu:128): zero normal */
        const float3 nrm = (type == RTB_PLANE) ? size : make_float3(0.f, 0.f, 0.f);
        const float4 q0 = make_float4(pos.x, pos.y, pos.z, __uint_as_float(tag));
        const float4 q1 = make_float4(nrm.x, nrm.y, nrm.z, 0.f);
        recs[3 * i] = q0; recs[3 * i + 1] = q1; recs[3 * i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        planes[2 * my_plane] = q0;
        planes[2 * my_plane + 1] = q1;
        return;
    }
    const int k = i - planes_before;
    bvh_src[k] = i;
    float3 lo, hi;
    if (type == RTB_SPHERE) {
        tag |= RT_TAG_SPHERE;
        const float r = size.x;
        recs[3 * i] = make_float4(pos.x, pos.y, pos.z, __uint_as_float(tag));
        recs[3 * i + 1] = make_float4(r, __fmul_rn(r, r), 0.f, 0.f);
        recs[3 * i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        const float ar = fabsf(r);
        lo = make_float3(pos.x - ar, pos.y - ar, pos.z - ar);
        hi = make_float3(pos.x + ar, pos.y + ar, pos.z + ar);
    } else {
        /* e1 = v1 - v0, e2 = v2 - v0: the same FADDs the reference does per test (kernel.cu:100-101) */
        recs[3 * i] = make_float4(pos.x, pos.y, pos.z, __uint_as_float(tag));
        recs[3 * i + 1] = make_float4(__fadd_rn(size.x, -pos.x), __fadd_rn(size.y, -pos.y), __fadd_rn(size.z, -pos.z), 0.f);
        recs[3 * i + 2] = make_float4(__fadd_rn(third.x, -pos.x), __fadd_rn(third.y, -pos.y), __fadd_rn(third.z, -pos.z), 0.f);
        lo = make_float3(fminf(pos.x, fminf(size.x, third.x)), fminf(pos.y, fminf(size.y, third.y)),
                         fminf(pos.z, fminf(size.z, third.z)));
        hi = make_float3(fmaxf(pos.x, fmaxf(size.x, third.x)), fmaxf(pos.y, fmaxf(size.y, third.y)),
                         fmaxf(pos.z, fmaxf(size.z, third.z)));
    }
    prim_lo[k] = make_float4(lo.x, lo.y, lo.z, 0.f);
    prim_hi[k] = make_float4(hi.x, hi.y, hi.z, 0.f);
    const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    const float ma = fmaxf(fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fabsf(lo.y), fabsf(hi.y))),
                           fmaxf(fabsf(lo.z), fabsf(hi.z)));
    bmin[0] = f2ord(cx); bmin[1] = f2ord(cy); bmin[2] = f2ord(cz);
    bmax[0] = f2ord(cx); bmax[1] = f2ord(cy); bmax[2] = f2ord(cz); bmax[3] = f2ord(ma);
}

/* warp reduction of the centroid bounds, then one atomic per warp and bound (bounds: global or shared) */
__device__ __forceinline__ void rt_merge_bounds(unsigned int bmin[3], unsigned int bmax[4], unsigned int* bounds) {
#pragma unroll
    for (int a = 0; a < 3; a++) bmin[a] = __reduce_min_sync(0xFFFFFFFFu, bmin[a]);
#pragma unroll
    for (int a = 0; a < 4; a++) bmax[a] = __reduce_max_sync(0xFFFFFFFFu, bmax[a]);
    if ((threadIdx.x & 31) == 0 && bmin[0] != 0xFFFFFFFFu) {
        atomicMin(&bounds[0], bmin[0]); atomicMin(&bounds[1], bmin[1]); atomicMin(&bounds[2], bmin[2]);
        atomicMax(&bounds[3], bmax[0]); atomicMax(&bounds[4], bmax[1]); atomicMax(&bounds[5], bmax[2]);
        atomicMax(&bounds[6], bmax[3]);
    }
}

__global__ void rt_setup_kernel(const float* __restrict__ objs, int n_objects, const int* __restrict__ plane_idx,
                                int n_planes, float4* __restrict__ mats, float4* __restrict__ recs,
                                float4* __restrict__ planes, int* __restrict__ bvh_src,
                                float4* __restrict__ prim_lo, float4* __restrict__ prim_hi,
                                unsigned int* __restrict__ bounds) {
    /* the block's objects (68 B each, array of structures) through shared memory: coalesced loads, and a
     * stride of 17 words per thread is free of bank conflicts */
    __shared__ __align__(16) float s_obj[17 * RT_BUILD_THREADS];
    const int first = blockIdx.x * blockDim.x;
    const int here = min((int)blockDim.x, n_objects - first);
    if (here == RT_BUILD_THREADS && blockDim.x == RT_BUILD_THREADS && ((uintptr_t)objs & 15) == 0) {
        /* full block: 17,408 B = 1,088 float4 from a 16-byte aligned start, all of a thread's loads in flight at
         * once (the word-by-word loop below kept one load per thread in flight: 71 % of the kernel's stalls) */
        const float4* __restrict__ src = (const float4*)(objs + (size_t)first * 17);
        float4 v[5];
#pragma unroll
        for (int q = 0; q < 5; q++) {
            const int k = threadIdx.x + RT_BUILD_THREADS * q;
            if (k < 17 * RT_BUILD_THREADS / 4) v[q] = __ldg(src + k);
        }
#pragma unroll
        for (int q = 0; q < 5; q++) {
            const int k = threadIdx.x + RT_BUILD_THREADS * q;
            if (k < 17 * RT_BUILD_THREADS / 4) ((float4*)s_obj)[k] = v[q];
        }
    } else {
        for (int k = threadIdx.x; k < 17 * here; k += blockDim.x) s_obj[k] = __ldg(objs + (size_t)first * 17 + k);
    }
    __syncthreads();
    const int i = first + threadIdx.x;
    /* neutral elements for the warp-level bounds reduction */
    unsigned int bmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, bmax[4] = {0u, 0u, 0u, 0u};
    if (i < n_objects) rt_setup_object(i, s_obj + 17 * threadIdx.x, plane_idx, n_planes, mats, recs, planes, bvh_src, prim_lo, prim_hi, bmin, bmax);
    /* warps -> shared memory -> one set of seven atomics per block: a million primitives would otherwise queue
     * 220,000 atomics on one 32-byte sector (measured: the kernel spent most of its 167 us there) */
    __shared__ unsigned int s_bounds[8];
    if (threadIdx.x < 3) s_bounds[threadIdx.x] = 0xFFFFFFFFu;
    else if (threadIdx.x < 8) s_bounds[threadIdx.x] = 0u;
    __syncthreads();
    rt_merge_bounds(bmin, bmax, s_bounds);
    __syncthreads();
    if (threadIdx.x < 3) { if (s_bounds[threadIdx.x] != 0xFFFFFFFFu) atomicMin(&bounds[threadIdx.x], s_bounds[threadIdx.x]); }
    else if (threadIdx.x < 7) atomicMax(&bounds[threadIdx.x], s_bounds[threadIdx.x]);
}

__device__ __forceinline__ unsigned int expand10(unsigned int v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void rt_morton_kernel(const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi, int n,
                                 const unsigned int* __restrict__ bounds, unsigned int* __restrict__ keys,
                                 unsigned int* __restrict__ vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const float4 lo = prim_lo[k], hi = prim_hi[k];
    const float mnx = ord2f(bounds[0]), mny = ord2f(bounds[1]), mnz = ord2f(bounds[2]);
    const float ex = ord2f(bounds[3]) - mnx, ey = ord2f(bounds[4]) - mny, ez = ord2f(bounds[5]) - mnz;
    const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    const float ux = ex > 0.f ? (cx - mnx) / ex : 0.f;
    const float uy = ey > 0.f ? (cy - mny) / ey : 0.f;
    const float uz = ez > 0.f ? (cz - mnz) / ez : 0.f;
    const unsigned int qx = (unsigned int)fminf(fmaxf(ux * 1024.f, 0.f), 1023.f);
    const unsigned int qy = (unsigned int)fminf(fmaxf(uy * 1024.f, 0.f), 1023.f);
    const unsigned int qz = (unsigned int)fminf(fmaxf(uz * 1024.f, 0.f), 1023.f);
    keys[k] = (expand10(qx) << 2) | (expand10(qy) << 1) | expand10(qz);
    vals[k] = (unsigned int)k;
}

/* ------------------------------------------------------------------ sort */

#define RS_THREADS 256
#ifndef RS_KEYS_PER_THREAD
#define RS_KEYS_PER_THREAD 8          /* 2,048 keys per block: a 1M-key sort has 490 blocks (16: 245 blocks, 1.65 waves of
                                         long serial ranking loops -- 118 us for the four passes against 101 with 8) */
#endif
#define RS_TILE (RS_THREADS * RS_KEYS_PER_THREAD)
#define RS_WARPS (RS_THREADS / 32)

/* per-block digit histogram -> hist[digit * n_blocks + block] */
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const unsigned int* __restrict__ keys, int n, int shift, unsigned int* __restrict__ hist, int n_blocks) {
    __shared__ unsigned int s_h[256];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
    /* all of a thread's keys in flight before the first shared-memory atomic (one round trip, not RS_KEYS_PER_THREAD) */
    unsigned int k[RS_KEYS_PER_THREAD];
#pragma unroll
    for (int s = 0; s < RS_KEYS_PER_THREAD; s++) {
        const int g = base + s * RS_THREADS + threadIdx.x;
        k[s] = g < n ? __ldg(keys + g) : 0u;
    }
#pragma unroll
    for (int s = 0; s < RS_KEYS_PER_THREAD; s++)
        if (base + s * RS_THREADS + (int)threadIdx.x < n) atomicAdd(&s_h[(k[s] >> shift) & 255u], 1u);
    __syncthreads();
    hist[threadIdx.x * n_blocks + blockIdx.x] = s_h[threadIdx.x];
}

/* exclusive scan of every bin's row hist[bin][0 .. n_blocks) in place, bin totals to hist[256 n_blocks + bin]; the
 * scatter kernel adds the exclusive scan of the 256 totals itself.  RS_SCAN_WARPS warps share a bin: each sums its
 * contiguous quarter of the row, the quarters' sums meet in shared memory, then each scans its quarter from its
 * offset (the re-read hits L1) -- two short round trips for a 1M-key sort's 490 counters per row where one warp per
 * bin made four.  (History: a flat single-CTA scan with four barriers per 1024 counters took 61 us for the 62,720
 * counters of a 1M-key sort; one CTA with 8 bins per warp and two walks over every row 21 us; one warp per bin
 * 5.4 us per pass.) */
#define RS_SCAN_LOADS 4                /* 32-wide loads in flight per step */
#define RS_SCAN_WARPS 4                /* warps per bin */
__global__ void __launch_bounds__(256) rs_scan_kernel(unsigned int* __restrict__ hist, int n_blocks) {
    __shared__ unsigned int s_part[256 / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x * (8 / RS_SCAN_WARPS) + warp / RS_SCAN_WARPS, q4 = warp % RS_SCAN_WARPS;
    unsigned int* __restrict__ row = hist + (size_t)b * n_blocks;
    /* this warp's part of the row: [j_lo, j_hi), a multiple of 32 long except at the end */
    const int seg = ((n_blocks + RS_SCAN_WARPS - 1) / RS_SCAN_WARPS + 31) & ~31;
    const int j_lo = min(q4 * seg, n_blocks), j_hi = min(j_lo + seg, n_blocks);
    unsigned int sum = 0;
    for (int j = j_lo + lane; j < j_hi; j += 32) sum += row[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
    if (lane == 0) s_part[warp] = sum;
    __syncthreads();
    unsigned int carry = 0;
    for (int w = 0; w < RS_SCAN_WARPS; w++)
        if (w < q4) carry += s_part[warp - q4 + w];
    for (int j0 = j_lo; j0 < j_hi; j0 += 32 * RS_SCAN_LOADS) {
        unsigned int v[RS_SCAN_LOADS], s[RS_SCAN_LOADS];
#pragma unroll
        for (int q = 0; q < RS_SCAN_LOADS; q++) {
            const int j = j0 + 32 * q + lane;
            v[q] = j < j_hi ? row[j] : 0u;
        }
#pragma unroll
        for (int q = 0; q < RS_SCAN_LOADS; q++) {
            s[q] = v[q];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, s[q], o);
                if (lane >= o) s[q] += t;
            }
        }
#pragma unroll
        for (int q = 0; q < RS_SCAN_LOADS; q++) {
            const int j = j0 + 32 * q + lane;
            if (j < j_hi) row[j] = carry + s[q] - v[q];
            carry += __shfl_sync(0xFFFFFFFFu, s[q], 31);
        }
    }
    if (q4 == RS_SCAN_WARPS - 1 && lane == 0) hist[(size_t)256 * n_blocks + b] = carry;
}

/* stable ranked scatter.  Warp w owns the contiguous chunk [w*32*RS_KEYS_PER_THREAD, (w+1)*32*RS_KEYS_PER_THREAD) of the block's
 * tile and walks it in RS_KEYS_PER_THREAD steps of 32 keys, so rank order == memory order. */
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const unsigned int* __restrict__ keys_in, const unsigned int* __restrict__ vals_in,
                  unsigned int* __restrict__ keys_out, unsigned int* __restrict__ vals_out, int n, int shift,
                  const unsigned int* __restrict__ hist, int n_blocks) {
    __shared__ unsigned int s_cnt[RS_WARPS][256];
    __shared__ unsigned int s_tot[RS_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&s_cnt[0][0])[i] = 0;
    __syncthreads();

    const int chunk = blockIdx.x * RS_TILE + warp * (32 * RS_KEYS_PER_THREAD);
    unsigned int key[RS_KEYS_PER_THREAD], val[RS_KEYS_PER_THREAD], rank[RS_KEYS_PER_THREAD];
    /* every load of the thread in flight before the ranking loop (its ballots and barriers would pin each load to
     * its own step) */
#pragma unroll
    for (int s = 0; s < RS_KEYS_PER_THREAD; s++) {
        const int g = chunk + s * 32 + lane;
        key[s] = g < n ? __ldg(keys_in + g) : 0xFFFFFFFFu;
        val[s] = g < n ? __ldg(vals_in + g) : 0u;
    }
#pragma unroll
    for (int s = 0; s < RS_KEYS_PER_THREAD; s++) {
        const int g = chunk + s * 32 + lane;
        const bool valid = g < n;
        const unsigned int digit = (key[s] >> shift) & 255u;
        const unsigned int vmask = __ballot_sync(0xFFFFFFFFu, valid);
        const unsigned int peers = __match_any_sync(0xFFFFFFFFu, digit) & vmask;
        const unsigned int below = peers & ((1u << lane) - 1u);
        unsigned int base = 0;
        const int leader = __ffs(peers) - 1;
        if (valid && lane == leader) {
            base = s_cnt[warp][digit];
            s_cnt[warp][digit] = base + __popc(peers);
        }
        base = __shfl_sync(0xFFFFFFFFu, base, leader < 0 ? 0 : leader);
        rank[s] = base + __popc(below);
        __syncwarp();
    }
    __syncthreads();
    /* digit d = threadIdx.x: exclusive prefix over warps + global base = (keys with a smaller digit: exclusive
     * scan of the 256 bin totals, 8 warp scans + their totals through s_tot) + (this digit in earlier blocks) */
    {
        const unsigned int d = threadIdx.x;
        const unsigned int tot = hist[(size_t)256 * n_blocks + d];
        unsigned int incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_tot[warp] = incl;
        __syncthreads();
        unsigned int before = incl - tot;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) before += w < warp ? s_tot[w] : 0u;
        unsigned int run = before + hist[d * n_blocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const unsigned int c = s_cnt[w][d];
            s_cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < RS_KEYS_PER_THREAD; s++) {
        const int g = chunk + s * 32 + lane;
        if (g < n) {
            const unsigned int digit = (key[s] >> shift) & 255u;
            const unsigned int dst = s_cnt[warp][digit] + rank[s];
            keys_out[dst] = key[s];
            vals_out[dst] = val[s];
        }
    }
}

/* single-CTA bitonic sort of (key,val) for small scenes; 64-bit composite keeps it stable */
#define SS_MAX 2048
__global__ void __launch_bounds__(1024) rs_small_sort_kernel(unsigned int* __restrict__ keys,
                                                             unsigned int* __restrict__ vals, int n) {
    __shared__ unsigned long long s[SS_MAX];
    int m = 1;
    while (m < n) m <<= 1;
    for (int i = threadIdx.x; i < m; i += blockDim.x)
        s[i] = i < n ? (((unsigned long long)keys[i] << 32) | vals[i]) : 0xFFFFFFFFFFFFFFFFull;
    __syncthreads();
    for (int k = 2; k <= m; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < m; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long a = s[i], b = s[ixj];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { s[i] = b; s[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        keys[i] = (unsigned int)(s[i] >> 32);
        vals[i] = (unsigned int)(s[i] & 0xFFFFFFFFull);
    }
}

/* sorts in place (result ends in keys/vals); tmp_* and hist are scratch */
extern "C" cudaError_t rt_radix_sort_pairs(unsigned int* keys, unsigned int* vals, unsigned int* tmp_keys,
                                           unsigned int* tmp_vals, unsigned int* hist, int n, int force_radix,
                                           cudaStream_t stream) {
    if (n <= 1) return cudaSuccess;
    if (n <= SS_MAX && !force_radix) {
        rs_small_sort_kernel<<<1, 1024, 0, stream>>>(keys, vals, n);
        RT_LAUNCHED(1);
        return cudaGetLastError();
    }
    const int n_blocks = (n + RS_TILE - 1) / RS_TILE;
    unsigned int *ki = keys, *vi = vals, *ko = tmp_keys, *vo = tmp_vals;
    for (int pass = 0; pass < 4; pass++) {
        const int shift = pass * 8;
        rs_hist_kernel<<<n_blocks, RS_THREADS, 0, stream>>>(ki, n, shift, hist, n_blocks);
        rs_scan_kernel<<<256 * RS_SCAN_WARPS / 8, 256, 0, stream>>>(hist, n_blocks);
        rs_scatter_kernel<<<n_blocks, RS_THREADS, 0, stream>>>(ki, vi, ko, vo, n, shift, hist, n_blocks);
        RT_LAUNCHED(3);
        unsigned int* t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
    }
    return cudaGetLastError();   /* 4 passes: result is back in keys/vals */
}

extern "C" int rt_radix_hist_entries(int n) { return 256 * ((n + RS_TILE - 1) / RS_TILE) + 256; }   /* rows + bin totals */

/* ---- the same sort with its scratch in a file-scope device array: allocates nothing (rt_sort_pairs_static, api.cu).
 * Layout in uints: [tmp keys, RT_SORT_STATIC_MAX][tmp values, RT_SORT_STATIC_MAX][histogram rows + bin totals].
 * 8.5 MiB that every GPU which loads the library holds until the process exits.  Nothing in it has to start at
 * zero: each of the four passes writes every histogram row, every bin total and all n temporary keys / values before
 * it reads them, so the previous call's contents are never seen and no clear is needed.  All calls share the array:
 * two sorts that overlap on one GPU (different streams, different contexts, different threads) race on it. */
#define RS_STATIC_HIST (256 * ((RT_SORT_STATIC_MAX + RS_TILE - 1) / RS_TILE) + 256)
static __device__ __align__(256) unsigned int g_sort_scratch[2 * RT_SORT_STATIC_MAX + RS_STATIC_HIST];

/* `g_sort_scratch` in HOST code is a host-side stand-in for the symbol, not a device address: the address a kernel
 * can use comes from cudaGetSymbolAddress, looked up once per device with that device current. */
#define RS_STATIC_MAX_DEVICES 64
static unsigned int* g_sort_scratch_dev[RS_STATIC_MAX_DEVICES];

/* device = the CURRENT device (the caller has made it so).  *why gets a static message on failure. */
extern "C" cudaError_t rt_radix_sort_pairs_static(unsigned int* keys, unsigned int* vals, int n, int device,
                                                  cudaStream_t stream, const char** why) {
    *why = "";
    if (n <= 1) return cudaSuccess;
    if (n > RT_SORT_STATIC_MAX) {
        *why = "n exceeds RT_SORT_STATIC_MAX, the capacity of the static scratch array";
        return cudaErrorInvalidValue;
    }
    if (device < 0 || device >= RS_STATIC_MAX_DEVICES) {
        *why = "device index beyond the per-device address table";
        return cudaErrorInvalidDevice;
    }
    if (!g_sort_scratch_dev[device]) {
        unsigned int* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_sort_scratch);
        if (e != cudaSuccess || !p) {
            *why = "cudaGetSymbolAddress(g_sort_scratch) failed";
            return e != cudaSuccess ? e : cudaErrorInvalidSymbol;
        }
        g_sort_scratch_dev[device] = p;
    }
    unsigned int* const base = g_sort_scratch_dev[device];
    const cudaError_t e = rt_radix_sort_pairs(keys, vals, base, base + RT_SORT_STATIC_MAX, base + 2 * RT_SORT_STATIC_MAX, n,
                                              1 /* always the multi-pass radix sort: the variant that uses scratch */, stream);
    if (e != cudaSuccess) *why = "sort kernel launch failed";
    return e;
}

/* ------------------------------------------------------------------ hierarchy */

/* common-prefix length of sorted keys i and j; index breaks ties between equal codes */
__device__ __forceinline__ int rt_delta(const unsigned int* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned int a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((unsigned int)i ^ (unsigned int)j);
    return __clz(a ^ b);
}

/* children: >=0 internal node, <0 leaf slot (~c).  parent of internal node i in node_parent[i],
 * of leaf slot j in leaf_parent[j].  Root is internal node 0. */
__global__ void rt_karras_kernel(const unsigned int* __restrict__ keys, int n, int2* __restrict__ children,
                                 int* __restrict__ node_parent, int* __restrict__ leaf_parent,
                                 int* __restrict__ node_other) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    if (i == 0) node_parent[0] = -1;
    const int d = (rt_delta(keys, n, i, i + 1) - rt_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = rt_delta(keys, n, i, i - d);
    /* (Probing two steps of each search per round trip -- 3 or 4 independent loads, decisions in order -- was
     * measured: 45 instead of 33 us for a million leaves; the kernel is bound by the probes of ALL nodes, not by the
     * long searches near the root.) */
    int lmax = 2;
    while (rt_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (rt_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = rt_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (rt_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    node_other[i] = j;                       /* node i covers the sorted leaves [min(i, j), max(i, j)] */
    int c0, c1;
    if (lo == gamma) { c0 = ~gamma; leaf_parent[gamma] = i; }
    else             { c0 = gamma; node_parent[gamma] = i; }
    if (hi == gamma + 1) { c1 = ~(gamma + 1); leaf_parent[gamma + 1] = i; }
    else                 { c1 = gamma + 1; node_parent[gamma + 1] = i; }
    children[i] = make_int2(c0, c1);
}

/* A child's box goes to two places: exact (lo,hi) into the build-side `aux` array, from which the sibling
 * forms the union, and centre / half-extent into the traversal node record (rt_device.h), rounded outward
 * so that [c-e, c+e] contains [lo,hi].  BLOCK: aux lives in shared memory (single-CTA refit). */
template <bool BLOCK>
__device__ __forceinline__ void rt_store_child_box(float* __restrict__ node, float* aux, int which,
                                                   const float3& lo, const float3& hi) {
    float* a = aux + 6 * which;
    if (BLOCK) {
        a[0] = lo.x; a[1] = lo.y; a[2] = lo.z; a[3] = hi.x; a[4] = hi.y; a[5] = hi.z;
    } else {
        __stcg(a + 0, lo.x); __stcg(a + 1, lo.y); __stcg(a + 2, lo.z);
        __stcg(a + 3, hi.x); __stcg(a + 4, hi.y); __stcg(a + 5, hi.z);
    }
    const float grow = 1.0f + 4.76837158203125e-07f;   /* 1 + 2^-21 */
    const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    const float ex = fmaxf(hi.x - cx, cx - lo.x) * grow, ey = fmaxf(hi.y - cy, cy - lo.y) * grow,
                ez = fmaxf(hi.z - cz, cz - lo.z) * grow;
    float* r = node + 6 * which;
    __stcg(r + 0, cx); __stcg(r + 1, cy); __stcg(r + 2, cz);
    __stcg(r + 3, ex); __stcg(r + 4, ey); __stcg(r + 5, ez);
}

/* Leaf slot j: copy its record into Morton order, pad its box and carry unions up while this thread is the
 * second arrival at each node.  BLOCK: flags / aux in shared memory, block-scope fences (single CTA). */
template <bool BLOCK>
__device__ __forceinline__ void rt_refit_leaf(int j, int n, const unsigned int* __restrict__ sorted_vals,
                                              const int* __restrict__ bvh_src, const float4* recs,
                                              const float4* prim_lo, const float4* prim_hi, float extent, float maxabs,
                                              const int2* __restrict__ children, const int* __restrict__ node_parent,
                                              const int* __restrict__ leaf_parent, unsigned int* flags,
                                              float* __restrict__ nodes, float* node_aux, float4* __restrict__ prims,
                                              int* __restrict__ depth_out) {
    const int k = (int)sorted_vals[j];
    const int obj = bvh_src[k];
    prims[3 * j] = recs[3 * obj];
    prims[3 * j + 1] = recs[3 * obj + 1];
    prims[3 * j + 2] = recs[3 * obj + 2];
    if (n < 2) {
        *depth_out = 0;
        return;
    }
    const float4 plo = prim_lo[k], phi = prim_hi[k];
    const float global_pad = 9.5367431640625e-07f * fmaxf(extent, maxabs);   /* 2^-20 */
    const float kRel = 3.0517578125e-05f;                                     /* 2^-15 */
    float3 lo, hi;
    {
        const float px = kRel * fmaxf(fabsf(plo.x), fabsf(phi.x)) + global_pad;
        const float py = kRel * fmaxf(fabsf(plo.y), fabsf(phi.y)) + global_pad;
        const float pz = kRel * fmaxf(fabsf(plo.z), fabsf(phi.z)) + global_pad;
        lo = make_float3(plo.x - px, plo.y - py, plo.z - pz);
        hi = make_float3(phi.x + px, phi.y + py, phi.z + pz);
    }
    int child = ~j;
    int node = leaf_parent[j];
    unsigned int height = 0;                 /* of the subtree being carried up (leaf = 0) */
    while (node >= 0) {
        const int2 ch = children[node];
        const int which = (ch.x == child) ? 0 : 1;
        float* rec = nodes + 16 * (size_t)node;
        float* aux = node_aux + 12 * (size_t)node;
        rt_store_child_box<BLOCK>(rec, aux, which, lo, hi);
        if (which == 0) __stcg(rec + 12, __int_as_float(ch.x));
        else            __stcg(rec + 13, __int_as_float(ch.y));
        if (BLOCK) __threadfence_block(); else __threadfence();
        /* bit 0..7: arrivals, bits 8..: height of the first arrival's subtree */
        const unsigned int old = atomicAdd(&flags[node], 1u + (height << 8));
        if (old == 0u) return;                           /* first arrival: the sibling will carry on */
        height = max(height, old >> 8) + 1u;
        /* second arrival: the sibling's box is visible (written before its fence + atomic) */
        const float* oa = aux + 6 * (which ^ 1);
        float olx, oly, olz, ohx, ohy, ohz;
        if (BLOCK) {
            const volatile float* va = oa;
            olx = va[0]; oly = va[1]; olz = va[2]; ohx = va[3]; ohy = va[4]; ohz = va[5];
        } else {
            olx = __ldcg(oa + 0); oly = __ldcg(oa + 1); olz = __ldcg(oa + 2);
            ohx = __ldcg(oa + 3); ohy = __ldcg(oa + 4); ohz = __ldcg(oa + 5);
        }
        lo = make_float3(fminf(lo.x, olx), fminf(lo.y, oly), fminf(lo.z, olz));
        hi = make_float3(fmaxf(hi.x, ohx), fmaxf(hi.y, ohy), fmaxf(hi.z, ohz));
        child = node;
        node = node_parent[node];
    }
    *depth_out = (int)height;                /* only the thread that reached the root gets here */
}

/* 128-bit exchange with global memory, and a 128-bit atomic read (compare-and-swap of a value with itself) */
__device__ __forceinline__ uint4 rt_exch128(uint4* p, const uint4 v) {
    uint4 o;
    asm volatile("{\n.reg .b128 a, b;\nmov.b128 a, {%4, %5, %6, %7};\natom.global.exch.b128 b, [%8], a;\n"
                 "mov.b128 {%0, %1, %2, %3}, b;\n}"
                 : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w) : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(p) : "memory");
    return o;
}
__device__ __forceinline__ uint4 rt_read128(uint4* p) {
    uint4 o;
    asm volatile("{\n.reg .b128 z, b;\nmov.b128 z, {%4, %4, %4, %4};\natom.global.cas.b128 b, [%5], z, z;\n"
                 "mov.b128 {%0, %1, %2, %3}, b;\n}"
                 : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w) : "r"(0u), "l"(p) : "memory");
    return o;
}

/* Full-size refit.  A Karras node covers a contiguous range of sorted leaves (node_other: its other end), so a
 * block of RT_REFIT_THREADS consecutive leaves owns every node whose range lies inside its own: those (255 of 256)
 * are joined through shared memory -- both arrivals park their box there, block-scope fence, shared atomic --
 * and only the nodes that straddle blocks take the global protocol.  There the child's box TRAVELS IN THE ARRIVAL
 * itself: a node has two 16-byte slots in `node_aux`, (lo.xyz, stamp) and (hi.xyz, stamp); an arrival exchanges its
 * halves into them with two 128-bit atomics.  Whoever finds slot 0 without this refit's stamp is first and leaves; the
 * second has the sibling's lo in hand, and its hi as soon as the sibling's second exchange has landed (it normally
 * has: else the slot is re-read until it carries the sibling's stamp).  No fence, no flag word, no re-read of the
 * sibling's box through memory -- one round trip per level where the fence + counter + six loads made three (the
 * round-1 protocol, which held the ~16,000 straddling nodes' serial chain at 2.6 us per level).  stamp = epoch << 12 |
 * height << 1 | which child; older epochs count as empty (RtBuildBuffers.refit_epoch).  node_other[i] < 0 (PLOC
 * topology: ranges unknown) sends a node to the global protocol. */
#ifndef RT_REFIT_THREADS
#define RT_REFIT_THREADS 256
#endif
__global__ void __launch_bounds__(RT_REFIT_THREADS)
rt_refit_kernel(const unsigned int* __restrict__ sorted_vals, int n,
                const int* __restrict__ bvh_src, const float4* __restrict__ recs,
                const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi,
                const unsigned int* __restrict__ bounds, const int2* __restrict__ children,
                const int* __restrict__ node_parent, const int* __restrict__ leaf_parent,
                const int* __restrict__ node_other, unsigned int* __restrict__ flags, float* __restrict__ nodes,
                float* __restrict__ node_aux, float4* __restrict__ prims, int* __restrict__ depth_out,
                unsigned int epoch) {
    __shared__ unsigned int s_flags[RT_REFIT_THREADS];
    __shared__ float s_box[RT_REFIT_THREADS][12];    /* per block-owned node: both children's exact (lo, hi) */
    const int b0 = blockIdx.x * blockDim.x, b1 = min(b0 + (int)blockDim.x, n);
    const int j = b0 + threadIdx.x;
    s_flags[threadIdx.x] = 0u;
    __syncthreads();
    if (j >= n) return;
    const float extent = fmaxf(fmaxf(ord2f(bounds[3]) - ord2f(bounds[0]), ord2f(bounds[4]) - ord2f(bounds[1])),
                               ord2f(bounds[5]) - ord2f(bounds[2]));
    const float maxabs = ord2f(bounds[6]);
    const int k = (int)sorted_vals[j];
    const int obj = bvh_src[k];
    prims[3 * j] = recs[3 * obj];
    prims[3 * j + 1] = recs[3 * obj + 1];
    prims[3 * j + 2] = recs[3 * obj + 2];
    if (n < 2) {
        *depth_out = 0;
        return;
    }
    const float4 plo = prim_lo[k], phi = prim_hi[k];
    const float global_pad = 9.5367431640625e-07f * fmaxf(extent, maxabs);   /* 2^-20 */
    const float kRel = 3.0517578125e-05f;                                     /* 2^-15 */
    float3 lo, hi;
    {
        const float px = kRel * fmaxf(fabsf(plo.x), fabsf(phi.x)) + global_pad;
        const float py = kRel * fmaxf(fabsf(plo.y), fabsf(phi.y)) + global_pad;
        const float pz = kRel * fmaxf(fabsf(plo.z), fabsf(phi.z)) + global_pad;
        lo = make_float3(plo.x - px, plo.y - py, plo.z - pz);
        hi = make_float3(phi.x + px, phi.y + py, phi.z + pz);
    }
    int child = ~j;
    int node = leaf_parent[j];
    unsigned int height = 0;                 /* of the subtree being carried up (leaf = 0) */
    while (node >= 0) {
        /* the three topology words of this level travel together (one L2 round trip per level, not two) */
        const int2 ch = __ldg(children + node);
        const int other = __ldg(node_other + node);
        const int parent = __ldg(node_parent + node);
        const int which = (ch.x == child) ? 0 : 1;
        const bool local = other >= 0 && min(node, other) >= b0 && max(node, other) < b1;
        float* rec = nodes + 16 * (size_t)node;
        if (local) {
            /* traversal record of this child: centre / half-extent rounded outward (rt_store_child_box) + link;
             * plain stores, read by nobody before the kernel ends */
            const float grow = 1.0f + 4.76837158203125e-07f;   /* 1 + 2^-21 */
            const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
            const float ex = fmaxf(hi.x - cx, cx - lo.x) * grow, ey = fmaxf(hi.y - cy, cy - lo.y) * grow,
                        ez = fmaxf(hi.z - cz, cz - lo.z) * grow;
            float2* r2 = (float2*)(rec + 6 * which);             /* 24-byte offset: 8-byte aligned */
            r2[0] = make_float2(cx, cy); r2[1] = make_float2(cz, ex); r2[2] = make_float2(ey, ez);
            rec[12 + which] = __int_as_float(which == 0 ? ch.x : ch.y);
            const int ln = node - b0;        /* a node is an end of its own range: inside the block too */
            volatile float* sb = s_box[ln] + 6 * which;
            sb[0] = lo.x; sb[1] = lo.y; sb[2] = lo.z; sb[3] = hi.x; sb[4] = hi.y; sb[5] = hi.z;
            __threadfence_block();
            /* bit 0..7: arrivals, bits 8..: height of the first arrival's subtree */
            const unsigned int old = atomicAdd(&s_flags[ln], 1u + (height << 8));
            if (old == 0u) return;                           /* first arrival: the sibling will carry on */
            height = max(height, old >> 8) + 1u;
            volatile const float* ob = s_box[ln] + 6 * (which ^ 1);
            lo = make_float3(fminf(lo.x, ob[0]), fminf(lo.y, ob[1]), fminf(lo.z, ob[2]));
            hi = make_float3(fmaxf(hi.x, ob[3]), fmaxf(hi.y, ob[4]), fmaxf(hi.z, ob[5]));
        } else {
            /* traversal record of this child, as above (plain stores, read by nobody before the kernel ends) */
            const float grow = 1.0f + 4.76837158203125e-07f;   /* 1 + 2^-21 */
            const float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
            const float ex = fmaxf(hi.x - cx, cx - lo.x) * grow, ey = fmaxf(hi.y - cy, cy - lo.y) * grow,
                        ez = fmaxf(hi.z - cz, cz - lo.z) * grow;
            float2* r2 = (float2*)(rec + 6 * which);
            r2[0] = make_float2(cx, cy); r2[1] = make_float2(cz, ex); r2[2] = make_float2(ey, ez);
            rec[12 + which] = __int_as_float(which == 0 ? ch.x : ch.y);
            uint4* slot = (uint4*)(node_aux + 12 * (size_t)node);      /* 48-byte stride: 16-byte aligned */
            const unsigned int stamp = (epoch << 12) | (height << 1) | (unsigned int)which;
            const uint4 o0 = rt_exch128(slot, make_uint4(__float_as_uint(lo.x), __float_as_uint(lo.y), __float_as_uint(lo.z), stamp));
            uint4 o1 = rt_exch128(slot + 1, make_uint4(__float_as_uint(hi.x), __float_as_uint(hi.y), __float_as_uint(hi.z), stamp));
            if ((o0.w >> 12) != epoch) return;               /* first arrival: the sibling will carry on */
            while ((o1.w >> 12) != epoch || (o1.w & 1u) == (unsigned int)which) o1 = rt_read128(slot + 1);
            height = max(height, (o0.w >> 1) & 0x7FFu) + 1u;
            lo = make_float3(fminf(lo.x, __uint_as_float(o0.x)), fminf(lo.y, __uint_as_float(o0.y)), fminf(lo.z, __uint_as_float(o0.z)));
            hi = make_float3(fmaxf(hi.x, __uint_as_float(o1.x)), fmaxf(hi.y, __uint_as_float(o1.y)), fmaxf(hi.z, __uint_as_float(o1.z)));
        }
        child = node;
        node = parent;
    }
    *depth_out = (int)height;                /* only the thread that reached the root gets here */
}

/* Small scenes (<= RT_FUSED_MAX objects), topology unchanged: bounds + setup + refit in ONE single-CTA launch,
 * arrival counters and the union scratch in shared memory (the per-frame path of launchKernel: the
 * reference re-uploads its 133 objects every frame, kernel.cu:448). */
#define RT_FUSED_MAX 1024
__device__ __forceinline__ void
rt_fused_refit_body(const float* objs, int n_objects, const int* __restrict__ plane_idx, int n_planes,
                    float4* mats, float4* recs, float4* planes, int* bvh_src, float4* prim_lo, float4* prim_hi,
                    unsigned int* __restrict__ bounds, const unsigned int* __restrict__ sorted_vals, int n,
                    const int2* __restrict__ children, const int* __restrict__ node_parent,
                    const int* __restrict__ leaf_parent, float* __restrict__ nodes, float4* __restrict__ prims,
                    int* __restrict__ depth_out, unsigned char* fused_smem) {
    __shared__ unsigned int s_bounds[8];
    unsigned int* s_flags = (unsigned int*)fused_smem;                   /* [n-1] */
    float* s_aux = (float*)(s_flags + (n > 1 ? n - 1 : 0));             /* [n-1][12] */
    const int i = threadIdx.x;
    if (i < 3) s_bounds[i] = 0xFFFFFFFFu;
    else if (i < 8) s_bounds[i] = 0u;
    for (int k = i; k < n - 1; k += blockDim.x) s_flags[k] = 0u;
    __syncthreads();
    unsigned int bmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, bmax[4] = {0u, 0u, 0u, 0u};
    if (i < n_objects) rt_setup_object(i, objs + (size_t)i * 17, plane_idx, n_planes, mats, recs, planes, bvh_src, prim_lo, prim_hi, bmin, bmax);
    rt_merge_bounds(bmin, bmax, s_bounds);
    __syncthreads();                         /* recs / prim boxes written above are visible to the whole CTA */
    if (i < 8) bounds[i] = s_bounds[i];
    if (i < n) {
        const float extent = fmaxf(fmaxf(ord2f(s_bounds[3]) - ord2f(s_bounds[0]), ord2f(s_bounds[4]) - ord2f(s_bounds[1])),
                                   ord2f(s_bounds[5]) - ord2f(s_bounds[2]));
        rt_refit_leaf<true>(i, n, sorted_vals, bvh_src, recs, prim_lo, prim_hi, extent, ord2f(s_bounds[6]), children,
                            node_parent, leaf_parent, s_flags, nodes, s_aux, prims, depth_out);
    }
}

__global__ void __launch_bounds__(RT_FUSED_MAX)
rt_fused_refit_kernel(const float* __restrict__ objs, int n_objects, const int* __restrict__ plane_idx, int n_planes,
                      float4* mats, float4* recs, float4* planes, int* bvh_src, float4* prim_lo, float4* prim_hi,
                      unsigned int* __restrict__ bounds, const unsigned int* __restrict__ sorted_vals, int n,
                      const int2* __restrict__ children, const int* __restrict__ node_parent,
                      const int* __restrict__ leaf_parent, float* __restrict__ nodes, float4* __restrict__ prims,
                      int* __restrict__ depth_out) {
    extern __shared__ __align__(16) unsigned char fused_smem[];
    rt_fused_refit_body(objs, n_objects, plane_idx, n_planes, mats, recs, planes, bvh_src, prim_lo, prim_hi, bounds,
                        sorted_vals, n, children, node_parent, leaf_parent, nodes, prims, depth_out, fused_smem);
}

/* The same with the objects passed BY VALUE in the kernel parameters (up to RT_BYVALUE_MAX objects = 20.7 KB;
 * sm_70+ takes 32,764 parameter bytes): the per-frame path of launchKernel needs no host-to-device copy at all,
 * and because nothing then stands between the previous frame's last kernel and this one on the stream, it is
 * launched programmatically dependent: with wait_first = 0 (the host knows that every reader of the scene
 * buffers finished before that kernel started) it refits WHILE that kernel -- the previous frame's FXAA --
 * runs, and waits for it only at its end, so that the completion of this grid implies the completion of that
 * one for everything downstream. */
struct RtObjectsByValue {
    float v[17 * RT_BYVALUE_MAX];
};
__global__ void __launch_bounds__(RT_FUSED_MAX)
rt_fused_refit_byvalue_kernel(const __grid_constant__ RtObjectsByValue blk, int n_objects,
                              const int* __restrict__ plane_idx, int n_planes, float4* mats, float4* recs,
                              float4* planes, int* bvh_src, float4* prim_lo, float4* prim_hi,
                              unsigned int* __restrict__ bounds, const unsigned int* __restrict__ sorted_vals, int n,
                              const int2* __restrict__ children, const int* __restrict__ node_parent,
                              const int* __restrict__ leaf_parent, float* __restrict__ nodes,
                              float4* __restrict__ prims, int* __restrict__ depth_out, int wait_first) {
    extern __shared__ __align__(16) unsigned char fused_smem[];
    if (wait_first) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    /* parameter space -> shared memory (behind the refit scratch) */
    const size_t scratch = ((size_t)(n > 1 ? n - 1 : 0) * (sizeof(unsigned int) + 12 * sizeof(float)) + 15) & ~(size_t)15;
    float* s_objs = (float*)(fused_smem + scratch);
    for (int k = threadIdx.x; k < 17 * n_objects; k += blockDim.x) s_objs[k] = blk.v[k];
    __syncthreads();
    rt_fused_refit_body(s_objs, n_objects, plane_idx, n_planes, mats, recs, planes, bvh_src, prim_lo, prim_hi, bounds,
                        sorted_vals, n, children, node_parent, leaf_parent, nodes, prims, depth_out, fused_smem);
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

/* ------------------------------------------------------------------ PLOC hierarchy (small scenes)
 *
 * Parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton-sorted primitives:
 * every cluster looks RADIUS positions left and right for the neighbour whose union with it has the
 * smallest surface area; mutual nearest neighbours merge; survivors are compacted; repeat until one
 * cluster is left.  Unlike the Karras split (which only looks at the Morton keys) the merge order sees
 * the real boxes, which matters for the built-in scene: 440-unit mountain triangles next to 0.1-unit
 * spheres.  One CTA, everything in shared memory, deterministic node numbering (ids are handed out by a
 * prefix sum, descending, so the last merge -- the root -- is node 0).  Emits the same topology arrays
 * as rt_karras_kernel; rt_refit_kernel then fills the boxes. */
#define PLOC_MAX 1024
#ifndef PLOC_RADIUS
#define PLOC_RADIUS 40
#endif

__device__ __forceinline__ float ploc_area(const float* a, const float* b) {
    const float dx = fmaxf(a[3], b[3]) - fminf(a[0], b[0]);
    const float dy = fmaxf(a[4], b[4]) - fminf(a[1], b[1]);
    const float dz = fmaxf(a[5], b[5]) - fminf(a[2], b[2]);
    return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(PLOC_MAX)
rt_ploc_small_kernel(const unsigned int* __restrict__ sorted_vals, int n, const float4* __restrict__ prim_lo,
                     const float4* __restrict__ prim_hi, int2* __restrict__ children, int* __restrict__ node_parent,
                     int* __restrict__ leaf_parent) {
    __shared__ float s_box[PLOC_MAX][6];
    __shared__ int s_cid[PLOC_MAX];
    __shared__ int s_nn[PLOC_MAX];
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned int s_total;
    const int i = threadIdx.x, lane = i & 31, warp = i >> 5;
    if (i < n) {
        const int k = (int)sorted_vals[i];
        const float4 lo = prim_lo[k], hi = prim_hi[k];
        s_box[i][0] = lo.x; s_box[i][1] = lo.y; s_box[i][2] = lo.z;
        s_box[i][3] = hi.x; s_box[i][4] = hi.y; s_box[i][5] = hi.z;
        s_cid[i] = ~i;
    }
    if (i == 0) node_parent[0] = -1;
    int count = n, next_node = n - 2;
    __syncthreads();
    while (count > 1) {
        /* 1. nearest neighbour within the window */
        if (i < count) {
            float best = __int_as_float(0x7f800000);
            int bj = -1;
            const int j0 = max(0, i - PLOC_RADIUS), j1 = min(count - 1, i + PLOC_RADIUS);
            for (int j = j0; j <= j1; j++) {
                if (j == i) continue;
                const float a = ploc_area(s_box[i], s_box[j]);
                if (a < best) { best = a; bj = j; }
            }
            s_nn[i] = bj;
        }
        __syncthreads();
        /* 2. mutual pairs merge into the lower position; the upper one dies */
        bool alive = false, leader = false;
        int j = -1;
        if (i < count) {
            j = s_nn[i];
            const bool mutual = s_nn[j] == i;
            leader = mutual && i < j;
            alive = !mutual || leader;
        }
        /* 3. one block scan carries both counts: survivors (low 16 bits) and merges (high 16 bits) */
        const unsigned int v = (alive ? 1u : 0u) | (leader ? 0x10000u : 0u);
        unsigned int s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) s_warp[warp] = s;
        __syncthreads();
        if (warp == 0) {
            unsigned int ws = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, ws, o);
                if (lane >= o) ws += t;
            }
            s_warp[lane] = ws;
            if (lane == 31) s_total = ws;
        }
        __syncthreads();
        const unsigned int excl = (warp ? s_warp[warp - 1] : 0u) + s - v;
        const unsigned int total = s_total;
        /* 4. read my (possibly merged) cluster into registers, then write it to its compacted slot */
        float box[6];
        int cid = 0;
        if (alive) {
#pragma unroll
            for (int a = 0; a < 6; a++) box[a] = s_box[i][a];
            cid = s_cid[i];
            if (leader) {
                const int id = next_node - (int)(excl >> 16);
                const int cj = s_cid[j];
                children[id] = make_int2(cid, cj);
                if (cid < 0) leaf_parent[~cid] = id; else node_parent[cid] = id;
                if (cj < 0) leaf_parent[~cj] = id; else node_parent[cj] = id;
#pragma unroll
                for (int a = 0; a < 3; a++) {
                    box[a] = fminf(box[a], s_box[j][a]);
                    box[a + 3] = fmaxf(box[a + 3], s_box[j][a + 3]);
                }
                cid = id;
            }
        }
        __syncthreads();
        if (alive) {
            const int dst = (int)(excl & 0xFFFFu);
#pragma unroll
            for (int a = 0; a < 6; a++) s_box[dst][a] = box[a];
            s_cid[dst] = cid;
        }
        count = (int)(total & 0xFFFFu);
        next_node -= (int)(total >> 16);
        __syncthreads();
    }
}

/* mode 0: full build; mode 1: refit only (topology from the previous build).
 * builder 0: PLOC when n <= PLOC_MAX, else Karras; 1: Karras; 2: PLOC (falls back to Karras above PLOC_MAX) */
/* does rt_build_bvh(b, mode) take the objects from b->objs_host in the kernel parameters (no upload needed)? */
extern "C" int rt_build_by_value(const RtBuildBuffers* b, int mode) {
    const size_t scratch = (size_t)(b->n_bvh > 1 ? b->n_bvh - 1 : 0) * (sizeof(unsigned int) + 12 * sizeof(float));
    return mode == 1 && b->n_objects <= RT_BYVALUE_MAX && b->n_bvh >= 1 &&
           scratch + 16 + sizeof(float) * 17 * (size_t)b->n_objects <= 48 * 1024;
}

extern "C" cudaError_t rt_build_bvh(const RtBuildBuffers* b, int mode, int force_radix, int builder, cudaStream_t stream) {
    const int T = RT_BUILD_THREADS;
    const size_t fused_smem = (size_t)(b->n_bvh > 1 ? b->n_bvh - 1 : 0) * (sizeof(unsigned int) + 12 * sizeof(float));
    if (mode == 1 && b->n_objects <= RT_FUSED_MAX && b->n_bvh >= 1 && fused_smem <= 40 * 1024) {
        const int n = b->n_bvh;
        const size_t smem = fused_smem;
        if (b->objs_host && rt_build_by_value(b, mode)) {
            RtObjectsByValue blk;                /* copied into the launch by cudaLaunchKernelEx */
            memcpy(blk.v, b->objs_host, sizeof(float) * 17 * (size_t)b->n_objects);
            cudaLaunchAttribute pdl[1];
            pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            pdl[0].val.programmaticStreamSerializationAllowed = 1;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(1, 1, 1);
            cfg.blockDim = dim3(RT_FUSED_MAX, 1, 1);
            cfg.dynamicSmemBytes = ((smem + 15) & ~(size_t)15) + sizeof(float) * 17 * (size_t)b->n_objects;
            cfg.stream = stream;
            cfg.attrs = pdl;
            cfg.numAttrs = 1;
            RT_LAUNCHED(1);
            return cudaLaunchKernelEx(&cfg, rt_fused_refit_byvalue_kernel, blk, b->n_objects, (const int*)b->plane_idx,
                                      b->n_planes, b->mats, b->recs, b->planes, b->bvh_src, b->prim_lo, b->prim_hi,
                                      b->bounds, (const unsigned int*)b->vals, n, (const int2*)b->children,
                                      (const int*)b->node_parent, (const int*)b->leaf_parent, b->nodes, b->prims,
                                      b->depth, b->after_own_post ? 0 : 1);
        }
        rt_fused_refit_kernel<<<1, RT_FUSED_MAX, smem, stream>>>(b->objs, b->n_objects, b->plane_idx, b->n_planes, b->mats,
                                                               b->recs, b->planes, b->bvh_src, b->prim_lo, b->prim_hi,
                                                               b->bounds, b->vals, n, b->children, b->node_parent,
                                                               b->leaf_parent, b->nodes, b->prims, b->depth);
        RT_LAUNCHED(1);
        return cudaGetLastError();
    }
    rt_bounds_init_kernel<<<1, 32, 0, stream>>>(b->bounds);
    RT_LAUNCHED(1 + (b->n_objects > 0 ? 1 : 0));
    if (b->n_objects > 0)
        rt_setup_kernel<<<(b->n_objects + T - 1) / T, T, 0, stream>>>(b->objs, b->n_objects, b->plane_idx, b->n_planes,
                                                                      b->mats, b->recs, b->planes, b->bvh_src,
                                                                      b->prim_lo, b->prim_hi, b->bounds);
    const int n = b->n_bvh;
    if (n <= 0) return cudaGetLastError();
    if (mode == 0) {
        rt_morton_kernel<<<(n + T - 1) / T, T, 0, stream>>>(b->prim_lo, b->prim_hi, n, b->bounds, b->keys, b->vals);
        RT_LAUNCHED(1 + (n >= 2 ? 1 : 0));       /* Morton codes + hierarchy (PLOC or Karras) */
        cudaError_t e = rt_radix_sort_pairs(b->keys, b->vals, b->tmp_keys, b->tmp_vals, b->hist, n, force_radix, stream);
        if (e != cudaSuccess) return e;
        if (n >= 2) {
            if (builder != 1 && n <= PLOC_MAX) {
                /* PLOC nodes cover no contiguous leaf range: every join of the refit takes the global protocol */
                e = cudaMemsetAsync(b->node_other, 0xFF, sizeof(int) * (size_t)(n - 1), stream);
                if (e != cudaSuccess) return e;
                rt_ploc_small_kernel<<<1, PLOC_MAX, 0, stream>>>(b->vals, n, b->prim_lo, b->prim_hi, b->children,
                                                                 b->node_parent, b->leaf_parent);
            } else
                rt_karras_kernel<<<(n - 1 + T - 1) / T, T, 0, stream>>>(b->keys, n, b->children, b->node_parent,
                                                                        b->leaf_parent, b->node_other);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    RT_LAUNCHED(1);
    rt_refit_kernel<<<(n + RT_REFIT_THREADS - 1) / RT_REFIT_THREADS, RT_REFIT_THREADS, 0, stream>>>(b->vals, n, b->bvh_src, b->recs, b->prim_lo, b->prim_hi, b->bounds,
                                                      b->children, b->node_parent, b->leaf_parent, b->node_other, b->flags,
                                                      b->nodes, b->node_aux, b->prims, b->depth, b->refit_epoch);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------ 4-wide view of the hierarchy
 * Wide node i = binary node i with each internal child replaced by that child's two children: 2..4 boxes and links
 * in one 128-byte line (layout: rt_device.h, RT_WIDE_F4).  The boxes are copied bit for bit from the binary nodes, a
 * link >= 0 names a binary node = its wide node.  Only the wide nodes of the root, its grandchildren, their
 * grandchildren ... are ever visited; all n - 1 are written (one thread each, no dependency between them), which
 * costs memory, not time.  Re-run after every build or refit. */
__global__ void __launch_bounds__(256)
rt_collapse_wide_kernel(const float* __restrict__ nodes, int n_internal, float4* __restrict__ wide) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal) return;
    const float4* __restrict__ me = (const float4*)(nodes + 16 * (size_t)i);
    float4 q[4] = {__ldg(me), __ldg(me + 1), __ldg(me + 2), __ldg(me + 3)};
    const float* mine = (const float*)q;
    float box[24];
    int link[4] = {RT_NODE_SENTINEL, RT_NODE_SENTINEL, RT_NODE_SENTINEL, RT_NODE_SENTINEL};
#pragma unroll
    for (int k = 0; k < 24; k++) box[k] = 0.0f;
    int at = 0;
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const int l = __float_as_int(mine[12 + c]);
        if (l >= 0) {
            const float4* __restrict__ ch = (const float4*)(nodes + 16 * (size_t)l);
            float4 r[4] = {__ldg(ch), __ldg(ch + 1), __ldg(ch + 2), __ldg(ch + 3)};
            const float* g = (const float*)r;
#pragma unroll
            for (int k = 0; k < 12; k++) box[6 * at + k] = g[k];
            link[at] = __float_as_int(g[12]);
            link[at + 1] = __float_as_int(g[13]);
            at += 2;
        } else {
#pragma unroll
            for (int k = 0; k < 6; k++) box[6 * at + k] = mine[6 * c + k];
            link[at] = l;
            at += 1;
        }
    }
    float4* __restrict__ w = wide + (size_t)RT_WIDE_F4 * i;
#pragma unroll
    for (int k = 0; k < 6; k++) w[k] = make_float4(box[4 * k], box[4 * k + 1], box[4 * k + 2], box[4 * k + 3]);
    w[6] = make_float4(__int_as_float(link[0]), __int_as_float(link[1]), __int_as_float(link[2]), __int_as_float(link[3]));
    w[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

extern "C" cudaError_t rt_collapse_wide(const float* nodes, int n_prims, float4* wide, cudaStream_t stream) {
    if (n_prims < 2) return cudaSuccess;
    const int n_internal = n_prims - 1;
    RT_LAUNCHED(1);
    rt_collapse_wide_kernel<<<(n_internal + 255) / 256, 256, 0, stream>>>(nodes, n_internal, wide);
    return cudaGetLastError();
}

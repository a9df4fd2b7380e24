/*
 * light_grid.cu -- light bins for scenes that do not fit the shared-memory tracer (the HBM tracer, trace.cu).
 *
 * Same idea and the same conservative geometry as the light bins of small scenes (light_bins.cuh): every shadow
 * ray of light L (reference kernel.cu:177-193) lies on a line through L, so the primitives it can meet depend on
 * its direction seen from L only.  Per light, a cube map of RT_LGB_G x RT_LGB_G (2048 x 2048) direction bins per face; here the
 * lists are unbounded (CSR: offsets + entries) and the build is primitive-centric: one thread per primitive and
 * face descends a quadtree over the face -- whole face, quadrants, ... single bins -- with the cull test of
 * light_bins.cuh at every level (conservative at every level, so the leaves it reaches are a superset of the bins
 * whose lines can hit the primitive).  Two passes (count, fill) around an exclusive scan.
 *
 * One-sided triangles (`det < 0.001`, kernel.cu:104) can only occlude a light they face AWAY from; emissive
 * primitives never occlude (kernel.cu:189).  For a terrain under a high sun that leaves a few percent of the
 * triangles in the grid, and a shadow ray tests the handful in its bin instead of walking the LBVH.  The exact
 * leaf tests decide, so the image equals the traversal's bit for bit (tests/test_gpu_fxaa_shard_bvh.py).
 */
#include "light_bins.cuh"

#if RT_LGB_G == 2048
#define RT_LGB_LEVELS 11                     /* log2(RT_LGB_G) */
#elif RT_LGB_G == 1024
#define RT_LGB_LEVELS 10
#else
#define RT_LGB_LEVELS 9
#endif
static_assert((1 << RT_LGB_LEVELS) == RT_LGB_G, "RT_LGB_G must be 2^RT_LGB_LEVELS (512, 1024 or 2048)");
#define RT_LGB_MAX_BINS_PER_PRIM 16384       /* a primitive spread over more bins of one face: no grid for this light */

/* PASS 0: counts[bin]++; PASS 1: entries[offsets[bin] + cursor[bin]++] = slot.  `fail` is raised when a primitive
 * covers too many bins (it sits next to the light): the host then leaves this light to the traversal. */
template <int PASS>
__global__ void __launch_bounds__(256)
rt_lgb_bin_kernel(const RtSceneDev sc, float lx, float ly, float lz, unsigned int* __restrict__ counts,
                  const unsigned int* __restrict__ offsets, unsigned int* __restrict__ entries, unsigned int cap,
                  unsigned int* __restrict__ fail) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= sc.n_bvh_prims) return;
    const float lpos[3] = {lx, ly, lz};
    const vec3 lp = v3(lx, ly, lz);
    const float pad = rs_cull_pad(sc, lpos);
    for (int face = 0; face < 6; face++) {
        RsLgPrim p;
        if (!rs_lg_prim(face, lp, sc.prims + 3 * (size_t)j, pad, p)) break;     /* emissive / facing the light: no face */
        unsigned int stack[3 * RT_LGB_LEVELS + 4];
        int sp = 0;
        int emitted = 0;
        stack[sp++] = 0u;                                                        /* level 0, cell (0, 0) */
        while (sp > 0) {
            const unsigned int e = stack[--sp];
            const int level = (int)(e >> 24), cy = (int)((e >> 12) & 4095u), cx = (int)(e & 4095u);
            const int size = RT_LGB_G >> level;
            const RsLgRect r = rs_lg_rect<RT_LGB_G>((float)(cx * size), (float)(cy * size), (float)(cx * size + size),
                                                   (float)(cy * size + size));
            if (!rs_lg_may_hit(p, r, pad)) continue;
            if (level == RT_LGB_LEVELS) {
                const unsigned int bin = ((unsigned int)face * RT_LGB_G + (unsigned int)cy) * RT_LGB_G + (unsigned int)cx;
                if (PASS == 0) {
                    atomicAdd(&counts[bin], 1u);
                } else {
                    const unsigned int at = __ldg(offsets + bin) + atomicAdd(&counts[bin], 1u);
                    if (at < cap) entries[at] = (unsigned int)j;
                }
                if (++emitted > RT_LGB_MAX_BINS_PER_PRIM) { atomicOr(fail, 1u); return; }
                continue;
            }
            const unsigned int nl = (unsigned int)(level + 1) << 24;
            stack[sp++] = nl | ((unsigned int)(2 * cy) << 12) | (unsigned int)(2 * cx);
            stack[sp++] = nl | ((unsigned int)(2 * cy) << 12) | (unsigned int)(2 * cx + 1);
            stack[sp++] = nl | ((unsigned int)(2 * cy + 1) << 12) | (unsigned int)(2 * cx);
            stack[sp++] = nl | ((unsigned int)(2 * cy + 1) << 12) | (unsigned int)(2 * cx + 1);
        }
    }
}

/* ---- exclusive scan of n counters (n = 6 G^2 per light): block sums, scan of the sums, rescan with offsets */
#define LGB_SCAN_THREADS 256
#define LGB_SCAN_PER_THREAD 8
#define LGB_SCAN_TILE (LGB_SCAN_THREADS * LGB_SCAN_PER_THREAD)

__device__ __forceinline__ unsigned int lgb_block_exclusive(unsigned int v, unsigned int* s_warp, unsigned int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned int before = 0, all = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) {
        const unsigned int c = s_warp[w];
        if (w < warp) before += c;
        all += c;
    }
    __syncthreads();
    total = all;
    return before + incl - v;
}

__global__ void __launch_bounds__(LGB_SCAN_THREADS)
rt_lgb_sums_kernel(const unsigned int* __restrict__ counts, unsigned int n, unsigned int* __restrict__ sums) {
    __shared__ unsigned int s_warp[32];
    const unsigned int base = blockIdx.x * LGB_SCAN_TILE + threadIdx.x * LGB_SCAN_PER_THREAD;
    unsigned int v = 0;
#pragma unroll
    for (int k = 0; k < LGB_SCAN_PER_THREAD; k++) v += base + k < n ? counts[base + k] : 0u;
    unsigned int total;
    lgb_block_exclusive(v, s_warp, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

/* one block: exclusive scan of the block sums in place; sums[n_sums] = grand total */
__global__ void __launch_bounds__(1024) rt_lgb_scan_sums_kernel(unsigned int* __restrict__ sums, unsigned int n_sums) {
    __shared__ unsigned int s_warp[32];
    unsigned int carry = 0;
    for (unsigned int i0 = 0; i0 < n_sums; i0 += blockDim.x) {
        const unsigned int i = i0 + threadIdx.x;
        const unsigned int v = i < n_sums ? sums[i] : 0u;
        unsigned int total;
        const unsigned int ex = lgb_block_exclusive(v, s_warp, total);
        if (i < n_sums) sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[n_sums] = carry;
}

__global__ void __launch_bounds__(LGB_SCAN_THREADS)
rt_lgb_offsets_kernel(const unsigned int* __restrict__ counts, unsigned int n, const unsigned int* __restrict__ sums,
                      unsigned int n_sums, unsigned int* __restrict__ offsets) {
    __shared__ unsigned int s_warp[32];
    const unsigned int base = blockIdx.x * LGB_SCAN_TILE + threadIdx.x * LGB_SCAN_PER_THREAD;
    unsigned int c[LGB_SCAN_PER_THREAD], v = 0;
#pragma unroll
    for (int k = 0; k < LGB_SCAN_PER_THREAD; k++) { c[k] = base + k < n ? counts[base + k] : 0u; v += c[k]; }
    unsigned int total;
    unsigned int run = sums[blockIdx.x] + lgb_block_exclusive(v, s_warp, total);
#pragma unroll
    for (int k = 0; k < LGB_SCAN_PER_THREAD; k++) {
        if (base + k < n) offsets[base + k] = run;
        run += c[k];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) offsets[n] = sums[n_sums];
}

extern "C" size_t rt_lgb_bins(void) { return (size_t)6 * RT_LGB_G * RT_LGB_G; }
extern "C" size_t rt_lgb_sums(void) { return (rt_lgb_bins() + LGB_SCAN_TILE - 1) / LGB_SCAN_TILE + 1; }

/* counts: [bins], offsets: [bins + 1], sums: [rt_lgb_sums()], fail: 1 uint.  Count pass + scan; the grand total ends
 * in offsets[bins]. */
extern "C" cudaError_t rt_lgb_count(const RtSceneDev* sc, const float* light_pos, unsigned int* counts,
                                    unsigned int* offsets, unsigned int* sums, unsigned int* fail, cudaStream_t stream) {
    const unsigned int bins = (unsigned int)rt_lgb_bins();
    const unsigned int n_sums = (bins + LGB_SCAN_TILE - 1) / LGB_SCAN_TILE;
    cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(unsigned int) * bins, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(fail, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const int n = sc->n_bvh_prims;
    rt_lgb_bin_kernel<0><<<(n + 255) / 256, 256, 0, stream>>>(*sc, light_pos[0], light_pos[1], light_pos[2], counts,
                                                            nullptr, nullptr, 0u, fail);
    rt_lgb_sums_kernel<<<n_sums, LGB_SCAN_THREADS, 0, stream>>>(counts, bins, sums);
    rt_lgb_scan_sums_kernel<<<1, 1024, 0, stream>>>(sums, n_sums);
    rt_lgb_offsets_kernel<<<n_sums, LGB_SCAN_THREADS, 0, stream>>>(counts, bins, sums, n_sums, offsets);
    RT_LAUNCHED(4);
    return cudaGetLastError();
}

/* fill pass: counts is reused as the per-bin cursor */
extern "C" cudaError_t rt_lgb_fill(const RtSceneDev* sc, const float* light_pos, unsigned int* counts,
                                   const unsigned int* offsets, unsigned int* entries, unsigned int cap,
                                   unsigned int* fail, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(unsigned int) * rt_lgb_bins(), stream);
    if (e != cudaSuccess) return e;
    const int n = sc->n_bvh_prims;
    RT_LAUNCHED(1);
    rt_lgb_bin_kernel<1><<<(n + 255) / 256, 256, 0, stream>>>(*sc, light_pos[0], light_pos[1], light_pos[2], counts,
                                                            offsets, entries, cap, fail);
    return cudaGetLastError();
}

/* Verdict of one light's build, left ON THE DEVICE so that the host need not wait for it: state[0] = 1 when no
 * primitive overflowed its bin budget and all `total` entries fitted the `cap` the fill pass ran with (only then may
 * the tracer use this light's lists, rt_light_grid_occluded), state[1] = total, state[2] = fail. */
__global__ void rt_lgb_finish_kernel(const unsigned int* __restrict__ total, const unsigned int* __restrict__ fail,
                                     unsigned int cap, unsigned int* __restrict__ state) {
    const unsigned int t = *total, f = *fail;
    state[1] = t;
    state[2] = f;
    state[0] = (f == 0u && t <= cap) ? 1u : 0u;
}

extern "C" cudaError_t rt_lgb_finish(const unsigned int* total, const unsigned int* fail, unsigned int cap,
                                     unsigned int* state, cudaStream_t stream) {
    RT_LAUNCHED(1);
    rt_lgb_finish_kernel<<<1, 1, 0, stream>>>(total, fail, cap, state);
    return cudaGetLastError();
}

/*
 * fxaa.cu -- the FXAA post-pass (replaces the reference's `antialiasing` kernel,
 * reference kernel.cu:262-403).
 *
 * The filter is a pure function of each interior pixel's 3x3 neighbourhood in the
 * pre-FXAA image; border pixels pass through (kernel.cu:330, 399-402).  The
 * reference's out-of-bounds halo loads (operator-precedence slip at
 * kernel.cu:318-319) only ever feed border pixels, so a clean version needs none.
 *
 * HBM-bound: 4 B read + 4 B written per pixel.  Two implementations:
 *   rt_fxaa_tma_kernel  : 64x32-pixel tiles (+1 halo) staged into shared memory by TMA
 *                         (cp.async.bulk.tensor.2d, 72x34 box, zero OOB fill) on mbarriers;
 *                         luminance computed once per staged texel; coalesced 128-byte row
 *                         stores.  Two shapes: one tile per CTA (frames of up to 148 x 8
 *                         tiles: the whole grid is resident at once) or persistent CTAs
 *                         looping over their tiles with the next load in flight.  With a peer
 *                         frame as the output the kernel carries the hand-shake with the
 *                         frame's owner itself (RtPeerSync).
 *   rt_fxaa_smem_kernel : plain shared-memory tile, any width, any shard tile size.
 *
 * Arithmetic is spelled with explicit rounding intrinsics in the fused shape the
 * rebuilt reference has in SASS (luminance = min(255, fma(b,c3, fma(r,c1, g*c2))) * (1/255)f,
 * blend = fma(1-b, centre, b*neighbour), smoothstep = (3-(f+f))*(f*f), ...), so the two
 * kernels and the reference agree bit for bit on identical input.
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_device.h"
#include "fxaa_core.cuh"
#include "peer_sync.cuh"

/* ------------------------------------------------------------------ plain shared-memory kernel */

#define FX_BW 32
#define FX_BH 8

/* one thread of the CTA waits for the owner's release of the peer buffer (RtPeerSync); everybody learns the outcome */
__device__ __forceinline__ bool fx_peer_acquire(const RtPeerSync& peer, bool leader) {
    __shared__ int s_peer_ok;
    if (leader)
        s_peer_ok = rt_peer_spin(peer.flags, RT_PEER_FLAG_RELEASED, peer.acquire_seq, peer.timeout_ms,
                                 0x100u + (unsigned int)peer.rank) ? 1 : 0;
    __syncthreads();
    return s_peer_ok != 0;
}

__global__ void __launch_bounds__(FX_BW * FX_BH)
rt_fxaa_smem_kernel(const unsigned int* __restrict__ in, unsigned int* __restrict__ out, int w, int h,
                    int tile_w, int tile_h, int n_shards, int shard_id, int stride, const RtPeerSync peer) {
    __shared__ unsigned int s_px[FX_BH + 2][FX_BW + 2];
    __shared__ float s_lum[FX_BH + 2][FX_BW + 2];
    const int x0 = blockIdx.x * FX_BW - 1, y0 = blockIdx.y * FX_BH - 1;
    const bool leader = threadIdx.x == 0 && threadIdx.y == 0;
    const unsigned int n_ctas = gridDim.x * gridDim.y;
    if (n_shards > 1) {
        /* blocks that own no pixel of this shard leave before touching memory */
        const int px = blockIdx.x * FX_BW + threadIdx.x, py = blockIdx.y * FX_BH + threadIdx.y;
        const bool mine = px < w && py < h && rt_tile_owner(px, py, tile_w, tile_h, n_shards, stride) == shard_id;
        if (!__syncthreads_or(mine)) {
            if (peer.flags) rt_peer_publish(peer, n_ctas, leader);
            return;
        }
    }
    for (int i = threadIdx.y * FX_BW + threadIdx.x; i < (FX_BH + 2) * (FX_BW + 2); i += FX_BW * FX_BH) {
        const int ly = i / (FX_BW + 2), lx = i % (FX_BW + 2);
        const int gx = x0 + lx, gy = y0 + ly;
        unsigned int c = 0;
        if (gx >= 0 && gx < w && gy >= 0 && gy < h) c = __ldg(in + (size_t)gy * w + gx);
        s_px[ly][lx] = c;
        s_lum[ly][lx] = fx_lum(c);
    }
    __syncthreads();
    const bool may_store = peer.flags ? fx_peer_acquire(peer, leader) : true;
    const int x = blockIdx.x * FX_BW + threadIdx.x, y = blockIdx.y * FX_BH + threadIdx.y;
    if (may_store && x < w && y < h && (n_shards <= 1 || rt_tile_owner(x, y, tile_w, tile_h, n_shards, stride) == shard_id)) {
        const int lx = threadIdx.x + 1, ly = threadIdx.y + 1;
        const unsigned int centre = s_px[ly][lx];
        unsigned int result = centre;
        if (x > 0 && y > 0 && x + 1 < w && y + 1 < h) {
            float b;
            int dx, dy;
            if (fx_decide(s_lum[ly][lx], s_lum[ly][lx + 1], s_lum[ly][lx - 1], s_lum[ly - 1][lx], s_lum[ly + 1][lx],
                          s_lum[ly - 1][lx + 1], s_lum[ly - 1][lx - 1], s_lum[ly + 1][lx + 1], s_lum[ly + 1][lx - 1],
                          b, dx, dy))
                result = fx_blend(centre, s_px[ly + dy][lx + dx], b);
        }
        out[(size_t)y * w + x] = result;
    }
    if (peer.flags) rt_peer_publish(peer, n_ctas, leader);
}

/* ------------------------------------------------------------------ TMA-staged persistent kernel */

#define FX_TW 64
#define FX_TH 32
/* TMA needs the box's inner start coordinate at a 16-byte boundary (measured on this B200: a start at
 * x0-1 raises "illegal instruction", x0-4 works; negative / past-the-end coordinates are zero-filled), so
 * the halo column is fetched as a 4-pixel group on each side: 4 + 64 + 4 = 72 pixels per row. */
#define FX_HALO_X 4
#define FX_BOXW (FX_TW + 2 * FX_HALO_X)
#define FX_BOXH 34
#define FX_STAGE_BYTES (FX_BOXW * FX_BOXH * 4)
#define FX_STAGE_WORDS 2464   /* 9792 B rounded up to a multiple of 128 B: every TMA destination is 128-byte aligned */
#define FX_THREADS 256
/* Two shapes of the same kernel.  FX_CTAS_WAVE CTAs per SM (32 registers, one stage): frames whose tiles all fit in
 * ONE wave (1920x1080 = 1,020 tiles on 1,184 slots) -- one tile per CTA, no loop; every CTA is resident with its
 * barrier initialised while the tracer's tail still runs (programmatic dependent launch).  FX_CTAS_LOOP CTAs per SM
 * (64 registers, two stages): larger frames, each CTA loops over its tiles with the next load in flight (faster at
 * 3840x2160: 25.0 vs 27.0 us).  profiles/r02_ab_fxaa_residency.txt has the measurements, and how they led to the
 * shared-memory carve-out (rt_fxaa_init), which is where the time was. */
#ifndef FX_CTAS_WAVE
#define FX_CTAS_WAVE 8
#endif
#ifndef FX_WAVE_STAGES
#define FX_WAVE_STAGES 1
#endif
#define FX_CTAS_LOOP 4

__device__ __forceinline__ uint32_t fx_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void fx_mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(fx_smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void fx_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fx_smem_addr(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void fx_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(fx_smem_addr(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void fx_tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
            "r"(fx_smem_addr(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(fx_smem_addr(bar))
        : "memory");
}

/* FXAA tile k of this shard -> tile coordinates.  Whole frame: row-major over the 64x32 FXAA tiles.  Shards: the
 * shard's screen tiles are kx x ky FXAA tiles each (tile_w = 64 kx, tile_h = 32 ky); slot k is FXAA tile
 * k % (kx ky) of the shard's screen tile k / (kx ky).  Returns false past the last slot; a slot outside the
 * frame comes back with tx >= tiles_x (the caller skips it). */
__device__ __forceinline__ bool fx_tile_origin(int k, int tiles_x, int tiles_y, int n_shards, int shard_id,
                                               int stride, int kx, int ky, int& tx, int& ty) {
    if (n_shards <= 1) {
        if (k >= tiles_x * tiles_y) return false;
        tx = k % tiles_x;
        ty = k / tiles_x;
        return true;
    }
    /* screen-tile slots: per tile row, ceil(stiles_x / n) candidates i = i0 + m*n with i0 = (shard - j*stride) mod n */
    const int stiles_x = (tiles_x + kx - 1) / kx, stiles_y = (tiles_y + ky - 1) / ky;
    const int per_row = (stiles_x + n_shards - 1) / n_shards;
    const int sub = k % (kx * ky), ks = k / (kx * ky);
    if (ks >= per_row * stiles_y) return false;
    const int sty = ks / per_row;
    const int m = ks % per_row;
    int i0 = (shard_id - sty * stride) % n_shards;
    if (i0 < 0) i0 += n_shards;
    const int stx = i0 + m * n_shards;
    tx = stx * kx + sub % kx;
    ty = sty * ky + sub / kx;
    if (stx >= stiles_x || ty >= tiles_y) tx = tiles_x;
    return true;
}

/* the 64x32 pixels of tile (tx, ty) from the staged box: 256 threads, each walks 8 CONSECUTIVE rows of one
 * column with the 3x3 luminance window in registers (3 shared-memory loads per pixel instead of 9); a warp
 * still covers 32 adjacent columns of one row per step, so loads are conflict-free and stores coalesced */
template <bool INNER>
__device__ __forceinline__ void fx_filter_tile(const unsigned int* __restrict__ px, const float (*lum)[FX_BOXW],
                                               unsigned int* __restrict__ out, int tx, int ty, int w, int h, int tid) {
    const int cx = tid & 63, row0 = (tid >> 6) * (FX_TH / 4);
    const int x = tx * FX_TW + cx;
    const int lx = cx + FX_HALO_X;
    unsigned int* o = out + (size_t)(ty * FX_TH + row0) * w + x;
    float a0 = lum[row0][lx - 1], a1 = lum[row0][lx], a2 = lum[row0][lx + 1];              /* the row above */
    float b0 = lum[row0 + 1][lx - 1], b1 = lum[row0 + 1][lx], b2 = lum[row0 + 1][lx + 1];  /* this row */
#pragma unroll
    for (int pass = 0; pass < FX_TH / 4; pass++, o += w) {
        const int ly = row0 + pass + 1;
        const int y = ty * FX_TH + row0 + pass;
        const float c0 = lum[ly + 1][lx - 1], c1 = lum[ly + 1][lx], c2 = lum[ly + 1][lx + 1];   /* the row below */
        if (INNER || (x < w && y < h)) {
            const unsigned int centre = px[ly * FX_BOXW + lx];
            unsigned int result = centre;
            if (INNER || (x > 0 && y > 0 && x + 1 < w && y + 1 < h)) {
                float b;
                int dx, dy;
                if (fx_decide(b1, b2, b0, a1, c1, a2, a0, c2, c0, b, dx, dy))
                    result = fx_blend(centre, px[(ly + dy) * FX_BOXW + lx + dx], b);
            }
            *o = result;
        }
        a0 = b0; a1 = b1; a2 = b2;
        b0 = c0; b1 = c1; b2 = c2;
    }
}

#ifdef RS_TIMELINE
/* debug variant: [slot][0] first CTA entry  [1] first CTA past griddepcontrol.wait  [2] last CTA exit (see trace_small.cu) */
__device__ unsigned long long g_tl_fx[8][4];
__device__ unsigned int g_tl_fx_no, g_tl_fx_ticket;
__device__ __forceinline__ unsigned long long fx_tl_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
extern "C" cudaError_t rt_debug_timeline_fxaa(unsigned long long* out, int reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out, g_tl_fx, sizeof(g_tl_fx));
    if (e == cudaSuccess && reset) {
        unsigned long long init[8][4];
        for (int s = 0; s < 8; s++)
            for (int k = 0; k < 4; k++) init[s][k] = k < 2 ? ~0ull : 0ull;
        const unsigned int zero = 0;
        e = cudaMemcpyToSymbol(g_tl_fx, init, sizeof(init));
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_tl_fx_no, &zero, sizeof(zero));
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_tl_fx_ticket, &zero, sizeof(zero));
    }
    return e;
}
#endif

/* SHARD = false: the whole frame on one GPU (no shard arithmetic, no peer handshake: fewer live registers) */
template <bool SHARD, int CTAS>
__global__ void __launch_bounds__(FX_THREADS, CTAS)
rt_fxaa_tma_kernel(const __grid_constant__ CUtensorMap in_map, const unsigned int* __restrict__ in,
                   unsigned int* __restrict__ out, int w, int h, int n_shards, int shard_id, int stride, int kx,
                   int ky, const RtPeerSync peer) {
    constexpr bool LOOP = CTAS != FX_CTAS_WAVE;     /* one wave: grid == tiles, one tile per CTA, one stage */
    __shared__ __align__(128) unsigned int s_px[LOOP ? 2 : FX_WAVE_STAGES][FX_STAGE_WORDS];
    __shared__ float s_lum[FX_BOXH][FX_BOXW];
    __shared__ __align__(8) uint64_t s_bar[2];

    const int tiles_x = (w + FX_TW - 1) / FX_TW, tiles_y = (h + FX_TH - 1) / FX_TH;
    const int tid = threadIdx.x;
#ifdef RS_TIMELINE
    unsigned long long* const tl = g_tl_fx[*(volatile unsigned int*)&g_tl_fx_no & 7u];
    if (tid == 0) atomicMin(tl + 0, fx_tl_now());
#endif
    if (tid == 0) {
        fx_mbar_init(&s_bar[0], 1);
        fx_mbar_init(&s_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    /* programmatic dependent launch: the tracer has completed from here on; the next frame's pre-pass may start */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#ifdef RS_TIMELINE
    if (tid == 0) atomicMin(tl + 1, fx_tl_now());
#endif

    int k = blockIdx.x;
    int tx, ty;
    bool have = fx_tile_origin(k, tiles_x, tiles_y, SHARD ? n_shards : 1, shard_id, stride, kx, ky, tx, ty);
    if (tid == 0 && have && tx < tiles_x) {
        fx_mbar_expect_tx(&s_bar[0], FX_STAGE_BYTES);
        fx_tma_load_2d(&s_px[0][0], &in_map, tx * FX_TW - FX_HALO_X, ty * FX_TH - 1, &s_bar[0]);
    }
    /* peer frame: the owner must have released the buffer before the first store; the first tile is already on its
     * way into shared memory meanwhile.  A wait that gave up leaves the buffer alone (error word set). */
    const bool may_store = (SHARD && peer.flags) ? fx_peer_acquire(peer, tid == 0) : true;
    uint32_t phases = 0;   /* bit s = parity to wait for on stage s */
    int stage = 0;
    /* whole frames: the next tile by addition (k advances by gridDim.x), no division in the loop */
    const int step_x = (int)gridDim.x % tiles_x, step_y = (int)gridDim.x / tiles_x;
    while (have) {
        const int kn = k + gridDim.x;
        int ntx, nty;
        bool have_next;
        if (!LOOP) {
            ntx = nty = 0;
            have_next = false;
        } else if (!SHARD) {
            ntx = tx + step_x;
            nty = ty + step_y;
            if (ntx >= tiles_x) { ntx -= tiles_x; nty++; }
            have_next = kn < tiles_x * tiles_y;
        } else {
            have_next = fx_tile_origin(kn, tiles_x, tiles_y, n_shards, shard_id, stride, kx, ky, ntx, nty);
        }
        if (LOOP && tid == 0 && have_next && ntx < tiles_x) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            fx_mbar_expect_tx(&s_bar[stage ^ 1], FX_STAGE_BYTES);
            fx_tma_load_2d(&s_px[LOOP ? stage ^ 1 : 0][0], &in_map, ntx * FX_TW - FX_HALO_X, nty * FX_TH - 1, &s_bar[stage ^ 1]);
        }
        if (tx < tiles_x) {
            fx_mbar_wait(&s_bar[stage], (phases >> stage) & 1u);
            phases ^= 1u << stage;
            for (int i = tid; i < FX_BOXH * FX_BOXW; i += FX_THREADS)
                (&s_lum[0][0])[i] = fx_lum(s_px[stage][i]);
            __syncthreads();
            /* tiles that touch no frame border skip the per-pixel bounds and border tests */
            const bool inner = tx > 0 && ty > 0 && (tx + 1) * FX_TW < w && (ty + 1) * FX_TH < h;
            if (!may_store) ;
            else if (inner) fx_filter_tile<true>(s_px[stage], s_lum, out, tx, ty, w, h, tid);
            else            fx_filter_tile<false>(s_px[stage], s_lum, out, tx, ty, w, h, tid);
        }
        if (!LOOP) break;
        __syncthreads();   /* everyone is done with s_px[stage] and s_lum before they are refilled */
        k = kn;
        tx = ntx;
        ty = nty;
        have = have_next;
        stage ^= 1;
    }
    if (SHARD && peer.flags) rt_peer_publish(peer, gridDim.x, tid == 0);
#ifdef RS_TIMELINE
    __syncthreads();
    if (tid == 0) {
        atomicMax(tl + 2, fx_tl_now());
        __threadfence();
        if (atomicAdd(&g_tl_fx_ticket, 1u) == gridDim.x - 1) { g_tl_fx_ticket = 0; g_tl_fx_no++; }
    }
#endif
}

/* pass-through when AA is off (reference kernel.cu:275-278), shard-aware */
__global__ void rt_copy_kernel(const unsigned int* __restrict__ in, unsigned int* __restrict__ out, int w, int h,
                               int tile_w, int tile_h, int n_shards, int shard_id, int stride, const RtPeerSync peer) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    const bool leader = threadIdx.x == 0 && threadIdx.y == 0;
    const bool mine = x < w && y < h && (n_shards <= 1 || rt_tile_owner(x, y, tile_w, tile_h, n_shards, stride) == shard_id);
    if (!peer.flags) {
        if (mine) out[(size_t)y * w + x] = in[(size_t)y * w + x];
        return;
    }
    const unsigned int v = mine ? in[(size_t)y * w + x] : 0u;
    if (__syncthreads_or(mine)) {
        if (fx_peer_acquire(peer, leader) && mine) out[(size_t)y * w + x] = v;
    }
    rt_peer_publish(peer, gridDim.x * gridDim.y, leader);
}

/* The reference's `factor` knob (main.cpp:40-47): the frame is rendered at factor x the window size and the
 * fixed-function blit shows it through a GL_NEAREST texture (main.cpp:317-318), i.e. window pixel (x, y) takes
 * texel (x f + f/2, y f + f/2).  mode 0 restates that; mode 1 averages the f x f block instead (per channel,
 * rounded to nearest), the supersampling the knob was meant for. */
__global__ void rt_present_kernel(const unsigned int* __restrict__ in, unsigned int* __restrict__ out, int w, int h,
                                  int factor, int mode) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    const int sw = w * factor;
    if (mode == 0) {
        out[(size_t)y * w + x] = __ldg(in + (size_t)(y * factor + factor / 2) * sw + x * factor + factor / 2);
        return;
    }
    unsigned int r = 0, g = 0, b = 0;
    for (int j = 0; j < factor; j++)
        for (int i = 0; i < factor; i++) {
            const unsigned int c = __ldg(in + (size_t)(y * factor + j) * sw + x * factor + i);
            r += c & 255u;
            g += (c >> 8) & 255u;
            b += (c >> 16) & 255u;
        }
    const unsigned int n = (unsigned int)(factor * factor), half = n / 2;
    out[(size_t)y * w + x] = ((r + half) / n) | (((g + half) / n) << 8) | (((b + half) / n) << 16);
}

/* gather / scatter of the tiles a shard owns (multi-GPU compositing) */
__global__ void rt_pack_tiles_kernel(const unsigned int* __restrict__ frame, unsigned int* __restrict__ packed,
                                     int w, int h, int tile_w, int tile_h, int n_shards, int shard_id, int stride,
                                     int unpack) {
    const int tiles_x = (w + tile_w - 1) / tile_w, tiles_y = (h + tile_h - 1) / tile_h;
    const int per_row = (tiles_x + n_shards - 1) / n_shards;
    const int slot = blockIdx.x;
    const int ty = slot / per_row, m = slot % per_row;
    if (ty >= tiles_y) return;
    int i0 = (shard_id - ty * stride) % n_shards;
    if (i0 < 0) i0 += n_shards;
    const int tx = i0 + m * n_shards;
    if (tx >= tiles_x) return;
    unsigned int* dst = packed + (size_t)slot * tile_w * tile_h;
    /* blockIdx.y strides over the tile so that large tiles still fill the GPU */
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < tile_w * tile_h; i += blockDim.x * gridDim.y) {
        const int lx = i % tile_w, ly = i / tile_w;
        const int x = tx * tile_w + lx, y = ty * tile_h + ly;
        if (x < w && y < h) {
            if (unpack) ((unsigned int*)frame)[(size_t)y * w + x] = dst[i];
            else        dst[i] = frame[(size_t)y * w + x];
        }
    }
}

static RtPeerSync fx_peer_arg(const RtPeerSync* peer) {
    RtPeerSync p = {nullptr, nullptr, 0u, 0u, 0, 0u};
    if (peer) p = *peer;
    return p;
}

/* per device, once (rt_create).  The FXAA / copy kernel sits between two tracers in the stream, chained by
 * programmatic dependent launch in both directions; it asks for the tracer's shared-memory carve-out so that its CTAs
 * can become resident while the tracer's last CTAs still run, and the next frame's CTAs while its own do
 * (measured through the CTA's shared-memory size: profiles/r02_ab_fxaa_residency.txt, runs 4 and 5). */
extern "C" cudaError_t rt_fxaa_init(void) {
    const int most = cudaSharedmemCarveoutMaxShared;
    cudaError_t e = cudaFuncSetAttribute(rt_fxaa_tma_kernel<false, FX_CTAS_WAVE>, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_fxaa_tma_kernel<false, FX_CTAS_LOOP>, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_fxaa_tma_kernel<true, FX_CTAS_WAVE>, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_fxaa_tma_kernel<true, FX_CTAS_LOOP>, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_fxaa_smem_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_copy_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, most);
    return e;
}

extern "C" cudaError_t rt_launch_fxaa_smem(const unsigned int* in, unsigned int* out, int w, int h,
                                           const rt_shard* sh, const RtPeerSync* peer, cudaStream_t stream) {
    dim3 block(FX_BW, FX_BH, 1), grid((w + FX_BW - 1) / FX_BW, (h + FX_BH - 1) / FX_BH, 1);
    RT_LAUNCHED(1);
    if (sh && sh->n_shards > 1)
        rt_fxaa_smem_kernel<<<grid, block, 0, stream>>>(in, out, w, h, sh->tile_w, sh->tile_h, sh->n_shards,
                                                        sh->shard_id, sh->stride, fx_peer_arg(peer));
    else
        rt_fxaa_smem_kernel<<<grid, block, 0, stream>>>(in, out, w, h, 1, 1, 1, 0, 0, fx_peer_arg(peer));
    return cudaGetLastError();
}

extern "C" cudaError_t rt_launch_fxaa_tma(const CUtensorMap* map, const unsigned int* in, unsigned int* out,
                                          int w, int h, const rt_shard* sh, int n_sm, const RtPeerSync* peer,
                                          cudaStream_t stream) {
    const int tiles_x = (w + FX_TW - 1) / FX_TW, tiles_y = (h + FX_TH - 1) / FX_TH;
    int n = 1, id = 0, st = 0, kx = 1, ky = 1;
    int slots = tiles_x * tiles_y;
    if (sh && sh->n_shards > 1) {     /* screen tiles that are whole multiples of the 64x32 FXAA tile */
        n = sh->n_shards; id = sh->shard_id; st = sh->stride;
        kx = sh->tile_w / FX_TW; ky = sh->tile_h / FX_TH;
        const int stiles_x = (tiles_x + kx - 1) / kx, stiles_y = (tiles_y + ky - 1) / ky;
        slots = ((stiles_x + n - 1) / n) * stiles_y * kx * ky;
    }
    const bool one_wave = slots <= n_sm * FX_CTAS_WAVE;
    int grid = n_sm * (one_wave ? FX_CTAS_WAVE : FX_CTAS_LOOP);
    if (grid > slots) grid = slots;
    if (grid < 1) grid = 1;
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(FX_THREADS, 1, 1);
    cfg.stream = stream;
    cfg.attrs = pdl;
    cfg.numAttrs = 1;
    RT_LAUNCHED(1);
    const bool shard = n > 1 || (peer && peer->flags);
    auto kernel = shard ? (one_wave ? rt_fxaa_tma_kernel<true, FX_CTAS_WAVE> : rt_fxaa_tma_kernel<true, FX_CTAS_LOOP>)
                        : (one_wave ? rt_fxaa_tma_kernel<false, FX_CTAS_WAVE> : rt_fxaa_tma_kernel<false, FX_CTAS_LOOP>);
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, *map, in, out, w, h, n, id, st, kx, ky, fx_peer_arg(peer));
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

extern "C" cudaError_t rt_launch_copy(const unsigned int* in, unsigned int* out, int w, int h, const rt_shard* sh,
                                      const RtPeerSync* peer, cudaStream_t stream) {
    dim3 block(32, 8, 1), grid((w + 31) / 32, (h + 7) / 8, 1);
    RT_LAUNCHED(1);
    if (sh && sh->n_shards > 1)
        rt_copy_kernel<<<grid, block, 0, stream>>>(in, out, w, h, sh->tile_w, sh->tile_h, sh->n_shards, sh->shard_id,
                                                   sh->stride, fx_peer_arg(peer));
    else
        rt_copy_kernel<<<grid, block, 0, stream>>>(in, out, w, h, 1, 1, 1, 0, 0, fx_peer_arg(peer));
    return cudaGetLastError();
}

extern "C" cudaError_t rt_launch_present(const unsigned int* in, unsigned int* out, int w, int h, int factor, int mode,
                                         cudaStream_t stream) {
    dim3 block(32, 8, 1), grid((w + 31) / 32, (h + 7) / 8, 1);
    rt_present_kernel<<<grid, block, 0, stream>>>(in, out, w, h, factor, mode);
    RT_LAUNCHED(1);
    return cudaGetLastError();
}

extern "C" cudaError_t rt_launch_pack_tiles(const unsigned int* frame, unsigned int* packed, int w, int h,
                                            const rt_shard* sh, int unpack, cudaStream_t stream) {
    const int tiles_x = (w + sh->tile_w - 1) / sh->tile_w, tiles_y = (h + sh->tile_h - 1) / sh->tile_h;
    const int per_row = (tiles_x + sh->n_shards - 1) / sh->n_shards;
    int chunks = (sh->tile_w * sh->tile_h + 2047) / 2048;
    if (chunks > 64) chunks = 64;
    RT_LAUNCHED(1);
    rt_pack_tiles_kernel<<<dim3(per_row * tiles_y, chunks, 1), 256, 0, stream>>>(frame, packed, w, h, sh->tile_w, sh->tile_h,
                                                                sh->n_shards, sh->shard_id, sh->stride, unpack);
    return cudaGetLastError();
}

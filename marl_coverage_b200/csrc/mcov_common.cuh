// mcov_common.cuh — shared device helpers for the batched coverage-env kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mcov.h"

namespace mcov {

// ---- per-env header record (32-byte head + pos[N] + cov[N] + adj[N]) ---------------------------
struct __align__(16) EnvHead {
    double done_thresh;   // curriculum threshold, persists across episodes
    double ep_return;     // running episode return (statistics only)
    int32_t currstep;
    int32_t map_idx;
    uint32_t flags;       // Dec: "shown" bit per robot (layer-0 quirk A.3.4); Super: bit0 = first step pending
    uint32_t episode;     // episodes started in this slot (device RNG counter)
};
static_assert(sizeof(EnvHead) == 32, "EnvHead must stay one 32-byte sector");

struct Geometry {
    int W, L;             // env-coordinate map size (Dec: bordered)
    int strips, Wp, PW;   // plane geometry
};

__host__ __device__ inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

__host__ __device__ inline Geometry make_geometry(int W, int L) {
    Geometry g;
    g.W = W; g.L = L;
    g.strips = (L + 31) / 32;
    g.Wp = round_up(W, 4);
    g.PW = g.strips * g.Wp;
    return g;
}

// ---- bit-window access on strip-major planes ----------------------------------------------------
__device__ __forceinline__ uint32_t low_mask(int width) {
    return width >= 32 ? 0xffffffffu : ((1u << width) - 1u);
}

// Cells (x, y0 .. y0+width-1) of a plane as bits 0..width-1; cells outside the plane read 0.
// Caller guarantees 0 <= x < W and width <= 32.
__device__ __forceinline__ uint32_t read_window(const uint32_t* __restrict__ plane, int Wp,
                                                int strips, int x, int y0, int width) {
    const int s0 = y0 >> 5;              // arithmetic shift: floor for negative y0
    const int sh = y0 & 31;
    uint32_t lo = 0, hi = 0;
    if (s0 >= 0 && s0 < strips) lo = __ldg(plane + s0 * Wp + x);
    if (sh + width > 32 && s0 + 1 >= 0 && s0 + 1 < strips) hi = __ldg(plane + (s0 + 1) * Wp + x);
    return __funnelshift_r(lo, hi, sh) & low_mask(width);
}

__device__ __forceinline__ int get_bit(const uint32_t* __restrict__ plane, int Wp, int x, int y) {
    return (__ldg(plane + (y >> 5) * Wp + x) >> (y & 31)) & 1u;
}

// Window-relative bits -> the two absolute words they land in (strip s0 = y0>>5 and s0+1).
__device__ __forceinline__ void split_words(uint32_t rel, int y0, uint32_t& lo, uint32_t& hi) {
    const int sh = y0 & 31;
    const uint64_t v = (uint64_t)rel << sh;
    lo = (uint32_t)v;
    hi = (uint32_t)(v >> 32);
}

// ---- Philox4x32-10 (device RNG for reset draws; keyed by seed, counter = env/episode/draw) ------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

// unbiased-enough bounded draw: floor(u * n / 2^32)
__device__ __forceinline__ int bounded(uint32_t u, int n) { return (int)__umulhi(u, (uint32_t)n); }

// ---- packed poses ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_xy(int x, int y) {
    return (uint32_t)(x & 0xffff) | ((uint32_t)(y & 0xffff) << 16);
}
__device__ __forceinline__ int unpack_x(uint32_t v) { return (int)(int16_t)(v & 0xffffu); }
__device__ __forceinline__ int unpack_y(uint32_t v) { return (int)(int16_t)(v >> 16); }

// Episode start draw (dec_grid_rl.py:462-502, super_grid_rl.py:354-388): uniform map, then N
// distinct walkable cells by rejection over the whole (bordered) map.  Same distribution as the
// reference, not numpy's Mersenne-Twister stream: Philox keyed by `seed`, counter = (global env
// id, episode, draw).  `map_idx_in` / `start_xy_in` (nullable) inject the draw instead.
__device__ inline void draw_starts(const uint32_t* __restrict__ wall_table, const Geometry& g, int N,
                                   int n_maps, const int32_t* map_idx_in, const int16_t* start_xy_in,
                                   uint64_t seed, uint32_t gid, int e, uint32_t episode,
                                   uint32_t* posS, int* map_out) {
    const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
    int mi;
    if (map_idx_in) {
        // caller-supplied device values are clamped, never trusted as addresses (host-side numpy inputs are
        // range-checked by the Python stub before they get here; a device tensor cannot be without a sync)
        mi = min(max(map_idx_in[e], 0), n_maps - 1);
    } else {
        const uint4 r = philox4x32_10(make_uint4(gid, episode, 0u, 0x4D434F56u), key);
        mi = bounded(r.x, n_maps);
    }
    *map_out = mi;
    if (start_xy_in) {
        const int16_t* s = start_xy_in + (size_t)e * N * 2;
        for (int j = 0; j < N; ++j)
            posS[j] = pack_xy(min(max((int)s[2 * j], 0), g.W - 1), min(max((int)s[2 * j + 1], 0), g.L - 1));
        return;
    }
    const uint32_t* wall = wall_table + (size_t)mi * g.PW;
    int placed = 0;
    for (uint32_t k = 1; placed < N && k < 8192u; ++k) {
        const uint4 r = philox4x32_10(make_uint4(gid, episode, k, 0x4D434F56u), key);
        const uint32_t cand[4] = {r.x, r.y, r.z, r.w};
        for (int h = 0; h < 2 && placed < N; ++h) {
            const int x = bounded(cand[2 * h], g.W), y = bounded(cand[2 * h + 1], g.L);
            if (get_bit(wall, g.Wp, x, y)) continue;
            bool taken = false;
            for (int j = 0; j < placed; ++j) taken |= (posS[j] == pack_xy(x, y));
            if (!taken) posS[placed++] = pack_xy(x, y);
        }
    }
    // pathological map (almost no walkable cell): deterministic scan so the kernel always ends
    for (int x = 0; x < g.W && placed < N; ++x)
        for (int y = 0; y < g.L && placed < N; ++y) {
            if (get_bit(wall, g.Wp, x, y)) continue;
            bool taken = false;
            for (int j = 0; j < placed; ++j) taken |= (posS[j] == pack_xy(x, y));
            if (!taken) posS[placed++] = pack_xy(x, y);
        }
    for (; placed < N; ++placed) posS[placed] = pack_xy(0, 0);
}

// ---- async bulk (TMA engine) store of a contiguous shared-memory block --------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- async bulk load global -> shared with mbarrier completion (map-window staging, mcov_dec_cta.cu) ----
// The destination is named in the shared::cluster window: without a cluster launch a CTA's own shared::cta
// addresses are valid shared::cluster addresses of itself (PTX ISA, "shared memory state space").
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void bulk_load_g2s(void* sdst, const void* gsrc, uint32_t bytes,
                                              uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(sdst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

}  // namespace mcov

// mcov_dec_shared.cuh — small device helpers shared by the two DecGridRL step kernels:
// the lane-group kernel (mcov_dec.cu, large batches) and the CTA-per-env kernel (mcov_dec_cta.cu,
// small batches / large maps).
#pragma once
#include "mcov_internal.h"

namespace mcov {

namespace {

constexpr unsigned FULL = 0xffffffffu;
enum { PASS_DONE = 0, PASS_STORE = 1, PASS_AGAIN = 2 };   // env pass results (PASS_STORE: the caller stores the block; PASS_AGAIN: re-seed the finished env in place)

// floor(k / d) for k < 2^26, d <= 64, magic = ceil(2^32 / d)
// (d == 1 gives magic = 2^32 = 0 in 32 bits: the quotient is k itself)
__device__ __forceinline__ int fast_div(int k, uint32_t magic) {
    return magic ? (int)__umulhi((uint32_t)k, magic) : k;
}

// Draws map + start poses on the device (see draw_starts in mcov_common.cuh).
__device__ void draw_reset(const DecParams& p, int e, uint32_t episode, uint32_t* posS, int* map_out) {
    draw_starts(p.wall, p.g, p.N, p.n_maps, p.map_idx_in, p.start_xy_in, p.seed,
                (uint32_t)(p.env_id_base + e), e, episode, posS, map_out);
}

__device__ __forceinline__ void prefetch_l1(const void* ptr) {
    asm volatile("prefetch.global.L1 [%0];" :: "l"(ptr));
}
__device__ __forceinline__ void prefetch_l2(const void* ptr) {
    asm volatile("prefetch.global.L2 [%0];" :: "l"(ptr));
}
#ifndef MCOV_ST_MODE
#define MCOV_ST_MODE 0                       // observation store policy: 0 .cs  1 default  2 .cg  3 .wt
#endif
__device__ __forceinline__ void obs_store(float4* ptr, const float4 v) {
#if MCOV_ST_MODE == 0
    __stcs(ptr, v);
#elif MCOV_ST_MODE == 1
    *ptr = v;
#elif MCOV_ST_MODE == 2
    __stcg(ptr, v);
#else
    __stwt(ptr, v);
#endif
}
// ORs `v` into the bit stream at bit offset `off` (shared-memory atomics:
// neighbouring rows share words).
__device__ __forceinline__ void stream_or(uint32_t* bits, int off, uint32_t v) {
    const int w = off >> 5, sh = off & 31;
    const uint32_t lo = v << sh, hi = __funnelshift_l(v, 0u, sh);   // bits of v pushed past bit 31
    if (lo) atomicOr(bits + w, lo);
    if (hi) atomicOr(bits + w + 1, hi);
}

// Sensor-frame helpers.  A robot at (xj, yj) with range R touches, in map row x (|x - xj| <= R),
// the cells y in [yj-R, yj+R]; `ymask_in_strip` is that interval cut to absolute strip s.
__device__ __forceinline__ uint32_t ymask_in_strip(int yj, int R, int S, int s) {
    const int y0 = yj - R, s0 = y0 >> 5;
    const uint64_t m = (uint64_t)low_mask(S) << (y0 & 31);
    return s == s0 ? (uint32_t)m : (s == s0 + 1 ? (uint32_t)(m >> 32) : 0u);
}


// Integer beam table walk, cell by cell (lidar.py:30-65 on the proved start-independent cell list, mcov_set_lidar_beams).
// `ww` / `wf` / `wo`: the robot's truth-wall, free-hit and obstacle-hit windows in shared memory, S + 1 words each — row S
// is a dummy the caller keeps all-wall (ww) / uses as a sink (wo), so a beam's unused slots end the walk like an obstacle.
__device__ __forceinline__ void walk_beam_cells(const uint16_t* __restrict__ cells, int l, const uint32_t* ww, uint32_t* wf,
                                                uint32_t* wo) {
    const uint4* tp = reinterpret_cast<const uint4*>(cells + l * 16);
    const uint4 ta = __ldg(tp), tb = __ldg(tp + 1);
    const uint32_t c[8] = {ta.x, ta.y, ta.z, ta.w, tb.x, tb.y, tb.z, tb.w};
    const unsigned char* wwb = reinterpret_cast<const unsigned char*>(ww);
    unsigned char* wfb = reinterpret_cast<unsigned char*>(wf);
    unsigned char* wob = reinterpret_cast<unsigned char*>(wo);
#pragma unroll
    for (int k = 1; k < 16; ++k) {
        const uint32_t cell = (k & 1) ? (c[k >> 1] >> 16) : (c[k >> 1] & 0xffffu);
        const uint32_t off = cell >> 8;                              // byte offset of the row word
        const uint32_t m = __funnelshift_l(0u, 1u, cell);            // 1 << (cell & 31): the y bit
        if (*reinterpret_cast<const uint32_t*>(wwb + off) & m) { atomicOr(reinterpret_cast<uint32_t*>(wob + off), m); break; }
        atomicOr(reinterpret_cast<uint32_t*>(wfb + off), m);
    }
}

// Integer beam table walk (lidar.py:30-65 restated on the proved start-independent cell list, mcov_set_lidar_beams): the
// cells of beam l after the robot's own one, as same-row runs.  `ww` / `wf` / `wo` are the robot's truth-wall, free-hit and
// obstacle-hit windows in shared memory (one word per window row).  Per run: no wall in it -> all its cells are free hits;
// otherwise the first wall in march order is the obstacle hit, the cells before it are free hits, and the beam ends.
__device__ __forceinline__ void walk_beam_runs(const uint2* __restrict__ runs, int l, const uint32_t* ww, uint32_t* wf,
                                               uint32_t* wo) {
    const uint2* tb = runs + l * 16;
    uint2 r = __ldg(tb);
#pragma unroll 1
    for (int k = 1; r.x != 0u; ++k) {
        const uint2 nx = k < 16 ? __ldg(tb + k) : make_uint2(0u, 0u);        // next run, in flight behind this one
        const uint32_t row = (r.y & 0x7fffffffu) >> 2;
        const uint32_t hit = ww[row] & r.x;
        if (hit) {
            uint32_t first, freeb;
            if (r.y >> 31) { first = 0x80000000u >> __clz(hit); freeb = r.x & ~((first << 1) - 1u); }
            else { first = hit & (0u - hit); freeb = r.x & (first - 1u); }
            if (freeb) atomicOr(wf + row, freeb);
            atomicOr(wo + row, first);
            break;
        }
        atomicOr(wf + row, r.x);
        r = nx;
    }
}

// ---- episode statistics: per-group (or per-CTA) sums in shared memory -> one row per CTA in global scratch ->
// the last-arriving CTA folds all rows into stats_out (single launch, graph-capturable).  `rows` holds `nrows`
// records of MCOV_STATS_LEN doubles, `row_stride` bytes apart; `red` is blockDim.x doubles of shared memory that
// may overlay `rows` (every thread has passed the first barrier before it is written).
__device__ __forceinline__ void fold_episode_stats(const DecParams& p, const unsigned char* rows, int nrows,
                                                   size_t row_stride, double* red, int* last_flag) {
    if (!p.stats_out) return;
    __syncthreads();
    if (threadIdx.x < MCOV_STATS_LEN) {
        double s = 0.0;
        for (int w = 0; w < nrows; ++w)
            s += reinterpret_cast<const double*>(rows + (size_t)w * row_stride)[threadIdx.x];
        p.stats_partial[(size_t)blockIdx.x * MCOV_STATS_LEN + threadIdx.x] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *last_flag = (atomicAdd(p.counter, 1u) == gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (*last_flag) {
        __threadfence();
        const int k = threadIdx.x & (MCOV_STATS_LEN - 1), r0 = threadIdx.x >> 3, rstep = blockDim.x >> 3;
        double s = 0.0;
#pragma unroll 8
        for (int r = r0; r < (int)gridDim.x; r += rstep)           // loads of 8 rows in flight, summed in row order
            s += __ldcg(p.stats_partial + (size_t)r * MCOV_STATS_LEN + k);
        __syncthreads();
        red[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x < MCOV_STATS_LEN) {
            double t = 0.0;
            for (int r = 0; r < rstep; ++r) t += red[r * MCOV_STATS_LEN + threadIdx.x];
            p.stats_out[threadIdx.x] = t;
        }
        if (threadIdx.x == 0) *p.counter = 0u;
    }
}

inline uint32_t magic_for(int d) { return (uint32_t)((0x100000000ull + (uint64_t)d - 1) / (uint64_t)d); }

}  // namespace

}  // namespace mcov

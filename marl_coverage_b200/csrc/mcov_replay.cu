// mcov_replay.cu — GPU replay-buffer rows for batched rollouts.
//
// Reference: Policies/replaybuffer.py:3-72 — float32 ring arrays _state / _action / _reward /
// _nextstate / _done, addtransition() writes one row at _ptr and advances it modulo the capacity,
// samplebatch() gathers rows by random index.  Here one launch writes the E transitions of a batched
// env step into E consecutive ring rows (wrapping), straight from the observation buffers the step
// kernel filled, and one launch gathers a minibatch of all five arrays.  Pure HBM copies: every
// thread moves 16 bytes, rows are contiguous, grid is a multiple of the SM count.
#include "mcov_internal.h"

namespace mcov {
namespace {

// Measured on B200 and NOT the default: 131 072 rows of 363 floats take 0.217 ms this way (0.275 ms when both source
// vectors are loaded instead of shuffled) against 0.156 ms for plain 4-byte accesses, whose warp-wide 128-byte
// transactions are already full lines; -DMCOV_REPLAY_SHIFT=1 rebuilds the variant for an A/B.
#ifndef MCOV_REPLAY_SHIFT
#define MCOV_REPLAY_SHIFT 0
#endif

// One warp copies `row` floats s -> d when neither side is 16-byte aligned (the reference's rows of
// 3*11*11 = 363 floats): stores are aligned float4 on the DESTINATION, every one assembled from the two
// aligned source float4 around it (the shift sa is the same for the whole row).  Only chunks whose two
// source vectors lie inside the row are vectorised, so nothing outside [s, s+row) is ever touched; the
// <= 7 floats in front and <= 10 behind go one lane each.
template <bool STREAM_LD>
__device__ __forceinline__ void copy_row_shifted(const float* __restrict__ s, float* __restrict__ d, int row, int lane) {
    const int h = (4 - (int)((reinterpret_cast<uintptr_t>(d) >> 2) & 3)) & 3;      // floats up to d's alignment
    const int sa = (int)((reinterpret_cast<uintptr_t>(s + h) >> 2) & 3);           // source shift in floats
    const int jbeg = sa > h ? 1 : 0;                                               // first chunk with lo >= s
    const int need = sa ? 8 - sa : 4;                                              // floats from chunk start to end of hi
    const int room = row - h - need;
    const int jend = room >= 0 ? room / 4 + 1 : 0;                                 // chunks [jbeg, jend)
    const int body0 = h + 4 * jbeg, body1 = jend > jbeg ? h + 4 * jend : body0;
    if (lane < body0 && lane < row) d[lane] = STREAM_LD ? __ldcs(s + lane) : __ldg(s + lane);
    if (body1 + lane < row) d[body1 + lane] = STREAM_LD ? __ldcs(s + body1 + lane) : __ldg(s + body1 + lane);
    const float4* sv = reinterpret_cast<const float4*>(s + h - sa);                // aligned: chunk j's lo is sv[j]
    float4* dv = reinterpret_cast<float4*>(d + h);
    // hi of chunk j is lo of chunk j+1: taken from the next lane by shuffle (every 16-byte vector is loaded once);
    // only lane 31 and the lane holding the last chunk load their own hi.
    for (int jb = jbeg; jb < jend; jb += 128) {                                    // warp-uniform trip count
        float4 lo[4];
        float hx[4], hy[4], hz[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = jb + lane + 32 * u;
            lo[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < jend) lo[u] = STREAM_LD ? __ldcs(sv + j) : __ldg(sv + j);
            hx[u] = hy[u] = hz[u] = 0.f;
            if (sa && j < jend && (lane == 31 || j + 1 >= jend)) {
                const float4 t = STREAM_LD ? __ldcs(sv + j + 1) : __ldg(sv + j + 1);
                hx[u] = t.x; hy[u] = t.y; hz[u] = t.z;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = jb + lane + 32 * u;
            const bool own = lane == 31 || j + 1 >= jend;
            float4 o = lo[u];
            if (sa) {                                                              // warp-uniform
                const float nx = __shfl_down_sync(0xffffffffu, lo[u].x, 1);
                const float x = own ? hx[u] : nx;
                if (sa == 1) o = make_float4(lo[u].y, lo[u].z, lo[u].w, x);
                else {
                    const float ny = __shfl_down_sync(0xffffffffu, lo[u].y, 1);
                    const float y = own ? hy[u] : ny;
                    if (sa == 2) o = make_float4(lo[u].z, lo[u].w, x, y);
                    else {
                        const float nz = __shfl_down_sync(0xffffffffu, lo[u].z, 1);
                        o = make_float4(lo[u].w, x, y, own ? hz[u] : nz);
                    }
                }
            }
            if (j < jend) __stcs(dv + j, o);
        }
    }
}

// VT = float4 (rows and strides multiples of 16 bytes) or float (anything else, e.g. the reference's
// state[0] rows of 3*11*11 = 363 floats).  src_stride: elements between consecutive envs' rows in the
// source buffers (>= row: lets the caller push obs[:, 0] without a slicing copy).
template <class VT>
__global__ void __launch_bounds__(256)
replay_push_kernel(const VT* __restrict__ state, const VT* __restrict__ next_state, int row, long long src_stride,
                   const uint8_t* __restrict__ act_u8, const float* __restrict__ act_f32, int act_dim,
                   const float* __restrict__ reward, const uint8_t* __restrict__ done, int E,
                   VT* __restrict__ r_state, VT* __restrict__ r_next, float* __restrict__ r_action,
                   float* __restrict__ r_reward, float* __restrict__ r_done, long long ptr, long long capacity) {
    // one warp per transition row: no per-element division, 8 independent 4/16-byte loads in flight per lane
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31, nwarps = (int)(stride >> 5);
    for (int e = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); e < E; e += nwarps) {
        const VT* s0 = state + e * src_stride;
        const VT* s1 = next_state + e * src_stride;
        const long long dst = ((ptr + e) % capacity) * row;
        if (MCOV_REPLAY_SHIFT && sizeof(VT) == 4 && row >= 16) {
            copy_row_shifted<true>(reinterpret_cast<const float*>(s0), reinterpret_cast<float*>(r_state + dst), row, lane);
            copy_row_shifted<true>(reinterpret_cast<const float*>(s1), reinterpret_cast<float*>(r_next + dst), row, lane);
            continue;
        }
        for (int c = lane; c < row; c += 128) {
            VT a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c + 32 * u < row) { a[u] = __ldcs(s0 + c + 32 * u); b[u] = __ldcs(s1 + c + 32 * u); }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c + 32 * u < row) { __stcs(r_state + dst + c + 32 * u, a[u]); __stcs(r_next + dst + c + 32 * u, b[u]); }
        }
    }
    const long long small = (long long)E * act_dim;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < small; k += stride) {
        const int e = (int)(k / act_dim);
        const int c = (int)(k - (long long)e * act_dim);
        r_action[((ptr + e) % capacity) * act_dim + c] = act_u8 ? (float)act_u8[k] : act_f32[k];
    }
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
        const long long r = (ptr + e) % capacity;
        r_reward[r] = reward[e];
        r_done[r] = done[e] ? 1.0f : 0.0f;
    }
}

template <class VT>
__global__ void __launch_bounds__(256)
replay_gather_kernel(const long long* __restrict__ idx, long long capacity, int B, int row, int act_dim,
                     const VT* __restrict__ r_state, const VT* __restrict__ r_next,
                     const float* __restrict__ r_action, const float* __restrict__ r_reward,
                     const float* __restrict__ r_done, VT* __restrict__ o_state, VT* __restrict__ o_next,
                     float* __restrict__ o_action, float* __restrict__ o_reward, float* __restrict__ o_done) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31, nwarps = (int)(stride >> 5);
    // sampled indices are device values nobody has looked at: clamped into the ring, never used raw
    auto ring_row = [&](int b) { return min(max(idx[b], 0ll), capacity - 1); };
    for (int b = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); b < B; b += nwarps) {
        const long long src = ring_row(b) * row, dst = (long long)b * row;
        if (MCOV_REPLAY_SHIFT && sizeof(VT) == 4 && row >= 16) {
            copy_row_shifted<false>(reinterpret_cast<const float*>(r_state + src), reinterpret_cast<float*>(o_state + dst), row, lane);
            copy_row_shifted<false>(reinterpret_cast<const float*>(r_next + src), reinterpret_cast<float*>(o_next + dst), row, lane);
            continue;
        }
        for (int c = lane; c < row; c += 128) {
            VT a[4], d[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c + 32 * u < row) { a[u] = __ldg(r_state + src + c + 32 * u); d[u] = __ldg(r_next + src + c + 32 * u); }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c + 32 * u < row) { __stcs(o_state + dst + c + 32 * u, a[u]); __stcs(o_next + dst + c + 32 * u, d[u]); }
        }
    }
    const long long small = (long long)B * act_dim;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < small; k += stride) {
        const int b = (int)(k / act_dim);
        o_action[k] = r_action[ring_row(b) * act_dim + (k - (long long)b * act_dim)];
    }
    for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < B; b += stride) {
        o_reward[b] = r_reward[ring_row((int)b)];
        o_done[b] = r_done[ring_row((int)b)];
    }
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int grid_for(long long items) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long blocks = (items + 255) / 256;
#ifndef MCOV_REPLAY_CAP
#define MCOV_REPLAY_CAP 64
#endif
    const long long cap = (long long)sms * MCOV_REPLAY_CAP;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

cudaError_t launch_replay_push(const float* state, const float* next_state, int row_floats, long long src_stride,
                               const uint8_t* act_u8, const float* act_f32, int act_dim, const float* reward,
                               const uint8_t* done, int E, float* r_state, float* r_next, float* r_action,
                               float* r_reward, float* r_done, long long ptr, long long capacity, cudaStream_t st) {
    const bool vec = row_floats % 4 == 0 && src_stride % 4 == 0 && al16(state) && al16(next_state) &&
                     al16(r_state) && al16(r_next);
    if (vec) {
        const int row = row_floats / 4;
        replay_push_kernel<float4><<<grid_for((long long)E * 32), 256, 0, st>>>(
            reinterpret_cast<const float4*>(state), reinterpret_cast<const float4*>(next_state), row, src_stride / 4,
            act_u8, act_f32, act_dim, reward, done, E, reinterpret_cast<float4*>(r_state),
            reinterpret_cast<float4*>(r_next), r_action, r_reward, r_done, ptr, capacity);
    } else {
        replay_push_kernel<float><<<grid_for((long long)E * 32), 256, 0, st>>>(
            state, next_state, row_floats, src_stride, act_u8, act_f32, act_dim, reward, done, E, r_state, r_next,
            r_action, r_reward, r_done, ptr, capacity);
    }
    return cudaGetLastError();
}

cudaError_t launch_replay_gather(const long long* idx, long long capacity, int B, int row_floats, int act_dim, const float* r_state,
                                 const float* r_next, const float* r_action, const float* r_reward,
                                 const float* r_done, float* o_state, float* o_next, float* o_action,
                                 float* o_reward, float* o_done, cudaStream_t st) {
    const bool vec = row_floats % 4 == 0 && al16(r_state) && al16(r_next) && al16(o_state) && al16(o_next);
    if (vec) {
        const int row = row_floats / 4;
        replay_gather_kernel<float4><<<grid_for((long long)B * 32), 256, 0, st>>>(
            idx, capacity, B, row, act_dim, reinterpret_cast<const float4*>(r_state), reinterpret_cast<const float4*>(r_next),
            r_action, r_reward, r_done, reinterpret_cast<float4*>(o_state), reinterpret_cast<float4*>(o_next),
            o_action, o_reward, o_done);
    } else {
        replay_gather_kernel<float><<<grid_for((long long)B * 32), 256, 0, st>>>(
            idx, capacity, B, row_floats, act_dim, r_state, r_next, r_action, r_reward, r_done, o_state, o_next, o_action,
            o_reward, o_done);
    }
    return cudaGetLastError();
}

}  // namespace mcov

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
replay_gather_kernel(const long long* __restrict__ idx, int B, int row, int act_dim,
                     const VT* __restrict__ r_state, const VT* __restrict__ r_next,
                     const float* __restrict__ r_action, const float* __restrict__ r_reward,
                     const float* __restrict__ r_done, VT* __restrict__ o_state, VT* __restrict__ o_next,
                     float* __restrict__ o_action, float* __restrict__ o_reward, float* __restrict__ o_done) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31, nwarps = (int)(stride >> 5);
    for (int b = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); b < B; b += nwarps) {
        const long long src = idx[b] * row, dst = (long long)b * row;
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
        o_action[k] = r_action[idx[b] * act_dim + (k - (long long)b * act_dim)];
    }
    for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < B; b += stride) {
        o_reward[b] = r_reward[idx[b]];
        o_done[b] = r_done[idx[b]];
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

cudaError_t launch_replay_gather(const long long* idx, int B, int row_floats, int act_dim, const float* r_state,
                                 const float* r_next, const float* r_action, const float* r_reward,
                                 const float* r_done, float* o_state, float* o_next, float* o_action,
                                 float* o_reward, float* o_done, cudaStream_t st) {
    const bool vec = row_floats % 4 == 0 && al16(r_state) && al16(r_next) && al16(o_state) && al16(o_next);
    if (vec) {
        const int row = row_floats / 4;
        replay_gather_kernel<float4><<<grid_for((long long)B * 32), 256, 0, st>>>(
            idx, B, row, act_dim, reinterpret_cast<const float4*>(r_state), reinterpret_cast<const float4*>(r_next),
            r_action, r_reward, r_done, reinterpret_cast<float4*>(o_state), reinterpret_cast<float4*>(o_next),
            o_action, o_reward, o_done);
    } else {
        replay_gather_kernel<float><<<grid_for((long long)B * 32), 256, 0, st>>>(
            idx, B, row_floats, act_dim, r_state, r_next, r_action, r_reward, r_done, o_state, o_next, o_action,
            o_reward, o_done);
    }
    return cudaGetLastError();
}

}  // namespace mcov

// mcov_aux.cu — the two callers either side of the env step that SURVEY.md §8f ranks 3 and 4 name:
//
//   sim_lookahead_kernel  SuperGrid_Sim.step (Environments/sim_env.py:57-190): a stateless one-robot,
//                         one-step lookahead on a caller-supplied 4-plane float state (position one-hot,
//                         observed obstacles, not-yet-covered, distance channel).  DRQN / HA* call it once
//                         per candidate action; here one launch evaluates a batch of (state, action) pairs.
//   render_kernel         render() of both envs (super_grid_rl.py:423-463, dec_grid_rl.py:554-616) into a
//                         headless uint8 RGB buffer straight from the bit planes; no window.
//
// Both are HBM streams (lookahead: copy 3 planes + write 4; render: write the frame), so they follow the
// same rules as the step kernels: 128-bit streaming accesses, grids sized in multiples of the SM count,
// no atomics to global.
#include "mcov_internal.h"

namespace mcov {
namespace {

constexpr int SIM_THREADS = 256;
constexpr int SIM_MAX_CELLS = (2 * MCOV_MAX_RADIUS + 1) * (2 * MCOV_MAX_RADIUS + 1);

// footprint cell classes, in the order of the reference's if / elif chain (sim_env.py:150-170)
enum { CELL_NONE = 0, CELL_NEW = 1, CELL_AGAIN = 2, CELL_OBST = 3 };

__device__ __forceinline__ int block_sum(int v, int* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
    for (int w = 0; w < SIM_THREADS / 32; ++w) t += red[w];
    __syncthreads();
    return t;
}

__global__ void __launch_bounds__(SIM_THREADS, 4) sim_lookahead_kernel(SimParams p) {
    __shared__ int s_first;                    // raster index of the robot in the position plane
    __shared__ int s_pos[2];                   // pose after the move
    __shared__ int s_red[SIM_THREADS / 32];
    __shared__ double s_reward;
    __shared__ int s_newly;
    __shared__ uint8_t s_cls[SIM_MAX_CELLS];
    __shared__ uint8_t s_val[SIM_MAX_CELLS];   // grid value of a newly covered cell (0 or 1)

    const Geometry g = p.g;
    const int W = g.W, L = g.L, cells = W * L, S = 2 * p.R + 1;
    const int tid = threadIdx.x;

    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        const float* in = p.state + (size_t)b * 4 * cells;
        float* out = p.next_state + (size_t)b * 4 * cells;
        const bool inplace = (in == out);
        const int mi = p.map_idx ? p.map_idx[b] : 0;
        const uint32_t* wall = p.wall + (size_t)mi * g.PW;
        const uint32_t* posm = p.pos + (size_t)mi * g.PW;

        // ---- where is the robot: np.nonzero(pos_img) (sim_env.py:78), or the caller's pose ----------
        if (tid == 0) s_first = p.pos_xy ? (int)p.pos_xy[2 * b] * L + (int)p.pos_xy[2 * b + 1] : 0x7fffffff;
        __syncthreads();
        if (!p.pos_xy) {
            int first = 0x7fffffff;
            for (int k = tid; k < cells; k += SIM_THREADS)
                if (__ldg(in + k) != 0.0f) { first = k; break; }        // own stride: ascending
            first = __reduce_min_sync(0xffffffffu, first);
            if ((tid & 31) == 0 && first != 0x7fffffff) atomicMin(&s_first, first);   // shared memory
            __syncthreads();
        }

        // ---- the move (sim_env.py:92-143) and the footprint classes (sim_env.py:150-170) --------------
        if (tid == 0) {
            const int idx = s_first == 0x7fffffff ? 0 : s_first;
            int x = idx / L, y = idx - x * L;
            const int a = p.actions[b] & 3;
            const int nx = x + (a == 0 ? -1 : a == 1 ? 1 : 0);
            const int ny = y + (a == 2 ? 1 : a == 3 ? -1 : 0);
            double r = 0.0;
            if (nx >= 0 && nx < W && ny >= 0 && ny < L && !get_bit(wall, g.Wp, nx, ny)) {
                x = nx; y = ny;
                if (p.dist_reward) r += (double)__ldg(in + 3 * (size_t)cells + x * L + y);
            } else {
                r -= p.cp;
            }
            s_pos[0] = x; s_pos[1] = y;
            s_reward = r;
        }
        __syncthreads();
        const int px = s_pos[0], py = s_pos[1];
        for (int k = tid; k < S * S; k += SIM_THREADS) {
            const int j = px - p.R + k / S, kk = py - p.R + k % S;
            int cls = CELL_NONE, val = 0;
            if (j >= 0 && j < W && kk >= 0 && kk < L) {
                if (!get_bit(wall, g.Wp, j, kk)) {
                    const float f = __ldg(in + 2 * (size_t)cells + j * L + kk);
                    if (f == 1.0f) { cls = CELL_NEW; val = get_bit(posm, g.Wp, j, kk); }
                    else if (f == 0.0f) cls = CELL_AGAIN;
                } else if (__ldg(in + (size_t)cells + j * L + kk) == 0.0f) {
                    cls = CELL_OBST;
                }
            }
            s_cls[k] = (uint8_t)cls;
            s_val[k] = (uint8_t)val;
        }
        __syncthreads();
        if (tid == 0) {              // float64 accumulation in the reference's raster order
            double r = s_reward;
            int newly = 0;
            for (int k = 0; k < S * S; ++k) {
                const int cls = s_cls[k];
                if (cls == CELL_NEW) { r += (double)s_val[k]; ++newly; }
                else if (cls == CELL_AGAIN) r -= p.free_penalty;
            }
            s_reward = r;
            s_newly = newly;
        }

        // ---- stream the planes: out[0] = 0, out[1..3] = in[1..3]; count covered cells on the way --------
        int covered = 0;
        if (p.vec_ok) {
            const int q = cells >> 2;
            const float4* in4 = reinterpret_cast<const float4*>(in);
            float4* out4 = reinterpret_cast<float4*>(out);
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            if (!inplace) {
                for (int k = tid; k < q; k += SIM_THREADS) {
                    __stcs(out4 + k, z);
                    __stcs(out4 + q + k, __ldcs(in4 + q + k));
                    __stcs(out4 + 3 * q + k, __ldcs(in4 + 3 * q + k));
                }
            }
            for (int k = tid; k < q; k += SIM_THREADS) {
                const float4 v = inplace ? in4[2 * q + k] : __ldcs(in4 + 2 * q + k);
                covered += (v.x < 1.0f) + (v.y < 1.0f) + (v.z < 1.0f) + (v.w < 1.0f);
                if (!inplace) __stcs(out4 + 2 * q + k, v);
            }
        } else {
            for (int k = tid; k < cells; k += SIM_THREADS) {
                const float v = in[2 * (size_t)cells + k];
                covered += (v < 1.0f);
                if (!inplace) {
                    out[k] = 0.0f;
                    out[(size_t)cells + k] = in[(size_t)cells + k];
                    out[2 * (size_t)cells + k] = v;
                    out[3 * (size_t)cells + k] = in[3 * (size_t)cells + k];
                }
            }
        }
        covered = block_sum(covered, s_red);     // barriers inside: the bulk stores precede the patches

        // ---- patches: <= (2r+1)^2 cells and the pose ---------------------------------------------------
        for (int k = tid; k < S * S; k += SIM_THREADS) {
            const int cls = s_cls[k];
            if (cls == CELL_NEW || cls == CELL_OBST) {
                const int j = px - p.R + k / S, kk = py - p.R + k % S;
                if (cls == CELL_NEW) out[2 * (size_t)cells + j * L + kk] = 0.0f;
                else out[(size_t)cells + j * L + kk] = 1.0f;
            }
        }
        if (tid == 0) {
            if (inplace && s_first != 0x7fffffff) out[s_first] = 0.0f;
            out[px * L + py] = 1.0f;
            // percent_covered (sim_env.py:248-259) and the terminal bonus (sim_env.py:187-188)
            const int total = __ldg(p.free_counts + mi);
            const double pct = (double)(covered + s_newly) / (double)total;
            const bool term = fmin(p.done_thresh, 1.0) <= pct;
            double r = s_reward;
            if (term) r += p.terminal;
            if (p.reward64) p.reward64[b] = r;
            if (p.reward) p.reward[b] = (float)r;
            if (p.terminal_out) p.terminal_out[b] = term ? 1 : 0;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------
// render
// ------------------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t render_byte(const RenderParams& p, int e, int pixel, int ch) {
    const int i = pixel / p.out_w;
    int j = pixel - i * p.out_w;
    if (p.flip) j = p.out_w - 1 - j;                     // cv2.flip(image, 1), dec_grid_rl.py:593
    if (i >= p.g.W * p.scale || j >= p.g.L * p.scale) return 0u;     // copyMakeBorder zeros
    const int x = i / p.scale, y = j / p.scale;          // INTER_NEAREST with an integer factor
    const int w = (y >> 5) * p.g.Wp + x;
    const uint32_t bit = 1u << (y & 31);
    uint32_t obst = 0, cov = 0;
    if (p.env_type == MCOV_ENV_DEC) {
        cov = p.plane_a[(size_t)e * p.g.PW + w] & bit;                   // _visited
        const uint2* fo = p.fo + (size_t)e * p.N * p.g.PW;
        for (int r = 0; r < p.N; ++r) obst |= fo[(size_t)r * p.g.PW + w].y & bit;   // union of obst_i
    } else {
        cov = p.plane_a[(size_t)e * p.g.PW + w] & bit;                   // 1 - _free
        obst = p.plane_b[(size_t)e * p.g.PW + w] & bit;                  // _observed_obstacles
    }
    const int16_t* pos = reinterpret_cast<const int16_t*>(p.hdr + (size_t)e * p.hdr_stride + p.off_pos);
    uint32_t robot = 0;
    for (int r = 0; r < p.n_shown; ++r) robot |= (pos[2 * r] == x && pos[2 * r + 1] == y);
    uint32_t v;
    if (ch == 0) v = (obst ? 200u : 0u) + (robot ? 255u : 0u);
    else if (ch == 1) v = cov ? 225u : 0u;
    else v = (obst ? 255u : 0u) + (cov ? 255u : 0u);
    return v & 0xffu;
}

__global__ void __launch_bounds__(256) render_kernel(RenderParams p) {
    const long long frame_bytes = (long long)p.out_h * p.out_w * 3;
    const long long total = frame_bytes * p.n;
    const long long words = total >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    uint32_t* out32 = reinterpret_cast<uint32_t*>(p.out);
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < words; t += stride) {
        const long long b0 = t << 2;
        int f = (int)(b0 / frame_bytes);
        int r = (int)(b0 - (long long)f * frame_bytes);
        uint32_t word = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (r >= frame_bytes) { r = 0; ++f; }
            const int e = p.env_ids ? p.env_ids[f] : f;
            const int pixel = r / 3;
            word |= render_byte(p, e, pixel, r - pixel * 3) << (8 * k);
            ++r;
        }
        __stcs(out32 + t, word);
    }
    if (blockIdx.x == 0 && threadIdx.x < (int)(total & 3)) {     // <= 3 tail bytes
        const long long b = (words << 2) + threadIdx.x;
        const int f = (int)(b / frame_bytes);
        const int r = (int)(b - (long long)f * frame_bytes);
        const int e = p.env_ids ? p.env_ids[f] : f;
        p.out[b] = (uint8_t)render_byte(p, e, r / 3, r % 3);
    }
}

}  // namespace

cudaError_t launch_sim_lookahead(const SimParams& p, int num_sms, cudaStream_t st) {
    int grid = num_sms * 4;
    if (grid > p.B) grid = p.B;
    sim_lookahead_kernel<<<grid, SIM_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_render(const RenderParams& p, int num_sms, cudaStream_t st) {
    const long long words = ((long long)p.out_h * p.out_w * 3 * p.n) >> 2;
    long long grid = (words + 255) / 256;
    const long long cap = (long long)num_sms * 16;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    render_kernel<<<(int)grid, 256, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace mcov

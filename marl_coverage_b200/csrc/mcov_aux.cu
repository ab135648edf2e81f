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

#ifndef MCOV_SIM_MINB
#define MCOV_SIM_MINB 6
#endif
__global__ void __launch_bounds__(SIM_THREADS, MCOV_SIM_MINB) sim_lookahead_kernel(SimParams p) {
    __shared__ int s_first;                    // raster index of the robot in the position plane
    __shared__ int s_pos[2];                   // pose after the move
    __shared__ int s_red[SIM_THREADS / 32];
    __shared__ double s_reward;
    __shared__ uint8_t s_cls[SIM_MAX_CELLS];
    __shared__ uint8_t s_val[SIM_MAX_CELLS];   // grid value of a newly covered cell (0 or 1)

    const Geometry g = p.g;
    const int W = g.W, L = g.L, cells = W * L, S = 2 * p.R + 1;
    const int tid = threadIdx.x;

    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        const float* in = p.state + (size_t)b * 4 * cells;
        float* out = p.next_state + (size_t)b * 4 * cells;
        const bool inplace = (in == out);
        const int mi = p.map_idx ? min(max(p.map_idx[b], 0), p.n_maps - 1) : 0;       // device value: clamped
        const uint32_t* wall = p.wall + (size_t)mi * g.PW;
        const uint32_t* posm = p.pos + (size_t)mi * g.PW;

        // ---- stream the planes first (nothing below depends on the pose): out[0] = 0, out[1..3] = in[1..3];
        //      count covered cells and find the robot (np.nonzero(pos_img), sim_env.py:78) on the way -------
        if (tid == 0)
            s_first = p.pos_xy ? min(max((int)p.pos_xy[2 * b], 0), W - 1) * L + min(max((int)p.pos_xy[2 * b + 1], 0), L - 1)
                               : 0x7fffffff;
        int covered = 0, first = 0x7fffffff;
        if (p.vec_ok) {
            const int q = cells >> 2;
            const float4* in4 = reinterpret_cast<const float4*>(in);
            float4* out4 = reinterpret_cast<float4*>(out);
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = tid; k < q; k += SIM_THREADS) {
                if (!p.pos_xy) {
                    const float4 v0 = __ldcs(in4 + k);
                    const int hit = v0.x != 0.f ? 0 : v0.y != 0.f ? 1 : v0.z != 0.f ? 2 : v0.w != 0.f ? 3 : -1;
                    if (hit >= 0) first = min(first, 4 * k + hit);
                }
                const float4 v2 = inplace ? in4[2 * q + k] : __ldcs(in4 + 2 * q + k);
                covered += (v2.x < 1.0f) + (v2.y < 1.0f) + (v2.z < 1.0f) + (v2.w < 1.0f);
                if (!inplace) {
                    const float4 v1 = __ldcs(in4 + q + k), v3 = __ldcs(in4 + 3 * q + k);
                    __stcs(out4 + k, z);
                    __stcs(out4 + q + k, v1);
                    __stcs(out4 + 2 * q + k, v2);
                    __stcs(out4 + 3 * q + k, v3);
                }
            }
        } else {
            for (int k = tid; k < cells; k += SIM_THREADS) {
                if (!p.pos_xy && in[k] != 0.0f) first = min(first, k);
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
        __syncthreads();                                                  // s_first initialised
        if (!p.pos_xy) {
            first = __reduce_min_sync(0xffffffffu, first);
            if ((tid & 31) == 0 && first != 0x7fffffff) atomicMin(&s_first, first);   // shared memory
        }
        covered = block_sum(covered, s_red);     // barriers inside: the bulk stores precede the patches

        // ---- the move (sim_env.py:92-143) -------------------------------------------------------------
        if (tid == 0) {
            const int idx = s_first == 0x7fffffff ? 0 : s_first;
            int x = idx / L, y = idx - x * L;
            const int a = p.actions[b] & 3;
            const int nx = x + (a == 0 ? -1 : a == 1 ? 1 : 0);
            const int ny = y + (a == 2 ? 1 : a == 3 ? -1 : 0);
            double r = 0.0;
            if (nx >= 0 && nx < W && ny >= 0 && ny < L && !get_bit(wall, g.Wp, nx, ny)) {
                x = nx; y = ny;
                if (p.dist_reward) r += (double)in[3 * (size_t)cells + x * L + y];
            } else {
                r -= p.cp;
            }
            s_pos[0] = x; s_pos[1] = y;
            s_reward = r;
        }
        __syncthreads();
        // ---- footprint classes in the order of the reference's if / elif chain (sim_env.py:150-170) ------
        const int px = s_pos[0], py = s_pos[1];
        for (int k = tid; k < S * S; k += SIM_THREADS) {
            const int j = px - p.R + k / S, kk = py - p.R + k % S;
            int cls = CELL_NONE, val = 0;
            if (j >= 0 && j < W && kk >= 0 && kk < L) {
                if (!get_bit(wall, g.Wp, j, kk)) {
                    const float f = in[2 * (size_t)cells + j * L + kk];
                    if (f == 1.0f) { cls = CELL_NEW; val = get_bit(posm, g.Wp, j, kk); }
                    else if (f == 0.0f) cls = CELL_AGAIN;
                } else if (in[(size_t)cells + j * L + kk] == 0.0f) {
                    cls = CELL_OBST;
                }
            }
            s_cls[k] = (uint8_t)cls;
            s_val[k] = (uint8_t)val;
        }
        __syncthreads();
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
            double r = s_reward;         // float64 accumulation in the reference's raster order
            int newly = 0;
            for (int k = 0; k < S * S; ++k) {
                const int cls = s_cls[k];
                if (cls == CELL_NEW) {                       // reward += grid[j][k] (sim_env.py:159)
                    r += p.cell_values ? __ldg(p.cell_values + ((size_t)mi * W + (px - p.R + k / S)) * L + (py - p.R + k % S))
                                       : (double)s_val[k];
                    ++newly;
                }
                else if (cls == CELL_AGAIN) r -= p.free_penalty;
            }
            if (inplace && s_first != 0x7fffffff) out[s_first] = 0.0f;
            out[px * L + py] = 1.0f;
            // percent_covered (sim_env.py:248-259) and the terminal bonus (sim_env.py:187-188)
            const int total = __ldg(p.free_counts + mi);
            const double pct = (double)(covered + newly) / (double)total;
            const bool term = fmin(p.done_thresh, 1.0) <= pct;
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

// Work item = (frame f, cell row x): the s output rows i in [x*s, x*s+s) are identical (INTER_NEAREST with
// an integer factor replicates pixels), so the L cell colours of the row are composed once into shared
// memory and then expanded through a per-column cell table that is the same for every row of every frame.
// Item x == W is the zero border below the map (copyMakeBorder).  Rows are written as 32-bit words of the
// flat byte stream (row starts are not 4-byte aligned in general: head / tail bytes go out singly).
constexpr int RENDER_THREADS = 256;

// One output row = the expanded row image in shared memory (`rowbuf`, or zeros), shifted to the row's own
// 4-byte phase: word k of the row is bytes [head + 4k, head + 4k + 4) of the image, i.e. a funnel shift of two
// aligned shared-memory words.
__device__ __forceinline__ void render_store_row(uint8_t* out, long long row_start, int row_bytes,
                                                 const uint8_t* rowbuf, bool zero) {
    const int head = (int)((4 - (row_start & 3)) & 3);
    for (int r = threadIdx.x; r < head && r < row_bytes; r += RENDER_THREADS)
        out[row_start + r] = zero ? (uint8_t)0 : rowbuf[r];
    const int words = (row_bytes - head) >> 2;
    uint32_t* out32 = reinterpret_cast<uint32_t*>(out + row_start + head);
    const uint32_t* rb32 = reinterpret_cast<const uint32_t*>(rowbuf);
    const int sh = 8 * head;                                   // head < 4: image words are read at byte phase `head`
    for (int w = threadIdx.x; w < words; w += RENDER_THREADS) {
        uint32_t v = 0;
        if (!zero) v = head ? __funnelshift_r(rb32[w], rb32[w + 1], sh) : rb32[w];
        __stcs(out32 + w, v);
    }
    for (int r = head + 4 * words + threadIdx.x; r < row_bytes; r += RENDER_THREADS)
        out[row_start + r] = zero ? (uint8_t)0 : rowbuf[r];
}

__global__ void __launch_bounds__(RENDER_THREADS) render_kernel(RenderParams p) {
    extern __shared__ __align__(16) unsigned char rsm[];
    uint16_t* ymap = reinterpret_cast<uint16_t*>(rsm);                     // [out_w] column -> cell y (0xffff: border)
    uint8_t* col = rsm + ((2 * p.out_w + 15) & ~15);                       // [L][3] colours of the current cell row
    uint8_t* rowbuf = col + ((3 * p.g.L + 15) & ~15);                      // [row_bytes + 8] the expanded pixel row
    const int W = p.g.W, L = p.g.L, s = p.scale;
    for (int j = threadIdx.x; j < p.out_w; j += RENDER_THREADS) {
        const int jj = p.flip ? p.out_w - 1 - j : j;                       // cv2.flip(image, 1), dec_grid_rl.py:593
        ymap[j] = jj < L * s ? (uint16_t)(jj / s) : (uint16_t)0xffffu;
    }
    const int row_bytes = p.out_w * 3;
    const int items = p.n * (W + 1);
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int f = item / (W + 1), x = item - f * (W + 1);
        const int e = p.env_ids ? p.env_ids[f] : f;
        const long long frame0 = (long long)f * p.out_h * row_bytes;
        __syncthreads();                                                   // ymap ready / previous row image consumed
        if (x == W) {
            for (int i = W * s; i < p.out_h; ++i)
                render_store_row(p.out, frame0 + (long long)i * row_bytes, row_bytes, rowbuf, true);
            continue;
        }
        const int16_t* pos = reinterpret_cast<const int16_t*>(p.hdr + (size_t)e * p.hdr_stride + p.off_pos);
        for (int y = threadIdx.x; y < L; y += RENDER_THREADS) {
            const int w = (y >> 5) * p.g.Wp + x;
            const uint32_t bit = 1u << (y & 31);
            uint32_t obst = 0, cov;
            if (p.env_type == MCOV_ENV_DEC) {
                cov = p.plane_a[(size_t)e * p.g.PW + w] & bit;                           // _visited
                const uint2* fo = p.fo + (size_t)e * p.N * p.g.PW;
                for (int r = 0; r < p.N; ++r) obst |= fo[(size_t)r * p.g.PW + w].y & bit;   // union of obst_i
            } else {
                cov = p.plane_a[(size_t)e * p.g.PW + w] & bit;                           // 1 - _free
                obst = p.plane_b[(size_t)e * p.g.PW + w] & bit;                          // _observed_obstacles
            }
            uint32_t robot = 0;
            for (int r = 0; r < p.n_shown; ++r) robot |= (pos[2 * r] == x && pos[2 * r + 1] == y);
            col[3 * y + 0] = (uint8_t)((obst ? 200u : 0u) + (robot ? 255u : 0u));
            col[3 * y + 1] = (uint8_t)(cov ? 225u : 0u);
            col[3 * y + 2] = (uint8_t)((obst ? 255u : 0u) + (cov ? 255u : 0u));
        }
        __syncthreads();
        // expand once: pixel j takes the colour of cell ymap[j]; the s output rows then only copy
        for (int j = threadIdx.x; j < p.out_w + 3; j += RENDER_THREADS) {
            const uint32_t c = j < p.out_w ? ymap[j] : 0xffffu;
            const bool in = c != 0xffffu;
            rowbuf[3 * j + 0] = in ? col[3 * c + 0] : (uint8_t)0;
            rowbuf[3 * j + 1] = in ? col[3 * c + 1] : (uint8_t)0;
            rowbuf[3 * j + 2] = in ? col[3 * c + 2] : (uint8_t)0;
        }
        __syncthreads();
        for (int i = x * s; i < x * s + s; ++i)
            render_store_row(p.out, frame0 + (long long)i * row_bytes, row_bytes, rowbuf, false);
    }
}

}  // namespace

cudaError_t launch_sim_lookahead(const SimParams& p, int num_sms, cudaStream_t st) {
    int grid = num_sms * MCOV_SIM_MINB;
    if (grid > p.B) grid = p.B;
    sim_lookahead_kernel<<<grid, SIM_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_render(const RenderParams& p, int num_sms, cudaStream_t st) {
    const int items = p.n * (p.g.W + 1);
    int grid = num_sms * 8;
    if (grid > items) grid = items;
    const size_t smem = ((2 * (size_t)p.out_w + 15) & ~(size_t)15) + ((3 * (size_t)p.g.L + 15) & ~(size_t)15) +
                        3 * ((size_t)p.out_w + 3) + 16;
    render_kernel<<<grid, RENDER_THREADS, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace mcov

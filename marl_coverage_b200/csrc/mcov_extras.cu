// mcov_extras.cu — gated DecGridRL layers and reward term (config switches of the reference):
//   dist_reward    get_distance_map + reward term   Environments/dec_grid_rl.py:260-282, 222-223, 240, 349-352
//   dijkstra_input dijkstra_path_map                Environments/dijkstra.py:112-187 via dec_grid_rl.py:354-358
//   mini_map_rad   cv2.resize(INTER_LINEAR) layers  Environments/dec_grid_rl.py:360-370
//
// One CTA per env, robots in index order (the float32 reward term is a sequential sum over robots).
// The reference runs these on maps padded by pad = max(egoradius, mini_map_rad) cells of "unknown";
// here nothing is padded: for the L1 transform a cell outside the map is exactly
// d(clamped cell) + taxicab distance to the clamped cell, and a BFS through known-free cells can
// never leave the obstacle-bordered map.
#include "mcov_internal.h"
#include "mcov_dt.cuh"

namespace mcov {

namespace {

constexpr int XT = 256;
constexpr unsigned FULL = 0xffffffffu;
constexpr uint32_t NO_GOAL = 0xffffffffu;

struct ExtrasParams {
    int E, N;
    Geometry g;
    int ego, ow, mini, pad, C;
    int dist, dijkstra;
    int pre_sense;               // 1: only the float32 distance reward term, from the pre-sensing memory
    const unsigned char* hdr;
    int hdr_stride, off_pos;
    const uint2* fo;
    float* obs;
    float* dist_term;            // [E]
    unsigned char* scratch;      // null -> shared memory
    int pitch;                   // uint16 per x row of the level / distance buffer
    size_t work_bytes;           // bytes of one CTA's work area
};

struct ExtrasSmem {
    int red0;
    uint32_t goal;
    int flag;
};

// distance value at map coordinates (x, y), possibly outside the map (the reference's padding)
__device__ __forceinline__ int padded_distance(const uint16_t* d, int pitch, int W, int L, int x, int y) {
    const int cx = min(max(x, 0), W - 1), cy = min(max(y, 0), L - 1);
    return (int)d[(size_t)cx * pitch + cy] + abs(x - cx) + abs(y - cy);
}

__global__ void __launch_bounds__(XT)
dec_extras_kernel(const ExtrasParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ ExtrasSmem sm;
    const int W = p.g.W, L = p.g.L, Wp = p.g.Wp, strips = p.g.strips, PW = p.g.PW;
    const int N = p.N, ego = p.ego, ow = p.ow, pad = p.pad, pitch = p.pitch;
    const int lastbits = L & 31;
    // dynamic shared memory: [transform scratch ints][work area unless it lives in global scratch]
    int* red = reinterpret_cast<int*>(smem_raw);
    const int head_bytes = (dt_scratch_ints(L, XT) * 4 + 15) & ~15;
    unsigned char* work = p.scratch ? p.scratch + (size_t)blockIdx.x * p.work_bytes : smem_raw + head_bytes;
    uint16_t* dbuf = reinterpret_cast<uint16_t*>(work);                       // [W][pitch]
    uint32_t* planes = reinterpret_cast<uint32_t*>(work + (((size_t)W * pitch * 2 + 15) & ~(size_t)15));
    uint32_t* visited = planes;
    uint32_t* cur = planes + PW;
    uint32_t* nxt = planes + 2 * PW;

    for (int e = blockIdx.x; e < p.E; e += gridDim.x) {
        const unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
        const uint32_t* hpos = reinterpret_cast<const uint32_t*>(hrec + p.off_pos);
        float acc = 0.0f;
        for (int i = 0; i < N; ++i) {
            const uint32_t pv = hpos[i];
            const int xi = unpack_x(pv), yi = unpack_y(pv);
            const uint2* plane = p.fo + ((size_t)e * N + i) * PW;
            float* obs_i = p.obs ? p.obs + (((size_t)e * N + i) * p.C) * ow * ow : nullptr;
            __syncthreads();

            const bool want_dist = p.dist && (p.pre_sense || (p.mini == 0 && !p.dijkstra));
            if (want_dist) {
                // zero pixels = cells this robot knows to be free (dec_grid_rl.py:271-274)
                const DtResult r = l1_transform(
                    [&](int ss, int x) -> uint32_t {
                        const uint32_t valid = (ss == strips - 1 && lastbits) ? low_mask(lastbits) : FULL;
                        return plane[ss * Wp + x].x & valid;
                    }, p.g, dbuf, pitch, red, whole_cta());
                // max over the PADDED image: edge cells push their distance outwards by `pad`
                int mx = r.dmax;
                if (r.any_zero && pad > 0) {
                    int m2 = 0;
                    for (int k = threadIdx.x; k < 2 * (W + L); k += XT) {
                        int x, y;
                        if (k < 2 * W) { x = k >> 1; y = (k & 1) ? L - 1 : 0; }
                        else { const int t = k - 2 * W; y = t >> 1; x = (t & 1) ? W - 1 : 0; }
                        const int ex = (x == 0 || x == W - 1) ? pad : 0, ey = (y == 0 || y == L - 1) ? pad : 0;
                        m2 = max(m2, (int)dbuf[(size_t)x * pitch + y] + ex + ey);
                    }
                    m2 = __reduce_max_sync(FULL, m2);
                    if (threadIdx.x == 0) sm.red0 = 0;
                    __syncthreads();
                    if ((threadIdx.x & 31) == 0) atomicMax(&sm.red0, m2);
                    __syncthreads();
                    mx = max(mx, sm.red0);
                }
                if (p.pre_sense) {
                    // quirk A.3.11: the padded map is indexed with UNpadded coordinates; float32 sum
                    if (threadIdx.x == 0) {
                        const int dv = padded_distance(dbuf, pitch, W, L, xi - pad, yi - pad);
                        acc = __fadd_rn(acc, dist_value(dv, r.any_zero, mx));
                    }
                } else {
                    for (int k = threadIdx.x; k < ow * ow; k += XT) {
                        const int a = k / ow, b = k - a * ow;
                        const int dv = padded_distance(dbuf, pitch, W, L, xi - ego + a, yi - ego + b);
                        obs_i[3 * ow * ow + k] = dist_value(dv, r.any_zero, mx);
                    }
                }
            }
            if (p.pre_sense) continue;

            if (p.dijkstra && p.mini == 0) {
                // ---- level-synchronous BFS on bit planes (dijkstra.py:112-160): expand through known
                // free cells until the first level that holds an unknown cell; the goal is the
                // lexicographically smallest (x, y) of that level (PriorityQueue order (cost,(x,y)))
                for (int w = threadIdx.x; w < PW; w += XT) { visited[w] = 0; cur[w] = 0; }
                if (threadIdx.x == 0) sm.goal = NO_GOAL;
                __syncthreads();
                if (threadIdx.x == 0) {
                    const int w = (yi >> 5) * Wp + xi;
                    cur[w] = visited[w] = 1u << (yi & 31);
                    dbuf[(size_t)xi * pitch + yi] = 0;
                }
                __syncthreads();
                int level = 0;
                bool found = false;
                while (true) {
                    for (int idx = threadIdx.x; idx < W * strips; idx += XT) {
                        const int s = idx / W, x = idx - s * W, w = s * Wp + x;
                        const uint2 f = plane[w];
                        const uint32_t cand = cur[w] & ~f.x & ~f.y;
                        if (cand) atomicMin(&sm.goal, ((uint32_t)x << 16) | (uint32_t)(32 * s + __ffs(cand) - 1));
                    }
                    __syncthreads();
                    if (sm.goal != NO_GOAL) { found = true; break; }
                    int any = 0;
                    for (int idx = threadIdx.x; idx < W * strips; idx += XT) {
                        const int s = idx / W, x = idx - s * W, w = s * Wp + x;
                        const uint32_t c = cur[w];
                        uint32_t n = (c << 1) | (c >> 1);
                        if (s > 0) n |= cur[w - Wp] >> 31;
                        if (s + 1 < strips) n |= cur[w + Wp] << 31;
                        if (x > 0) n |= cur[w - 1];
                        if (x + 1 < W) n |= cur[w + 1];
                        const uint32_t valid = (s == strips - 1 && lastbits) ? low_mask(lastbits) : FULL;
                        n &= ~plane[w].y & ~visited[w] & valid;
                        nxt[w] = n;
                        any |= (n != 0u);
                        uint32_t t = n;
                        while (t) {
                            const int b = __ffs(t) - 1;
                            t &= t - 1;
                            dbuf[(size_t)x * pitch + 32 * s + b] = (uint16_t)(level + 1);
                        }
                    }
                    any = __syncthreads_or(any);
                    if (!any) break;
                    for (int idx = threadIdx.x; idx < W * strips; idx += XT) {
                        const int s = idx / W, x = idx - s * W, w = s * Wp + x;
                        visited[w] |= nxt[w];
                    }
                    uint32_t* t = cur; cur = nxt; nxt = t;
                    ++level;
                    __syncthreads();
                }
                // ---- back-trace (dijkstra.py:162-187): first neighbour in order x+1, x-1, y+1, y-1
                // that was finalised one level earlier; path cells inside the window become 1.0
                if (found && threadIdx.x == 0) {
                    int x = (int)(sm.goal >> 16), y = (int)(sm.goal & 0xffffu), c = level;
                    while (true) {
                        const int a = x - (xi - ego), b = y - (yi - ego);
                        if (a >= 0 && a < ow && b >= 0 && b < ow) obs_i[3 * ow * ow + a * ow + b] = 1.0f;
                        if (c == 0) break;
                        const int dx[4] = {1, -1, 0, 0}, dy[4] = {0, 0, 1, -1};
                        bool stepped = false;
                        for (int k = 0; k < 4 && !stepped; ++k) {
                            const int u = x + dx[k], v = y + dy[k];
                            if (u < 0 || u >= W || v < 0 || v >= L) continue;
                            if (!((visited[(v >> 5) * Wp + u] >> (v & 31)) & 1u)) continue;
                            if ((int)dbuf[(size_t)u * pitch + v] != c - 1) continue;
                            x = u; y = v; --c;
                            stepped = true;
                        }
                        if (!stepped) break;                       // cannot happen on a consistent BFS
                    }
                }
                __syncthreads();
            }

            if (p.mini > 0) {
                // ---- mini-map layers 3 and 4 (dec_grid_rl.py:360-370): (2m+1)^2 windows of free_i and
                // obst_i, bilinear to (2e+1)^2 with OpenCV's pixel-centre convention, float64 products
                const int m = p.mini, n_in = 2 * m + 1;
                const double scale = (double)n_in / (double)ow;
                for (int k = threadIdx.x; k < 2 * ow * ow; k += XT) {
                    const int layer = k / (ow * ow), rem = k - layer * ow * ow;
                    const int r = rem / ow, c = rem - r * ow;
                    int ir, ic;
                    double fr, fc;
                    {
                        double f = ((double)r + 0.5) * scale - 0.5;
                        int s = (int)floor(f);
                        f -= (double)s;
                        if (s < 0) { s = 0; f = 0.0; }
                        if (s >= n_in - 1) { s = n_in - 1; f = 0.0; }
                        ir = s; fr = (double)(float)f;
                        f = ((double)c + 0.5) * scale - 0.5;
                        s = (int)floor(f);
                        f -= (double)s;
                        if (s < 0) { s = 0; f = 0.0; }
                        if (s >= n_in - 1) { s = n_in - 1; f = 0.0; }
                        ic = s; fc = (double)(float)f;
                    }
                    const int ir1 = min(ir + 1, n_in - 1), ic1 = min(ic + 1, n_in - 1);
                    auto src = [&](int j, int kk) -> double {
                        const int x = xi - m + j, y = yi - m + kk;
                        if (x < 0 || x >= W || y < 0 || y >= L) return 0.0;
                        const uint2 f = plane[(y >> 5) * Wp + x];
                        return (double)(((layer ? f.y : f.x) >> (y & 31)) & 1u);
                    };
                    const double gr = 1.0 - fr, gc = 1.0 - fc;
                    const double rowA = __dadd_rn(__dmul_rn(src(ir, ic), gr), __dmul_rn(src(ir1, ic), fr));
                    const double rowB = __dadd_rn(__dmul_rn(src(ir, ic1), gr), __dmul_rn(src(ir1, ic1), fr));
                    const double v = __dadd_rn(__dmul_rn(rowA, gc), __dmul_rn(rowB, fc));
                    obs_i[(3 + layer) * ow * ow + rem] = (float)v;
                }
            }
        }
        if (p.pre_sense && threadIdx.x == 0) p.dist_term[e] = acc;
    }
}

struct ExtrasShape {
    int grid, pitch;
    size_t work_bytes, smem;
    bool global_work;
};

inline ExtrasShape extras_shape(const mcov_env* env) {
    ExtrasShape s;
    s.pitch = dt_pitch_of(env->g);
    const size_t lev = ((size_t)env->g.W * s.pitch * 2 + 15) & ~(size_t)15;
    s.work_bytes = lev + (size_t)3 * env->g.PW * 4;
    s.work_bytes = (s.work_bytes + 255) & ~(size_t)255;
    const size_t head = ((size_t)dt_scratch_ints(env->g.L, XT) * 4 + 15) & ~(size_t)15;
    s.global_work = s.work_bytes + head > 200 * 1024;
    s.smem = head + (s.global_work ? 0 : s.work_bytes);
    int per_sm = s.global_work ? 2 : (int)((220 * 1024) / (s.smem + 2048));
    if (per_sm > 8) per_sm = 8;
    if (per_sm < 1) per_sm = 1;
    int grid = env->num_sms * per_sm;
    if (grid > env->cfg.num_envs) grid = env->cfg.num_envs;
    s.grid = grid;
    return s;
}

}  // namespace

int64_t dec_extras_scratch_bytes(const mcov_env* env) {
    const mcov_config& c = env->cfg;
    if (!(c.dist_reward || c.dijkstra_input || c.mini_map_rad > 0)) return 0;
    const ExtrasShape s = extras_shape(env);
    return s.global_work ? (int64_t)s.grid * (int64_t)s.work_bytes : 0;
}

cudaError_t launch_dec_extras(mcov_env* env, float* obs, float* dist_term, unsigned char* scratch,
                              int pre_sense, cudaStream_t st) {
    const mcov_config& c = env->cfg;
    const mcov_layout& l = env->lay;
    const ExtrasShape s = extras_shape(env);
    ExtrasParams p;
    p.E = c.num_envs;
    p.N = c.num_robots;
    p.g = env->g;
    p.ego = c.egoradius;
    p.ow = 2 * c.egoradius + 1;
    p.mini = c.mini_map_rad;
    p.pad = c.egoradius > c.mini_map_rad ? c.egoradius : c.mini_map_rad;     // dec_grid_rl.py:78
    p.C = env->obs_C;
    p.dist = c.dist_reward;
    p.dijkstra = c.dijkstra_input;
    p.pre_sense = pre_sense;
    p.hdr = env->arena + l.off_hdr;
    p.hdr_stride = l.hdr_stride;
    p.off_pos = l.hdr_off_pos;
    p.fo = reinterpret_cast<const uint2*>(env->arena + l.off_fo);
    p.obs = obs;
    p.dist_term = dist_term;
    p.scratch = s.global_work ? scratch : nullptr;
    p.pitch = s.pitch;
    p.work_bytes = s.work_bytes;
    if (s.global_work && !scratch) return cudaErrorInvalidValue;
    if (s.smem > 40 * 1024) {
        cudaError_t err = cudaFuncSetAttribute(dec_extras_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)s.smem);
        if (err != cudaSuccess) return err;
    }
    dec_extras_kernel<<<s.grid, XT, s.smem, st>>>(p);
    env->launches++;
    return cudaGetLastError();
}

}  // namespace mcov

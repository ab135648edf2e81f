// mcov_api.cu — extern "C" boundary of libmcov_b200.so (see include/mcov.h).
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <new>

#include "mcov_internal.h"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    return fail(MCOV_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

int64_t align_up64(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

int obs_channels(const mcov_config& c) {
    if (c.env_type == MCOV_ENV_SUPER) return c.use_scanning ? 4 : c.num_robots + 3;
    int layers = c.mini_map_rad > 0 ? 5 : 3;          // dec_grid_rl.py:322-332
    if (c.dist_reward) layers += 1;
    if (c.dijkstra_input) layers += 1;
    return layers;
}

// ---- map packer --------------------------------------------------------------------------------
// One thread per output word; 32 cells along y.  cells: int8 [n][W][L]; output geometry is
// (W + 2b, L + 2b) with b = add_border (obstacle ring, dec_grid_rl.py:471-472).
__global__ void pack_maps_kernel(const int8_t* __restrict__ cells, int n_maps, int W, int L, int b,
                                 int strips, int Wp, uint32_t* __restrict__ wall,
                                 uint32_t* __restrict__ pos) {
    const int PW = strips * Wp;
    const long long total = (long long)n_maps * PW;
    const int Wo = W + 2 * b, Lo = L + 2 * b;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / PW);
        const int w = (int)(t % PW);
        const int s = w / Wp, x = w % Wp;
        uint32_t wb = 0, pb = 0;
        if (x < Wo) {
            for (int k = 0; k < 32; ++k) {
                const int y = s * 32 + k;
                if (y >= Lo) break;
                int v = -1;                                           // border cell
                const int xi = x - b, yi = y - b;
                if (xi >= 0 && xi < W && yi >= 0 && yi < L) v = cells[((size_t)m * W + xi) * L + yi];
                wb |= (uint32_t)(v < 0) << k;
                pb |= (uint32_t)(v > 0) << k;
            }
        }
        wall[t] = wb;
        pos[t] = pb;
    }
}

// gridgen on the device: same word loop as the packer, the cell value drawn instead of read.  Cell (m, xi, yi)
// takes lane yi & 3 of Philox(counter = (m, xi, yi >> 2, tag), key = seed): independent of n_maps and layout.
__global__ void gen_maps_kernel(int n_maps, int W, int L, int b, int strips, int Wp, unsigned long long thresh,
                                uint2 key, uint32_t* __restrict__ wall, uint32_t* __restrict__ pos) {
    const int PW = strips * Wp;
    const long long total = (long long)n_maps * PW;
    const int Wo = W + 2 * b, Lo = L + 2 * b;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / PW);
        const int w = (int)(t % PW);
        const int s = w / Wp, x = w % Wp;
        uint32_t wb = 0, pb = 0;
        if (x < Wo) {
            const int xi = x - b;
            int have = -1;                                            // yi >> 2 of the draw held in r
            uint4 r = make_uint4(0u, 0u, 0u, 0u);
            for (int k = 0; k < 32; ++k) {
                const int y = s * 32 + k;
                if (y >= Lo) break;
                const int yi = y - b;
                bool obst = true;                                     // border ring
                if (xi >= 0 && xi < W && yi >= 0 && yi < L) {
                    if ((yi >> 2) != have) {
                        have = yi >> 2;
                        r = mcov::philox4x32_10(make_uint4((uint32_t)m, (uint32_t)xi, (uint32_t)have, 0x47454E4Du), key);
                    }
                    const uint32_t u = (yi & 3) == 0 ? r.x : (yi & 3) == 1 ? r.y : (yi & 3) == 2 ? r.z : r.w;
                    obst = (unsigned long long)u < thresh;
                }
                wb |= (uint32_t)obst << k;
                pb |= (uint32_t)(!obst) << k;
            }
        }
        wall[t] = wb;
        pos[t] = pb;
    }
}

__global__ void count_free_kernel(const uint32_t* __restrict__ pos, int n_maps, int PW, int32_t* out) {
    __shared__ int red[32];
    const int m = blockIdx.x;
    int c = 0;
    for (int w = threadIdx.x; w < PW; w += blockDim.x) c += __popc(pos[(size_t)m * PW + w]);
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += red[k];
        out[m] = s;
    }
}

int64_t scratch_stats_bytes(const mcov_env* env) {
    return align_up64((int64_t)env->num_sms * 32 * MCOV_STATS_LEN * 8, 256);
}
unsigned char* scratch_counter(const mcov_env* env) {
    return env->arena + env->lay.off_scratch + scratch_stats_bytes(env);
}
unsigned char* scratch_dist_term(const mcov_env* env) { return scratch_counter(env) + 256; }
unsigned char* scratch_moves(const mcov_env* env) {
    return scratch_dist_term(env) + align_up64((int64_t)env->cfg.num_envs * 4, 256);
}
unsigned char* scratch_super_dt(const mcov_env* env) {
    return scratch_moves(env) + align_up64((int64_t)env->cfg.num_envs * 8, 256);
}
unsigned char* scratch_extras(const mcov_env* env) { return scratch_super_dt(env); }   // Dec: same slot

// one handle <-> one device: its buffers, map table and kernel attributes belong to the device that was current
// at mcov_create, so a call with another device current is refused instead of launching there
int check_device(const mcov_env* env) {
    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaGetDevice");
    if (dev != env->device)
        return fail(MCOV_ERR_STATE, "handle was created on device %d, the current device is %d", env->device, dev);
    return MCOV_OK;
}

int sm_count() {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms > 0 ? sms : 148;
}

int check_ready(const mcov_env* env) {
    if (!env) return fail(MCOV_ERR_NULL, "env handle is NULL");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    if (!env->bound) return fail(MCOV_ERR_STATE, "mcov_bind_state has not been called");
    return check_device(env);
}

void fill_dec_params(mcov_env* env, mcov::DecParams& p) {
    const mcov_config& c = env->cfg;
    const mcov_layout& l = env->lay;
    memset(&p, 0, sizeof(p));
    p.E = c.num_envs;
    p.N = c.num_robots;
    p.g = env->g;
    p.R = c.sensor_range;
    p.S = 2 * c.sensor_range + 1;
    p.ego = c.egoradius;
    p.ow = 2 * c.egoradius + 1;
    p.C = env->obs_C;
    p.sensor = c.sensor_type;
    p.n_lasers = c.num_lasers;
    p.beams = env->beams_dev;
    p.lidar_cells = env->lidar_cells_dev;
    p.lidar_runs = env->lidar_runs_dev;
    p.single_tool = c.single_square_tool;
    p.comm_radius = c.comm_radius;
    p.want_adj = 1;
    p.dist_reward = c.dist_reward;
    p.maxsteps = c.maxsteps;
    p.truncate_steps = c.truncate_steps;
    p.auto_reset = c.auto_reset;
    p.env_id_base = c.env_id_base;
    p.cp = c.collision_penalty;
    p.terminal = c.terminal_reward;
    p.done_incr = c.done_incr;
    p.wall = env->wall;
    p.pos = env->pos;
    p.free_counts = env->free_counts;
    p.n_maps = env->n_maps;
    p.hdr = env->arena + l.off_hdr;
    p.hdr_stride = l.hdr_stride;
    p.off_pos = l.hdr_off_pos;
    p.off_cov = l.hdr_off_cov;
    p.off_adj = l.hdr_off_adj;
    p.visited = reinterpret_cast<uint32_t*>(env->arena + l.off_visited);
    p.fo = reinterpret_cast<uint2*>(env->arena + l.off_fo);
    p.needs_reset = env->arena + l.off_needs_reset;
    p.obs_env_floats = p.N * p.C * p.ow * p.ow;
}

void fill_super_params(mcov_env* env, mcov::SuperParams& p) {
    const mcov_config& c = env->cfg;
    const mcov_layout& l = env->lay;
    memset(&p, 0, sizeof(p));
    p.E = c.num_envs;
    p.N = c.num_robots;
    p.g = env->g;
    p.R = c.senseradius;
    p.C = env->obs_C;
    p.use_scanning = c.use_scanning;
    p.dist_reward = c.dist_reward;
    p.truncate_steps = c.truncate_steps;
    p.auto_reset = c.auto_reset;
    p.env_id_base = c.env_id_base;
    p.cp = c.collision_penalty;
    p.terminal = c.terminal_reward;
    p.done_incr = c.done_incr;
    p.free_penalty = c.free_penalty;
    p.wall = env->wall;
    p.pos = env->pos;
    p.free_counts = env->free_counts;
    p.cell_values = env->cell_values;
    p.n_maps = env->n_maps;
    p.hdr = env->arena + l.off_hdr;
    p.hdr_stride = l.hdr_stride;
    p.off_pos = l.hdr_off_pos;
    p.off_cov = l.hdr_off_cov;
    p.covered = reinterpret_cast<uint32_t*>(env->arena + l.off_covered);
    p.seen_obst = reinterpret_cast<uint32_t*>(env->arena + l.off_seen_obst);
    p.needs_reset = env->arena + l.off_needs_reset;
}

}  // namespace

extern "C" {

int mcov_abi_version(void) { return MCOV_ABI_VERSION; }

const char* mcov_last_error(void) { return g_err; }

int mcov_plane_geometry(int W, int L, int* strips_out, int* wp_out, int* words_out) {
    if (W <= 0 || L <= 0) return fail(MCOV_ERR_BAD_CONFIG, "map size %d x %d", W, L);
    const mcov::Geometry g = mcov::make_geometry(W, L);
    if (strips_out) *strips_out = g.strips;
    if (wp_out) *wp_out = g.Wp;
    if (words_out) *words_out = g.PW;
    return MCOV_OK;
}

int mcov_create(const mcov_config* cfg, mcov_env** out) {
    if (!cfg || !out) return fail(MCOV_ERR_NULL, "cfg/out is NULL");
    if (cfg->struct_size != (int32_t)sizeof(mcov_config))
        return fail(MCOV_ERR_ABI, "mcov_config.struct_size %d != %d (ABI %d)", cfg->struct_size,
                    (int)sizeof(mcov_config), MCOV_ABI_VERSION);
    const mcov_config& c = *cfg;
    if (c.env_type != MCOV_ENV_DEC && c.env_type != MCOV_ENV_SUPER)
        return fail(MCOV_ERR_BAD_CONFIG, "env_type %d", c.env_type);
    if (c.num_envs < 1) return fail(MCOV_ERR_BAD_CONFIG, "num_envs %d < 1", c.num_envs);
    if (c.num_robots < 1 || c.num_robots > MCOV_MAX_ROBOTS)
        return fail(MCOV_ERR_BAD_CONFIG, "numrobot %d outside [1, %d]", c.num_robots, MCOV_MAX_ROBOTS);
    if (c.env_type == MCOV_ENV_DEC) {
        if (c.sensor_type != MCOV_SENSOR_SQUARE && c.sensor_type != MCOV_SENSOR_LIDAR)
            return fail(MCOV_ERR_BAD_CONFIG, "sensor_type %d", c.sensor_type);
        if (c.sensor_range < 0 || c.sensor_range > MCOV_MAX_RADIUS)
            return fail(MCOV_ERR_BAD_CONFIG, "sensor range %d outside [0, %d]", c.sensor_range, MCOV_MAX_RADIUS);
        if (c.egoradius < 0 || c.egoradius > MCOV_MAX_RADIUS)
            return fail(MCOV_ERR_BAD_CONFIG, "egoradius %d outside [0, %d]", c.egoradius, MCOV_MAX_RADIUS);
        if (c.mini_map_rad < 0 || c.mini_map_rad > MCOV_MAX_RADIUS)
            return fail(MCOV_ERR_BAD_CONFIG, "mini_map_rad %d outside [0, %d]", c.mini_map_rad, MCOV_MAX_RADIUS);
        if (c.sensor_type == MCOV_SENSOR_LIDAR) {
            if (c.num_lasers < 1 || c.num_lasers > MCOV_MAX_LASERS || (c.num_lasers % 2) == 0)
                return fail(MCOV_ERR_BAD_CONFIG, "num_lasers %d must be odd and in [1, %d] (lidar.py:11)",
                            c.num_lasers, MCOV_MAX_LASERS);
        }
        if (c.comm_radius < 0) return fail(MCOV_ERR_BAD_CONFIG, "comm_radius %d < 0", c.comm_radius);
    } else {
        if (c.senseradius < 0 || c.senseradius > MCOV_MAX_RADIUS)
            return fail(MCOV_ERR_BAD_CONFIG, "senseradius %d outside [0, %d]", c.senseradius, MCOV_MAX_RADIUS);
    }
    mcov_env* e = new (std::nothrow) mcov_env();
    if (!e) return fail(MCOV_ERR_CUDA, "out of host memory");
    memset(e, 0, sizeof(*e));
    e->cfg = c;
    cudaError_t ce = cudaGetDevice(&e->device);
    if (ce != cudaSuccess) { delete e; return cuda_fail(ce, "cudaGetDevice"); }
    ce = cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, e->device);
    if (ce != cudaSuccess) { delete e; return cuda_fail(ce, "cudaDeviceGetAttribute"); }
    e->obs_C = obs_channels(c);
    e->debug = getenv("MCOV_DEBUG") ? atoi(getenv("MCOV_DEBUG")) : 0;
    auto pin = [](const char* name) { const char* v = getenv(name); return v && *v ? atoi(v) : -1; };
    e->pin_group = pin("MCOV_DEC_GROUP");
    e->pin_store = pin("MCOV_DEC_STORE");
    e->pin_grid = pin("MCOV_DEC_GRID");
    e->pin_nbuf = pin("MCOV_DEC_NBUF");
    e->pin_shape = pin("MCOV_DEC_SHAPE");
    e->pin_sw = pin("MCOV_DEC_SW");
    e->pin_rot = pin("MCOV_DEC_ROT");
    e->pin_tma = pin("MCOV_DEC_TMA");
    e->pin_team = pin("MCOV_DEC_TEAM");
    e->pin_warps = pin("MCOV_DEC_WARPS");
    e->pin_balance = pin("MCOV_DEC_BALANCE");
    e->pin_smt = pin("MCOV_DEC_SMT");
    *out = e;
    return MCOV_OK;
}

int mcov_destroy(mcov_env* env) {
    if (!env) return MCOV_OK;
    if (env->beams_dev) cudaFree(env->beams_dev);
    if (env->lidar_cells_dev) cudaFree(env->lidar_cells_dev);
    if (env->lidar_runs_dev) cudaFree(env->lidar_runs_dev);
    delete env;
    return MCOV_OK;
}

int mcov_set_lidar_beams(mcov_env* env, const double* beams_host, int num_lasers) {
    if (!env || !beams_host) return fail(MCOV_ERR_NULL, "env/beams is NULL");
    if (num_lasers != env->cfg.num_lasers)
        return fail(MCOV_ERR_BAD_CONFIG, "num_lasers %d != config %d", num_lasers, env->cfg.num_lasers);
    if (!env->beams_dev) {
        cudaError_t ce = cudaMalloc(&env->beams_dev, sizeof(double) * 3 * num_lasers);
        if (ce != cudaSuccess) return cuda_fail(ce, "cudaMalloc(beams)");
    }
    cudaError_t ce = cudaMemcpy(env->beams_dev, beams_host, sizeof(double) * 3 * num_lasers,
                                cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaMemcpy(beams)");

    // Integer beam table.  The reference marches cx += xinc in float64 from the robot cell and
    // truncates (lidar.py:50-56).  For integer starts below 2^15 the accumulated rounding error is
    // < 1e-10, so whenever every partial sum k*inc stays further than 1e-7 from an integer (or the
    // increment is itself an integer: exact arithmetic) the visited cells are start + floor(k*inc)
    // for EVERY start, and the march becomes a table walk.  Otherwise the float64 march stays.
    if (env->lidar_cells_dev) { cudaFree(env->lidar_cells_dev); env->lidar_cells_dev = nullptr; }
    if (env->lidar_runs_dev) { cudaFree(env->lidar_runs_dev); env->lidar_runs_dev = nullptr; }
    const int R = env->cfg.sensor_range;
    const char* force = getenv("MCOV_LIDAR_FP64");
    bool ok = !(force && atoi(force) != 0);
    uint16_t* table = ok ? new (std::nothrow) uint16_t[(size_t)num_lasers * 16] : nullptr;
    if (!table) ok = false;
    for (int l = 0; ok && l < num_lasers; ++l) {
        const double xinc = beams_host[3 * l], yinc = beams_host[3 * l + 1], dinc = beams_host[3 * l + 2];
        if (!(dinc >= 1.0)) { ok = false; break; }               // steps shorter than a cell: not this sensor
        const bool xi_exact = xinc == floor(xinc), yi_exact = yinc == floor(yinc);
        double sx = 0.0, sy = 0.0, dist = 0.0;
        int k = 0;
        for (;; ++k) {
            if (k >= 16) { ok = false; break; }
            const double fx = floor(sx), fy = floor(sy);
            const double mx = fmin(sx - fx, fx + 1.0 - sx), my = fmin(sy - fy, fy + 1.0 - sy);
            if ((!xi_exact && k > 0 && mx < 1e-7) || (!yi_exact && k > 0 && my < 1e-7)) { ok = false; break; }
            const int dx = (int)fx, dy = (int)fy;
            if (dx < -R || dx > R || dy < -R || dy > R) { ok = false; break; }
            table[l * 16 + k] = (uint16_t)((dx + R) | ((dy + R) << 8));
            if (!(dist < (double)R)) break;                      // this was the end cell (lidar.py:58-65)
            sx += xinc;
            sy += yinc;
            dist += dinc;
        }
        for (int j = k + 1; ok && j < 16; ++j) table[l * 16 + j] = 0xffffu;
    }
    // The same cells grouped into same-row runs (what the kernels walk): consecutive cells of a beam that stay in one
    // window row become one {mask, row} entry; inside a run y moves strictly one way, so the first wall of the run is
    // the lowest (or highest) set bit of wall & mask and the free cells before it are one more mask.
    uint2* runs = ok ? new (std::nothrow) uint2[(size_t)num_lasers * 16] : nullptr;
    if (ok && !runs) ok = false;
    for (int l = 0; ok && l < num_lasers; ++l) {
        const uint16_t* c = table + (size_t)l * 16;
        int nr = 0, k = 1;
        while (ok && k < 16 && c[k] != 0xffffu) {
            const int rx = c[k] & 0xff;
            int prev = c[k] >> 8, dir = 0;
            uint32_t mask = 1u << prev;
            for (++k; k < 16 && c[k] != 0xffffu && (c[k] & 0xff) == rx; ++k) {
                const int ry = c[k] >> 8;
                const int d = ry > prev ? 1 : (ry < prev ? -1 : 0);
                if (d == 0 || (dir != 0 && d != dir)) { ok = false; break; }   // not monotone: keep the float64 march
                dir = d;
                mask |= 1u << ry;
                prev = ry;
            }
            if (nr >= 15) ok = false;
            if (ok) runs[(size_t)l * 16 + nr++] = make_uint2(mask, (uint32_t)(rx * 4) | (dir < 0 ? 0x80000000u : 0u));
        }
        for (int j = nr; ok && j < 16; ++j) runs[(size_t)l * 16 + j] = make_uint2(0u, 0u);
    }
    if (ok) {
        // device format of a cell: low byte = y bit of the cell in its window row, high byte = byte offset of the row
        // word (4 * row); the slots behind a beam's last cell point at row S, a dummy row the kernels keep all-wall, so
        // the walk needs no end-of-list test: it stops there like at any obstacle
        const int S = 2 * R + 1;
        for (size_t k = 0; k < (size_t)num_lasers * 16; ++k) {
            const uint16_t c = table[k];
            table[k] = c == 0xffffu ? (uint16_t)((4 * S) << 8) : (uint16_t)((c >> 8) | ((4 * (c & 0xff)) << 8));
        }
        ce = cudaMalloc(&env->lidar_cells_dev, sizeof(uint16_t) * 16 * num_lasers);
        if (ce == cudaSuccess)
            ce = cudaMemcpy(env->lidar_cells_dev, table, sizeof(uint16_t) * 16 * num_lasers, cudaMemcpyHostToDevice);
        if (ce == cudaSuccess) ce = cudaMalloc(&env->lidar_runs_dev, sizeof(uint2) * 16 * num_lasers);
        if (ce == cudaSuccess)
            ce = cudaMemcpy(env->lidar_runs_dev, runs, sizeof(uint2) * 16 * num_lasers, cudaMemcpyHostToDevice);
        if (ce != cudaSuccess) { delete[] table; delete[] runs; return cuda_fail(ce, "cudaMalloc/cudaMemcpy(lidar cells)"); }
    }
    delete[] runs;
    delete[] table;
    return MCOV_OK;
}

int mcov_pack_maps(const int8_t* cells_dev, int n_maps, int W, int L, int add_border,
                   uint32_t* wall_out_dev, uint32_t* pos_out_dev, int32_t* free_counts_out_dev,
                   void* stream) {
    if (!cells_dev || !wall_out_dev || !pos_out_dev || !free_counts_out_dev)
        return fail(MCOV_ERR_NULL, "mcov_pack_maps: NULL buffer");
    if (n_maps < 1 || W < 1 || L < 1) return fail(MCOV_ERR_BAD_CONFIG, "n_maps/W/L %d/%d/%d", n_maps, W, L);
    const int b = add_border ? 1 : 0;
    const mcov::Geometry g = mcov::make_geometry(W + 2 * b, L + 2 * b);
    if (g.W > 32767 || g.L > 32767) return fail(MCOV_ERR_BAD_CONFIG, "map larger than int16 coordinates");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long total = (long long)n_maps * g.PW;
    int grid = (int)((total + 255) / 256);
    if (grid > sm_count() * 16) grid = sm_count() * 16;
    pack_maps_kernel<<<grid, 256, 0, st>>>(cells_dev, n_maps, W, L, b, g.strips, g.Wp, wall_out_dev, pos_out_dev);
    count_free_kernel<<<n_maps, 256, 0, st>>>(pos_out_dev, n_maps, g.PW, free_counts_out_dev);
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) return cuda_fail(ce, "pack_maps launch");
    return MCOV_OK;
}

int mcov_gen_maps(int n_maps, int W, int L, int add_border, double prob_obst, uint64_t seed,
                  uint32_t* wall_out_dev, uint32_t* pos_out_dev, int32_t* free_counts_out_dev, void* stream) {
    if (!wall_out_dev || !pos_out_dev || !free_counts_out_dev)
        return fail(MCOV_ERR_NULL, "mcov_gen_maps: NULL buffer");
    if (n_maps < 1 || W < 1 || L < 1) return fail(MCOV_ERR_BAD_CONFIG, "n_maps/W/L %d/%d/%d", n_maps, W, L);
    if (!(prob_obst >= 0.0 && prob_obst <= 1.0))
        return fail(MCOV_ERR_BAD_CONFIG, "prob_obst %g outside [0, 1]", prob_obst);
    const int b = add_border ? 1 : 0;
    const mcov::Geometry g = mcov::make_geometry(W + 2 * b, L + 2 * b);
    if (g.W > 32767 || g.L > 32767) return fail(MCOV_ERR_BAD_CONFIG, "map larger than int16 coordinates");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long total = (long long)n_maps * g.PW;
    int grid = (int)((total + 255) / 256);
    if (grid > sm_count() * 16) grid = sm_count() * 16;
    // a cell is an obstacle iff its 32-bit draw < prob_obst * 2^32 (exact for p = 0 and p = 1)
    const unsigned long long thresh = (unsigned long long)llround(prob_obst * 4294967296.0);
    gen_maps_kernel<<<grid, 256, 0, st>>>(n_maps, W, L, b, g.strips, g.Wp, thresh,
                                          make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), wall_out_dev, pos_out_dev);
    count_free_kernel<<<n_maps, 256, 0, st>>>(pos_out_dev, n_maps, g.PW, free_counts_out_dev);
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) return cuda_fail(ce, "gen_maps launch");
    return MCOV_OK;
}

int mcov_set_maps(mcov_env* env, const uint32_t* wall_dev, const uint32_t* pos_dev,
                  const int32_t* free_counts_dev, int n_maps, int W, int L) {
    if (!env || !wall_dev || !pos_dev || !free_counts_dev) return fail(MCOV_ERR_NULL, "mcov_set_maps: NULL");
    if (n_maps < 1 || W < 1 || L < 1 || W > 32767 || L > 32767)
        return fail(MCOV_ERR_BAD_CONFIG, "n_maps/W/L %d/%d/%d", n_maps, W, L);
    if (!aligned(wall_dev, 16) || !aligned(pos_dev, 16))
        return fail(MCOV_ERR_ALIGN, "map planes must be 16-byte aligned");
    {
        const mcov::Geometry gq = mcov::make_geometry(W, L);
        const unsigned long long words = (unsigned long long)env->cfg.num_envs * env->cfg.num_robots * gq.PW;
        if (env->cfg.env_type == MCOV_ENV_DEC && words >= 0xffffffffull)
            return fail(MCOV_ERR_BAD_CONFIG, "num_envs x numrobot x plane words = %llu exceeds the kernel's 32-bit "
                        "plane offsets: shard the batch over more handles (<= %llu envs of this map size each)",
                        words, 0xfffffffeull / ((unsigned long long)env->cfg.num_robots * gq.PW));
    }
    env->g = mcov::make_geometry(W, L);
    env->wall = wall_dev;
    env->pos = pos_dev;
    env->free_counts = free_counts_dev;
    env->n_maps = n_maps;
    env->maps_set = true;
    env->bound = false;
    env->cell_values = nullptr;                                   // belongs to the previous table

    // ---- arena layout ---------------------------------------------------------------------------
    const mcov_config& c = env->cfg;
    mcov_layout& l = env->lay;
    memset(&l, 0, sizeof(l));
    l.struct_size = (int32_t)sizeof(mcov_layout);
    l.W = W; l.L = L;
    l.strips = env->g.strips; l.Wp = env->g.Wp; l.plane_words = env->g.PW;
    const int N = c.num_robots;
    l.hdr_off_pos = (int32_t)sizeof(mcov::EnvHead);
    l.hdr_off_cov = l.hdr_off_pos + 4 * N;
    l.hdr_off_adj = l.hdr_off_cov + 4 * N;
    l.hdr_stride = mcov::round_up(l.hdr_off_adj + 4 * N, 32);
    const int64_t E = c.num_envs;
    int64_t off = 0;
    l.off_hdr = off; off = align_up64(off + E * l.hdr_stride, 256);
    if (c.env_type == MCOV_ENV_DEC) {
        l.off_visited = off; off = align_up64(off + E * (int64_t)l.plane_words * 4, 256);
        l.off_fo = off; off = align_up64(off + E * N * (int64_t)l.plane_words * 8, 256);
        l.off_covered = l.off_seen_obst = -1;
    } else {
        l.off_covered = off; off = align_up64(off + E * (int64_t)l.plane_words * 4, 256);
        l.off_seen_obst = off; off = align_up64(off + E * (int64_t)l.plane_words * 4, 256);
        l.off_visited = l.off_fo = -1;
    }
    l.off_needs_reset = off; off = align_up64(off + E, 256);
    // scratch: stats partial rows | CTA counter | per-env float32 dist term | per-env move scratch
    l.off_scratch = off;
    off = align_up64(off + scratch_stats_bytes(env), 256);
    off = align_up64(off + 256, 256);
    off = align_up64(off + E * 4, 256);
    off = align_up64(off + E * 8, 256);
    if (c.env_type == MCOV_ENV_SUPER) off = align_up64(off + mcov::super_dt_scratch_bytes(env), 256);
    else off = align_up64(off + mcov::dec_extras_scratch_bytes(env), 256);
    l.total_bytes = off;
    return MCOV_OK;
}

int mcov_set_cell_values(mcov_env* env, const double* values_dev) {
    if (!env) return fail(MCOV_ERR_NULL, "env handle is NULL");
    if (env->cfg.env_type != MCOV_ENV_SUPER)
        return fail(MCOV_ERR_BAD_CONFIG, "cell values only enter SuperGridRL's reward (super_grid_rl.py:189); "
                    "DecGridRL's is a count (dec_grid_rl.py:243-256)");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    if (values_dev && !aligned(values_dev, 8)) return fail(MCOV_ERR_ALIGN, "cell values must be 8-byte aligned");
    env->cell_values = values_dev;
    return MCOV_OK;
}

int mcov_state_layout(const mcov_env* env, mcov_layout* out) {
    if (!env || !out) return fail(MCOV_ERR_NULL, "env/out is NULL");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    if (out->struct_size != 0 && out->struct_size != (int32_t)sizeof(mcov_layout))
        return fail(MCOV_ERR_ABI, "mcov_layout.struct_size %d != %d", out->struct_size, (int)sizeof(mcov_layout));
    *out = env->lay;
    return MCOV_OK;
}

int mcov_bind_state(mcov_env* env, void* arena_dev, size_t bytes, void* stream) {
    if (!env || !arena_dev) return fail(MCOV_ERR_NULL, "env/arena is NULL");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    if (!aligned(arena_dev, 256)) return fail(MCOV_ERR_ALIGN, "state arena must be 256-byte aligned");
    if ((int64_t)bytes < env->lay.total_bytes)
        return fail(MCOV_ERR_BAD_CONFIG, "arena %zu bytes < required %lld", bytes, (long long)env->lay.total_bytes);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    env->arena = static_cast<unsigned char*>(arena_dev);
    cudaError_t ce = cudaMemsetAsync(arena_dev, 0, (size_t)env->lay.total_bytes, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaMemsetAsync(arena)");
    ce = mcov::launch_init_headers(env, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "init_headers launch");
    env->bound = true;
    return MCOV_OK;
}

int mcov_obs_shape(const mcov_env* env, int32_t dims_out[5], int32_t* ndims_out) {
    if (!env || !dims_out || !ndims_out) return fail(MCOV_ERR_NULL, "NULL argument");
    const mcov_config& c = env->cfg;
    if (c.env_type == MCOV_ENV_DEC) {
        dims_out[0] = c.num_envs; dims_out[1] = c.num_robots; dims_out[2] = env->obs_C;
        dims_out[3] = dims_out[4] = 2 * c.egoradius + 1;
        *ndims_out = 5;
    } else {
        if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
        dims_out[0] = c.num_envs; dims_out[1] = env->obs_C; dims_out[2] = env->g.W; dims_out[3] = env->g.L;
        dims_out[4] = 0;
        *ndims_out = 4;
    }
    return MCOV_OK;
}

static int run_dec(mcov_env* env, int mode, const void* actions, int act_fmt, const uint8_t* mask,
                   int mask_from_flags, const int32_t* map_idx, const int16_t* start_xy, uint64_t seed,
                   float* obs, float* reward, double* reward64, uint8_t* done, double* stats,
                   cudaStream_t st) {
    mcov::DecParams p;
    fill_dec_params(env, p);
    p.mode = mode;
    p.actions = actions;
    p.act_fmt = act_fmt;
    p.mask = mask;
    p.mask_from_flags = mask_from_flags;
    p.map_idx_in = map_idx;
    p.start_xy_in = start_xy;
    p.seed = seed;
    p.obs = obs;
    p.reward = reward;
    p.reward64 = reward64;
    p.done = done;
    p.vec_ok = ((p.obs_env_floats * 4) % 16 == 0) && aligned(obs, 16);
    p.phases = mcov::PH_MOVE | mcov::PH_SENSE;
    p.stats_partial = reinterpret_cast<double*>(env->arena + env->lay.off_scratch);
    p.stats_out = (mode == mcov::MODE_STEP) ? stats : nullptr;
    p.counter = reinterpret_cast<unsigned int*>(scratch_counter(env));
    p.move_scratch = reinterpret_cast<uint2*>(scratch_moves(env));
    const mcov_config& c = env->cfg;
    const mcov_layout& l = env->lay;
    const bool extras = c.dist_reward || c.dijkstra_input || c.mini_map_rad > 0;
    cudaError_t ce;
    if (mode == mcov::MODE_RESET && !mask && !mask_from_flags) {
        // whole-batch reset: zero every plane with one memset instead of per-warp wipes
        const size_t nb = (size_t)c.num_envs * l.plane_words * 4;
        ce = cudaMemsetAsync(env->arena + l.off_visited, 0, nb, st);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(env->arena + l.off_fo, 0, nb * 2 * c.num_robots, st);
        if (ce != cudaSuccess) return cuda_fail(ce, "cudaMemsetAsync(planes)");
        p.skip_wipe = 1;
    }
    if (mode == mcov::MODE_STEP && c.map_sharing) {
        ce = mcov::launch_dec_share(env, st);                     // dec_grid_rl.py:124-125
        if (ce != cudaSuccess) return cuda_fail(ce, "dec_share launch");
    }
    if (c.dist_reward && mode == mcov::MODE_STEP) {
        // row D11: the distance term is read from the PRE-sensing memory at the POST-move pose, so
        // the step is split: moves -> per-robot distance maps -> sensing/reward/observation.
        p.phases = mcov::PH_MOVE;
        ce = mcov::launch_dec(env, p, st);
        if (ce != cudaSuccess) return cuda_fail(ce, "dec move launch");
        p.dist_term = reinterpret_cast<float*>(scratch_dist_term(env));
        ce = mcov::launch_dec_extras(env, nullptr, p.dist_term, scratch_extras(env), 1, st);
        if (ce != cudaSuccess) return cuda_fail(ce, "dec dist-term launch");
        p.phases = mcov::PH_SENSE;
    }
    ce = mcov::launch_dec(env, p, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "dec kernel launch");
    if (extras) {
        ce = mcov::launch_dec_extras(env, obs, nullptr, scratch_extras(env), 0, st);
        if (ce != cudaSuccess) return cuda_fail(ce, "dec extras launch");
    }
    return MCOV_OK;
}

static int run_super(mcov_env* env, int mode, const void* actions, int act_fmt, const uint8_t* mask,
                     int mask_from_flags, const int32_t* map_idx, const int16_t* start_xy,
                     uint64_t seed, float* obs, float* reward, double* reward64, uint8_t* done,
                     double* stats, cudaStream_t st) {
    mcov::SuperParams p;
    fill_super_params(env, p);
    p.mode = mode;
    p.actions = actions;
    p.act_fmt = act_fmt;
    p.mask = mask;
    p.mask_from_flags = mask_from_flags;
    p.map_idx_in = map_idx;
    p.start_xy_in = start_xy;
    p.seed = seed;
    p.obs = obs;
    p.reward = reward;
    p.reward64 = reward64;
    p.done = done;
    p.stats_partial = reinterpret_cast<double*>(env->arena + env->lay.off_scratch);
    p.stats_out = (mode == mcov::MODE_STEP) ? stats : nullptr;
    p.counter = reinterpret_cast<unsigned int*>(scratch_counter(env));
    p.dt_scratch = mcov::super_dt_scratch_bytes(env) > 0
                       ? reinterpret_cast<uint16_t*>(scratch_super_dt(env)) : nullptr;
    p.vec_ok = (env->g.L % 4 == 0) && aligned(obs, 16);
    cudaError_t ce = mcov::launch_super(env, p, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "super kernel launch");
    return MCOV_OK;
}

int mcov_reset(mcov_env* env, const uint8_t* mask_dev, const int32_t* map_idx_dev,
               const int16_t* start_xy_dev, uint64_t seed, float* obs_out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc) return rc;
    if (!obs_out_dev) return fail(MCOV_ERR_NULL, "obs_out is NULL");
    if (!aligned(obs_out_dev, 16)) return fail(MCOV_ERR_ALIGN, "obs_out must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    env->seed = seed;
    if (env->cfg.env_type == MCOV_ENV_DEC)
        return run_dec(env, mcov::MODE_RESET, nullptr, 0, mask_dev, 0, map_idx_dev, start_xy_dev, seed,
                       obs_out_dev, nullptr, nullptr, nullptr, nullptr, st);
    return run_super(env, mcov::MODE_RESET, nullptr, 0, mask_dev, 0, map_idx_dev, start_xy_dev, seed,
                     obs_out_dev, nullptr, nullptr, nullptr, nullptr, st);
}

int mcov_step(mcov_env* env, const void* actions_dev, int action_format, float* obs_out_dev,
              float* reward_out_dev, double* reward64_out_dev, uint8_t* done_out_dev,
              double* stats_out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc) return rc;
    if (!actions_dev || !obs_out_dev || !reward_out_dev || !done_out_dev)
        return fail(MCOV_ERR_NULL, "mcov_step: actions/obs/reward/done is NULL");
    if (action_format != MCOV_ACT_U8_PER_ROBOT && action_format != MCOV_ACT_I64_BASE4)
        return fail(MCOV_ERR_BAD_CONFIG, "action_format %d", action_format);
    if (action_format == MCOV_ACT_I64_BASE4 && env->cfg.num_robots > 31)
        return fail(MCOV_ERR_BAD_CONFIG, "base-4 int64 actions hold at most 31 robots");
    if (action_format == MCOV_ACT_I64_BASE4 && !aligned(actions_dev, 8))
        return fail(MCOV_ERR_ALIGN, "int64 actions must be 8-byte aligned");
    if (!aligned(obs_out_dev, 16)) return fail(MCOV_ERR_ALIGN, "obs_out must be 16-byte aligned");
    if (!aligned(reward_out_dev, 4) || (reward64_out_dev && !aligned(reward64_out_dev, 8)))
        return fail(MCOV_ERR_ALIGN, "reward buffers misaligned");
    if (stats_out_dev && !aligned(stats_out_dev, 8)) return fail(MCOV_ERR_ALIGN, "stats_out misaligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool dec = env->cfg.env_type == MCOV_ENV_DEC;
    // finished / truncated envs restart in place when auto_reset is on; their obs rows become the
    // new episode's first observation (gym auto-reset convention).  Seed: derived from the seed of
    // the last mcov_reset; the per-slot episode counter and the GLOBAL env id key the draw inside
    // the kernel, so a batch gives the same episodes however it is sharded over GPUs.
    const uint64_t seed = 0x9E3779B97F4A7C15ull ^ env->seed;
    if (dec)   // DecGridRL: the re-seed is fused into the step kernel (one launch per step)
        return run_dec(env, mcov::MODE_STEP, actions_dev, action_format, nullptr, 0, nullptr, nullptr, seed,
                       obs_out_dev, reward_out_dev, reward64_out_dev, done_out_dev, stats_out_dev, st);
    // SuperGridRL: same single-launch scheme
    return run_super(env, mcov::MODE_STEP, actions_dev, action_format, nullptr, 0, nullptr, nullptr, seed,
                     obs_out_dev, reward_out_dev, reward64_out_dev, done_out_dev, stats_out_dev, st);
}

int mcov_observe(mcov_env* env, float* obs_out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc) return rc;
    if (!obs_out_dev) return fail(MCOV_ERR_NULL, "obs_out is NULL");
    if (!aligned(obs_out_dev, 16)) return fail(MCOV_ERR_ALIGN, "obs_out must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (env->cfg.env_type == MCOV_ENV_DEC)
        return run_dec(env, mcov::MODE_OBSERVE, nullptr, 0, nullptr, 0, nullptr, nullptr, 0, obs_out_dev,
                       nullptr, nullptr, nullptr, nullptr, st);
    return run_super(env, mcov::MODE_OBSERVE, nullptr, 0, nullptr, 0, nullptr, nullptr, 0, obs_out_dev,
                     nullptr, nullptr, nullptr, nullptr, st);
}

int mcov_percent_covered(mcov_env* env, double* out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc) return rc;
    if (!out_dev) return fail(MCOV_ERR_NULL, "out is NULL");
    cudaError_t ce = mcov::launch_percent_covered(env, out_dev, static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return cuda_fail(ce, "percent_covered launch");
    return MCOV_OK;
}

int mcov_replay_push(const float* state_dev, const float* next_state_dev, int row_floats,
                     int64_t src_stride_floats, const void* actions_dev, int actions_are_u8, int act_dim,
                     const float* reward_dev, const uint8_t* done_dev, int num_envs, float* ring_state_dev,
                     float* ring_next_dev, float* ring_action_dev, float* ring_reward_dev,
                     float* ring_done_dev, int64_t ptr, int64_t capacity, void* stream) {
    if (!state_dev || !next_state_dev || !actions_dev || !reward_dev || !done_dev || !ring_state_dev ||
        !ring_next_dev || !ring_action_dev || !ring_reward_dev || !ring_done_dev)
        return fail(MCOV_ERR_NULL, "mcov_replay_push: NULL buffer");
    if (row_floats < 1 || act_dim < 1 || num_envs < 1 || capacity < 1 || src_stride_floats < row_floats)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_replay_push: row %d stride %lld act_dim %d envs %d capacity %lld",
                    row_floats, (long long)src_stride_floats, act_dim, num_envs, (long long)capacity);
    if (num_envs > capacity) return fail(MCOV_ERR_BAD_CONFIG, "mcov_replay_push: %d envs > capacity %lld",
                                         num_envs, (long long)capacity);
    if (ptr < 0 || ptr >= capacity) return fail(MCOV_ERR_BAD_CONFIG, "mcov_replay_push: ptr %lld", (long long)ptr);
    cudaError_t ce = mcov::launch_replay_push(
        state_dev, next_state_dev, row_floats, src_stride_floats,
        actions_are_u8 ? static_cast<const uint8_t*>(actions_dev) : nullptr,
        actions_are_u8 ? nullptr : static_cast<const float*>(actions_dev), act_dim, reward_dev, done_dev, num_envs,
        ring_state_dev, ring_next_dev, ring_action_dev, ring_reward_dev, ring_done_dev, ptr, capacity,
        static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return cuda_fail(ce, "replay push launch");
    return MCOV_OK;
}

int mcov_replay_gather(const int64_t* idx_dev, int64_t capacity, int batch, int row_floats, int act_dim,
                       const float* ring_state_dev, const float* ring_next_dev, const float* ring_action_dev,
                       const float* ring_reward_dev, const float* ring_done_dev, float* out_state_dev,
                       float* out_next_dev, float* out_action_dev, float* out_reward_dev, float* out_done_dev,
                       void* stream) {
    if (!idx_dev || !ring_state_dev || !ring_next_dev || !ring_action_dev || !ring_reward_dev || !ring_done_dev ||
        !out_state_dev || !out_next_dev || !out_action_dev || !out_reward_dev || !out_done_dev)
        return fail(MCOV_ERR_NULL, "mcov_replay_gather: NULL buffer");
    if (batch < 1 || row_floats < 1 || act_dim < 1 || capacity < 1)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_replay_gather: batch %d row %d act_dim %d capacity %lld", batch,
                    row_floats, act_dim, (long long)capacity);
    cudaError_t ce = mcov::launch_replay_gather(
        reinterpret_cast<const long long*>(idx_dev), (long long)capacity, batch, row_floats, act_dim, ring_state_dev, ring_next_dev,
        ring_action_dev, ring_reward_dev, ring_done_dev, out_state_dev, out_next_dev, out_action_dev, out_reward_dev,
        out_done_dev, static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return cuda_fail(ce, "replay gather launch");
    return MCOV_OK;
}

int mcov_sim_step(mcov_env* env, const float* state_dev, const uint8_t* actions_dev, const int32_t* map_idx_dev,
                  const int16_t* pos_xy_dev, int batch, double done_thresh, float* next_state_dev,
                  float* reward_out_dev, double* reward64_out_dev, uint8_t* terminal_out_dev, void* stream) {
    if (!env) return fail(MCOV_ERR_NULL, "env handle is NULL");
    if (env->cfg.env_type != MCOV_ENV_SUPER)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_sim_step needs a SuperGridRL handle (sim_env.py:21-22)");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    if (!state_dev || !actions_dev || !next_state_dev) return fail(MCOV_ERR_NULL, "mcov_sim_step: NULL buffer");
    if (batch < 1) return fail(MCOV_ERR_BAD_CONFIG, "mcov_sim_step: batch %d", batch);
    if (int rc = check_device(env)) return rc;
    if (!aligned(state_dev, 4) || !aligned(next_state_dev, 4))
        return fail(MCOV_ERR_ALIGN, "mcov_sim_step: state buffers must be float aligned");
    const mcov_config& c = env->cfg;
    mcov::SimParams p;
    memset(&p, 0, sizeof(p));
    p.B = batch;
    p.g = env->g;
    p.R = c.senseradius;
    p.dist_reward = c.dist_reward;
    p.cp = c.collision_penalty;
    p.terminal = c.terminal_reward;
    p.free_penalty = c.free_penalty;
    p.done_thresh = done_thresh;
    p.wall = env->wall;
    p.pos = env->pos;
    p.free_counts = env->free_counts;
    p.cell_values = env->cell_values;
    p.n_maps = env->n_maps;
    p.state = state_dev;
    p.actions = actions_dev;
    p.map_idx = map_idx_dev;
    p.pos_xy = pos_xy_dev;
    p.next_state = next_state_dev;
    p.reward = reward_out_dev;
    p.reward64 = reward64_out_dev;
    p.terminal_out = terminal_out_dev;
    p.vec_ok = ((env->g.W * env->g.L) % 4 == 0) && aligned(state_dev, 16) && aligned(next_state_dev, 16);
    cudaError_t ce = mcov::launch_sim_lookahead(p, env->num_sms, static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return cuda_fail(ce, "sim lookahead launch");
    env->launches += 1;
    return MCOV_OK;
}

static int render_geometry(const mcov_env* env, int* scale, int* h, int* w) {
    // scaling = max(min(B // W, B // L), 1), B = 1075 (super_grid_rl.py:449) or 1024 (dec_grid_rl.py:583);
    // copyMakeBorder(image, 0, 1075 - shape[1], 0, 1075 - shape[0]) adds rows by the COLUMN deficit and
    // columns by the ROW deficit (super_grid_rl.py:452-454, dec_grid_rl.py:587-589)
    const int base = env->cfg.env_type == MCOV_ENV_DEC ? 1024 : 1075;
    int s = base / env->g.W < base / env->g.L ? base / env->g.W : base / env->g.L;
    if (s < 1) s = 1;
    const int bottom = 1075 - env->g.L * s, right = 1075 - env->g.W * s;
    if (bottom < 0 || right < 0)
        return fail(MCOV_ERR_BAD_CONFIG, "map %dx%d does not fit the 1075 canvas (cv2.copyMakeBorder fails too)",
                    env->g.W, env->g.L);
    *scale = s;
    *h = env->g.W * s + bottom;
    *w = env->g.L * s + right;
    return MCOV_OK;
}

int mcov_render_shape(const mcov_env* env, int32_t* height_out, int32_t* width_out) {
    if (!env || !height_out || !width_out) return fail(MCOV_ERR_NULL, "mcov_render_shape: NULL");
    if (!env->maps_set) return fail(MCOV_ERR_STATE, "mcov_set_maps has not been called");
    int s, h, w;
    int rc = render_geometry(env, &s, &h, &w);
    if (rc != MCOV_OK) return rc;
    *height_out = h;
    *width_out = w;
    return MCOV_OK;
}

int mcov_render(mcov_env* env, const int32_t* env_ids_dev, int n_frames, uint8_t* rgb_out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc != MCOV_OK) return rc;
    if (!rgb_out_dev) return fail(MCOV_ERR_NULL, "mcov_render: output is NULL");
    if (n_frames < 1 || (!env_ids_dev && n_frames > env->cfg.num_envs))
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_render: n_frames %d", n_frames);
    if (!aligned(rgb_out_dev, 4)) return fail(MCOV_ERR_ALIGN, "mcov_render: output must be 4-byte aligned");
    mcov::RenderParams p;
    memset(&p, 0, sizeof(p));
    rc = render_geometry(env, &p.scale, &p.out_h, &p.out_w);
    if (rc != MCOV_OK) return rc;
    const mcov_config& c = env->cfg;
    const mcov_layout& l = env->lay;
    p.n = n_frames;
    p.env_ids = env_ids_dev;
    p.g = env->g;
    p.env_type = c.env_type;
    p.N = c.num_robots;
    p.hdr = env->arena + l.off_hdr;
    p.hdr_stride = l.hdr_stride;
    p.off_pos = l.hdr_off_pos;
    if (c.env_type == MCOV_ENV_DEC) {
        p.n_shown = c.num_robots;                       // _robot_pos_map holds every robot
        p.flip = 1;
        p.plane_a = reinterpret_cast<const uint32_t*>(env->arena + l.off_visited);
        p.fo = reinterpret_cast<const uint2*>(env->arena + l.off_fo);
    } else {
        p.n_shown = c.use_scanning ? c.num_robots : 1;  // get_pos_image()[0] (super_grid_rl.py:442)
        p.flip = 0;
        p.plane_a = reinterpret_cast<const uint32_t*>(env->arena + l.off_covered);
        p.plane_b = reinterpret_cast<const uint32_t*>(env->arena + l.off_seen_obst);
    }
    p.out = rgb_out_dev;
    cudaError_t ce = mcov::launch_render(p, env->num_sms, static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return cuda_fail(ce, "render launch");
    env->launches += 1;
    return MCOV_OK;
}

int64_t mcov_launch_count(const mcov_env* env) { return env ? env->launches : 0; }

}  // extern "C"

// ---- mcov_step_static_scratch: the DecGridRL step with its launch scratch in a static device array ----------------
// No counterpart in the reference; added for callers that cannot let a library
// touch the allocator on the step path (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  mcov_step keeps its per-launch scratch — one row of MCOV_STATS_LEN doubles per CTA for the episode
// statistics fold, and the CTA arrival counter — inside the caller's state arena; this variant keeps them in the array
// below instead, so the step needs nothing from the arena's scratch region.  Every GPU that loads the library holds the
// array (MCOV_STATIC_SCRATCH_BYTES = 128 KiB + 256 B) until the process exits.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the kernels get the address that
// cudaGetSymbolAddress returns for the device that is current (resolved once per device, below).
#define MCOV_STATIC_SCRATCH_BYTES (MCOV_STATIC_SCRATCH_MAX_ROWS * MCOV_STATS_LEN * 8 + 256)
static __device__ __align__(256) unsigned char g_mcov_static_scratch[MCOV_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::atomic<unsigned char*> g_static_scratch_ptr[kMaxDevices];   // device address of the array, per device ordinal

// Device address of g_mcov_static_scratch on `device`, which must be the current device.  Two threads racing here
// store the same value.  Allocates nothing (the array is part of the loaded module).
int static_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) return fail(MCOV_ERR_BAD_CONFIG, "device ordinal %d outside [0, %d)", device, kMaxDevices);
    unsigned char* ptr = g_static_scratch_ptr[device].load(std::memory_order_acquire);
    if (!ptr) {
        void* sym = nullptr;
        cudaError_t ce = cudaGetSymbolAddress(&sym, g_mcov_static_scratch);
        if (ce != cudaSuccess) return cuda_fail(ce, "cudaGetSymbolAddress(static scratch)");
        if (!sym) return fail(MCOV_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        ptr = static_cast<unsigned char*>(sym);
        g_static_scratch_ptr[device].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return MCOV_OK;
}

}  // namespace

extern "C" int mcov_step_static_scratch(mcov_env* env, const void* actions_dev, int action_format, float* obs_out_dev,
                                        float* reward_out_dev, double* reward64_out_dev, uint8_t* done_out_dev,
                                        double* stats_out_dev, void* stream) {
    int rc = check_ready(env);
    if (rc) return rc;
    const mcov_config& c = env->cfg;
    if (c.env_type != MCOV_ENV_DEC)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_step_static_scratch: DecGridRL handles only (use mcov_step)");
    if (c.map_sharing || c.dist_reward || c.dijkstra_input || c.mini_map_rad > 0)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_step_static_scratch: map_sharing / dist_reward / dijkstra_input / mini_map_rad "
                                         "steps take further launches with arena scratch (use mcov_step)");
    if (!actions_dev || !obs_out_dev || !reward_out_dev || !done_out_dev)
        return fail(MCOV_ERR_NULL, "mcov_step_static_scratch: actions/obs/reward/done is NULL");
    if (action_format != MCOV_ACT_U8_PER_ROBOT && action_format != MCOV_ACT_I64_BASE4)
        return fail(MCOV_ERR_BAD_CONFIG, "action_format %d", action_format);
    if (action_format == MCOV_ACT_I64_BASE4 && c.num_robots > 31)
        return fail(MCOV_ERR_BAD_CONFIG, "base-4 int64 actions hold at most 31 robots");
    if (action_format == MCOV_ACT_I64_BASE4 && !aligned(actions_dev, 8))
        return fail(MCOV_ERR_ALIGN, "int64 actions must be 8-byte aligned");
    if (!aligned(obs_out_dev, 16)) return fail(MCOV_ERR_ALIGN, "obs_out must be 16-byte aligned");
    if (!aligned(reward_out_dev, 4) || (reward64_out_dev && !aligned(reward64_out_dev, 8)))
        return fail(MCOV_ERR_ALIGN, "reward buffers misaligned");
    if (stats_out_dev && !aligned(stats_out_dev, 8)) return fail(MCOV_ERR_ALIGN, "stats_out misaligned");
    // the size that matters: one statistics row per CTA of the launch
    const int rows = mcov::dec_launch_grid_max(env);
    if (rows > MCOV_STATIC_SCRATCH_MAX_ROWS)
        return fail(MCOV_ERR_BAD_CONFIG, "mcov_step_static_scratch: this batch launches up to %d CTAs, the static scratch holds "
                                         "%d statistics rows (MCOV_STATIC_SCRATCH_MAX_ROWS); use mcov_step", rows,
                    MCOV_STATIC_SCRATCH_MAX_ROWS);
    int cur = -1;
    cudaError_t ce = cudaGetDevice(&cur);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaGetDevice");
    if (cur != env->device)
        return fail(MCOV_ERR_STATE, "mcov_step_static_scratch: handle lives on device %d, current device is %d", env->device, cur);
    unsigned char* scratch = nullptr;
    rc = static_scratch_address(cur, &scratch);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char* counter = scratch + (size_t)MCOV_STATIC_SCRATCH_MAX_ROWS * MCOV_STATS_LEN * 8;
    // the array keeps whatever the last call (of any handle) left: the rows are overwritten before they are read, the
    // arrival counter must start at zero
    ce = cudaMemsetAsync(counter, 0, 256, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaMemsetAsync(static scratch counter)");

    mcov::DecParams p;
    fill_dec_params(env, p);
    p.mode = mcov::MODE_STEP;
    p.actions = actions_dev;
    p.act_fmt = action_format;
    p.mask = nullptr;
    p.mask_from_flags = 0;
    p.map_idx_in = nullptr;
    p.start_xy_in = nullptr;
    p.seed = 0x9E3779B97F4A7C15ull ^ env->seed;                   // as mcov_step: fused auto-resets draw from it
    p.obs = obs_out_dev;
    p.reward = reward_out_dev;
    p.reward64 = reward64_out_dev;
    p.done = done_out_dev;
    p.vec_ok = ((p.obs_env_floats * 4) % 16 == 0) && aligned(obs_out_dev, 16);
    p.phases = mcov::PH_MOVE | mcov::PH_SENSE;
    p.stats_partial = reinterpret_cast<double*>(scratch);
    p.stats_out = stats_out_dev;
    p.counter = reinterpret_cast<unsigned int*>(counter);
    p.move_scratch = nullptr;                                      // only the dist_reward pipeline uses it (refused above)
    ce = mcov::launch_dec(env, p, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "dec kernel launch (static scratch)");
    return MCOV_OK;
}

// mcov_internal.h — host-side handle and kernel parameter blocks (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mcov.h"
#include "mcov_common.cuh"

namespace mcov {

enum { MODE_STEP = 0, MODE_RESET = 1, MODE_OBSERVE = 2 };
enum { PH_MOVE = 1, PH_SENSE = 2 };   // step phases (split only for the dist_reward pipeline)

// Parameter block of the DecGridRL kernels (passed by value).
struct DecParams {
    int E, N;
    Geometry g;
    int R, S;                 // sensor range, S = 2R+1 rows/bits of the sensor window
    int ego, ow;              // egoradius, ow = 2*ego+1
    int C;                    // observation channels (3 + extras)
    int sensor;               // MCOV_SENSOR_*
    int n_lasers;
    const double* beams;      // [n_lasers][3] device
    const uint16_t* lidar_cells;  // [n_lasers][16] window cells of each beam in march order: y bit | (4 * row) << 8; slots behind
                                  // the last cell name the dummy all-wall row S; slot 0 (the robot's own cell) is not walked; or null
    const uint2* lidar_runs;      // [n_lasers][16] the same cells grouped into same-row runs: {bit mask of the run's cells,
                                  // row byte offset | bit 31 = cells are marched towards lower y}, mask 0 ends the beam
    uint32_t l_magic;         // ceil(2^32 / n_lasers)
    uint32_t l_mult;          // beam permutation of the table walk: l -> l * l_mult mod n_lasers (coprime)
    int single_tool, comm_radius, want_adj, dist_reward;
    int maxsteps, truncate_steps, auto_reset;
    int env_id_base;
    double cp, terminal, done_incr;
    // maps
    const uint32_t* wall;
    const uint32_t* pos;
    const int32_t* free_counts;
    int n_maps;
    // state
    unsigned char* hdr;
    int hdr_stride, off_pos, off_cov, off_adj;
    uint32_t* visited;
    uint2* fo;
    uint8_t* needs_reset;
    float* dist_term;         // [E] float32 distance-reward term prepared by the dist kernel (or null)
    // io
    int mode;
    const void* actions;
    int act_fmt;
    const uint8_t* mask;      // reset: which envs (null = all); MODE_RESET with from_flags uses needs_reset
    int mask_from_flags;
    const int32_t* map_idx_in;
    const int16_t* start_xy_in;
    uint64_t seed;
    float* obs;
    float* reward;
    double* reward64;
    uint8_t* done;
    double* stats_partial;    // [gridDim][MCOV_STATS_LEN] scratch rows
    double* stats_out;        // [MCOV_STATS_LEN] or null: folded by the last CTA of the launch
    unsigned int* counter;    // CTA arrival counter for that fold (zero between launches)
    int obs_env_floats;       // N*C*ow*ow
    int vec_ok;               // env obs block is a multiple of 16 bytes and obs is 16B aligned
    int phases;               // PH_MOVE | PH_SENSE
    uint2* move_scratch;      // [E] (collisions, shown mask) handed from a PH_MOVE to a PH_SENSE launch
    int skip_wipe;            // reset: planes were already zeroed by a memset
    // launch-derived (filled by launch_dec)
    int rad, RW;              // rad = max(ego, R), RW = 2*rad+1 rows of the fused band
    int bits_words, grp_smem; // obs bit-stream words, shared-memory bytes per lane group
    uint32_t rw_magic, s_magic, n_magic;
    int debug;                // MCOV_DEBUG env bits (profiling aids; 0 in normal use)
    // store-warp variant (Spec::ST, mcov_dec.cu): layout of the CTA's dynamic shared memory behind the group areas
    int st_slot_words;        // words per bit-stream slot (two per lane group); 0 = direct stores
    int st_nbuf;              // float32 staging buffers per store warp
    int st_warps;             // store warps per CTA
    int store_team;           // lanes that store an observation block together (>= lanes per env; direct-store path)
    int smt_map_off, smt_bar_off;   // Spec::SMT: byte offsets of the shared-memory truth-map table and its mbarrier
    int st_rot_div;           // store warp index = (blockIdx / st_rot_div) % 4 (0: the CTA's last warps)
    int st_bar_off, st_stage_off, st_stage_pitch;
};

struct SuperParams {
    int E, N;
    Geometry g;
    int R;                    // senseradius
    int C;                    // N+3, or 4 with use_scanning
    int use_scanning, dist_reward;
    int truncate_steps, auto_reset, env_id_base;
    double cp, terminal, done_incr, free_penalty;
    const uint32_t* wall;
    const uint32_t* pos;
    const int32_t* free_counts;
    const double* cell_values;  // [n_maps][W][L] reward of covering a cell (super_grid_rl.py:189), or null = 1.0
    int n_maps;
    unsigned char* hdr;
    int hdr_stride, off_pos, off_cov;
    uint32_t* covered;
    uint32_t* seen_obst;
    uint8_t* needs_reset;
    int mode;
    const void* actions;
    int act_fmt;
    const uint8_t* mask;
    int mask_from_flags;
    const int32_t* map_idx_in;
    const int16_t* start_xy_in;
    uint64_t seed;
    float* obs;
    float* reward;
    double* reward64;
    uint8_t* done;
    double* stats_partial;    // [gridDim][MCOV_STATS_LEN] scratch rows
    double* stats_out;        // [MCOV_STATS_LEN] or null: folded by the last CTA of the launch
    unsigned int* counter;    // CTA arrival counter for that fold (zero between launches)
    // distance-transform work buffer: uint16 [W][dt_pitch] per CTA, in shared memory when it fits,
    // otherwise one slice per CTA of a global scratch area inside the arena
    uint16_t* dt_scratch;     // null -> shared memory
    int dt_pitch;             // uint16 elements per x row (2 * odd number of words)
    int vec_ok;               // L % 4 == 0 and obs 16-byte aligned: float4 stores
};

// SuperGrid_Sim.step lookahead batch (mcov_aux.cu).
struct SimParams {
    int B;
    Geometry g;
    int R, dist_reward;
    double cp, terminal, free_penalty, done_thresh;
    const uint32_t* wall;
    const uint32_t* pos;
    const int32_t* free_counts;
    const double* cell_values; // [n_maps][W][L] or null (sim_env.py:159 reward += grid[j][k])
    const float* state;        // [B][4][W][L]
    const uint8_t* actions;    // [B]
    int n_maps;
    const int32_t* map_idx;    // [B] or null (map 0)
    const int16_t* pos_xy;     // [B][2] or null (scan the position plane)
    float* next_state;         // [B][4][W][L]; == state: patch in place
    float* reward;
    double* reward64;
    uint8_t* terminal_out;
    int vec_ok;
};

// render() frames (mcov_aux.cu).
struct RenderParams {
    int n;
    const int32_t* env_ids;    // [n] or null (envs 0..n-1)
    Geometry g;
    int env_type, N, n_shown;  // n_shown: robots drawn (Super without scanning draws robot 0 only)
    int scale, out_h, out_w, flip;
    const unsigned char* hdr;
    int hdr_stride, off_pos;
    const uint32_t* plane_a;   // Dec: visited; Super: covered
    const uint32_t* plane_b;   // Super: seen_obst
    const uint2* fo;           // Dec: per-robot {free, obst}
    uint8_t* out;              // [n][out_h][out_w][3]
};

}  // namespace mcov

struct mcov_env {
    mcov_config cfg;
    int device;
    int num_sms;
    bool maps_set, bound;
    mcov::Geometry g;
    const uint32_t* wall;
    const uint32_t* pos;
    const int32_t* free_counts;
    int n_maps;
    mcov_layout lay;
    unsigned char* arena;
    double* beams_dev;        // owned
    uint16_t* lidar_cells_dev; // owned; null when the integer beam table could not be proven exact
    uint2* lidar_runs_dev;     // owned; same table as same-row runs (null with lidar_cells_dev)
    int64_t launches;
    uint64_t seed;            // seed of the last mcov_reset (auto-reset draws derive from it)
    int obs_C;
    int debug;                // MCOV_DEBUG at mcov_create (profiling aids; 0 in normal use)
    // tuning / test pins read from the environment once, at mcov_create (-1 = not pinned):
    // MCOV_DEC_GROUP (lanes per env), MCOV_DEC_STORE (store-warp path 0|1), MCOV_DEC_GRID (grid cap),
    // MCOV_DEC_NBUF (staging buffers per store warp), MCOV_DEC_SW (store warps per CTA), MCOV_DEC_SHAPE (0 lane groups, 1 CTA per env), MCOV_DEC_TMA (CTA shape: 0 plain loads, 1 bulk-copy staging), MCOV_DEC_SMT (lane groups: truth-map table in shared memory)
    int pin_group, pin_store, pin_grid, pin_nbuf, pin_shape, pin_sw, pin_rot, pin_tma, pin_team, pin_warps, pin_balance, pin_smt;
    size_t dec_smem_configured, super_smem_configured;   // dynamic shared memory opted into on this handle's device
    const void* dec_smem_func;  // ... and for which instantiation (MCOV_DEC_GROUP can change it between launches)
    const double* cell_values;  // Super: float64 [n_maps][W][L] reward of covering a cell, or null (= 1.0)
};

// kernel launchers (defined in the .cu files)
namespace mcov {
cudaError_t launch_dec(mcov_env* env, DecParams& p, cudaStream_t st);
bool want_dec_cta(const mcov_env* env);                 // CTA-per-env shape (mcov_dec_cta.cu) instead of lane groups
cudaError_t launch_dec_cta(mcov_env* env, DecParams& p, cudaStream_t st);
cudaError_t launch_dec_share(mcov_env* env, cudaStream_t st);
cudaError_t launch_dec_extras(mcov_env* env, float* obs, float* dist_term, unsigned char* scratch,
                              int pre_sense, cudaStream_t st);
int64_t dec_extras_scratch_bytes(const mcov_env* env);  // 0 when the work area lives in shared memory
cudaError_t launch_super(mcov_env* env, SuperParams& p, cudaStream_t st);
cudaError_t launch_stats_reduce(mcov_env* env, const double* partial, int rows, double* out,
                                cudaStream_t st);
cudaError_t launch_init_headers(mcov_env* env, cudaStream_t st);
cudaError_t launch_percent_covered(mcov_env* env, double* out, cudaStream_t st);
cudaError_t launch_replay_push(const float* state, const float* next_state, int row_floats, long long src_stride,
                               const uint8_t* act_u8, const float* act_f32, int act_dim, const float* reward,
                               const uint8_t* done, int E, float* r_state, float* r_next, float* r_action,
                               float* r_reward, float* r_done, long long ptr, long long capacity, cudaStream_t st);
cudaError_t launch_replay_gather(const long long* idx, long long capacity, int B, int row_floats, int act_dim, const float* r_state,
                                 const float* r_next, const float* r_action, const float* r_reward,
                                 const float* r_done, float* o_state, float* o_next, float* o_action,
                                 float* o_reward, float* o_done, cudaStream_t st);
cudaError_t launch_sim_lookahead(const SimParams& p, int num_sms, cudaStream_t st);
cudaError_t launch_render(const RenderParams& p, int num_sms, cudaStream_t st);
int dec_grid_size(const mcov_env* env);
int dec_cta_grid_size(const mcov_env* env);             // grid of the CTA-per-env shape (mcov_dec_cta.cu)
int dec_launch_grid_max(const mcov_env* env);            // upper bound of the statistics rows a Dec step launch writes
int super_grid_size(const mcov_env* env);
int64_t super_dt_scratch_bytes(const mcov_env* env);   // 0 when the buffer lives in shared memory
}  // namespace mcov

// Gogoro scooter (registered variant) env-step hot path for B200 (sm_100a).
//
// Replaces, behind the C ABI of include/thormang_b200.h, the TorchScript functions and method
// bodies of /root/reference/isaacgymenvs/tasks/gogoro_new.py:
//   pre_physics_step :347-369, post_physics_step :373-389, compute_obs_rwd :424-462,
//   reset_idx :505-590 (+ randomize :474-482, generate_spawn_r :485-492, euler_to_quaternion :496-502),
//   compute_gogoro_observations :692-723, shortest_angle_distance :687-689, compute_gogoro_reward :645-684.
//
// 192 B / env-step of algorithmic traffic (172 B with in-kernel Philox noise).  One CTA = four warps = 128 consecutive
// envs, thread e owns env e.  The AoS rows that are not a multiple of 16 bytes (root 52 B, action history 20 B, supplied
// sensor noise 20 B, the two 24 B obs rows) move as cp.async.bulk (TMA) tile copies through shared memory (row strides
// 13 / 5 / 6 words: conflict-free or 2-way), everything else is a coalesced 4/8-byte access per lane.
// 128-env CTAs, not 32: a B200 dispatches ~0.85 CTA / ns (tools/launch_floor.py: an EMPTY 2048-CTA grid takes 2.4 us, a
// 4096-CTA grid 4.8 us), so with one warp per CTA the 65536-env launch spent 2.4 us and the 4 M-env launch 153 us --
// more than its 113 us of HBM time -- on CTA dispatch alone.
#include "tk_common.cuh"

#include <cstdlib>
#include <new>

namespace tk {
namespace gogoro {

constexpr int kRoot = 13;
constexpr int kObs = 6;
constexpr int kHist = 5;
constexpr int kSensorNoise = 5;
constexpr int kResetNoise = 10;
constexpr int kTile = 128;     // envs per CTA (tile slices of every array are 16-byte aligned bulk-copy operands)
constexpr int kWarps = kTile / 32;

struct KParams {
    float* root;
    float* state_dof;
    float* obs_buf;
    float* buffer_obs;
    float* rew;
    long long* reset;
    long long* progress;
    uint8_t* timeout;
    float* yaw_command;
    float* curent_command;
    float* curent_speed;
    float* steer_offsets;
    float* imu_offsets;
    float* curent_speed_offset;
    float* action_history;
    float* config_vector;
    const float* thormang_pose;
    const float* root_reset;
    float* pos_cmd;
    float* vel_cmd;
    float* ep_return;
    double* partials;
    const int* tile_base_reset;   // COMPACT: resets before this tile
    const int* tile_base_speed;   // COMPACT: speed resamples before this tile
    const float* reset_unit;
    const float* sensor_unit;
    const float* speed_unit;
    const float* yaw_unit;
    long long reset_rows;
    unsigned long long seed, step;
    unsigned long long* step_dev;   // device-side noise epoch (upper 32 bits of the Philox step), see tk_common.cuh
    long long env_off;
    int noise_mode;
    int n, num_dof, steering_joint, rear_wheel_joint, max_episode_length, speed_freq, yaw_freq;
    int vec_ok;
    int prefetch;                 // L2 prefetch of the tile inputs ahead of griddepcontrol.wait (thin grids, host decides)
    int ldg_tiles;                // developer knob TK_GOGORO_LDG: input tiles through the load/store unit (measured slower)
    int dbg_reset;                // developer ablation mask (TK_GOGORO_DBG_RESET): 1 no resets, 2 no scalar stores, 4 no row fetch/stores, 8 no Philox
    long long* dbg_ts;            // developer timeline probe (tools/tile_timeline.py --task gogoro); NULL in normal use
    float max_steering, max_steering_change;
    // (float)lo / (float)mean, (float)(hi - lo) / (float)std with the span formed in double (get_randoms :466-467)
    float imu_filter[2], imu[2], speed_sensor[2], steer_action[2], speed[2], seat_x[2], seat_y[2], seat_z[2], seat_xr[2],
        steer_off[2], speed_off[2];
};

// philox stream ids per (env, step); see tk_gogoro_philox_noise for the materialised layout
enum { kStreamSensor = 0, kStreamMisc = 1, kStreamResetU = 2, kStreamResetN0 = 3, kStreamResetN1 = 4 };

// Per (env, step): stream kStreamSensor -> the four LIVE sensor normals (gogoro_new.py:451-453,460; the draw of
// :455 is dead because :457-458 overwrite channel 3), stream kStreamMisc -> steer normal, the dead draw, and the two
// resample uniforms (needed only on the rare steps where progress hits the update frequency).
struct StepNoise {
    float sensor[kSensorNoise];
};
__device__ __forceinline__ StepNoise philox_sensor_noise(unsigned long long seed, unsigned long long step, long long env) {
    StepNoise s;
    Philox4 a = philox_env(seed, step, env, kStreamSensor);
    float2 n01 = normal_pair(a.x, a.y), n24 = normal_pair(a.z, a.w);
    s.sensor[0] = n01.x; s.sensor[1] = n01.y; s.sensor[2] = n24.x; s.sensor[3] = 0.0f; s.sensor[4] = n24.y;
    return s;
}
struct MiscNoise { float steer, dead3, speed_u, yaw_u; };
__device__ __forceinline__ MiscNoise philox_misc_noise(unsigned long long seed, unsigned long long step, long long env) {
    Philox4 b = philox_env(seed, step, env, kStreamMisc);
    float2 n = normal_pair(b.x, b.y);
    return MiscNoise{n.x, n.y, u01(b.z), u01(b.w)};
}
// The reset draws run in a handful of warps per launch, i.e. from a cold instruction cache: the three Philox blocks and the
// three Box-Muller pairs go through ONE out-of-line copy each (3 KB of code instead of 10 KB fetched on the critical path
// of the tiles that hold a reset -- those tiles set the length of a small launch).
__device__ __noinline__ Philox4 philox_env_cold(unsigned long long seed, unsigned long long step, long long env, unsigned stream_id) {
    unsigned k0 = static_cast<unsigned>(seed), k1 = static_cast<unsigned>(seed >> 32);
    const unsigned long long hi = (step << 8) | stream_id;
    Philox4 c{static_cast<unsigned>(env), static_cast<unsigned>(static_cast<unsigned long long>(env) >> 32), static_cast<unsigned>(hi),
              static_cast<unsigned>(hi >> 32)};
#pragma unroll 1
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = mulhi32(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const unsigned hi1 = mulhi32(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = Philox4{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __noinline__ float2 normal_pair_cold(unsigned a, unsigned b) { return normal_pair(a, b); }
// Warp-collective: lanes 0 / 1 / 2 evaluate the three blocks (streams ResetU / ResetN0 / ResetN1) side by side and the ten
// values are broadcast -- one Philox block + one Box-Muller pair deep instead of three of each.  Same functions, same
// bits as philox_reset_noise.
__device__ __forceinline__ void philox_reset_noise_cold(unsigned long long seed, unsigned long long step, long long env, int lane,
                                                        float* u) {
    const unsigned sid = kStreamResetU + static_cast<unsigned>(min(lane, 2));
    const Philox4 blk = philox_env_cold(seed, step, env, sid);
    const float2 pa = normal_pair_cold(blk.x, blk.y), pb = normal_pair_cold(blk.z, blk.w);
    const unsigned full = 0xffffffffu;
    u[0] = __shfl_sync(full, u01(blk.x), 0); u[2] = __shfl_sync(full, u01(blk.y), 0);
    u[3] = __shfl_sync(full, u01(blk.z), 0); u[4] = __shfl_sync(full, u01(blk.w), 0);
    u[1] = __shfl_sync(full, pa.x, 1); u[5] = __shfl_sync(full, pa.y, 1);
    u[6] = __shfl_sync(full, pb.x, 1); u[7] = __shfl_sync(full, pb.y, 1);
    u[8] = __shfl_sync(full, pa.x, 2); u[9] = __shfl_sync(full, pa.y, 2);
}
__device__ __forceinline__ void philox_reset_noise(unsigned long long seed, unsigned long long step, long long env, float* u) {
    Philox4 a = philox_env(seed, step, env, kStreamResetU);
    Philox4 b = philox_env(seed, step, env, kStreamResetN0);
    Philox4 c = philox_env(seed, step, env, kStreamResetN1);
    float2 p0 = normal_pair(b.x, b.y), p1 = normal_pair(b.z, b.w), p2 = normal_pair(c.x, c.y);
    u[0] = u01(a.x); u[2] = u01(a.y); u[3] = u01(a.z); u[4] = u01(a.w);
    u[1] = p0.x; u[5] = p0.y; u[6] = p1.x; u[7] = p1.y; u[8] = p2.x; u[9] = p2.y;
}

// compute_gogoro_observations :692-723
__device__ __forceinline__ void observe(const float* r, float desired_yaw, float last_command, float* o) {
    Q4 q{r[3], r[4], r[5], r[6]};
    float roll, yaw;
    euler_roll_yaw_ref(q, &roll, &yaw);                                             // :696
    V3 lin = quat_rotate_ref<-1>(q, V3{r[7], r[8], r[9]});                          // :698
    V3 ang = quat_rotate_ref<-1>(q, V3{r[10], r[11], r[12]});                       // :699
    roll = (roll > kPi) ? roll - kTwoPi : roll;                                     // :704
    roll = (roll < -kPi) ? roll + kTwoPi : roll;                                    // :705
    yaw = (yaw > kPi) ? yaw - kTwoPi : yaw;                                         // :707
    yaw = (yaw < -kPi) ? yaw + kTwoPi : yaw;                                        // :708
    float delta = floor_mod_two_pi_near(desired_yaw - yaw + kPi) - kPi;             // :710, :689
    o[0] = roll; o[1] = ang.x; o[2] = ang.z; o[3] = lin.x; o[4] = delta; o[5] = last_command;
}

// compute_gogoro_reward :645-684; hist = 5 action-history values
__device__ __forceinline__ float reward_of(const float* o, const float* hist, bool* felt_out) {
    const float tilt = o[0], dtilt = o[1], yaw_e = o[4];
    // divisions by constants as multiplications by the rounded reciprocal (<= 1.5 ulp; the masks do not use them)
    float tilt_err = fminf(fmaxf(tilt * (1.0f / 0.30f), -1.0f), 1.0f);              // :655-656
    float yaw_err = fminf(fmaxf(yaw_e * (1.0f / kPi), -1.0f), 1.0f);                // :658-659
    float dtilt_err = fminf(fmaxf(dtilt * (1.0f / 0.3f), -1.0f), 1.0f);             // :661-662
    float t30 = yaw_err * 30.0f;
    float reward1 = 1.0f / (1.0f + t30 * t30);                                      // :667
    float reward2 = 1.0f - tilt_err * tilt_err;                                     // :670
    float reward4 = 1.0f - dtilt_err * dtilt_err;                                   // :671
    float ce = 0.0f;
#pragma unroll
    for (int i = 0; i < kHist; ++i) ce += 1.0f - hist[i] * hist[i];                 // :673
    float reward = reward1 * 5.0f + reward2 * 0.2f + reward4 * 0.3f + ce * 0.5f;    // :675
    const bool felt = fabsf(tilt) >= 0.30f;                                         // :677
    *felt_out = felt;
    return felt ? -100.0f : reward;                                                 // :682
}

__device__ __forceinline__ float uniform_of(const float* pair, float u) { return pair[0] + u * pair[1]; }   // get_randoms
__device__ __forceinline__ float normal_of(const float* pair, float z) { return pair[0] + z * pair[1]; }    // get_randoms_norm

// reset_idx :505-590 for env `ne`, executed by ONE thread (rare, divergent).  u = the 10 unit draws.
// Writes global state; returns the new root row in r[13] and zeroed history / command in the caller's registers.
__device__ __forceinline__ void reset_env(const KParams& p, long long ne, const float* u, float* r, float* yaw_cmd) {
    p.curent_speed[ne] = uniform_of(p.speed, u[0]);                                 // :479
    // :481 steer_offsets <- N(steering_offset) is overwritten at :575, dead
    p.curent_speed_offset[ne] = uniform_of(p.speed_off, u[2]);                      // :482
    const float initial_target = (u[3] * 2.0f - 1.0f) * kPi;                        // :486
    const float initial_rot = initial_target + (-1.57f + u[4] * static_cast<float>(1.57 - (-1.57)));   // :487
    const float half = initial_rot / 2.0f;
    float sy, cy;
    sincosf(half, &sy, &cy);
    const float sr = 0.0f, cr = 1.0f, sp = 0.0f, cp = 1.0f;                         // sin/cos of roll/2 = pitch/2 = 0
    const float qx = sr * cp * cy - cr * sp * sy;                                   // :497-500
    const float qy = cr * sp * cy + sr * cp * sy;
    const float qz = cr * cp * sy - sr * sp * cy;
    const float qw = cr * cp * cy + sr * sp * sy;
#pragma unroll
    for (int c = 0; c < kRoot; ++c) r[c] = p.root_reset[ne * kRoot + c];            // :518
    r[3] = qx; r[4] = qy; r[5] = qz; r[6] = qw;                                     // :519
#pragma unroll
    for (int c = 7; c < kRoot; ++c) r[c] = 0.0f;                                    // :520
    r[2] = 0.03f;                                                                   // :537
#pragma unroll
    for (int c = 0; c < kRoot; ++c) p.root[ne * kRoot + c] = r[c];                  // :547
    float2* ds = reinterpret_cast<float2*>(p.state_dof) + ne * p.num_dof;
    for (int j = 0; j < p.num_dof; ++j) ds[j] = make_float2(p.thormang_pose[ne * p.num_dof + j], 0.0f);   // :549-550
    float cfgv[5];
    cfgv[0] = normal_of(p.seat_x, u[5]);                                            // :554
    cfgv[1] = normal_of(p.seat_y, u[6]);                                            // :555
    cfgv[2] = normal_of(p.seat_z, u[7]);                                            // :556
    cfgv[3] = normal_of(p.seat_xr, u[8]);                                           // :557
    cfgv[4] = normal_of(p.steer_off, u[9]);                                         // :559
#pragma unroll
    for (int c = 0; c < 5; ++c) p.config_vector[ne * 5 + c] = cfgv[c];
    p.imu_offsets[ne] = cfgv[3];                                                    // :574
    p.steer_offsets[ne] = cfgv[4];                                                  // :575
    *yaw_cmd = initial_target;                                                      // :589
}

struct Smem {
    alignas(16) float root[kTile * kRoot];
    alignas(16) float hist[kTile * kHist];
    alignas(16) float noise[kTile * kSensorNoise];
    alignas(16) float clean[kTile * kObs];
    alignas(16) float noisy[kTile * kObs];
    alignas(8) uint64_t bar;
    unsigned ballots[2][kWarps];
    float wsum[kWarps][TK_STATS];
};

__device__ __forceinline__ void tile_load(float* dst, const float* src, int cnt, bool vec, int tid) {
    if (vec) {
        const int c4 = cnt >> 2;
        for (int i = tid; i < c4; i += kTile) reinterpret_cast<float4*>(dst)[i] = ld_stream4(src + 4 * i);
        for (int i = (c4 << 2) + tid; i < cnt; i += kTile) dst[i] = src[i];
    } else {
        for (int i = tid; i < cnt; i += kTile) dst[i] = src[i];
    }
}
__device__ __forceinline__ void tile_store(float* dst, const float* src, int cnt, bool vec, int tid) {
    if (vec) {
        const int c4 = cnt >> 2;
        for (int i = tid; i < c4; i += kTile) stg_stream4(dst + 4 * i, reinterpret_cast<const float4*>(src)[i]);
        for (int i = (c4 << 2) + tid; i < cnt; i += kTile) dst[i] = src[i];
    } else {
        for (int i = tid; i < cnt; i += kTile) dst[i] = src[i];
    }
}

// reset_idx :505-590 for env `ne` of the fused kernel, executed by ONE WARP (lane-cooperative; the single-thread
// reset_env above serves the explicit-id-list entry point).  Every lane holds the same 10 unit draws `u`, so every lane
// derives the same scalars with the expressions of reset_env; the rows are spread over the lanes -- lane j the (pos, vel)
// pair of DOF j, j + 32, ..., lanes 0..12 the root row, lanes 0..14 one per-env scalar each -- and ALL global reads are
// requested before any of them is consumed: one memory round trip (~1 us) where the single thread walked 39 + 13 + 5
// rows one after the other (~3 us, and the tiles that held a reset set the length of the whole launch).
// Returns the new root value of column `lane` (lanes 0..12) for the patch of the staged tile.
struct ResetScalars { float yaw_cmd, imu_off; };
struct ResetRows { float pose0, pose1, rv; };
// the rows a reset reads, requested as early as possible (they depend on the env id only): the noise draws are computed
// under their round trip
__device__ __forceinline__ ResetRows reset_env_fetch(const KParams& p, long long ne, int lane) {
    const int d = p.num_dof;
    ResetRows r;
    r.pose0 = (lane < d) ? p.thormang_pose[ne * d + lane] : 0.0f;
    r.pose1 = (lane + 32 < d) ? p.thormang_pose[ne * d + lane + 32] : 0.0f;
    r.rv = (lane < kRoot) ? p.root_reset[ne * kRoot + lane] : 0.0f;                 // :518
    return r;
}
// Part 1: everything of reset_idx that does not depend on the fetched rows -- the per-env scalars, the spawn quaternion.
// Returns the new root value of column `lane` for lanes 2..12 (z, quaternion, zero velocities); columns 0 / 1 (x, y) come
// from root_reset and are filled in by reset_env_rows.  compute_gogoro_observations reads no position column (:696-699),
// so the staged tile is patched with this value alone and the row fetch stays off the tile's critical path.
__device__ __forceinline__ float reset_env_scalars(const KParams& p, long long ne, const float* u, int lane, ResetScalars* out) {
    const float initial_target = (u[3] * 2.0f - 1.0f) * kPi;                        // :486
    const float initial_rot = initial_target + (-1.57f + u[4] * static_cast<float>(1.57 - (-1.57)));   // :487
    const float half = initial_rot / 2.0f;
    float sy, cy;
    sincosf(half, &sy, &cy);
    const float sr = 0.0f, cr = 1.0f, sp = 0.0f, cp = 1.0f;                         // sin/cos of roll/2 = pitch/2 = 0
    const float qx = sr * cp * cy - cr * sp * sy;                                   // :497-500
    const float qy = cr * sp * cy + sr * cp * sy;
    const float qz = cr * cp * sy - sr * sp * cy;
    const float qw = cr * cp * cy + sr * sp * sy;
    float cfgv[5];
    cfgv[0] = normal_of(p.seat_x, u[5]);                                            // :554
    cfgv[1] = normal_of(p.seat_y, u[6]);                                            // :555
    cfgv[2] = normal_of(p.seat_z, u[7]);                                            // :556
    cfgv[3] = normal_of(p.seat_xr, u[8]);                                           // :557
    cfgv[4] = normal_of(p.steer_off, u[9]);                                         // :559
    // root row: [0:3] <- root_reset (z <- 0.03 :537), [3:7] <- spawn quaternion :519, [7:13] <- 0 :520
    float rv = 0.0f;
    rv = (lane == 2) ? 0.03f : rv;
    rv = (lane == 3) ? qx : rv; rv = (lane == 4) ? qy : rv; rv = (lane == 5) ? qz : rv; rv = (lane == 6) ? qw : rv;
    // per-env scalars, one lane each
    float sv = 0.0f;
    float* sp_ = nullptr;
    switch (lane) {
        case 0: sp_ = p.curent_speed + ne; sv = uniform_of(p.speed, u[0]); break;           // :479
        case 1: sp_ = p.curent_speed_offset + ne; sv = uniform_of(p.speed_off, u[2]); break;   // :482 (:481 is overwritten at :575, dead)
        case 2: case 3: case 4: case 5: case 6: sp_ = p.config_vector + ne * 5 + (lane - 2); break;
        case 7: sp_ = p.imu_offsets + ne; sv = cfgv[3]; break;                              // :574
        case 8: sp_ = p.steer_offsets + ne; sv = cfgv[4]; break;                            // :575
        case 9: sp_ = p.curent_command + ne; break;                                         // :587
        case 10: case 11: case 12: case 13: case 14: sp_ = p.action_history + ne * kHist + (lane - 10); break;   // :590
        default: break;
    }
    if (lane >= 2 && lane <= 6) {
        sv = cfgv[0];
        sv = (lane == 3) ? cfgv[1] : sv; sv = (lane == 4) ? cfgv[2] : sv; sv = (lane == 5) ? cfgv[3] : sv; sv = (lane == 6) ? cfgv[4] : sv;
    }
    if (sp_ && !(p.dbg_reset & 2)) *sp_ = sv;
    out->yaw_cmd = initial_target;                                                  // :589
    out->imu_off = cfgv[3];
    return rv;
}
// Part 2: the two row writes that consume the fetched rows (root :547, the 39 locked joints :549-550).  Nothing in the
// step reads them back, so the fused kernel issues them at its very end, when the fetch has long landed.
__device__ __forceinline__ void reset_env_rows(const KParams& p, long long ne, int lane, const ResetRows& rows, float rv_fixed) {
    const int d = p.num_dof;
    const float rv = (lane < 2) ? rows.rv : rv_fixed;
    if (lane < kRoot) p.root[ne * kRoot + lane] = rv;                               // :547
    float2* ds = reinterpret_cast<float2*>(p.state_dof) + ne * d;
    if (lane < d) ds[lane] = make_float2(rows.pose0, 0.0f);                         // :549-550
    if (lane + 32 < d) ds[lane + 32] = make_float2(rows.pose1, 0.0f);
    for (int j = lane + 64; j < d; j += 32) ds[j] = make_float2(p.thormang_pose[ne * d + j], 0.0f);
}

// the 10 unit draws of one reset (same value in every lane of the calling warp)
__device__ __forceinline__ void reset_units(const KParams& p, long long ne, long long row, bool philox, unsigned long long step_now,
                                            int lane, float* u) {
    if (philox && (p.dbg_reset & 8)) {
#pragma unroll
        for (int c = 0; c < kResetNoise; ++c) u[c] = 0.25f;
    } else if (philox) {
        philox_reset_noise_cold(p.seed, step_now, p.env_off + ne, lane, u);
    } else {
#pragma unroll
        for (int c = 0; c < kResetNoise; ++c)
            u[c] = (p.noise_mode != TK_NOISE_NONE && row < p.reset_rows) ? p.reset_unit[row * kResetNoise + c]
                                                                         : ((c == 0 || (c >= 2 && c <= 4)) ? 0.5f : 0.0f);
    }
}

// SPEC variant (thin grids, in-kernel Philox): every thread evaluates the reset of ITS OWN env speculatively -- the draws are a
// pure function of (seed, step, global env id) -- ahead of griddepcontrol.wait (eager launches) or under the tile's round trip
// (CUDA-graph replays, where the epoch word must be read first).  A lone warp walking the ~500 dependent instructions of
// three Philox blocks, three Box-Muller pairs and a sincosf takes 1.5-2.5 us, and in the lazy scheme that walk sat between
// "reset flag landed" and "math may start" of exactly the tiles that end a thin launch (tools/tile_timeline.py: 61 resets in
// 65536 envs made the launch 5.4 us long instead of 3.3 us; with resets disabled the step took 5.8 us instead of 8.4 us).
// Same functions, same bits as philox_reset_noise / reset_env_scalars.
struct SpecReset { float speed, speed_off, cfg[5], qx, qy, qz, qw, yaw_target; };
__device__ __forceinline__ SpecReset spec_reset(const KParams& p, unsigned long long step_now, long long env_global) {
    float u[kResetNoise];
    philox_reset_noise(p.seed, step_now, env_global, u);
    SpecReset r;
    r.speed = uniform_of(p.speed, u[0]);                                            // :479
    r.speed_off = uniform_of(p.speed_off, u[2]);                                    // :482
    const float initial_target = (u[3] * 2.0f - 1.0f) * kPi;                        // :486
    const float initial_rot = initial_target + (-1.57f + u[4] * static_cast<float>(1.57 - (-1.57)));   // :487
    const float half = initial_rot / 2.0f;
    float sy, cy;
    sincosf(half, &sy, &cy);
    const float sr = 0.0f, cr = 1.0f, sp = 0.0f, cp = 1.0f;
    r.qx = sr * cp * cy - cr * sp * sy;                                             // :497-500
    r.qy = cr * sp * cy + sr * cp * sy;
    r.qz = cr * cp * sy - sr * sp * cy;
    r.qw = cr * cp * cy + sr * sp * sy;
    r.cfg[0] = normal_of(p.seat_x, u[5]);                                           // :554
    r.cfg[1] = normal_of(p.seat_y, u[6]);                                           // :555
    r.cfg[2] = normal_of(p.seat_z, u[7]);                                           // :556
    r.cfg[3] = normal_of(p.seat_xr, u[8]);                                          // :557
    r.cfg[4] = normal_of(p.steer_off, u[9]);                                        // :559
    r.yaw_target = initial_target;
    return r;
}
// the per-env scalars of reset_idx, written by the env's own thread from its speculative values
__device__ __forceinline__ void spec_reset_store(const KParams& p, long long ne, const SpecReset& r) {
    p.curent_speed[ne] = r.speed;                                                   // :479
    p.curent_speed_offset[ne] = r.speed_off;                                        // :482
#pragma unroll
    for (int c = 0; c < 5; ++c) p.config_vector[ne * 5 + c] = r.cfg[c];
    p.imu_offsets[ne] = r.cfg[3];                                                   // :574
    p.steer_offsets[ne] = r.cfg[4];                                                 // :575
    p.curent_command[ne] = 0.0f;                                                    // :587
#pragma unroll
    for (int c = 0; c < kHist; ++c) p.action_history[ne * kHist + c] = 0.0f;        // :590
}

// MODE 0: fused post_physics_step; 1: compute_obs_rwd only; 2: observations only (obs_out = clean [N,6])
template <int MODE, bool SPEC = false>
__global__ void __launch_bounds__(kTile) gogoro_tile_kernel(const __grid_constant__ KParams p, float* __restrict__ obs_out) {
    static_assert(!SPEC || MODE == 0, "the speculative reset belongs to the fused step");
    __shared__ Smem s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int env0 = blockIdx.x * kTile;
    const int nt = min(kTile, p.n - env0);
    const long long n = env0 + tid;
    const bool valid = tid < nt;
    const bool vec = p.vec_ok != 0;
    const bool philox = p.noise_mode == TK_NOISE_PHILOX;
    long long* ts = p.dbg_ts ? p.dbg_ts + static_cast<size_t>(blockIdx.x) * 8 : nullptr;
#define TK_GSTAMP(i) do { if (ts && tid == 0) { long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g) :: "memory"); ts[i] = g; } } while (0)
    // programmatic dependent launch (no-ops without the launch attribute): everything up to griddepcontrol.wait touches
    // no global memory and may overlap the drain of the previous kernel in the stream
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const bool use_noise = (MODE != 2) && !philox && p.sensor_unit != nullptr;
    const bool bulk = vec && nt == kTile;
    if (bulk && tid == 0) {
        mbar_init(&s.bar, 1);
        fence_proxy_async_smem();
    }
    // The sensor noise of this step is a pure function of (seed, step, global env id): it is computed right here, under the
    // previous kernel's drain.
    StepNoise sn;
#pragma unroll
    for (int c = 0; c < kSensorNoise; ++c) sn.sensor[c] = 0.0f;
    // L2 prefetch of the tile's inputs ahead of the dependency wait (thin grids; a hint: nothing is consumed, L2 is the
    // point of coherence, stream order is intact).  With the state sets of bench.py rotating over > 2x the L2 -- more than
    // the 256 MB the TLB reaches -- the first touch of a tile costs a page walk plus a DRAM round trip (2.3 us from the
    // wait to the first scalar, 1.3 us when L2-resident: tools/tile_timeline.py --task gogoro --sets 26); both move under
    // the previous kernel's tail.
    if (p.prefetch && bulk) {
        if (tid == 0) bulk_prefetch_l2(p.root + static_cast<size_t>(env0) * kRoot, kTile * kRoot * 4);
        else if (tid == 1 && MODE != 2) bulk_prefetch_l2(p.action_history + static_cast<size_t>(env0) * kHist, kTile * kHist * 4);
        else if (tid == 2 && use_noise) bulk_prefetch_l2(p.sensor_unit + static_cast<size_t>(env0) * kSensorNoise, kTile * kSensorNoise * 4);
        else if (warp == 1) {               // the six per-env scalar arrays: 4 + 4 + 4 + 4 + 8 + 8 lines of 128 bytes
            const char* a = nullptr;
            const int l = lane;
            if (l < 4) a = reinterpret_cast<const char*>(p.yaw_command + env0) + 128 * l;
            else if (l < 8) a = reinterpret_cast<const char*>(p.curent_command + env0) + 128 * (l - 4);
            else if (l < 12) a = (MODE != 2) ? reinterpret_cast<const char*>(p.imu_offsets + env0) + 128 * (l - 8) : nullptr;
            else if (l < 16) a = (MODE == 0 && p.ep_return) ? reinterpret_cast<const char*>(p.ep_return + env0) + 128 * (l - 12) : nullptr;
            else if (l < 24) a = (MODE != 2) ? reinterpret_cast<const char*>(p.progress + env0) + 128 * (l - 16) : nullptr;
            else a = (MODE == 0) ? reinterpret_cast<const char*>(p.reset + env0) + 128 * (l - 24) : nullptr;
            if (a) prefetch_l2(a);
        } else if (warp == 2 && MODE == 0) {
            // one line of every array the tile writes or (on a reset) reads late: not for the data, for the TRANSLATION --
            // a first touch of a page the TLB has dropped is a page walk of several dependent DRAM reads (~2 us measured
            // on the reset rows), and a store that misses holds back the completion of the whole grid
            const void* a = nullptr;
            switch (lane) {
                case 0: a = p.thormang_pose + static_cast<size_t>(env0) * p.num_dof; break;
                case 1: a = p.root_reset + static_cast<size_t>(env0) * kRoot; break;
                case 2: a = p.state_dof + static_cast<size_t>(env0) * p.num_dof * 2; break;
                case 3: a = p.obs_buf + static_cast<size_t>(env0) * kObs; break;
                case 4: a = p.buffer_obs + static_cast<size_t>(env0) * kObs; break;
                case 5: a = p.rew + env0; break;
                case 6: a = p.timeout ? p.timeout + env0 : nullptr; break;
                case 7: a = p.curent_speed + env0; break;             // the arrays only a reset / a resample writes
                case 8: a = p.curent_speed_offset + env0; break;
                case 9: a = p.config_vector + static_cast<size_t>(env0) * 5; break;
                case 10: a = p.steer_offsets + env0; break;
                default: break;
            }
            if (a) prefetch_l2(a);
        }
    }
    // The noise epoch (CUDA-graph replays) is read AHEAD of the wait as well: the word is written only by
    // tk_gogoro_advance_epoch's single-thread kernel, a plain launch that never triggers its dependents early -- whatever
    // follows it in the stream starts after it has completed, and this grid starts no earlier than its own predecessor.
    const bool early_noise = (MODE != 2) && philox;
    unsigned long long step_now = p.step;
    if (early_noise && p.step_dev) {
        unsigned long long ep;
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(ep) : "l"(p.step_dev) : "memory");
        step_now = p.step + (ep << 32);
    }
    if (early_noise) sn = philox_sensor_noise(p.seed, step_now, p.env_off + n);
    SpecReset spec;
    if (SPEC && early_noise) spec = spec_reset(p, step_now, p.env_off + n);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    TK_GSTAMP(0);

    // full, 16-byte aligned tile: the bulk-copy (TMA) engine moves root / history / noise rows in and both obs rows out
    // Developer knob (p.ldg_tiles, TK_GOGORO_LDG=1): the two input tiles through the load/store unit instead -- coalesced 128-bit
    // loads held in registers across the scalar requests, stored to shared memory in front of the CTA barrier below.  Tried
    // because a lone plain load lands 0.3-0.5 us after the dependency wait and the bulk copies 1.0-1.5 us after it
    // (tools/probe_first_load.cu, tools/tile_timeline.py); measured SLOWER (see fill_kparams): the bulk copy is not the
    // bottleneck of the thin launch.
    const bool ldg = bulk && p.ldg_tiles && !use_noise && MODE == 0;
    constexpr int kRoot4 = kTile * kRoot / 4, kHist4 = kTile * kHist / 4;      // 416 and 160 float4 per tile
    float4 rt[(kRoot4 + kTile - 1) / kTile], ht[(kHist4 + kTile - 1) / kTile];
    if (ldg) {
        const float4* gr = reinterpret_cast<const float4*>(p.root + static_cast<size_t>(env0) * kRoot);
        const float4* gh = reinterpret_cast<const float4*>(p.action_history + static_cast<size_t>(env0) * kHist);
#pragma unroll
        for (int k = 0; k < (kRoot4 + kTile - 1) / kTile; ++k)
            if (tid + kTile * k < kRoot4) rt[k] = gr[tid + kTile * k];
#pragma unroll
        for (int k = 0; k < (kHist4 + kTile - 1) / kTile; ++k)
            if (tid + kTile * k < kHist4) ht[k] = gh[tid + kTile * k];
    } else if (bulk) {
        if (tid == 0) mbar_expect_tx(&s.bar, (kRoot + (MODE != 2 ? kHist : 0) + (use_noise ? kSensorNoise : 0)) * kTile * 4);
        __syncwarp();
        if (tid == 0) bulk_load(s.root, p.root + static_cast<size_t>(env0) * kRoot, kTile * kRoot * 4, &s.bar);
        else if (tid == 1 && MODE != 2) bulk_load(s.hist, p.action_history + static_cast<size_t>(env0) * kHist, kTile * kHist * 4, &s.bar);
        else if (tid == 2 && use_noise) bulk_load(s.noise, p.sensor_unit + static_cast<size_t>(env0) * kSensorNoise, kTile * kSensorNoise * 4, &s.bar);
    } else {
        tile_load(s.root, p.root + static_cast<size_t>(env0) * kRoot, nt * kRoot, false, tid);
        if (MODE != 2) {
            tile_load(s.hist, p.action_history + static_cast<size_t>(env0) * kHist, nt * kHist, false, tid);
            if (use_noise) tile_load(s.noise, p.sensor_unit + static_cast<size_t>(env0) * kSensorNoise, nt * kSensorNoise, false, tid);
        }
    }

    long long prog = 0;
    bool do_reset = false, do_speed = false;
    float yaw_cmd = 0.f, cur_cmd = 0.f, imu_off = 0.f, ret_old = 0.f;
    if (valid) {
        // every per-env scalar is requested up front, together with the tile copies: one memory round trip per tile
        yaw_cmd = p.yaw_command[n];
        cur_cmd = p.curent_command[n];
        if (MODE != 2) imu_off = p.imu_offsets[n];
        if (MODE == 0 && p.ep_return) ret_old = p.ep_return[n];
        if (MODE != 2) prog = p.progress[n];
        if (MODE == 0) do_reset = p.reset[n] != 0 && !(p.dbg_reset & 1);            // :377
    }
    if (MODE == 0 && valid) {
        prog += 1;                                                                  // :375
        if (do_reset) prog = 0;                                                     // :583
        do_speed = prog == p.speed_freq;                                            // :384 (after the reset)
    }
    unsigned reset_bits = 0;
    if (MODE == 0) {
        reset_bits = __ballot_sync(0xffffffffu, do_reset);
        const unsigned b1 = __ballot_sync(0xffffffffu, do_speed);
        if (lane == 0) { s.ballots[0][warp] = reset_bits; s.ballots[1][warp] = b1; }
    }
    if (ldg) {
#pragma unroll
        for (int k = 0; k < (kRoot4 + kTile - 1) / kTile; ++k)
            if (tid + kTile * k < kRoot4) reinterpret_cast<float4*>(s.root)[tid + kTile * k] = rt[k];
#pragma unroll
        for (int k = 0; k < (kHist4 + kTile - 1) / kTile; ++k)
            if (tid + kTile * k < kHist4) reinterpret_cast<float4*>(s.hist)[tid + kTile * k] = ht[k];
    }
    TK_GSTAMP(1);
    __syncthreads();                       // ballots, mbarrier init, cooperative copies visible to the whole CTA

    // ---- reset prologue (rare): this warp's flagged envs, lane-cooperative, BEFORE the tile is waited for -- the new
    // state does not depend on the staged rows; only the patch of the staged copy comes after the wait
    constexpr int kPre = 2;
    float pre_root[kPre];
    ResetRows pre_rows[kPre];
    int pre_env[kPre];
    bool hist_zero = false;
    if (SPEC && reset_bits) {              // only the rows are fetched here; everything else is already in registers
        unsigned bits = reset_bits;
#pragma unroll
        for (int q = 0; q < kPre; ++q) {
            pre_root[q] = 0.0f;
            pre_env[q] = -1;
            pre_rows[q] = ResetRows{0.0f, 0.0f, 0.0f};
            if (bits) {
                pre_env[q] = __ffs(bits) - 1;
                bits &= bits - 1;
                pre_rows[q] = reset_env_fetch(p, env0 + warp * 32 + pre_env[q], lane);
            }
        }
        if (do_reset) { yaw_cmd = spec.yaw_target; imu_off = spec.cfg[3]; cur_cmd = 0.0f; hist_zero = true; }   // :589,574,587,590
    }
    if (!SPEC && MODE == 0 && reset_bits) {
        int rank = 0;
        if (p.noise_mode == TK_NOISE_COMPACT) {
            rank = p.tile_base_reset[blockIdx.x];
            for (int w = 0; w < warp; ++w) rank += __popc(s.ballots[0][w]);
        }
        unsigned bits = reset_bits;
#pragma unroll
        for (int q = 0; q < kPre; ++q) {
            pre_root[q] = 0.0f;
            pre_env[q] = -1;
            pre_rows[q] = ResetRows{0.0f, 0.0f, 0.0f};
            if (bits) {
                const int e = __ffs(bits) - 1;
                bits &= bits - 1;
                const long long ne = env0 + warp * 32 + e;
                float u[kResetNoise];
                pre_env[q] = e;
                reset_units(p, ne, (p.noise_mode == TK_NOISE_COMPACT) ? rank + q : ne, philox, step_now, lane, u);
                ResetScalars rs;
                pre_root[q] = reset_env_scalars(p, ne, u, lane, &rs);
                if (lane == e) { yaw_cmd = rs.yaw_cmd; imu_off = rs.imu_off; cur_cmd = 0.0f; hist_zero = true; }   // :589,574,587,590
            }
        }
        // the rows reset_env_rows writes back at the end of the kernel: requested AFTER the out-of-line Philox calls (a
        // value in flight cannot live across a call: it would be waited for and spilled)
#pragma unroll
        for (int q = 0; q < kPre; ++q)
            if (pre_env[q] >= 0 && !(p.dbg_reset & 4)) pre_rows[q] = reset_env_fetch(p, env0 + warp * 32 + pre_env[q], lane);
    }
    TK_GSTAMP(2);
    if (bulk && !ldg) mbar_wait(&s.bar, 0);        // bulk loads landed
    if (!SPEC && MODE == 0 && reset_bits) {         // patch the staged rows of the envs that were reset
        int rank = 0;
        if (p.noise_mode == TK_NOISE_COMPACT) {
            rank = p.tile_base_reset[blockIdx.x];
            for (int w = 0; w < warp; ++w) rank += __popc(s.ballots[0][w]);
        }
        unsigned bits = reset_bits;
        for (int q = 0; bits; ++q) {
            const int e = __ffs(bits) - 1;
            bits &= bits - 1;
            float rv;
            if (q < kPre) {
                rv = (q == 0) ? pre_root[0] : pre_root[kPre - 1];
            } else {                       // more than kPre resets in one warp: the rest after the wait
                const long long ne = env0 + warp * 32 + e;
                float u[kResetNoise];
                const ResetRows rows = reset_env_fetch(p, ne, lane);
                reset_units(p, ne, (p.noise_mode == TK_NOISE_COMPACT) ? rank + q : ne, philox, step_now, lane, u);
                ResetScalars rs;
                rv = reset_env_scalars(p, ne, u, lane, &rs);
                reset_env_rows(p, ne, lane, rows, rv);
                if (lane == e) { yaw_cmd = rs.yaw_cmd; imu_off = rs.imu_off; cur_cmd = 0.0f; hist_zero = true; }
            }
            // columns 2..12 only: x / y of the staged row are not read by the observation (and for q < kPre not known yet)
            if (lane >= 2 && lane < kRoot) s.root[(warp * 32 + e) * kRoot + lane] = rv;
        }
        __syncwarp();
    }
    TK_GSTAMP(3);

    float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_tilt = 0.f, st_yaw = 0.f;
    bool out_reset = false, out_timeout = false;
    if (valid) {
        float r[kRoot], hist[kHist];
#pragma unroll
        for (int c = 0; c < kRoot; ++c) r[c] = s.root[tid * kRoot + c];
        if (SPEC && do_reset) {             // the new root row (:518-520): position columns are not read by the observation
            r[3] = spec.qx; r[4] = spec.qy; r[5] = spec.qz; r[6] = spec.qw;
#pragma unroll
            for (int c = 7; c < kRoot; ++c) r[c] = 0.0f;
        }
        if (MODE != 2) {
#pragma unroll
            for (int c = 0; c < kHist; ++c) hist[c] = hist_zero ? 0.0f : s.hist[tid * kHist + c];
            if (!philox) {
#pragma unroll
                for (int c = 0; c < kSensorNoise; ++c) sn.sensor[c] = p.sensor_unit ? s.noise[tid * kSensorNoise + c] : 0.0f;
            }
        }

        float o[kObs];
        observe(r, yaw_cmd, cur_cmd, o);                                            // :430-434
        if (MODE == 2) {
#pragma unroll
            for (int c = 0; c < kObs; ++c) s.clean[tid * kObs + c] = o[c];
        } else {
#pragma unroll
            for (int c = 0; c < kObs; ++c) s.clean[tid * kObs + c] = o[c];          // :436-437 (buff_size 1: plain overwrite)
            bool felt;
            float reward = reward_of(o, hist, &felt);                               // :439-444
            const bool finished = prog >= p.max_episode_length - 1;                 // :678
            out_reset = felt || finished;                                           // :679-680
            p.rew[n] = reward;
            p.reset[n] = out_reset ? 1 : 0;
            // sensor-noise model on a copy :449-460
            float z[kObs];
            z[0] = o[0] + (normal_of(p.imu_filter, sn.sensor[0]) + imu_off);             // :451
            z[1] = o[1] + normal_of(p.imu, sn.sensor[1]);                           // :452
            z[2] = o[2] + normal_of(p.imu, sn.sensor[2]);                           // :453
            z[3] = rintf(o[4]);                                                     // :455-458: channel 3 <- round(channel 4), sic
            z[4] = o[4] + normal_of(p.imu_filter, sn.sensor[4]);                    // :460
            z[5] = o[5];
#pragma unroll
            for (int c = 0; c < kObs; ++c) s.noisy[tid * kObs + c] = z[c];          // :462
            st_rew = reward; st_tilt = fabsf(o[0]); st_yaw = fabsf(o[4]);
        }
        if (MODE == 0) {
            p.progress[n] = prog;
            // command resampling :384-389
            const bool do_yaw = prog == p.yaw_freq;
            float su = 0.5f, yu = 0.5f;
            if (do_speed || do_yaw) {                                               // rare: second Philox block
                if (philox) {
                    MiscNoise mn = philox_misc_noise(p.seed, step_now, p.env_off + n);
                    su = mn.speed_u; yu = mn.yaw_u;
                } else if (p.yaw_unit) {
                    yu = p.yaw_unit[n];
                }
            }
            if (do_speed) {
                if (!philox && p.speed_unit) {
                    long long row = n;
                    if (p.noise_mode == TK_NOISE_COMPACT) {
                        int rank = p.tile_base_speed[blockIdx.x];
                        for (int w = 0; w < warp; ++w) rank += __popc(s.ballots[1][w]);
                        rank += __popc(s.ballots[1][warp] & ((1u << lane) - 1u));
                        row = rank;
                    }
                    su = p.speed_unit[row];
                }
                p.curent_speed[n] = uniform_of(p.speed, su);                        // :386
            }
            if (do_yaw) yaw_cmd = -kPi + yu * kTwoPi;                               // :387
            yaw_cmd = (yaw_cmd > kPi) ? yaw_cmd - kTwoPi : yaw_cmd;                 // :388
            yaw_cmd = (yaw_cmd < -kPi) ? yaw_cmd + kTwoPi : yaw_cmd;                // :389
            p.yaw_command[n] = yaw_cmd;
            out_timeout = (prog >= p.max_episode_length - 1) && out_reset;          // vec_task.py:345
            if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
            if (p.ep_return) {
                float ret = (do_reset ? 0.0f : ret_old) + st_rew;
                if (out_reset) st_ret = ret;
                p.ep_return[n] = ret;
            }
            if (out_reset) st_len = static_cast<float>(prog);
        }
    }

    TK_GSTAMP(4);
    if (MODE == 0) {
        const unsigned m_reset = __ballot_sync(0xffffffffu, out_reset);
        const unsigned m_to = __ballot_sync(0xffffffffu, out_timeout);
        const unsigned m_valid = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_ret += __shfl_xor_sync(0xffffffffu, st_ret, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            st_tilt += __shfl_xor_sync(0xffffffffu, st_tilt, o);
            st_yaw += __shfl_xor_sync(0xffffffffu, st_yaw, o);
        }
        if (lane == 0) {
            float* w = s.wsum[warp];
            w[0] = st_rew; w[1] = static_cast<float>(__popc(m_reset)); w[2] = static_cast<float>(__popc(m_to));
            w[3] = st_ret; w[4] = st_len; w[5] = st_tilt; w[6] = st_yaw; w[7] = static_cast<float>(__popc(m_valid));
        }
    }
    fence_proxy_async_smem();              // obs rows written via the generic proxy, read by the bulk-copy engine
    __syncthreads();
    TK_GSTAMP(5);
    if (MODE == 0 && tid < TK_STATS) {
        double acc = 0.0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) acc += static_cast<double>(s.wsum[w][tid]);
        // running sum since the last tk_*_stats (slot scheme: tk_common.cuh)
        atomicAdd(p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS + tid, acc);
    }
    TK_GSTAMP(7);
    float* out_clean = (MODE == 2) ? obs_out + static_cast<size_t>(env0) * kObs : p.buffer_obs + static_cast<size_t>(env0) * kObs;
    const bool bulk_out = bulk && (MODE != 2 || (reinterpret_cast<uintptr_t>(obs_out) & 15) == 0);
    if (bulk_out) {
        if (tid == 0) bulk_store(out_clean, s.clean, kTile * kObs * 4);
        if (tid == 32 && MODE != 2) bulk_store(p.obs_buf + static_cast<size_t>(env0) * kObs, s.noisy, kTile * kObs * 4);
        if (tid == 0 || (tid == 32 && MODE != 2)) bulk_commit();
    } else {
        tile_store(out_clean, s.clean, nt * kObs, false, tid);
        if (MODE != 2) tile_store(p.obs_buf + static_cast<size_t>(env0) * kObs, s.noisy, nt * kObs, false, tid);
    }
    if (SPEC && reset_bits) {              // per-env scalars by the env's own thread, rows warp-cooperatively (lane e holds env e's quaternion)
        if (do_reset) spec_reset_store(p, n, spec);
        unsigned bits = reset_bits;
        for (int q = 0; bits; ++q) {
            const int e = __ffs(bits) - 1;
            bits &= bits - 1;
            const float qx = __shfl_sync(0xffffffffu, spec.qx, e), qy = __shfl_sync(0xffffffffu, spec.qy, e);
            const float qz = __shfl_sync(0xffffffffu, spec.qz, e), qw = __shfl_sync(0xffffffffu, spec.qw, e);
            float rv = 0.0f;
            rv = (lane == 2) ? 0.03f : rv;                                          // :537
            rv = (lane == 3) ? qx : rv; rv = (lane == 4) ? qy : rv; rv = (lane == 5) ? qz : rv; rv = (lane == 6) ? qw : rv;
            const long long ne = env0 + warp * 32 + e;
            ResetRows rows;
            if (q == 0) rows = pre_rows[0];
            else if (q == 1) rows = pre_rows[kPre - 1];
            else rows = reset_env_fetch(p, ne, lane);                               // more than kPre resets in one warp
            reset_env_rows(p, ne, lane, rows, rv);
        }
    }
    if (!SPEC && MODE == 0 && reset_bits) {         // the deferred row writes of this warp's first kPre resets
#pragma unroll
        for (int q = 0; q < kPre; ++q)
            if (pre_env[q] >= 0 && !(p.dbg_reset & 4)) reset_env_rows(p, env0 + warp * 32 + pre_env[q], lane, pre_rows[q], pre_root[q]);
    }
    if (bulk_out && (tid == 0 || (tid == 32 && MODE != 2))) bulk_wait_read<0>();   // shared memory must stay alive until the engine has read it
    TK_GSTAMP(6);
#undef TK_GSTAMP
}

// per-tile counts of the two COMPACT predicates: reset != 0 ; reset == 0 && progress + 1 == speed_freq
__global__ void tile_count_kernel(const long long* __restrict__ reset, const long long* __restrict__ progress, int n, int tiles,
                                  int speed_freq, int* __restrict__ c_reset, int* __restrict__ c_speed) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= tiles) return;
    int a = 0, b = 0;
    for (int i = lane; i < kTile; i += 32) {
        int e = warp * kTile + i;
        if (e < n) {
            const bool rs = reset[e] != 0;
            a += rs ? 1 : 0;
            b += (!rs && progress[e] + 1 == speed_freq) ? 1 : 0;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (lane == 0) { c_reset[warp] = a; c_speed[warp] = b; }
}

__global__ void exclusive_scan_kernel(const int* __restrict__ counts, int tiles, int* __restrict__ base) {
    // single CTA, sequential over chunks of blockDim.x tiles
    __shared__ int warp_tot[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int start = 0; start < tiles; start += blockDim.x) {
        const int i = start + threadIdx.x;
        const int v = (i < tiles) ? counts[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int t = (lane < (blockDim.x >> 5)) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            warp_tot[lane] = t;
        }
        __syncthreads();
        const int before = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + (x - v);
        if (i < tiles) base[i] = before;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = before + v;
        __syncthreads();
    }
}

// pre_physics_step :347-369, one env per thread
__global__ void __launch_bounds__(256) gogoro_pre_kernel(const __grid_constant__ KParams p, const float* __restrict__ actions,
                                                          const float* __restrict__ steer_unit, int zero_fill) {
    const long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (zero_fill) {
        // fresh torch.zeros targets of :360,366 (376 B/env mode): flat coalesced fill by the whole grid
        const long long total = static_cast<long long>(p.n) * p.num_dof;
        // (the two live columns are written by a second launch of this kernel with zero_fill = 0)
        for (long long i = n; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) { p.pos_cmd[i] = 0.0f; p.vel_cmd[i] = 0.0f; }
        return;
    }
    if (n >= p.n) return;
    const float a = actions[n];                                                     // actions[:,0], num_acts = 1
    float* h = p.action_history + n * kHist;
    float h1 = h[1], h2 = h[2], h3 = h[3], h4 = h[4];
    h[0] = h1; h[1] = h2; h[2] = h3; h[3] = h4; h[4] = a;                           // :349-350
    float a2 = a * p.max_steering_change;                                           // :353
    a2 = fminf(fmaxf(a2, -p.max_steering_change), p.max_steering_change);           // :354
    float cmd = p.curent_command[n] + a2;                                           // :355
    cmd = fminf(fmaxf(cmd, -p.max_steering), p.max_steering);                       // :359
    p.curent_command[n] = cmd;
    float z = steer_unit ? steer_unit[n] : philox_misc_noise(p.seed, effective_step(p.step, p.step_dev), p.env_off + n).steer;
    float noise = normal_of(p.steer_action, z);                                     // :362
    p.pos_cmd[n * p.num_dof + p.steering_joint] = cmd + p.steer_offsets[n] + noise; // :363
    p.vel_cmd[n * p.num_dof + p.rear_wheel_joint] = p.curent_speed[n];              // :367
}

// compute_gogoro_reward on an explicit obs tensor (method-level entry)
__global__ void gogoro_reward_kernel(const __grid_constant__ KParams p, const float* __restrict__ obs) {
    const long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (n >= p.n) return;
    float o[kObs], hist[kHist];
#pragma unroll
    for (int c = 0; c < kObs; ++c) o[c] = obs[n * kObs + c];
#pragma unroll
    for (int c = 0; c < kHist; ++c) hist[c] = p.action_history[n * kHist + c];
    bool felt;
    float reward = reward_of(o, hist, &felt);
    const bool finished = p.progress[n] >= p.max_episode_length - 1;
    p.rew[n] = reward;
    p.reset[n] = (felt || finished) ? 1 : 0;
}

// reset_idx :505-590 for an explicit id list, one thread per listed env
__global__ void gogoro_reset_idx_kernel(const __grid_constant__ KParams p, const long long* __restrict__ ids, int k,
                                        const float* __restrict__ unit) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    const long long ne = ids[i];
    if (ne < 0 || ne >= p.n) return;
    float u[kResetNoise], r[kRoot], yaw_cmd;
#pragma unroll
    for (int c = 0; c < kResetNoise; ++c) u[c] = unit[static_cast<size_t>(i) * kResetNoise + c];
    reset_env(p, ne, u, r, &yaw_cmd);
    p.progress[ne] = 0;                                                             // :583
    p.reset[ne] = 0;                                                                // :584
#pragma unroll
    for (int c = 0; c < kObs; ++c) { p.obs_buf[ne * kObs + c] = 0.0f; p.buffer_obs[ne * kObs + c] = 0.0f; }   // :585-586
    p.curent_command[ne] = 0.0f;                                                    // :587
    p.yaw_command[ne] = yaw_cmd;                                                    // :589
#pragma unroll
    for (int c = 0; c < kHist; ++c) p.action_history[ne * kHist + c] = 0.0f;        // :590
}

__global__ void gogoro_philox_kernel(int n, unsigned long long seed, unsigned long long step, long long env_off,
                                     const unsigned long long* __restrict__ epoch, float* steer,
                                     float* sensor, float* reset_unit, float* speed, float* yaw) {
    step = effective_step(step, epoch);   // the draws of the NEXT pre/post-physics launches
    const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= n) return;
    StepNoise sn = philox_sensor_noise(seed, step, env_off + e);
    MiscNoise mn = philox_misc_noise(seed, step, env_off + e);
    sn.sensor[3] = mn.dead3;   // the draw of gogoro_new.py:455 never reaches an output
    if (steer) steer[e] = mn.steer;
    if (sensor) for (int c = 0; c < kSensorNoise; ++c) sensor[e * kSensorNoise + c] = sn.sensor[c];
    if (speed) speed[e] = mn.speed_u;
    if (yaw) yaw[e] = mn.yaw_u;
    if (reset_unit) {
        float u[kResetNoise];
        philox_reset_noise(seed, step, env_off + e, u);
        for (int c = 0; c < kResetNoise; ++c) reset_unit[e * kResetNoise + c] = u[c];
    }
}

__global__ void stats_finalize_kernel(double* __restrict__ partials, int rows, double* __restrict__ out, int accumulate) {
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k >= TK_STATS) return;
    double acc = 0.0;
    for (int r = lane; r < rows; r += 32) {
        acc += partials[static_cast<size_t>(r) * TK_STATS + k];
        partials[static_cast<size_t>(r) * TK_STATS + k] = 0.0;   // the partials restart from zero
    }
    acc = warp_sum(acc);
    if (lane == 0) out[k] = accumulate ? out[k] + acc : acc;
}

}  // namespace gogoro
}  // namespace tk

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace tk;
using namespace tk::gogoro;

struct TkGogoro {
    TkGogoroParams params;
    TkGogoroBuffers buf;
    bool bound;
    int tiles;
    double* partials;
    int* counts;   // [2, tiles]
    int* bases;    // [2, tiles]
    int last_grid;
    unsigned long long* step_dev;
    bool epoch_used;      // tk_gogoro_advance_epoch was called: the kernels must read the device-side epoch (NULL otherwise)
    long long* dbg_ts;
    int prefetch;         // TK_GOGORO_PREFETCH: L2 prefetch ahead of griddepcontrol.wait (default: thin grids only)
    int knob_ldg, knob_spec, dbg_reset;   // developer knobs, read once in tk_gogoro_create
};

static void pair_uniform(const double* in, float* out) { out[0] = static_cast<float>(in[0]); out[1] = static_cast<float>(in[1] - in[0]); }
static void pair_normal(const double* in, float* out) { out[0] = static_cast<float>(in[0]); out[1] = static_cast<float>(in[1]); }

// The device-side epoch word, or NULL while nobody can have bumped it (saves the eager path one dependent load).  A launch
// that is being CAPTURED always gets the word: tk_gogoro_advance_epoch may be recorded after it in the same graph, or be
// called between replays, and a NULL baked into the graph would make every replay draw the same Philox noise.
static unsigned long long* epoch_word(const TkGogoro* h, cudaStream_t st) {
    if (h->epoch_used) return h->step_dev;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return h->step_dev; }
    return cs == cudaStreamCaptureStatusNone ? nullptr : h->step_dev;
}

static void fill_kparams(const TkGogoro* h, KParams& k, cudaStream_t st) {
    const TkGogoroParams& P = h->params;
    const TkGogoroBuffers& B = h->buf;
    k.root = B.root_tensor; k.state_dof = B.state_dof; k.obs_buf = B.obs_buf; k.buffer_obs = B.buffer_obs; k.rew = B.rew_buf;
    k.reset = reinterpret_cast<long long*>(B.reset_buf); k.progress = reinterpret_cast<long long*>(B.progress_buf);
    k.timeout = B.timeout_buf; k.yaw_command = B.yaw_command; k.curent_command = B.curent_command;
    k.curent_speed = B.curent_speed; k.steer_offsets = B.steer_offsets; k.imu_offsets = B.imu_offsets;
    k.curent_speed_offset = B.curent_speed_offset; k.action_history = B.action_history; k.config_vector = B.config_vector;
    k.thormang_pose = B.thormang_pose; k.root_reset = B.root_reset_tensor; k.pos_cmd = B.dof_pos_command;
    k.vel_cmd = B.dof_speed_command; k.ep_return = B.episode_return;
    k.partials = h->partials; k.tile_base_reset = h->bases; k.tile_base_speed = h->bases + h->tiles;
    k.reset_unit = nullptr; k.sensor_unit = nullptr; k.speed_unit = nullptr; k.yaw_unit = nullptr; k.reset_rows = 0;
    k.seed = 0; k.step = 0; k.env_off = 0; k.noise_mode = TK_NOISE_NONE; k.step_dev = epoch_word(h, st); k.dbg_ts = h->dbg_ts;
    k.n = P.num_envs; k.num_dof = P.num_dof; k.steering_joint = P.steering_joint; k.rear_wheel_joint = P.rear_wheel_joint;
    k.max_episode_length = P.max_episode_length; k.speed_freq = P.speed_freq_update; k.yaw_freq = P.yaw_freq_update;
    k.vec_ok = aligned_to(B.root_tensor, 16) && aligned_to(B.action_history, 16) && aligned_to(B.obs_buf, 16) &&
               aligned_to(B.buffer_obs, 16);
    k.prefetch = h->prefetch;
    k.ldg_tiles = 0;              // measured slower: 6.7 vs 6.3 us eager, 6.1 vs 5.8 us from a graph at 65536 envs; equal at 131072 / 262144
    if (h->knob_ldg != kKnobUnset) k.ldg_tiles = h->knob_ldg;
    k.dbg_reset = h->dbg_reset;
    k.max_steering = P.max_steering; k.max_steering_change = P.max_steering_change;
    pair_normal(P.imu_filter_noise, k.imu_filter); pair_normal(P.imu_noise, k.imu); pair_normal(P.speed_sensor_noise, k.speed_sensor);
    pair_normal(P.steering_action_noise, k.steer_action); pair_uniform(P.speed_range, k.speed);
    pair_normal(P.seat_offset_x_range, k.seat_x); pair_normal(P.seat_offset_y_range, k.seat_y);
    pair_normal(P.seat_offset_z_range, k.seat_z); pair_normal(P.seat_offset_xr_range, k.seat_xr);
    pair_normal(P.steering_offset, k.steer_off); pair_uniform(P.speed_sensor_offset, k.speed_off);
}

static int check_bound(const TkGogoro* h) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(h->bound, TK_ERR_UNBOUND, "tk_gogoro_bind was not called");
    return TK_OK;
}

extern "C" {

int tk_gogoro_create(const TkGogoroParams* params, TkGogoro** out) {
    TK_REQUIRE(params && out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(params->num_envs > 0, TK_ERR_INVALID, "num_envs must be positive (got %d)", params->num_envs);
    TK_REQUIRE(params->num_dof > 0 && params->num_dof <= 4096, TK_ERR_INVALID, "bad num_dof %d", params->num_dof);
    TK_REQUIRE(params->steering_joint >= 0 && params->steering_joint < params->num_dof && params->rear_wheel_joint >= 0 &&
                   params->rear_wheel_joint < params->num_dof,
               TK_ERR_INVALID, "joint index out of range");
    TK_CUDA(cudaSetDevice(params->device));
    TkGogoro* h = new (std::nothrow) TkGogoro();
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "out of host memory");
    h->params = *params;
    h->bound = false;
    h->tiles = (params->num_envs + kTile - 1) / kTile;
    h->last_grid = 0;
    h->dbg_ts = nullptr;
    h->epoch_used = false;
    h->knob_ldg = knob_int("TK_GOGORO_LDG");
    h->knob_spec = knob_int("TK_GOGORO_SPEC");
    h->dbg_reset = ablation_knob("TK_GOGORO_DBG_RESET");
    {   // the prefetch pays while the whole grid is resident at once (the CTAs sit at the wait while the previous kernel
        // drains); on deep grids it would only double the requests of a bandwidth-bound stream
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, params->device);
        // measured (HBM-cold, prefetch + speculative reset on / off): 65536 envs 6.2 / 8.4 us, 131072 7.1 / 9.8, 262144 13.5 / 14.7,
        // 1048576 47.6 / 40.0 -- the switch sits at two waves of resident CTAs
        h->prefetch = (h->tiles <= 16 * sms) ? 1 : 0;
        if (const char* env = knob("TK_GOGORO_PREFETCH")) h->prefetch = atoi(env);
    }
    h->partials = nullptr; h->counts = nullptr; h->bases = nullptr; h->step_dev = nullptr;
    cudaError_t e = cudaMalloc(&h->partials, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMalloc(&h->counts, sizeof(int) * 2 * h->tiles);
    if (e == cudaSuccess) e = cudaMalloc(&h->bases, sizeof(int) * 2 * h->tiles);
    if (e == cudaSuccess) e = cudaMemset(h->partials, 0, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMemset(h->bases, 0, sizeof(int) * 2 * h->tiles);
    if (e == cudaSuccess) e = cudaMalloc(&h->step_dev, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->step_dev, 0, sizeof(unsigned long long));
    if (e != cudaSuccess) {
        cudaFree(h->partials); cudaFree(h->counts); cudaFree(h->bases); cudaFree(h->step_dev);
        delete h;
        return cuda_fail(e, "tk_gogoro_create");
    }
    *out = h;
    return TK_OK;
}

int tk_gogoro_destroy(TkGogoro* h) {
    if (!h) return TK_OK;
    cudaSetDevice(h->params.device);
    cudaFree(h->partials); cudaFree(h->counts); cudaFree(h->bases); cudaFree(h->step_dev);
    delete h;
    return TK_OK;
}

int tk_gogoro_bind(TkGogoro* h, const TkGogoroBuffers* b) {
    TK_REQUIRE(h && b, TK_ERR_INVALID, "null argument");
    const void* req[] = {b->root_tensor, b->state_dof, b->obs_buf, b->buffer_obs, b->rew_buf, b->reset_buf, b->progress_buf,
                         b->yaw_command, b->curent_command, b->curent_speed, b->steer_offsets, b->imu_offsets,
                         b->curent_speed_offset, b->action_history, b->config_vector, b->thormang_pose, b->root_reset_tensor,
                         b->dof_pos_command, b->dof_speed_command};
    const char* names[] = {"root_tensor", "state_dof", "obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf",
                           "yaw_command", "curent_command", "curent_speed", "steer_offsets", "imu_offsets",
                           "curent_speed_offset", "action_history", "config_vector", "thormang_pose", "root_reset_tensor",
                           "dof_pos_command", "dof_speed_command"};
    for (unsigned i = 0; i < sizeof(req) / sizeof(req[0]); ++i) {
        TK_REQUIRE(req[i] != nullptr, TK_ERR_UNBOUND, "buffer %s is NULL", names[i]);
        TK_REQUIRE(aligned_to(req[i], 4), TK_ERR_ALIGN, "buffer %s is not 4-byte aligned", names[i]);
    }
    TK_REQUIRE(aligned_to(b->reset_buf, 8) && aligned_to(b->progress_buf, 8), TK_ERR_ALIGN, "int64 buffers must be 8-byte aligned");
    TK_REQUIRE(aligned_to(b->state_dof, 8), TK_ERR_ALIGN, "state_dof must be 8-byte aligned ((pos,vel) pairs)");
    h->buf = *b;
    h->bound = true;
    return TK_OK;
}

int tk_gogoro_pre_physics(TkGogoro* h, const float* actions, const float* steer_unit, uint64_t seed, uint64_t step,
                          int64_t env_id_offset, int32_t zero_fill, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(actions != nullptr, TK_ERR_INVALID, "null actions");
    TK_CUDA(cudaSetDevice(h->params.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    k.seed = seed; k.step = step; k.env_off = env_id_offset;
    const int blocks = (k.n + 255) / 256;
    if (zero_fill) {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->params.device);
        const long long want = (static_cast<long long>(k.n) * k.num_dof + 255) / 256;
        const int zb = static_cast<int>(want < static_cast<long long>(sms) * 8 ? (want < 1 ? 1 : want) : static_cast<long long>(sms) * 8);
        gogoro_pre_kernel<<<zb, 256, 0, st>>>(k, actions, steer_unit, 1);
    }
    gogoro_pre_kernel<<<blocks, 256, 0, st>>>(k, actions, steer_unit, 0);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_compute_observations(TkGogoro* h, float* obs_out, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(obs_out != nullptr, TK_ERR_INVALID, "null obs_out");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    gogoro_tile_kernel<2><<<h->tiles, kTile, 0, static_cast<cudaStream_t>(stream)>>>(k, obs_out);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_compute_reward(TkGogoro* h, const float* cur_obs, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(cur_obs != nullptr, TK_ERR_INVALID, "null cur_obs");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    gogoro_reward_kernel<<<(k.n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(k, cur_obs);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_compute_obs_rwd(TkGogoro* h, const float* sensor_unit, uint64_t seed, uint64_t step, int64_t env_id_offset,
                              void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    k.sensor_unit = sensor_unit;
    k.noise_mode = sensor_unit ? TK_NOISE_DENSE : TK_NOISE_PHILOX;
    k.seed = seed; k.step = step; k.env_off = env_id_offset;
    k.vec_ok = k.vec_ok && (!sensor_unit || aligned_to(sensor_unit, 16));
    gogoro_tile_kernel<1><<<h->tiles, kTile, 0, static_cast<cudaStream_t>(stream)>>>(k, nullptr);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_reset_idx(TkGogoro* h, const int64_t* env_ids, int32_t kcount, const float* unit, void* stream) {
    if (int rc = check_bound(h)) return rc;
    if (kcount == 0) return TK_OK;
    TK_REQUIRE(kcount > 0 && env_ids && unit, TK_ERR_INVALID, "bad reset_idx arguments");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    gogoro_reset_idx_kernel<<<(kcount + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        k, reinterpret_cast<const long long*>(env_ids), kcount, unit);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_post_physics(TkGogoro* h, const TkGogoroNoise* noise, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(noise != nullptr, TK_ERR_INVALID, "null noise descriptor");
    TK_CUDA(cudaSetDevice(h->params.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KParams k; fill_kparams(h, k, static_cast<cudaStream_t>(stream));
    const int mode = noise->reset.mode;
    TK_REQUIRE(mode >= TK_NOISE_NONE && mode <= TK_NOISE_PHILOX, TK_ERR_INVALID, "bad noise mode %d", mode);
    k.noise_mode = mode;
    k.reset_unit = noise->reset.unit_a; k.reset_rows = noise->reset.rows;
    k.seed = noise->reset.seed; k.step = noise->reset.step; k.env_off = noise->reset.env_id_offset;
    k.sensor_unit = noise->sensor_unit; k.speed_unit = noise->speed_unit; k.yaw_unit = noise->yaw_unit;
    if (mode == TK_NOISE_COMPACT || mode == TK_NOISE_DENSE)
        TK_REQUIRE(noise->reset.unit_a != nullptr, TK_ERR_INVALID, "reset noise rows missing");
    k.vec_ok = k.vec_ok && (!k.sensor_unit || aligned_to(k.sensor_unit, 16));
    if (mode == TK_NOISE_COMPACT) {
        const int wpb = 8;
        tile_count_kernel<<<(h->tiles + wpb - 1) / wpb, wpb * 32, 0, st>>>(k.reset, k.progress, k.n, h->tiles, k.speed_freq,
                                                                          h->counts, h->counts + h->tiles);
        exclusive_scan_kernel<<<1, 1024, 0, st>>>(h->counts, h->tiles, h->bases);
        exclusive_scan_kernel<<<1, 1024, 0, st>>>(h->counts + h->tiles, h->tiles, h->bases + h->tiles);
    }
    {
        // programmatic dependent launch: CTAs may be scheduled while the previous kernel in the stream drains; the
        // kernel executes griddepcontrol.wait before it touches global memory, so stream order is kept
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(h->tiles);
        cfg.blockDim = dim3(kTile);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        float* no_out = nullptr;
        // thin grids with in-kernel noise: the speculative-reset variant (see spec_reset); TK_GOGORO_SPEC=0/1 overrides
        bool spec = mode == TK_NOISE_PHILOX && h->prefetch != 0;
        if (h->knob_spec != kKnobUnset) spec = mode == TK_NOISE_PHILOX && h->knob_spec != 0;
        if (spec) cudaLaunchKernelEx(&cfg, gogoro_tile_kernel<0, true>, k, no_out);
        else cudaLaunchKernelEx(&cfg, gogoro_tile_kernel<0, false>, k, no_out);
    }
    TK_CUDA(cudaGetLastError());
    h->last_grid = kStatSlots;
    return TK_OK;
}

// developer probe, not part of the public header: per-tile globaltimer stamps [tiles, 8] (tools/tile_timeline.py --task gogoro)
int tk_debug_gogoro_timeline(TkGogoro* h, long long* device_buffer) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    h->dbg_ts = device_buffer;
    return TK_OK;
}

int tk_gogoro_advance_epoch(TkGogoro* h, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    h->epoch_used = true;
    advance_epoch_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(h->step_dev);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_stats(TkGogoro* h, double* stats_out, int32_t accumulate, void* stream) {
    TK_REQUIRE(h && stats_out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(h->last_grid > 0, TK_ERR_INVALID, "tk_gogoro_post_physics has not run yet");
    TK_CUDA(cudaSetDevice(h->params.device));
    tk::gogoro::stats_finalize_kernel<<<1, TK_STATS * 32, 0, static_cast<cudaStream_t>(stream)>>>(h->partials, h->last_grid,
                                                                                                  stats_out, accumulate);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_gogoro_philox_noise(TkGogoro* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* steer_unit,
                           float* sensor_unit, float* reset_unit, float* speed_unit, float* yaw_unit, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    gogoro_philox_kernel<<<(h->params.num_envs + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        h->params.num_envs, seed, step, env_id_offset, epoch_word(h, static_cast<cudaStream_t>(stream)), steer_unit, sensor_unit, reset_unit,
        speed_unit, yaw_unit);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

}  // extern "C"

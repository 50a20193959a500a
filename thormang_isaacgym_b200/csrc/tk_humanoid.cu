// Humanoid env-step hot path for B200 (sm_100a).
//
// Replaces, behind the C ABI of include/thormang_b200.h, the TorchScript functions and method
// bodies of /root/reference/isaacgymenvs/tasks/humanoid.py:
//   pre_physics_step :281-285, post_physics_step :287-296, reset_idx :253-279,
//   compute_humanoid_observations :379-417, compute_humanoid_reward :323-376, compute_reward :425-472.
//
// Data layout in HBM is Isaac Gym's (AoS): root_states [N,13], dof_state [N*D,2], dof_force [N,D],
// sensors [N,12], actions [N,D], obs_buf [N,110].  The fused post-physics kernel is a pure
// streaming kernel (924 B / env-step, ~0.3 flop/B): one CTA owns a tile of 32 consecutive
// envs, whose slice of EVERY array is one contiguous, 16-byte aligned byte range.
//   * the tile's inputs arrive as five cp.async.bulk (TMA) global->shared copies on one mbarrier;
//   * lane e pulls ITS env's inputs out of the staged tile into registers (conflict-free row strides), computes the
//     96 scaled copies and the torso math, and writes its obs row IN PLACE over the staging area;
//   * the finished [32,110] tile leaves as ONE cp.async.bulk (TMA) shared->global store;
//   * per-task constants (limits, 1/range, efforts) ride in the __grid_constant__ parameter block and are read as
//     constant-bank operands with compile-time indices: not staged in shared memory;
//   * per-env flags/potentials/reward are coalesced 4/8-byte accesses, one env per lane.
// Two kernels share that tile layout: humanoid_warp_kernel (tk_humanoid_tile.cuh; one warp per tile, 12 CTAs / SM:
// the bandwidth shape for grids of several waves) and humanoid_quad_kernel (tk_humanoid_quad.cuh; four warps per tile
// with the torso math split over the warps and column-parallel copies: the latency shape for single-wave grids).
// launch_tile() picks by tiles-per-SM.
// No tensor cores: nothing on this path is a contraction.
#include "tk_common.cuh"

#include <cstdlib>
#include <new>

namespace tk {
namespace humanoid {

constexpr int kRoot = 13;
constexpr int kVarHumanoid = TK_VARIANT_HUMANOID, kVarAnt = TK_VARIANT_ANT;

// Column layout of one obs row.  Humanoid (humanoid.py:409-414): 12 torso columns, dof_pos_scaled, dof_vel, dof_force,
// 12 sensor floats, actions, the two potentials = 110.  Ant (ant.py:399-404): no dof_force block, 24 sensor floats, no
// potential columns = 60.
template <int D_, int NSENSOR, bool HAS_FORCE, int VARIANT>
struct Spec {
    static constexpr int D = D_;
    static constexpr int kSensor = NSENSOR;
    static constexpr bool kHasForce = HAS_FORCE;
    static constexpr int kVariant = VARIANT;
    static constexpr int kPos = 12;
    static constexpr int kVel = kPos + D;
    static constexpr int kForce = kVel + D;
    static constexpr int kSens = kForce + (HAS_FORCE ? D : 0);
    static constexpr int kAct = kSens + NSENSOR;
    static constexpr int kPot = kAct + D;
    static constexpr int kObs = kPot + (VARIANT == kVarHumanoid ? 2 : 0);
};
using HumanoidSpec = Spec<21, 12, true, kVarHumanoid>;   // 110 observations (humanoid.py:67)
using AntSpec = Spec<8, 24, false, kVarAnt>;             // 60 observations (ant.py:67)
static_assert(HumanoidSpec::kObs == 110 && AntSpec::kObs == 60, "obs widths of the reference");
template <int D>
using Layout = Spec<D, 12, true, kVarHumanoid>;

struct KParams {
    float* root;
    float* dof_state;
    const float* dof_force;
    const float* sensors;
    float* actions;
    float* obs;
    float* obs2;                // optional second destination of the obs tile (tk_humanoid_set_obs_copy)
    float* rew;
    long long* reset;
    long long* progress;
    uint8_t* timeout;
    long long* randomize;
    float* pot;
    float* prev_pot;
    float* up_vec;
    float* heading_vec;
    const float* init_root;
    const float* init_dof_pos;
    float* ep_return;
    double* partials;           // [kStatSlots, TK_STATS]
    const int* tile_rank_base;  // COMPACT noise: resets in envs before this tile
    const float* unit_a;
    const float* unit_b;
    long long noise_rows;
    unsigned long long seed;
    unsigned long long step;
    long long env_off;
    int noise_mode;
    int n;
    int vec_ok;                 // all tile bases 16-byte aligned -> vector loads + bulk store
    float dt, dof_vel_scale, ang_vel_scale, contact_scale, power_scale;
    float heading_weight, up_weight, actions_cost_scale, energy_cost_scale, joints_cost_scale;
    float death_cost, termination_height, max_episode_length, clip_actions, max_motor_effort;
    float tgt[3], isr[4], b0[3], b1[3];
    float lo[TK_MAX_DOF], hi[TK_MAX_DOF], effort[TK_MAX_DOF];
    float rinv[TK_MAX_DOF];     // 1 / (hi - lo), rounded to fp32 on the host
    unsigned long long* step_dev;   // device-side noise epoch (upper 32 bits of the Philox step), see tk_common.cuh
    long long* dbg_ts;          // developer timeline probe (tools/tile_timeline.py); NULL in normal use
    int stagger_cycles;         // first-wave stagger: resident slot s of an SM delays its bulk loads by s * cycles (0 = off)
    int stagger_ctas;           // CTAs of the first wave (resident CTAs per SM x SMs)
    int prefetch_ctas;          // prefetch mode 2: the CTAs (lowest block ids = first resident slots of every SM) that prefetch
    int stagger_free;           // the first `stagger_free` resident slots of every SM start at once (fills the memory pipe), the rest are spaced
    int sm_count;
    int prefetch;               // L2 prefetch of the tile inputs ahead of griddepcontrol.wait
    int warm;                   // warm_ahead_of_wait (tk_humanoid_tile.cuh): 1 translations, 2 initial rows of the tile
    int dbg_reset;              // developer ablation (TK_HUMANOID_DBG_RESET=1): reset flags ignored
    int coop_store;             // developer knob (TK_HUMANOID_COOP_STORE=1): four-warp kernel stores its obs tile with STG.128
};

// ---------------------------------------------------------------------------------------------
// torso math of one env: compute_humanoid_observations :386-405 (+ compute_heading_and_up, compute_rot)
// ---------------------------------------------------------------------------------------------
struct Head {
    float c[12];      // obs columns 0..11
    float pot;        // potentials'
    V3 up, heading;   // up_vec, heading_vec
};

template <int VARIANT>
__device__ __forceinline__ Head torso_math(const float* r, const KParams& p) {
    Head h;
    V3 pos{r[0], r[1], r[2]};
    Q4 q{r[3], r[4], r[5], r[6]};
    V3 lin{r[7], r[8], r[9]};
    V3 ang{r[10], r[11], r[12]};

    V3 tt{p.tgt[0] - pos.x, p.tgt[1] - pos.y, 0.0f};              // :391-392
    float nrm = norm3_ref(tt);
    h.pot = (-nrm) / p.dt;                                        // :395

    float nclamp = fmaxf(nrm, 1e-9f);                             // normalize(): x / clamp(norm, min=eps)
    V3 td{tt.x / nclamp, tt.y / nclamp, tt.z / nclamp};
    Q4 tq = quat_mul_ref(q, Q4{p.isr[0], p.isr[1], p.isr[2], p.isr[3]});
    h.up = quat_rotate_ref<1>(tq, V3{p.b1[0], p.b1[1], p.b1[2]});
    h.heading = quat_rotate_ref<1>(tq, V3{p.b0[0], p.b0[1], p.b0[2]});
    float up_proj = h.up.z;
    float heading_proj = dot_ref(h.heading, td);

    V3 vel_loc = quat_rotate_ref<-1>(tq, lin);
    V3 angvel_loc = quat_rotate_ref<-1>(tq, ang);
    float roll, yaw;
    euler_roll_yaw_ref(tq, &roll, &yaw);
    float walk = atan2f(p.tgt[2] - pos.z, p.tgt[0] - pos.x);      // z and x, sic (compute_rot)
    float att = walk - yaw;

    h.c[0] = pos.z;
    h.c[1] = vel_loc.x; h.c[2] = vel_loc.y; h.c[3] = vel_loc.z;
    if (VARIANT == kVarAnt) {   // ant.py:399: raw angular velocity and raw Euler angles in [0, 2*pi]
        h.c[4] = angvel_loc.x; h.c[5] = angvel_loc.y; h.c[6] = angvel_loc.z;
        h.c[7] = yaw; h.c[8] = roll; h.c[9] = att;
    } else {                    // humanoid.py:403-405,410
        h.c[4] = angvel_loc.x * p.ang_vel_scale; h.c[5] = angvel_loc.y * p.ang_vel_scale; h.c[6] = angvel_loc.z * p.ang_vel_scale;
        h.c[7] = normalize_angle_fast(yaw);
        h.c[8] = normalize_angle_fast(roll);
        h.c[9] = normalize_angle_fast(att);
    }
    h.c[10] = up_proj;
    h.c[11] = heading_proj;
    return h;
}

// The same arithmetic cut into four independent pieces, one per warp of the four-warp tile kernel
// (tk_humanoid_quad.cuh); every piece re-derives the torso quaternion (9 multiplies) and evaluates its columns with
// exactly the expressions of torso_math, so the two kernels agree bit for bit.
struct HeadA { float height, up_proj, heading_proj, pot; V3 up, heading; };
template <int VARIANT>
__device__ __forceinline__ HeadA torso_part_a(const float* r, const KParams& p) {     // potentials, up / heading projections
    HeadA h;
    V3 pos{r[0], r[1], r[2]};
    Q4 q{r[3], r[4], r[5], r[6]};
    V3 tt{p.tgt[0] - pos.x, p.tgt[1] - pos.y, 0.0f};              // :391-392
    float nrm = norm3_ref(tt);
    h.pot = (-nrm) / p.dt;                                        // :395
    float nclamp = fmaxf(nrm, 1e-9f);
    V3 td{tt.x / nclamp, tt.y / nclamp, tt.z / nclamp};
    Q4 tq = quat_mul_ref(q, Q4{p.isr[0], p.isr[1], p.isr[2], p.isr[3]});
    h.up = quat_rotate_ref<1>(tq, V3{p.b1[0], p.b1[1], p.b1[2]});
    h.heading = quat_rotate_ref<1>(tq, V3{p.b0[0], p.b0[1], p.b0[2]});
    h.up_proj = h.up.z;
    h.heading_proj = dot_ref(h.heading, td);
    h.height = pos.z;
    return h;
}
template <int VARIANT>
__device__ __forceinline__ void torso_part_b(const float* r, const KParams& p, float* c1to6) {   // local velocities
    Q4 q{r[3], r[4], r[5], r[6]};
    Q4 tq = quat_mul_ref(q, Q4{p.isr[0], p.isr[1], p.isr[2], p.isr[3]});
    V3 vel_loc = quat_rotate_ref<-1>(tq, V3{r[7], r[8], r[9]});
    V3 angvel_loc = quat_rotate_ref<-1>(tq, V3{r[10], r[11], r[12]});
    c1to6[0] = vel_loc.x; c1to6[1] = vel_loc.y; c1to6[2] = vel_loc.z;
    if (VARIANT == kVarAnt) {
        c1to6[3] = angvel_loc.x; c1to6[4] = angvel_loc.y; c1to6[5] = angvel_loc.z;
    } else {
        c1to6[3] = angvel_loc.x * p.ang_vel_scale; c1to6[4] = angvel_loc.y * p.ang_vel_scale; c1to6[5] = angvel_loc.z * p.ang_vel_scale;
    }
}
template <int VARIANT>
__device__ __forceinline__ void torso_part_c(const float* r, const KParams& p, float* yaw_out, float* roll_out) {   // columns 7, 8
    Q4 q{r[3], r[4], r[5], r[6]};
    Q4 tq = quat_mul_ref(q, Q4{p.isr[0], p.isr[1], p.isr[2], p.isr[3]});
    float roll, yaw;
    euler_roll_yaw_ref(tq, &roll, &yaw);
    *yaw_out = (VARIANT == kVarAnt) ? yaw : normalize_angle_fast(yaw);
    *roll_out = (VARIANT == kVarAnt) ? roll : normalize_angle_fast(roll);
}
template <int VARIANT>
__device__ __forceinline__ float torso_part_d(const float* r, const KParams& p) {                // column 9: angle_to_target
    Q4 q{r[3], r[4], r[5], r[6]};
    Q4 tq = quat_mul_ref(q, Q4{p.isr[0], p.isr[1], p.isr[2], p.isr[3]});
    float roll, yaw;
    euler_roll_yaw_ref(tq, &roll, &yaw);                          // roll is dead here: the compiler drops its atan2
    float walk = atan2f(p.tgt[2] - r[2], p.tgt[0] - r[0]);        // z and x, sic (compute_rot)
    float att = walk - yaw;
    return (VARIANT == kVarAnt) ? att : normalize_angle_fast(att);
}

// unscale (humanoid.py:406): (2x - hi - lo) / (hi - lo).  The numerator is formed exactly as the reference does;
// the division is a multiplication by the host-rounded reciprocal (<= 1.5 ulp from the IEEE quotient, i.e. 2e-7
// relative: far inside the 1e-5 parity tolerance, and no threshold of the live reward reads these columns).
__device__ __forceinline__ float unscale_fast(float x, float lo, float hi, float rinv) { return (2.0f * x - hi - lo) * rinv; }
__device__ __forceinline__ float unscale_ref(float x, float lo, float hi) { return (2.0f * x - hi - lo) / (hi - lo); }

}  // namespace humanoid
}  // namespace tk

#include "tk_humanoid_tile.cuh"
#include "tk_humanoid_quad.cuh"

namespace tk {
namespace humanoid {

// ---------------------------------------------------------------------------------------------
// COMPACT noise order: exclusive prefix of the per-tile reset counts (two tiny kernels)
// ---------------------------------------------------------------------------------------------
__global__ void tile_reset_count_kernel(const long long* __restrict__ reset, int n, int tiles, int tile, int* __restrict__ counts) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= tiles) return;
    int c = 0;
    for (int i = lane; i < tile; i += 32) {
        int e = warp * tile + i;
        if (e < n && reset[e] != 0) ++c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) counts[warp] = c;
}

__global__ void tile_exclusive_scan_kernel(const int* __restrict__ counts, int tiles, int* __restrict__ base) {
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
            warp_tot[lane] = t;   // inclusive over warps
        }
        __syncthreads();
        const int before = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + (x - v);
        if (i < tiles) base[i] = before;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = before + v;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// pre_physics_step :281-285 (+ clamp vec_task.py:327): flat stream over N*D floats
// ---------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256) humanoid_pre_kernel(const __grid_constant__ KParams p, const float* __restrict__ in,
                                                            float* __restrict__ forces, int clip, int vec) {
    __shared__ float eff[D];
    if (threadIdx.x < D) eff[threadIdx.x] = p.effort[threadIdx.x];
    __syncthreads();
    const long long total = static_cast<long long>(p.n) * D;
    const float c = p.clip_actions, ps = p.power_scale;
    const long long gtid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    const long long gstride = static_cast<long long>(gridDim.x) * blockDim.x;
    if (vec) {
        const long long t4 = total >> 2;
        for (long long i = gtid; i < t4; i += gstride) {
            float4 a = ldg_stream4(in + 4 * i);
            float av[4] = {a.x, a.y, a.z, a.w}, fv[4];
            const int j0 = static_cast<int>((4 * i) % D);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float x = av[k];
                if (clip) x = fminf(fmaxf(x, -c), c);
                int j = j0 + k; if (j >= D) j -= D;
                av[k] = x;
                fv[k] = x * eff[j] * ps;                                           // :283
            }
            stg_stream4(p.actions + 4 * i, make_float4(av[0], av[1], av[2], av[3]));
            stg_stream4(forces + 4 * i, make_float4(fv[0], fv[1], fv[2], fv[3]));
        }
        for (long long e = (t4 << 2) + gtid; e < total; e += gstride) {
            float x = in[e];
            if (clip) x = fminf(fmaxf(x, -c), c);
            p.actions[e] = x;
            forces[e] = x * eff[e % D] * ps;
        }
    } else {
        for (long long e = gtid; e < total; e += gstride) {
            float x = in[e];
            if (clip) x = fminf(fmaxf(x, -c), c);
            p.actions[e] = x;
            forces[e] = x * eff[e % D] * ps;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// split entry points (method-level drop-in; not the hot configuration)
// ---------------------------------------------------------------------------------------------
template <int D>
__global__ void humanoid_reward_kernel(const __grid_constant__ KParams p) {
    using L = Layout<D>;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.n) return;
    const float* o = p.obs + static_cast<size_t>(n) * L::kObs;
    float up_reward = (o[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;               // :349-350
    float total = 2.0f + up_reward;                                                // :362,365
    const bool fallen = o[0] < p.termination_height;
    if (fallen) total = 1.0f * p.death_cost;                                       // :370
    long long reset = fallen ? 1 : p.reset[n];                                     // :373
    if (static_cast<float>(p.progress[n]) >= p.max_episode_length - 1.0f) reset = 1;   // :374
    p.rew[n] = total;
    p.reset[n] = reset;
}

// compute_ant_reward (ant.py:318-364) as a method-level entry point: one env per thread, sums in column order
__global__ void ant_reward_kernel(const __grid_constant__ KParams p) {
    using L = AntSpec;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.n) return;
    const float* o = p.obs + static_cast<size_t>(n) * L::kObs;
    const float* a = p.actions + static_cast<size_t>(n) * L::D;
    const float hp = o[11];
    const float heading_reward = (hp > 0.8f) ? (1.0f * p.heading_weight) : (p.heading_weight * hp / 0.8f);   // :337-338
    const float up_reward = (o[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;                                   // :341-342
    float actions_cost = 0.0f, electricity_cost = 0.0f;
    int at_limit = 0;
#pragma unroll
    for (int j = 0; j < L::D; ++j) {
        actions_cost += a[j] * a[j];                                               // :345
        electricity_cost += fabsf(a[j] * o[L::kVel + j]);                          // :346
        at_limit += (o[L::kPos + j] > 0.99f) ? 1 : 0;                              // :347
    }
    const float progress_reward = p.pot[n] - p.prev_pot[n];                        // :351
    float total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                  p.energy_cost_scale * electricity_cost - static_cast<float>(at_limit) * p.joints_cost_scale;   // :353-354
    const bool fallen = o[0] < p.termination_height;
    if (fallen) total = 1.0f * p.death_cost;                                       // :357
    long long reset = fallen ? 1 : p.reset[n];                                     // :360
    if (static_cast<float>(p.progress[n]) >= p.max_episode_length - 1.0f) reset = 1;   // :361
    p.rew[n] = total;
    p.reset[n] = reset;
}

// all terms of :344-363, one warp per env (lane = DOF), sequential-order sums via shuffles
template <int D>
__global__ void humanoid_terms_kernel(const __grid_constant__ KParams p, float* __restrict__ terms) {
    using L = Layout<D>;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= p.n) return;
    const float* o = p.obs + static_cast<size_t>(warp) * L::kObs;
    float a2 = 0.f, lim = 0.f, el = 0.f;
    if (lane < D) {
        float a = p.actions[static_cast<size_t>(warp) * D + lane];
        float ratio = p.effort[lane] / p.max_motor_effort;                         // :355
        float pa = fabsf(o[L::kPos + lane]);
        float scaled = p.joints_cost_scale * (pa - 0.98f) / 0.02f;                 // :356
        a2 = a * a;                                                                // :352
        lim = ((pa > 0.98f) ? 1.0f : 0.0f) * scaled * ratio;                       // :357
        el = fabsf(a * o[L::kVel + lane]) * ratio;                                 // :359
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        a2 += __shfl_xor_sync(0xffffffffu, a2, off);
        lim += __shfl_xor_sync(0xffffffffu, lim, off);
        el += __shfl_xor_sync(0xffffffffu, el, off);
    }
    if (lane == 0) {
        float hp = o[11];
        float* t = terms + static_cast<size_t>(warp) * 7;
        t[0] = (hp > 0.8f) ? (1.0f * p.heading_weight) : (p.heading_weight * hp / 0.8f);   // :345-346
        t[1] = (o[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;                       // :349-350
        t[2] = a2; t[3] = lim; t[4] = el;
        t[5] = 2.0f;                                                               // :362
        t[6] = p.pot[warp] - p.prev_pot[warp];                                     // :363
    }
}

// compute_reward :425-472 over obs[M,110]
template <int D>
__global__ void humanoid_wm_reward_kernel(const __grid_constant__ KParams p, const float* __restrict__ obs, long long m,
                                          float* __restrict__ out) {
    using L = Layout<D>;
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const float* o = obs + i * L::kObs;
    float hp = o[11];
    float heading_reward = (hp > 0.8f) ? (1.0f * p.heading_weight) : (p.heading_weight * hp / 0.8f);
    float up_reward = (o[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;
    float progress_reward = o[L::kPot] - o[L::kPot + 1];
    float total = progress_reward + 2.0f + up_reward + heading_reward;             // :467
    if (o[0] < p.termination_height) total = 1.0f * p.death_cost;                  // :470
    out[i] = total;
}

// reset_idx :253-279 for an explicit id list; one warp per listed env
template <int D>
__global__ void humanoid_reset_idx_kernel(const __grid_constant__ KParams p, const long long* __restrict__ ids, int k,
                                          const float* __restrict__ unit_pos, const float* __restrict__ unit_vel) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= k) return;
    const long long ne = ids[warp];
    if (ne < 0 || ne >= p.n) return;
    for (int j = lane; j < D; j += 32) {
        float positions = 0.4f * unit_pos[static_cast<size_t>(warp) * D + j] + (-0.2f);
        float velocities = 0.2f * unit_vel[static_cast<size_t>(warp) * D + j] + (-0.1f);
        float np_ = p.init_dof_pos[ne * D + j] + positions;
        np_ = fmaxf(fminf(np_, p.hi[j]), p.lo[j]);
        reinterpret_cast<float2*>(p.dof_state)[ne * D + j] = make_float2(np_, velocities);
    }
    if (lane < kRoot) p.root[ne * kRoot + lane] = p.init_root[ne * kRoot + lane];
    if (lane == 0) {
        V3 tt{p.tgt[0] - p.init_root[ne * kRoot + 0], p.tgt[1] - p.init_root[ne * kRoot + 1], 0.0f};   // :273-274
        float v = (-norm3_ref(tt)) / p.dt;                                          // :275
        p.prev_pot[ne] = v;
        p.pot[ne] = v;                                                             // :276
        p.progress[ne] = 0;                                                        // :278
        p.reset[ne] = 0;                                                           // :279
    }
}

template <int D>
__global__ void humanoid_philox_kernel(int n, unsigned long long seed, unsigned long long step, long long env_off,
                                       const unsigned long long* __restrict__ epoch,
                                       float* __restrict__ ua, float* __restrict__ ub) {
    step = effective_step(step, epoch);
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= static_cast<long long>(n) * D) return;
    const long long e = i / D;
    const int j = static_cast<int>(i - e * D);
    Philox4 r = philox_env(seed, step, env_off + e, static_cast<uint32_t>(j >> 1));
    ua[i] = u01((j & 1) ? r.z : r.x);
    ub[i] = u01((j & 1) ? r.w : r.y);
}

__global__ void stats_finalize_kernel(double* __restrict__ partials, int rows, double* __restrict__ out, int accumulate) {
    // one warp per statistic, fixed order
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

}  // namespace humanoid
}  // namespace tk

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace tk;
using namespace tk::humanoid;

struct TkHumanoid {
    TkHumanoidParams params;
    TkHumanoidBuffers buf;
    int variant;          // TK_VARIANT_HUMANOID | TK_VARIANT_ANT: selects the kernel instantiation
    bool bound;
    int tiles;
    double* partials;     // [kStatSlots, TK_STATS]
    int* tile_counts;     // [tiles]
    int* tile_base;       // [tiles]
    int last_grid;
    float* obs_copy;      // tk_humanoid_set_obs_copy
    int use_pdl;          // programmatic dependent launch (TK_HUMANOID_PDL=0 disables)
    int stagger_cycles;   // TK_HUMANOID_STAGGER (SM cycles per resident slot), see the kernel
    int stagger_slots;    // TK_HUMANOID_STAGGER_SLOTS
    int stagger_free;     // TK_HUMANOID_STAGGER_FREE
    int smem_pad;         // TK_HUMANOID_SMEM_PAD: extra dynamic shared memory per CTA (caps the resident CTAs per SM)
    int prefetch;         // TK_HUMANOID_PREFETCH: L2 prefetch ahead of griddepcontrol.wait (default: small grids only)
    int warm;             // TK_HUMANOID_WARM: see KParams::warm
    int sm_count;
    int quad;             // TK_HUMANOID_QUAD: four-warp tile kernel (tk_humanoid_quad.cuh) 1 | one-warp kernel 0; default by grid size
    unsigned long long* step_dev;
    long long* dbg_ts;
    int dbg_reset, coop_store, knob_prefetch_slots;   // developer knobs, read once in tk_humanoid_create
};

// fallback for the four-warp kernel's occupancy when the query fails (see tk_humanoid_create); measured B200, us / step,
// four-warp vs one-warp without prefetch: 6.7 / 8.3 at 16384 envs, 9.3 / 9.9 at 32768, 15.0 / 13.3 at 65536, 25.9 / 22.6 at 131072
static constexpr int kQuadMaxTilesPerSm = 7;
static constexpr int kD = HumanoidSpec::D;
static constexpr int kDAnt = AntSpec::D;

template <bool FUSED>
static void launch_tile(const TkHumanoid* h, const KParams& k, cudaStream_t st) {
    // Programmatic dependent launch: the CTAs of this launch may be scheduled while the previous kernel in the stream
    // drains; the kernel itself executes griddepcontrol.wait before it touches global memory, so stream order is kept.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(h->tiles);
    cfg.blockDim = dim3(h->quad ? kQuadThreads : kTile);
    cfg.dynamicSmemBytes = h->quad ? 0 : h->smem_pad;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = h->use_pdl ? 1 : 0;
    if (h->quad) {      // thin / medium grids: four warps per tile (tk_humanoid_quad.cuh)
        if (h->variant == kVarAnt) cudaLaunchKernelEx(&cfg, humanoid_quad_kernel<AntSpec, FUSED>, k);
        else cudaLaunchKernelEx(&cfg, humanoid_quad_kernel<HumanoidSpec, FUSED>, k);
        return;
    }
    if (h->variant == kVarAnt) cudaLaunchKernelEx(&cfg, humanoid_warp_kernel<AntSpec, FUSED>, k);
    else cudaLaunchKernelEx(&cfg, humanoid_warp_kernel<HumanoidSpec, FUSED>, k);
}

static void fill_kparams(const TkHumanoid* h, KParams& k) {
    const TkHumanoidParams& P = h->params;
    const TkHumanoidBuffers& B = h->buf;
    k.root = B.root_states; k.dof_state = B.dof_state; k.dof_force = B.dof_force; k.sensors = B.sensors;
    k.actions = B.actions; k.obs = B.obs_buf; k.obs2 = h->obs_copy; k.rew = B.rew_buf;
    k.reset = reinterpret_cast<long long*>(B.reset_buf); k.progress = reinterpret_cast<long long*>(B.progress_buf);
    k.timeout = B.timeout_buf; k.randomize = reinterpret_cast<long long*>(B.randomize_buf);
    k.pot = B.potentials; k.prev_pot = B.prev_potentials; k.up_vec = B.up_vec; k.heading_vec = B.heading_vec;
    k.init_root = B.initial_root_states; k.init_dof_pos = B.initial_dof_pos; k.ep_return = B.episode_return;
    k.partials = h->partials; k.tile_rank_base = h->tile_base; k.dbg_ts = h->dbg_ts; k.step_dev = h->step_dev;
    k.prefetch = h->prefetch;
    k.warm = h->warm;
    k.dbg_reset = h->dbg_reset;
    k.coop_store = h->coop_store;
    k.stagger_cycles = h->stagger_cycles; k.sm_count = h->sm_count; k.stagger_ctas = h->stagger_slots * h->sm_count; k.stagger_free = h->stagger_free;
    k.prefetch_ctas = k.stagger_ctas;
    if (h->knob_prefetch_slots != kKnobUnset) k.prefetch_ctas = h->knob_prefetch_slots * h->sm_count;
    k.unit_a = nullptr; k.unit_b = nullptr; k.noise_rows = 0; k.seed = 0; k.step = 0; k.env_off = 0;
    k.noise_mode = TK_NOISE_NONE;
    k.n = P.num_envs;
    k.vec_ok = aligned_to(B.root_states, 16) && aligned_to(B.dof_state, 16) && aligned_to(B.dof_force, 16) &&
               aligned_to(B.sensors, 16) && aligned_to(B.actions, 16) && aligned_to(B.obs_buf, 16);   // NULL dof_force (Ant) passes
    k.dt = P.dt; k.dof_vel_scale = P.dof_vel_scale; k.ang_vel_scale = P.angular_velocity_scale;
    k.contact_scale = P.contact_force_scale; k.power_scale = P.power_scale; k.heading_weight = P.heading_weight;
    k.up_weight = P.up_weight; k.actions_cost_scale = P.actions_cost_scale; k.energy_cost_scale = P.energy_cost_scale;
    k.joints_cost_scale = P.joints_at_limit_cost_scale; k.death_cost = P.death_cost;
    k.termination_height = P.termination_height; k.max_episode_length = P.max_episode_length;
    k.clip_actions = P.clip_actions; k.max_motor_effort = P.max_motor_effort;
    for (int i = 0; i < 3; ++i) { k.tgt[i] = P.targets[i]; k.b0[i] = P.basis_vec0[i]; k.b1[i] = P.basis_vec1[i]; }
    for (int i = 0; i < 4; ++i) k.isr[i] = P.inv_start_rot[i];
    for (int i = 0; i < TK_MAX_DOF; ++i) { k.lo[i] = P.dof_limits_lower[i]; k.hi[i] = P.dof_limits_upper[i]; k.effort[i] = P.motor_efforts[i];
        const float span = P.dof_limits_upper[i] - P.dof_limits_lower[i];
        k.rinv[i] = (span != 0.0f) ? 1.0f / span : 0.0f; }
}

static int check_bound(const TkHumanoid* h) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(h->bound, TK_ERR_UNBOUND, "tk_humanoid_bind was not called");
    return TK_OK;
}

extern "C" {

int tk_humanoid_create(const TkHumanoidParams* params, TkHumanoid** out) {
    TK_REQUIRE(params && out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(params->num_envs > 0, TK_ERR_INVALID, "num_envs must be positive (got %d)", params->num_envs);
    TK_REQUIRE(params->variant == kVarHumanoid || params->variant == kVarAnt, TK_ERR_INVALID, "unknown variant %d", params->variant);
    if (params->variant == kVarAnt)
        TK_REQUIRE(params->num_dof == kDAnt, TK_ERR_INVALID, "Ant kernels are built for num_dof = %d (got %d)", kDAnt, params->num_dof);
    else
        TK_REQUIRE(params->num_dof == kD, TK_ERR_INVALID, "Humanoid kernels are built for num_dof = %d (got %d)", kD, params->num_dof);
    TK_REQUIRE(params->dt > 0.f, TK_ERR_INVALID, "dt must be positive");
    TK_CUDA(cudaSetDevice(params->device));
    TkHumanoid* h = new (std::nothrow) TkHumanoid();
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "out of host memory");
    h->params = *params;
    h->variant = params->variant;
    h->bound = false;
    h->dbg_ts = nullptr;
    h->use_pdl = 1;
    h->obs_copy = nullptr;
    if (const char* env = knob("TK_HUMANOID_PDL")) h->use_pdl = atoi(env) != 0;
    h->dbg_reset = ablation_knob("TK_HUMANOID_DBG_RESET");
    { const int v = knob_int("TK_HUMANOID_COOP_STORE"); h->coop_store = v == kKnobUnset ? 0 : v; }
    h->knob_prefetch_slots = knob_int("TK_HUMANOID_PREFETCH_SLOTS");
    // First-wave stagger (see the kernel): 500 SM cycles per resident slot; anything in 400..700 measured the same
    // (B200: 16.5 -> 14.0 us at 65536 envs, 25.4 -> 23.0 us at 131072).  The slot count is the kernel's real occupancy.
    h->stagger_cycles = 500; h->stagger_slots = 12; h->smem_pad = 0; h->sm_count = 148;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, params->device);
    {
        int occ = 0;
        cudaError_t oe = (params->variant == kVarAnt)
            ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_warp_kernel<AntSpec, true>, kTile, 0)
            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_warp_kernel<HumanoidSpec, true>, kTile, 0);
        if (oe == cudaSuccess && occ > 0) h->stagger_slots = occ;
        else cudaGetLastError();
    }
    h->prefetch = -1;   // auto, resolved below
    if (const char* env = knob("TK_HUMANOID_PREFETCH")) h->prefetch = atoi(env);
    if (const char* env = knob("TK_HUMANOID_STAGGER")) h->stagger_cycles = atoi(env);
    if (const char* env = knob("TK_HUMANOID_STAGGER_SLOTS")) h->stagger_slots = atoi(env);
    h->stagger_free = 0;
    if (const char* env = knob("TK_HUMANOID_STAGGER_FREE")) h->stagger_free = atoi(env);
    if (const char* env = knob("TK_HUMANOID_SMEM_PAD")) {
        h->smem_pad = atoi(env);
        cudaFuncSetAttribute(humanoid_warp_kernel<HumanoidSpec, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_pad);
        cudaFuncSetAttribute(humanoid_warp_kernel<AntSpec, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_pad);
    }
    h->tiles = (params->num_envs + kTile - 1) / kTile;
    // L2 prefetch ahead of griddepcontrol.wait pays only while the whole grid is resident at once and thin (measured:
    // 10.1 -> 8.3 us at 16384 envs, but +1.8 us at 65536 where it re-synchronises the staggered first wave)
    // four-warp tiles while the WHOLE grid is resident at once (one wave of the four-warp kernel's own occupancy: 8 CTAs / SM
    // for the Humanoid = 37888 envs on 148 SMs), one-warp tiles (most bytes in flight per SM) beyond.  Measured, us / step,
    // four-warp + prefetch vs the previous choice: 8.2 vs 9.5 at 32768 envs, 8.7 vs 10.0 at 36864; one CTA past the wave the
    // four-warp kernel collapses (14.3 vs 10.3 at 40960).
    int quad_occ = kQuadMaxTilesPerSm;
    {
        int occ = 0;
        cudaError_t oe = (params->variant == kVarAnt)
            ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_quad_kernel<AntSpec, true>, kQuadThreads, 0)
            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_quad_kernel<HumanoidSpec, true>, kQuadThreads, 0);
        if (oe == cudaSuccess && occ > 0) quad_occ = occ < 8 ? occ : 8;      // measured up to 8 per SM (Humanoid); Ant's smaller tiles would allow more
        else cudaGetLastError();
    }
    h->quad = (h->tiles <= quad_occ * h->sm_count) ? 1 : 0;
    // Ant (60 obs, 48 copy columns, every reward term live in warp 0) does not profit from the four-warp split: measured, us / step,
    // four-warp + prefetch vs one-warp + prefetch: 9.05 / 7.47 at 16384 envs, 9.13 / 7.97 at 32768.  Its one-warp kernel takes the
    // bulk-engine prefetch while the grid fits one wave (10.9 -> 10.4 us at 65536; 14.1 -> 14.9 at 98304: off beyond)
    if (params->variant == kVarAnt) {
        h->quad = 0;
        if (h->prefetch < 0) h->prefetch = (h->tiles <= 14 * h->sm_count) ? 1 : 0;
    }
    if (const char* env = knob("TK_HUMANOID_QUAD")) h->quad = atoi(env) != 0;
    // the prefetch goes with the single wave: there the CTAs of launch k+1 become resident while launch k runs (or drains)
    // and have slack ahead of the wait; measured with the four-warp kernel 7.5 -> 6.65 us at 16384 envs, 9.0 -> 7.9 at 24576,
    // 9.55 -> 8.2 at 32768 (the one-warp kernel on such grids: 8.6 -> 8.0, 9.5 -> 8.6, 10.7 -> 10.5)
    // Between the four-warp kernel's wave and the ONE-warp kernel's own (12 CTAs / SM: 37888 .. 56832 Humanoid envs) the one-warp
    // kernel takes the bulk-engine prefetch: 10.3 -> 8.7 us at 40960 envs, 11.2 -> 10.3 at 49152, 12.1 -> 11.3 at 56832 (71 % of
    // peak); one tile past the wave it costs (65536: 13.3 -> 15.5+ us), so the switch is the wave itself.
    if (h->prefetch < 0) {
        if (h->quad || h->tiles <= 4 * h->sm_count) h->prefetch = 2;
        else h->prefetch = (params->variant != kVarAnt && h->tiles <= h->stagger_slots * h->sm_count) ? 1 : 0;
    }
    h->warm = 0;          // measured: no gain at 16384 envs (6.55 vs 6.58 / 7.09 us with 1 / 3), -3 % at 32768, +8 % at 65536
    if (const char* env = knob("TK_HUMANOID_WARM")) h->warm = atoi(env);
    // the four-warp kernel only runs single-wave grids, where spacing the CTAs merely delays them (6.5 us with 0 cycles,
    // 6.8 with 500, 7.4 with 1000 at 16384 envs)
    if (h->quad && !knob("TK_HUMANOID_STAGGER")) h->stagger_cycles = 0;
    if (h->quad && !knob("TK_HUMANOID_STAGGER_SLOTS")) {      // the stagger counts the resident CTAs of the kernel that runs
        int occ = 0;
        cudaError_t oe = (params->variant == kVarAnt)
            ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_quad_kernel<AntSpec, true>, kQuadThreads, 0)
            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, humanoid_quad_kernel<HumanoidSpec, true>, kQuadThreads, 0);
        if (oe == cudaSuccess && occ > 0) h->stagger_slots = occ;
        else cudaGetLastError();
    }
    h->last_grid = 0;
    h->partials = nullptr; h->tile_counts = nullptr; h->tile_base = nullptr; h->step_dev = nullptr;
    cudaError_t e = cudaMalloc(&h->partials, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMalloc(&h->tile_counts, sizeof(int) * h->tiles);
    if (e == cudaSuccess) e = cudaMalloc(&h->tile_base, sizeof(int) * h->tiles);
    if (e == cudaSuccess) e = cudaMemset(h->partials, 0, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMemset(h->tile_base, 0, sizeof(int) * h->tiles);
    if (e == cudaSuccess) e = cudaMalloc(&h->step_dev, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->step_dev, 0, sizeof(unsigned long long));
    if (e != cudaSuccess) {
        cudaFree(h->partials); cudaFree(h->tile_counts); cudaFree(h->tile_base); cudaFree(h->step_dev);
        delete h;
        return cuda_fail(e, "tk_humanoid_create");
    }
    *out = h;
    return TK_OK;
}

int tk_humanoid_destroy(TkHumanoid* h) {
    if (!h) return TK_OK;
    cudaSetDevice(h->params.device);
    cudaFree(h->partials); cudaFree(h->tile_counts); cudaFree(h->tile_base); cudaFree(h->step_dev);
    delete h;
    return TK_OK;
}

int tk_humanoid_bind(TkHumanoid* h, const TkHumanoidBuffers* b) {
    TK_REQUIRE(h && b, TK_ERR_INVALID, "null argument");
    const void* req[] = {b->root_states, b->dof_state, b->dof_force, b->sensors, b->actions, b->obs_buf, b->rew_buf,
                         b->reset_buf, b->progress_buf, b->potentials, b->prev_potentials, b->initial_root_states,
                         b->initial_dof_pos};
    const char* names[] = {"root_states", "dof_state", "dof_force", "sensors", "actions", "obs_buf", "rew_buf",
                           "reset_buf", "progress_buf", "potentials", "prev_potentials", "initial_root_states",
                           "initial_dof_pos"};
    for (unsigned i = 0; i < sizeof(req) / sizeof(req[0]); ++i) {
        if (i == 2 && h->variant == kVarAnt) continue;   // Ant has no dof_force block (ant.py:399-404)
        TK_REQUIRE(req[i] != nullptr, TK_ERR_UNBOUND, "buffer %s is NULL", names[i]);
        TK_REQUIRE(aligned_to(req[i], 4), TK_ERR_ALIGN, "buffer %s is not 4-byte aligned", names[i]);
    }
    TK_REQUIRE(aligned_to(b->reset_buf, 8) && aligned_to(b->progress_buf, 8), TK_ERR_ALIGN, "int64 buffers must be 8-byte aligned");
    TK_REQUIRE(aligned_to(b->dof_state, 8), TK_ERR_ALIGN, "dof_state must be 8-byte aligned ((pos,vel) pairs)");
    TK_REQUIRE(!b->randomize_buf || aligned_to(b->randomize_buf, 8), TK_ERR_ALIGN, "randomize_buf must be 8-byte aligned");
    h->buf = *b;
    h->bound = true;
    return TK_OK;
}

int tk_humanoid_pre_physics(TkHumanoid* h, const float* actions_in, float* forces_out, int32_t clip, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(actions_in && forces_out, TK_ERR_INVALID, "null actions_in / forces_out");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    const int vec = aligned_to(actions_in, 16) && aligned_to(forces_out, 16) && aligned_to(h->buf.actions, 16);
    const int dof = h->params.num_dof;
    const long long total4 = (static_cast<long long>(k.n) * dof + 3) / 4;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->params.device);
    long long blocks = (total4 + 255) / 256;
    if (blocks > static_cast<long long>(sms) * 8) blocks = static_cast<long long>(sms) * 8;
    if (blocks < 1) blocks = 1;
    if (h->variant == kVarAnt)   // ant.py:273-277: forces = actions * joint_gears * power_scale
        humanoid_pre_kernel<kDAnt><<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(k, actions_in, forces_out, clip, vec);
    else
        humanoid_pre_kernel<kD><<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(k, actions_in, forces_out, clip, vec);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_compute_observations(TkHumanoid* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    launch_tile<false>(h, k, static_cast<cudaStream_t>(stream));
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_compute_reward(TkHumanoid* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    if (h->variant == kVarAnt) ant_reward_kernel<<<(k.n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(k);
    else humanoid_reward_kernel<kD><<<(k.n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(k);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_reset_idx(TkHumanoid* h, const int64_t* env_ids, int32_t kcount, const float* unit_pos, const float* unit_vel,
                          void* stream) {
    if (int rc = check_bound(h)) return rc;
    if (kcount == 0) return TK_OK;
    TK_REQUIRE(kcount > 0 && env_ids && unit_pos && unit_vel, TK_ERR_INVALID, "bad reset_idx arguments");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    const int wpb = 8;
    if (h->variant == kVarAnt)
        humanoid_reset_idx_kernel<kDAnt><<<(kcount + wpb - 1) / wpb, wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            k, reinterpret_cast<const long long*>(env_ids), kcount, unit_pos, unit_vel);
    else
        humanoid_reset_idx_kernel<kD><<<(kcount + wpb - 1) / wpb, wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            k, reinterpret_cast<const long long*>(env_ids), kcount, unit_pos, unit_vel);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_post_physics(TkHumanoid* h, const TkResetNoise* noise, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KParams k; fill_kparams(h, k);
    if (noise) {
        k.noise_mode = noise->mode;
        k.unit_a = noise->unit_a; k.unit_b = noise->unit_b; k.noise_rows = noise->rows;
        k.seed = noise->seed; k.step = noise->step; k.env_off = noise->env_id_offset;
        if (noise->mode == TK_NOISE_COMPACT || noise->mode == TK_NOISE_DENSE)
            TK_REQUIRE(noise->unit_a && noise->unit_b && noise->rows >= 0, TK_ERR_INVALID, "noise arrays missing");
        TK_REQUIRE(noise->mode >= TK_NOISE_NONE && noise->mode <= TK_NOISE_PHILOX, TK_ERR_INVALID, "bad noise mode %d", noise->mode);
    }
    if (k.noise_mode == TK_NOISE_COMPACT) {
        const int wpb = 8;
        tile_reset_count_kernel<<<(h->tiles + wpb - 1) / wpb, wpb * 32, 0, st>>>(k.reset, k.n, h->tiles, kTile, h->tile_counts);
        tile_exclusive_scan_kernel<<<1, 1024, 0, st>>>(h->tile_counts, h->tiles, h->tile_base);
    }
    launch_tile<true>(h, k, st);
    TK_CUDA(cudaGetLastError());
    h->last_grid = kStatSlots;
    return TK_OK;
}

int tk_humanoid_reward_terms(TkHumanoid* h, float* terms, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(terms != nullptr, TK_ERR_INVALID, "null terms");
    TK_REQUIRE(h->variant == kVarHumanoid, TK_ERR_INVALID, "reward_terms is a Humanoid entry point (Ant's terms are all live in rew_buf)");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    const int wpb = 8;
    humanoid_terms_kernel<kD><<<(k.n + wpb - 1) / wpb, wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(k, terms);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_wm_reward(TkHumanoid* h, const float* obs, int64_t m, float* reward_out, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(obs && reward_out && m >= 0, TK_ERR_INVALID, "bad wm_reward arguments");
    TK_REQUIRE(h->variant == kVarHumanoid, TK_ERR_INVALID, "wm_reward is a Humanoid entry point (humanoid.py:425-472)");
    if (m == 0) return TK_OK;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    humanoid_wm_reward_kernel<kD><<<static_cast<unsigned>((m + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        k, obs, m, reward_out);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_philox_noise(TkHumanoid* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* unit_pos,
                             float* unit_vel, void* stream) {
    TK_REQUIRE(h && unit_pos && unit_vel, TK_ERR_INVALID, "null argument");
    TK_CUDA(cudaSetDevice(h->params.device));
    const long long total = static_cast<long long>(h->params.num_envs) * h->params.num_dof;
    if (h->variant == kVarAnt)
        humanoid_philox_kernel<kDAnt><<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
            h->params.num_envs, seed, step, env_id_offset, h->step_dev, unit_pos, unit_vel);
    else
        humanoid_philox_kernel<kD><<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
            h->params.num_envs, seed, step, env_id_offset, h->step_dev, unit_pos, unit_vel);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

// developer probe, not part of the public header: per-tile globaltimer stamps [tiles, 8] (tools/tile_timeline.py)
int tk_debug_humanoid_timeline(TkHumanoid* h, long long* device_buffer) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    h->dbg_ts = device_buffer;
    return TK_OK;
}

int tk_humanoid_set_obs_copy(TkHumanoid* h, float* obs_copy) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(aligned_to(obs_copy, 16), TK_ERR_ALIGN, "obs_copy must be 16-byte aligned");
    h->obs_copy = obs_copy;
    return TK_OK;
}

int tk_humanoid_advance_epoch(TkHumanoid* h, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    advance_epoch_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(h->step_dev);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_humanoid_stats(TkHumanoid* h, double* stats_out, int32_t accumulate, void* stream) {
    TK_REQUIRE(h && stats_out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(h->last_grid > 0, TK_ERR_INVALID, "tk_humanoid_post_physics has not run yet");
    TK_CUDA(cudaSetDevice(h->params.device));
    stats_finalize_kernel<<<1, TK_STATS * 32, 0, static_cast<cudaStream_t>(stream)>>>(h->partials, h->last_grid, stats_out, accumulate);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

}  // extern "C"

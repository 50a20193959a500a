// Anymal env-step hot path for B200 (sm_100a) -- SURVEY 8(f) row N2 (the PD-target write + contact-force termination).
//
// Replaces, behind the C ABI of include/thormang_b200.h, the TorchScript functions and method bodies of
// /root/reference/isaacgymenvs/tasks/anymal.py (this fork's copy, 61 observations):
//   pre_physics_step :229-232, post_physics_step :234-242, compute_reward :244-259, compute_observations :262-283,
//   reset_idx :285-311, compute_anymal_reward :323-372, compute_anymal_observations_V2 :399-431.
//
// 645 B / env-step: reads root 52 + dof_state 96 + torques 48 + contact forces 156 + commands 12 + progress 8 + reset 8,
// writes obs 244 + reward 4 + reset 8 + progress 8 + timeout 1.  Same shape as the other fused kernels: one warp = one
// CTA = 32 consecutive envs, the five AoS inputs arrive as cp.async.bulk (TMA) tile copies on one mbarrier, lane e owns
// env e (row strides 13/24/12/39/3/61 words: odd or conflict-free), the [32,61] obs tile leaves as one bulk store.
#include "tk_common.cuh"

#include <cstdlib>
#include <new>

namespace tk {
namespace anymal {

constexpr int kRoot = 13;
constexpr int kDof = 12;
constexpr int kBodies = 13;
constexpr int kContact = kBodies * 3;
constexpr int kObs = 61;
constexpr int kTile = 32;
constexpr int kMaxKnee = 8;

struct KParams {
    float* root;
    float* dof_state;
    const float* torques;
    const float* contact;
    float* commands;
    float* actions;
    float* obs;
    float* rew;
    long long* reset;
    long long* progress;
    uint8_t* timeout;
    float* ep_return;
    double* partials;
    const float* unit_pos;
    const float* unit_vel;
    const float* unit_cmd;
    unsigned long long seed, step;
    unsigned long long* step_dev;
    long long env_off;
    int noise_mode;
    int n, max_episode_length, base_index, num_knee;
    int knee[kMaxKnee];
    int vec_ok;
    int prefetch;      // L2 prefetch of the tile ahead of griddepcontrol.wait (TK_ANYMAL_PREFETCH overrides)
    int stagger_cycles, stagger_ctas, sm_count;   // first-wave stagger (same finding as in the Humanoid kernel)
    int dbg_reset;     // developer ablation (TK_ANYMAL_DBG_RESET=1): reset flags ignored
    float dof_pos_scale, dof_vel_scale, action_scale, clip_actions;
    float k_lin, k_ang, k_torque;           // (float)(scale * dt), anymal.py:96-97
    float cmd_x[2], cmd_y[2], cmd_yaw[2];   // lo, (hi - lo) formed in double
    float default_dof_pos[kDof];            // row of the replicated [N,12] tensor (:131-136)
    float init_root[kRoot];                 // row of initial_root_states (:140-141)
    float gravity[3];                       // row of gravity_vec (:142)
};

enum { kStreamPos = 32, kStreamVel = 35, kStreamCmd = 38 };

__device__ __forceinline__ float rand_float(float lo, float span, float u) { return span * u + lo; }   // torch_rand_float

struct Smem {
    alignas(128) float obs[kTile * kObs];
    alignas(16) float root[kTile * kRoot];
    alignas(16) float dof[kTile * kDof * 2];
    alignas(16) float torques[kTile * kDof];
    alignas(16) float contact[kTile * kContact];
    alignas(16) float commands[kTile * 3];
    alignas(8) uint64_t bar;
};

// MODE 0: fused post_physics_step; 1: compute_observations; 2: compute_reward
template <int MODE>
__global__ void __launch_bounds__(kTile) anymal_tile_kernel(const __grid_constant__ KParams p) {
    __shared__ Smem s;
    const int lane = threadIdx.x;
    const int env0 = blockIdx.x * kTile;
    const int nt = min(kTile, p.n - env0);
    const long long n = env0 + lane;
    const bool valid = lane < nt;
    const bool bulk = p.vec_ok && nt == kTile;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (bulk && lane == 0) mbar_init(&s.bar, 1);
    // L2 prefetch of the tile ahead of the dependency wait (grids of about one wave, host decides: p.prefetch): a hint, nothing is
    // consumed, stream order intact (see tk_humanoid_tile.cuh)
    if (bulk && p.prefetch) {
        const size_t e0p = static_cast<size_t>(env0);
        if (lane == 0) bulk_prefetch_l2(p.contact + e0p * kContact, kTile * kContact * 4);
        else if (lane == 1 && MODE != 2) bulk_prefetch_l2(p.dof_state + e0p * kDof * 2, kTile * kDof * 2 * 4);
        else if (lane == 2) bulk_prefetch_l2(p.root + e0p * kRoot, kTile * kRoot * 4);
        else if (lane == 3) bulk_prefetch_l2(p.torques + e0p * kDof, kTile * kDof * 4);
        else if (lane == 4) bulk_prefetch_l2(p.commands + e0p * 3, kTile * 3 * 4);
        else if (lane == 5 && MODE != 1) prefetch_l2(p.progress + env0);
        else if (lane == 6 && MODE != 1) prefetch_l2(p.progress + env0 + 16);
        else if (lane == 7 && MODE == 0) prefetch_l2(p.reset + env0);
        else if (lane == 8 && MODE == 0) prefetch_l2(p.reset + env0 + 16);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // First-wave stagger: the CTAs of the first wave start together, so every SM would run load-all / math-all / store-all
    // in lockstep and HBM would idle during the math; resident slot s of the first wave waits s * stagger_cycles.
    if (MODE == 0 && p.stagger_cycles > 0 && blockIdx.x < static_cast<unsigned>(p.stagger_ctas)) {
        const long long until = clock64() + static_cast<long long>(blockIdx.x / p.sm_count) * p.stagger_cycles;
        while (clock64() < until) __nanosleep(32);
    }
    const size_t e0 = static_cast<size_t>(env0);
    if (bulk) {
        if (lane == 0) {
            fence_proxy_async_smem();
            mbar_expect_tx(&s.bar, (kRoot + (MODE != 2 ? 2 * kDof : 0) + kDof + kContact + 3) * kTile * 4);
        }
        __syncwarp();
        if (lane == 0) bulk_load(s.contact, p.contact + e0 * kContact, kTile * kContact * 4, &s.bar);
        else if (lane == 1 && MODE != 2) bulk_load(s.dof, p.dof_state + e0 * kDof * 2, kTile * kDof * 2 * 4, &s.bar);
        else if (lane == 2) bulk_load(s.root, p.root + e0 * kRoot, kTile * kRoot * 4, &s.bar);
        else if (lane == 3) bulk_load(s.torques, p.torques + e0 * kDof, kTile * kDof * 4, &s.bar);
        else if (lane == 4) bulk_load(s.commands, p.commands + e0 * 3, kTile * 3 * 4, &s.bar);
    } else {
        for (int i = lane; i < nt * kRoot; i += kTile) s.root[i] = p.root[e0 * kRoot + i];
        if (MODE != 2) for (int i = lane; i < nt * kDof * 2; i += kTile) s.dof[i] = p.dof_state[e0 * kDof * 2 + i];
        for (int i = lane; i < nt * kDof; i += kTile) s.torques[i] = p.torques[e0 * kDof + i];
        for (int i = lane; i < nt * kContact; i += kTile) s.contact[i] = p.contact[e0 * kContact + i];
        for (int i = lane; i < nt * 3; i += kTile) s.commands[i] = p.commands[e0 * 3 + i];
    }
    long long prog = 0;
    bool do_reset = false;
    float ret_old = 0.f;
    if (valid && MODE != 1) {
        prog = p.progress[n];
        if (MODE == 0) {
            prog += 1;                                                              // :235
            do_reset = p.reset[n] != 0 && !p.dbg_reset;                             // :237
            if (do_reset) prog = 0;                                                 // :309
            if (p.ep_return) ret_old = p.ep_return[n];
        }
    }
    __syncwarp();
    if (bulk) mbar_wait(&s.bar, 0);

    float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_e1 = 0.f, st_e2 = 0.f;
    bool out_reset = false, out_timeout = false;
    if (valid) {
        float r[kRoot], cmd[3];
#pragma unroll
        for (int c = 0; c < kRoot; ++c) r[c] = s.root[lane * kRoot + c];
#pragma unroll
        for (int c = 0; c < 3; ++c) cmd[c] = s.commands[lane * 3 + c];
        float2* dofp = reinterpret_cast<float2*>(s.dof) + lane * kDof;
        if (MODE == 0 && do_reset) {                                                // reset_idx :285-311, this env's lane
            const bool philox = p.noise_mode == TK_NOISE_PHILOX, dense = p.noise_mode == TK_NOISE_DENSE;
            const unsigned long long step_now = philox ? effective_step(p.step, p.step_dev) : 0ull;
            float up[kDof], uv[kDof], uc[3];
            if (philox) {
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    Philox4 a = philox_env(p.seed, step_now, p.env_off + n, kStreamPos + b);
                    Philox4 c = philox_env(p.seed, step_now, p.env_off + n, kStreamVel + b);
                    up[4 * b] = u01(a.x); up[4 * b + 1] = u01(a.y); up[4 * b + 2] = u01(a.z); up[4 * b + 3] = u01(a.w);
                    uv[4 * b] = u01(c.x); uv[4 * b + 1] = u01(c.y); uv[4 * b + 2] = u01(c.z); uv[4 * b + 3] = u01(c.w);
                }
                Philox4 d = philox_env(p.seed, step_now, p.env_off + n, kStreamCmd);
                uc[0] = u01(d.x); uc[1] = u01(d.y); uc[2] = u01(d.z);
            } else {
#pragma unroll
                for (int j = 0; j < kDof; ++j) {
                    up[j] = (dense && p.unit_pos) ? p.unit_pos[n * kDof + j] : 0.5f;
                    uv[j] = (dense && p.unit_vel) ? p.unit_vel[n * kDof + j] : 0.5f;
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) uc[c] = (dense && p.unit_cmd) ? p.unit_cmd[n * 3 + c] : 0.5f;
            }
            float2* gd = reinterpret_cast<float2*>(p.dof_state) + n * kDof;
#pragma unroll
            for (int j = 0; j < kDof; ++j) {
                const float pos = p.default_dof_pos[j] * rand_float(0.5f, 1.0f, up[j]);          // :288,291
                const float vel = rand_float(-0.1f, static_cast<float>(0.1 - (-0.1)), uv[j]);    // :289,292
                dofp[j] = make_float2(pos, vel);
                gd[j] = make_float2(pos, vel);
            }
#pragma unroll
            for (int c = 0; c < kRoot; ++c) { r[c] = p.init_root[c]; p.root[n * kRoot + c] = r[c]; }   // :296-298
            cmd[0] = rand_float(p.cmd_x[0], p.cmd_x[1], uc[0]);                     // :305
            cmd[1] = rand_float(p.cmd_y[0], p.cmd_y[1], uc[1]);                     // :306
            cmd[2] = rand_float(p.cmd_yaw[0], p.cmd_yaw[1], uc[2]);                 // :307
#pragma unroll
            for (int c = 0; c < 3; ++c) p.commands[n * 3 + c] = cmd[c];
        }
        const Q4 q{r[3], r[4], r[5], r[6]};
        const V3 lin = quat_rotate_ref<-1>(q, V3{r[7], r[8], r[9]});                // :342 / :411
        const V3 ang = quat_rotate_ref<-1>(q, V3{r[10], r[11], r[12]});             // :343 / :412
        const float* tq = s.torques + lane * kDof;
        const float* cf = s.contact + lane * kContact;
        if (MODE != 2) {                                                            // compute_anymal_observations_V2 :399-431
            float* o = s.obs + lane * kObs;
            o[0] = lin.x; o[1] = lin.y; o[2] = lin.z; o[3] = ang.x; o[4] = ang.y; o[5] = ang.z;
#pragma unroll
            for (int j = 0; j < kDof; ++j) o[6 + j] = tq[j];
            const V3 g = quat_rotate_ref<1>(q, V3{p.gravity[0], p.gravity[1], p.gravity[2]});    // :413
            o[18] = g.x; o[19] = g.y; o[20] = g.z;
#pragma unroll
            for (int b = 0; b < kBodies; ++b) o[21 + b] = norm3_ref(V3{cf[3 * b], cf[3 * b + 1], cf[3 * b + 2]});   // :417
#pragma unroll
            for (int j = 0; j < kDof; ++j) {
                const float2 pv = dofp[j];
                o[34 + j] = (pv.x - p.default_dof_pos[j]) * p.dof_pos_scale;       // :414
                o[46 + j] = pv.y * p.dof_vel_scale;
            }
            o[58] = cmd[0]; o[59] = cmd[1]; o[60] = cmd[2];
        }
        if (MODE != 1) {                                                            // compute_anymal_reward :323-372
            const float dx = cmd[0] - lin.x, dy = cmd[1] - lin.y, dz = cmd[2] - ang.z;
            const float e1 = dx * dx + dy * dy;                                     // :346
            const float e2 = dz * dz;                                               // :348
            const float r1 = expf(-e1 / 0.25f) * p.k_lin;                           // :349
            const float r2 = expf(-e2 / 0.25f) * p.k_ang;                           // :350
            float tsum = 0.0f;
#pragma unroll
            for (int j = 0; j < kDof; ++j) tsum += tq[j] * tq[j];
            const float r3 = tsum * p.k_torque;                                     // :352
            const float total = fmaxf(r1 + r2 + r3, 0.0f);                          // :356-357
            // the fused step has just written the 13 contact norms into this lane's obs row (columns 21..33, same expression):
            // the termination test reads them back instead of evaluating 1 + num_knee more IEEE square roots
            const float* on = s.obs + lane * kObs + 21;
            const int bi = p.base_index * 3;
            bool rs = (MODE == 0 ? on[p.base_index] : norm3_ref(V3{cf[bi], cf[bi + 1], cf[bi + 2]})) > 1.0f;      // :359
            for (int k = 0; k < p.num_knee; ++k) {
                const int ki = p.knee[k] * 3;
                rs = rs || ((MODE == 0 ? on[p.knee[k]] : norm3_ref(V3{cf[ki], cf[ki + 1], cf[ki + 2]})) > 1.0f);   // :360
            }
            const bool time_out = prog >= p.max_episode_length - 1;                 // :361
            out_reset = rs || time_out;                                             // :362
            p.rew[n] = total;
            p.reset[n] = out_reset ? 1 : 0;
            st_rew = total; st_e1 = e1; st_e2 = e2;
            if (MODE == 0) {
                p.progress[n] = prog;
                out_timeout = time_out && out_reset;                                // vec_task.py:345
                if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
                if (p.ep_return) {
                    const float ret = (do_reset ? 0.0f : ret_old) + total;
                    if (out_reset) st_ret = ret;
                    p.ep_return[n] = ret;
                }
                if (out_reset) st_len = static_cast<float>(prog);
            }
        }
    }
    if (MODE == 0) {
        const unsigned m_reset = __ballot_sync(0xffffffffu, out_reset), m_to = __ballot_sync(0xffffffffu, out_timeout);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_ret += __shfl_xor_sync(0xffffffffu, st_ret, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            st_e1 += __shfl_xor_sync(0xffffffffu, st_e1, o);
            st_e2 += __shfl_xor_sync(0xffffffffu, st_e2, o);
        }
        double* pr = p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS;
        const float vals[TK_STATS] = {st_rew, static_cast<float>(__popc(m_reset)), static_cast<float>(__popc(m_to)), st_ret,
                                      st_len, st_e1, st_e2, static_cast<float>(nt)};
        if (lane < TK_STATS) atomicAdd(pr + lane, static_cast<double>(vals[lane]));
    }
    if (MODE == 2) return;
    float* out = p.obs + e0 * kObs;
    if (bulk) {
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            bulk_store(out, s.obs, kTile * kObs * 4);
            bulk_commit();
            bulk_wait_read<0>();
        }
    } else {
        __syncwarp();
        for (int i = lane; i < nt * kObs; i += kTile) out[i] = s.obs[i];
    }
}

// pre_physics_step :229-232 (+ clamp vec_task.py:327): flat stream over N*12 floats
__global__ void __launch_bounds__(256) anymal_pre_kernel(const __grid_constant__ KParams p, const float* __restrict__ actions_in,
                                                          float* __restrict__ targets, int clip) {
    const long long total = static_cast<long long>(p.n) * kDof;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float a = actions_in[i];
        if (clip) a = fminf(fmaxf(a, -p.clip_actions), p.clip_actions);
        p.actions[i] = a;                                                           // :230
        targets[i] = p.action_scale * a + p.default_dof_pos[i % kDof];              // :231
    }
}

// reset_idx :285-311 for an explicit id list (the reference leaves reset_buf = 1, :310)
__global__ void anymal_reset_idx_kernel(const __grid_constant__ KParams p, const long long* __restrict__ ids, int k,
                                        const float* __restrict__ unit_pos, const float* __restrict__ unit_vel,
                                        const float* __restrict__ unit_cmd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    const long long ne = ids[i];
    if (ne < 0 || ne >= p.n) return;
    float2* gd = reinterpret_cast<float2*>(p.dof_state) + ne * kDof;
    for (int j = 0; j < kDof; ++j) {
        const float pos = p.default_dof_pos[j] * rand_float(0.5f, 1.0f, unit_pos[static_cast<size_t>(i) * kDof + j]);
        const float vel = rand_float(-0.1f, static_cast<float>(0.1 - (-0.1)), unit_vel[static_cast<size_t>(i) * kDof + j]);
        gd[j] = make_float2(pos, vel);
    }
    for (int c = 0; c < kRoot; ++c) p.root[ne * kRoot + c] = p.init_root[c];
    p.commands[ne * 3 + 0] = rand_float(p.cmd_x[0], p.cmd_x[1], unit_cmd[static_cast<size_t>(i) * 3 + 0]);
    p.commands[ne * 3 + 1] = rand_float(p.cmd_y[0], p.cmd_y[1], unit_cmd[static_cast<size_t>(i) * 3 + 1]);
    p.commands[ne * 3 + 2] = rand_float(p.cmd_yaw[0], p.cmd_yaw[1], unit_cmd[static_cast<size_t>(i) * 3 + 2]);
    p.progress[ne] = 0;                                                             // :309
    p.reset[ne] = 1;                                                                // :310
}

__global__ void anymal_philox_kernel(int n, unsigned long long seed, unsigned long long step, long long env_off,
                                     const unsigned long long* __restrict__ epoch, float* up, float* uv, float* uc) {
    step = effective_step(step, epoch);
    const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= n) return;
    for (int b = 0; b < 3; ++b) {
        Philox4 a = philox_env(seed, step, env_off + e, kStreamPos + b), c = philox_env(seed, step, env_off + e, kStreamVel + b);
        if (up) { float* o = up + e * kDof + 4 * b; o[0] = u01(a.x); o[1] = u01(a.y); o[2] = u01(a.z); o[3] = u01(a.w); }
        if (uv) { float* o = uv + e * kDof + 4 * b; o[0] = u01(c.x); o[1] = u01(c.y); o[2] = u01(c.z); o[3] = u01(c.w); }
    }
    if (uc) {
        Philox4 d = philox_env(seed, step, env_off + e, kStreamCmd);
        uc[e * 3] = u01(d.x); uc[e * 3 + 1] = u01(d.y); uc[e * 3 + 2] = u01(d.z);
    }
}

__global__ void anymal_stats_finalize_kernel(double* __restrict__ partials, int rows, double* __restrict__ out, int accumulate) {
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k >= TK_STATS) return;
    double acc = 0.0;
    for (int r = lane; r < rows; r += 32) {
        acc += partials[static_cast<size_t>(r) * TK_STATS + k];
        partials[static_cast<size_t>(r) * TK_STATS + k] = 0.0;
    }
    acc = warp_sum(acc);
    if (lane == 0) out[k] = accumulate ? out[k] + acc : acc;
}

}  // namespace anymal
}  // namespace tk

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace tk;
using namespace tk::anymal;

struct TkAnymal {
    TkAnymalParams params;
    TkAnymalBuffers buf;
    bool bound, ran;
    int tiles, use_pdl, sm_count, stagger_cycles, stagger_slots;
    double* partials;
    unsigned long long* step_dev;
    int knob_prefetch, dbg_reset;      // developer knobs, read once in tk_anymal_create
};

static void fill_kparams(const TkAnymal* h, KParams& k) {
    const TkAnymalParams& P = h->params;
    const TkAnymalBuffers& B = h->buf;
    k.root = B.root_states; k.dof_state = B.dof_state; k.torques = B.torques; k.contact = B.contact_forces;
    k.commands = B.commands; k.actions = B.actions; k.obs = B.obs_buf; k.rew = B.rew_buf;
    k.reset = reinterpret_cast<long long*>(B.reset_buf); k.progress = reinterpret_cast<long long*>(B.progress_buf);
    k.timeout = B.timeout_buf; k.ep_return = B.episode_return; k.partials = h->partials;
    k.unit_pos = nullptr; k.unit_vel = nullptr; k.unit_cmd = nullptr; k.seed = 0; k.step = 0; k.step_dev = h->step_dev;
    k.env_off = 0; k.noise_mode = TK_NOISE_NONE;
    k.n = P.num_envs; k.max_episode_length = P.max_episode_length; k.base_index = P.base_index; k.num_knee = P.num_knee;
    for (int i = 0; i < kMaxKnee; ++i) k.knee[i] = P.knee_indices[i];
    k.stagger_cycles = h->stagger_cycles; k.sm_count = h->sm_count; k.stagger_ctas = h->stagger_slots * h->sm_count;
    // pays while the whole grid is resident at once (one wave: 11 CTAs / SM = 52096 envs): 7.0 -> 6.5 us at 16384 envs, 8.3 -> 7.4 at
    // 32768, 9.4 -> 8.8 at 49152; beyond the wave it costs (11.8 -> 12.4 us at 65536)
    k.prefetch = (h->tiles <= h->stagger_slots * h->sm_count) ? 1 : 0;
    if (h->knob_prefetch != kKnobUnset) k.prefetch = h->knob_prefetch;
    k.dbg_reset = h->dbg_reset;
    k.vec_ok = aligned_to(B.root_states, 16) && aligned_to(B.dof_state, 16) && aligned_to(B.torques, 16) &&
               aligned_to(B.contact_forces, 16) && aligned_to(B.commands, 16) && aligned_to(B.obs_buf, 16);
    k.dof_pos_scale = P.dof_pos_scale; k.dof_vel_scale = P.dof_vel_scale; k.action_scale = P.action_scale;
    k.clip_actions = P.clip_actions;
    k.k_lin = static_cast<float>(P.rew_lin_vel_xy * P.dt); k.k_ang = static_cast<float>(P.rew_ang_vel_z * P.dt);
    k.k_torque = static_cast<float>(P.rew_torque * P.dt);
    k.cmd_x[0] = static_cast<float>(P.command_x_range[0]); k.cmd_x[1] = static_cast<float>(P.command_x_range[1] - P.command_x_range[0]);
    k.cmd_y[0] = static_cast<float>(P.command_y_range[0]); k.cmd_y[1] = static_cast<float>(P.command_y_range[1] - P.command_y_range[0]);
    k.cmd_yaw[0] = static_cast<float>(P.command_yaw_range[0]);
    k.cmd_yaw[1] = static_cast<float>(P.command_yaw_range[1] - P.command_yaw_range[0]);
    for (int j = 0; j < kDof; ++j) k.default_dof_pos[j] = P.default_dof_pos[j];
    for (int c = 0; c < kRoot; ++c) k.init_root[c] = P.base_init_state[c];
    for (int c = 0; c < 3; ++c) k.gravity[c] = P.gravity_vec[c];
}

static int check_bound(const TkAnymal* h) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(h->bound, TK_ERR_UNBOUND, "tk_anymal_bind was not called");
    return TK_OK;
}

template <int MODE>
static void launch_tile(const TkAnymal* h, const KParams& k, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(h->tiles);
    cfg.blockDim = dim3(kTile);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (MODE == 0) ? h->use_pdl : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, anymal_tile_kernel<MODE>, k);
}

extern "C" {

int tk_anymal_create(const TkAnymalParams* params, TkAnymal** out) {
    TK_REQUIRE(params && out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(params->num_envs > 0, TK_ERR_INVALID, "num_envs must be positive (got %d)", params->num_envs);
    TK_REQUIRE(params->num_dof == kDof && params->num_bodies == kBodies, TK_ERR_INVALID,
               "this build serves the 12-DOF / 13-body ANYmal of anymal.py (got %d DOFs, %d bodies)", params->num_dof,
               params->num_bodies);
    TK_REQUIRE(params->num_knee >= 0 && params->num_knee <= kMaxKnee, TK_ERR_INVALID, "bad num_knee %d", params->num_knee);
    TK_REQUIRE(params->base_index >= 0 && params->base_index < kBodies, TK_ERR_INVALID, "bad base_index");
    for (int i = 0; i < params->num_knee; ++i)
        TK_REQUIRE(params->knee_indices[i] >= 0 && params->knee_indices[i] < kBodies, TK_ERR_INVALID, "knee index out of range");
    TK_CUDA(cudaSetDevice(params->device));
    TkAnymal* h = new (std::nothrow) TkAnymal();
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "out of host memory");
    h->params = *params;
    h->bound = false; h->ran = false;
    h->use_pdl = 1;
    if (const char* env = knob("TK_ANYMAL_PDL")) h->use_pdl = atoi(env) != 0;
    h->knob_prefetch = knob_int("TK_ANYMAL_PREFETCH");
    h->dbg_reset = ablation_knob("TK_ANYMAL_DBG_RESET");
    h->tiles = (params->num_envs + kTile - 1) / kTile;
    h->sm_count = 148;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, params->device);
    h->stagger_cycles = 500; h->stagger_slots = 11;
    {
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, anymal_tile_kernel<0>, kTile, 0) == cudaSuccess && occ > 0) h->stagger_slots = occ;
        else cudaGetLastError();
    }
    if (h->tiles <= h->sm_count * 2) h->stagger_cycles = 0;      // a thin single wave is only delayed by it
    if (const char* env = knob("TK_ANYMAL_STAGGER")) h->stagger_cycles = atoi(env);
    h->partials = nullptr; h->step_dev = nullptr;
    cudaError_t e = cudaMalloc(&h->partials, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMemset(h->partials, 0, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMalloc(&h->step_dev, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->step_dev, 0, sizeof(unsigned long long));
    if (e != cudaSuccess) {
        cudaFree(h->partials); cudaFree(h->step_dev);
        delete h;
        return cuda_fail(e, "tk_anymal_create");
    }
    *out = h;
    return TK_OK;
}

int tk_anymal_destroy(TkAnymal* h) {
    if (!h) return TK_OK;
    cudaSetDevice(h->params.device);
    cudaFree(h->partials); cudaFree(h->step_dev);
    delete h;
    return TK_OK;
}

int tk_anymal_bind(TkAnymal* h, const TkAnymalBuffers* b) {
    TK_REQUIRE(h && b, TK_ERR_INVALID, "null argument");
    const void* req[] = {b->root_states, b->dof_state, b->torques, b->contact_forces, b->commands, b->actions, b->obs_buf,
                         b->rew_buf, b->reset_buf, b->progress_buf};
    const char* names[] = {"root_states", "dof_state", "torques", "contact_forces", "commands", "actions", "obs_buf",
                           "rew_buf", "reset_buf", "progress_buf"};
    for (unsigned i = 0; i < sizeof(req) / sizeof(req[0]); ++i) {
        TK_REQUIRE(req[i] != nullptr, TK_ERR_UNBOUND, "buffer %s is NULL", names[i]);
        TK_REQUIRE(aligned_to(req[i], 4), TK_ERR_ALIGN, "buffer %s is not 4-byte aligned", names[i]);
    }
    TK_REQUIRE(aligned_to(b->reset_buf, 8) && aligned_to(b->progress_buf, 8), TK_ERR_ALIGN, "int64 buffers must be 8-byte aligned");
    TK_REQUIRE(aligned_to(b->dof_state, 8), TK_ERR_ALIGN, "dof_state must be 8-byte aligned ((pos,vel) pairs)");
    h->buf = *b;
    h->bound = true;
    return TK_OK;
}

int tk_anymal_pre_physics(TkAnymal* h, const float* actions_in, float* targets_out, int32_t clip, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(actions_in && targets_out, TK_ERR_INVALID, "null argument");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    const long long total = static_cast<long long>(k.n) * kDof;
    long long blocks = (total + 255) / 256;
    if (blocks > static_cast<long long>(h->sm_count) * 16) blocks = static_cast<long long>(h->sm_count) * 16;
    anymal_pre_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(k, actions_in, targets_out, clip);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_compute_observations(TkAnymal* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    launch_tile<1>(h, k, static_cast<cudaStream_t>(stream));
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_compute_reward(TkAnymal* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    launch_tile<2>(h, k, static_cast<cudaStream_t>(stream));
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_reset_idx(TkAnymal* h, const int64_t* env_ids, int32_t kcount, const float* unit_pos, const float* unit_vel,
                        const float* unit_cmd, void* stream) {
    if (int rc = check_bound(h)) return rc;
    if (kcount == 0) return TK_OK;
    TK_REQUIRE(kcount > 0 && env_ids && unit_pos && unit_vel && unit_cmd, TK_ERR_INVALID, "bad reset_idx arguments");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    anymal_reset_idx_kernel<<<(kcount + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        k, reinterpret_cast<const long long*>(env_ids), kcount, unit_pos, unit_vel, unit_cmd);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_post_physics(TkAnymal* h, const TkAnymalNoise* noise, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(noise != nullptr, TK_ERR_INVALID, "null noise descriptor");
    TK_REQUIRE(noise->mode == TK_NOISE_NONE || noise->mode == TK_NOISE_DENSE || noise->mode == TK_NOISE_PHILOX, TK_ERR_INVALID,
               "noise mode %d: Anymal takes TK_NOISE_NONE, TK_NOISE_DENSE (row n = env n) or TK_NOISE_PHILOX", noise->mode);
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    k.noise_mode = noise->mode;
    if (noise->mode == TK_NOISE_DENSE) { k.unit_pos = noise->unit_pos; k.unit_vel = noise->unit_vel; k.unit_cmd = noise->unit_cmd; }
    k.seed = noise->seed; k.step = noise->step; k.env_off = noise->env_id_offset;
    launch_tile<0>(h, k, static_cast<cudaStream_t>(stream));
    TK_CUDA(cudaGetLastError());
    h->ran = true;
    return TK_OK;
}

int tk_anymal_advance_epoch(TkAnymal* h, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    advance_epoch_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(h->step_dev);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_stats(TkAnymal* h, double* stats_out, int32_t accumulate, void* stream) {
    TK_REQUIRE(h && stats_out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(h->ran, TK_ERR_INVALID, "tk_anymal_post_physics has not run yet");
    TK_CUDA(cudaSetDevice(h->params.device));
    anymal_stats_finalize_kernel<<<1, TK_STATS * 32, 0, static_cast<cudaStream_t>(stream)>>>(h->partials, kStatSlots, stats_out, accumulate);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_anymal_philox_noise(TkAnymal* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* unit_pos, float* unit_vel,
                           float* unit_cmd, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    anymal_philox_kernel<<<(h->params.num_envs + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        h->params.num_envs, seed, step, env_id_offset, h->step_dev, unit_pos, unit_vel, unit_cmd);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

}  // extern "C"

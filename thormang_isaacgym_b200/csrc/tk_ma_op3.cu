// MA_OP3 env-step hot path for B200 (sm_100a) -- SURVEY 8(f) row N3 (two OP3 humanoids + one table per env).
//
// Replaces, behind the C ABI of include/thormang_b200.h, the TorchScript functions and method bodies of
// /root/reference/isaacgymenvs/tasks/MA_OP3.py: compute_op3_observations :662-708, compute_op3_reward :550-658,
// compute_objective_reward :496-547, the contact terms :319-358, compute_observations :411-438, compute_reward :360-409,
// reset_idx :440-489.  The reference's step hooks are stubs (:309-317); the fused step is the anymal.py:229-242 ordering
// around those methods (oracle/ma_op3_oracle.py states the assumptions).
//
// One CTA = one tile of 16 consecutive envs = 32 (env, agent) rows.  The fused step runs 64 threads per CTA: lanes 0..31 build
// the 32 observation rows, lanes 32..63 the rewards / reset bits of the same rows (row = 2 * env + agent, so a lane's row of
// the torque / action tiles starts 88 B after its neighbour's and its dof_state row 176 B after: the 64- and 128-bit shared
// loads touch every bank once per phase, and the two agents of an env sit in neighbouring lanes -- the table terms need
// both: one shuffle, no barrier).  root / dof_state / torques / prev_torques / actions / last_actions AND the tile's whole
// contact block (16 x 51 x 12 B) arrive as cp.async.bulk (TMA) tile copies on one mbarrier; per-env scalars are requested
// by the lanes before the tile wait.  An observation row is built in registers and leaves as 11 streaming 256-bit stores
// (352-byte rows; 128-bit stores when the base is only 16-byte aligned); prev_torques / last_actions are refreshed by
// bulk-storing the torque / action tiles already resident in shared memory.  7 CTAs per SM (register target).
// History: v1 staged padded obs rows and all six tiles (64 KB, 3 CTAs / SM): 46 % of HBM peak, stalled on the tile wait;
// v3 (warp per agent, stride-88 row reads: 58 % of the shared wavefronts conflicted, 166 registers): 62 %; this layout: 82 %
// at 524288 envs (profiles/r02_n3_ma_op3_tile_kernel_524288.json).
//
// Algorithmic bytes / env-step (F = 4 feet): reads root 156 + dof 352 + torques 176 + prev_torques 176 + actions 176 +
// last_actions 176 + foot contacts 96 + gripper x-forces 16 + commands 12 + potentials 8 + table potential 4 +
// air time 32 + last contacts 8 + progress 8 + reset 8 = 1404; writes obs 704 + prev_torques 176 + last_actions 176 +
// rew 8 + reset 8 + progress 8 + potentials 16 + prev table potential 4 + air time 32 + last contacts 8 + timeout 1
// = 1141  =>  2545 B.
#include "tk_common.cuh"

#include <cstdlib>
#include <new>

namespace tk {
namespace op3 {

constexpr int kA = TK_OP3_AGENTS;
constexpr int kRoot = 13;
constexpr int kActors = 3;
constexpr int kDof = TK_OP3_DOF;
constexpr int kObs = TK_OP3_OBS;
#ifndef TK_OP3_TILE
#define TK_OP3_TILE 16          // envs per CTA (16 or 32): one warp of observation lanes + one of reward lanes per 16 envs
#endif
constexpr int kTile = TK_OP3_TILE;
constexpr int kMaxContactTileBytes = 24 * 1024;     // a tile's whole contact block is staged when it is at most this big
constexpr int kRows = kTile * kA;            // (env, agent) rows of a tile
constexpr int kMaxFeet = TK_OP3_MAX_FEET;

struct KParams {
    float* root;
    const float* init_root;
    float* dof_state;
    const float* torques;
    float* prev_torques;
    const float* contact;
    float* commands;
    float* actions;
    float* last_actions;
    float* air_time;
    uint8_t* last_contacts;
    float* potentials;
    float* prev_potentials;
    float* table_pot;
    float* prev_table_pot;
    float* obs;
    float* rew;
    long long* reset;
    long long* progress;
    uint8_t* timeout;
    float* ep_return;
    double* partials;
    unsigned long long* grip_counts;         // [agents][2]
    float* pre_rew;                          // [N,2] un-clipped per-agent totals before the hold term (fused step only)
    const float* unit_cmd;
    unsigned long long seed, step;
    unsigned long long* step_dev;
    long long env_off;
    int noise_mode;
    int n, nb, max_episode_length, num_feet;
    int feet[kA][kMaxFeet];
    int grip[kA][2];
    int vec_ok;                              // 16-byte aligned tile arrays: cp.async.bulk tile copies
    int stagger_cycles, stagger_ctas, sm_count;   // first-wave stagger (same finding as in the Humanoid kernel)
    int contact_tile;                        // the tile's contact_forces block [kTile, nb, 3] is bulk-copied as well
    int obs32, vec_feet, contact8;           // 32-byte aligned obs rows; 16-/4-byte aligned air_time / last_contacts; 8-byte contact
    float lin_vel_scale, ang_vel_scale, dof_pos_scale, dof_vel_scale, action_scale, clip_actions, dt;
    float k_torque, k_up, k_air, k_syns, k_nofly, k_stand, k_rate;     // (float)(scale * dt), MA_OP3.py:80-81
    float cmd_x[2], cmd_y[2], cmd_yaw[2];   // lo, (hi - lo) formed in double
    float default_dof_pos[kDof];
    float inv_start_rot[4], gravity[3], up[3], targets[3], goal[3];
    long long* dbg_ts;                       // developer timeline probe (tools/op3_timeline.py): [tiles, 4 warps, 8]; NULL in normal use
};

enum { kStreamCmd = 48 };

__device__ __forceinline__ float rand_float(float lo, float span, float u) { return span * u + lo; }   // torch_rand_float
// torch.norm(dim=1) of two elements
__device__ __forceinline__ float norm2_ref(float x, float y) { return sqrtf(fmaf(y, y, x * x)); }

struct Smem {
    alignas(128) float root[kTile * kActors * kRoot];
    alignas(16) float dof[kTile * kA * kDof * 2];
    alignas(16) float tq[kTile * kA * kDof];
    alignas(16) float act[kTile * kA * kDof];
    alignas(16) float ptq[kTile * kA * kDof];     // prev_torques / last_actions tiles: bulk-loaded like the others.  Per-lane
    alignas(16) float la[kTile * kA * kDof];      // row reads straight from global (88 B rows, 176 B lane stride) lived on L1
                                                  // sector reuse: the step time followed the L1 size left by the shared-memory
                                                  // carve-out (336 us with 60 KB of L1, 400 us with 28 KB, 524288 envs)
    alignas(16) float sums[kRows * 4];            // fused step: the four 22-DOF sums of a row, observation lane -> reward lane
    alignas(8) uint64_t bar;
};

// obs row piece c (columns 4c .. 4c+3) is complete after iteration `obs_piece_ready(c)` of the per-DOF loop (-1: before it):
// columns 0..8 head, 9..30 dof_pos (j = col - 9), 31..52 dof_vel (j = col - 31), 53..74 actions (j = col - 53), 75..87 table
__host__ __device__ constexpr int obs_col_ready(int col) {
    return col < 9 ? -1 : col < 31 ? col - 9 : col < 53 ? col - 31 : col < 75 ? col - 53 : -1;
}
__host__ __device__ constexpr int obs_piece_ready(int c) {      // piece = 8 columns = one 32-byte sector of the row
    int r = -1;
    for (int k = 0; k < 8; ++k) r = obs_col_ready(8 * c + k) > r ? obs_col_ready(8 * c + k) : r;
    return r;
}
__device__ __forceinline__ void store_obs_piece(float* row, int c, const float* v, bool obs32) {
    if (obs32) {
        stg_stream8(row + 8 * c, v + 8 * c);
    } else {
        stg_stream4(row + 8 * c, make_float4(v[8 * c], v[8 * c + 1], v[8 * c + 2], v[8 * c + 3]));
        stg_stream4(row + 8 * c + 4, make_float4(v[8 * c + 4], v[8 * c + 5], v[8 * c + 6], v[8 * c + 7]));
    }
}

// MODE 0: fused post_physics_step; 1: compute_observations; 2: compute_reward
#ifndef TK_OP3_MIN_CTAS
#define TK_OP3_MIN_CTAS (TK_OP3_TILE == 16 ? 7 : 3)        // resident CTAs per SM the register allocation targets (developer knob)
#endif

template <int MODE> __host__ __device__ constexpr int threads_of() { return MODE == 0 ? 2 * kRows : kRows; }

// compute_observations :411-438 / compute_op3_observations :662-708 for one (env, agent) row: potentials bookkeeping
__device__ __forceinline__ void potentials_step(const KParams& p, long long n, int a, const float* r, const float* tb,
                                                float pot_old, float& pot_new, float& ppot_new, float& ptpot) {
    ppot_new = pot_old;                                                             // :688
    pot_new = -norm3_ref(V3{p.goal[0] - r[0], p.goal[1] - r[1], 0.0f}) / p.dt;      // :680-689
    p.potentials[n * kA + a] = pot_new;
    p.prev_potentials[n * kA + a] = ppot_new;
    if (a == 0) {
        ptpot = -norm3_ref(V3{p.targets[0] - tb[0], p.targets[1] - tb[1], p.targets[2] - tb[2]}) / p.dt;   // :418-419
        p.prev_table_pot[n] = ptpot;
    }
}

// NF > 0: number of feet known at compile time (4 for the OP3 of MA_OP3.py:237-238); NF = 0: p.num_feet at run time
template <int MODE, int NF>
__global__ void __launch_bounds__(threads_of<MODE>(), TK_OP3_MIN_CTAS) op3_tile_kernel(const __grid_constant__ KParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    // thread -> ((env, agent) row, role): rows are consecutive over the lanes of a warp (lane = 2 * env + agent within the
    // warp's 16 envs), so a lane's row of the torque / action tiles starts 88 B after its neighbour's and its dof_state row
    // 176 B after: 64-bit / 128-bit shared loads of those rows touch every bank once per phase (no conflicts), and the two
    // agents of an env sit in neighbouring lanes (the table terms need both: one shuffle, no barrier).  In the fused step
    // threads 0..63 build the observation rows and threads 64..127 the rewards / resets of the same 64 rows.
    const int row = tid & (kRows - 1), le = row >> 1, a = row & 1;
    const bool role_rew = MODE == 2 || (MODE == 0 && tid >= kRows);
    const bool role_obs = !role_rew;
    const int env0 = blockIdx.x * kTile;
    const int nt = min(kTile, p.n - env0);
    const long long n = env0 + le;
    const bool valid = le < nt;
    const bool bulk = p.vec_ok && nt == kTile;
    const size_t e0 = static_cast<size_t>(env0);
    constexpr bool kNeedTq = MODE != 1;
    constexpr int kThreads = threads_of<MODE>();
    // the finalize grid that follows (programmatic dependent launch) may be scheduled as soon as every CTA got here; it
    // still waits for this grid's completion and memory flush in its own griddepcontrol.wait
    if (MODE == 0) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    long long* ts = (MODE == 0 && p.dbg_ts) ? p.dbg_ts + (static_cast<size_t>(blockIdx.x) * 4 + (tid >> 5)) * 8 : nullptr;
#define TK_OSTAMP(i) do { if (ts && lane == 0) { long long g_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_) :: "memory"); ts[i] = g_; } } while (0)
    TK_OSTAMP(0);
    // First-wave stagger: the CTAs of the first wave start together, so every SM would run load-all / math-all / store-all
    // in lockstep and HBM would idle during the math; resident slot s of the first wave waits s * stagger_cycles.
    if (MODE == 0 && p.stagger_cycles > 0 && blockIdx.x < static_cast<unsigned>(p.stagger_ctas)) {
        const long long until = clock64() + static_cast<long long>(blockIdx.x / p.sm_count) * p.stagger_cycles;
        while (clock64() < until) __nanosleep(32);
    }
    constexpr uint32_t kRootB = kTile * kActors * kRoot * 4, kDofB = kTile * kA * kDof * 2 * 4, kRowB = kTile * kA * kDof * 4;
    // The feet / gripper rows an env needs are 12 of its nb = 51 contact rows, 12 bytes each at a 612-byte stride: fetched by
    // the lanes they cost one 64-byte DRAM atom per row (768 B per env, more than the 612-byte block) and, worse, one L1 miss
    // per row and lane -- the reward warps spent 6 us of their 12 us just getting those loads accepted.  The block comes
    // with the other tiles instead.
    float* s_contact = reinterpret_cast<float*>(smem_raw + sizeof(Smem));
    const bool ctile = kNeedTq && bulk && p.contact_tile;
    const uint32_t contact_bytes = static_cast<uint32_t>(kTile) * static_cast<uint32_t>(p.nb) * 12u;
    if (bulk) {
        if (tid == 0) {                                 // the issuing thread needs no barrier between init and use
            mbar_init(&s.bar, 1);
            fence_proxy_async_smem();
            mbar_expect_tx(&s.bar, kRootB + kDofB + kRowB + (kNeedTq ? 3 * kRowB : 0) + (ctile ? contact_bytes : 0u));
            if (ctile) bulk_load(s_contact, p.contact + e0 * p.nb * 3, contact_bytes, &s.bar);
            bulk_load(s.dof, p.dof_state + e0 * kA * kDof * 2, kDofB, &s.bar);
            bulk_load(s.root, p.root + e0 * kActors * kRoot, kRootB, &s.bar);
            bulk_load(s.act, p.actions + e0 * kA * kDof, kRowB, &s.bar);
            if (kNeedTq) {
                bulk_load(s.tq, p.torques + e0 * kA * kDof, kRowB, &s.bar);
                bulk_load(s.ptq, p.prev_torques + e0 * kA * kDof, kRowB, &s.bar);
                bulk_load(s.la, p.last_actions + e0 * kA * kDof, kRowB, &s.bar);
            }
        }
    } else {
        for (int i = tid; i < nt * kActors * kRoot; i += kThreads) s.root[i] = p.root[e0 * kActors * kRoot + i];
        for (int i = tid; i < nt * kA * kDof * 2; i += kThreads) s.dof[i] = p.dof_state[e0 * kA * kDof * 2 + i];
        for (int i = tid; i < nt * kA * kDof; i += kThreads) {
            s.act[i] = p.actions[e0 * kA * kDof + i];
            if (kNeedTq) {
                s.tq[i] = p.torques[e0 * kA * kDof + i];
                s.ptq[i] = p.prev_torques[e0 * kA * kDof + i];
                s.la[i] = p.last_actions[e0 * kA * kDof + i];
            }
        }
    }
    // ---- per-env scalars and the foot contact rows: issued before the tile wait ----
    long long prog = 0, rb = 0, reset_raw = 0;
    bool do_reset = false;
    float cmd[3] = {0.f, 0.f, 0.f}, pot_old = 0.f, ppot_old = 0.f, tpot = 0.f, ptpot = 0.f, ret_old = 0.f;
    float cf[kMaxFeet][3], air[kMaxFeet];
    unsigned lc_raw = 0;                        // last_contacts of this lane's feet: the four bytes as loaded (four-foot build),
                                                // else one bit per foot
    float gx0 = 0.f, gx1 = 0.f;                 // x contact force on this agent's left / right gripper
    const int nf = NF > 0 ? NF : p.num_feet;
    const bool feet_vec = NF == 4 && p.vec_feet;
    if (valid) {
        if (MODE != 1) reset_raw = p.reset[n];
        if (role_rew || MODE == 1) pot_old = p.potentials[n * kA + a];
        if (role_rew) {
            prog = p.progress[n];
#pragma unroll
            for (int c = 0; c < 3; ++c) cmd[c] = p.commands[n * 3 + c];
            if (MODE == 2) ppot_old = p.prev_potentials[n * kA + a];
            tpot = p.table_pot[n];
            if (MODE == 2) ptpot = p.prev_table_pot[n];
            if (MODE == 0 && !ctile) {
                gx0 = p.contact[(static_cast<size_t>(n) * p.nb + p.grip[a][0]) * 3];
                gx1 = p.contact[(static_cast<size_t>(n) * p.nb + p.grip[a][1]) * 3];
            }
            if (feet_vec) {                             // this lane's four feet: one 128-bit and one 32-bit load
                const float4 at = *reinterpret_cast<const float4*>(p.air_time + (n * kA + a) * 4);
                lc_raw = *reinterpret_cast<const unsigned*>(p.last_contacts + (n * kA + a) * 4);
                air[0] = at.x; air[1] = at.y; air[2] = at.z; air[3] = at.w;
            }
#pragma unroll
            for (int f = 0; f < kMaxFeet; ++f) {
                if (f < nf) {
                    if (!ctile) {
                        const float* crow = p.contact + (static_cast<size_t>(n) * p.nb + p.feet[a][f]) * 3;
                        if (p.contact8) ldg_row3(crow, cf[f][0], cf[f][1], cf[f][2]);
                        else { cf[f][0] = crow[0]; cf[f][1] = crow[1]; cf[f][2] = crow[2]; }
                    }
                    if (!feet_vec) {
                        air[f] = p.air_time[(n * kA + a) * nf + f];
                        lc_raw |= (p.last_contacts[(n * kA + a) * nf + f] != 0 ? 1u : 0u) << f;
                    }
                }
            }
        }
    }
    // Every thread has its reset flag in hand before any thread of the CTA can write reset_buf (the reward lanes do, at the
    // end): the barrier orders the two roles; it also publishes the mbarrier.  The other loads above stay in flight.
    // (__syncthreads_or takes the flag as an operand, so the load has landed when the thread arrives.)  Nothing that was
    // loaded above may be looked at before this point: the compiler otherwise turns loaded values into predicates as soon as
    // it can (prog + 1, gx > 0.1, flag != 0 ...), and every such use is a full memory latency in front of the loads behind it
    // (7 us from kernel entry to the barrier in the first build of this layout).  The empty asm statements pin that order.
    asm volatile("" : "+l"(reset_raw) :: "memory");
    TK_OSTAMP(1);
    if (MODE == 0) (void)__syncthreads_or(reset_raw != 0 ? 1 : 0);
    else __syncthreads();
    TK_OSTAMP(2);
    do_reset = MODE == 0 && reset_raw != 0;
    rb = reset_raw;
    // A resetting env takes its robot / table rows from initial_root_states (:468-470): those global loads are issued
    // here, under the tile wait, instead of in the middle of the compute phase (18 % of the stall samples in the v3 profile).
    float r[kRoot], tb[kRoot];
    if (MODE == 0 && valid && do_reset) {
        const float* ir = p.init_root + static_cast<size_t>(n) * kActors * kRoot;
#pragma unroll
        for (int c = 0; c < kRoot; ++c) { r[c] = ir[a * kRoot + c]; tb[c] = ir[2 * kRoot + c]; }
    }
    if (bulk) mbar_wait(&s.bar, 0);
    else __syncthreads();
    TK_OSTAMP(3);
    if (role_rew) {
        if (ctile && valid) {
            const float* cb = s_contact + static_cast<size_t>(le) * p.nb * 3;
            if (MODE == 0) { gx0 = cb[p.grip[a][0] * 3]; gx1 = cb[p.grip[a][1] * 3]; }
#pragma unroll
            for (int f = 0; f < kMaxFeet; ++f) {
                if (f < nf) {
                    const float* crow = cb + p.feet[a][f] * 3;
                    cf[f][0] = crow[0]; cf[f][1] = crow[1]; cf[f][2] = crow[2];
                }
            }
        }
        asm volatile("" : "+l"(prog), "+f"(gx0), "+f"(gx1), "+r"(lc_raw), "+f"(pot_old), "+f"(tpot));
#pragma unroll
        for (int f = 0; f < kMaxFeet; ++f)
            if (f < nf) asm volatile("" : "+f"(cf[f][0]), "+f"(cf[f][1]), "+f"(cf[f][2]), "+f"(air[f]));
        if (MODE == 0) {
            prog = do_reset ? 0 : prog + 1;                                         // :488
            // the episode return is only needed at the very end: its load rides under the row math
            if (valid && a == 0 && p.ep_return) ret_old = p.ep_return[n];
            // :351-352, this warp's share of the batch-global gripper counts (the two forces die here)
            const unsigned m0 = __ballot_sync(0xffffffffu, valid && gx0 > 0.1f), m1 = __ballot_sync(0xffffffffu, valid && gx1 > 0.1f);
            if (lane < 4) {                             // lane = agent * 2 + gripper; agent a sits in the lanes of parity a
                const unsigned m = ((lane & 1) ? m1 : m0) & ((lane >> 1) ? 0xaaaaaaaau : 0x55555555u);
                if (m) atomicAdd(p.grip_counts + lane, static_cast<unsigned long long>(__popc(m)));
            }
        }
    }

    float my_rew = 0.f, my_dist = 0.f;
    int my_flag = 0;                        // bit 0: up_proj < 0.90, bit 1: distance to the table > 0.40
    // Fused step: the reward lane was the long pole of a tile (row math 4.1 us against the observation lane's 1.5 us,
    // tools/op3_timeline.py, and the CTA lives as long as its slowest warp).  The four sums over the 22 DOFs (:619-630) read
    // what the observation lane walks anyway, so THAT lane accumulates them (same order, same bits) and hands them over
    // through shared memory; the reward lane keeps its other terms in registers across one CTA barrier and adds everything
    // up in the reference's order afterwards.
    constexpr bool kSplitSums = MODE == 0;
    float t_progress = 0.f, t_up = 0.f, t_air = 0.f, t_angz = 0.f, t_step = 0.f, t_nofly = 0.f, t_dist = 0.f, t_hold = 0.f;
    bool t_still = false;
    if (valid) {
        if (!(MODE == 0 && do_reset)) {
#pragma unroll
            for (int c = 0; c < kRoot; ++c) r[c] = s.root[((le * kActors) + a) * kRoot + c];
#pragma unroll
            for (int c = 0; c < kRoot; ++c) tb[c] = s.root[((le * kActors) + 2) * kRoot + c];
        }
        const float4* dof4 = reinterpret_cast<const float4*>(s.dof) + row * (kDof / 2);    // (pos_j, vel_j, pos_j+1, vel_j+1)
        const float2* act2 = reinterpret_cast<const float2*>(s.act) + row * (kDof / 2);
        const bool rst = MODE == 0 && do_reset;
        const Q4 q{r[3], r[4], r[5], r[6]};
        const V3 ang_raw = quat_rotate_ref<-1>(q, V3{r[10], r[11], r[12]});         // :579 / :693 before scaling
        if (role_obs) {                                                             // compute_observations :411-438
            if (rst) {
#pragma unroll
                for (int c = 0; c < kRoot; ++c) {                                   // :468-470 (rows preloaded above)
                    p.root[(n * kActors + a) * kRoot + c] = r[c];
                    if (a == 0) p.root[(n * kActors + 2) * kRoot + c] = tb[c];
                }
            }
            if (MODE == 1) {
                float pot_new, ppot_new;
                potentials_step(p, n, a, r, tb, pot_old, pot_new, ppot_new, ptpot);
            }
            const V3 lin = quat_rotate_ref<-1>(q, V3{r[7], r[8], r[9]});            // :692
            const V3 g = quat_rotate_ref<1>(q, V3{p.gravity[0], p.gravity[1], p.gravity[2]});   // :694
            float* o = p.obs + (static_cast<size_t>(n) * kA + a) * kObs;            // 352-byte rows = 11 sectors
            float v[kObs];
            v[0] = lin.x * p.lin_vel_scale; v[1] = lin.y * p.lin_vel_scale; v[2] = lin.z * p.lin_vel_scale;
            v[3] = ang_raw.x * p.ang_vel_scale; v[4] = ang_raw.y * p.ang_vel_scale; v[5] = ang_raw.z * p.ang_vel_scale;
            v[6] = g.x; v[7] = g.y; v[8] = g.z;
#pragma unroll
            for (int c = 0; c < 10; ++c) v[75 + c] = tb[c];                         // object pose + linvel
            v[85] = p.targets[0]; v[86] = p.targets[1]; v[87] = p.targets[2];
            // Every 32-byte piece of the row is stored as soon as its eight columns exist (obs_piece_ready: the j at which
            // the last of them is produced), not after the whole row: the row never sits in 88 registers.  One 256-bit
            // store per piece: a lane's row costs the L1 11 sector writes, not 22 half-sector ones.
#pragma unroll
            for (int c = 0; c < kObs / 8; ++c)
                if (obs_piece_ready(c) < 0) store_obs_piece(o, c, v, p.obs32);
            const float2* otq2 = reinterpret_cast<const float2*>(s.tq) + row * (kDof / 2);
            const float2* optq2 = reinterpret_cast<const float2*>(s.ptq) + row * (kDof / 2);
            const float2* ola2 = reinterpret_cast<const float2*>(s.la) + row * (kDof / 2);
            float tsum = 0.0f, legsum = 0.0f, ratesum = 0.0f, standsum = 0.0f;
#pragma unroll
            for (int jj = 0; jj < kDof / 2; ++jj) {
                float4 d = lds128(dof4 + jj);
                const float2 ac = act2[jj];
                if (rst) d = make_float4(p.default_dof_pos[2 * jj] + 0.0f, 0.0f, p.default_dof_pos[2 * jj + 1] + 0.0f, 0.0f);   // :456-457
                if (kSplitSums) {                                                   // compute_op3_reward :619-630 for the reward lane
                    const float2 t = otq2[jj], pt = optq2[jj];
                    float2 l = ola2[jj];
                    if (rst) l = make_float2(0.0f, 0.0f);                           // :486 last_actions of a reset env
                    const float px[2] = {d.x, d.z};
                    const float tv[2] = {t.x, t.y}, ptv[2] = {pt.x, pt.y}, av[2] = {ac.x, ac.y}, lv[2] = {l.x, l.y};
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int j = 2 * jj + h;
                        tsum += fabsf(ptv[h] - tv[h]);                              // :619
                        const float dev = fabsf(px[h] - p.default_dof_pos[j]);
                        standsum += dev;                                            // :630
                        if ((j >= 2 && j <= 7) || (j >= 11 && j <= 17) || j == 21) legsum += dev;   // :622-623
                        const float da = lv[h] - av[h];
                        ratesum += da * da;                                         // :627
                    }
                }
                v[9 + 2 * jj] = (d.x - p.default_dof_pos[2 * jj]) * p.dof_pos_scale;                // :695
                v[10 + 2 * jj] = (d.z - p.default_dof_pos[2 * jj + 1]) * p.dof_pos_scale;
                v[31 + 2 * jj] = d.y * p.dof_vel_scale;
                v[32 + 2 * jj] = d.w * p.dof_vel_scale;
                v[53 + 2 * jj] = ac.x;
                v[54 + 2 * jj] = ac.y;
#pragma unroll
                for (int c = 0; c < kObs / 8; ++c)
                    if (obs_piece_ready(c) >= 0 && (obs_piece_ready(c) >> 1) == jj) store_obs_piece(o, c, v, p.obs32);
            }
            if (kSplitSums) reinterpret_cast<float4*>(s.sums)[row] = make_float4(tsum, legsum, ratesum, standsum);
        } else {                                                                    // compute_reward :360-409, this agent
            bool zero_last = false;
            if (rst) {                                                              // reset_idx :440-489, this env's lanes
                // 64-bit stores: tk_ma_op3_bind guarantees dof_state 8-byte alignment only (a 128-bit store here faulted on an
                // 8-byte aligned base: found by tests/test_gpu_guard_bands.py)
                float2* gd = reinterpret_cast<float2*>(p.dof_state) + (n * kA + a) * kDof;
#pragma unroll
                for (int j = 0; j < kDof; ++j) gd[j] = make_float2(p.default_dof_pos[j] + 0.0f, 0.0f);   // :456-457 (offsets are zeros)
                float uc[3] = {0.5f, 0.5f, 0.5f};
                if (p.noise_mode == TK_NOISE_PHILOX) {
                    Philox4 d = philox_env(p.seed, effective_step(p.step, p.step_dev), p.env_off + n, kStreamCmd);
                    uc[0] = u01(d.x); uc[1] = u01(d.y); uc[2] = u01(d.z);
                } else if (p.noise_mode == TK_NOISE_DENSE && p.unit_cmd) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) uc[c] = p.unit_cmd[n * 3 + c];
                }
                cmd[0] = rand_float(p.cmd_x[0], p.cmd_x[1], uc[0]);                 // :472
                cmd[1] = rand_float(p.cmd_y[0], p.cmd_y[1], uc[1]);                 // :473
                cmd[2] = rand_float(p.cmd_yaw[0], p.cmd_yaw[1], uc[2]);             // :474
                pot_old = -norm3_ref(V3{p.goal[0] - r[0], p.goal[1] - r[1], 0.0f}) / p.dt;      // :476-479
                tpot = -norm3_ref(V3{p.targets[0] - tb[0], p.targets[1] - tb[1], 0.0f}) / p.dt; // :481-484
                if (a == 0) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) p.commands[n * 3 + c] = cmd[c];
                    p.table_pot[n] = tpot;
                }
                zero_last = true;                                                   // :486
#pragma unroll
                for (int f = 0; f < kMaxFeet; ++f) air[f] = 0.0f;                   // :487
            }
            float pot_r = pot_old, ppot_r = ppot_old;
            if (MODE == 0) potentials_step(p, n, a, r, tb, pot_old, pot_r, ppot_r, ptpot);
            const bool moving = norm2_ref(cmd[0], cmd[1]) > 0.1f, still = norm2_ref(cmd[0], cmd[1]) < 0.1f;
            // _reward_air_time :319-329, _reward_feet_contact_forces :331-334, _reward_no_fly :336-340
            float air_sum = 0.0f, step_sum = 0.0f;
            unsigned lc_new = 0;
            int touching = 0;
#pragma unroll
            for (int f = 0; f < kMaxFeet; ++f) {
                if (f < nf) {
                    const bool contact = cf[f][2] > 1.1f;
                    const bool filt = contact || (feet_vec ? ((lc_raw >> (8 * (f & 3))) & 0xffu) : ((lc_raw >> f) & 1u)) != 0;
                    const bool first = (air[f] > 0.0f) && filt;
                    const float t = air[f] + p.dt;
                    air_sum += (t - 0.5f) * (first ? 1.0f : 0.0f);
                    const float t2 = t * (filt ? 0.0f : 1.0f);
                    if (feet_vec) {
                        air[f] = t2;
                        lc_new |= (contact ? 1u : 0u) << (8 * (f & 3));
                    } else {
                        p.air_time[(n * kA + a) * nf + f] = t2;
                        p.last_contacts[(n * kA + a) * nf + f] = contact ? 1 : 0;
                    }
                    step_sum += fmaxf(norm3_ref(V3{cf[f][0], cf[f][1], cf[f][2]}) - 450.0f, 0.0f);
                    touching += cf[f][2] > 0.1f ? 1 : 0;
                }
            }
            if (feet_vec) {
                *reinterpret_cast<float4*>(p.air_time + (n * kA + a) * 4) = make_float4(air[0], air[1], air[2], air[3]);
                *reinterpret_cast<unsigned*>(p.last_contacts + (n * kA + a) * 4) = lc_new;
            }
            const float rew_airtime = (air_sum * (moving ? 1.0f : 0.0f)) * p.k_air;
            const float rew_no_fly = (touching == 1 ? 1.0f : 0.0f) * p.k_nofly;
            // _reward_gripper_contact_hold :346-358: batch-global counts (tk_ma_op3_grip_counts, all-reduced when sharded)
            // The fused step counts while it runs and assumes hold = 0; op3_finalize_kernel adds the term afterwards in the
            // (one env in the whole batch) case where it is 1.  compute_reward (MODE 2) reads the counts of the pre-pass.
            const float hold = (MODE == 2 && p.grip_counts[a * 2] == 1ull && p.grip_counts[a * 2 + 1] == 1ull) ? 1.0f : 0.0f;
            // compute_op3_reward :550-658
            const Q4 torso = quat_mul_ref(q, Q4{p.inv_start_rot[0], p.inv_start_rot[1], p.inv_start_rot[2], p.inv_start_rot[3]});   // :593
            const float up_proj = quat_rotate_ref<1>(torso, V3{p.up[0], p.up[1], p.up[2]}).z;    // :601-602
            const float rew_up = up_proj > 0.95f ? 0.0f + p.k_up : 0.0f;           // :605-606
            const float progress_reward = (pot_r - ppot_r) * 5.0f;                  // :612
            const float2* tq2 = reinterpret_cast<const float2*>(s.tq) + row * (kDof / 2);
            const float2* ptq2 = reinterpret_cast<const float2*>(s.ptq) + row * (kDof / 2);
            const float2* la2 = reinterpret_cast<const float2*>(s.la) + row * (kDof / 2);
            float tsum = 0.0f, legsum = 0.0f, ratesum = 0.0f, standsum = 0.0f;
#pragma unroll
            for (int jj = 0; jj < (kSplitSums ? 0 : kDof / 2); ++jj) {              // fused step: the observation lane has them
                const float4 d = lds128(dof4 + jj);                                  // only the positions are used here
                const float2 t = tq2[jj], pt = ptq2[jj], ac = act2[jj];
                float2 l = la2[jj];
                if (zero_last) l = make_float2(0.0f, 0.0f);
                const float px[2] = {rst ? p.default_dof_pos[2 * jj] + 0.0f : d.x, rst ? p.default_dof_pos[2 * jj + 1] + 0.0f : d.z};
                const float tv[2] = {t.x, t.y}, ptv[2] = {pt.x, pt.y}, av[2] = {ac.x, ac.y}, lv[2] = {l.x, l.y};
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int j = 2 * jj + h;
                    tsum += fabsf(ptv[h] - tv[h]);                                  // :619
                    const float dev = fabsf(px[h] - p.default_dof_pos[j]);
                    standsum += dev;                                                // :630
                    if ((j >= 2 && j <= 7) || (j >= 11 && j <= 17) || j == 21) legsum += dev;   // :622-623
                    const float da = lv[h] - av[h];
                    ratesum += da * da;                                             // :627
                }
            }
            const float dz = p.targets[2] - ang_raw.z;
            const float rew_ang_vel_z = expf(-(dz * dz) / 0.1f) * 0.1f;             // :632-633
            my_dist = norm3_ref(V3{r[0] - tb[0], r[1] - tb[1], r[2] - tb[2]});      // :644
            const float rew_distobject = expf(-my_dist / 0.32f) * 0.5f;             // :645
            my_flag = (up_proj < 0.90f ? 1 : 0) | (my_dist > 0.40f ? 2 : 0);        // :654-655
            if (kSplitSums) {                   // the total is formed behind the CTA barrier below, once the sums are there
                t_progress = progress_reward; t_up = rew_up; t_air = rew_airtime; t_angz = rew_ang_vel_z; t_step = step_sum;
                t_nofly = rew_no_fly; t_dist = rew_distobject; t_hold = hold; t_still = still;
            } else {
                const float rew_torque = tsum * p.k_torque;
                const float rew_syns = legsum * p.k_syns;                           // :624
                const float rew_action_rate = ratesum * p.k_rate;
                const float rew_stand_still = (standsum * (still ? 1.0f : 0.0f)) * p.k_stand;
                float total = progress_reward + 2.0f;                               // :648-649, left to right (rew_orient is -0)
                total += rew_torque; total += rew_up; total += rew_airtime; total += rew_ang_vel_z; total += step_sum;
                total += rew_no_fly; total += rew_stand_still; total += rew_action_rate; total += rew_syns;
                total += rew_distobject;
                total += hold;
                my_rew = fmaxf(total, 0.0f);                                        // :650
            }
        }
    }
    if (kSplitSums) {
        __syncthreads();                                // the observation lanes' sums are in shared memory
        if (role_rew && valid) {
            const float4 sm = reinterpret_cast<const float4*>(s.sums)[row];         // (tsum, legsum, ratesum, standsum)
            const float rew_torque = sm.x * p.k_torque;
            const float rew_syns = sm.y * p.k_syns;                                 // :624
            const float rew_action_rate = sm.z * p.k_rate;
            const float rew_stand_still = (sm.w * (t_still ? 1.0f : 0.0f)) * p.k_stand;
            float total = t_progress + 2.0f;                                        // :648-649, left to right (rew_orient is -0)
            total += rew_torque; total += t_up; total += t_air; total += t_angz; total += t_step;
            total += t_nofly; total += rew_stand_still; total += rew_action_rate; total += rew_syns;
            total += t_dist;
            p.pre_rew[n * kA + a] = total;
            total += t_hold;
            my_rew = fmaxf(total, 0.0f);                                            // :650
        }
    }
    TK_OSTAMP(4);
    if (role_rew) {                                     // warp-uniform: whole warps of reward lanes from here on
        // the other agent of this env sits in the neighbouring lane
        const float rew1 = __shfl_xor_sync(0xffffffffu, my_rew, 1), dist1 = __shfl_xor_sync(0xffffffffu, my_dist, 1);
        const int flag1 = __shfl_xor_sync(0xffffffffu, my_flag, 1);
        float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_d0 = 0.f, st_d1 = 0.f;
        bool out_reset = false, out_timeout = false;
        if (a == 0 && valid) {
            // reset_buf through the two compute_op3_reward calls (:653-656) ...
            long long rv = rb;
            const bool over = prog >= p.max_episode_length;
            if (over) rv = 1;
            if (my_flag) rv = 1;
            if (prog == 0) rv = 0;
            if (over) rv = 1;
            if (flag1) rv = 1;
            if (prog == 0) rv = 0;
            // ... and compute_objective_reward :496-547
            const Q4 tq{tb[3], tb[4], tb[5], tb[6]};
            const float tup = quat_rotate_ref<1>(tq, V3{p.up[0], p.up[1], p.up[2]}).z;  // :528-529
            const float rew_up = tup > 0.98f ? 0.0f + 0.1f : 0.0f;                      // :532-533
            const float hz = tb[2] - 0.29f;
            const float rew_h = (hz * hz) * -0.001f;                                    // :536
            const float dist_rew = (tpot - ptpot) * 5.0f;                               // :540
            const float r0 = ((my_rew + dist_rew) + rew_up) + rew_h;                    // :542
            const float r1 = ((rew1 + dist_rew) + rew_up) + rew_h;
            if (tup < 0.90f) rv = 1;                                                    // :544
            if (tb[2] < 0.25f) rv = 1;                                                  // :545
            *reinterpret_cast<float2*>(p.rew + n * kA) = make_float2(r0, r1);
            p.reset[n] = rv;
            out_reset = rv != 0;
            st_rew = r0 + r1; st_d0 = my_dist; st_d1 = dist1;
            if (MODE == 0) {
                p.progress[n] = prog;
                out_timeout = (prog >= p.max_episode_length - 1) && out_reset;          // multi_vec_task.py:193
                if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
                if (p.ep_return) {
                    const float ret = (do_reset ? 0.0f : ret_old) + (r0 + r1);
                    if (out_reset) st_ret = ret;
                    p.ep_return[n] = ret;
                }
                if (out_reset) st_len = static_cast<float>(prog);
            }
        }
        if (MODE == 0) {
            const unsigned m_reset = __ballot_sync(0xffffffffu, out_reset), m_to = __ballot_sync(0xffffffffu, out_timeout);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
                st_ret += __shfl_xor_sync(0xffffffffu, st_ret, o);
                st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
                st_d0 += __shfl_xor_sync(0xffffffffu, st_d0, o);
                st_d1 += __shfl_xor_sync(0xffffffffu, st_d1, o);
            }
            const int warp_envs = min(max(nt - (row >> 5) * 16, 0), 16);      // a warp = 16 envs x 2 agents
            double* pr = p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS;
            const float vals[TK_STATS] = {st_rew, static_cast<float>(__popc(m_reset)), static_cast<float>(__popc(m_to)), st_ret,
                                          st_len, st_d0, st_d1, static_cast<float>(warp_envs)};
            if (lane < TK_STATS) atomicAdd(pr + lane, static_cast<double>(vals[lane]));
        }
    }
    TK_OSTAMP(5);
    // ---- prev_torques / last_actions are refreshed from the torque / action tiles that are resident (TMA wrote them; no
    //      thread modifies them), by the thread that issued the loads ----
    if (bulk) {
        if (MODE == 0 && tid == 0) {
            bulk_store(p.prev_torques + e0 * kA * kDof, s.tq, kRowB);                // end-of-step refresh (see header)
            bulk_store(p.last_actions + e0 * kA * kDof, s.act, kRowB);
            bulk_commit();
            bulk_wait_read<0>();
        }
    } else {
        if (MODE == 0)
            for (int i = tid; i < nt * kA * kDof; i += kThreads) {
                p.prev_torques[e0 * kA * kDof + i] = s.tq[i];
                p.last_actions[e0 * kA * kDof + i] = s.act[i];
            }
    }
    TK_OSTAMP(6);
#undef TK_OSTAMP
}

// un-stubbed pre_physics_step (anymal.py:229-232 + clamp vec_task.py:327): flat stream over N*2*22 floats
__global__ void __launch_bounds__(256) op3_pre_kernel(const __grid_constant__ KParams p, const float* __restrict__ actions_in,
                                                       float* __restrict__ targets, int clip) {
    const long long total = static_cast<long long>(p.n) * kA * kDof;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float a = actions_in[i];
        if (clip) a = fminf(fmaxf(a, -p.clip_actions), p.clip_actions);
        p.actions[i] = a;
        targets[i] = p.action_scale * a + p.default_dof_pos[i % kDof];
    }
}

// _reward_gripper_contact_hold :351-352: how many envs press the left / right gripper (x force > 0.1), per agent
__global__ void __launch_bounds__(256) op3_grip_count_kernel(const __grid_constant__ KParams p, unsigned long long* __restrict__ counts) {
    int c[kA * 2] = {0, 0, 0, 0};
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < p.n;
         e += static_cast<long long>(gridDim.x) * blockDim.x) {
#pragma unroll
        for (int k = 0; k < kA * 2; ++k)
            c[k] += p.contact[(static_cast<size_t>(e) * p.nb + p.grip[k >> 1][k & 1]) * 3] > 0.1f ? 1 : 0;
    }
#pragma unroll
    for (int k = 0; k < kA * 2; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_xor_sync(0xffffffffu, c[k], o);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < kA * 2; ++k)
            if (c[k]) atomicAdd(counts + k, static_cast<unsigned long long>(c[k]));
    }
}

// Second half of the fused step: add _reward_gripper_contact_hold (:346-358) where the batch-global counts say it is 1.
// Exits at once otherwise (the common case: exactly one env of the whole batch must press each gripper).
__global__ void __launch_bounds__(256) op3_finalize_kernel(const __grid_constant__ KParams p) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const bool h0 = p.grip_counts[0] == 1ull && p.grip_counts[1] == 1ull, h1 = p.grip_counts[2] == 1ull && p.grip_counts[3] == 1ull;
    if (!h0 && !h1) return;
    float d_rew = 0.f, d_ret = 0.f;
    for (long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; n < p.n;
         n += static_cast<long long>(gridDim.x) * blockDim.x) {
        const float* tb = p.root + (n * kActors + 2) * kRoot;
        const float tup = quat_rotate_ref<1>(Q4{tb[3], tb[4], tb[5], tb[6]}, V3{p.up[0], p.up[1], p.up[2]}).z;
        const float rew_up = tup > 0.98f ? 0.0f + 0.1f : 0.0f;
        const float hz = tb[2] - 0.29f;
        const float rew_h = (hz * hz) * -0.001f;
        const float dist_rew = (p.table_pot[n] - p.prev_table_pot[n]) * 5.0f;
        float delta = 0.f;
#pragma unroll
        for (int a = 0; a < kA; ++a) {
            if (a == 0 ? h0 : h1) {
                const float agent = fmaxf(p.pre_rew[n * kA + a] + 1.0f, 0.0f);      // :648-650 with rew_gripper_hold = 1
                const float r = ((agent + dist_rew) + rew_up) + rew_h;              // :542
                delta += r - p.rew[n * kA + a];
                p.rew[n * kA + a] = r;
            }
        }
        d_rew += delta;
        if (p.ep_return) {
            p.ep_return[n] += delta;
            if (p.reset[n] != 0) d_ret += delta;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        d_rew += __shfl_xor_sync(0xffffffffu, d_rew, o);
        d_ret += __shfl_xor_sync(0xffffffffu, d_ret, o);
    }
    if ((threadIdx.x & 31) == 0) {
        double* pr = p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS;
        atomicAdd(pr + 0, static_cast<double>(d_rew));
        atomicAdd(pr + 3, static_cast<double>(d_ret));
    }
}

// reset_idx :440-489 for an explicit id list
__global__ void op3_reset_idx_kernel(const __grid_constant__ KParams p, const long long* __restrict__ ids, int k,
                                     const float* __restrict__ unit_cmd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    const long long ne = ids[i];
    if (ne < 0 || ne >= p.n) return;
    float2* gd = reinterpret_cast<float2*>(p.dof_state) + ne * kA * kDof;
    for (int j = 0; j < kA * kDof; ++j) gd[j] = make_float2(p.default_dof_pos[j % kDof] + 0.0f, 0.0f);     // :456-457
    const float* ir = p.init_root + static_cast<size_t>(ne) * kActors * kRoot;
    for (int c = 0; c < kActors * kRoot; ++c) p.root[ne * kActors * kRoot + c] = ir[c];                   // :468-470
    p.commands[ne * 3 + 0] = rand_float(p.cmd_x[0], p.cmd_x[1], unit_cmd[static_cast<size_t>(i) * 3 + 0]);
    p.commands[ne * 3 + 1] = rand_float(p.cmd_y[0], p.cmd_y[1], unit_cmd[static_cast<size_t>(i) * 3 + 1]);
    p.commands[ne * 3 + 2] = rand_float(p.cmd_yaw[0], p.cmd_yaw[1], unit_cmd[static_cast<size_t>(i) * 3 + 2]);
    for (int a = 0; a < kA; ++a) {
        const float pot = -norm3_ref(V3{p.goal[0] - ir[a * kRoot], p.goal[1] - ir[a * kRoot + 1], 0.0f}) / p.dt;   // :476-479
        p.prev_potentials[ne * kA + a] = pot;
        p.potentials[ne * kA + a] = pot;
        for (int j = 0; j < kDof; ++j) p.last_actions[(ne * kA + a) * kDof + j] = 0.0f;                   // :486
        for (int f = 0; f < p.num_feet; ++f) p.air_time[(ne * kA + a) * p.num_feet + f] = 0.0f;          // :487
    }
    const float tpot = -norm3_ref(V3{p.targets[0] - ir[2 * kRoot], p.targets[1] - ir[2 * kRoot + 1], 0.0f}) / p.dt;   // :481-484
    p.prev_table_pot[ne] = tpot;
    p.table_pot[ne] = tpot;
    p.progress[ne] = 0;                                                             // :488
    p.reset[ne] = 0;                                                                // :489
}

__global__ void op3_philox_kernel(int n, unsigned long long seed, unsigned long long step, long long env_off,
                                  const unsigned long long* __restrict__ epoch, float* uc) {
    step = effective_step(step, epoch);
    const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= n) return;
    Philox4 d = philox_env(seed, step, env_off + e, kStreamCmd);
    uc[e * 3] = u01(d.x); uc[e * 3 + 1] = u01(d.y); uc[e * 3 + 2] = u01(d.z);
}

__global__ void op3_stats_finalize_kernel(double* __restrict__ partials, int rows, double* __restrict__ out, int accumulate) {
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

}  // namespace op3
}  // namespace tk

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace tk;
using namespace tk::op3;

struct TkOp3 {
    TkOp3Params params;
    TkOp3Buffers buf;
    bool bound, ran, counted;
    int tiles, sm_count;
    double* partials;
    unsigned long long* step_dev;
    float* pre_rew;
    long long* dbg_ts;
};

static void fill_kparams(const TkOp3* h, KParams& k) {
    const TkOp3Params& P = h->params;
    const TkOp3Buffers& B = h->buf;
    k.root = B.root_states; k.init_root = B.initial_root_states; k.dof_state = B.dof_state; k.torques = B.torques;
    k.prev_torques = B.prev_torques; k.contact = B.contact_forces; k.commands = B.commands; k.actions = B.actions;
    k.last_actions = B.last_actions; k.air_time = B.feet_air_time; k.last_contacts = B.last_contacts;
    k.potentials = B.potentials; k.prev_potentials = B.prev_potentials; k.table_pot = B.table_potentials;
    k.prev_table_pot = B.prev_table_potentials; k.obs = B.obs_buf; k.rew = B.rew_buf;
    k.reset = reinterpret_cast<long long*>(B.reset_buf); k.progress = reinterpret_cast<long long*>(B.progress_buf);
    k.timeout = B.timeout_buf; k.ep_return = B.episode_return; k.partials = h->partials;
    k.grip_counts = reinterpret_cast<unsigned long long*>(B.grip_counts); k.pre_rew = h->pre_rew;
    k.dbg_ts = h->dbg_ts;
    k.unit_cmd = nullptr; k.seed = 0; k.step = 0; k.step_dev = h->step_dev; k.env_off = 0; k.noise_mode = TK_NOISE_NONE;
    k.n = P.num_envs; k.nb = P.num_bodies_total; k.max_episode_length = P.max_episode_length; k.num_feet = P.num_feet;
    for (int a = 0; a < kA; ++a) {
        for (int f = 0; f < kMaxFeet; ++f) k.feet[a][f] = P.feet_indices[a][f];
        k.grip[a][0] = P.gripper_indices[a][0]; k.grip[a][1] = P.gripper_indices[a][1];
    }
    k.vec_ok = aligned_to(B.root_states, 16) && aligned_to(B.dof_state, 16) && aligned_to(B.torques, 16) &&
               aligned_to(B.prev_torques, 16) && aligned_to(B.actions, 16) && aligned_to(B.last_actions, 16) &&
               aligned_to(B.obs_buf, 16);
    k.obs32 = aligned_to(B.obs_buf, 32);
    k.vec_feet = aligned_to(B.feet_air_time, 16) && aligned_to(B.last_contacts, 4);
    k.contact8 = aligned_to(B.contact_forces, 8);
    k.contact_tile = aligned_to(B.contact_forces, 16) && static_cast<long long>(kTile) * P.num_bodies_total * 12 <= kMaxContactTileBytes;
    k.lin_vel_scale = P.lin_vel_scale; k.ang_vel_scale = P.ang_vel_scale; k.dof_pos_scale = P.dof_pos_scale;
    k.dof_vel_scale = P.dof_vel_scale; k.action_scale = P.action_scale; k.clip_actions = P.clip_actions;
    k.dt = static_cast<float>(P.dt);
    k.k_torque = static_cast<float>(P.rew_torque * P.dt); k.k_up = static_cast<float>(P.rew_up * P.dt);
    k.k_air = static_cast<float>(P.rew_air_time * P.dt); k.k_syns = static_cast<float>(P.rew_syns_hip * P.dt);
    k.k_nofly = static_cast<float>(P.rew_no_fly * P.dt); k.k_stand = static_cast<float>(P.rew_stand * P.dt);
    k.k_rate = static_cast<float>(P.rew_action_rate * P.dt);
    k.cmd_x[0] = static_cast<float>(P.command_x_range[0]); k.cmd_x[1] = static_cast<float>(P.command_x_range[1] - P.command_x_range[0]);
    k.cmd_y[0] = static_cast<float>(P.command_y_range[0]); k.cmd_y[1] = static_cast<float>(P.command_y_range[1] - P.command_y_range[0]);
    k.cmd_yaw[0] = static_cast<float>(P.command_yaw_range[0]);
    k.cmd_yaw[1] = static_cast<float>(P.command_yaw_range[1] - P.command_yaw_range[0]);
    for (int j = 0; j < kDof; ++j) k.default_dof_pos[j] = P.default_dof_pos[j];
    for (int c = 0; c < 4; ++c) k.inv_start_rot[c] = P.inv_start_rot[c];
    for (int c = 0; c < 3; ++c) {
        k.gravity[c] = P.gravity_vec[c]; k.up[c] = P.up_vec[c]; k.targets[c] = P.targets[c]; k.goal[c] = P.goal_pos[c];
    }
}

static int check_bound(const TkOp3* h) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_REQUIRE(h->bound, TK_ERR_UNBOUND, "tk_ma_op3_bind was not called");
    return TK_OK;
}

template <int MODE, int NF>
static cudaError_t launch_tile_nf(const TkOp3* h, const KParams& k, cudaStream_t st) {
    static int opted = 0;    // dynamic shared memory this instantiation has been opted in for
    static int pad = -1;     // TK_OP3_SMEM_PAD (developer knob): extra dynamic shared memory per CTA = fewer resident CTAs per SM
    if (pad < 0) {
        const char* env = knob("TK_OP3_SMEM_PAD");
        pad = env ? atoi(env) : 0;
    }
    const int bytes = static_cast<int>(sizeof(Smem)) + pad + ((MODE != 1 && k.contact_tile) ? kTile * k.nb * 12 : 0);
    if (bytes > opted) {
        cudaError_t e = cudaFuncSetAttribute(op3_tile_kernel<MODE, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e != cudaSuccess) return e;
        opted = bytes;
    }
    KParams kk = k;
    kk.stagger_cycles = 0; kk.stagger_ctas = 0; kk.sm_count = h->sm_count;
    if (MODE == 0) {
        static int slots = -1, cycles = -1;
        if (slots < 0) {
            int occ = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, op3_tile_kernel<MODE, NF>, threads_of<MODE>(), bytes) != cudaSuccess || occ <= 0) {
                cudaGetLastError();
                occ = 6;
            }
            slots = occ;
            const char* env = knob("TK_OP3_STAGGER");
            cycles = env ? atoi(env) : 0;      // measured: no effect at 65536 .. 524288 envs (16-env tiles: many waves)
        }
        kk.stagger_ctas = slots * h->sm_count;
        kk.stagger_cycles = (h->tiles <= 2 * h->sm_count) ? 0 : cycles;      // a thin single wave is only delayed by it
    }
    op3_tile_kernel<MODE, NF><<<h->tiles, threads_of<MODE>(), bytes, st>>>(kk);
    return cudaGetLastError();
}

template <int MODE>
static cudaError_t launch_tile(const TkOp3* h, const KParams& k, cudaStream_t st) {
    return k.num_feet == 4 ? launch_tile_nf<MODE, 4>(h, k, st) : launch_tile_nf<MODE, 0>(h, k, st);
}

static int count_grips(TkOp3* h, const KParams& k, cudaStream_t st) {
    unsigned long long* counts = reinterpret_cast<unsigned long long*>(h->buf.grip_counts);
    TK_CUDA(cudaMemsetAsync(counts, 0, sizeof(unsigned long long) * kA * 2, st));
    int blocks = (k.n + 255) / 256;
    if (blocks > h->sm_count * 8) blocks = h->sm_count * 8;
    op3_grip_count_kernel<<<blocks, 256, 0, st>>>(k, counts);
    TK_CUDA(cudaGetLastError());
    h->counted = true;
    return TK_OK;
}

extern "C" {

int tk_ma_op3_create(const TkOp3Params* params, TkOp3** out) {
    TK_REQUIRE(params && out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(params->num_envs > 0, TK_ERR_INVALID, "num_envs must be positive (got %d)", params->num_envs);
    TK_REQUIRE(params->num_agents == kA && params->num_dof == kDof, TK_ERR_INVALID,
               "this build serves the 2-agent / 22-DOF OP3 of MA_OP3.py (got %d agents, %d DOFs)", params->num_agents,
               params->num_dof);
    TK_REQUIRE(params->num_feet >= 0 && params->num_feet <= kMaxFeet, TK_ERR_INVALID, "bad num_feet %d", params->num_feet);
    TK_REQUIRE(params->num_grippers >= 2 && params->num_grippers <= TK_OP3_MAX_GRIPPERS, TK_ERR_INVALID,
               "num_grippers must be 2..%d (got %d): _reward_gripper_contact_hold reads grippers [0] and [1]",
               TK_OP3_MAX_GRIPPERS, params->num_grippers);
    TK_REQUIRE(params->num_bodies_total > 0, TK_ERR_INVALID, "num_bodies_total must be positive");
    for (int a = 0; a < kA; ++a) {
        for (int f = 0; f < params->num_feet; ++f)
            TK_REQUIRE(params->feet_indices[a][f] >= 0 && params->feet_indices[a][f] < params->num_bodies_total, TK_ERR_INVALID,
                       "feet index out of range");
        for (int g = 0; g < 2; ++g)
            TK_REQUIRE(params->gripper_indices[a][g] >= 0 && params->gripper_indices[a][g] < params->num_bodies_total,
                       TK_ERR_INVALID, "gripper index out of range");
    }
    TK_CUDA(cudaSetDevice(params->device));
    TkOp3* h = new (std::nothrow) TkOp3();
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "out of host memory");
    h->params = *params;
    h->bound = false; h->ran = false; h->counted = false;
    h->tiles = (params->num_envs + kTile - 1) / kTile;
    h->sm_count = 148;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, params->device);
    h->partials = nullptr; h->step_dev = nullptr; h->pre_rew = nullptr; h->dbg_ts = nullptr;
    cudaError_t e = cudaMalloc(&h->partials, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMemset(h->partials, 0, sizeof(double) * TK_STATS * kStatSlots);
    if (e == cudaSuccess) e = cudaMalloc(&h->step_dev, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->step_dev, 0, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&h->pre_rew, sizeof(float) * kA * static_cast<size_t>(params->num_envs));
    if (e != cudaSuccess) {
        cudaFree(h->partials); cudaFree(h->step_dev); cudaFree(h->pre_rew);
        delete h;
        return cuda_fail(e, "tk_ma_op3_create");
    }
    *out = h;
    return TK_OK;
}

int tk_ma_op3_destroy(TkOp3* h) {
    if (!h) return TK_OK;
    cudaSetDevice(h->params.device);
    cudaFree(h->partials); cudaFree(h->step_dev); cudaFree(h->pre_rew);
    delete h;
    return TK_OK;
}

int tk_ma_op3_bind(TkOp3* h, const TkOp3Buffers* b) {
    TK_REQUIRE(h && b, TK_ERR_INVALID, "null argument");
    const void* req[] = {b->root_states, b->initial_root_states, b->dof_state, b->torques, b->prev_torques, b->contact_forces,
                         b->commands, b->actions, b->last_actions, b->feet_air_time, b->potentials, b->prev_potentials,
                         b->table_potentials, b->prev_table_potentials, b->obs_buf, b->rew_buf, b->reset_buf, b->progress_buf};
    const char* names[] = {"root_states", "initial_root_states", "dof_state", "torques", "prev_torques", "contact_forces",
                           "commands", "actions", "last_actions", "feet_air_time", "potentials", "prev_potentials",
                           "table_potentials", "prev_table_potentials", "obs_buf", "rew_buf", "reset_buf", "progress_buf"};
    for (unsigned i = 0; i < sizeof(req) / sizeof(req[0]); ++i) {
        TK_REQUIRE(req[i] != nullptr, TK_ERR_UNBOUND, "buffer %s is NULL", names[i]);
        TK_REQUIRE(aligned_to(req[i], 4), TK_ERR_ALIGN, "buffer %s is not 4-byte aligned", names[i]);
    }
    TK_REQUIRE(b->last_contacts != nullptr, TK_ERR_UNBOUND, "buffer last_contacts is NULL");
    TK_REQUIRE(b->grip_counts != nullptr && aligned_to(b->grip_counts, 8), TK_ERR_UNBOUND, "buffer grip_counts is NULL or misaligned");
    TK_REQUIRE(aligned_to(b->reset_buf, 8) && aligned_to(b->progress_buf, 8), TK_ERR_ALIGN, "int64 buffers must be 8-byte aligned");
    TK_REQUIRE(aligned_to(b->dof_state, 8) && aligned_to(b->rew_buf, 8) && aligned_to(b->prev_torques, 8) &&
               aligned_to(b->last_actions, 8), TK_ERR_ALIGN,
               "dof_state, rew_buf, prev_torques and last_actions are accessed as 8-byte pairs");
    TK_REQUIRE(aligned_to(b->obs_buf, 16), TK_ERR_ALIGN, "obs_buf rows are written as 128-bit stores: 16-byte alignment");
    h->buf = *b;
    h->bound = true;
    return TK_OK;
}

int tk_ma_op3_pre_physics(TkOp3* h, const float* actions_in, float* targets_out, int32_t clip, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(actions_in && targets_out, TK_ERR_INVALID, "null argument");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    const long long total = static_cast<long long>(k.n) * kA * kDof;
    long long blocks = (total + 255) / 256;
    if (blocks > static_cast<long long>(h->sm_count) * 16) blocks = static_cast<long long>(h->sm_count) * 16;
    op3_pre_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(k, actions_in, targets_out, clip);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_ma_op3_grip_counts(TkOp3* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    return count_grips(h, k, static_cast<cudaStream_t>(stream));
}

int tk_ma_op3_compute_observations(TkOp3* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    TK_CUDA(launch_tile<1>(h, k, static_cast<cudaStream_t>(stream)));
    return TK_OK;
}

int tk_ma_op3_compute_reward(TkOp3* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(h->counted, TK_ERR_INVALID, "tk_ma_op3_grip_counts must run before tk_ma_op3_compute_reward");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    TK_CUDA(launch_tile<2>(h, k, static_cast<cudaStream_t>(stream)));
    return TK_OK;
}

int tk_ma_op3_reset_idx(TkOp3* h, const int64_t* env_ids, int32_t kcount, const float* unit_cmd, void* stream) {
    if (int rc = check_bound(h)) return rc;
    if (kcount == 0) return TK_OK;
    TK_REQUIRE(kcount > 0 && env_ids && unit_cmd, TK_ERR_INVALID, "bad reset_idx arguments");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    op3_reset_idx_kernel<<<(kcount + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        k, reinterpret_cast<const long long*>(env_ids), kcount, unit_cmd);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_ma_op3_post_physics(TkOp3* h, const TkOp3Noise* noise, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(noise != nullptr, TK_ERR_INVALID, "null noise descriptor");
    TK_REQUIRE(noise->mode == TK_NOISE_NONE || noise->mode == TK_NOISE_DENSE || noise->mode == TK_NOISE_PHILOX, TK_ERR_INVALID,
               "noise mode %d: MA_OP3 takes TK_NOISE_NONE, TK_NOISE_DENSE (row n = env n) or TK_NOISE_PHILOX", noise->mode);
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    TK_CUDA(cudaMemsetAsync(k.grip_counts, 0, sizeof(unsigned long long) * kA * 2, static_cast<cudaStream_t>(stream)));
    h->counted = true;
    k.noise_mode = noise->mode;
    if (noise->mode == TK_NOISE_DENSE) k.unit_cmd = noise->unit_cmd;
    k.seed = noise->seed; k.step = noise->step; k.env_off = noise->env_id_offset;
    TK_CUDA(launch_tile<0>(h, k, static_cast<cudaStream_t>(stream)));
    h->ran = true;
    return TK_OK;
}

int tk_ma_op3_finalize_reward(TkOp3* h, void* stream) {
    if (int rc = check_bound(h)) return rc;
    TK_REQUIRE(h->ran, TK_ERR_INVALID, "tk_ma_op3_post_physics has not run yet");
    TK_CUDA(cudaSetDevice(h->params.device));
    KParams k; fill_kparams(h, k);
    int blocks = (k.n + 255) / 256;
    if (blocks > h->sm_count * 8) blocks = h->sm_count * 8;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TK_CUDA(cudaLaunchKernelEx(&cfg, op3_finalize_kernel, k));
    return TK_OK;
}

// developer probe, not part of the public header: per-warp globaltimer stamps [tiles, 4, 8] (tools/op3_timeline.py)
int tk_debug_ma_op3_timeline(TkOp3* h, long long* device_buffer) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    h->dbg_ts = device_buffer;
    return TK_OK;
}

int tk_ma_op3_advance_epoch(TkOp3* h, void* stream) {
    TK_REQUIRE(h != nullptr, TK_ERR_INVALID, "null handle");
    TK_CUDA(cudaSetDevice(h->params.device));
    advance_epoch_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(h->step_dev);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_ma_op3_stats(TkOp3* h, double* stats_out, int32_t accumulate, void* stream) {
    TK_REQUIRE(h && stats_out, TK_ERR_INVALID, "null argument");
    TK_REQUIRE(h->ran, TK_ERR_INVALID, "tk_ma_op3_post_physics has not run yet");
    TK_CUDA(cudaSetDevice(h->params.device));
    op3_stats_finalize_kernel<<<1, TK_STATS * 32, 0, static_cast<cudaStream_t>(stream)>>>(h->partials, kStatSlots, stats_out, accumulate);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

int tk_ma_op3_philox_noise(TkOp3* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* unit_cmd, void* stream) {
    TK_REQUIRE(h != nullptr && unit_cmd != nullptr, TK_ERR_INVALID, "null argument");
    TK_CUDA(cudaSetDevice(h->params.device));
    op3_philox_kernel<<<(h->params.num_envs + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        h->params.num_envs, seed, step, env_id_offset, h->step_dev, unit_cmd);
    TK_CUDA(cudaGetLastError());
    return TK_OK;
}

}  // extern "C"

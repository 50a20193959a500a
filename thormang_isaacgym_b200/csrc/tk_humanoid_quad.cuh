// Fused Humanoid / Ant post-physics kernel, four-warp variant: one CTA = 128 threads = one tile of 32 consecutive envs
// (included by tk_humanoid.cu after tk_humanoid_tile.cuh; same KParams, Spec, torso_math, reset_env_values).
//
// The one-warp kernel (tk_humanoid_tile.cuh) keeps the most tiles in flight per SM and wins when the grid is many
// waves deep.  On thin and medium grids (<= ~4 waves: 16384 ... 131072 envs) the launch is bound by the LIFE of a tile,
// not by bandwidth: one warp pulls 109 floats per lane, runs ~1400 dependent instructions and writes 55 STS.64 before
// the store can start (1.5 us of a 4 us tile life, tools/tile_timeline.py).  Here the same tile is worked by four warps:
//
//   warp w, lane e     : a quarter of env e's torso math (13 root floats -> the 12 head columns + the 2 potential
//                        columns): warp 0 potentials + up / heading projections + reward + reset / timeout masks +
//                        per-env scalars + statistics partials, warp 1 the two local velocities, warp 2 yaw + roll,
//                        warp 3 angle_to_target -- four ~150-instruction chains instead of one of ~600 (3 atan2f, 1 sqrt,
//                        7 IEEE divisions), same expressions, same bits;
//   warps 1-3, thread c: obs column 12 + c of ALL 32 envs (96 copy columns for the Humanoid = exactly 96 threads; 48 for
//                        Ant): 32 independent LDS -> scale -> STS triples with immediate offsets, no dependence chain;
//                        one branch-free form ((A*x - B) - C) * R covers unscale / vel / force / sensor / action columns
//                        bit-for-bit (A = 1, B = C = 0 leave x exact; no FMA contraction: -fmad=false);
//   resets (rare)      : the q-th flagged env of the tile belongs to warp q & 3 (lane = DOF), the first four before the
//                        tile lands.
//
// Inputs and the obs image live in SEPARATE shared-memory buffers (13.6 + 13.75 KB): nothing has to be staged in
// registers, so there is no "everything in registers" barrier and 72 registers per thread suffice => 7 CTAs / SM
// (shared-memory limited), 98 KB of loads in flight per SM -- still 2x what Little's law needs at 6.5 TB/s.
// Bulk (TMA) copies both ways, PDL and the L2 prefetch ahead of griddepcontrol.wait are those of the one-warp kernel.
#pragma once

namespace tk {
namespace humanoid {

constexpr int kQuadThreads = 128;
#ifndef TK_HUMANOID_QUAD_MIN_CTAS
#define TK_HUMANOID_QUAD_MIN_CTAS 7
#endif

template <class S, bool FUSED>
__global__ void __launch_bounds__(kQuadThreads, TK_HUMANOID_QUAD_MIN_CTAS) humanoid_quad_kernel(const __grid_constant__ KParams p) {
    using L = S;
    using WT = WarpTile<S>;
    constexpr int D = S::D;
    constexpr int kSensor = S::kSensor;
    constexpr int OBS = L::kObs;
    constexpr int kCopy = (S::kHasForce ? 4 : 3) * D + kSensor;      // copy columns 12 .. 12 + kCopy
    static_assert(kCopy <= 96, "three copy warps");
    __shared__ __align__(128) float in[WT::kIn];
    __shared__ __align__(128) float out[WT::kOut];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned s_reset_bits;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int env0 = blockIdx.x * kTile;
    const int nt = min(kTile, p.n - env0);
    const int n = env0 + lane;
    const bool valid = lane < nt;
    const bool bulk = (p.vec_ok != 0) && (nt == kTile);
    long long* ts = p.dbg_ts ? p.dbg_ts + static_cast<size_t>(blockIdx.x) * 8 : nullptr;
#define TK_QSTAMP(i) do { if (ts && tid == 0) { long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ts[i] = g; } } while (0)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const float* g_dof = p.dof_state + static_cast<size_t>(env0) * D * 2;
    const float* g_force = S::kHasForce ? p.dof_force + static_cast<size_t>(env0) * D : nullptr;
    const float* g_act = p.actions + static_cast<size_t>(env0) * D;
    const float* g_sens = p.sensors + static_cast<size_t>(env0) * kSensor;
    const float* g_root = p.root + static_cast<size_t>(env0) * kRoot;

    if (bulk && tid == 0) {
        mbar_init(&bar, 1);
        fence_proxy_async_smem();          // the init (generic proxy) must be visible to the bulk-copy engine (async proxy)
    }
    // L2 prefetch of the tile ahead of griddepcontrol.wait (a hint: consumes nothing, stream order intact -- see the
    // one-warp kernel).  Thin grids only (host decides): there the CTA is resident while the previous kernel drains.
    if (bulk && p.prefetch && warp == 0) {
        if (lane == 0) bulk_prefetch_l2(g_dof, kTile * D * 2 * 4);
        else if (S::kHasForce && lane == 1) bulk_prefetch_l2(g_force, kTile * D * 4);
        else if (lane == 2) bulk_prefetch_l2(g_act, kTile * D * 4);
        else if (lane == 3) bulk_prefetch_l2(g_root, kTile * kRoot * 4);
        else if (lane == 4) bulk_prefetch_l2(g_sens, kTile * kSensor * 4);
        else if (lane == 5) prefetch_l2(p.pot + env0);
        else if (FUSED && lane == 6) prefetch_l2(p.progress + env0);
        else if (FUSED && lane == 7) prefetch_l2(p.progress + env0 + 16);
        else if (FUSED && lane == 8) prefetch_l2(p.reset + env0);
        else if (FUSED && lane == 9) prefetch_l2(p.reset + env0 + 16);
        else if (FUSED && lane == 10 && p.ep_return) prefetch_l2(p.ep_return + env0);
    }

    // ---- per-thread constants of the copy columns (parameter-bank reads: allowed before the dependency resolves)
    const int c = tid - 32;                                   // copy column index (warps 1-3)
    int src = 0, sstride = 0;
    float cA = 1.0f, cB = 0.0f, cC = 0.0f, cR = 1.0f, cDen = 1.0f;
    bool is_pos = false;
    if (c >= 0 && c < kCopy) {
        int j = c;
        if (j < D) {                                          // dof_pos_scaled (humanoid.py:406 | ant.py:389)
            src = WT::oDof + 2 * j; sstride = 2 * D; is_pos = true;
            cA = 2.0f; cB = p.hi[j]; cC = p.lo[j]; cR = p.rinv[j]; cDen = p.hi[j] - p.lo[j];
        } else if ((j -= D) < D) {                            // dof_vel * dof_vel_scale
            src = WT::oDof + 2 * j + 1; sstride = 2 * D; cR = p.dof_vel_scale;
        } else if (S::kHasForce && (j -= D) < D) {            // dof_force * contact_force_scale
            src = WT::oForce + j; sstride = D; cR = p.contact_scale;
        } else if ((j -= D) < kSensor) {                      // sensors * contact_force_scale
            src = WT::oSens + j; sstride = kSensor; cR = p.contact_scale;
        } else {                                              // actions (raw copy)
            j -= kSensor;
            src = WT::oAct + j; sstride = D;
        }
    }
    const float warm = p.dt + p.tgt[0] + p.isr[3] + p.b0[0] + p.b1[2] + p.termination_height;   // constant-bank lines of the torso math
    if (bulk && warp >= 2) warm_ahead_of_wait<D, OBS>(p, env0, tid - 64, FUSED);
    const unsigned long long step_now = FUSED ? step_ahead_of_wait(p) : 0ull;

    asm volatile("griddepcontrol.wait;" ::: "memory");
    TK_QSTAMP(0);
    if (warm == -1.2345e-30f && p.dbg_ts) p.dbg_ts[0] = 0;
    // first-wave stagger (see the one-warp kernel): resident slot s of the first wave issues its loads s * stagger_cycles late,
    // so that the CTAs of an SM are at different points of load / math / store instead of in lockstep
    if (p.stagger_cycles > 0 && blockIdx.x < static_cast<unsigned>(p.stagger_ctas) && warp == 0) {
        const long long until = clock64() + static_cast<long long>(max(0, static_cast<int>(blockIdx.x / p.sm_count) - p.stagger_free)) * p.stagger_cycles;
        while (clock64() < until) __nanosleep(32);
    }

    // ---- per-env scalars (coalesced, warp 0), requested ahead of the bulk copies (issuing one holds its lane ~0.25 us and the
    // scalars were the last thing a tile waited for).  The reset set of the tile is published
    // through shared memory: all four warps must agree on it, and warp 0 overwrites reset_buf at the end of the step.
    long long prog = 0;
    bool do_reset = false;
    float pot_old = 0.0f, ret_old = 0.0f;
    if (warp == 0) {
        if (valid) {
            pot_old = p.pot[n];
            if (FUSED) {
                prog = p.progress[n] + 1;                         // humanoid.py:288
                do_reset = p.reset[n] != 0 && !p.dbg_reset;                       // :291
                if (p.ep_return) ret_old = p.ep_return[n];
                if (p.randomize) p.randomize[n] += 1;             // :289
            }
        }
    }
    if (bulk) {
        if (warp == 0) {
            if (lane == 0) mbar_expect_tx(&bar, WT::kIn * 4);
            __syncwarp();
            if (lane == 0) bulk_load(in + WT::oDof, g_dof, kTile * D * 2 * 4, &bar);
            else if (S::kHasForce && lane == 1) bulk_load(in + WT::oForce, g_force, kTile * D * 4, &bar);
            else if (lane == 2) bulk_load(in + WT::oAct, g_act, kTile * D * 4, &bar);
            else if (lane == 3) bulk_load(in + WT::oRoot, g_root, kTile * kRoot * 4, &bar);
            else if (lane == 4) bulk_load(in + WT::oSens, g_sens, kTile * kSensor * 4, &bar);
        }
    } else {
        auto coop = [&](float* dst, const float* s, int cnt) { for (int i = tid; i < cnt; i += kQuadThreads) dst[i] = s[i]; };
        coop(in + WT::oDof, g_dof, nt * D * 2);
        if (S::kHasForce) coop(in + WT::oForce, g_force, nt * D);
        coop(in + WT::oAct, g_act, nt * D);
        coop(in + WT::oSens, g_sens, nt * kSensor);
        coop(in + WT::oRoot, g_root, nt * kRoot);
    }

    if (warp == 0 && FUSED) {
        const unsigned bits = __ballot_sync(0xffffffffu, do_reset);
        if (lane == 0) s_reset_bits = bits;
    }
    TK_QSTAMP(1);
    __syncthreads();   // s_reset_bits, the mbarrier init and (generic path) the cooperative copies are visible to every warp

    // ---- reset prologue (rare): reset_idx :253-279, the q-th flagged env of the tile is done by warp q & 3, lane = DOF;
    // its first flagged env per warp before the tile lands (the new state does not depend on the staged tile)
    const unsigned reset_bits = FUSED ? s_reset_bits : 0u;
    int rank0 = 0, pre_e = -1;
    float2 pre_pv = make_float2(0.0f, 0.0f);
    float pre_root = 0.0f;
    if (FUSED && reset_bits) {
        rank0 = (p.noise_mode == TK_NOISE_COMPACT) ? p.tile_rank_base[blockIdx.x] : 0;
        unsigned bits = reset_bits;
        for (int q = 0; q < 4 && bits; ++q) {
            const int e = __ffs(bits) - 1;
            bits &= bits - 1;
            if (q == warp) {
                pre_e = e;
                reset_env_values<D>(p, env0 + e, rank0 + q, lane, &pre_pv, &pre_root, step_now);
            }
        }
    }
    TK_QSTAMP(2);
    if (bulk) mbar_wait(&bar, 0);          // bulk loads landed (every thread observes the phase: data visible to it)
    TK_QSTAMP(3);

    if (FUSED && reset_bits) {             // patch the staged rows, then one CTA barrier (reset_bits is CTA-uniform)
        if (pre_e >= 0) {
            if (lane < D) reinterpret_cast<float2*>(in + WT::oDof)[pre_e * D + lane] = pre_pv;
            if (lane < kRoot) in[WT::oRoot + pre_e * kRoot + lane] = pre_root;
        }
        unsigned bits = reset_bits;
        for (int q = 0; bits; ++q) {
            const int e = __ffs(bits) - 1;
            bits &= bits - 1;
            if (q >= 4 && (q & 3) == warp) {   // more than four resets in one tile: the rest after the wait
                float2 pv;
                float rv;
                reset_env_values<D>(p, env0 + e, rank0 + q, lane, &pv, &rv, step_now);
                if (lane < D) reinterpret_cast<float2*>(in + WT::oDof)[e * D + lane] = pv;
                if (lane < kRoot) in[WT::oRoot + e * kRoot + lane] = rv;
            }
        }
        __syncthreads();
    }

    float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_up = 0.f, st_head = 0.f;
    bool out_reset = false, out_timeout = false;
    float r[kRoot];
    {
        const float* rr = in + WT::oRoot + (valid ? lane : 0) * kRoot;   // stride 13 words: bank-conflict free
#pragma unroll
        for (int k = 0; k < kRoot; ++k) r[k] = rr[k];
    }
    float* row = out + lane * OBS;
    if (warp != 0) {
        // ---- copy columns: thread c moves column 12 + c of every env of the tile
        if (c < kCopy) {
            const float* s = in + src;
            float* d = out + L::kPos + c;
            if (nt == kTile) {
#pragma unroll
                for (int e = 0; e < kTile; ++e) {
                    const float x = s[e * sstride];
                    const float num = (cA * x - cB) - cC;
                    // Ant counts `dof_pos_scaled > 0.99` in its reward (ant.py:347): the true IEEE quotient keeps that
                    // threshold bit-exact; Humanoid's live reward never reads these columns -> reciprocal multiply
                    d[e * OBS] = (S::kVariant == kVarAnt && is_pos) ? num / cDen : num * cR;
                }
            } else {
                for (int e = 0; e < nt; ++e) {
                    const float x = s[e * sstride];
                    const float num = (cA * x - cB) - cC;
                    d[e * OBS] = (S::kVariant == kVarAnt && is_pos) ? num / cDen : num * cR;
                }
            }
        }
        // ---- this warp's share of env `lane`'s torso math (tk_humanoid.cu: torso_part_*)
        if (valid) {
            if (warp == 1) {
                float v[6];
                torso_part_b<S::kVariant>(r, p, v);
#pragma unroll
                for (int k = 0; k < 6; ++k) row[1 + k] = v[k];
            } else if (warp == 2) {
                float yaw, roll;
                torso_part_c<S::kVariant>(r, p, &yaw, &roll);
                row[7] = yaw; row[8] = roll;
            } else {
                row[9] = torso_part_d<S::kVariant>(r, p);
            }
        }
    } else if (valid) {
        // ---- warp 0: potentials, projections, reward and masks of env `lane`
        const HeadA h = torso_part_a<S::kVariant>(r, p);
        // after a reset potentials = prev_potentials = -|targets - initial_pos| / dt (:273-276); the root row now IS the
        // initial row, so that value equals the freshly computed potential
        const float prev_new = (FUSED && do_reset) ? h.pot : pot_old;              // :394
        row[0] = h.height; row[10] = h.up_proj; row[11] = h.heading_proj;
        if (S::kVariant == kVarHumanoid) {
            row[L::kPot] = h.pot / 60000.0f;                                       // :414
            row[L::kPot + 1] = prev_new / 60000.0f;
        }
        p.pot[n] = h.pot;
        p.prev_pot[n] = prev_new;
        if (p.up_vec) { p.up_vec[n * 3 + 0] = h.up.x; p.up_vec[n * 3 + 1] = h.up.y; p.up_vec[n * 3 + 2] = h.up.z; }
        if (p.heading_vec) {
            p.heading_vec[n * 3 + 0] = h.heading.x; p.heading_vec[n * 3 + 1] = h.heading.y; p.heading_vec[n * 3 + 2] = h.heading.z;
        }
        if (FUSED) {
            if (do_reset) prog = 0;                                                // :278 (and reset_buf -> 0, :279)
            const float up_reward = (h.up_proj > 0.93f) ? (0.0f + p.up_weight) : 0.0f;   // :349-350 | ant.py:341-342
            float total;
            if (S::kVariant == kVarAnt) {
                // compute_ant_reward, ant.py:337-354: every term live; the columns it reads are recomputed from the staged
                // inputs with the very expressions of the copy threads (same bits), so no barrier is needed here
                const float hp = h.heading_proj;
                const float heading_reward = (hp > 0.8f) ? (1.0f * p.heading_weight) : (p.heading_weight * hp / 0.8f);
                float actions_cost = 0.0f, electricity_cost = 0.0f;
                int at_limit = 0;
                const float2* dof2 = reinterpret_cast<const float2*>(in + WT::oDof) + lane * D;
                const float* ac = in + WT::oAct + lane * D;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    const float2 pv = dof2[j];
                    const float a = ac[j];
                    actions_cost += a * a;                                         // :345
                    electricity_cost += fabsf(a * (pv.y * p.dof_vel_scale));       // :346 (obs[:, 20:28] = scaled dof_vel)
                    at_limit += (unscale_ref(pv.x, p.lo[j], p.hi[j]) > 0.99f) ? 1 : 0;   // :347
                }
                const float progress_reward = h.pot - prev_new;                    // :351
                total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                        p.energy_cost_scale * electricity_cost -
                        static_cast<float>(at_limit) * p.joints_cost_scale;        // :353-354
            } else {
                total = 2.0f + up_reward;                                          // humanoid.py:362,365
            }
            const bool fallen = h.height < p.termination_height;
            if (fallen) total = 1.0f * p.death_cost;                               // :370 | ant.py:357
            const bool finished = static_cast<float>(prog) >= p.max_episode_length - 1.0f;   // :374
            out_reset = fallen || finished;                                        // :373-374 (reset_buf is 0 here)
            out_timeout = finished && out_reset;                                   // vec_task.py:345
            p.rew[n] = total;
            p.reset[n] = out_reset ? 1 : 0;
            p.progress[n] = prog;
            if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
            st_rew = total; st_up = h.up_proj; st_head = h.heading_proj;
            if (p.ep_return) {
                const float ret = (do_reset ? 0.0f : ret_old) + total;
                if (out_reset) st_ret = ret;
                p.ep_return[n] = ret;
            }
            if (out_reset) st_len = static_cast<float>(prog);
        }
    }
    TK_QSTAMP(4);
    fence_proxy_async_smem();   // obs image was written through the generic proxy, the bulk engine reads it
    __syncthreads();
    TK_QSTAMP(5);

    // ---- obs tile -> global
    float* dst = p.obs + static_cast<size_t>(env0) * OBS;
    if (bulk && p.coop_store) {
        // developer knob (TK_HUMANOID_COOP_STORE): the obs tile leaves as 128-bit stores of all 128 threads instead of one bulk store
        const float4* o4 = reinterpret_cast<const float4*>(out);
        for (int i = tid; i < WT::kOut / 4; i += kQuadThreads) stg_stream4(dst + 4 * i, o4[i]);
        if (p.obs2) for (int i = tid; i < WT::kOut / 4; i += kQuadThreads) stg_stream4(p.obs2 + static_cast<size_t>(env0) * OBS + 4 * i, o4[i]);
    } else if (bulk) {
        if (tid == 0) {
            bulk_store(dst, out, WT::kOut * 4);
            // second destination of the same tile (tk_humanoid_set_obs_copy): e.g. slot t of a [horizon, N, obs] rollout buffer
            if (p.obs2) bulk_store(p.obs2 + static_cast<size_t>(env0) * OBS, out, WT::kOut * 4);
            bulk_commit();
        }
    } else {
        for (int i = tid; i < nt * OBS; i += kQuadThreads) dst[i] = out[i];
        if (p.obs2) for (int i = tid; i < nt * OBS; i += kQuadThreads) p.obs2[static_cast<size_t>(env0) * OBS + i] = out[i];
    }

    // ---- statistics partials (warp 0 holds the per-env values)
    if (FUSED && warp == 0) {
        const unsigned m_reset = __ballot_sync(0xffffffffu, out_reset);
        const unsigned m_to = __ballot_sync(0xffffffffu, out_timeout);
        const unsigned m_valid = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, off);
            st_ret += __shfl_xor_sync(0xffffffffu, st_ret, off);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, off);
            st_up += __shfl_xor_sync(0xffffffffu, st_up, off);
            st_head += __shfl_xor_sync(0xffffffffu, st_head, off);
        }
        if (lane < TK_STATS) {
            float v = st_rew;
            v = (lane == 1) ? static_cast<float>(__popc(m_reset)) : v;
            v = (lane == 2) ? static_cast<float>(__popc(m_to)) : v;
            v = (lane == 3) ? st_ret : v;
            v = (lane == 4) ? st_len : v;
            v = (lane == 5) ? st_up : v;
            v = (lane == 6) ? st_head : v;
            v = (lane == 7) ? static_cast<float>(__popc(m_valid)) : v;
            atomicAdd(p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS + lane, static_cast<double>(v));
        }
    }
    if (bulk && !p.coop_store && tid == 0) bulk_wait_read<0>();   // shared memory must stay alive until the engine has read it
    TK_QSTAMP(6);
#undef TK_QSTAMP
}

}  // namespace humanoid
}  // namespace tk

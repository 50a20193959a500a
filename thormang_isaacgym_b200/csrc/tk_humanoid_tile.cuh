// Fused Humanoid post-physics tile kernel (included by tk_humanoid.cu; needs KParams, Layout, torso_math).
//
// One CTA of T threads owns T consecutive envs.  Its slice of every AoS array is one contiguous,
// 16-byte aligned byte range, so the whole tile moves with the bulk-copy (TMA) engine:
//
//   thread 0:  mbarrier.expect_tx + 5 x cp.async.bulk global->shared   (root, dof_state, dof_force, actions, sensors)
//   all:       coalesced per-env scalars (progress, reset, potentials) while the bulk loads fly
//   all:       wait(mbarrier)  ->  scaled copies shared->shared into the [T,110] obs image (96 of 110 columns)
//   (rare)     reset prologue patches the staged rows and writes the new state to global memory
//   all:       torso math, one env per thread (root rows at stride 13 words: bank-conflict free)
//   thread 0:  fence.proxy.async + ONE cp.async.bulk shared->global for the finished obs tile
//
// No thread ever issues a global load/store for the 876 B/env of bulk data; registers only hold
// the env being computed.  Shared memory per env: 436 B staged input + 440 B obs image.
#pragma once

namespace tk {
namespace humanoid {

template <int D, int T>
struct TileSmem {
    static constexpr int kObs = Layout<D>::kObs;
    alignas(128) float obs[T * kObs];
    alignas(16) float root[T * kRoot];
    alignas(16) float dof[T * D * 2];
    alignas(16) float force[T * D];
    alignas(16) float act[T * D];
    alignas(16) float sens[T * kSensor];
    alignas(8) uint64_t bar;
    float lo[D], hi[D];
    unsigned ballots[(T + 31) / 32];
    float wsum[(T + 31) / 32][TK_STATS];
};

template <int T>
__device__ __forceinline__ void coop_copy(float* dst, const float* src, int cnt, int tid) {
    for (int i = tid; i < cnt; i += T) dst[i] = src[i];
}

template <int D, int T, bool FUSED>
__global__ void __launch_bounds__(T) humanoid_tile_kernel(const __grid_constant__ KParams p) {
    using L = Layout<D>;
    constexpr int OBS = L::kObs;
    constexpr int W = (T + 31) / 32;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    TileSmem<D, T>& s = *reinterpret_cast<TileSmem<D, T>*>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int env0 = blockIdx.x * T;
    const int nt = min(T, p.n - env0);
    const int n = env0 + tid;
    const bool valid = tid < nt;
    const bool bulk = (p.vec_ok != 0) && (nt == T);   // full, 16-byte aligned tile -> TMA both ways

    const float* g_root = p.root + static_cast<size_t>(env0) * kRoot;
    const float* g_dof = p.dof_state + static_cast<size_t>(env0) * D * 2;
    const float* g_force = p.dof_force + static_cast<size_t>(env0) * D;
    const float* g_act = p.actions + static_cast<size_t>(env0) * D;
    const float* g_sens = p.sensors + static_cast<size_t>(env0) * kSensor;

    if (bulk) {
        if (tid == 0) {
            mbar_init(&s.bar, 1);
            mbar_fence_init();
            constexpr uint32_t kBytes = T * (kRoot + 4 * D + kSensor) * 4;
            mbar_expect_tx(&s.bar, kBytes);
            bulk_load(s.dof, g_dof, T * D * 2 * 4, &s.bar);
            bulk_load(s.force, g_force, T * D * 4, &s.bar);
            bulk_load(s.act, g_act, T * D * 4, &s.bar);
            bulk_load(s.sens, g_sens, T * kSensor * 4, &s.bar);
            bulk_load(s.root, g_root, T * kRoot * 4, &s.bar);
        }
    } else {
        coop_copy<T>(s.dof, g_dof, nt * D * 2, tid);
        coop_copy<T>(s.force, g_force, nt * D, tid);
        coop_copy<T>(s.act, g_act, nt * D, tid);
        coop_copy<T>(s.sens, g_sens, nt * kSensor, tid);
        coop_copy<T>(s.root, g_root, nt * kRoot, tid);
    }
    if (tid < D) { s.lo[tid] = p.lo[tid]; s.hi[tid] = p.hi[tid]; }

    // ---- per-env scalars (coalesced), overlapped with the bulk loads
    long long prog = 0;
    bool do_reset = false;
    float pot_old = 0.0f;
    if (valid) {
        pot_old = p.pot[n];
        if (FUSED) {
            prog = p.progress[n] + 1;                              // humanoid.py:288
            do_reset = p.reset[n] != 0;                            // :291
            if (p.randomize) p.randomize[n] += 1;                  // :289
        }
    }
    if (FUSED) {
        unsigned b = __ballot_sync(0xffffffffu, do_reset);
        if (lane == 0) s.ballots[warp] = b;
    }
    __syncthreads();                       // mbarrier init, lo/hi, ballots, cooperative copies visible
    if (bulk) mbar_wait(&s.bar, 0);        // bulk loads landed

    // ---- scaled copies shared -> obs image
    {
        const int cnt = nt * D;
        const float dvs = p.dof_vel_scale, cfs = p.contact_scale;
        const float2* dof2 = reinterpret_cast<const float2*>(s.dof);
#pragma unroll 4
        for (int e = tid; e < cnt; e += T) {
            const int nn = e / D, j = e - nn * D;
            const float2 pv = dof2[e];
            float* o = &s.obs[nn * OBS + j];
            o[L::kPos] = unscale_ref(pv.x, s.lo[j], s.hi[j]);      // :406
            o[L::kVel] = pv.y * dvs;                               // :413
            o[L::kForce] = s.force[e] * cfs;
            o[L::kAct] = s.act[e];
        }
        const int cs = nt * kSensor;
#pragma unroll 2
        for (int e = tid; e < cs; e += T) {
            const int nn = e / kSensor, k = e - nn * kSensor;
            s.obs[nn * OBS + L::kSens + k] = s.sens[e] * cfs;      // :414
        }
    }

    // ---- reset prologue (rare): reset_idx :253-279 for this tile's flagged envs, one warp per env
    if (FUSED) {
        __syncthreads();
        const unsigned mine = s.ballots[warp];
        if (mine) {
            int rank = (p.noise_mode == TK_NOISE_COMPACT) ? p.tile_rank_base[blockIdx.x] : 0;
            for (int w = 0; w < warp; ++w) rank += __popc(s.ballots[w]);
            unsigned bits = mine;
            while (bits) {
                const int le = __ffs(bits) - 1;
                bits &= bits - 1;
                const int e = warp * 32 + le;
                const int ne = env0 + e;
                const long long row = (p.noise_mode == TK_NOISE_COMPACT) ? rank : ne;
                ++rank;
                for (int j = lane; j < D; j += 32) {
                    float ua = 0.5f, ub = 0.5f;
                    if (p.noise_mode == TK_NOISE_PHILOX) {
                        Philox4 r = philox_env(p.seed, p.step, p.env_off + ne, static_cast<uint32_t>(j >> 1));
                        ua = u01((j & 1) ? r.z : r.x);
                        ub = u01((j & 1) ? r.w : r.y);
                    } else if (p.noise_mode != TK_NOISE_NONE && row < p.noise_rows) {
                        ua = p.unit_a[row * D + j];
                        ub = p.unit_b[row * D + j];
                    }
                    const float positions = 0.4f * ua + (-0.2f);                   // :258 torch_rand_float(-0.2, 0.2)
                    const float velocities = 0.2f * ub + (-0.1f);                  // :259 torch_rand_float(-0.1, 0.1)
                    float np_ = p.init_dof_pos[static_cast<size_t>(ne) * D + j] + positions;
                    np_ = fmaxf(fminf(np_, s.hi[j]), s.lo[j]);                     // :261 tensor_clamp
                    reinterpret_cast<float2*>(p.dof_state)[static_cast<size_t>(ne) * D + j] = make_float2(np_, velocities);
                    s.obs[e * OBS + L::kPos + j] = unscale_ref(np_, s.lo[j], s.hi[j]);
                    s.obs[e * OBS + L::kVel + j] = velocities * p.dof_vel_scale;
                }
                if (lane < kRoot) {                                                // :265-267
                    const float v = p.init_root[static_cast<size_t>(ne) * kRoot + lane];
                    p.root[static_cast<size_t>(ne) * kRoot + lane] = v;
                    s.root[e * kRoot + lane] = v;
                }
            }
        }
        __syncthreads();
    }

    // ---- torso math, one env per thread
    float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_up = 0.f, st_head = 0.f;
    bool out_reset = false, out_timeout = false;
    if (valid) {
        float r[kRoot];
#pragma unroll
        for (int c = 0; c < kRoot; ++c) r[c] = s.root[tid * kRoot + c];
        const Head h = torso_math(r, p);
        // after a reset potentials = prev_potentials = -|targets - initial_pos| / dt (:273-276); the root row now IS the
        // initial row, so that value equals the freshly computed potential
        const float prev_new = (FUSED && do_reset) ? h.pot : pot_old;              // :394
        float* o = &s.obs[tid * OBS];
#pragma unroll
        for (int c = 0; c < 12; ++c) o[c] = h.c[c];
        o[L::kPot] = h.pot / 60000.0f;                                             // :414
        o[L::kPot + 1] = prev_new / 60000.0f;
        p.pot[n] = h.pot;
        p.prev_pot[n] = prev_new;
        if (p.up_vec) { p.up_vec[n * 3 + 0] = h.up.x; p.up_vec[n * 3 + 1] = h.up.y; p.up_vec[n * 3 + 2] = h.up.z; }
        if (p.heading_vec) {
            p.heading_vec[n * 3 + 0] = h.heading.x; p.heading_vec[n * 3 + 1] = h.heading.y; p.heading_vec[n * 3 + 2] = h.heading.z;
        }
        if (FUSED) {
            if (do_reset) prog = 0;                                                // :278 (and reset_buf -> 0, :279)
            const float up_reward = (h.c[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;   // :349-350
            float total = 2.0f + up_reward;                                        // :362,365
            const bool fallen = h.c[0] < p.termination_height;
            if (fallen) total = 1.0f * p.death_cost;                               // :370
            const bool finished = static_cast<float>(prog) >= p.max_episode_length - 1.0f;   // :374
            out_reset = fallen || finished;                                        // :373-374 (reset_buf is 0 here)
            out_timeout = finished && out_reset;                                   // vec_task.py:345
            p.rew[n] = total;
            p.reset[n] = out_reset ? 1 : 0;
            p.progress[n] = prog;
            if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
            st_rew = total; st_up = h.c[10]; st_head = h.c[11];
            if (p.ep_return) {
                const float ret = (do_reset ? 0.0f : p.ep_return[n]) + total;
                if (out_reset) st_ret = ret;
                p.ep_return[n] = ret;
            }
            if (out_reset) st_len = static_cast<float>(prog);
        }
    }

    // ---- statistics partials (fixed reduction order -> deterministic)
    if (FUSED) {
        const unsigned m_reset = __ballot_sync(0xffffffffu, out_reset);
        const unsigned m_to = __ballot_sync(0xffffffffu, out_timeout);
        const unsigned m_valid = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            st_rew += __shfl_xor_sync(0xffffffffu, st_rew, o);
            st_ret += __shfl_xor_sync(0xffffffffu, st_ret, o);
            st_len += __shfl_xor_sync(0xffffffffu, st_len, o);
            st_up += __shfl_xor_sync(0xffffffffu, st_up, o);
            st_head += __shfl_xor_sync(0xffffffffu, st_head, o);
        }
        if (lane == 0) {
            float* w = s.wsum[warp];
            w[0] = st_rew; w[1] = static_cast<float>(__popc(m_reset)); w[2] = static_cast<float>(__popc(m_to));
            w[3] = st_ret; w[4] = st_len; w[5] = st_up; w[6] = st_head; w[7] = static_cast<float>(__popc(m_valid));
        }
    }
    fence_proxy_async_smem();   // obs image was written through the generic proxy, the bulk engine reads it
    __syncthreads();

    if (FUSED && tid < TK_STATS) {
        double acc = 0.0;
#pragma unroll
        for (int w = 0; w < W; ++w) acc += static_cast<double>(s.wsum[w][tid]);
        p.partials[static_cast<size_t>(blockIdx.x) * TK_STATS + tid] += acc;   // running sum since the last tk_humanoid_stats
    }

    // ---- obs tile -> global
    float* dst = p.obs + static_cast<size_t>(env0) * OBS;
    if (bulk) {
        if (tid == 0) {
            bulk_store(dst, s.obs, T * OBS * 4);
            bulk_commit();
            bulk_wait_read<0>();   // shared memory must stay alive until the engine has read it
        }
    } else {
        coop_copy<T>(dst, s.obs, nt * OBS, tid);
    }
}

}  // namespace humanoid
}  // namespace tk

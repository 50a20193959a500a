// Fused Humanoid post-physics kernel: one warp = one CTA = one tile of 32 consecutive envs
// (included by tk_humanoid.cu; needs KParams, Spec, torso_math, unscale_fast).  The same kernel, instantiated with
// AntSpec, is the Ant post-physics step (ant.py:279-287 with D = 8, 60 observations, every reward term live).
//
// The slice of every AoS array that belongs to 32 consecutive envs is one contiguous, 16-byte
// aligned byte range, so the whole tile moves with the bulk-copy (TMA) engine, and the SAME
// 14 KB of shared memory holds first the staged inputs (32 x 109 floats) and then the image of
// the 32 obs rows (32 x 110 floats):
//
//   all:     preamble (parameters, addresses, mbarrier.init) may overlap the previous kernel's drain (PDL), then
//            griddepcontrol.wait -- nothing of global memory is touched before the previous kernel has completed
//   lanes 0-4: mbarrier.expect_tx + one cp.async.bulk global->shared each (dof_state, dof_force, actions, root, sensors)
//   all:     coalesced per-env scalars (progress, reset, potentials, episode return) while the bulk loads fly
//   (rare)   reset_idx for flagged envs, lane = DOF: computed and written to global memory BEFORE the wait
//   all:     wait(mbarrier); patch the staged rows of the envs that were reset
//   lane e:  reads ITS env's 109 input floats into registers (immediate offsets, LDS.64/.128 where aligned)
//   __syncwarp  -- every input of the tile is now in registers, the buffer may be overwritten in place
//   lane e:  scaled copies + torso math in registers, writes ITS obs row with 55 conflict-free STS.64
//   lane 0:  fence.proxy.async + ONE cp.async.bulk shared->global for the finished [32,110] tile
//
// No thread issues a global load/store for the 876 B/env of bulk data and there is no index
// arithmetic per element: every shared-memory access is [lane base + immediate].  14 KB of shared
// memory and 160 registers per thread => 12 resident single-warp CTAs per SM (register-limited), each
// at a different point of load / math / store: 168 KB of loads in flight per SM, far more than
// the ~22 KB Little's law asks for at 6.5 TB/s (the kernel is a pure stream: 924 B/env-step, ~0.3 flop/B).
// Alternatives that were built and measured slower (three warps per tile, persistent CTAs with a
// double-buffered TMA prefetch) are recorded in DESIGN.md 3.1.
#pragma once

namespace tk {
namespace humanoid {

constexpr int kTile = 32;   // envs per CTA == one warp

template <class S>
struct WarpTile {
    static constexpr int T = kTile;
    static constexpr int D = S::D;
    static constexpr int OBS = S::kObs;
    // float offsets of the staged inputs inside the tile buffer (each a multiple of 4 floats = 16 bytes)
    static constexpr int oDof = 0;
    static constexpr int oForce = oDof + T * 2 * D;
    static constexpr int oAct = oForce + (S::kHasForce ? T * D : 0);
    static constexpr int oSens = oAct + T * D;
    static constexpr int oRoot = oSens + T * S::kSensor;
    static constexpr int kIn = oRoot + T * kRoot;
    static constexpr int kOut = T * OBS;                       // the obs image of the tile (what the bulk store moves)
    static constexpr int kBuf = kIn > kOut ? kIn : kOut;       // Humanoid: 109 inputs < 110 outputs; Ant: 61 > 60
    static_assert(oForce % 4 == 0 && oAct % 4 == 0 && oSens % 4 == 0 && oRoot % 4 == 0, "16-byte aligned regions");
    static_assert((T * OBS) % 4 == 0 && S::kSensor % 4 == 0 && OBS % 2 == 0, "vector accesses");
};

__device__ __forceinline__ void warp_copy(float* dst, const float* src, int cnt, int lane) {
    for (int i = lane; i < cnt; i += 32) dst[i] = src[i];
}

// reset_idx :253-279 for env `ne` (lane = DOF): draws the noise, writes the new (pos, vel) pairs and the initial root
// row to GLOBAL memory and returns this lane's values so that the caller can patch its staged copy.
// `rank` = index of this env among the envs being reset (COMPACT noise rows follow the reference's nonzero() order).
// split in two so that a tile with several flagged envs can request ALL their rows first (one DRAM round trip, not one
// per env) and only then draw the noise and write
template <int D>
__device__ __forceinline__ void reset_env_fetch(const KParams& p, int ne, int lane, float* idp, float* rv) {
    *idp = (lane < D) ? p.init_dof_pos[static_cast<size_t>(ne) * D + lane] : 0.0f;
    *rv = (lane < kRoot) ? p.init_root[static_cast<size_t>(ne) * kRoot + lane] : 0.0f;   // :265-267 root row <- initial_root_states
}
template <int D>   // reset_idx of humanoid.py:253-279 and ant.py:247-271 (identical bodies)
__device__ __forceinline__ void reset_env_apply(const KParams& p, int ne, int rank, int lane, float idp, float rv, float2* pv_out,
                                                unsigned long long step_now) {
    const long long row = (p.noise_mode == TK_NOISE_COMPACT) ? rank : ne;
    float2 pv = make_float2(0.0f, 0.0f);
    if (lane < D) {
        const int j = lane;
        float ua = 0.5f, ub = 0.5f;
        if (p.noise_mode == TK_NOISE_PHILOX) {
            Philox4 r = philox_env(p.seed, step_now, p.env_off + ne, static_cast<uint32_t>(j >> 1));
            ua = u01((j & 1) ? r.z : r.x);
            ub = u01((j & 1) ? r.w : r.y);
        } else if (p.noise_mode != TK_NOISE_NONE && row < p.noise_rows) {
            ua = p.unit_a[row * D + j];
            ub = p.unit_b[row * D + j];
        }
        const float positions = 0.4f * ua + (-0.2f);                   // :258 torch_rand_float(-0.2, 0.2)
        const float velocities = 0.2f * ub + (-0.1f);                  // :259 torch_rand_float(-0.1, 0.1)
        float np_ = idp + positions;
        np_ = fmaxf(fminf(np_, p.hi[j]), p.lo[j]);                     // :261 tensor_clamp
        pv = make_float2(np_, velocities);
        reinterpret_cast<float2*>(p.dof_state)[static_cast<size_t>(ne) * D + j] = pv;   // :261-262
    }
    if (lane < kRoot) p.root[static_cast<size_t>(ne) * kRoot + lane] = rv;
    *pv_out = pv;
}
template <int D>
__device__ __forceinline__ void reset_env_values(const KParams& p, int ne, int rank, int lane, float2* pv_out, float* root_out,
                                                 unsigned long long step_now) {
    float idp, rv;
    reset_env_fetch<D>(p, ne, lane, &idp, &rv);    // both global reads are requested before the noise is computed
    reset_env_apply<D>(p, ne, rank, lane, idp, rv, pv_out, step_now);
    *root_out = rv;
}

// The Philox step of this launch, resolved AHEAD of griddepcontrol.wait: the device-side noise epoch (CUDA-graph replays) is
// written only by tk_humanoid_advance_epoch's single-thread kernel, a plain launch that never triggers its dependents
// early -- whatever follows it in the stream starts after it has completed, and this grid starts no earlier than its own
// predecessor.  (Read inside the reset path it was a dependent DRAM round trip on exactly the tiles that end a thin launch.)
__device__ __forceinline__ unsigned long long step_ahead_of_wait(const KParams& p) {
    unsigned long long step_now = p.step;
    if (p.noise_mode == TK_NOISE_PHILOX && p.step_dev) {
        unsigned long long ep;
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(ep) : "l"(p.step_dev) : "memory");
        step_now = p.step + (ep << 32);
    }
    return step_now;
}
// Ahead of the wait as well (p.warm, host decides by grid shape): bit 0 -- one line of every array a reset reads late or the
// tile writes, not for the data but for the TRANSLATION (bench.py rotates its state sets over more than the 256 MB the TLB
// reaches: the first touch of a page is a walk of several dependent DRAM reads, and a store that misses holds back the
// completion of the grid); bit 1 -- the tile's rows of initial_dof_pos / initial_root_states in full (+136 B / env of
// reads: thin grids only, where the launch is bound by the life of its slowest tile -- one holding a reset -- not by bytes).
template <int D, int OBS>
__device__ __forceinline__ void warm_ahead_of_wait(const KParams& p, int env0, int t, bool fused) {
    if (!p.warm) return;
    if (p.warm & 2) {
        constexpr int kDofLines = (kTile * D * 4 + 127) / 128, kRootLines = (kTile * kRoot * 4 + 127) / 128;
        if (fused && t < kDofLines) prefetch_l2(reinterpret_cast<const char*>(p.init_dof_pos + static_cast<size_t>(env0) * D) + 128 * t);
        if (fused && t < kRootLines) prefetch_l2(reinterpret_cast<const char*>(p.init_root + static_cast<size_t>(env0) * kRoot) + 128 * t);
    }
    if (p.warm & 1) {
        const void* a = nullptr;
        switch (t) {
            case 16: a = fused ? p.init_dof_pos + static_cast<size_t>(env0) * D : nullptr; break;
            case 17: a = fused ? p.init_root + static_cast<size_t>(env0) * kRoot : nullptr; break;
            case 18: a = p.obs + static_cast<size_t>(env0) * OBS; break;
            case 19: a = fused ? p.rew + env0 : nullptr; break;
            case 20: a = p.prev_pot + env0; break;
            case 21: a = (fused && p.timeout) ? p.timeout + env0 : nullptr; break;
            case 22: a = p.obs2 ? p.obs2 + static_cast<size_t>(env0) * OBS : nullptr; break;
            default: break;
        }
        if (a) prefetch_l2(a);
    }
}

#ifndef TK_HUMANOID_MIN_CTAS
#define TK_HUMANOID_MIN_CTAS 12
#endif

#ifdef TK_HUMANOID_MAXNREG      // developer knob: a register cap instead of a CTA-count bound (the two cannot be combined)
#define TK_HUMANOID_BOUNDS __maxnreg__(TK_HUMANOID_MAXNREG)
#else
#define TK_HUMANOID_BOUNDS __launch_bounds__(kTile, TK_HUMANOID_MIN_CTAS)
#endif
template <class S, bool FUSED>
__global__ void TK_HUMANOID_BOUNDS humanoid_warp_kernel(const __grid_constant__ KParams p) {
    using L = S;
    using WT = WarpTile<S>;
    constexpr int D = S::D;
    constexpr int kSensor = S::kSensor;
    constexpr int OBS = L::kObs;
    __shared__ __align__(128) float buf[WT::kBuf];
    __shared__ __align__(8) uint64_t bar;

    const int lane = threadIdx.x;
    const int env0 = blockIdx.x * kTile;
    const int nt = min(kTile, p.n - env0);
    const int n = env0 + lane;
    const bool valid = lane < nt;
    const bool bulk = (p.vec_ok != 0) && (nt == kTile);   // full, 16-byte aligned tile -> TMA both ways
    long long* ts = p.dbg_ts ? p.dbg_ts + static_cast<size_t>(blockIdx.x) * 8 : nullptr;
#define TK_STAMP(i) do { if (ts && lane == 0) { long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ts[i] = g; } } while (0)
    // programmatic dependent launch: let the next kernel in the stream start scheduling, and do not touch global
    // memory before the previous kernel has completed and flushed (no-ops when launched without the attribute)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const float* g_dof = p.dof_state + static_cast<size_t>(env0) * D * 2;
    const float* g_force = S::kHasForce ? p.dof_force + static_cast<size_t>(env0) * D : nullptr;
    const float* g_act = p.actions + static_cast<size_t>(env0) * D;
    const float* g_sens = p.sensors + static_cast<size_t>(env0) * kSensor;
    const float* g_root = p.root + static_cast<size_t>(env0) * kRoot;

    if (bulk && lane == 0) mbar_init(&bar, 1);
    // L2 prefetch of this tile's inputs, issued BEFORE griddepcontrol.wait: under programmatic dependent launch this
    // CTA becomes resident while the previous kernel in the stream is still draining, and the HBM reads of its 14 KB
    // overlap that drain instead of starting after it.  A prefetch consumes nothing, so stream order is intact: L2 is
    // the point of coherence and any later write of the previous kernel lands in the resident line; the real loads
    // below still come after the wait.
    if (bulk && p.prefetch == 1) {          // through the bulk-copy engine
        if (lane == 0) bulk_prefetch_l2(g_dof, kTile * D * 2 * 4);
        else if (S::kHasForce && lane == 1) bulk_prefetch_l2(g_force, kTile * D * 4);
        else if (lane == 2) bulk_prefetch_l2(g_act, kTile * D * 4);
        else if (lane == 3) bulk_prefetch_l2(g_root, kTile * kRoot * 4);
        else if (lane == 4) bulk_prefetch_l2(g_sens, kTile * kSensor * 4);
    } else if (bulk && p.prefetch == 2 && blockIdx.x < static_cast<unsigned>(p.prefetch_ctas)) {
        // through the load/store unit, one 128-byte line per instruction, first-wave CTAs only (later CTAs start
        // after the dependency has resolved: a prefetch would only add requests)
        auto pf = [&](const float* base, int bytes) {
            for (int o = lane * 128; o < bytes; o += 32 * 128) prefetch_l2(reinterpret_cast<const char*>(base) + o);
        };
        pf(g_dof, kTile * D * 2 * 4);
        if (S::kHasForce) pf(g_force, kTile * D * 4);
        pf(g_act, kTile * D * 4);
        pf(g_root, kTile * kRoot * 4);
        pf(g_sens, kTile * kSensor * 4);
    }
    if (bulk && p.prefetch) {
        if (lane == 5) prefetch_l2(p.pot + env0);
        else if (FUSED && lane == 6) prefetch_l2(p.progress + env0);
        else if (FUSED && lane == 7) prefetch_l2(p.progress + env0 + 16);
        else if (FUSED && lane == 8) prefetch_l2(p.reset + env0);
        else if (FUSED && lane == 9) prefetch_l2(p.reset + env0 + 16);
        else if (FUSED && lane == 10 && p.ep_return) prefetch_l2(p.ep_return + env0);
    }
    // touch both ends of the per-DOF constant tables so that their constant-cache lines are fetched now, not in the
    // middle of the math (the result only feeds a never-true predicate)
    const float warm = p.lo[0] + p.lo[D - 1] + p.hi[0] + p.hi[D - 1] + p.rinv[0] + p.rinv[D - 1] + p.dt + p.tgt[0];
    if (bulk) warm_ahead_of_wait<D, OBS>(p, env0, lane, FUSED);
    const unsigned long long step_now = FUSED ? step_ahead_of_wait(p) : 0ull;
    // everything above (parameter fetch, address arithmetic, mbarrier init) may overlap the drain of the previous kernel
    asm volatile("griddepcontrol.wait;" ::: "memory");
    TK_STAMP(0);
    if (warm == -1.2345e-30f && p.dbg_ts) p.dbg_ts[0] = 0;
    // First-wave stagger.  CTAs of the first wave all start at once, so without this every SM runs load-all / math-all /
    // store-all phases in lockstep and HBM idles during the math; later CTAs start as slots free up and are staggered
    // by construction.  Resident slot s (~ blockIdx / #SMs) of the first wave waits s * stagger_cycles.
    if (p.stagger_cycles > 0 && blockIdx.x < static_cast<unsigned>(p.stagger_ctas)) {
        const long long until = clock64() + static_cast<long long>(max(0, static_cast<int>(blockIdx.x / p.sm_count) - p.stagger_free)) * p.stagger_cycles;
        while (clock64() < until) __nanosleep(32);
    }
    if (bulk) {
        if (lane == 0) {
            fence_proxy_async_smem();      // the init (generic proxy) must be visible to the bulk-copy engine (async proxy)
            mbar_expect_tx(&bar, WT::kIn * 4);
        }
        __syncwarp();
        // one bulk load per lane: issuing a cp.async.bulk holds the issuing thread for ~0.25 us (tools/tile_timeline.py),
        // five lanes pay that once instead of one lane paying it five times
        if (lane == 0) bulk_load(buf + WT::oDof, g_dof, kTile * D * 2 * 4, &bar);
        else if (S::kHasForce && lane == 1) bulk_load(buf + WT::oForce, g_force, kTile * D * 4, &bar);
        else if (lane == 2) bulk_load(buf + WT::oAct, g_act, kTile * D * 4, &bar);
        else if (lane == 3) bulk_load(buf + WT::oRoot, g_root, kTile * kRoot * 4, &bar);
        else if (lane == 4) bulk_load(buf + WT::oSens, g_sens, kTile * kSensor * 4, &bar);
    } else {
        warp_copy(buf + WT::oDof, g_dof, nt * D * 2, lane);
        if (S::kHasForce) warp_copy(buf + WT::oForce, g_force, nt * D, lane);
        warp_copy(buf + WT::oAct, g_act, nt * D, lane);
        warp_copy(buf + WT::oSens, g_sens, nt * kSensor, lane);
        warp_copy(buf + WT::oRoot, g_root, nt * kRoot, lane);
    }

    // ---- per-env scalars (coalesced), overlapped with the bulk loads
    long long prog = 0;
    bool do_reset = false;
    float pot_old = 0.0f, ret_old = 0.0f;
    if (valid) {
        pot_old = p.pot[n];
        if (FUSED) {
            prog = p.progress[n] + 1;                              // humanoid.py:288
            do_reset = p.reset[n] != 0 && !p.dbg_reset;                            // :291
            if (p.ep_return) ret_old = p.ep_return[n];
            if (p.randomize) p.randomize[n] += 1;                  // :289
        }
    }
    TK_STAMP(1);

    // ---- reset prologue (rare): reset_idx :253-279 for this tile's flagged envs, lane = DOF.  The new state does not
    // depend on the staged tile, so for the first kPre flagged envs it is computed and written to global memory WHILE
    // the bulk loads are still in flight; only the patch of the staged copy waits for them.
    // Why the overlap of these global writes with the in-flight bulk load of the SAME rows is benign: (1) the bulk copy may
    // deliver the old or the new bytes of a flagged env's rows into shared memory, or a mix -- and exactly those staged rows
    // (all D (pos, vel) pairs and all 13 root columns of every flagged env) are overwritten after the mbarrier wait with the
    // values that were written to global memory, so nothing downstream ever reads what the copy delivered for them; (2) the
    // rows of an env belong to ONE tile, no other CTA of this launch reads or writes them; (3) the next launch is ordered
    // behind this one by the stream (griddepcontrol.wait).  tests: test_fused_every_env_of_a_tile_resets (all 32 envs of a
    // tile flagged, COMPACT and Philox noise), test_quad_kernel_matches_warp_kernel.
    constexpr int kPre = 4;                // 1 % of the envs resetting: 0.4 % of the tiles have 3+ resets, 0.003 % have 5+
    unsigned reset_bits = 0;
    int rank0 = 0;
    float2 pre_pv[kPre];
    float pre_root[kPre];
    if (FUSED) {
        reset_bits = __ballot_sync(0xffffffffu, do_reset);
        if (reset_bits) {
            rank0 = (p.noise_mode == TK_NOISE_COMPACT) ? p.tile_rank_base[blockIdx.x] : 0;
            // the rows of ALL (up to kPre) flagged envs are requested first: one DRAM round trip instead of one per env
            // (tiles with 3-4 resets used to finish them one after the other behind the wait and were the last ~1 us of
            // every launch: tools/tile_timeline.py prints the stragglers)
            float idp[kPre];
            unsigned bits = reset_bits;
#pragma unroll
            for (int q = 0; q < kPre; ++q) {
                idp[q] = 0.0f; pre_root[q] = 0.0f;
                if (bits) {
                    const int e = __ffs(bits) - 1;
                    bits &= bits - 1;
                    reset_env_fetch<D>(p, env0 + e, lane, &idp[q], &pre_root[q]);
                }
            }
            bits = reset_bits;
#pragma unroll
            for (int q = 0; q < kPre; ++q) {
                if (bits) {
                    const int e = __ffs(bits) - 1;
                    bits &= bits - 1;
                    reset_env_apply<D>(p, env0 + e, rank0 + q, lane, idp[q], pre_root[q], &pre_pv[q], step_now);
                }
            }
        }
    }
    TK_STAMP(2);
    __syncwarp();                          // mbarrier init / cooperative copies visible to the warp
    if (bulk) mbar_wait(&bar, 0);          // bulk loads landed
    TK_STAMP(3);

    if (FUSED && reset_bits) {             // patch the staged rows
        unsigned bits = reset_bits;
#pragma unroll
        for (int q = 0; q < kPre; ++q) {
            if (bits) {
                const int e = __ffs(bits) - 1;
                bits &= bits - 1;
                if (lane < D) reinterpret_cast<float2*>(buf + WT::oDof)[e * D + lane] = pre_pv[q];
                if (lane < kRoot) buf[WT::oRoot + e * kRoot + lane] = pre_root[q];
            }
        }
        for (int q = kPre; bits; ++q) {    // more than kPre resets in one tile: the rest after the wait
            const int e = __ffs(bits) - 1;
            bits &= bits - 1;
            float2 pv;
            float rv;
            reset_env_values<D>(p, env0 + e, rank0 + q, lane, &pv, &rv, step_now);
            if (lane < D) reinterpret_cast<float2*>(buf + WT::oDof)[e * D + lane] = pv;
            if (lane < kRoot) buf[WT::oRoot + e * kRoot + lane] = rv;
        }
        __syncwarp();
    }

    // ---- lane e pulls ITS env's inputs into registers
    float o[OBS];
    float r[kRoot];
    {
        const float2* dof2 = reinterpret_cast<const float2*>(buf + WT::oDof) + lane * D;
        const float* fo = buf + WT::oForce + lane * D;
        const float* ac = buf + WT::oAct + lane * D;
        const float4* se = reinterpret_cast<const float4*>(buf + WT::oSens) + lane * (kSensor / 4);
        const float* rr = buf + WT::oRoot + lane * kRoot;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            const float2 pv = dof2[j];
            o[L::kPos + j] = pv.x;
            o[L::kVel + j] = pv.y;
            if (S::kHasForce) o[L::kForce + j] = fo[j];
            o[L::kAct + j] = ac[j];
        }
#pragma unroll
        for (int k = 0; k < kSensor / 4; ++k) {
            const float4 v = se[k];
            o[L::kSens + 4 * k + 0] = v.x; o[L::kSens + 4 * k + 1] = v.y;
            o[L::kSens + 4 * k + 2] = v.z; o[L::kSens + 4 * k + 3] = v.w;
        }
#pragma unroll
        for (int c = 0; c < kRoot; ++c) r[c] = rr[c];
    }
    __syncwarp();   // the whole tile is in registers: from here on the buffer is the obs image
    TK_STAMP(4);

    float st_rew = 0.f, st_ret = 0.f, st_len = 0.f, st_up = 0.f, st_head = 0.f;
    bool out_reset = false, out_timeout = false;
    if (valid) {
        // ---- scaled copies (humanoid.py:406, 413-414 | ant.py:389, 402-404), constants straight from the parameter bank
        const float dvs = p.dof_vel_scale, cfs = p.contact_scale;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            // Ant counts `dof_pos_scaled > 0.99` in its reward (ant.py:347): the true IEEE quotient keeps that
            // threshold bit-exact; Humanoid's live reward never reads these columns -> reciprocal multiply
            o[L::kPos + j] = (S::kVariant == kVarAnt) ? unscale_ref(o[L::kPos + j], p.lo[j], p.hi[j])
                                                      : unscale_fast(o[L::kPos + j], p.lo[j], p.hi[j], p.rinv[j]);
            o[L::kVel + j] *= dvs;
            if (S::kHasForce) o[L::kForce + j] *= cfs;
        }
#pragma unroll
        for (int k = 0; k < kSensor; ++k) o[L::kSens + k] *= cfs;

        // ---- torso math
        const Head h = torso_math<S::kVariant>(r, p);
        // after a reset potentials = prev_potentials = -|targets - initial_pos| / dt (:273-276); the root row now IS the
        // initial row, so that value equals the freshly computed potential
        const float prev_new = (FUSED && do_reset) ? h.pot : pot_old;              // :394
#pragma unroll
        for (int c = 0; c < 12; ++c) o[c] = h.c[c];
        if (S::kVariant == kVarHumanoid) {
            o[L::kPot] = h.pot / 60000.0f;                                         // :414
            o[L::kPot + 1] = prev_new / 60000.0f;
        }

        float2* row = reinterpret_cast<float2*>(buf + lane * OBS);                 // 440-byte rows: 8-byte aligned
#pragma unroll
        for (int k = 0; k < OBS / 2; ++k) row[k] = make_float2(o[2 * k], o[2 * k + 1]);

        p.pot[n] = h.pot;
        p.prev_pot[n] = prev_new;
        if (p.up_vec) { p.up_vec[n * 3 + 0] = h.up.x; p.up_vec[n * 3 + 1] = h.up.y; p.up_vec[n * 3 + 2] = h.up.z; }
        if (p.heading_vec) {
            p.heading_vec[n * 3 + 0] = h.heading.x; p.heading_vec[n * 3 + 1] = h.heading.y; p.heading_vec[n * 3 + 2] = h.heading.z;
        }
        if (FUSED) {
            if (do_reset) prog = 0;                                                // :278 (and reset_buf -> 0, :279)
            const float up_reward = (h.c[10] > 0.93f) ? (0.0f + p.up_weight) : 0.0f;   // :349-350 | ant.py:341-342
            float total;
            if (S::kVariant == kVarAnt) {
                // compute_ant_reward, ant.py:337-354: every term live
                const float hp = h.c[11];
                const float heading_reward = (hp > 0.8f) ? (1.0f * p.heading_weight) : (p.heading_weight * hp / 0.8f);
                float actions_cost = 0.0f, electricity_cost = 0.0f;
                int at_limit = 0;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    const float a = o[L::kAct + j];
                    actions_cost += a * a;                                         // :345
                    electricity_cost += fabsf(a * o[L::kVel + j]);                 // :346 (obs[:, 20:28] = scaled dof_vel)
                    at_limit += (o[L::kPos + j] > 0.99f) ? 1 : 0;                  // :347
                }
                const float progress_reward = h.pot - prev_new;                    // :351
                total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                        p.energy_cost_scale * electricity_cost -
                        static_cast<float>(at_limit) * p.joints_cost_scale;        // :353-354
            } else {
                total = 2.0f + up_reward;                                          // humanoid.py:362,365
            }
            const bool fallen = h.c[0] < p.termination_height;
            if (fallen) total = 1.0f * p.death_cost;                               // :370 | ant.py:357
            const bool finished = static_cast<float>(prog) >= p.max_episode_length - 1.0f;   // :374
            out_reset = fallen || finished;                                        // :373-374 (reset_buf is 0 here)
            out_timeout = finished && out_reset;                                   // vec_task.py:345
            p.rew[n] = total;
            p.reset[n] = out_reset ? 1 : 0;
            p.progress[n] = prog;
            if (p.timeout) p.timeout[n] = out_timeout ? 1 : 0;
            st_rew = total; st_up = h.c[10]; st_head = h.c[11];
            if (p.ep_return) {
                const float ret = (do_reset ? 0.0f : ret_old) + total;
                if (out_reset) st_ret = ret;
                p.ep_return[n] = ret;
            }
            if (out_reset) st_len = static_cast<float>(prog);
        }
    }
    TK_STAMP(5);
    fence_proxy_async_smem();   // obs image was written through the generic proxy, the bulk engine reads it
    __syncwarp();

    // ---- obs tile -> global
    float* dst = p.obs + static_cast<size_t>(env0) * OBS;
    if (bulk) {
        if (lane == 0) {
            bulk_store(dst, buf, WT::kOut * 4);
            // second destination of the same tile (tk_humanoid_set_obs_copy): e.g. slot t of a [horizon, N, obs] rollout buffer
            if (p.obs2) bulk_store(p.obs2 + static_cast<size_t>(env0) * OBS, buf, WT::kOut * 4);
            bulk_commit();
        }
    } else {
        warp_copy(dst, buf, nt * OBS, lane);
        if (p.obs2) warp_copy(p.obs2 + static_cast<size_t>(env0) * OBS, buf, nt * OBS, lane);
    }

    // ---- statistics partials: tiles that share a slot (tile & 255) add with fp64 RED.ADD in arrival order, so the last bits of the
    //      sums can differ from run to run (counts are exact); tests compare them at rtol 1e-9
    if (FUSED) {
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
            // running sum since the last tk_humanoid_stats (slot scheme: tk_common.cuh)
            atomicAdd(p.partials + static_cast<size_t>(blockIdx.x & (kStatSlots - 1)) * TK_STATS + lane, static_cast<double>(v));
        }
    }
    if (bulk && lane == 0) bulk_wait_read<0>();   // shared memory must stay alive until the engine has read it
    TK_STAMP(6);
#undef TK_STAMP
}

}  // namespace humanoid
}  // namespace tk

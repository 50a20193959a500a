// Shared device/host helpers for the env-step kernels (sm_100a only).
//
// Floating-point policy: the translation units are compiled with -fmad=false, so `a*b+c`
// is two IEEE operations exactly like the separate torch CPU kernels of the reference.
// Where the reference's CPU kernels DO fuse (torch.cross, torch.norm -- measured, see
// DESIGN.md "Numerics") the fused form is written explicitly with fmaf().
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include "../../include/thormang_b200.h"

namespace tk {

// ---------------------------------------------------------------------------------------------
// host-side error plumbing
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
const char* last_error();

// Developer knobs (TK_* environment variables of the tuning tools).  They are read when a handle is CREATED and kept in
// it -- never on the launch path -- and a knob that is set announces itself once on stderr, so a stray variable cannot
// change a trainer's kernels unnoticed.  The *_DBG_RESET ablations, which switch the reset prologue off and thereby break
// parity with the reference, exist only in builds made with -DTK_DEVELOPER_KNOBS (TK_NVCC_EXTRA of build.py).
constexpr int kKnobUnset = -2147483647 - 1;
const char* knob(const char* name);                  // getenv + one warning per name
int knob_int(const char* name);                      // kKnobUnset when the variable is not set
int ablation_knob(const char* name);                 // 0 unless built with -DTK_DEVELOPER_KNOBS

#define TK_CUDA(call)                                                   \
    do {                                                                \
        cudaError_t _e = (call);                                        \
        if (_e != cudaSuccess) return tk::cuda_fail(_e, #call);         \
    } while (0)

#define TK_REQUIRE(cond, code, ...)                                     \
    do {                                                                \
        if (!(cond)) {                                                  \
            tk::set_error(__VA_ARGS__);                                 \
            return (code);                                              \
        }                                                               \
    } while (0)

static inline bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// ---------------------------------------------------------------------------------------------
// constants (fp32 values of the python doubles the reference passes to torch)
// ---------------------------------------------------------------------------------------------
constexpr float kPi = 3.14159265358979323846f;        // (float)np.pi
constexpr float kTwoPi = 6.28318530717958647692f;     // (float)(2*np.pi)

// ---------------------------------------------------------------------------------------------
// PTX wrappers: cache-hinted vector access, bulk (TMA) copies, mbarrier
// ---------------------------------------------------------------------------------------------
// read-only streaming loads (non-coherent path, no L1 allocation).  Not volatile: the compiler may batch them.
__device__ __forceinline__ float4 ldg_stream4(const float* p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
// same through the coherent path, evict-first: for arrays this kernel may also write (root_states, dof_state)
__device__ __forceinline__ float4 ld_stream4(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stg_stream4(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
// One whole 32-byte sector per lane in one request (sm_100 has 256-bit global accesses: STG.E.EF.ENL2.256); p 32-byte aligned
__device__ __forceinline__ void stg_stream8(float* p, const float* v) {
    asm volatile("st.global.cs.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
// 128-bit shared load the compiler may not narrow: a lane that uses two of the four words would otherwise get two 32-bit
// loads, and rows that are conflict-free as 128-bit accesses (odd multiples of 16 bytes apart) are 4-way conflicted as words
__device__ __forceinline__ float4 lds128(const float4* p) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))));
    return v;
}
// A 12-byte row at a 4-byte aligned address as one 64-bit and one 32-bit load (two requests instead of three)
__device__ __forceinline__ void ldg_row3(const float* row, float& x, float& y, float& z) {
    if ((reinterpret_cast<uintptr_t>(row) & 4u) == 0) {
        const float2 t = *reinterpret_cast<const float2*>(row);
        x = t.x; y = t.y; z = row[2];
    } else {
        x = row[0];
        const float2 t = *reinterpret_cast<const float2*>(row + 1);
        y = t.x; z = t.y;
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// generic-proxy writes to shared memory -> visible to the async (TMA) proxy
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// shared -> global bulk copy (UBLKCP); dst/src 16-byte aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// global -> shared bulk copy completing on an mbarrier
__device__ __forceinline__ void bulk_load(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// L2 prefetch of a contiguous global range through the bulk-copy engine (no destination, no completion to wait for);
// src 16-byte aligned, bytes % 16 == 0.  A hint only: L2 is the point of coherence, later writes by other kernels
// update the resident lines.
__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* gsrc) { asm volatile("prefetch.global.L2 [%0];" ::"l"(gsrc)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}

// ---------------------------------------------------------------------------------------------
// quaternion / angle arithmetic, operation order = oracle/isaac_helpers.py (restated
// isaacgym.torch_utils) evaluated by torch's CPU kernels
// ---------------------------------------------------------------------------------------------
struct V3 { float x, y, z; };
struct Q4 { float x, y, z, w; };

// torch.cross CPU kernel: a1*b2 - a2*b1 compiles to fma(a1, b2, -(a2*b1))
__device__ __forceinline__ V3 cross_ref(V3 a, V3 b) {
    V3 r;
    r.x = fmaf(a.y, b.z, -(a.z * b.y));
    r.y = fmaf(a.z, b.x, -(a.x * b.z));
    r.z = fmaf(a.x, b.y, -(a.y * b.x));
    return r;
}
// torch.bmm [N,1,3]x[N,3,1]: (a0*b0 + a1*b1) + a2*b2, unfused
__device__ __forceinline__ float dot_ref(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// torch.norm(p=2, dim=-1) on 3 elements: sqrt(fma(z,z,fma(y,y,x*x)))
__device__ __forceinline__ float norm3_ref(V3 a) { return sqrtf(fmaf(a.z, a.z, fmaf(a.y, a.y, a.x * a.x))); }

// quat_mul, 9-multiply factorisation
__device__ __forceinline__ Q4 quat_mul_ref(Q4 a, Q4 b) {
    float ww = (a.z + a.x) * (b.x + b.y);
    float yy = (a.w - a.y) * (b.w + b.z);
    float zz = (a.w + a.y) * (b.w - b.z);
    float xx = ww + yy + zz;
    float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
    Q4 r;
    r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
    r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
    r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
    r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
    return r;
}

// quat_rotate (sign = +1) / quat_rotate_inverse (sign = -1): a +- b + c
template <int SIGN>
__device__ __forceinline__ V3 quat_rotate_ref(Q4 q, V3 v) {
    V3 qv{q.x, q.y, q.z};
    float s = 2.0f * (q.w * q.w) - 1.0f;
    V3 cr = cross_ref(qv, v);
    float d = dot_ref(qv, v);
    V3 r;
    float ax = v.x * s, ay = v.y * s, az = v.z * s;
    float bx = cr.x * q.w * 2.0f, by = cr.y * q.w * 2.0f, bz = cr.z * q.w * 2.0f;
    float cx = qv.x * d * 2.0f, cy = qv.y * d * 2.0f, cz = qv.z * d * 2.0f;
    if (SIGN > 0) {
        r.x = ax + bx + cx; r.y = ay + by + cy; r.z = az + bz + cz;
    } else {
        r.x = ax - bx + cx; r.y = ay - by + cy; r.z = az - bz + cz;
    }
    return r;
}

// torch `x % (2*pi)` for x = atan2(..) in [-pi, pi]: fmod is the identity there, the floor-mod
// fix-up adds the divisor to negative non-zero values (result can round to exactly 2*pi)
__device__ __forceinline__ float mod_two_pi_of_atan2(float x) { return (x != 0.0f && x < 0.0f) ? x + kTwoPi : x; }
// general torch floor-mod by 2*pi
__device__ __forceinline__ float floor_mod_two_pi(float a) {
    float m = fmodf(a, kTwoPi);
    if (m != 0.0f && m < 0.0f) m += kTwoPi;
    return m;
}
// same result, cheap for |a| < 4*pi (all the hot path produces): fmod is the identity below 2*pi and an exact
// subtraction (Sterbenz) in [2*pi, 4*pi); anything else takes the library fmodf
__device__ __forceinline__ float floor_mod_two_pi_near(float a) {
    float m;
    if (fabsf(a) < kTwoPi) m = a;
    else if (a >= kTwoPi && a < 2.0f * kTwoPi) m = a - kTwoPi;
    else m = fmodf(a, kTwoPi);
    if (m != 0.0f && m < 0.0f) m += kTwoPi;
    return m;
}
__device__ __forceinline__ float normalize_angle_ref(float x) {
    float s, c;
    sincosf(x, &s, &c);
    return atan2f(s, c);
}
// normalize_angle (atan2(sin x, cos x)) for |x| < 3*pi without the three transcendentals: wrap by +-2*pi.
// Differs from the sin/cos/atan2 sequence by the rounding of fl(2*pi) (1.7e-7) plus one ulp -- the same order as
// the libm-vs-CUDA difference of that sequence itself; at exactly +-pi the representative may differ by 2*pi
// (parity compares angles modulo 2*pi).
__device__ __forceinline__ float normalize_angle_fast(float x) {
    x = (x > kPi) ? x - kTwoPi : x;
    x = (x < -kPi) ? x + kTwoPi : x;
    return x;
}
// roll and yaw of get_euler_xyz (pitch is never consumed on the hot path)
__device__ __forceinline__ void euler_roll_yaw_ref(Q4 q, float* roll, float* yaw) {
    float sinr_cosp = 2.0f * (q.w * q.x + q.y * q.z);
    float cosr_cosp = q.w * q.w - q.x * q.x - q.y * q.y + q.z * q.z;
    float siny_cosp = 2.0f * (q.w * q.z + q.x * q.y);
    float cosy_cosp = q.w * q.w + q.x * q.x - q.y * q.y - q.z * q.z;
    *roll = mod_two_pi_of_atan2(atan2f(sinr_cosp, cosr_cosp));
    *yaw = mod_two_pi_of_atan2(atan2f(siny_cosp, cosy_cosp));
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 counter RNG (in-kernel reset / sensor noise); u in [0,1)
// ---------------------------------------------------------------------------------------------
struct Philox4 { uint32_t x, y, z, w; };
__device__ __host__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
    return static_cast<uint32_t>((static_cast<uint64_t>(a) * b) >> 32);
}
__device__ __host__ inline Philox4 philox4x32_10(uint64_t seed, uint64_t ctr_lo, uint64_t ctr_hi) {
    uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
    Philox4 c{static_cast<uint32_t>(ctr_lo), static_cast<uint32_t>(ctr_lo >> 32), static_cast<uint32_t>(ctr_hi),
              static_cast<uint32_t>(ctr_hi >> 32)};
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = mulhi32(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = mulhi32(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        Philox4 n{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
        c = n;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __host__ __forceinline__ float u01(uint32_t bits) { return static_cast<float>(bits >> 8) * (1.0f / 16777216.0f); }
// standard normal pair by Box-Muller from two 32-bit words.  Hardware-approximate log / sin / cos: this is noise,
// and parity checks materialise the very same values through tk_*_philox_noise.
__device__ __forceinline__ float2 normal_pair(uint32_t a, uint32_t b) {
    float u1 = (static_cast<float>(a >> 8) + 1.0f) * (1.0f / 16777216.0f);   // (0,1]
    float u2 = u01(b);
    float r = sqrtf(-2.0f * __logf(u1));
    float s, c;
    __sincosf(kTwoPi * u2, &s, &c);
    return make_float2(r * c, r * s);
}
// counter layout: ctr_lo = global env id, ctr_hi = (step << 8) | stream-id
__device__ __host__ __forceinline__ Philox4 philox_env(uint64_t seed, uint64_t step, int64_t env, uint32_t stream_id) {
    return philox4x32_10(seed, static_cast<uint64_t>(env), (step << 8) | stream_id);
}

// ---------------------------------------------------------------------------------------------
// Noise epoch.  The Philox streams are keyed by (seed, step, global env id); `step` comes from the host with every
// call, so it is frozen inside a captured CUDA graph.  A device-side epoch, bumped by tk_*_advance_epoch (one
// single-thread kernel, captured once per graph), is folded into the upper 32 bits of the step: every replay of the
// graph draws fresh noise.  (A per-launch device counter was tried first: one same-address RED per CTA cost 20 % of
// the kernel time at 32768 CTAs.)
// ---------------------------------------------------------------------------------------------
// `epoch` may be NULL: a handle whose epoch was never advanced passes NULL, so that eager launches neither read global
// memory for it nor make their first Philox block wait on that read.
__device__ __forceinline__ unsigned long long effective_step(unsigned long long step, const unsigned long long* epoch) {
    return epoch ? step + (*epoch << 32) : step;
}
__global__ void advance_epoch_kernel(unsigned long long* epoch);

// ---------------------------------------------------------------------------------------------
// Episode-statistics partials: kStatSlots rows of TK_STATS fp64 accumulators per handle.  A tile adds its eight warp
// sums to row (tile & (kStatSlots-1)) with RED.ADD.F64 (no return value); tk_*_stats sums the rows in a fixed order
// (the adds INTO a row arrive in scheduling order: sums are reproducible to fp64 rounding, not bit for bit).
// 256 rows keep same-address contention negligible (one address hit by every CTA cost 20 % at 32768 CTAs) and keep
// the finalize kernel at ~2 us (a row per tile made it a 40 us single-CTA walk over 4096 rows).  fp64 adds of
// warp-level fp32 sums: the order of the REDs within a row is not fixed, results agree to ~1e-13 relative.
// ---------------------------------------------------------------------------------------------
constexpr int kStatSlots = 256;

// ---------------------------------------------------------------------------------------------
// warp helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace tk
